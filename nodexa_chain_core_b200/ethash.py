"""Mirror of the reference's `ethash::` namespace for the KawPow path
(src/crypto/ethash/include/ethash/ethash.hpp:24-172, ethash.h:20-131)."""
from __future__ import annotations

import ctypes as C

from ._lib import check, inb, lib, outb

epoch_length = 7500  # ethash.h:29
light_cache_item_size = 64
full_dataset_item_size = 128


def get_epoch_number(block_number: int) -> int:
    """ethash.hpp:75-78"""
    return block_number // epoch_length if block_number else 0


def calculate_light_cache_num_items(epoch_number: int) -> int:
    return lib.nx_light_cache_num_items(epoch_number)


def calculate_full_dataset_num_items(epoch_number: int) -> int:
    return lib.nx_full_dataset_num_items(epoch_number)


def calculate_epoch_seed(epoch_number: int) -> bytes:
    o = outb(32)
    lib.nx_epoch_seed(epoch_number, o)
    return bytes(o)


def find_epoch_number(seed: bytes) -> int:
    """ethash.cpp:57-97; -1 when the seed is not found within 30000 epochs."""
    return lib.nx_find_epoch_number(inb(seed))


def keccak256(data: bytes) -> bytes:
    o = outb(32)
    lib.nx_keccak256(inb(data) if data else outb(1), len(data), o)
    return bytes(o)


def keccak512(data: bytes) -> bytes:
    o = outb(64)
    lib.nx_keccak512(inb(data) if data else outb(1), len(data), o)
    return bytes(o)


class EpochContext:
    """Owns one epoch's light cache, L1 cache and full DAG on one GPU.

    Public attributes follow struct ethash_epoch_context (ethash.h:35-42): epoch_number,
    light_cache_num_items, light_cache, l1_cache, full_dataset_num_items.  The reference
    distinguishes a light and a full context; on B200 both hold the resident DAG.
    """

    def __init__(self, epoch_number: int, device: int = 0, rank: int = 0, world: int = 1, deferred: bool = False,
                 light_cache: bytes = None, shell: bool = False):
        h = C.c_void_p()
        if shell:
            # allocations only: the light cache comes from build_light_cache() (one rank) or adopt_light_cache() (the others)
            check(lib.nx_epoch_create_shell(epoch_number, device, rank, world, C.byref(h)))
        elif deferred:
            # light_cache: the epoch's light cache if the caller already has it (one serial chain for several contexts)
            lc = C.cast(C.c_char_p(light_cache), C.c_void_p) if light_cache is not None else None
            check(lib.nx_epoch_create_deferred(epoch_number, device, rank, world, lc, C.byref(h)))
        elif world == 1:
            check(lib.nx_epoch_create(epoch_number, device, C.byref(h)))
        else:
            check(lib.nx_epoch_create_sharded(epoch_number, device, rank, world, C.byref(h)))
        self._h = h
        self.rank, self.world = rank, world

    # -- reference-visible fields --
    @property
    def epoch_number(self) -> int:
        return lib.nx_epoch_number(self._h)

    @property
    def light_cache_num_items(self) -> int:
        return lib.nx_epoch_light_items(self._h)

    @property
    def full_dataset_num_items(self) -> int:
        return lib.nx_epoch_dataset_items(self._h)

    @property
    def light_cache(self) -> bytes:
        return C.string_at(lib.nx_epoch_light_cache(self._h), self.light_cache_num_items * 64)

    @property
    def l1_cache(self):
        p = lib.nx_epoch_l1_cache(self._h)
        return [p[i] for i in range(4096)]

    # -- device side --
    @property
    def device(self) -> int:
        return lib.nx_epoch_device(self._h)

    @property
    def dag_ptr(self) -> int:
        return lib.nx_epoch_dag_ptr(self._h)

    @property
    def dag_bytes(self) -> int:
        return lib.nx_epoch_dag_bytes(self._h)

    @property
    def dag_padded_bytes(self) -> int:
        return lib.nx_epoch_dag_padded_bytes(self._h)

    @property
    def dag_share_bytes(self) -> int:
        return lib.nx_epoch_dag_share_bytes(self._h)

    @property
    def dag_num_items(self) -> int:
        """256-byte items addressable by KawPow = full_dataset_num_items // 2 (progpow.cpp:182)."""
        return self.dag_bytes // 256

    def build_times_ms(self):
        a, b = C.c_double(), C.c_double()
        lib.nx_epoch_build_times(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def ipc_handle(self) -> bytes:
        o = outb(64)
        check(lib.nx_epoch_ipc_handle(self._h, o))
        return bytes(o)

    def generate_fused_ipc(self, handles_by_rank=None) -> None:
        """handles_by_rank: list of `world` 64-byte blobs from every rank's ipc_handle() (own entry ignored);
        None once open_ipc_peers() has mapped them."""
        if handles_by_rank is None:
            check(lib.nx_epoch_generate_fused_ipc(self._h, None))
        else:
            check(lib.nx_epoch_generate_fused_ipc(self._h, inb(b"".join(handles_by_rank))))

    def generate_slice(self) -> None:
        """Shell / deferred contexts on the NCCL route: this rank's slice into its own buffer; all-gather + finalize follow."""
        check(lib.nx_epoch_generate_slice(self._h))

    def open_ipc_peers(self, handles_by_rank) -> None:
        check(lib.nx_epoch_open_ipc_peers(self._h, inb(b"".join(handles_by_rank))))

    def build_light_cache(self) -> None:
        """Shell contexts: run the serial chain here (the one rank of a job that does)."""
        check(lib.nx_epoch_build_light_cache(self._h))

    def adopt_light_cache(self) -> None:
        """Shell contexts: light_tensor() has been filled by a broadcast from the rank that ran the chain."""
        check(lib.nx_epoch_adopt_light_cache(self._h))

    def light_tensor(self):
        """The light cache on the device as a torch uint8 CUDA tensor aliasing the library's buffer (for NCCL)."""
        return self._alias_tensor(lib.nx_epoch_light_ptr(self._h), self.light_cache_num_items * 64)

    def generate_fused_peers(self, dag_ptrs_by_rank) -> None:
        """dag_ptrs_by_rank: `world` device pointers (dag_ptr of every rank's context in THIS process; own entry ignored)."""
        arr = (C.c_void_p * len(dag_ptrs_by_rank))(*[C.c_void_p(p) for p in dag_ptrs_by_rank])
        check(lib.nx_epoch_generate_fused_peers(self._h, arr))

    def finalize(self):
        """After the caller's in-place all-gather of a sharded DAG."""
        check(lib.nx_epoch_finalize(self._h))

    def dag_tensor(self):
        """The padded DAG as a torch uint8 CUDA tensor aliasing the library's buffer (for NCCL)."""
        return self._alias_tensor(self.dag_ptr, self.dag_padded_bytes)

    def _alias_tensor(self, ptr: int, nbytes: int):
        import torch

        class _Arr:
            pass

        a = _Arr()
        a.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        a._keepalive = self
        return torch.as_tensor(a, device=f"cuda:{self.device}")

    def read_items(self, first: int, count: int) -> bytes:
        o = outb(256 * count)
        check(lib.nx_dag_read_items(self._h, first, count, o))
        return bytes(o)

    def generate_range(self, first: int, count: int, out_device_ptr: int):
        check(lib.nx_dag_generate_range(self._h, first, count, C.c_void_p(out_device_ptr)))

    def checksum(self) -> int:
        v = C.c_uint64()
        check(lib.nx_dag_checksum(self._h, C.byref(v)))
        return v.value

    def last_timing(self):
        ms, n = C.c_float(), C.c_int()
        lib.nx_kawpow_last_timing(self._h, C.byref(ms), C.byref(n))
        return ms.value, n.value

    def close(self):
        if getattr(self, "_h", None):
            lib.nx_epoch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def create_epoch_context(epoch_number: int, device: int = 0) -> EpochContext:
    """ethash.hpp:112-115 / ethash.cpp:395-398"""
    return EpochContext(epoch_number, device)


def create_epoch_context_full(epoch_number: int, device: int = 0) -> EpochContext:
    """ethash.hpp:117-120 / ethash.cpp:400-403"""
    return EpochContext(epoch_number, device)


def hash(context: EpochContext, header_hash: bytes, nonce: int):  # noqa: A001
    """ethash::hash / ethash_hash (ethash.cpp:306-324,416-422): classic Ethash, returns (final_hash, mix_hash)."""
    from ._lib import NxResult
    r = NxResult()
    check(lib.nx_ethash_hash(context._h, inb(header_hash), nonce, C.byref(r)))
    return bytes(r.final_hash), bytes(r.mix_hash)


def hash_batch(context: EpochContext, header_hash: bytes, start_nonce: int, count: int) -> bytes:
    out = outb(64 * count)
    check(lib.nx_ethash_hash_batch(context._h, inb(header_hash), start_nonce, count, out))
    return bytes(out)


def verify_final_hash(header_hash: bytes, mix_hash: bytes, nonce: int, boundary: bytes) -> bool:
    """ethash_verify_final_hash (ethash.cpp:424-429)"""
    return bool(lib.nx_ethash_verify_final_hash(inb(header_hash), inb(mix_hash), nonce, inb(boundary)))


def verify(context: EpochContext, header_hash: bytes, mix_hash: bytes, nonce: int, boundary: bytes) -> bool:
    """ethash_verify (ethash.cpp:431-440): boundary test first, then the mix is recomputed."""
    if not verify_final_hash(header_hash, mix_hash, nonce, boundary):
        return False
    return hash(context, header_hash, nonce)[1] == mix_hash
