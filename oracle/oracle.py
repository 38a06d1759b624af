"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the CPU oracle.

`Oracle`  -> oracle/libnxoracle.so  (our plain-C restatement, oracle/nx_oracle.c; kind "port")
`Ref`     -> oracle/_ref/libnxref.so (the reference's own sources compiled in place by
             `make -C oracle ref`; kind "reference").  Present only where it was built
             (this container; it travels to the GPU box as a prebuilt .so).

Nothing in nodexa_chain_core_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libnxoracle.so")
REF_SO = os.path.join(HERE, "_ref", "libnxref.so")

u8p = C.POINTER(C.c_uint8)


def build_oracle() -> str:
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    return ORACLE_SO


def build_ref() -> str | None:
    """Compile the reference TUs where they lie (only where /root/reference exists)."""
    if not os.path.isdir("/root/reference/src/crypto/ethash"):
        return REF_SO if os.path.exists(REF_SO) else None
    subprocess.check_call(["make", "-s", "-C", HERE, "ref"])
    return REF_SO


def _buf(n):
    return (C.c_uint8 * n)()


def _in(b: bytes):
    return (C.c_uint8 * len(b)).from_buffer_copy(b)


class _Ctx(C.Structure):
    _fields_ = [("epoch", C.c_int), ("light_items", C.c_int), ("dataset_items", C.c_int),
                ("light", C.c_void_p), ("l1", C.c_uint32 * 4096), ("dag", C.c_void_p)]


class Oracle:
    kind = "port"

    def __init__(self):
        srcs = [os.path.join(HERE, f) for f in ("nx_oracle.c", "nx_equihash_oracle.c")]
        if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(f) for f in srcs):
            build_oracle()
        L = self.L = C.CDLL(ORACLE_SO)
        L.nxo_ctx_create.restype = C.POINTER(_Ctx)
        L.nxo_ctx_create.argtypes = [C.c_int]
        L.nxo_ctx_destroy.argtypes = [C.POINTER(_Ctx)]
        L.nxo_ctx_set_dag.argtypes = [C.POINTER(_Ctx), C.c_void_p]
        L.nxo_kawpow_hash.argtypes = [C.POINTER(_Ctx), C.c_int, u8p, C.c_uint64, u8p, u8p]
        L.nxo_kawpow_hash_range.argtypes = [C.POINTER(_Ctx), C.c_int, u8p, C.c_uint64, C.c_uint64, u8p, C.c_int]
        L.nxo_kawpow_verify.argtypes = [C.POINTER(_Ctx), C.c_int, u8p, u8p, C.c_uint64, u8p]
        L.nxo_kawpow_hash_no_verify.argtypes = [C.c_int, u8p, u8p, C.c_uint64, u8p]
        L.nxo_kawpow_search_mt.argtypes = [C.POINTER(_Ctx), C.c_int, u8p, u8p, C.c_uint64, C.c_uint64,
                                           C.POINTER(C.c_uint64), u8p, u8p, C.c_int]
        L.nxo_kawpow_program.argtypes = [C.c_uint64, C.POINTER(C.c_uint32 * 6)]
        L.nxo_dataset_item_512.argtypes = [C.POINTER(_Ctx), C.c_int64, u8p]
        L.nxo_dataset_item_2048.argtypes = [C.POINTER(_Ctx), C.c_uint32, u8p]
        L.nxo_dataset_items_2048.argtypes = [C.POINTER(_Ctx), C.c_uint32, C.c_uint32, u8p, C.c_int]
        L.nxo_ethash_hash.argtypes = [C.POINTER(_Ctx), u8p, C.c_uint64, u8p, u8p]
        L.nxo_ethash_verify_final.argtypes = [u8p, u8p, C.c_uint64, u8p]
        L.nxo_epoch_seed.argtypes = [C.c_int, u8p]
        L.nxo_find_epoch_number.argtypes = [u8p]
        L.nxo_keccak256.argtypes = [u8p, C.c_size_t, u8p]
        L.nxo_keccak512.argtypes = [u8p, C.c_size_t, u8p]
        L.nxo_keccakf800.argtypes = [C.POINTER(C.c_uint32)]
        L.nxo_keccakf1600.argtypes = [C.POINTER(C.c_uint64)]
        L.nxo_sha256.argtypes = [u8p, C.c_size_t, u8p]
        L.nxo_kawpow_header_hash.argtypes = [C.c_int32, u8p, u8p, C.c_uint32, C.c_uint32, C.c_uint32, u8p]
        L.nxo_compact_to_boundary.argtypes = [C.c_uint32, u8p]
        L.nxo_less_or_equal.argtypes = [u8p, u8p]
        L.nxo_build_light_cache.argtypes = [u8p, C.c_int, u8p]
        # Equihash self-consistency oracle (nx_equihash_oracle.c; NOT reference-derived)
        L.nxeq_blake2b.argtypes = [u8p, C.c_size_t, u8p, u8p, C.c_size_t]
        L.nxeq_solution_bytes.restype = C.c_size_t
        L.nxeq_solution_bytes.argtypes = [C.c_int, C.c_int]
        L.nxeq_solve.argtypes = [C.c_int, C.c_int, u8p, u8p, C.c_size_t, u8p, C.c_int]
        L.nxeq_validate.argtypes = [C.c_int, C.c_int, u8p, u8p, C.c_size_t, u8p]
        L.nxeq_decode.argtypes = [C.c_int, C.c_int, u8p, C.POINTER(C.c_uint32)]
        L.nxeq_encode.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_uint32), u8p]
        self._ctx = {}

    # -- Equihash (self-consistency only) --
    def blake2b(self, data: bytes, personal: bytes, outlen: int) -> bytes:
        o = _buf(outlen)
        self.L.nxeq_blake2b(_in(data) if data else _buf(1), len(data), _in(personal), o, outlen)
        return bytes(o)

    def equihash_solution_bytes(self, n, k) -> int:
        return self.L.nxeq_solution_bytes(n, k)

    def equihash_solve(self, n, k, tag: bytes, header: bytes, max_solutions: int = 64):
        sb = self.equihash_solution_bytes(n, k)
        out = _buf(sb * max_solutions)
        cnt = self.L.nxeq_solve(n, k, _in(tag), _in(header), len(header), out, max_solutions)
        if cnt < 0:
            raise ValueError("bad Equihash parameters")
        raw = bytes(out)
        return [raw[i * sb:(i + 1) * sb] for i in range(min(cnt, max_solutions))]

    def equihash_validate(self, n, k, tag: bytes, header: bytes, solution: bytes) -> bool:
        return bool(self.L.nxeq_validate(n, k, _in(tag), _in(header), len(header), _in(solution)))

    def equihash_decode(self, n, k, solution: bytes):
        idx = (C.c_uint32 * (1 << k))()
        self.L.nxeq_decode(n, k, _in(solution), idx)
        return list(idx)

    # -- contexts are cached per epoch (light cache build is ~1 s) --
    def ctx(self, epoch: int):
        if epoch not in self._ctx:
            c = self.L.nxo_ctx_create(epoch)
            if not c:
                raise MemoryError("oracle context")
            self._ctx[epoch] = c
        return self._ctx[epoch]

    def light_items(self, epoch):
        return self.L.nxo_light_cache_num_items(epoch)

    def dataset_items(self, epoch):
        return self.L.nxo_full_dataset_num_items(epoch)

    def light_cache(self, epoch) -> bytes:
        c = self.ctx(epoch)
        return C.string_at(c.contents.light, c.contents.light_items * 64)

    def l1(self, epoch):
        return list(self.ctx(epoch).contents.l1)

    def set_dag(self, epoch, ptr):
        self.L.nxo_ctx_set_dag(self.ctx(epoch), ptr)

    def epoch_seed(self, epoch) -> bytes:
        o = _buf(32)
        self.L.nxo_epoch_seed(epoch, o)
        return bytes(o)

    def find_epoch_number(self, seed: bytes) -> int:
        return self.L.nxo_find_epoch_number(_in(seed))

    def keccak256(self, d: bytes) -> bytes:
        o = _buf(32)
        self.L.nxo_keccak256(_in(d) if d else _buf(1), len(d), o)
        return bytes(o)

    def keccak512(self, d: bytes) -> bytes:
        o = _buf(64)
        self.L.nxo_keccak512(_in(d) if d else _buf(1), len(d), o)
        return bytes(o)

    def keccakf800(self, words):
        a = (C.c_uint32 * 25)(*words)
        self.L.nxo_keccakf800(a)
        return list(a)

    def keccakf1600(self, words):
        a = (C.c_uint64 * 25)(*words)
        self.L.nxo_keccakf1600(a)
        return list(a)

    def kawpow_hash(self, epoch, height, header: bytes, nonce: int):
        f, m = _buf(32), _buf(32)
        self.L.nxo_kawpow_hash(self.ctx(epoch), height, _in(header), nonce, f, m)
        return bytes(f), bytes(m)

    def kawpow_hash_range(self, epoch, height, header: bytes, start: int, count: int, threads: int = 0) -> bytes:
        out = _buf(64 * count)
        self.L.nxo_kawpow_hash_range(self.ctx(epoch), height, _in(header), start, count, out,
                                     threads or os.cpu_count() or 1)
        return bytes(out)

    def kawpow_verify(self, epoch, height, header, mix, nonce, boundary) -> bool:
        return bool(self.L.nxo_kawpow_verify(self.ctx(epoch), height, _in(header), _in(mix), nonce, _in(boundary)))

    def kawpow_hash_no_verify(self, height, header, mix, nonce) -> bytes:
        o = _buf(32)
        self.L.nxo_kawpow_hash_no_verify(height, _in(header), _in(mix), nonce, o)
        return bytes(o)

    def kawpow_search(self, epoch, height, header, boundary, start, iterations, threads: int = 0):
        n = C.c_uint64(0)
        f, m = _buf(32), _buf(32)
        found = self.L.nxo_kawpow_search_mt(self.ctx(epoch), height, _in(header), _in(boundary), start, iterations,
                                            C.byref(n), f, m, threads or os.cpu_count() or 1)
        return bool(found), n.value, bytes(f), bytes(m)

    def kawpow_program(self, prog_seed: int):
        ops = ((C.c_uint32 * 6) * 33)()
        n = self.L.nxo_kawpow_program(prog_seed, ops)
        return [tuple(ops[i]) for i in range(n)]

    def dataset_item_512(self, epoch, index) -> bytes:
        o = _buf(64)
        self.L.nxo_dataset_item_512(self.ctx(epoch), index, o)
        return bytes(o)

    def dataset_item_2048(self, epoch, index) -> bytes:
        o = _buf(256)
        self.L.nxo_dataset_item_2048(self.ctx(epoch), index, o)
        return bytes(o)

    def dataset_items_2048(self, epoch, first, count, threads: int = 0) -> bytes:
        o = _buf(256 * count)
        self.L.nxo_dataset_items_2048(self.ctx(epoch), first, count, o, threads or os.cpu_count() or 1)
        return bytes(o)

    def ethash_hash(self, epoch, header, nonce):
        f, m = _buf(32), _buf(32)
        self.L.nxo_ethash_hash(self.ctx(epoch), _in(header), nonce, f, m)
        return bytes(f), bytes(m)

    def ethash_verify_final(self, header, mix, nonce, boundary) -> bool:
        return bool(self.L.nxo_ethash_verify_final(_in(header), _in(mix), nonce, _in(boundary)))

    def sha256(self, d: bytes) -> bytes:
        o = _buf(32)
        self.L.nxo_sha256(_in(d) if d else _buf(1), len(d), o)
        return bytes(o)

    def kawpow_header_hash(self, version, prev, merkle, time, bits, height) -> bytes:
        o = _buf(32)
        self.L.nxo_kawpow_header_hash(version, _in(prev), _in(merkle), time, bits, height, o)
        return bytes(o)

    def compact_to_boundary(self, bits):
        o = _buf(32)
        ok = self.L.nxo_compact_to_boundary(bits, o)
        return bool(ok), bytes(o)

    def less_or_equal(self, a, b) -> bool:
        return bool(self.L.nxo_less_or_equal(_in(a), _in(b)))


class Ref:
    """The reference's own compiled code (oracle/_ref/libnxref.so)."""
    kind = "reference"

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def __init__(self):
        if not os.path.exists(REF_SO):
            if build_ref() is None or not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO)
        L = self.L = C.CDLL(REF_SO)
        vp = C.c_void_p
        L.nxref_ctx_create.restype = vp
        L.nxref_ctx_create.argtypes = [C.c_int]
        L.nxref_ctx_destroy.argtypes = [vp]
        L.nxref_ctx_light_items.argtypes = [vp]
        L.nxref_ctx_dataset_items.argtypes = [vp]
        L.nxref_ctx_light_cache.restype = vp
        L.nxref_ctx_light_cache.argtypes = [vp]
        L.nxref_ctx_l1.restype = C.POINTER(C.c_uint32)
        L.nxref_ctx_l1.argtypes = [vp]
        L.nxref_kawpow_hash.argtypes = [vp, C.c_int, u8p, C.c_uint64, u8p, u8p]
        L.nxref_kawpow_hash_range.argtypes = [vp, C.c_int, u8p, C.c_uint64, C.c_uint64, u8p, C.c_int]
        L.nxref_kawpow_verify.argtypes = [vp, C.c_int, u8p, u8p, C.c_uint64, u8p]
        L.nxref_kawpow_hash_no_verify.argtypes = [C.c_int, u8p, u8p, C.c_uint64, u8p]
        L.nxref_kawpow_search_light.argtypes = [vp, C.c_int, u8p, u8p, C.c_uint64, C.c_uint64,
                                                C.POINTER(C.c_uint64), u8p, u8p]
        L.nxref_ctx_create_full.restype = vp
        L.nxref_ctx_create_full.argtypes = [C.c_int]
        L.nxref_ctx_destroy_full.argtypes = [vp]
        L.nxref_ctx_full_dataset.restype = vp
        L.nxref_ctx_full_dataset.argtypes = [vp]
        L.nxref_kawpow_hash_range_full.argtypes = [vp, C.c_int, u8p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int]
        L.nxref_dataset_item_512.argtypes = [vp, C.c_int64, u8p]
        L.nxref_dataset_item_2048.argtypes = [vp, C.c_uint32, u8p]
        L.nxref_dataset_items_2048.argtypes = [vp, C.c_uint32, C.c_uint32, u8p, C.c_int]
        L.nxref_ethash_hash.argtypes = [vp, u8p, C.c_uint64, u8p, u8p]
        L.nxref_epoch_seed.argtypes = [C.c_int, u8p]
        L.nxref_find_epoch_number.argtypes = [u8p]
        L.nxref_keccak256.argtypes = [u8p, C.c_size_t, u8p]
        L.nxref_keccak512.argtypes = [u8p, C.c_size_t, u8p]
        L.nxref_keccakf800.argtypes = [C.POINTER(C.c_uint32)]
        L.nxref_keccakf1600.argtypes = [C.POINTER(C.c_uint64)]
        self._ctx = {}

    def ctx(self, epoch):
        if epoch not in self._ctx:
            c = self.L.nxref_ctx_create(epoch)
            if not c:
                raise MemoryError("reference context")
            self._ctx[epoch] = c
        return self._ctx[epoch]

    # -- full-DAG context (reference's best-case CPU path) --
    def full_ctx_create(self, epoch):
        c = self.L.nxref_ctx_create_full(epoch)
        if not c:
            raise MemoryError("reference full context")
        return c

    def full_ctx_dataset_ptr(self, c) -> int:
        return self.L.nxref_ctx_full_dataset(c)

    def full_ctx_destroy(self, c):
        self.L.nxref_ctx_destroy_full(c)

    def kawpow_hash_range_full(self, c, height, header, start, count, threads: int = 0, want_output=True):
        out = _buf(64 * count) if want_output else None
        self.L.nxref_kawpow_hash_range_full(c, height, _in(header), start, count, out, threads or os.cpu_count() or 1)
        return bytes(out) if want_output else None

    def light_items(self, epoch):
        return self.L.nxref_light_cache_num_items(epoch)

    def dataset_items(self, epoch):
        return self.L.nxref_full_dataset_num_items(epoch)

    def light_cache(self, epoch) -> bytes:
        c = self.ctx(epoch)
        return C.string_at(self.L.nxref_ctx_light_cache(c), self.L.nxref_ctx_light_items(c) * 64)

    def l1(self, epoch):
        p = self.L.nxref_ctx_l1(self.ctx(epoch))
        return [p[i] for i in range(4096)]

    def epoch_seed(self, epoch) -> bytes:
        o = _buf(32)
        self.L.nxref_epoch_seed(epoch, o)
        return bytes(o)

    def find_epoch_number(self, seed: bytes) -> int:
        return self.L.nxref_find_epoch_number(_in(seed))

    def keccak256(self, d: bytes) -> bytes:
        o = _buf(32)
        self.L.nxref_keccak256(_in(d) if d else _buf(1), len(d), o)
        return bytes(o)

    def keccak512(self, d: bytes) -> bytes:
        o = _buf(64)
        self.L.nxref_keccak512(_in(d) if d else _buf(1), len(d), o)
        return bytes(o)

    def keccakf800(self, words):
        a = (C.c_uint32 * 25)(*words)
        self.L.nxref_keccakf800(a)
        return list(a)

    def keccakf1600(self, words):
        a = (C.c_uint64 * 25)(*words)
        self.L.nxref_keccakf1600(a)
        return list(a)

    def kawpow_hash(self, epoch, height, header, nonce):
        f, m = _buf(32), _buf(32)
        self.L.nxref_kawpow_hash(self.ctx(epoch), height, _in(header), nonce, f, m)
        return bytes(f), bytes(m)

    def kawpow_hash_range(self, epoch, height, header, start, count, threads: int = 0) -> bytes:
        out = _buf(64 * count)
        self.L.nxref_kawpow_hash_range(self.ctx(epoch), height, _in(header), start, count, out,
                                       threads or os.cpu_count() or 1)
        return bytes(out)

    def kawpow_verify(self, epoch, height, header, mix, nonce, boundary) -> bool:
        return bool(self.L.nxref_kawpow_verify(self.ctx(epoch), height, _in(header), _in(mix), nonce, _in(boundary)))

    def kawpow_hash_no_verify(self, height, header, mix, nonce) -> bytes:
        o = _buf(32)
        self.L.nxref_kawpow_hash_no_verify(height, _in(header), _in(mix), nonce, o)
        return bytes(o)

    def kawpow_search(self, epoch, height, header, boundary, start, iterations, threads: int = 0):
        n = C.c_uint64(0)
        f, m = _buf(32), _buf(32)
        found = self.L.nxref_kawpow_search_light(self.ctx(epoch), height, _in(header), _in(boundary), start,
                                                 iterations, C.byref(n), f, m)
        return bool(found), n.value, bytes(f), bytes(m)

    def dataset_item_512(self, epoch, index) -> bytes:
        o = _buf(64)
        self.L.nxref_dataset_item_512(self.ctx(epoch), index, o)
        return bytes(o)

    def dataset_item_2048(self, epoch, index) -> bytes:
        o = _buf(256)
        self.L.nxref_dataset_item_2048(self.ctx(epoch), index, o)
        return bytes(o)

    def dataset_items_2048(self, epoch, first, count, threads: int = 0) -> bytes:
        o = _buf(256 * count)
        self.L.nxref_dataset_items_2048(self.ctx(epoch), first, count, o, threads or os.cpu_count() or 1)
        return bytes(o)

    def ethash_hash(self, epoch, header, nonce):
        f, m = _buf(32), _buf(32)
        self.L.nxref_ethash_hash(self.ctx(epoch), _in(header), nonce, f, m)
        return bytes(f), bytes(m)


# ---------------------------------------------------------------------------------------------------------------------
# Node-side glue, reference-compiled (oracle/ref_node_shim.cpp -> oracle/_ref/libnxref_node.so; `make -C oracle ref_node`).
REF_NODE_SO = os.path.join(HERE, "_ref", "libnxref_node.so")
REF_NODE_DROPIN_SO = os.path.join(HERE, "_ref", "libnxref_node_dropin.so")
REF_NODE_EXPLICIT_SO = os.path.join(HERE, "_ref", "libnxref_node_explicit.so")


class _NodeFields(C.Structure):
    _fields_ = [("version", C.c_int32), ("prev_block", C.c_uint8 * 32), ("merkle_root", C.c_uint8 * 32),
                ("time", C.c_uint32), ("bits", C.c_uint32), ("nonce", C.c_uint32), ("height", C.c_uint32),
                ("nonce64", C.c_uint64), ("mix_hash", C.c_uint8 * 32)]


class _NodeDiskIndex(C.Structure):
    _fields_ = [("height", C.c_int32), ("status", C.c_uint32), ("tx_count", C.c_uint32), ("file", C.c_int32),
                ("data_pos", C.c_uint32), ("undo_pos", C.c_uint32), ("header", _NodeFields)]


class RefNode:
    """The reference's own CBlockHeader / KAWPOWHash / arith_uint256 / CDiskBlockIndex code.  uint256 arguments and
    results are raw storage bytes (what `begin()` points at), i.e. the byte-reverse of GetHex()."""
    kind = "reference"

    def __init__(self, activation_time: int = 1651444217, dropin: bool = False, explicit: bool = False):  # chainparams.cpp:264
        """dropin=True loads libnxref_node_dropin.so instead: the SAME reference block.cpp / hash.cpp, compiled against this
        repository's headers and linked with libnxkawpow.so - the reference's glue running over the product (needs a GPU
        for the full hash, none for the mix-only one)."""
        so, target = (REF_NODE_DROPIN_SO, "ref_node_dropin") if dropin else (REF_NODE_SO, "ref_node")
        if explicit:  # KAWPOWHash / _OnlyMix from examples/node_patch/hash_kawpow_nx.cpp
            so, target = REF_NODE_EXPLICIT_SO, "ref_node_explicit"
        if not os.path.exists(so):
            if not os.path.isdir("/root/reference/src"):
                raise FileNotFoundError(so)
            subprocess.check_call(["make", "-s", "-C", HERE, target])
        L = self.L = C.CDLL(so)
        F, D = C.POINTER(_NodeFields), C.POINTER(_NodeDiskIndex)
        L.nxrefnode_set_activation_time.argtypes = [C.c_uint32]
        L.nxrefnode_header_hash.argtypes = [F, u8p, u8p]
        L.nxrefnode_kawpow_hash.argtypes = [F, u8p, u8p]
        L.nxrefnode_block_hash.argtypes = [F, u8p]
        L.nxrefnode_block_hash_full.argtypes = [F, u8p, u8p]
        L.nxrefnode_serialize_header.argtypes = [F, u8p, C.c_int]
        L.nxrefnode_parse_header.argtypes = [u8p, C.c_int, F]
        L.nxrefnode_serialize_kawpow_input.argtypes = [F, u8p]
        L.nxrefnode_set_compact.argtypes = [C.c_uint32, u8p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.nxrefnode_get_compact.argtypes = [u8p]
        L.nxrefnode_get_compact.restype = C.c_uint32
        L.nxrefnode_check_pow.argtypes = [u8p, C.c_uint32, u8p]
        L.nxrefnode_get_hex.argtypes = [u8p, C.c_char_p]
        L.nxrefnode_set_hex.argtypes = [C.c_char_p, u8p]
        L.nxrefnode_write_varint.argtypes = [C.c_uint64, u8p]
        L.nxrefnode_read_varint.argtypes = [u8p, C.c_int, C.POINTER(C.c_uint64)]
        L.nxrefnode_diskindex_serialize.argtypes = [D, u8p, C.c_int]
        L.nxrefnode_diskindex_parse.argtypes = [u8p, C.c_int, D, C.POINTER(C.c_int32), u8p]
        self.activation_time = activation_time
        L.nxrefnode_set_activation_time(activation_time)

    def set_activation_time(self, t: int):
        self.activation_time = t
        self.L.nxrefnode_set_activation_time(t)

    @staticmethod
    def fields(version, prev, merkle, time, bits, height, nonce64=0, mix=bytes(32), nonce=0) -> _NodeFields:
        return _NodeFields(version, (C.c_uint8 * 32)(*prev), (C.c_uint8 * 32)(*merkle), time, bits, nonce, height, nonce64,
                           (C.c_uint8 * 32)(*mix))

    def header_hash(self, f):
        """(GetKAWPOWHeaderHash() raw bytes, the hash256 hash.cpp:269-270 hands to progpow::hash)"""
        a, b = _buf(32), _buf(32)
        self.L.nxrefnode_header_hash(C.byref(f), a, b)
        return bytes(a), bytes(b)

    def kawpow_hash(self, f):
        """KAWPOWHash (hash.cpp:258-279) -> (final, mix) as uint256 raw bytes"""
        a, b = _buf(32), _buf(32)
        self.L.nxrefnode_kawpow_hash(C.byref(f), a, b)
        return bytes(a), bytes(b)

    def block_hash(self, f) -> bytes:
        a = _buf(32)
        self.L.nxrefnode_block_hash(C.byref(f), a)
        return bytes(a)

    def block_hash_full(self, f):
        a, b = _buf(32), _buf(32)
        self.L.nxrefnode_block_hash_full(C.byref(f), a, b)
        return bytes(a), bytes(b)

    def serialize_header(self, f) -> bytes:
        o = _buf(256)
        n = self.L.nxrefnode_serialize_header(C.byref(f), o, 256)
        return bytes(o[:n])

    def parse_header(self, raw: bytes):
        f = _NodeFields()
        used = self.L.nxrefnode_parse_header(_in(raw), len(raw), C.byref(f))
        return (f, used) if used >= 0 else (None, -1)

    def kawpow_input(self, f) -> bytes:
        o = _buf(80)
        assert self.L.nxrefnode_serialize_kawpow_input(C.byref(f), o) == 80
        return bytes(o)

    def set_compact(self, bits: int):
        """arith_uint256::SetCompact -> (value as 32 big-endian bytes, negative, overflow)"""
        o, neg, ovf = _buf(32), C.c_int(), C.c_int()
        self.L.nxrefnode_set_compact(bits, o, C.byref(neg), C.byref(ovf))
        return bytes(o), bool(neg.value), bool(ovf.value)

    def get_compact(self, be: bytes) -> int:
        return self.L.nxrefnode_get_compact(_in(be))

    def check_pow(self, hash_raw: bytes, bits: int, limit_raw: bytes) -> bool:
        return bool(self.L.nxrefnode_check_pow(_in(hash_raw), bits, _in(limit_raw)))

    def get_hex(self, raw: bytes) -> str:
        o = C.create_string_buffer(65)
        self.L.nxrefnode_get_hex(_in(raw), o)
        return o.value.decode()

    def set_hex(self, text: str) -> bytes:
        o = _buf(32)
        self.L.nxrefnode_set_hex(text.encode(), o)
        return bytes(o)

    def write_varint(self, n: int) -> bytes:
        o = _buf(10)
        return bytes(o[:self.L.nxrefnode_write_varint(n, o)])

    def read_varint(self, data: bytes):
        v = C.c_uint64()
        used = self.L.nxrefnode_read_varint(_in(data) if data else _buf(1), len(data), C.byref(v))
        return (v.value, used) if used >= 0 else (None, -1)

    def diskindex_serialize(self, height, status, tx_count, file, data_pos, undo_pos, f) -> bytes:
        d = _NodeDiskIndex(height, status, tx_count, file, data_pos, undo_pos, f)
        o = _buf(256)
        n = self.L.nxrefnode_diskindex_serialize(C.byref(d), o, 256)
        return bytes(o[:n])

    def diskindex_parse(self, raw: bytes):
        """-> (dict of fields, client_version, GetBlockHash() raw bytes or None, bytes consumed) or None if it throws"""
        d, cv, h = _NodeDiskIndex(), C.c_int32(), _buf(32)
        used = self.L.nxrefnode_diskindex_parse(_in(raw) if raw else _buf(1), len(raw), C.byref(d), C.byref(cv), h)
        if used < 0:
            return None
        f = d.header
        rec = {"height": d.height, "status": d.status, "tx_count": d.tx_count, "file": d.file, "data_pos": d.data_pos,
               "undo_pos": d.undo_pos, "version": f.version, "prev_block": bytes(f.prev_block), "merkle_root": bytes(f.merkle_root),
               "time": f.time, "bits": f.bits, "nonce": f.nonce, "nonce64": f.nonce64, "mix_hash": bytes(f.mix_hash)}
        return rec, cv.value, (bytes(h) if f.time >= self.activation_time else None), used

    def client_version(self) -> int:
        return self.L.nxrefnode_client_version()


# ---------------------------------------------------------------------------------------------------------------------
# Plain-Python restatement of the block-index record format (small inputs only), pinned to tests/golden/node_glue.json.

def varint_write(n: int) -> bytes:
    """WriteVarInt (serialize.h:330-345): MSB base-128, all digits but the last carry bit 7 and have one subtracted."""
    tmp = []
    while True:
        tmp.append((n & 0x7F) | (0x80 if tmp else 0x00))
        if n <= 0x7F:
            break
        n = (n >> 7) - 1
    return bytes(reversed(tmp))


def varint_read(data: bytes, max_value: int = 2**64 - 1):
    """ReadVarInt<I> (serialize.h:347-366) with I's maximum given: (value, bytes used), or None where it throws."""
    n = 0
    for i, c in enumerate(data):
        if n > (max_value >> 7):
            return None
        n = (n << 7) | (c & 0x7F)
        if c & 0x80:
            if n == max_value:
                return None
            n += 1
        else:
            return n, i + 1
    return None  # the stream ran dry: ser_readdata8 throws


def disk_block_index_parse(data: bytes, activation_time: int):
    """CDiskBlockIndex::SerializationOp, read direction (chain.h:398-431): (dict, bytes used) or None where it throws.
    Field types: _nVersion/nHeight/nFile int, nStatus uint32_t, nTx/nDataPos/nUndoPos unsigned int (chain.h:185-209)."""
    import struct
    I32, U32 = 2**31 - 1, 2**32 - 1
    p, rec = 0, {"file": 0, "data_pos": 0, "undo_pos": 0, "nonce": 0, "nonce64": 0, "mix_hash": bytes(32)}

    def take(name, mx):
        nonlocal p
        got = varint_read(data[p:], mx)
        if got is None:
            return False
        rec[name], used = got
        p += used
        return True

    if not (take("client_version", I32) and take("height", I32) and take("status", U32) and take("tx_count", U32)):
        return None
    if rec["status"] & (8 | 16) and not take("file", I32):   # BLOCK_HAVE_DATA | BLOCK_HAVE_UNDO (chain.h:156-157)
        return None
    if rec["status"] & 8 and not take("data_pos", U32):
        return None
    if rec["status"] & 16 and not take("undo_pos", U32):
        return None
    if len(data) - p < 76:
        return None
    rec["version"], rec["prev_block"], rec["merkle_root"], rec["time"], rec["bits"] = struct.unpack_from("<i32s32sII", data, p)
    p += 76
    if rec["time"] < activation_time:
        if len(data) - p < 4:
            return None
        (rec["nonce"],) = struct.unpack_from("<I", data, p)
        p += 4
    else:
        if len(data) - p < 40:
            return None
        rec["nonce64"], rec["mix_hash"] = struct.unpack_from("<Q32s", data, p)
        p += 40
    return rec, p


def disk_block_index_serialize(rec: dict, activation_time: int) -> bytes:
    """CDiskBlockIndex::SerializationOp, write direction (chain.h:398-431)."""
    import struct
    out = varint_write(rec["client_version"]) + varint_write(rec["height"]) + varint_write(rec["status"]) + varint_write(rec["tx_count"])
    if rec["status"] & (8 | 16):
        out += varint_write(rec["file"])
    if rec["status"] & 8:
        out += varint_write(rec["data_pos"])
    if rec["status"] & 16:
        out += varint_write(rec["undo_pos"])
    out += struct.pack("<i32s32sII", rec["version"], rec["prev_block"], rec["merkle_root"], rec["time"], rec["bits"])
    if rec["time"] < activation_time:
        return out + struct.pack("<I", rec["nonce"])
    return out + struct.pack("<Q32s", rec["nonce64"], rec["mix_hash"])
