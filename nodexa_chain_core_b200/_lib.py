"""ctypes binding of libnxkawpow.so (the C ABI declared in include/nx_kawpow.h).

The library is built in-tree by `make lib` / `__graft_entry__.build()`.  There is no Python or CPU
fallback: if the shared object is missing, importing this module raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnxkawpow.so")

NX_OK = 0
u8p = C.POINTER(C.c_uint8)


class NxError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"nx error {code}: {msg}")
        self.code = code


class NxResult(C.Structure):
    _fields_ = [("final_hash", C.c_uint8 * 32), ("mix_hash", C.c_uint8 * 32)]


class NxSearchResult(C.Structure):
    _fields_ = [("found", C.c_int32), ("reserved", C.c_uint32), ("nonce", C.c_uint64),
                ("final_hash", C.c_uint8 * 32), ("mix_hash", C.c_uint8 * 32)]


class NxHashRecord(C.Structure):
    _fields_ = [("block_number", C.c_int32), ("pad", C.c_uint32), ("nonce", C.c_uint64), ("header_hash", C.c_uint8 * 32)]


class NxBlockHeader(C.Structure):
    _fields_ = [("version", C.c_int32), ("prev_block", C.c_uint8 * 32), ("merkle_root", C.c_uint8 * 32),
                ("time", C.c_uint32), ("bits", C.c_uint32), ("height", C.c_uint32), ("nonce64", C.c_uint64),
                ("mix_hash", C.c_uint8 * 32)]


class NxDiskBlockIndex(C.Structure):
    _fields_ = [("client_version", C.c_int32), ("height", C.c_int32), ("status", C.c_uint32), ("tx_count", C.c_uint32),
                ("file", C.c_int32), ("data_pos", C.c_uint32), ("undo_pos", C.c_uint32), ("version", C.c_int32),
                ("prev_block", C.c_uint8 * 32), ("merkle_root", C.c_uint8 * 32), ("time", C.c_uint32), ("bits", C.c_uint32),
                ("nonce", C.c_uint32), ("nonce64", C.c_uint64), ("mix_hash", C.c_uint8 * 32)]


class NxVerifyRecord(C.Structure):
    _fields_ = [("block_number", C.c_int32), ("pad", C.c_uint32), ("nonce", C.c_uint64),
                ("header_hash", C.c_uint8 * 32), ("mix_hash", C.c_uint8 * 32), ("boundary", C.c_uint8 * 32)]


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make lib` (or __graft_entry__.build()). "
            "nodexa_chain_core_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i, u32, u64, sz = C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_size_t
    sig = {
        "nx_last_error": (C.c_char_p, []),
        "nx_version": (C.c_char_p, []),
        "nx_device_count": (i, []),
        "nx_find_largest_prime": (i, [i]),
        "nx_light_cache_num_items": (i, [i]),
        "nx_full_dataset_num_items": (i, [i]),
        "nx_epoch_seed": (None, [i, u8p]),
        "nx_find_epoch_number": (i, [u8p]),
        "nx_keccakf1600": (None, [C.POINTER(u64)]),
        "nx_keccakf800": (None, [C.POINTER(u32)]),
        "nx_keccak256": (None, [u8p, sz, u8p]),
        "nx_keccak512": (None, [u8p, sz, u8p]),
        "nx_epoch_create": (i, [i, i, C.POINTER(vp)]),
        "nx_epoch_create_from_light": (i, [i, i, vp, C.POINTER(vp)]),
        "nx_epoch_create_sharded": (i, [i, i, i, i, C.POINTER(vp)]),
        "nx_epoch_finalize": (i, [vp]),
        "nx_epoch_create_deferred": (i, [i, i, i, i, vp, C.POINTER(vp)]),
        "nx_epoch_ipc_handle": (i, [vp, u8p]),
        "nx_epoch_generate_fused_ipc": (i, [vp, u8p]),
        "nx_epoch_generate_fused_peers": (i, [vp, C.POINTER(vp)]),
        "nx_epoch_create_shell": (i, [i, i, i, i, C.POINTER(vp)]),
        "nx_epoch_build_light_cache": (i, [vp]),
        "nx_epoch_light_ptr": (vp, [vp]),
        "nx_epoch_adopt_light_cache": (i, [vp]),
        "nx_epoch_open_ipc_peers": (i, [vp, u8p]),
        "nx_epoch_generate_slice": (i, [vp]),
        "nx_epoch_destroy": (None, [vp]),
        "nx_epoch_number": (i, [vp]),
        "nx_epoch_light_items": (i, [vp]),
        "nx_epoch_dataset_items": (i, [vp]),
        "nx_epoch_light_cache": (vp, [vp]),
        "nx_epoch_l1_cache": (C.POINTER(u32), [vp]),
        "nx_epoch_device": (i, [vp]),
        "nx_epoch_dag_ptr": (vp, [vp]),
        "nx_epoch_dag_bytes": (sz, [vp]),
        "nx_epoch_dag_padded_bytes": (sz, [vp]),
        "nx_epoch_dag_share_bytes": (sz, [vp]),
        "nx_epoch_build_times": (None, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "nx_dag_generate_range": (i, [vp, u32, u32, vp]),
        "nx_dag_read_items": (i, [vp, u32, u32, u8p]),
        "nx_dag_checksum": (i, [vp, C.POINTER(u64)]),
        "nx_dag_checksum_noalloc": (i, [vp, sz, vp, C.POINTER(u64)]),
        "nx_kawpow_prepare": (i, [vp, i]),
        "nx_kawpow_hash": (i, [vp, i, u8p, u64, C.POINTER(NxResult)]),
        "nx_kawpow_hash_batch": (i, [vp, i, u8p, u64, u64, vp]),
        "nx_kawpow_hash_batch_device": (i, [vp, i, u8p, u64, u64, vp]),
        "nx_kawpow_search": (i, [vp, i, u8p, u8p, u64, u64, C.POINTER(NxSearchResult)]),
        "nx_kawpow_search_enqueue": (i, [vp, i, u8p, u8p, u64, u64]),
        "nx_kawpow_search_sync": (i, [vp, C.POINTER(NxSearchResult)]),
        "nx_kawpow_last_timing": (i, [vp, C.POINTER(C.c_float), C.POINTER(i)]),
        "nx_kawpow_verify": (i, [vp, i, u8p, u8p, u64, u8p]),
        "nx_kawpow_hash_no_verify": (None, [i, u8p, u8p, u64, u8p]),
        "nx_kawpow_hash_many": (i, [vp, C.POINTER(NxHashRecord), sz, vp]),
        "nx_test_hash_many_guarded": (i, [vp, C.POINTER(NxHashRecord), sz, vp, C.POINTER(u64)]),
        "nx_kawpow_verify_batch": (i, [vp, C.POINTER(NxVerifyRecord), sz, u8p]),
        "nx_ethash_hash": (i, [vp, u8p, u64, C.POINTER(NxResult)]),
        "nx_ethash_hash_batch": (i, [vp, u8p, u64, u64, vp]),
        "nx_ethash_verify_final_hash": (i, [u8p, u8p, u64, u8p]),
        "nx_kawpow_header_hash": (None, [C.c_int32, u8p, u8p, u32, u32, u32, u8p]),
        "nx_sha256d": (None, [u8p, sz, u8p]),
        "nx_compact_to_boundary": (i, [u32, u8p]),
        "nx_check_proof_of_work": (i, [u8p, u32, u8p]),
        "nx_block_header_parse": (None, [u8p, C.POINTER(NxBlockHeader)]),
        "nx_block_header_serialize": (None, [C.POINTER(NxBlockHeader), u8p]),
        "nx_block_hash": (None, [u8p, u8p]),
        "nx_check_block_headers": (i, [vp, u8p, sz, u8p, i, u8p]),
        "nx_varint_write": (sz, [u64, u8p]),
        "nx_varint_read": (sz, [u8p, sz, C.POINTER(u64)]),
        "nx_disk_block_index_serialize": (sz, [C.POINTER(NxDiskBlockIndex), u32, u8p, sz]),
        "nx_disk_block_index_parse": (sz, [u8p, sz, u32, C.POINTER(NxDiskBlockIndex)]),
        "nx_disk_block_index_check": (i, [C.POINTER(NxDiskBlockIndex), u32, u8p, u8p]),
        "nx_miner_create": (i, [C.POINTER(i), i, C.POINTER(vp)]),
        "nx_miner_destroy": (None, [vp]),
        "nx_miner_device_count": (i, [vp]),
        "nx_miner_set_batch": (i, [vp, u64]),
        "nx_miner_set_epoch": (i, [vp, i]),
        "nx_miner_prepare_async": (i, [vp, i, i]),
        "nx_miner_wait_prepared": (i, [vp]),
        "nx_miner_epoch": (i, [vp]),
        "nx_miner_prepared_epoch": (i, [vp]),
        "nx_miner_context": (vp, [vp, i]),
        "nx_miner_search": (i, [vp, i, u8p, u8p, u64, u64, C.POINTER(i), C.POINTER(NxSearchResult)]),
        "nx_miner_last_failed_devices": (i, [vp]),
        "nx_miner_last_stats": (i, [vp, C.POINTER(u64), C.POINTER(C.c_double)]),
        "nx_miner_mine_header": (i, [vp, C.c_int32, u8p, u8p, u32, u32, u32, u64, u64, C.POINTER(i),
                                     C.POINTER(NxSearchResult)]),
        # gated Equihash engine (include/nx_equihash.h; no reference counterpart)
        "nx_equihash_solution_bytes": (sz, [i, i]),
        "nx_equihash_create": (i, [i, i, i, u8p, C.POINTER(vp)]),
        "nx_equihash_destroy": (None, [vp]),
        "nx_equihash_solve": (i, [vp, u8p, sz, u8p, i, C.POINTER(i)]),
        "nx_equihash_solve_batch": (i, [vp, u8p, sz, i, u8p, i, C.POINTER(i), C.POINTER(C.c_float)]),
        "nx_equihash_validate": (i, [vp, u8p, sz, u8p, i, u8p]),
        "nx_equihash_last_dropped": (i, [vp]),
        "nx_equihash_last_candidates": (i, [vp]),
        "nx_equihash_last_timing": (i, [vp, C.POINTER(C.c_float), C.POINTER(i)]),
        # introspection (nx_tools.cpp; not part of the drop-in surface)
        "nx_kawpow_program_source": (sz, [u64, u32, i, C.c_char_p, sz]),
        "nx_kawpow_program_ops": (i, [u64, C.POINTER(u32 * 6), i]),
        "nx_kawpow_compile_cubin": (C.c_long, [u64, u32, C.c_char_p, sz]),
        "nx_host_light_cache": (None, [i, vp]),
        "nx_host_light_cache_variant": (i, [i, i, vp]),
        "nx_host_light_cache_cached": (i, [i, vp]),
        "nx_keccakf1600_avx512": (i, [C.POINTER(u64)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    return L


lib = _load()


def check(rc: int) -> int:
    if rc < 0:
        raise NxError(rc, lib.nx_last_error().decode(errors="replace"))
    return rc


def inb(b: bytes):
    return (C.c_uint8 * len(b)).from_buffer_copy(b)


def outb(n: int):
    return (C.c_uint8 * n)()
