"""Single-process multi-GPU miner front end (ctypes over the nx_miner_* C ABI).

Replaces the node's nonce loops (reference src/miner.cpp:656-713, src/rpc/mining.cpp:117-173): one call
scans a nonce range over all GPUs and returns the LOWEST qualifying nonce, like progpow::search."""
from __future__ import annotations

import ctypes as C

from ._lib import NxSearchResult, check, inb, lib
from .progpow import SearchResult, _sr


class Miner:
    def __init__(self, devices=None):
        h = C.c_void_p()
        if devices:
            arr = (C.c_int * len(devices))(*devices)
            check(lib.nx_miner_create(arr, len(devices), C.byref(h)))
        else:
            check(lib.nx_miner_create(None, 0, C.byref(h)))
        self._h = h
        self._cancel = C.c_int(0)

    @property
    def device_count(self) -> int:
        return lib.nx_miner_device_count(self._h)

    @property
    def epoch(self) -> int:
        return lib.nx_miner_epoch(self._h)

    @property
    def prepared_epoch(self) -> int:
        return lib.nx_miner_prepared_epoch(self._h)

    def set_batch(self, nonces: int):
        check(lib.nx_miner_set_batch(self._h, nonces))

    def set_epoch(self, epoch: int):
        check(lib.nx_miner_set_epoch(self._h, epoch))

    def prepare_async(self, block_number: int, lookahead_blocks: int = 0):
        check(lib.nx_miner_prepare_async(self._h, block_number, lookahead_blocks))

    def wait_prepared(self):
        check(lib.nx_miner_wait_prepared(self._h))

    def cancel(self):
        self._cancel.value = 1

    def search(self, block_number: int, header_hash: bytes, boundary: bytes, start_nonce: int, iterations: int) -> SearchResult:
        self._cancel.value = 0
        r = NxSearchResult()
        check(lib.nx_miner_search(self._h, block_number, inb(header_hash), inb(boundary), start_nonce, iterations,
                                  C.byref(self._cancel), C.byref(r)))
        return _sr(r)

    def mine_header(self, version: int, prev_block: bytes, merkle_root: bytes, time: int, bits: int, height: int,
                    start_nonce: int = 0, max_tries: int = 1 << 32) -> SearchResult:
        self._cancel.value = 0
        r = NxSearchResult()
        check(lib.nx_miner_mine_header(self._h, version, inb(prev_block), inb(merkle_root), time, bits, height,
                                       start_nonce, max_tries, C.byref(self._cancel), C.byref(r)))
        return _sr(r)

    @property
    def last_failed_devices(self) -> int:
        return lib.nx_miner_last_failed_devices(self._h)

    def last_stats(self):
        n, s = C.c_uint64(), C.c_double()
        lib.nx_miner_last_stats(self._h, C.byref(n), C.byref(s))
        return n.value, s.value

    def close(self):
        if getattr(self, "_h", None):
            lib.nx_miner_destroy(self._h)
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
