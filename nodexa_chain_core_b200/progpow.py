"""Mirror of the reference's `progpow::` namespace
(src/crypto/ethash/include/ethash/progpow.hpp:13-49, lib/ethash/progpow.cpp:298-579)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

from ._lib import NxHashRecord, NxResult, NxSearchResult, NxVerifyRecord, check, inb, lib, outb
from .ethash import EpochContext

revision = "0.9.4"
period_length = 3
num_regs = 32
num_lanes = 16
num_cache_accesses = 11
num_math_operations = 18
l1_cache_size = 16 * 1024
l1_cache_num_items = l1_cache_size // 4


@dataclass
class Result:
    """ethash_result (ethash.h:48-52)"""
    final_hash: bytes
    mix_hash: bytes


@dataclass
class SearchResult:
    """ethash::search_result (ethash.hpp:49-61); all-zero when nothing was found."""
    solution_found: bool = False
    nonce: int = 0
    final_hash: bytes = bytes(32)
    mix_hash: bytes = bytes(32)


def hash(context: EpochContext, block_number: int, header_hash: bytes, nonce: int) -> Result:  # noqa: A001
    """progpow::hash (progpow.cpp:298-355, :357-429)"""
    r = NxResult()
    check(lib.nx_kawpow_hash(context._h, block_number, inb(header_hash), nonce, C.byref(r)))
    return Result(bytes(r.final_hash), bytes(r.mix_hash))


def hash_batch(context: EpochContext, block_number: int, header_hash: bytes, start_nonce: int, count: int) -> bytes:
    """count x {final_hash[32], mix_hash[32]} for consecutive nonces (kernel K2); no reference
    counterpart beyond calling progpow::hash in a loop."""
    out = outb(64 * count)
    check(lib.nx_kawpow_hash_batch(context._h, block_number, inb(header_hash), start_nonce, count, out))
    return bytes(out)


def hash_batch_device(context: EpochContext, block_number: int, header_hash: bytes, start_nonce: int, count: int,
                      out_device_ptr: int) -> None:
    check(lib.nx_kawpow_hash_batch_device(context._h, block_number, inb(header_hash), start_nonce, count,
                                          C.c_void_p(out_device_ptr)))


def verify(context: EpochContext, block_number: int, header_hash: bytes, mix_hash: bytes, nonce: int,
           boundary: bytes) -> bool:
    """progpow::verify (progpow.cpp:431-495)"""
    return bool(check(lib.nx_kawpow_verify(context._h, block_number, inb(header_hash), inb(mix_hash), nonce,
                                           inb(boundary))))


def hash_no_verify(block_number: int, header_hash: bytes, mix_hash: bytes, nonce: int) -> bytes:
    """progpow::hash_no_verify (progpow.cpp:498-550)"""
    o = outb(32)
    lib.nx_kawpow_hash_no_verify(block_number, inb(header_hash), inb(mix_hash), nonce, o)
    return bytes(o)


def _sr(r: NxSearchResult) -> SearchResult:
    return SearchResult(bool(r.found), r.nonce, bytes(r.final_hash), bytes(r.mix_hash))


def search(context: EpochContext, block_number: int, header_hash: bytes, boundary: bytes, start_nonce: int,
           iterations: int) -> SearchResult:
    """progpow::search (progpow.cpp:567-579): lowest qualifying nonce in [start, start+iterations)."""
    r = NxSearchResult()
    check(lib.nx_kawpow_search(context._h, block_number, inb(header_hash), inb(boundary), start_nonce, iterations,
                               C.byref(r)))
    return _sr(r)


search_light = search  # progpow.cpp:553-565: same contract, only the context type differs in the reference


def search_enqueue(context: EpochContext, block_number: int, header_hash: bytes, boundary: bytes, start_nonce: int,
                   count: int) -> None:
    check(lib.nx_kawpow_search_enqueue(context._h, block_number, inb(header_hash), inb(boundary), start_nonce, count))


def search_sync(context: EpochContext) -> SearchResult:
    r = NxSearchResult()
    check(lib.nx_kawpow_search_sync(context._h, C.byref(r)))
    return _sr(r)


def prepare(context: EpochContext, block_number: int) -> None:
    check(lib.nx_kawpow_prepare(context._h, block_number))


def hash_many(context: EpochContext, records) -> list:
    """records = iterable of (block_number, header_hash, nonce), any mix of periods inside the context's epoch.
    One launch of the interpreted kernel (K6); returns [Result, ...]."""
    records = list(records)
    arr = (NxHashRecord * max(1, len(records)))()
    for i, (bn, hd, nonce) in enumerate(records):
        arr[i].block_number = bn
        arr[i].nonce = nonce
        C.memmove(arr[i].header_hash, hd, 32)
    out = outb(64 * len(records))
    check(lib.nx_kawpow_hash_many(context._h, arr, len(records), out))
    b = bytes(out)
    return [Result(b[64 * i:64 * i + 32], b[64 * i + 32:64 * i + 64]) for i in range(len(records))]


def verify_batch(context: EpochContext, records) -> list:
    """records = iterable of (block_number, header_hash, mix_hash, nonce, boundary); returns [bool, ...]
    with progpow::verify's semantics for each (header sync: validation.cpp:11652-11662)."""
    records = list(records)
    arr = (NxVerifyRecord * max(1, len(records)))()
    for i, (bn, hd, mix, nonce, boundary) in enumerate(records):
        arr[i].block_number = bn
        arr[i].nonce = nonce
        C.memmove(arr[i].header_hash, hd, 32)
        C.memmove(arr[i].mix_hash, mix, 32)
        C.memmove(arr[i].boundary, boundary, 32)
    ok = outb(max(1, len(records)))
    check(lib.nx_kawpow_verify_batch(context._h, arr, len(records), ok))
    return [bool(ok[i]) for i in range(len(records))]
