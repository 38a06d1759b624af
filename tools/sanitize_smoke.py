#!/usr/bin/env python
"""Small end-to-end pass over every kernel, meant to be run under compute-sanitizer on a GPU box:
  compute-sanitizer --tool memcheck  python tools/sanitize_smoke.py
  compute-sanitizer --tool racecheck python tools/sanitize_smoke.py
Checks results against known answers so that a sanitizer-clean run is also a correct run."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nodexa_chain_core_b200 as nx  # noqa: E402
from nodexa_chain_core_b200 import ethash, progpow  # noqa: E402

KAT_FINAL = "e601a7257a70dc48fccc97a7330d704d776047623b92883d77111fb36870f3d1"
ctx = nx.create_epoch_context_full(0, 0)                                   # K3 (+ host light cache)
assert ctx.l1_cache[0] == 2492749011
got = progpow.hash_batch(ctx, 0, bytes(32), 0, 300)                          # K2, ragged
assert got[:32].hex() == KAT_FINAL
many = progpow.hash_many(ctx, [(0, bytes(32), 0), (1000, bytes(32), 5), (7499, bytes(range(32)), 2**64 - 1)])  # K6
assert many[0].final_hash.hex() == KAT_FINAL
sr = progpow.search(ctx, 0, bytes(32), bytes.fromhex("00" + "ff" * 31), 300, 100)   # K1
assert sr.solution_found and sr.nonce == 395
assert progpow.verify(ctx, 0, bytes(32), sr.mix_hash, 395, sr.final_hash)
f, m = ethash.hash(ctx, bytes(32), 7)                                        # K7
assert ethash.verify(ctx, bytes(32), m, 7, f)
cs = ctx.checksum()                                                          # checksum kernels
out = C.c_uint64()
assert nx.lib.nx_dag_checksum_noalloc(C.c_void_p(ctx.dag_ptr), ctx.dag_bytes, None, C.byref(out)) == 0 and out.value == cs
ctx.close()
print("sanitize_smoke ok")
