#!/usr/bin/env python
"""Throughput of ONE progpow::search call (C ABI, host arguments in, result out) as a function of its `iterations`
argument - what a node gets when it swaps its inner nonce loop (reference src/rpc/mining.cpp:141-149 uses
nInnerLoopCount = 65536, :119) for a single call.  Prints one JSON object."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epoch", type=int, default=448)
    a = ap.parse_args()
    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200 import ethash, progpow
    ctx = nx.EpochContext(a.epoch, 0)
    height = a.epoch * 7500
    hd = ethash.keccak256(b"sizes")
    progpow.prepare(ctx, height)
    progpow.search(ctx, height, hd, bytes(32), 0, 1 << 22)  # warm-up
    rows = []
    for lg in (10, 12, 14, 16, 17, 18, 20, 22, 24, 26):
        n = 1 << lg
        reps = max(3, min(200, (1 << 25) // n))
        best = 1e9
        t_all0 = time.perf_counter()
        for r in range(reps):
            t0 = time.perf_counter()
            sr = progpow.search(ctx, height, hd, bytes(32), (r + 1) << 40, n)
            best = min(best, time.perf_counter() - t0)
            assert not sr.solution_found
        mean = (time.perf_counter() - t_all0) / reps
        rows.append({"iterations": n, "best_ms": best * 1e3, "mean_ms": mean * 1e3, "MHs_mean": n / mean / 1e6})
    print(json.dumps({"epoch": a.epoch, "call": "nx_kawpow_search via ctypes", "rows": rows}))


if __name__ == "__main__":
    main()
