#!/usr/bin/env python
"""Equihash solve throughput (Sol/s) of the gated engine on one GPU (its CPU oracle is timed by `bench.py --cpu-baseline-extras`).
NOT a comparison with the reference: the reference has no Equihash (SURVEY.md §0 Fact 1)."""
import argparse
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TAG = b"NodexaPW"


def header(nonce: int) -> bytes:
    base = hashlib.sha512(b"nodexa-equihash-bench").digest() * 3
    return base[:108] + nonce.to_bytes(32, "little")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nonces", type=int, default=64)
    a = ap.parse_args()
    import nodexa_chain_core_b200 as nx
    res = {"note": "self-consistency only: the reference contains no Equihash; (n,k) and tag are inputs", "runs": []}
    for n, k in ((200, 9), (144, 5)):
        with nx.equihash.Solver(n, k, TAG) as s:
            for w in range(3):
                s.solve(header(10**6 + w))
            sols, kernel_ms, dropped, cands = 0, 0.0, 0, 0
            t0 = time.perf_counter()
            for i in range(a.nonces):
                got = s.solve(header(i))
                sols += len(got)
                kernel_ms += s.last_timing()[0]
                dropped += s.last_dropped
                cands += s.last_candidates
            wall = time.perf_counter() - t0
            run = {"n": n, "k": k, "nonces": a.nonces, "solutions": sols, "sol_per_s_device": sols / (kernel_ms * 1e-3),
                   "sol_per_s_e2e": sols / wall, "ms_per_nonce_device": kernel_ms / a.nonces, "ms_per_nonce_e2e": wall / a.nonces * 1e3,
                   "launches_per_nonce": s.last_timing()[1], "rows_dropped": dropped, "candidates": cands}
            # the same headers as one batch: two solves in flight, host decoding overlapped (wall clock of the whole call)
            s.solve_batch([header(10**6 + w) for w in range(4)])
            t0 = time.perf_counter()
            got = s.solve_batch([header(i) for i in range(a.nonces)])
            wall_b = time.perf_counter() - t0
            run.update(batch_solutions=sum(len(x) for x in got), sol_per_s_batch_e2e=sum(len(x) for x in got) / wall_b,
                       ms_per_nonce_batch_e2e=wall_b / a.nonces * 1e3, batch_rows_dropped=s.last_dropped)
            res["runs"].append(run)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
