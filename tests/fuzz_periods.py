#!/usr/bin/env python
"""One-off differential fuzz (not collected by pytest; run by hand on a GPU box): N random ProgPoW periods of one epoch,
each compiled by NVRTC and run for a few nonces on the compiled kernel (K2) and the interpreted kernel (K6), every result
compared with the CPU oracle.  Hunts for code-generator corner cases that the fixed test heights do not reach.
  python tests/fuzz_periods.py --periods 400 --seed 1 > profiles/r1_fuzz_periods.json"""
import argparse
import json
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--periods", type=int, default=400)
    ap.add_argument("--epoch", type=int, default=0)
    ap.add_argument("--nonces", type=int, default=8)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200 import progpow
    from oracle.oracle import Oracle
    orc = Oracle()
    rnd = random.Random(a.seed)
    ctx = nx.EpochContext(a.epoch, 0)
    heights = sorted(rnd.sample(range(a.epoch * 7500, a.epoch * 7500 + 7500), a.periods))
    heights = sorted({h // 3: h for h in heights}.values())  # one height per period
    t0 = time.perf_counter()
    checked, mismatches = 0, []
    for h in heights:
        hd = rnd.randbytes(32)
        start = rnd.getrandbits(64) - a.nonces
        want = orc.kawpow_hash_range(a.epoch, h, hd, start, a.nonces)
        got = progpow.hash_batch(ctx, h, hd, start, a.nonces)
        interp = progpow.hash_many(ctx, [(h, hd, start + i) for i in range(a.nonces)])
        interp = b"".join(r.final_hash + r.mix_hash for r in interp)
        checked += a.nonces
        if got != want or interp != want:
            mismatches.append({"height": h, "header": hd.hex(), "start": start, "compiled_ok": got == want, "interpreted_ok": interp == want})
    ctx.close()
    print(json.dumps({"epoch": a.epoch, "periods": len(heights), "nonces_per_period": a.nonces, "hashes_checked_per_kernel": checked,
                      "seed": a.seed, "mismatches": mismatches, "seconds": time.perf_counter() - t0}))
    sys.exit(1 if mismatches else 0)


if __name__ == "__main__":
    main()
