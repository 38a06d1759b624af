#!/usr/bin/env python
"""Run bench.py once per kernel shape (NX_KAWPOW_TUNE) and print MH/s for each — on a GPU box."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TUNES = sys.argv[1:] or [
    "", "rotimad=1",
    "copies=8,threads=1024,minblocks=1", "copies=8,threads=1024,minblocks=1,rotimad=1",
    "copies=4,threads=512,minblocks=2", "copies=4,threads=512,minblocks=2,rotimad=1",
    "copies=2,threads=256,minblocks=4", "copies=2,threads=256,minblocks=4,rotimad=1",
    "copies=1,threads=512,minblocks=2", "copies=1,threads=128,minblocks=8",
    "copies=4,threads=256,minblocks=2,rotimad=1",
]
for t in TUNES:
    env = dict(os.environ, NX_KAWPOW_TUNE=t)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", os.environ.get("SWEEP_STEPS", "10"), "--warmup", "3",
                          "--no-cpu-baseline"], env=env, capture_output=True, text=True, cwd=ROOT)
    try:
        j = json.loads(out.stdout.strip().splitlines()[-1])
        print(f"{t or 'default':55s} {j['value']:8.2f} MH/s  frac {j['roofline']['frac']:.3f}  e2e {j['e2e']['value']:.1f}  "
              f"sm {j['clocks']['sm_mhz']} MHz", flush=True)
    except Exception:
        print(f"{t or 'default':55s} FAILED: {out.stderr[-300:]}", flush=True)
