#!/usr/bin/env python
"""Secondary paths of SURVEY.md §8 measured on one GPU (bench.py covers the headline search):
K3 DAG generation, K2 batch hash (results to host), K6 interpreted hash_many / verify_batch, K7 classic Ethash.
Prints one JSON object; CUDA-event times come from nx_kawpow_last_timing / nx_epoch_build_times."""
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
    ap.add_argument("--epoch", type=int, default=448)
    ap.add_argument("--records", type=int, default=1 << 16)
    a = ap.parse_args()
    import torch

    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200 import ethash, progpow
    res = {"epoch": a.epoch}
    t0 = time.perf_counter()
    ctx = nx.EpochContext(a.epoch, 0)
    res["epoch_create_s"] = time.perf_counter() - t0
    light_ms, dag_ms = ctx.build_times_ms()
    res["light_cache_ms_host"] = light_ms
    # K3 alone, warm, three times
    buf = torch.empty(ctx.dag_bytes, dtype=torch.uint8, device="cuda:0")
    times = []
    for _ in range(3):
        ctx.generate_range(0, ctx.dag_num_items, buf.data_ptr())
        times.append(ctx.last_timing()[0])
    ms = min(times)
    gather_bytes = ctx.dag_bytes // 64 * 512 * 64
    res["K3_dag_generate"] = {"ms": ms, "first_build_ms": dag_ms, "dag_bytes": ctx.dag_bytes, "write_GBps": ctx.dag_bytes / ms / 1e6,
                              "l2_gather_bytes": gather_bytes, "l2_gather_TBps": gather_bytes / ms / 1e9}
    del buf
    height = a.epoch * 7500
    hd = ethash.keccak256(b"paths")
    # K2: batch hash with results copied to the host
    n = 1 << 22
    progpow.prepare(ctx, height)
    progpow.hash_batch(ctx, height, hd, 0, 1 << 16)
    t0 = time.perf_counter()
    progpow.hash_batch(ctx, height, hd, 1 << 40, n)
    wall = time.perf_counter() - t0
    kms, _ = ctx.last_timing()
    res["K2_hash_batch"] = {"nonces": n, "kernel_MHs": n / kms / 1e3, "e2e_MHs_with_256MiB_D2H_and_python_copy": n / wall / 1e6}
    # K6: independent records, every one in a different period
    rnd = random.Random(1)
    recs = [(height + rnd.randrange(7500), rnd.randbytes(32), rnd.getrandbits(64)) for _ in range(a.records)]
    progpow.hash_many(ctx, recs[:256])
    t0 = time.perf_counter()
    out = progpow.hash_many(ctx, recs)
    wall = time.perf_counter() - t0
    kms, _ = ctx.last_timing()
    res["K6_hash_many"] = {"records": a.records, "distinct_periods": len({r[0] // 3 for r in recs}), "kernel_ms": kms,
                           "kernel_MHs": a.records / kms / 1e3, "e2e_MHs_incl_python_marshalling": a.records / wall / 1e6,
                           "dag_GBps": a.records * 16384 / kms / 1e6}
    vrecs = [(r[0], r[1], o.mix_hash, r[2], b"\xff" * 32) for r, o in zip(recs[:8192], out[:8192])]
    t0 = time.perf_counter()
    ok = progpow.verify_batch(ctx, vrecs)
    res["K6_verify_batch"] = {"records": len(vrecs), "all_valid": all(ok), "e2e_kHs": len(vrecs) / (time.perf_counter() - t0) / 1e3}
    # single-header latency: what the node's validation thread sees per block (progpow::hash at its own height)
    singles = [(height + rnd.randrange(7500), rnd.randbytes(32), rnd.getrandbits(64)) for _ in range(300)]
    progpow.hash(ctx, *singles[0])
    t0 = time.perf_counter()
    for r in singles:
        progpow.hash(ctx, *r)
    res["single_hash_latency_us"] = {"mean": (time.perf_counter() - t0) / len(singles) * 1e6, "calls": len(singles),
                                     "path": "nx_kawpow_hash -> interpreted kernel K6 (uncompiled period), via ctypes"}
    # K7 classic Ethash
    n7 = 1 << 20
    ethash.hash_batch(ctx, hd, 0, 1 << 12)
    ethash.hash_batch(ctx, hd, 1 << 32, n7)
    kms, _ = ctx.last_timing()
    res["K7_ethash_hash"] = {"nonces": n7, "kernel_MHs": n7 / kms / 1e3, "dag_GBps": n7 * 8192 / kms / 1e6}
    # the CPU side of the epoch build (reference code on the host cores) is timed by `bench.py --cpu-baseline-extras`
    print(json.dumps(res))
    ctx.close()


if __name__ == "__main__":
    main()
