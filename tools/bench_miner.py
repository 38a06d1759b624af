#!/usr/bin/env python
"""Hash rate of the single-process multi-GPU miner (nx_miner_*): what a node that links libnxkawpow.so gets.
Epoch 448, boundary 0 (no hit), all visible GPUs.  Prints one JSON object."""
import argparse
import json
import os
import struct
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epoch", type=int, default=448)
    ap.add_argument("--log2-nonces-per-gpu", type=int, default=27)
    a = ap.parse_args()
    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200 import ethash
    height = a.epoch * 7500
    with nx.Miner() as m:
        t0 = time.perf_counter()
        m.set_epoch(a.epoch)
        build_s = time.perf_counter() - t0
        n = m.device_count
        hd = ethash.keccak256(struct.pack("<Q", 1))
        m.set_batch(1 << 24)
        m.search(height, hd, bytes(32), 0, n << 24)  # warm: compiles the period program on every device
        total = n << a.log2_nonces_per_gpu
        r = m.search(height, hd, bytes(32), 1 << 40, total)
        hashes, secs = m.last_stats()
        assert not r.solution_found and hashes == total
        print(json.dumps({"devices": n, "epoch": a.epoch, "epoch_build_s_all_devices": build_s, "nonces": total,
                          "seconds_wall": secs, "MHs": hashes / secs / 1e6, "MHs_per_gpu": hashes / secs / 1e6 / n,
                          "failed_devices": m.last_failed_devices}))


if __name__ == "__main__":
    main()
