#!/usr/bin/env python
"""Mine a synthetic chain across ProgPoW period switches and an epoch boundary on every visible GPU, the way a node
would drive the library block after block (nx_miner_mine_header + nx_miner_prepare_async), then check every block
with the header-verification path (nx_check_block_headers); tests/test_gpu_miner.py runs the same chain against the CPU oracle.
Prints one JSON object: blocks, periods and epochs crossed, wall time per block, the worst stall, verification result."""
import argparse
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def mine_chain(nx, node, first_height, blocks, bits, lookahead, devices=None, batch=1 << 14):
    headers, secs = [], []
    prev = hashlib.sha256(b"nodexa-b200-soak").digest()
    with nx.Miner(devices) as m:
        m.set_batch(batch)
        m.set_epoch(first_height // 7500)
        for i in range(blocks):
            height = first_height + i
            h = node.BlockHeader(0x30000000, prev, hashlib.sha256(b"merkle %d" % height).digest(),
                                 node.MAINNET_KAWPOW_ACTIVATION_TIME + 60 * i, bits, height)
            t0 = time.perf_counter()
            m.prepare_async(height, lookahead)  # next period's program, next epoch's contexts: in the background
            sr = m.mine_header(h.version, h.prev_block, h.merkle_root, h.time, h.bits, h.height, start_nonce=i << 32,
                               max_tries=1 << 24)
            secs.append(time.perf_counter() - t0)
            if not sr.solution_found:
                raise RuntimeError(f"no nonce for height {height}")
            h.nonce64, h.mix_hash = sr.nonce, sr.mix_hash[::-1]
            prev = h.block_hash()
            headers.append((h, sr))
        devices_used = m.device_count
    return headers, secs, devices_used


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--first-height", type=int, default=7500 - 150)
    ap.add_argument("--blocks", type=int, default=300)
    ap.add_argument("--bits", type=lambda s: int(s, 0), default=0x1f00ffff)  # one nonce in 65536
    ap.add_argument("--lookahead", type=int, default=30)
    a = ap.parse_args()
    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200 import node
    t0 = time.perf_counter()
    headers, secs, ndev = mine_chain(nx, node, a.first_height, a.blocks, a.bits, a.lookahead)
    wall = time.perf_counter() - t0
    # verification: one batch per epoch through the interpreted kernel, exactly what headers-first sync would do
    verdicts = []
    by_epoch = {}
    for h, _ in headers:
        by_epoch.setdefault(h.height // 7500, []).append(h)
    t1 = time.perf_counter()
    for e, hs in sorted(by_epoch.items()):
        ctx = nx.EpochContext(e, 0)
        try:
            verdicts += node.check_block_headers(ctx, hs)
        finally:
            ctx.close()
    verify_s = time.perf_counter() - t1
    linked = all(headers[i + 1][0].prev_block == headers[i][0].block_hash() for i in range(len(headers) - 1))
    s = sorted(secs)
    print(json.dumps({
        "devices": ndev, "blocks": len(headers), "first_height": a.first_height, "bits": hex(a.bits),
        "periods_crossed": headers[-1][0].height // 3 - headers[0][0].height // 3,
        "epochs_crossed": headers[-1][0].height // 7500 - headers[0][0].height // 7500,
        "wall_s": wall, "block_ms_median": s[len(s) // 2] * 1e3, "block_ms_p95": s[int(len(s) * 0.95)] * 1e3,
        "block_ms_max": s[-1] * 1e3, "all_headers_ok": all(v == node.HEADER_OK for v in verdicts), "chain_linked": linked,
        "verify_s": verify_s}))


if __name__ == "__main__":
    main()
