#!/usr/bin/env python
"""Solo-mine against a node over its JSON-RPC with every visible GPU (the external-miner route):
  clore_blockchaind -miningaddress=<address> -rpcuser=u -rpcpassword=p ...
  python tools/nxminer.py --rpc-url http://127.0.0.1:8766/ --rpc-user u --rpc-password p
Polls getblocktemplate (pprpcheader / pprpcepoch / target / height), searches on the GPUs, submits with pprpcsb, and
prints one JSON line per submitted block and a hash-rate line every --report seconds."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rpc-url", default="http://127.0.0.1:8766/")
    ap.add_argument("--rpc-user", default=os.environ.get("NX_RPC_USER", ""))
    ap.add_argument("--rpc-password", default=os.environ.get("NX_RPC_PASSWORD", ""))
    ap.add_argument("--devices", default="", help="comma-separated GPU indices (default: all)")
    ap.add_argument("--window", type=int, default=1 << 28, help="nonces per search call")
    ap.add_argument("--start-nonce", type=lambda s: int(s, 0), default=0)
    ap.add_argument("--max-blocks", type=int, default=None)
    ap.add_argument("--max-seconds", type=float, default=None)
    ap.add_argument("--report", type=float, default=10.0)
    a = ap.parse_args()
    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200.rpc_miner import RpcMiner, http_rpc
    devices = [int(x) for x in a.devices.split(",") if x] or None
    m = RpcMiner(http_rpc(a.rpc_url, a.rpc_user, a.rpc_password), miner=nx.Miner(devices), window=a.window, start_nonce=a.start_nonce)
    t0 = last = time.perf_counter()
    seen = 0
    try:
        windows = 0
        while (a.max_blocks is None or len(m.submitted) < a.max_blocks) and (a.max_seconds is None or time.perf_counter() - t0 < a.max_seconds):
            if m.work is None or windows % 4 == 0:
                if m.poll():
                    print(json.dumps({"event": "work", "height": m.work.height, "epoch": m.work.epoch, "header": m.work.header_hash.hex()}), flush=True)
            m.step()
            windows += 1
            for s in m.submitted[seen:]:
                print(json.dumps({"event": "submitted", **s}), flush=True)
            seen = len(m.submitted)
            if time.perf_counter() - last >= a.report:
                last = time.perf_counter()
                print(json.dumps({"event": "rate", "MHs": m.hashrate / 1e6, "devices": m.miner.device_count, "blocks": seen}), flush=True)
    finally:
        m.close()


if __name__ == "__main__":
    main()
