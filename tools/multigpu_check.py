#!/usr/bin/env python
"""Run under torchrun (one rank per GPU): sharded DAG build + in-place NCCL all-gather, then a sharded
search; every rank checks the result against the single-GPU answer and the reference-made digests.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
         tools/multigpu_check.py [--epoch 0]
"""
import argparse
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epoch", type=int, default=0)
    ap.add_argument("--fused", action="store_true", help="DAG exchange inside the generating kernel (peer stores via CUDA IPC) instead of NCCL")
    ap.add_argument("--shell", action="store_true", help="one light-cache chain for the job: rank 0 builds it, NCCL broadcasts it "
                    "(nx_epoch_create_shell / build_light_cache / adopt_light_cache), then --fused or the NCCL all-gather")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist

    import nodexa_chain_core_b200 as nx
    from nodexa_chain_core_b200 import progpow, sharding

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    dev = f"cuda:{local}"

    if a.shell:
        ctx = nx.EpochContext(a.epoch, local, rank=rank, world=world, shell=True)
        handles = None
        if a.fused:
            handles = [None] * world
            dist.all_gather_object(handles, ctx.ipc_handle())
            ctx.open_ipc_peers(handles)
        if rank == 0:
            ctx.build_light_cache()
        dist.broadcast(ctx.light_tensor(), src=0)
        torch.cuda.synchronize()
        if rank != 0:
            ctx.adopt_light_cache()
        gold_e = json.load(open(os.path.join(ROOT, "tests", "golden", "epochs.json")))
        assert hashlib.sha256(ctx.light_cache).hexdigest() == gold_e[f"light_cache_sha256_epoch{a.epoch}"], rank
        if not a.fused:
            ctx.generate_slice()
    else:
        ctx = nx.EpochContext(a.epoch, local, rank=rank, world=world, deferred=a.fused)
    share = ctx.dag_share_bytes
    first, last = sharding.dag_slice(ctx.dag_num_items, rank, world)
    assert share == sharding.dag_share_items(ctx.dag_num_items, world) * 256
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if a.fused:
        if a.shell:
            handles = None  # peers already mapped
        else:
            handles = [None] * world
            dist.all_gather_object(handles, ctx.ipc_handle())
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        ctx.generate_fused_ipc(handles)   # one kernel: generate my slice, store it into every rank's DAG
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()                    # every rank's kernel has finished writing into my buffer
    else:
        dag = ctx.dag_tensor()
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        dist.all_gather_into_tensor(dag, dag[rank * share:(rank + 1) * share])
        e1.record()
        torch.cuda.synchronize()
    ctx.finalize()
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "dag_digest.json")))[str(a.epoch)]
    cs = ctx.checksum()
    assert cs == gold["weighted_sum"], (rank, cs, gold["weighted_sum"])
    items = json.load(open(os.path.join(ROOT, "tests", "golden", "dag_items.json")))[str(a.epoch)]
    for i, sha in items["shard_edges_sha256"].items():
        assert hashlib.sha256(ctx.read_items(int(i), 1)).hexdigest() == sha, (rank, i)

    # sharded search == single-GPU search (each rank holds the full DAG, so it can compute both)
    height = a.epoch * 7500
    hd = hashlib.sha256(b"multigpu").digest()
    boundary = ((1 << (256 - 16)) - 1).to_bytes(32, "big")
    n = 1 << 20
    single = progpow.search(ctx, height, hd, boundary, 1000, n)
    for batch in (1 << 14, 1 << 18):
        sh = sharding.search_sharded(ctx, height, hd, boundary, 1000, n, rank, world, batch=batch, dist=dist, device=dev)
        assert sh == single, (rank, batch, sh, single)
    miss = sharding.search_sharded(ctx, height, hd, bytes(32), 1000, 1 << 16, rank, world, batch=4096, dist=dist, device=dev)
    assert not miss.solution_found and miss.nonce == 0
    if a.epoch == 0:  # the reference's search KAT, split over the ranks
        kat = sharding.search_sharded(ctx, 0, bytes(32), bytes.fromhex("00" + "ff" * 31), 300, 100, rank, world, batch=16,
                                      dist=dist, device=dev)
        assert kat.solution_found and kat.nonce == 395
    light_ms, dag_ms = ctx.build_times_ms()
    t = torch.tensor([dag_ms, e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"multigpu_check": "ok", "exchange": "fused peer stores (IPC)" if a.fused else "nccl all-gather",
                          "world": world, "epoch": a.epoch, "dag_kernel_ms_max": t[0].item(),
                          "allgather_ms_max": t[1].item(), "allgather_GBps_per_gpu_recv": share * (world - 1) / t[1].item() / 1e6,
                          "found_nonce": single.nonce}))
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
