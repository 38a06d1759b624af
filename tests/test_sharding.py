"""N>1 host logic on the CPU: DAG slices tile the item range, nonce super-batches tile the nonce range, and
the lowest-nonce reduction over world_size-2 `gloo` picks the right rank's hit."""
import os
import socket

from nodexa_chain_core_b200 import sharding


def test_dag_slices_tile_the_dag():
    for m in (4194296, 18874358, 16777196, 7, 1):
        for world in (1, 2, 3, 4, 8):
            share = sharding.dag_share_items(m, world)
            assert share * world >= m and (share - 1) * world < m
            covered = []
            for r in range(world):
                a, b = sharding.dag_slice(m, r, world)
                assert 0 <= a <= b <= m and b - a <= share
                covered.append((a, b))
            assert covered[0][0] == 0 and covered[-1][1] == m
            for (a0, b0), (a1, b1) in zip(covered, covered[1:]):
                assert b0 == a1


def test_super_batches_tile_the_nonce_range():
    for world in (1, 2, 4, 8):
        for iters in (0, 1, 1000, 65536, 65536 * 8 + 17):
            plan = sharding.ShardedSearchPlan(start_nonce=2**64 - 300, iterations=iters, batch=4096, world=world)
            seen = 0
            expect_off = 0
            for k in range(plan.rounds):
                for r in range(world):
                    s, n = plan.window(k, r)
                    if n:
                        assert s == (2**64 - 300 + expect_off) % 2**64  # wraps like the reference's uint64_t
                        expect_off += n
                        seen += n
            assert seen == iters


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out = []
        # (a) nobody hits, (b) only rank 1 hits, (c) both hit: the lower offset wins whichever rank holds it
        cases = [(None, None), (None, 77), (5000, 4999), (12, 900)]
        for c in cases:
            out.append(sharding.reduce_lowest_offset(c[rank], dist))
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


def test_lowest_offset_reduction_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = dict(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] == got[1] == [None, 77, 4999, 12]
