"""K3 (DAG generation) / light cache / L1 cache parity on the GPU, through the C ABI.
Bit-exact against the oracle, the reference-made fixtures and whole-DAG digests."""
import hashlib
import random
import struct

import pytest
from conftest import golden

pytestmark = pytest.mark.gpu


def _dag_sha256(ctx, chunk_items=1 << 20):
    h = hashlib.sha256()
    m = ctx.dag_num_items
    for first in range(0, m, chunk_items):
        h.update(ctx.read_items(first, min(chunk_items, m - first)))
    return h.hexdigest()


def test_epoch0_context_fields_and_l1_kat(ctx_of, oracle):
    ctx = ctx_of(0)
    assert ctx.epoch_number == 0
    assert (ctx.light_cache_num_items, ctx.full_dataset_num_items) == (262139, 8388593)
    assert ctx.dag_num_items == 4194296
    assert ctx.l1_cache[:20] == golden("kawpow_kats.json")["l1_epoch0_first20"]  # kawpow_tests.cpp:20-39
    assert ctx.l1_cache == oracle.l1(0)
    e = golden("epochs.json")
    assert hashlib.sha256(ctx.light_cache).hexdigest() == e["light_cache_sha256_epoch0"]


@pytest.mark.parametrize("epoch", [0, 448])
def test_dag_items_vs_reference_fixture(ctx_of, epoch):
    ctx = ctx_of(epoch)
    g = golden("dag_items.json")[str(epoch)]
    assert ctx.dag_num_items == g["num_items_256"]
    for i, hx in g["items"].items():
        assert ctx.read_items(int(i), 1).hex() == hx, i
    for i, sha in g["shard_edges_sha256"].items():
        assert hashlib.sha256(ctx.read_items(int(i), 1)).hexdigest() == sha, i
    assert hashlib.sha256(ctx.read_items(0, 4096)).hexdigest() == g["first4096_sha256"]


@pytest.mark.parametrize("epoch", [0, 448])
def test_whole_dag_digest(ctx_of, epoch):
    """Every byte of the DAG: sha256 and the position-weighted sum, vs the reference-computed digests."""
    ctx = ctx_of(epoch)
    g = golden("dag_digest.json")[str(epoch)]
    assert ctx.dag_bytes == g["bytes"]
    assert ctx.checksum() == g["weighted_sum"]
    assert _dag_sha256(ctx) == g["sha256"]


def test_dag_items_vs_oracle_live(ctx_of, oracle):
    ctx = ctx_of(0)
    rnd = random.Random(11)
    m = ctx.dag_num_items
    for i in [0, 1, m - 1] + [rnd.randrange(m) for _ in range(20)]:
        assert ctx.read_items(i, 1) == oracle.dataset_item_2048(0, i)


def test_generate_range_edges(nx, ctx_of):
    """Ragged ranges (count not a multiple of the CTA's 32 items), first and last items, empty range."""
    import torch
    ctx = ctx_of(0)
    m = ctx.dag_num_items
    for first, count in [(0, 1), (5, 33), (m - 7, 7), (1000, 257), (m - 1, 1)]:
        buf = torch.zeros(count * 256 + 256, dtype=torch.uint8, device="cuda:0")
        ctx.generate_range(first, count, buf.data_ptr())
        got = bytes(buf.cpu().numpy().tobytes())
        assert got[:count * 256] == ctx.read_items(first, count)
        assert got[count * 256:] == bytes(256)  # nothing written past the range
    ctx.generate_range(10, 0, 0x1000)  # empty: no launch, no fault
    with pytest.raises(nx.NxError):
        ctx.generate_range(m - 1, 2, 0x1000)


@pytest.mark.parametrize("world", [2, 8])
def test_sharded_build_slices_equal_full_dag(nx, ctx_of, world):
    """Each rank of an N-way build writes only its slice, at its final offset; the slices tile the DAG.
    (Ranks emulated one after the other on cuda:0: no collective needed to check the slices.)"""
    full = ctx_of(0)
    m = full.dag_num_items
    share = (m + world - 1) // world
    for rank in sorted({0, 1, world - 1}):
        c = nx.EpochContext(0, 0, rank=rank, world=world)
        try:
            assert c.dag_share_bytes == share * 256 and c.dag_padded_bytes == share * world * 256
            first, last = rank * share, min(m, (rank + 1) * share)
            for a in (first, (first + last) // 2, last - 1):
                assert c.read_items(a, 1) == full.read_items(a, 1)
            h1 = hashlib.sha256(c.read_items(first, min(4096, last - first))).hexdigest()
            h2 = hashlib.sha256(full.read_items(first, min(4096, last - first))).hexdigest()
            assert h1 == h2
            with pytest.raises(nx.NxError):  # not finalized: hashing must refuse
                nx.progpow.hash(c, 0, bytes(32), 0)
        finally:
            c.close()


@pytest.mark.parametrize("kernel,epoch", [("warp", 0), ("thread", 0), ("warp", 448)])
def test_light_cache_gpu_kernels_match_the_reference_digest(nx, monkeypatch, kernel, epoch):
    """K4 (the serial chain on the device; "warp" = one Keccak-f[1600] spread over 25 lanes, "thread" = the first
    version) produces the light cache the reference's build_light_cache produces (ethash.cpp:101-129) - the whole
    cache by sha256 - and with it the same L1 cache / DAG.  Epoch 448 (72 MiB, 4.7 M permutations) is the bench's."""
    import time
    monkeypatch.setenv("NX_LIGHT_CACHE", "gpu")
    monkeypatch.setenv("NX_LIGHT_CACHE_KERNEL", kernel)
    t0 = time.perf_counter()
    c = nx.EpochContext(epoch, 0)
    try:
        light_ms, _ = c.build_times_ms()
        print(f"K4 {kernel} epoch {epoch}: light cache {light_ms:.0f} ms on the device (whole create {time.perf_counter() - t0:.2f} s)")
        e = golden("epochs.json")
        assert hashlib.sha256(c.light_cache).hexdigest() == e[f"light_cache_sha256_epoch{epoch}"]
        assert hashlib.sha256(struct.pack("<4096I", *c.l1_cache)).hexdigest() == e[f"l1_sha256_epoch{epoch}"]
    finally:
        c.close()


@pytest.mark.parametrize("route", [[], ["--fused"], ["--shell", "--fused"], ["--shell"]])
def test_two_gpu_sharded_build_and_search(nx, route):
    """Real N=2 paths when the box has two GPUs: the DAG slices exchanged by one NCCL all-gather or by peer stores inside
    the generating kernel, each with every rank running its own light-cache chain or with ONE chain whose result is
    broadcast (--shell); then a sharded search.  Every rank checks the reference-made digests."""
    import subprocess
    import sys

    from conftest import ROOT
    if nx.lib.nx_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", "tools/multigpu_check.py"] + route,
                         cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert '"multigpu_check": "ok"' in out.stdout


def test_epoch_that_does_not_fit_the_device_fails_cleanly(nx, gpu, ctx_of):
    """ethash.cpp:147-148: the reference returns nullptr when the context cannot be allocated.  Epoch 29999 wants a
    235 GiB DAG (more than a B200 has): NX_ERR_OOM with a message, nothing leaked, and the device still works."""
    from nodexa_chain_core_b200._lib import NxError

    with pytest.raises(NxError) as e:
        nx.create_epoch_context_full(29999, gpu)
    assert e.value.code == -4 and "cudaMalloc" in str(e.value)
    ctx = ctx_of(0)
    assert ctx.l1_cache[:20] == golden("kawpow_kats.json")["l1_epoch0_first20"]


def test_contexts_give_all_device_memory_back(nx, gpu):
    """A node changes epoch every 7500 blocks for as long as it runs: creating and destroying contexts (directly, and through
    the miner's epoch hand-over with its background prebuild) must not leak device memory."""
    import torch
    torch.cuda.init()
    torch.cuda.synchronize()

    def free_bytes():
        torch.cuda.synchronize()
        return torch.cuda.mem_get_info(0)[0]

    nx.EpochContext(0, gpu).close()   # first use: CUDA context, module loading, allocator pools settle
    base = free_bytes()
    for e in (0, 1, 2, 0, 1, 2):
        c = nx.EpochContext(e, gpu)
        assert free_bytes() < base - (1 << 30)          # the 1 GiB+ DAG is really resident
        c.close()
    assert abs(free_bytes() - base) <= (8 << 20), (base, free_bytes())
    with nx.Miner([gpu]) as m:
        hd = bytes(range(32))
        for height in (7499, 7500, 15000, 7499):
            m.prepare_async(height, lookahead_blocks=2)
            assert m.search(height, hd, b"\xff" * 32, 0, 1).solution_found
        m.wait_prepared()
    assert abs(free_bytes() - base) <= (8 << 20), (base, free_bytes())


@pytest.mark.parametrize("world", [2, 8])
def test_fused_exchange_on_one_device_gives_the_reference_dag(nx, world):
    """The fused multi-GPU build (every rank's K3 stores its slice into the DAG buffer of EVERY rank) with all `world`
    ranks on cuda:0 - a pointer on the same device is a valid "peer" - so that the path is pinned on a 1-GPU box:
    every rank's whole DAG must carry the reference-computed digest (ethash.cpp:209-251 item by item)."""
    want = golden("dag_digest.json")["0"]["weighted_sum"]
    first = nx.EpochContext(0, 0, rank=0, world=world, deferred=True)
    ctxs = [first]
    try:
        light = first.light_cache
        ctxs += [nx.EpochContext(0, 0, rank=r, world=world, deferred=True, light_cache=light) for r in range(1, world)]
        with pytest.raises(nx.NxError):  # nothing generated yet: finalize must refuse instead of marking the DAG ready
            ctxs[0].finalize()
        ptrs = [c.dag_ptr for c in ctxs]
        holes = list(ptrs)
        holes[world - 1] = 0
        with pytest.raises(nx.NxError):  # a rank without a buffer: refused, no store through a null pointer
            ctxs[0].generate_fused_peers(holes)
        for c in ctxs:
            c.generate_fused_peers(ptrs)
        for c in ctxs:
            c.finalize()
        for c in ctxs:
            assert c.checksum() == want, c.rank
        with pytest.raises(nx.NxError):  # a second generation on a finished context is a state error
            ctxs[0].generate_fused_peers(ptrs)
        kat = golden("kawpow_kats.json")["hash_cases"][0]
        r = nx.progpow.hash(ctxs[world - 1], kat["height"], bytes.fromhex(kat["header"]), kat["nonce"])
        assert r.final_hash.hex() == kat["final"] and r.mix_hash.hex() == kat["mix"]
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("world", [2, 8])
def test_sharded_slices_copied_together_give_the_reference_dag(nx, world):
    """The other exchange route (nx_epoch_create_sharded + an in-place all-gather, NCCL in production) with the gather
    done by device-to-device copies on cuda:0: rank 0's buffer, completed with every other rank's slice, has the digest."""
    import torch
    want = golden("dag_digest.json")["0"]["weighted_sum"]
    c0 = nx.EpochContext(0, 0, rank=0, world=world)
    try:
        dag0 = c0.dag_tensor()
        share = c0.dag_share_bytes
        for r in range(1, world):
            c = nx.EpochContext(0, 0, rank=r, world=world)
            try:
                dag0[r * share:(r + 1) * share].copy_(c.dag_tensor()[r * share:(r + 1) * share])
                torch.cuda.synchronize()
            finally:
                c.close()
        c0.finalize()
        assert c0.checksum() == want
    finally:
        c0.close()


def test_one_chain_for_the_job_shell_contexts_on_one_device(nx):
    """The multi-process epoch build in which ONE rank runs the light-cache chain (nx_epoch_create_shell /
    nx_epoch_build_light_cache / nx_epoch_adopt_light_cache), with both "ranks" on cuda:0 and a device-to-device copy
    standing in for the NCCL broadcast: same light cache (sha256 against the reference's), same DAG digest on both, through
    the fused route on one pair of contexts and the slice + gather route on another; and the state checks in between."""
    import torch
    e = golden("epochs.json")
    want = golden("dag_digest.json")["0"]["weighted_sum"]
    for route in ("fused", "slice"):
        a = nx.EpochContext(0, 0, rank=0, world=2, shell=True)
        b = nx.EpochContext(0, 0, rank=1, world=2, shell=True)
        try:
            with pytest.raises(nx.NxError):      # no light cache yet: nothing may be generated from it
                a.generate_slice()
            with pytest.raises(nx.NxError):
                a.generate_fused_peers([a.dag_ptr, b.dag_ptr])
            a.build_light_cache()                # the one chain of the job
            with pytest.raises(nx.NxError):      # a second light cache for the same context is a state error
                a.build_light_cache()
            b.light_tensor().copy_(a.light_tensor())   # stands for ncclBroadcast into nx_epoch_light_ptr()
            torch.cuda.synchronize()
            b.adopt_light_cache()
            with pytest.raises(nx.NxError):
                b.adopt_light_cache()
            for c in (a, b):
                assert hashlib.sha256(c.light_cache).hexdigest() == e["light_cache_sha256_epoch0"]
            if route == "fused":
                ptrs = [a.dag_ptr, b.dag_ptr]
                a.generate_fused_peers(ptrs)
                b.generate_fused_peers(ptrs)
            else:
                a.generate_slice()
                b.generate_slice()
                share = a.dag_share_bytes
                a.dag_tensor()[share:2 * share].copy_(b.dag_tensor()[share:2 * share])
                b.dag_tensor()[:share].copy_(a.dag_tensor()[:share])
                torch.cuda.synchronize()
            a.finalize()
            b.finalize()
            assert a.checksum() == want and b.checksum() == want
            assert hashlib.sha256(struct.pack("<4096I", *b.l1_cache)).hexdigest() == e["l1_sha256_epoch0"]
        finally:
            a.close()
            b.close()
