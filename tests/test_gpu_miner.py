"""nx_miner_*: the node's nonce loop as one call, over every visible GPU, through the C ABI."""
import hashlib
import struct

import pytest
from conftest import golden

from nodexa_chain_core_b200 import progpow

pytestmark = pytest.mark.gpu


def test_miner_search_equals_single_context_search(nx, ctx_of):
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"miner").digest()
    boundary = ((1 << (256 - 17)) - 1).to_bytes(32, "big")
    n = (1 << 21) + 777
    want = progpow.search(ctx, 600, hd, boundary, 5 << 40, n)
    with nx.Miner() as m:
        assert m.device_count == nx.lib.nx_device_count()
        for batch in (1 << 12, 1 << 16, 1 << 22):  # many small windows / few large ones: same lowest nonce
            m.set_batch(batch)
            got = m.search(600, hd, boundary, 5 << 40, n)
            assert got == want, batch
        hashes, secs = m.last_stats()
        assert 0 < hashes <= n and secs > 0
        assert m.epoch == 0
        # the reference's search KAT through the miner (kawpow_tests.cpp:109-140)
        m.set_batch(16)
        kat = m.search(0, bytes(32), bytes.fromhex("00" + "ff" * 31), 300, 100)
        assert kat.solution_found and kat.nonce == 395
        miss = m.search(0, bytes(32), bytes.fromhex("00" + "ff" * 31), 700, 100)
        assert not miss.solution_found and miss.nonce == 0 and miss.final_hash == bytes(32)


def test_mine_header_like_generate_blocks(nx, oracle):
    """rpc/mining.cpp:141-158: scan nNonce64 until CheckProofOfWork passes, keep the mix hash."""
    g = golden("kawpow_config1.json")
    raw = bytes.fromhex(g["header_input_80B"])
    version, prev, merkle, t, _bits, height = struct.unpack("<i32s32sIII", raw)
    bits = 0x2000ffff  # kawpowLimit-sized target: roughly one nonce in 256 qualifies
    with nx.Miner([0]) as m:
        sr = m.mine_header(version, prev, merkle, t, bits, height, start_nonce=0, max_tries=1 << 16)
        assert sr.solution_found
        hh = oracle.kawpow_header_hash(version, prev, merkle, t, bits, height)
        ok, boundary = oracle.compact_to_boundary(bits)
        assert ok
        f, mx = oracle.kawpow_hash(0, height, hh, sr.nonce)   # the oracle agrees with the block that was "found"
        assert (f, mx) == (sr.final_hash, sr.mix_hash) and f <= boundary
        ctx = nx.create_epoch_context(0, 0)
        try:
            lower = progpow.hash_batch(ctx, height, hh, 0, sr.nonce) if sr.nonce else b""
            assert all(lower[64 * i:64 * i + 32] > boundary for i in range(sr.nonce))  # no lower nonce qualified
        finally:
            ctx.close()
        with pytest.raises(nx.NxError):
            m.mine_header(version, prev, merkle, t, 0x01800000, height)  # negative target: refused like pow.cpp:190


def test_background_prepare_next_epoch(nx):
    with nx.Miner([0]) as m:
        m.set_epoch(0)
        m.prepare_async(7498, lookahead_blocks=4)   # two blocks before the epoch boundary
        m.wait_prepared()
        assert m.prepared_epoch == 1
        hd = bytes(range(32))
        r0 = m.search(7499, hd, b"\xff" * 32, 0, 1)
        assert m.epoch == 0 and r0.solution_found
        r1 = m.search(7500, hd, b"\xff" * 32, 0, 1)  # adopts the prebuilt epoch-1 contexts
        assert m.epoch == 1 and m.prepared_epoch == -1 and r1.solution_found
        assert r1.final_hash != r0.final_hash


def test_failed_device_drops_out_and_its_windows_are_reissued(nx, ctx_of, monkeypatch):
    """Failure handling: a device whose launch fails leaves the search, its window goes back to the pool, and the
    answer is still the lowest nonce.  Needs two GPUs to have a survivor; with one GPU the call must fail loudly."""
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"failover").digest()
    boundary = ((1 << (256 - 18)) - 1).to_bytes(32, "big")
    n = 1 << 20
    want = progpow.search(ctx, 900, hd, boundary, 1 << 50, n)
    ndev = nx.lib.nx_device_count()
    with nx.Miner() as m:
        m.set_batch(1 << 14)
        m.set_epoch(0)
        # The last device fails on its third window (two GPUs) or on its first (more GPUs: the first hit lies around
        # window 16, which eight devices reach before any of them has claimed a third window).
        monkeypatch.setenv("NX_MINER_INJECT_FAIL", f"{ndev - 1}:{2 if ndev <= 2 else 0}")
        if ndev >= 2:
            got = m.search(900, hd, boundary, 1 << 50, n)
            assert got == want
            assert m.last_failed_devices == 1
        else:
            with pytest.raises(nx.NxError):
                m.search(900, hd, bytes(32), 1 << 50, n)
            assert m.last_failed_devices == 1
        monkeypatch.delenv("NX_MINER_INJECT_FAIL")
        assert m.search(900, hd, boundary, 1 << 50, n) == want and m.last_failed_devices == 0


def test_cancel_flag_stops_a_long_search(nx):
    """miner.cpp:693-712: the outer loop must be able to abandon the scan (new tip, stale template)."""
    import threading
    import time
    with nx.Miner([0]) as m:
        m.set_epoch(0)
        m.set_batch(1 << 20)
        m.search(5, bytes(32), bytes(32), 0, 1 << 20)  # compile the period program first
        out = {}

        def run():
            t0 = time.perf_counter()
            out["r"] = m.search(5, bytes(32), bytes(32), 0, 1 << 40)  # ~90 minutes of work if left alone
            out["s"] = time.perf_counter() - t0

        th = threading.Thread(target=run)
        th.start()
        time.sleep(0.3)
        m.cancel()
        th.join(timeout=30)
        assert not th.is_alive() and out["s"] < 5.0
        assert not out["r"].solution_found
        hashes, _ = m.last_stats()
        assert 0 < hashes < 1 << 34


def test_program_cache_dir_persists_compiled_periods(tmp_path):
    """NX_PROGRAM_CACHE_DIR: the cubin of a period is written once and reused by the next process."""
    import os
    import subprocess
    import sys

    from conftest import ROOT
    code = ("import nodexa_chain_core_b200 as nx; from nodexa_chain_core_b200 import progpow;"
            "c = nx.create_epoch_context(0, 0); print(progpow.hash_batch(c, 300, bytes(32), 0, 4).hex()); c.close()")
    env = dict(os.environ, NX_PROGRAM_CACHE_DIR=str(tmp_path), NX_VERBOSE="1")
    first = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert first.returncode == 0, first.stderr[-2000:]
    files = sorted(os.listdir(tmp_path))
    assert len(files) == 1 and files[0].startswith("kawpow_sm_100a_100_4194296_") and files[0].endswith(".cubin")
    second = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert second.returncode == 0 and second.stdout == first.stdout
    assert sorted(os.listdir(tmp_path)) == files

    def compile_ms(stderr):
        import re
        return float(re.search(r": ([0-9.]+) ms, regs", stderr).group(1))
    assert compile_ms(second.stderr) < compile_ms(first.stderr)  # loading a cubin beats NVRTC
    # a damaged cache file (a crash mid-write under an older version, a foreign file) is ignored and replaced
    path = os.path.join(tmp_path, files[0])
    good = open(path, "rb").read()
    open(path, "wb").write(good[:1000].replace(b"\x7fELF", b"\x7fELG"))
    third = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert third.returncode == 0 and third.stdout == first.stdout, third.stderr[-2000:]
    assert sorted(os.listdir(tmp_path)) == files and len(open(path, "rb").read()) == len(good)


def test_next_period_is_compiled_in_the_background(tmp_path):
    """NX_PREFETCH=1: the use of a period starts the NVRTC run of the next one on a background thread (the period
    changes every 3 blocks, progpow.hpp:21), so that the first call of the next period costs a module load instead of a
    compilation.  Seen from outside through the cubin directory, and through the clock; a process that exits while the
    background run is still going waits for it first and exits cleanly."""
    import json
    import os
    import subprocess
    import sys

    from conftest import ROOT
    code = r"""
import json, sys, time
import nodexa_chain_core_b200 as nx
from nodexa_chain_core_b200 import progpow
c = nx.create_epoch_context(0, 0)
t0 = time.perf_counter(); progpow.prepare(c, 6000); cold = time.perf_counter() - t0    # period 2000: a full compile
assert len(progpow.hash_batch(c, 6000, bytes(32), 0, 4)) == 256                         # its next use starts 2001
if sys.argv[1] == "wait":
    time.sleep(2.0 * cold + 0.5)
    t0 = time.perf_counter(); progpow.prepare(c, 6003); warm = time.perf_counter() - t0
    print(json.dumps({"cold": cold, "warm": warm}))
c.close()
"""
    env = dict(os.environ, NX_PROGRAM_CACHE_DIR=str(tmp_path), NX_PREFETCH="1")
    run = subprocess.run([sys.executable, "-c", code, "wait"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, run.stderr[-2000:]
    t = json.loads(run.stdout.strip().splitlines()[-1])
    assert t["warm"] < 0.25 * t["cold"], t
    seeds = sorted(int(f.split("_")[3]) for f in os.listdir(tmp_path) if f.endswith(".cubin"))
    assert seeds[:2] == [2000, 2001], seeds
    # exit right behind the call that started the background run: the process must wait for it and leave with rc 0
    for f in os.listdir(tmp_path):
        os.remove(os.path.join(tmp_path, f))
    run = subprocess.run([sys.executable, "-c", code, "exit"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert run.returncode == 0, (run.returncode, run.stderr[-2000:])
    seeds = sorted(int(f.split("_")[3]) for f in os.listdir(tmp_path) if f.endswith(".cubin"))
    assert seeds == [2000, 2001], seeds
    # without the switch nothing is compiled ahead
    for f in os.listdir(tmp_path):
        os.remove(os.path.join(tmp_path, f))
    env.pop("NX_PREFETCH")
    run = subprocess.run([sys.executable, "-c", code, "exit"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert run.returncode == 0
    assert sorted(int(f.split("_")[3]) for f in os.listdir(tmp_path) if f.endswith(".cubin")) == [2000]


def test_chain_across_period_switches_and_an_epoch_boundary(nx, oracle):
    """Block after block like a node (tools/soak_chain.py): 24 linked blocks from height 7488 to 7511 - eight ProgPoW
    periods, epoch 0 -> 1 with the next epoch prebuilt in the background - every header then passes the batched
    CheckBlockHeader path and the oracle."""
    import importlib.util
    import os
    from nodexa_chain_core_b200 import node
    spec = importlib.util.spec_from_file_location("soak_chain", os.path.join(os.path.dirname(__file__), "..", "tools", "soak_chain.py"))
    soak = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(soak)
    headers, secs, _ = soak.mine_chain(nx, node, 7488, 24, 0x2000ffff, lookahead=6, devices=[0], batch=1 << 12)
    assert [h.height for h, _ in headers] == list(range(7488, 7512))
    for e in (0, 1):
        ctx = nx.EpochContext(e, 0)
        try:
            hs = [h for h, _ in headers if h.height // 7500 == e]
            assert len(hs) == 12 and node.check_block_headers(ctx, hs) == [node.HEADER_OK] * 12
        finally:
            ctx.close()
    for (h, sr), (nxt, _) in zip(headers, headers[1:]):
        assert nxt.prev_block == h.block_hash() == sr.final_hash[::-1]
    for h, sr in headers[::5]:
        assert oracle.kawpow_hash(h.height // 7500, h.height, h.kawpow_header_hash(), h.nonce64) == (sr.final_hash, sr.mix_hash)


def test_unbounded_try_count_scans_until_the_hit(nx):
    """max_tries = 2^64 - 1 is a natural "no limit": the window count must not wrap to zero (it once did), the scan
    starts at start_nonce and stops at the first qualifying nonce."""
    hd = hashlib.sha256(b"no-limit").digest()
    boundary = ((1 << (256 - 10)) - 1).to_bytes(32, "big")
    with nx.Miner([0]) as m:
        m.set_batch(1 << 16)
        want = m.search(0, hd, boundary, 1, 1 << 20)
        assert want.solution_found
        got = m.search(0, hd, boundary, 1, 2**64 - 2)  # start + iterations stays below 2^64: the reference's loop runs
        assert got == want


def test_fused_build_on_one_device_and_a_rank_that_fails(nx, monkeypatch):
    """The miner's fused multi-GPU DAG build (each device generates 1/N and stores it into every device's buffer) runs
    on a 1-GPU box when the same device is listed twice.  When one entry's context cannot be created, the whole build
    must be given up BEFORE the generating kernels run - they would store through a buffer that does not exist - with
    an error, no fault on the device, and a later build must work and hash correctly."""
    hd = bytes(32)
    kat_boundary = bytes.fromhex("00" + "ff" * 31)
    with nx.Miner([0, 0]) as m:
        assert m.device_count == 2
        monkeypatch.setenv("NX_MINER_INJECT_CREATE_FAIL", "1")
        with pytest.raises(nx.NxError) as e:
            m.set_epoch(0)
        assert "injected failure" in str(e.value)
        monkeypatch.delenv("NX_MINER_INJECT_CREATE_FAIL")
        m.set_epoch(0)                                         # the device is healthy: same miner, clean build
        m.set_batch(16)
        kat = m.search(0, hd, kat_boundary, 300, 100)          # kawpow_tests.cpp:109-140 over two contexts of one GPU
        assert kat.solution_found and kat.nonce == 395
    with nx.Miner([0]) as one:
        one.set_batch(16)
        assert one.search(0, hd, kat_boundary, 300, 100) == kat
