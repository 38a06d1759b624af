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
        monkeypatch.setenv("NX_MINER_INJECT_FAIL", f"{ndev - 1}:2")  # last device fails on its third window
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
