"""KawPow parity on the GPU through the C ABI: every mix and final hash bit-exact against the
reference's KATs, the reference-made fixtures (BASELINE config 1: 2^16 nonces) and the oracle live."""
import hashlib
import random

import pytest
from conftest import golden

from nodexa_chain_core_b200 import progpow

pytestmark = pytest.mark.gpu
KATS = golden("kawpow_kats.json")


def test_hash_and_verify_kats(ctx_of):
    """kawpow_tests.cpp:75-107 (13 vectors, each with its two negative verify cases) + :58-73"""
    for t in KATS["hash_cases"] + [KATS["hash_30000"]]:
        ctx = ctx_of(t["height"] // 7500)
        hd = bytes.fromhex(t["header"])
        r = progpow.hash(ctx, t["height"], hd, t["nonce"])
        assert r.mix_hash.hex() == t["mix"], t["height"]
        assert r.final_hash.hex() == t["final"], t["height"]
        assert progpow.verify(ctx, t["height"], hd, r.mix_hash, t["nonce"], r.final_hash)
        lower = bytearray(r.final_hash)
        lower[31] = (lower[31] - 1) & 0xff
        assert not progpow.verify(ctx, t["height"], hd, r.mix_hash, t["nonce"], bytes(lower))
        bad = bytearray(r.mix_hash)
        bad[7] = (bad[7] + 1) & 0xff
        assert not progpow.verify(ctx, t["height"], hd, bytes(bad), t["nonce"], r.final_hash)
        assert progpow.hash_no_verify(t["height"], hd, r.mix_hash, t["nonce"]) == r.final_hash


def test_hash_empty_descending_heights(ctx_of):
    """kawpow_tests.cpp:41-56 hashes heights 1000..1 with the epoch-0 context and checks the last one.
    (A subset of the heights: each distinct period costs one NVRTC compile.)"""
    ctx = ctx_of(0)
    r = None
    for h in list(range(1000, 940, -7)) + list(range(12, 0, -1)):
        r = progpow.hash(ctx, h, bytes(32), 0)
    assert r.mix_hash.hex() == KATS["hash_cases"][0]["mix"]
    assert r.final_hash.hex() == KATS["hash_cases"][0]["final"]


def test_search_kat(ctx_of):
    """kawpow_tests.cpp:109-140"""
    ctx = ctx_of(0)
    s = KATS["search"]
    hd, b = bytes.fromhex(s["header"]), bytes.fromhex(s["boundary"])
    sr = progpow.search(ctx, 0, hd, b, s["miss"]["start"], s["miss"]["iterations"])
    assert (sr.solution_found, sr.nonce, sr.final_hash, sr.mix_hash) == (False, 0, bytes(32), bytes(32))
    srl = progpow.search_light(ctx, 0, hd, b, s["miss"]["start"], s["miss"]["iterations"])
    assert srl == sr
    sr = progpow.search(ctx, 0, hd, b, s["hit"]["start"], s["hit"]["iterations"])
    assert sr.solution_found and sr.nonce == s["hit"]["nonce"]
    r = progpow.hash(ctx, 0, hd, 395)
    assert (sr.final_hash, sr.mix_hash) == (r.final_hash, r.mix_hash)
    assert sr.final_hash != bytes(32) and sr.mix_hash != bytes(32)


def test_config1_all_65536_nonces(ctx_of):
    """BASELINE.json configs[0]: 2^16 nonces over the fixed synthetic header at epoch 0, plus 2^12 nonces at
    heights 1, 2, 3, 7499 — every {final, mix} against the fixture produced by the reference's progpow::hash."""
    g = golden("kawpow_config1.json")
    hh = bytes.fromhex(g["header_hash"])
    ctx = ctx_of(0)
    for run in g["runs"]:
        buf = progpow.hash_batch(ctx, run["height"], hh, run["start"], run["count"])
        assert len(buf) == run["count"] * 64
        assert buf[:16 * 64].hex() == run["first16"]
        per = [hashlib.sha256(buf[i:i + 256 * 64]).hexdigest() for i in range(0, len(buf), 256 * 64)]
        bad = [i for i, (a, b) in enumerate(zip(per, run["sha256_per_256_nonces"])) if a != b]
        assert not bad, f"height {run['height']}: mismatching 256-nonce chunks {bad[:8]}"
        assert hashlib.sha256(buf).hexdigest() == run["sha256_all"]
    z = progpow.hash_batch(ctx, 0, bytes(32), 0, 1024)
    assert hashlib.sha256(z).hexdigest() == g["zero_header_first1024_sha256"]


def test_epoch448_vs_reference_fixture(ctx_of):
    """BASELINE.json configs[1]'s epoch (4.5 GiB DAG): 8 headers x 16 nonces made by the reference's code."""
    k = golden("kawpow_epoch448.json")
    ctx = ctx_of(448)
    for c in k["cases"]:
        buf = progpow.hash_batch(ctx, c["height"], bytes.fromhex(c["header"]), c["start"], c["count"])
        assert buf.hex() == c["out"], c["height"]


def test_random_inputs_vs_oracle_live(ctx_of, oracle):
    rnd = random.Random(20261106)
    ctx = ctx_of(0)
    for _ in range(6):
        hd = rnd.randbytes(32)
        height = rnd.randrange(0, 7500)
        start = rnd.getrandbits(64) - 40
        got = progpow.hash_batch(ctx, height, hd, start & (2**64 - 1), 24)
        want = oracle.kawpow_hash_range(0, height, hd, start & (2**64 - 1), 24)
        assert got == want


def test_ragged_and_empty_batches(ctx_of, oracle):
    ctx = ctx_of(0)
    hd = bytes(range(32))
    ref_1k = progpow.hash_batch(ctx, 5, hd, 1000, 1024)
    assert progpow.hash_batch(ctx, 5, hd, 1000, 0) == b""
    for count in (1, 15, 16, 17, 255, 256, 257, 1000):
        assert progpow.hash_batch(ctx, 5, hd, 1000, count) == ref_1k[:count * 64]
    # a window that does not start on a 16-nonce boundary relative to the first batch
    assert progpow.hash_batch(ctx, 5, hd, 1007, 300) == ref_1k[7 * 64:307 * 64]
    assert ref_1k[:8 * 64] == oracle.kawpow_hash_range(0, 5, hd, 1000, 8)
    # nonce arithmetic wraps modulo 2^64 like the reference's uint64_t
    top = progpow.hash_batch(ctx, 5, hd, 2**64 - 4, 4)
    assert top == oracle.kawpow_hash_range(0, 5, hd, 2**64 - 4, 4)


def test_hash_batch_device_equals_host(ctx_of):
    import torch
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"dev").digest()
    n = 5000
    out = torch.zeros(n * 64 + 64, dtype=torch.uint8, device="cuda:0")
    progpow.hash_batch_device(ctx, 77, hd, 123456789, n, out.data_ptr())
    got = out.cpu().numpy().tobytes()
    assert got[:n * 64] == progpow.hash_batch(ctx, 77, hd, 123456789, n)
    assert got[n * 64:] == bytes(64)


def test_search_returns_lowest_nonce_across_batches(ctx_of):
    """progpow.cpp:572-577: the FIRST hit of an ascending scan, not the first to be found.  Checked against
    a host-side scan of the batch-hash output, over a range spanning several internal launches."""
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"search").digest()
    n = (1 << 22) + 12345  # more than one internal batch, ragged tail
    start = 0x1234567800000000
    allh = progpow.hash_batch(ctx, 300, hd, start, n)
    finals = [allh[i * 64:i * 64 + 32] for i in range(n)]
    for bits in (12, 20):  # about n/2^bits hits
        boundary = ((1 << (256 - bits)) - 1).to_bytes(32, "big")
        hits = [i for i, f in enumerate(finals) if f <= boundary]
        sr = progpow.search(ctx, 300, hd, boundary, start, n)
        if hits:
            assert sr.solution_found and sr.nonce == start + hits[0]
            assert sr.final_hash == finals[hits[0]] and sr.mix_hash == allh[hits[0] * 64 + 32:hits[0] * 64 + 64]
            # a window that ends just before the first hit finds nothing; one that starts after finds the next
            sr0 = progpow.search(ctx, 300, hd, boundary, start, hits[0])
            assert not sr0.solution_found and sr0.nonce == 0
            if len(hits) > 1:
                sr1 = progpow.search(ctx, 300, hd, boundary, start + hits[0] + 1, n - hits[0] - 1)
                assert sr1.solution_found and sr1.nonce == start + hits[1]
        else:
            assert not sr.solution_found
    # exact-boundary case: final == boundary qualifies (is_less_or_equal)
    sr = progpow.search(ctx, 300, hd, finals[100], start + 100, 1)
    assert sr.solution_found and sr.nonce == start + 100
    below = (int.from_bytes(finals[100], "big") - 1).to_bytes(32, "big")
    assert not progpow.search(ctx, 300, hd, below, start + 100, 1).solution_found
    assert not progpow.search(ctx, 300, hd, boundary, start, 0).solution_found
    # uint64_t wrap-around: the reference's `nonce < start + iterations` loop does not run when the end wraps
    easy = b"\xff" * 32
    assert progpow.search(ctx, 300, hd, easy, 2**64 - 10, 9).nonce == 2**64 - 10
    assert not progpow.search(ctx, 300, hd, easy, 2**64 - 10, 10).solution_found
    assert not progpow.search(ctx, 300, hd, easy, 2**64 - 10, 500).solution_found


def test_search_enqueue_sync_pipelining(ctx_of):
    """Several enqueued batches, one sync: the lowest hit over all of them."""
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"pipe").digest()
    boundary = ((1 << (256 - 14)) - 1).to_bytes(32, "big")
    ref = progpow.search(ctx, 9, hd, boundary, 5000, 3 * 65536)
    for k in range(3):
        progpow.search_enqueue(ctx, 9, hd, boundary, 5000 + k * 65536, 65536)
    sr = progpow.search_sync(ctx)
    assert sr == ref
    ms, launches = ctx.last_timing()
    assert launches >= 3 and ms > 0
    assert progpow.search_sync(ctx) == progpow.SearchResult()  # nothing pending


def test_concurrent_host_threads_share_one_context(ctx_of):
    """hash / search / verify are re-entrant on a shared context (SURVEY.md §8b threading): several host
    threads hammer one context at once (ctypes releases the GIL) and every answer must still be right."""
    import threading
    ctx = ctx_of(0)
    boundary = ((1 << (256 - 13)) - 1).to_bytes(32, "big")
    jobs = []
    for t in range(6):
        hd = hashlib.sha256(b"thread%d" % t).digest()
        jobs.append((hd, progpow.hash_batch(ctx, 30 + t, hd, t << 20, 2000), progpow.search(ctx, 30 + t, hd, boundary, t << 20, 50000)))
    errors = []

    def work(t):
        hd, want_hashes, want_search = jobs[t]
        try:
            for _ in range(3):
                assert progpow.hash_batch(ctx, 30 + t, hd, t << 20, 2000) == want_hashes
                assert progpow.search(ctx, 30 + t, hd, boundary, t << 20, 50000) == want_search
                r = progpow.hash(ctx, 30 + t, hd, (t << 20) + 7)
                assert r.final_hash + r.mix_hash == want_hashes[7 * 64:8 * 64]
                assert progpow.verify(ctx, 30 + t, hd, r.mix_hash, (t << 20) + 7, r.final_hash)
        except Exception as e:  # noqa: BLE001
            errors.append((t, repr(e)))

    th = [threading.Thread(target=work, args=(t,)) for t in range(6)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errors, errors


def test_epoch448_1024_nonces_and_4096_dag_items_vs_reference_live(ctx_of, ref):
    """SURVEY.md §8(d) config 2's parity spot check at the bench's own epoch, against the reference's compiled
    code running on this box's host cores: 1024 consecutive nonces (light path) and 4096 random DAG items."""
    ctx = ctx_of(448)
    height = 448 * 7500 + 11
    hd = hashlib.sha256(b"epoch448-live").digest()
    start = 0x448 << 40
    want = ref.kawpow_hash_range(448, height, hd, start, 1024)
    assert progpow.hash_batch(ctx, height, hd, start, 1024) == want
    rnd = random.Random(448)
    m = ctx.dag_num_items
    idx = sorted(rnd.randrange(m) for _ in range(4096))
    for i in idx[::64]:  # 64 single-item reads spread over the DAG ...
        assert ctx.read_items(i, 1) == ref.dataset_item_2048(448, i)
    first = rnd.randrange(m - 4096)  # ... and one run of 4096 consecutive items
    assert ctx.read_items(first, 4096) == ref.dataset_items_2048(448, first, 4096)


def test_cached_period_is_not_blocked_by_a_compile_in_another_thread(ctx_of):
    """The miner compiles period p+1 in the background while period p is being searched: a lookup of the cached
    program must not queue behind that NVRTC run (nx_kawpow_rt.cpp keeps the map lock and the compile lock apart)."""
    import threading
    import time

    ctx = ctx_of(0)
    progpow.prepare(ctx, 3000)  # period 1000, now cached
    t0 = time.perf_counter()
    progpow.prepare(ctx, 3003)  # a fresh period: one full compile, for scale
    compile_s = time.perf_counter() - t0
    th = threading.Thread(target=progpow.prepare, args=(ctx, 3006))  # ctypes releases the GIL during the call
    th.start()
    time.sleep(min(0.05, compile_s / 4))  # let it get into NVRTC
    t0 = time.perf_counter()
    got = progpow.hash_batch(ctx, 3000, bytes(32), 0, 64)
    cached_s = time.perf_counter() - t0
    th.join()
    assert len(got) == 64 * 64
    assert cached_s < 0.5 * compile_s, (cached_s, compile_s)


@pytest.mark.parametrize("epoch", [1, 129, 300])
def test_other_epochs_live_vs_oracle(nx, oracle, epoch):
    """Epochs no fixture covers (other DAG sizes, other prime divisors for every `% n` on the path): DAG items, compiled
    KawPow (K2), interpreted KawPow (K6) and classic Ethash (K7) against the CPU oracle running on this box."""
    from nodexa_chain_core_b200 import ethash
    rnd = random.Random(epoch)
    ctx = nx.EpochContext(epoch, 0)
    try:
        assert ctx.dag_num_items == oracle.dataset_items(epoch) // 2 and ctx.light_cache_num_items == oracle.light_items(epoch)
        m = ctx.dag_num_items
        for first in (0, m - 8, rnd.randrange(m - 8)):
            assert ctx.read_items(first, 8) == oracle.dataset_items_2048(epoch, first, 8)
        height = epoch * 7500 + rnd.randrange(7500)
        hd = rnd.randbytes(32)
        start = rnd.getrandbits(64) - 64
        assert progpow.hash_batch(ctx, height, hd, start, 48) == oracle.kawpow_hash_range(epoch, height, hd, start, 48)
        recs = [(epoch * 7500 + rnd.randrange(7500), rnd.randbytes(32), rnd.getrandbits(64)) for _ in range(12)]
        got = progpow.hash_many(ctx, recs)
        for (bn, h, n), r in zip(recs, got):
            assert (r.final_hash, r.mix_hash) == oracle.kawpow_hash(epoch, bn, h, n)
        for n in (0, rnd.getrandbits(64)):
            assert ethash.hash(ctx, hd, n) == oracle.ethash_hash(epoch, hd, n)
    finally:
        ctx.close()


def test_small_call_shapes_are_bit_exact(ctx_of, oracle):
    """A call too small to give every resident warp 32 nonces is spread thinner (1..15 nonces per half-warp instead of
    16, nx_kawpow_rt.cpp kawpow_launch_shape): every such shape must give the hashes of the full-size launch."""
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"small-calls").digest()
    start = 0xfeed00000000
    big = progpow.hash_batch(ctx, 42, hd, start, 400000)  # 16 nonces per half-warp on any GPU up to 390 SMs
    assert big[:6 * 64] == oracle.kawpow_hash_range(0, 42, hd, start, 6)
    assert big[399990 * 64:] == oracle.kawpow_hash_range(0, 42, hd, start + 399990, 10)
    for count in (1, 2, 3, 31, 32, 33, 4735, 4736, 4737, 9472, 9473, 18945, 30000, 65536, 66305, 131072, 151552, 151553, 303105):
        assert progpow.hash_batch(ctx, 42, hd, start, count) == big[:count * 64], count
    # search over the same small windows: lowest qualifying nonce, as a host scan of the batch output finds it
    boundary = ((1 << (256 - 11)) - 1).to_bytes(32, "big")
    finals = [big[i * 64:i * 64 + 32] for i in range(70000)]
    hits = [i for i, f in enumerate(finals) if f <= boundary]
    assert len(hits) > 8
    for first, count in ((0, 65536), (hits[0] + 1, 9000), (hits[2] + 1, 70000 - hits[2] - 1), (hits[1], 1), (hits[1] + 1, hits[2] - hits[1] - 1)):
        want = next((h for h in hits if first <= h < first + count), None)
        sr = progpow.search(ctx, 42, hd, boundary, start + first, count)
        assert sr.solution_found == (want is not None), (first, count)
        if want is not None:
            assert sr.nonce == start + want and sr.final_hash == finals[want]


def test_search_stops_scanning_above_a_hit(ctx_of):
    """The reference returns at its first hit (progpow.cpp:572-577).  The kernel's warps look at the running minimum
    before each unit of work, so a hit near the start of a 2^26-nonce window (0.3 s of hashing) ends the call within
    milliseconds, with the lowest qualifying nonce."""
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"early-exit").digest()
    start = 77 << 40
    boundary = ((1 << (256 - 9)) - 1).to_bytes(32, "big")
    head = progpow.hash_batch(ctx, 600, hd, start, 8192)
    hits = [i for i in range(8192) if head[i * 64:i * 64 + 32] <= boundary]
    assert hits, "no hit among 8192 nonces at 9 bits: change the header"
    progpow.search(ctx, 600, hd, bytes(32), start, 1 << 20)  # warm: program loaded
    progpow.search(ctx, 600, hd, bytes(32), start, 1 << 24)
    full_ms, _ = ctx.last_timing()
    sr = progpow.search(ctx, 600, hd, boundary, start, 1 << 26)
    ms, launches = ctx.last_timing()
    assert sr.solution_found and sr.nonce == start + hits[0]
    assert ms < 5.0 and ms < full_ms, (ms, full_ms)
    # several queued launches (2^28 nonces = 4 launches): the ones behind the hit return at once
    sr = progpow.search(ctx, 600, hd, boundary, start, 1 << 28)
    ms, launches = ctx.last_timing()
    assert sr.solution_found and sr.nonce == start + hits[0] and ms < 5.0, (ms, launches)
    # a hit in the LAST unit of a window is still found, and it is the lowest one
    last = hits[-1]
    first = max(0, last - 4095)
    want = min(h for h in hits if h >= first)
    sr = progpow.search(ctx, 600, hd, boundary, start + first, last - first + 1)
    assert sr.solution_found and sr.nonce == start + want
    sr = progpow.search(ctx, 600, hd, boundary, start + last, 1)
    assert sr.solution_found and sr.nonce == start + last


def test_batch_kernel_writes_only_its_results(ctx_of):
    """Every launch shape of the compiled kernel (thin units for small counts, full units, a ragged last unit, warps whose
    unit lies past the end) stores exactly count x 64 bytes: guard zones on both sides of the device buffer stay intact."""
    import torch
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"guards").digest()
    guard = 8192
    for count in (1, 31, 33, 4737, 9473, 65536, 151553, 303105):
        buf = torch.full((guard + count * 64 + guard,), 0xA5, dtype=torch.uint8, device="cuda:0")
        progpow.hash_batch_device(ctx, 91, hd, 1 << 50, count, buf.data_ptr() + guard)
        got = buf.cpu().numpy().tobytes()
        assert got[:guard] == b"\xa5" * guard and got[guard + count * 64:] == b"\xa5" * guard, count
        if count <= 65536:
            assert got[guard:guard + count * 64] == progpow.hash_batch(ctx, 91, hd, 1 << 50, count), count
