"""Kernel K6 (interpreted KawPow for independent records) and batched verification, through the C ABI.
Every record may sit in a different ProgPoW period; nothing is NVRTC-compiled on this path."""
import hashlib
import random

import pytest
from conftest import golden

from nodexa_chain_core_b200 import progpow

pytestmark = pytest.mark.gpu
KATS = golden("kawpow_kats.json")


def test_hash_many_kats_and_oracle(ctx_of, oracle):
    ctx = ctx_of(0)
    kats = [t for t in KATS["hash_cases"] if t["height"] // 7500 == 0]
    recs = [(t["height"], bytes.fromhex(t["header"]), t["nonce"]) for t in kats]
    rnd = random.Random(77)
    extra = [(rnd.randrange(7500), rnd.randbytes(32), rnd.getrandbits(64)) for _ in range(21)]  # ragged: 25 records
    got = progpow.hash_many(ctx, recs + extra)
    for t, r in zip(kats, got):
        assert (r.mix_hash.hex(), r.final_hash.hex()) == (t["mix"], t["final"])
    for (h, hd, n), r in zip(extra, got[len(kats):]):
        assert (r.final_hash, r.mix_hash) == oracle.kawpow_hash(0, h, hd, n)
    assert progpow.hash_many(ctx, []) == []


def test_interpreted_equals_compiled(ctx_of):
    """Same period, 700 nonces: K6 (interpreter) against K2 (NVRTC-compiled) bit for bit."""
    ctx = ctx_of(0)
    hd = hashlib.sha256(b"k6-vs-k2").digest()
    n = 700
    compiled = progpow.hash_batch(ctx, 4242, hd, 10**12, n)
    interp = progpow.hash_many(ctx, [(4242, hd, 10**12 + i) for i in range(n)])
    assert b"".join(r.final_hash + r.mix_hash for r in interp) == compiled


def test_hash_empty_full_reference_loop(ctx_of):
    """kawpow_tests.cpp:41-56 exactly as written: heights 1000 down to 1 (334 periods) with the epoch-0
    context, checking the last result.  Single hashes of uncompiled periods take the interpreted path."""
    ctx = ctx_of(0)
    r = None
    for h in range(1000, 0, -1):
        r = progpow.hash(ctx, h, bytes(32), 0)
    assert r.mix_hash.hex() == KATS["hash_cases"][0]["mix"]
    assert r.final_hash.hex() == KATS["hash_cases"][0]["final"]


def test_epoch448_records(ctx_of):
    k = golden("kawpow_epoch448.json")
    ctx = ctx_of(448)
    recs, want = [], []
    for c in k["cases"]:
        out = bytes.fromhex(c["out"])
        for i in range(0, c["count"], 5):
            recs.append((c["height"], bytes.fromhex(c["header"]), c["start"] + i))
            want.append(out[64 * i:64 * i + 64])
    got = progpow.hash_many(ctx, recs)
    assert [r.final_hash + r.mix_hash for r in got] == want


def test_verify_batch_semantics(nx, ctx_of):
    ctx = ctx_of(0)
    rnd = random.Random(3)
    base = [(rnd.randrange(7500), rnd.randbytes(32), rnd.getrandbits(64)) for _ in range(40)]
    res = progpow.hash_many(ctx, base)
    recs, expect = [], []
    for i, ((h, hd, n), r) in enumerate(zip(base, res)):
        kind = i % 4
        if kind == 0:    # valid, boundary == final (<= holds)
            recs.append((h, hd, r.mix_hash, n, r.final_hash)); expect.append(True)
        elif kind == 1:  # valid with a generous boundary
            recs.append((h, hd, r.mix_hash, n, b"\xff" * 32)); expect.append(True)
        elif kind == 2:  # boundary one below the final hash
            lower = (int.from_bytes(r.final_hash, "big") - 1).to_bytes(32, "big")
            recs.append((h, hd, r.mix_hash, n, lower)); expect.append(False)
        else:            # wrong mix that still passes the boundary test
            bad = bytearray(r.mix_hash); bad[7] = (bad[7] + 1) & 0xff
            recs.append((h, hd, bytes(bad), n, b"\xff" * 32)); expect.append(False)
    assert progpow.verify_batch(ctx, recs) == expect
    for rec, e in zip(recs[:8], expect[:8]):
        assert progpow.verify(ctx, *rec) == e
    assert progpow.verify_batch(ctx, []) == []
    with pytest.raises(nx.NxError):  # a block outside the context's epoch is refused, not mis-hashed
        progpow.hash_many(ctx, [(7500, bytes(32), 0)])


def test_check_block_headers_like_the_node(nx, ctx_of, oracle):
    """validation.cpp:11638-11665 on wire headers: mine a few synthetic blocks on the GPU, serialise them as the
    node would (120 bytes), then check them in one batch - untouched, with a wrong nonce, with a wrong mix, with a
    target above powLimit, from another epoch, and through the below-checkpoint (mix-only) branch."""
    from nodexa_chain_core_b200 import node
    ctx = ctx_of(0)
    rnd = random.Random(2026)
    bits = 0x2000ffff  # target 00ffff00..: about one nonce in 256 qualifies
    good = []
    with nx.Miner([0]) as m:
        m.set_batch(1 << 12)
        for i in range(6):
            h = node.BlockHeader(0x30000000, rnd.randbytes(32), rnd.randbytes(32), node.MAINNET_KAWPOW_ACTIVATION_TIME + 60 * i,
                                 bits, 1000 + 7 * i)
            sr = m.mine_header(h.version, h.prev_block, h.merkle_root, h.time, h.bits, h.height, start_nonce=i << 32,
                               max_tries=1 << 16)
            assert sr.solution_found
            h.nonce64, h.mix_hash = sr.nonce, sr.mix_hash[::-1]  # uint256S(to_hex(mix)): stored byte-reversed
            # the oracle agrees with every field the node would store and compare
            f, mx = oracle.kawpow_hash(0, h.height, h.kawpow_header_hash(), h.nonce64)
            assert (f, mx) == (sr.final_hash, sr.mix_hash)
            assert h.block_hash() == f[::-1]          # GetHash() of a valid block equals the full PoW hash
            assert node.check_proof_of_work(f, bits)
            good.append(h)
    assert node.check_block_headers(ctx, good) == [node.HEADER_OK] * len(good)
    assert node.check_block_headers(ctx, good, mix_only=True) == [node.HEADER_OK] * len(good)
    assert node.check_block_headers(ctx, []) == []

    import copy
    bad_nonce = copy.copy(good[0]); bad_nonce.nonce64 ^= 1
    bad_mix = copy.copy(good[1]); bad_mix.mix_hash = bytes([good[1].mix_hash[0] ^ 1]) + good[1].mix_hash[1:]
    bad_bits = copy.copy(good[2]); bad_bits.bits = 0x207fffff        # regtest-sized target > kawpowLimit
    zero_bits = copy.copy(good[2]); zero_bits.bits = 0
    other_epoch = copy.copy(good[3]); other_epoch.height = 7500
    got = node.check_block_headers(ctx, [bad_nonce, bad_mix, bad_bits, zero_bits, other_epoch, good[4]])
    f_bad, _ = oracle.kawpow_hash(0, bad_nonce.height, bad_nonce.kawpow_header_hash(), bad_nonce.nonce64)
    ok_t, target = oracle.compact_to_boundary(bits)
    expect_bad_nonce = node.HEADER_INVALID_MIX if f_bad <= target else node.HEADER_HIGH_HASH
    assert got == [expect_bad_nonce, node.HEADER_INVALID_MIX, node.HEADER_BAD_BITS, node.HEADER_BAD_BITS,
                   node.HEADER_WRONG_EPOCH, node.HEADER_OK]
    # mix-only branch: a wrong mix changes the block id, which then (almost surely) misses the target
    got = node.check_block_headers(ctx, [bad_mix, good[5]], mix_only=True)
    fin = oracle.kawpow_hash_no_verify(bad_mix.height, bad_mix.kawpow_header_hash(), bad_mix.mix_hash[::-1], bad_mix.nonce64)
    assert got == [node.HEADER_OK if fin <= target else node.HEADER_HIGH_HASH, node.HEADER_OK]


def test_external_miner_rpc_surface(nx, ctx_of, oracle):
    """rpc/mining.cpp:722-737 (pprpcheader/pprpcepoch), :763-840 (getkawpowhash), :842-880 (pprpcsb's PoW gate):
    a template goes out as header hash + epoch, an external miner answers with mix + nonce in hex, the node re-hashes."""
    from nodexa_chain_core_b200 import node, progpow
    k = golden("kawpow_kats.json")["hash_30000"]  # kawpow_tests.cpp:58-73, epoch 4
    ctx = ctx_of(4)
    r = node.getkawpowhash(ctx, k["header"], k["mix"], "%x" % k["nonce"], k["height"], target="0x" + k["final"])
    assert r == {"result": "true", "digest": k["final"], "mix_hash": k["mix"], "info": "", "meets_target": "true"}
    lower = "%064x" % (int(k["final"], 16) - 1)
    r = node.getkawpowhash(ctx, k["header"], "00" + k["mix"][2:], "0x%x" % k["nonce"], k["height"], target=lower)
    assert (r["result"], r["meets_target"], r["digest"]) == ("false", "false", k["final"])
    assert "meets_target" not in node.getkawpowhash(ctx, k["header"], k["mix"], "%x" % k["nonce"], k["height"])
    for bad in ("xyz", "1" * 17, "-1"):
        with pytest.raises(ValueError):
            node.getkawpowhash(ctx, k["header"], k["mix"], bad, k["height"])

    ctx0 = ctx_of(0)
    tmpl = node.BlockHeader(0x30000000, bytes(range(32)), bytes(range(32, 64)), node.MAINNET_KAWPOW_ACTIVATION_TIME + 5,
                            0x2000ffff, 4242)
    out = node.pprpc_fields(tmpl)
    assert out == {"pprpcheader": oracle.kawpow_header_hash(tmpl.version, tmpl.prev_block, tmpl.merkle_root, tmpl.time,
                                                             tmpl.bits, tmpl.height).hex(), "pprpcepoch": 0}
    ok_t, target = oracle.compact_to_boundary(tmpl.bits)
    sr = progpow.search(ctx0, tmpl.height, bytes.fromhex(out["pprpcheader"]), target, 0, 1 << 14)  # the external miner
    assert sr.solution_found
    accepted, filled = node.pprpcsb_check(ctx0, tmpl, sr.mix_hash.hex(), "%x" % sr.nonce)
    assert accepted and node.check_block_headers(ctx0, [filled]) == [node.HEADER_OK]
    # a nonce that does not solve the block is turned away ("Block does not solve the boundary")
    miss = next(n for n in range(64) if oracle.kawpow_hash(0, tmpl.height, bytes.fromhex(out["pprpcheader"]), n)[0] > target)
    accepted, _ = node.pprpcsb_check(ctx0, tmpl, sr.mix_hash.hex(), "%x" % miss)
    assert not accepted
    # the right nonce with somebody else's mix passes this gate (GetHashFull ignores the claim) and fails the next one
    accepted, forged = node.pprpcsb_check(ctx0, tmpl, "ab" * 32, "%x" % sr.nonce)
    assert accepted and node.check_block_headers(ctx0, [forged])[0] in (node.HEADER_INVALID_MIX, node.HEADER_HIGH_HASH)


@pytest.mark.parametrize("periods", [1, 2, 33])
def test_per_record_kernel_stays_inside_its_buffers(nx, ctx_of, periods):
    """compute-sanitizer is closed on this pool, so the per-record kernel (K6: one CTA = 8 records, dead slots of the
    last CTA re-read record n-1) is run on buffers of exactly the bytes it may touch, each between guard zones: no guard
    byte and no input byte may change, for record counts around the CTA size, and the hashes must equal the regular
    entry point's (nx_kawpow_hash_many, and through it nx_kawpow_verify_batch / nx_check_block_headers)."""
    import ctypes as C

    from nodexa_chain_core_b200._lib import NxHashRecord, check, outb
    ctx = ctx_of(0)
    rnd = random.Random(800 + periods)
    heights = [3 * rnd.randrange(2500) for _ in range(periods)]
    for n in (1, 7, 8, 9, 17, 64, 65):
        recs = [(heights[i % periods] + rnd.randrange(3), rnd.randbytes(32), rnd.getrandbits(64)) for i in range(n)]
        arr = (NxHashRecord * n)()
        for i, (bn, hd, nonce) in enumerate(recs):
            arr[i].block_number, arr[i].nonce = bn, nonce
            C.memmove(arr[i].header_hash, hd, 32)
        out = outb(64 * n)
        changed = C.c_uint64(12345)
        check(nx.lib.nx_test_hash_many_guarded(ctx._h, arr, n, out, C.byref(changed)))
        assert changed.value == 0, (n, periods, changed.value)
        want = progpow.hash_many(ctx, recs)
        assert bytes(out) == b"".join(r.final_hash + r.mix_hash for r in want), (n, periods)
        # the batch-verify entry built on the same kernel agrees record by record
        vrecs = [(bn, hd, w.mix_hash, nonce, w.final_hash) for (bn, hd, nonce), w in zip(recs, want)]
        assert progpow.verify_batch(ctx, vrecs) == [True] * n
    # one record checked against the KAT so that "equal to hash_many" is not a closed loop
    k = golden("kawpow_kats.json")["hash_cases"][0]
    r = progpow.hash_many(ctx, [(k["height"], bytes.fromhex(k["header"]), k["nonce"])])[0]
    assert r.final_hash.hex() == k["final"]
