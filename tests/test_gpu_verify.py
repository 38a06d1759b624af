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
