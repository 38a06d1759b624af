#!/usr/bin/env python
"""TEST INFRASTRUCTURE: writes tests/golden/*.json from the REFERENCE's own code.

Runs only in the build container (needs /root/reference to compile oracle/_ref/libnxref.so and to
read the reference's KAT table).  The committed JSON files are what travels to the GPU box.

  python oracle/gen_golden.py            # everything except the whole-DAG digests (~3 min on 8 cores)
  python oracle/gen_golden.py --dag 0    # + whole-DAG digest for epoch 0   (~1 min)
  python oracle/gen_golden.py --dag 448  # + whole-DAG digest for epoch 448 (~5 min)
"""
import argparse
import hashlib
import json
import os
import re
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.oracle import Ref  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
REFSRC = "/root/reference/src"

MAINNET_KAWPOW_TIME = 1651444217  # src/chainparams.cpp:264


def synthetic_header_hash():
    """SHA256d of the 80-byte CKAWPOWInput, byte-reversed (src/hash.cpp:269-270)."""
    merkle = hashlib.sha256(b"nodexa-b200-golden").digest()
    raw = struct.pack("<i32s32sIII", 0x30000000, bytes(32), merkle, MAINNET_KAWPOW_TIME, 0x207fffff, 0)
    assert len(raw) == 80
    return hashlib.sha256(hashlib.sha256(raw).digest()).digest()[::-1], raw


def dag_weighted_sum(buf: bytes, first_word: int = 0) -> int:
    """sum_i w_i * (2*i+1) mod 2^64 over the u64 words of the DAG (the product's nx_dag_checksum)."""
    w = np.frombuffer(buf, dtype="<u8")
    i = np.arange(first_word, first_word + w.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        return int((w * (i * np.uint64(2) + np.uint64(1))).sum(dtype=np.uint64))


def write(name, obj):
    os.makedirs(GOLD, exist_ok=True)
    with open(os.path.join(GOLD, name), "w") as f:
        json.dump(obj, f, indent=1, sort_keys=True)
        f.write("\n")
    print("wrote", name)


def gen_kats(ref):
    """Every vector the reference's own tests hold for this path."""
    txt = open(os.path.join(REFSRC, "crypto/ethash/progpow_test_vectors.hpp")).read()
    cases = re.findall(r'\{(\d+),\s*"([0-9a-f]{64})",\s*"([0-9a-f]{16})",\s*"([0-9a-f]{64})",\s*"([0-9a-f]{64})"\}', txt)
    assert len(cases) == 13
    t = open(os.path.join(REFSRC, "test/kawpow_tests.cpp")).read()
    l1 = [int(x) for x in re.search(r"expected\{\s*\{([^}]*)\}\}", t).group(1).replace("\n", " ").split(",")]
    assert len(l1) == 20
    m30000 = re.search(r'to_hash256\("([0-9a-f]{64})"\);\s*const uint64_t nonce = (0x[0-9a-f]+);.*?mix_hex = "([0-9a-f]{64})";\s*'
                       r'const auto final_hex = "([0-9a-f]{64})"', t, re.S)
    obj = {
        "source": "src/crypto/ethash/progpow_test_vectors.hpp:22-62; src/test/kawpow_tests.cpp:29-33,52-53,60-69,115-139",
        "hash_cases": [{"height": int(h), "header": hd, "nonce": int(n, 16), "mix": mx, "final": fn} for h, hd, n, mx, fn in cases],
        "l1_epoch0_first20": l1,
        "hash_30000": {"height": 30000, "header": m30000.group(1), "nonce": int(m30000.group(2), 16),
                       "mix": m30000.group(3), "final": m30000.group(4)},
        "search": {"height": 0, "header": "00" * 32, "boundary": "00" + "ff" * 31,
                   "miss": {"start": 700, "iterations": 100}, "hit": {"start": 300, "iterations": 100, "nonce": 395}},
    }
    write("kawpow_kats.json", obj)


def chunks_sha(buf, per):
    return [hashlib.sha256(buf[i:i + per]).hexdigest() for i in range(0, len(buf), per)]


def gen_config1(ref):
    hh, raw = synthetic_header_hash()
    out = {"header_input_80B": raw.hex(), "header_hash": hh.hex(), "runs": []}
    for height, count in ((0, 65536), (1, 4096), (2, 4096), (3, 4096), (7499, 4096)):
        buf = ref.kawpow_hash_range(0, height, hh, 0, count)
        out["runs"].append({"height": height, "start": 0, "count": count, "layout": "count x {final32, mix32}",
                            "sha256_all": hashlib.sha256(buf).hexdigest(),
                            "sha256_per_256_nonces": chunks_sha(buf, 256 * 64),
                            "first16": buf[:16 * 64].hex()})
        print("config1 height", height, "done")
    # the all-zero header of the KATs, a short window that includes the search hit at 395
    buf = ref.kawpow_hash_range(0, 0, bytes(32), 0, 1024)
    out["zero_header_first1024_sha256"] = hashlib.sha256(buf).hexdigest()
    write("kawpow_config1.json", out)


def gen_epochs(ref):
    obj = {"sizes": {}, "seeds": {}, "find_epoch": {}}
    for e in (0, 1, 2, 3, 4, 7, 22, 100, 316, 384, 448, 512, 1000):
        obj["sizes"][str(e)] = [ref.light_items(e), ref.dataset_items(e)]
        s = ref.epoch_seed(e)
        obj["seeds"][str(e)] = s.hex()
        obj["find_epoch"][s.hex()] = ref.find_epoch_number(s)
    obj["find_epoch"]["ff" * 32] = ref.find_epoch_number(b"\xff" * 32)
    for e in (0, 448):
        lc = ref.light_cache(e)
        obj[f"light_cache_sha256_epoch{e}"] = hashlib.sha256(lc).hexdigest()
        obj[f"light_cache_first64_epoch{e}"] = lc[:64].hex()
        obj[f"light_cache_last64_epoch{e}"] = lc[-64:].hex()
        obj[f"l1_sha256_epoch{e}"] = hashlib.sha256(struct.pack("<4096I", *ref.l1(e))).hexdigest()
    write("epochs.json", obj)


def gen_dag_items(ref):
    rng = np.random.default_rng(20261106)
    obj = {}
    for e in (0, 448):
        m = ref.dataset_items(e) // 2
        idx = sorted(set([0, 1, 63, 64, m - 1, m // 2, 1150723 % m] + [int(x) for x in rng.integers(0, m, 57)]))
        obj[str(e)] = {"num_items_256": m,
                       "items": {str(i): ref.dataset_item_2048(e, i).hex() for i in idx}}
        # shard-boundary items for world sizes 2,4,8 (first/last item of each rank's slice)
        per = {}
        for n in (2, 4, 8):
            share = (m + n - 1) // n
            for g in range(n):
                for i in (g * share, min(m, (g + 1) * share) - 1):
                    per[str(i)] = hashlib.sha256(ref.dataset_item_2048(e, i)).hexdigest()
        obj[str(e)]["shard_edges_sha256"] = per
        first = ref.dataset_items_2048(e, 0, 4096)
        obj[str(e)]["first4096_sha256"] = hashlib.sha256(first).hexdigest()
        obj[str(e)]["first4096_weighted_sum"] = dag_weighted_sum(first)
    write("dag_items.json", obj)


def gen_epoch448(ref):
    """KawPow at the mainnet-sized epoch (config 2's parity spot check) + the paths the reference never pins."""
    height = 448 * 7500
    obj = {"epoch": 448, "height": height, "cases": []}
    for seed in range(1, 9):
        hh = ref.keccak256(struct.pack("<Q", seed))
        start = seed << 32
        buf = ref.kawpow_hash_range(448, height + seed, hh, start, 16)
        obj["cases"].append({"height": height + seed, "header": hh.hex(), "start": start, "count": 16, "out": buf.hex()})
    write("kawpow_epoch448.json", obj)

    misc = {"hash_no_verify": [], "ethash_hash": [], "keccak": {}}
    rng = np.random.default_rng(7)
    for _ in range(16):
        hd, mx = rng.bytes(32), rng.bytes(32)
        n = int(rng.integers(0, 2**63))
        misc["hash_no_verify"].append({"header": hd.hex(), "mix": mx.hex(), "nonce": n,
                                       "final": ref.kawpow_hash_no_verify(0, hd, mx, n).hex()})
    for e, n in ((0, 0), (0, 0x123456789abcdef0), (3, 77)):
        hd = rng.bytes(32)
        f, m = ref.ethash_hash(e, hd, n)
        misc["ethash_hash"].append({"epoch": e, "header": hd.hex(), "nonce": n, "final": f.hex(), "mix": m.hex()})
    for ln in (0, 1, 32, 64, 71, 72, 73, 135, 136, 137, 200, 300):
        d = bytes(range(256)) * 2
        misc["keccak"][str(ln)] = [ref.keccak256(d[:ln]).hex(), ref.keccak512(d[:ln]).hex()]
    st = list(range(1, 26))
    misc["keccakf800_of_1to25"] = ref.keccakf800(st)
    misc["keccakf1600_of_1to25"] = [str(x) for x in ref.keccakf1600(st)]
    write("misc.json", misc)


def gen_node(_ref):
    """Node-side glue from the reference's own block.cpp / hash.cpp / arith_uint256.cpp / chain.h (oracle/ref_node_shim.cpp)."""
    from oracle.oracle import RefNode
    rn = RefNode(MAINNET_KAWPOW_TIME)
    rng = np.random.default_rng(20261107)
    rb = lambda n: rng.bytes(n)  # noqa: E731
    out = {"activation_time": MAINNET_KAWPOW_TIME, "client_version": rn.client_version(), "headers": [], "kawpow_hash": [],
           "pre_activation_headers": [], "compact": [], "check_pow": [], "hex": [], "varint": [], "varint_bad": [],
           "disk_index": [], "disk_index_bad": []}

    def hdr_dict(f):
        return {"version": f.version, "prev_block": bytes(f.prev_block).hex(), "merkle_root": bytes(f.merkle_root).hex(),
                "time": f.time, "bits": f.bits, "nonce": f.nonce, "height": f.height, "nonce64": f.nonce64,
                "mix_hash": bytes(f.mix_hash).hex()}

    for i in range(12):
        f = rn.fields(int(rng.integers(-2**31, 2**31)) if i else 0x30000000, rb(32), rb(32),
                      MAINNET_KAWPOW_TIME + int(rng.integers(0, 10**8)), int(rng.integers(0, 2**32)), int(rng.integers(0, 2**31)),
                      int(rng.integers(0, 2**64, dtype=np.uint64)), rb(32))
        raw_hash, pp = rn.header_hash(f)
        d = hdr_dict(f)
        d.update(header_hash_uint256=raw_hash.hex(), header_hash_progpow=pp.hex(), kawpow_input=rn.kawpow_input(f).hex(),
                 wire=rn.serialize_header(f).hex(), block_hash=rn.block_hash(f).hex())
        pf, used = rn.parse_header(bytes.fromhex(d["wire"]))
        assert used == 120 and hdr_dict(pf) == {**hdr_dict(f), "nonce": 0}
        out["headers"].append(d)
    for height in (0, 1234, 7499, 7500):  # KAWPOWHash builds a light context per epoch (two epochs here)
        f = rn.fields(0x30000000, rb(32), rb(32), MAINNET_KAWPOW_TIME + 60 * height, 0x1e00ffff, height,
                      int(rng.integers(0, 2**64, dtype=np.uint64)), bytes(32))
        fin, mix = rn.kawpow_hash(f)
        fin2, mix2 = rn.block_hash_full(f)
        assert (fin, mix) == (fin2, mix2)
        d = hdr_dict(f)
        d.update(final_uint256=fin.hex(), mix_uint256=mix.hex())
        # with the right mix in the header, GetHash() (mix-only) equals the full hash
        g = rn.fields(f.version, bytes(f.prev_block), bytes(f.merkle_root), f.time, f.bits, f.height, f.nonce64, mix)
        assert rn.block_hash(g) == fin
        out["kawpow_hash"].append(d)
    for i in range(3):  # 80-byte wire form of the X16R era: only the serialization is on this path
        f = rn.fields(0x20000000, rb(32), rb(32), MAINNET_KAWPOW_TIME - 1 - i * 1000, 0x1c0fffff, 0, 0, bytes(32),
                      nonce=int(rng.integers(0, 2**32)))
        d = hdr_dict(f)
        d["wire"] = rn.serialize_header(f).hex()
        assert len(d["wire"]) == 160
        out["pre_activation_headers"].append(d)
    edge_bits = [0, 0x00123456, 0x01003456, 0x01123456, 0x01803456, 0x01fedcba, 0x02000056, 0x02008000, 0x02800056, 0x03000000,
                 0x03123456, 0x03800000, 0x04000000, 0x04123456, 0x04800000, 0x04923456, 0x05009234, 0x1b0404cb, 0x1c0fffff,
                 0x1d00ffff, 0x1e00ffff, 0x1f00ffff, 0x2000ffff, 0x20123456, 0x207fffff, 0x20800000, 0x21000001, 0x2100ffff,
                 0x21010000, 0x22000001, 0x220000ff, 0x22000100, 0x23000001, 0xff123456, 0xffffffff]
    for bits in edge_bits + [int(x) for x in rng.integers(0, 2**32, 40)]:
        v, neg, ovf = rn.set_compact(bits)
        out["compact"].append({"bits": bits, "value_be": v.hex(), "negative": neg, "overflow": ovf, "get_compact": rn.get_compact(v)})
    limit_raw = bytes.fromhex("00" + "ff" * 31)[::-1]  # kawpowLimit (chainparams.cpp:116-117) as uint256 storage bytes
    regtest_raw = bytes.fromhex("7f" + "ff" * 31)[::-1]
    for bits in (0x1b0404cb, 0x1f00ffff, 0x2000ffff, 0x207fffff, 0x01803456, 0, 0x23000001):
        target, _, _ = rn.set_compact(bits)
        t = int.from_bytes(target, "big")
        cands = [0, 1, max(t - 1, 0), t, t + 1, 2**256 - 1] + [int.from_bytes(rb(32), "big") >> int(rng.integers(0, 64)) for _ in range(4)]
        for hv in cands:
            hv %= 2**256
            hraw = hv.to_bytes(32, "big")[::-1]
            for lim_name, lim in (("main", limit_raw), ("regtest", regtest_raw)):
                out["check_pow"].append({"hash_uint256": hraw.hex(), "bits": bits, "limit": lim_name, "ok": rn.check_pow(hraw, bits, lim)})
    for _ in range(4):
        raw = rb(32)
        out["hex"].append({"raw": raw.hex(), "gethex": rn.get_hex(raw)})
    for text in ("0x1234", "  abcdef", "00" * 31 + "7f", "ff" * 40, "12xyz"):
        out["hex"].append({"sethex": text, "raw": rn.set_hex(text).hex()})
    for n in [0, 1, 127, 128, 255, 256, 16383, 16384, 16511, 16512, 65535, 2113663, 2113664, 2**31 - 1, 2**31, 2**32 - 1, 2**32,
              2**63, 2**64 - 1] + [int(x) for x in rng.integers(0, 2**63, 12)]:
        enc = rn.write_varint(n)
        assert rn.read_varint(enc) == (n, len(enc))
        out["varint"].append({"n": str(n), "bytes": enc.hex()})
    for bad in ("", "80", "ffff", "ff" * 10 + "00", "80" * 11, "fefefefefefefefefeff7f"):
        v, used = rn.read_varint(bytes.fromhex(bad))
        out["varint_bad"].append({"bytes": bad, "throws": used < 0, "value": None if used < 0 else str(v)})
    statuses = [0, 1, 5, 8 | 3, 16 | 4, 8 | 16 | 5, 8 | 16 | 32 | 5, 0x80 | 8 | 5]
    for i, st in enumerate(statuses + statuses[:3]):
        pre = i >= len(statuses)
        t = MAINNET_KAWPOW_TIME - 100 - i if pre else MAINNET_KAWPOW_TIME + int(rng.integers(0, 10**8))
        f = rn.fields(0x30000000, rb(32), rb(32), t, int(rng.integers(0, 2**32)), 0,
                      0 if pre else int(rng.integers(0, 2**64, dtype=np.uint64)), bytes(32) if pre else rb(32),
                      nonce=int(rng.integers(0, 2**32)) if pre else 0)
        height, ntx = int(rng.integers(0, 2**31)), int(rng.integers(0, 2**20))
        file_, dpos, upos = int(rng.integers(0, 5000)), int(rng.integers(0, 2**32)), int(rng.integers(0, 2**32))
        raw = rn.diskindex_serialize(height, st, ntx, file_, dpos, upos, f)
        rec, cv, bh, used = rn.diskindex_parse(raw)
        assert used == len(raw) and cv == rn.client_version()
        rec = {k: (v.hex() if isinstance(v, bytes) else v) for k, v in rec.items()}
        out["disk_index"].append({"raw": raw.hex(), "record": rec, "client_version": cv,
                                  "block_hash": bh.hex() if bh is not None else None,
                                  "pow_ok_main": None if bh is None else rn.check_pow(bh, f.bits, limit_raw)})
    good = bytes.fromhex(out["disk_index"][5]["raw"])
    big_height = rn.write_varint(4040402) + rn.write_varint(2**31) + good[6:]  # nHeight is an int: 2^31 does not fit
    tail = struct.pack("<i32s32sIIQ32s", 1, rb(32), rb(32), MAINNET_KAWPOW_TIME + 1, 0x1b0404cb, 7, rb(32))

    def scratch(cv, h, st, ntx, extra=b""):
        return rn.write_varint(cv) + rn.write_varint(h) + rn.write_varint(st) + rn.write_varint(ntx) + extra + tail

    limits = [scratch(4040402, 2**31 - 1, 0, 0), scratch(4040402, 2**31, 0, 0), scratch(2**31, 5, 0, 0),
              scratch(1, 5, 2**32 - 1, 0, rn.write_varint(1) + rn.write_varint(2) + rn.write_varint(3)), scratch(1, 5, 2**32, 0),
              scratch(1, 5, 0, 2**32 - 1), scratch(1, 5, 0, 2**32), scratch(1, 5, 8, 0, rn.write_varint(2**31 - 1) + rn.write_varint(2**32 - 1)),
              scratch(1, 5, 8, 0, rn.write_varint(2**31) + rn.write_varint(0)), scratch(1, 5, 16, 0, rn.write_varint(0) + rn.write_varint(2**32))]
    for bad in [good[:-1], good[:10], b"", big_height, rn.write_varint(4040402) + b"\xff" * 12] + limits:
        got = rn.diskindex_parse(bad)
        out["disk_index_bad"].append({"raw": bad.hex(), "throws": got is None, "consumed": None if got is None else got[3],
                                      "height": None if got is None else got[0]["height"]})
    write("node_glue.json", out)


def gen_dag_digest(ref, epoch):
    m = ref.dataset_items(epoch) // 2
    h = hashlib.sha256()
    wsum, step = 0, 1 << 16
    for first in range(0, m, step):
        cnt = min(step, m - first)
        buf = ref.dataset_items_2048(epoch, first, cnt)
        h.update(buf)
        wsum = (wsum + dag_weighted_sum(buf, first * 32)) & (2**64 - 1)
        if (first // step) % 16 == 0:
            print(f"epoch {epoch}: {first}/{m}", flush=True)
    path = os.path.join(GOLD, "dag_digest.json")
    obj = json.load(open(path)) if os.path.exists(path) else {}
    obj[str(epoch)] = {"num_items_256": m, "bytes": m * 256, "sha256": h.hexdigest(), "weighted_sum": wsum,
                       "weighted_sum_def": "sum_i u64word_i * (2*i+1) mod 2^64"}
    write("dag_digest.json", obj)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--dag", type=int, action="append", default=[])
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    ref = Ref()
    if a.dag:
        for e in a.dag:
            gen_dag_digest(ref, e)
    else:
        steps = {"kats": gen_kats, "epochs": gen_epochs, "dag_items": gen_dag_items, "epoch448": gen_epoch448,
                 "config1": gen_config1, "node": gen_node}
        for k, fn in steps.items():
            if not a.only or k in a.only.split(","):
                fn(ref)
