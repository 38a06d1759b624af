"""The oracle is pinned here: every KAT the reference's tests hold for the path
(SURVEY.md §8c), the fixtures generated from the reference's compiled code
(oracle/gen_golden.py), and — where oracle/_ref was built — the reference itself, live."""
import hashlib
import struct

import pytest
from conftest import golden

KATS = golden("kawpow_kats.json")


def test_keccak_and_permutations(oracle):
    m = golden("misc.json")
    d = bytes(range(256)) * 2
    for ln, (h256, h512) in m["keccak"].items():
        assert oracle.keccak256(d[:int(ln)]).hex() == h256
        assert oracle.keccak512(d[:int(ln)]).hex() == h512
    assert oracle.keccakf800(list(range(1, 26))) == m["keccakf800_of_1to25"]
    assert [str(x) for x in oracle.keccakf1600(list(range(1, 26)))] == m["keccakf1600_of_1to25"]


def test_epoch_geometry(oracle):
    e = golden("epochs.json")
    for k, (light, full) in e["sizes"].items():
        assert oracle.light_items(int(k)) == light
        assert oracle.dataset_items(int(k)) == full
    for k, s in e["seeds"].items():
        assert oracle.epoch_seed(int(k)).hex() == s
    for s, n in e["find_epoch"].items():
        if n <= 1000:
            assert oracle.find_epoch_number(bytes.fromhex(s)) == n
    # epoch 0 and the canonical mainnet-sized epoch 448
    assert (oracle.light_items(0), oracle.dataset_items(0)) == (262139, 8388593)
    assert (oracle.light_items(448), oracle.dataset_items(448)) == (1179641, 37748717)


def test_l1_cache_kat(oracle):
    """kawpow_tests.cpp:20-39"""
    assert oracle.l1(0)[:20] == KATS["l1_epoch0_first20"]
    e = golden("epochs.json")
    assert hashlib.sha256(struct.pack("<4096I", *oracle.l1(0))).hexdigest() == e["l1_sha256_epoch0"]
    assert hashlib.sha256(oracle.light_cache(0)).hexdigest() == e["light_cache_sha256_epoch0"]


def test_hash_kats_epoch0(oracle):
    """progpow_test_vectors.hpp:23-34 (the four epoch-0 cases), with the verify negatives of kawpow_tests.cpp:90-105"""
    for t in KATS["hash_cases"]:
        if t["height"] // 7500 != 0:
            continue
        hd = bytes.fromhex(t["header"])
        f, m = oracle.kawpow_hash(0, t["height"], hd, t["nonce"])
        assert m.hex() == t["mix"] and f.hex() == t["final"]
        assert oracle.kawpow_verify(0, t["height"], hd, m, t["nonce"], f)
        lower = bytearray(f)
        lower[31] = (lower[31] - 1) & 0xff
        assert not oracle.kawpow_verify(0, t["height"], hd, m, t["nonce"], bytes(lower))
        bad = bytearray(m)
        bad[7] = (bad[7] + 1) & 0xff
        assert not oracle.kawpow_verify(0, t["height"], hd, bytes(bad), t["nonce"], f)


@pytest.mark.parametrize("epoch", [3, 4, 7, 22])
def test_hash_kats_other_epochs(oracle, epoch):
    """progpow_test_vectors.hpp:35-61 + kawpow_tests.cpp:58-73 (height 30000, epoch 4)"""
    cases = [t for t in KATS["hash_cases"] if t["height"] // 7500 == epoch]
    if epoch == 4:
        cases.append(KATS["hash_30000"])
    assert cases
    for t in cases:
        f, m = oracle.kawpow_hash(epoch, t["height"], bytes.fromhex(t["header"]), t["nonce"])
        assert m.hex() == t["mix"] and f.hex() == t["final"]


def test_hash_empty_loop_last(oracle):
    """kawpow_tests.cpp:41-56 checks only the last iteration, i.e. height 1 (same period as height 0)."""
    f, m = oracle.kawpow_hash(0, 1, bytes(32), 0)
    assert m.hex() == KATS["hash_cases"][0]["mix"] and f.hex() == KATS["hash_cases"][0]["final"]


def test_search_kat(oracle):
    """kawpow_tests.cpp:109-140"""
    s = KATS["search"]
    hd, b = bytes.fromhex(s["header"]), bytes.fromhex(s["boundary"])
    found, nonce, f, m = oracle.kawpow_search(0, s["height"], hd, b, s["miss"]["start"], s["miss"]["iterations"])
    assert (found, nonce, f, m) == (False, 0, bytes(32), bytes(32))
    found, nonce, f, m = oracle.kawpow_search(0, s["height"], hd, b, s["hit"]["start"], s["hit"]["iterations"])
    assert found and nonce == s["hit"]["nonce"]
    assert (f, m) == oracle.kawpow_hash(0, 0, hd, 395)


def test_config1_prefix_and_chunks(oracle):
    """First 256 nonces of every config-1 run against the fixture made by the reference's code."""
    g = golden("kawpow_config1.json")
    hh = bytes.fromhex(g["header_hash"])
    raw = bytes.fromhex(g["header_input_80B"])
    assert oracle.kawpow_header_hash(*struct.unpack("<i32s32sIII", raw)) == hh
    for run in g["runs"]:
        buf = oracle.kawpow_hash_range(0, run["height"], hh, 0, 256)
        assert buf[:16 * 64].hex() == run["first16"]
        assert hashlib.sha256(buf).hexdigest() == run["sha256_per_256_nonces"][0]


def test_dag_items(oracle):
    g = golden("dag_items.json")["0"]
    for i, hx in g["items"].items():
        assert oracle.dataset_item_2048(0, int(i)).hex() == hx
    first = oracle.dataset_items_2048(0, 0, 4096)
    assert hashlib.sha256(first).hexdigest() == g["first4096_sha256"]


def test_epoch448(oracle):
    """The mainnet-sized epoch: light cache, a few DAG items and KawPow hashes vs the reference-made fixture."""
    e = golden("epochs.json")
    assert hashlib.sha256(oracle.light_cache(448)).hexdigest() == e["light_cache_sha256_epoch448"]
    g = golden("dag_items.json")["448"]
    for i, hx in list(g["items"].items())[:12]:
        assert oracle.dataset_item_2048(448, int(i)).hex() == hx
    k = golden("kawpow_epoch448.json")
    for c in k["cases"][:3]:
        buf = oracle.kawpow_hash_range(448, c["height"], bytes.fromhex(c["header"]), c["start"], 4)
        assert buf.hex() == c["out"][:4 * 128]


def test_unpinned_paths_vs_reference_fixture(oracle):
    m = golden("misc.json")
    for c in m["hash_no_verify"]:
        assert oracle.kawpow_hash_no_verify(0, bytes.fromhex(c["header"]), bytes.fromhex(c["mix"]),
                                            c["nonce"]).hex() == c["final"]
    for c in m["ethash_hash"][:2]:
        f, mx = oracle.ethash_hash(c["epoch"], bytes.fromhex(c["header"]), c["nonce"])
        assert (f.hex(), mx.hex()) == (c["final"], c["mix"])
        assert oracle.ethash_verify_final(bytes.fromhex(c["header"]), mx, c["nonce"], f)


def test_compact_and_compare(oracle):
    ok, b = oracle.compact_to_boundary(0x207fffff)  # regtest powLimit, chainparams.cpp:438
    assert ok and b.hex() == "7fffff" + "00" * 29
    ok, b = oracle.compact_to_boundary(0x1e00ffff)
    assert ok and b.hex() == "000000ffff" + "00" * 27
    ok, b = oracle.compact_to_boundary(0x2000ffff)  # kawpowLimit 00ff.. compact form
    assert ok and b.hex() == "00ffff" + "00" * 29
    assert not oracle.compact_to_boundary(0x01800000)[0]  # negative / zero mantissa cases
    assert not oracle.compact_to_boundary(0)[0]
    assert not oracle.compact_to_boundary(0x23000001)[0]  # overflow
    assert oracle.less_or_equal(bytes(32), bytes(32))
    assert oracle.less_or_equal(b"\x00" * 31 + b"\x01", b"\x01" + b"\x00" * 31)
    assert not oracle.less_or_equal(b"\x01" + b"\x00" * 31, b"\x00" + b"\xff" * 31)


def test_oracle_vs_reference_live(oracle, ref):
    """Where the reference's own objects were compiled (oracle/_ref), compare live on fresh inputs."""
    import random
    rnd = random.Random(5)
    for _ in range(6):
        hd = rnd.randbytes(32)
        n = rnd.getrandbits(64)
        h = rnd.randrange(0, 7500)
        assert oracle.kawpow_hash(0, h, hd, n) == ref.kawpow_hash(0, h, hd, n)
        mx = rnd.randbytes(32)
        assert oracle.kawpow_hash_no_verify(h, hd, mx, n) == ref.kawpow_hash_no_verify(h, hd, mx, n)
    for i in (0, 5, 4194295, 123456):
        assert oracle.dataset_item_2048(0, i) == ref.dataset_item_2048(0, i)
    assert oracle.light_cache(0) == ref.light_cache(0)
    assert oracle.ethash_hash(0, bytes(32), 7) == ref.ethash_hash(0, bytes(32), 7)
