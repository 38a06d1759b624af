"""Host-side product code (csrc/nx_host.cpp) against the golden fixtures and the oracle: epoch geometry,
Keccak, hash_no_verify, header hash, compact target.  No GPU needed."""
import ctypes as C
import random
import struct

from conftest import golden

from nodexa_chain_core_b200 import ethash, progpow
from nodexa_chain_core_b200._lib import inb, lib, outb


def test_epoch_geometry_vs_golden():
    e = golden("epochs.json")
    for k, (light, full) in e["sizes"].items():
        assert ethash.calculate_light_cache_num_items(int(k)) == light
        assert ethash.calculate_full_dataset_num_items(int(k)) == full
    for k, s in e["seeds"].items():
        assert ethash.calculate_epoch_seed(int(k)).hex() == s
    for s, n in e["find_epoch"].items():
        assert ethash.find_epoch_number(bytes.fromhex(s)) == n  # includes the not-found (-1) case
    assert ethash.get_epoch_number(0) == 0 and ethash.get_epoch_number(7499) == 0
    assert ethash.get_epoch_number(7500) == 1 and ethash.get_epoch_number(3360000) == 448


def test_primes_edge_cases(oracle):
    for ub in (-5, 0, 1, 2, 3, 4, 9, 10, 97, 100, 262144, 8388608):
        assert lib.nx_find_largest_prime(ub) == oracle.L.nxo_find_largest_prime(ub)
    assert lib.nx_find_largest_prime(1) == 0 and lib.nx_find_largest_prime(2) == 2


def test_keccak_vs_golden_and_oracle(oracle):
    m = golden("misc.json")
    d = bytes(range(256)) * 2
    for ln, (h256, h512) in m["keccak"].items():
        assert ethash.keccak256(d[:int(ln)]).hex() == h256
        assert ethash.keccak512(d[:int(ln)]).hex() == h512
    st = (C.c_uint32 * 25)(*range(1, 26))
    lib.nx_keccakf800(st)
    assert list(st) == m["keccakf800_of_1to25"]
    st64 = (C.c_uint64 * 25)(*range(1, 26))
    lib.nx_keccakf1600(st64)
    assert [str(x) for x in st64] == m["keccakf1600_of_1to25"]
    rnd = random.Random(1)
    for ln in (0, 5, 71, 72, 135, 136, 137, 1000):
        b = rnd.randbytes(ln)
        assert ethash.keccak256(b) == oracle.keccak256(b)
        assert ethash.keccak512(b) == oracle.keccak512(b)


def test_hash_no_verify_vs_golden_and_kats():
    for c in golden("misc.json")["hash_no_verify"]:
        assert progpow.hash_no_verify(0, bytes.fromhex(c["header"]), bytes.fromhex(c["mix"]),
                                      c["nonce"]).hex() == c["final"]
    # the final hash of every reference KAT is reproduced from its mix without the DAG
    for t in golden("kawpow_kats.json")["hash_cases"]:
        assert progpow.hash_no_verify(t["height"], bytes.fromhex(t["header"]), bytes.fromhex(t["mix"]),
                                      t["nonce"]).hex() == t["final"]


def test_header_hash_and_sha256d(oracle):
    g = golden("kawpow_config1.json")
    raw = bytes.fromhex(g["header_input_80B"])
    v, prev, merkle, t, bits, h = struct.unpack("<i32s32sIII", raw)
    o = outb(32)
    lib.nx_kawpow_header_hash(v, inb(prev), inb(merkle), t, bits, h, o)
    assert bytes(o).hex() == g["header_hash"]
    import hashlib
    rnd = random.Random(2)
    for ln in (0, 1, 55, 56, 63, 64, 80, 119, 120, 200):
        b = rnd.randbytes(ln)
        lib.nx_sha256d(inb(b) if b else outb(1), ln, o)
        assert bytes(o) == hashlib.sha256(hashlib.sha256(b).digest()).digest()


def test_compact_to_boundary_vs_oracle(oracle):
    rnd = random.Random(3)
    cases = [0, 0x207fffff, 0x2000ffff, 0x1e00ffff, 0x1d00ffff, 0x01003456, 0x01803456, 0x02008000, 0x03123456,
             0x04923456, 0x20123456, 0x21123456, 0x22123456, 0x23000001, 0xff123456, 0x1b0404cb]
    cases += [rnd.getrandbits(32) for _ in range(2000)]
    for bits in cases:
        o = outb(32)
        ok = lib.nx_compact_to_boundary(bits, o)
        ook, ob = oracle.compact_to_boundary(bits)
        assert (bool(ok), bytes(o)) == (ook, ob), hex(bits)


def test_check_proof_of_work():
    limit = bytes.fromhex("00" + "ff" * 31)  # kawpowLimit, chainparams.cpp:116-117
    o = outb(32)
    assert lib.nx_compact_to_boundary(0x1e00ffff, o)
    target = bytes(o)
    assert lib.nx_check_proof_of_work(inb(target), 0x1e00ffff, inb(limit)) == 1
    above = (int.from_bytes(target, "big") + 1).to_bytes(32, "big")
    assert lib.nx_check_proof_of_work(inb(above), 0x1e00ffff, inb(limit)) == 0
    assert lib.nx_check_proof_of_work(inb(bytes(32)), 0x207fffff, inb(limit)) == 0  # target above powLimit
    assert lib.nx_check_proof_of_work(inb(bytes(32)), 0, inb(limit)) == 0           # zero target


def test_host_light_cache_vs_reference_digest():
    """The product's own host light-cache chain (scalar and, where the CPU has it, AVX-512) for epoch 0 and the
    mainnet-sized epoch 448, against the digest of the reference's light cache."""
    import hashlib
    import time
    e = golden("epochs.json")
    for epoch in (0, 448):
        n = ethash.calculate_light_cache_num_items(epoch)
        for which, name in ((0, "dispatch"), (1, "scalar"), (2, "avx512")):
            if epoch == 448 and which == 0:
                continue  # same code as one of the other two
            buf = C.create_string_buffer(n * 64)
            t = time.time()
            rc = lib.nx_host_light_cache_variant(epoch, which, buf)
            dt = time.time() - t
            if rc == -1:
                assert which == 2  # CPU without AVX-512F
                continue
            assert hashlib.sha256(buf.raw).hexdigest() == e[f"light_cache_sha256_epoch{epoch}"], name
            assert buf.raw[:64].hex() == e[f"light_cache_first64_epoch{epoch}"]
            assert buf.raw[-64:].hex() == e[f"light_cache_last64_epoch{epoch}"]
            print(f"epoch {epoch}: host light cache ({name}) {dt:.2f} s")


def test_avx512_keccak_equals_scalar():
    rnd = random.Random(4)
    for _ in range(50):
        st = [rnd.getrandbits(64) for _ in range(25)]
        a, b = (C.c_uint64 * 25)(*st), (C.c_uint64 * 25)(*st)
        lib.nx_keccakf1600(a)
        if lib.nx_keccakf1600_avx512(b) == -1:
            return  # no AVX-512F on this CPU
        assert list(a) == list(b)
    z = (C.c_uint64 * 25)()
    lib.nx_keccakf1600_avx512(z)
    assert z[0] == 0xF1258F7940E1DDE7  # Keccak-f[1600] of the all-zero state, first lane


def test_wire_header_parse_serialize_and_block_hash(oracle):
    """120-byte KawPow header (primitives/block.h:61-74), CKAWPOWInput hash (block.cpp:94-99 + hash.cpp:269-270)
    and the block id (GetHash = KAWPOWHash_OnlyMix, hash.cpp:280-291), against the oracle's pieces."""
    from nodexa_chain_core_b200.node import BlockHeader
    from nodexa_chain_core_b200._lib import NxBlockHeader
    rnd = random.Random(12)
    for _ in range(20):
        h = BlockHeader(rnd.getrandbits(31), rnd.randbytes(32), rnd.randbytes(32), rnd.getrandbits(32), rnd.getrandbits(32),
                        rnd.getrandbits(31), rnd.getrandbits(64), rnd.randbytes(32))
        raw = h.serialize()
        assert len(raw) == 120 and BlockHeader.parse(raw) == h
        f = NxBlockHeader()
        lib.nx_block_header_parse(inb(raw), C.byref(f))
        assert (f.version, bytes(f.prev_block), bytes(f.merkle_root), f.time, f.bits, f.height, f.nonce64, bytes(f.mix_hash)) == \
               (h.version, h.prev_block, h.merkle_root, h.time, h.bits, h.height, h.nonce64, h.mix_hash)
        back = outb(120)
        lib.nx_block_header_serialize(C.byref(f), back)
        assert bytes(back) == raw
        hh = oracle.kawpow_header_hash(h.version, h.prev_block, h.merkle_root, h.time, h.bits, h.height)
        assert h.kawpow_header_hash() == hh
        fin = oracle.kawpow_hash_no_verify(h.height, hh, h.mix_hash[::-1], h.nonce64)
        assert h.block_hash() == fin[::-1]


def test_light_cache_file_cache(tmp_path, monkeypatch):
    """NX_CACHE_DIR: first call builds and writes, second call is served from disk, a corrupted or foreign file
    is detected and rebuilt - the bytes are always the reference's."""
    import hashlib
    e = golden("epochs.json")
    n = ethash.calculate_light_cache_num_items(0)
    want = e["light_cache_sha256_epoch0"]
    buf = C.create_string_buffer(n * 64)
    monkeypatch.delenv("NX_CACHE_DIR", raising=False)
    assert lib.nx_host_light_cache_cached(0, buf) == 0              # no directory configured: always built
    monkeypatch.setenv("NX_CACHE_DIR", str(tmp_path))
    assert lib.nx_host_light_cache_cached(0, buf) == 0              # first time: built + stored
    path = tmp_path / "light_cache_epoch0.nxlc"
    assert path.exists() and path.stat().st_size == n * 64 + 24
    buf2 = C.create_string_buffer(n * 64)
    assert lib.nx_host_light_cache_cached(0, buf2) == 1             # second time: from disk
    assert hashlib.sha256(buf2.raw).hexdigest() == want
    raw = bytearray(path.read_bytes())
    raw[24 + 12345] ^= 1                                            # flip one payload bit
    path.write_bytes(bytes(raw))
    buf3 = C.create_string_buffer(n * 64)
    assert lib.nx_host_light_cache_cached(0, buf3) == 0             # checksum mismatch: rebuilt (and rewritten)
    assert hashlib.sha256(buf3.raw).hexdigest() == want
    assert lib.nx_host_light_cache_cached(0, buf3) == 1
    path.write_bytes(path.read_bytes()[:1000])                      # truncated file
    assert lib.nx_host_light_cache_cached(0, buf3) == 0 and hashlib.sha256(buf3.raw).hexdigest() == want


def test_kawpow_launch_shape_covers_every_nonce(nx):
    """nx_kawpow_rt.cpp kawpow_launch_shape: a call is cut into units of 2 x per_hw nonces (one per warp at a time).  Whatever
    the count, the units must cover it exactly once, per_hw stays in 1..16, the persistent grid never exceeds the resident
    CTAs, and a 65 536-nonce call (the reference's nInnerLoopCount, src/rpc/mining.cpp:119) is spread over the whole GPU."""
    import ctypes as C
    import random
    f = nx.lib.nx_kawpow_launch_shape
    f.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    f.restype = None

    def shape(count, threads=1024, sms=148, minblocks=1, persist=1):
        k, u, g = C.c_uint32(), C.c_uint32(), C.c_uint32()
        f(count, threads, sms, minblocks, persist, C.byref(k), C.byref(u), C.byref(g))
        return k.value, u.value, g.value

    rnd = random.Random(5)
    counts = [1, 2, 31, 32, 33, 9472, 9473, 65536, 151552, 151553, 1 << 20, 1 << 24, 1 << 32] + [rnd.randrange(1, 1 << 27) for _ in range(200)]
    for threads, sms, mb in ((1024, 148, 1), (512, 148, 2), (256, 132, 4), (1024, 1, 1)):
        wpc = threads // 32
        for count in counts:
            k, units, grid = shape(count, threads, sms, mb)
            assert 1 <= k <= 16
            assert (units - 1) * 2 * k < count <= units * 2 * k          # covered, and no unit is wholly past the end
            assert 1 <= grid <= sms * mb and grid * wpc >= min(units, sms * mb * wpc - wpc + 1)
            if count >= 32 * sms * mb * wpc:
                assert k == 16                                            # large calls: the full 16-nonce walk
            _, u0, g0 = shape(count, threads, sms, mb, persist=0)
            assert g0 == -(-u0 // wpc)                                    # non-persistent: one unit per warp
    assert shape(65536) == (7, 4682, 147)
    assert shape(1 << 24) == (16, 1 << 19, 148)
