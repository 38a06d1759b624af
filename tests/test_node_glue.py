"""Node-side glue of the KawPow path against fixtures made by the reference's
OWN block.cpp / hash.cpp / arith_uint256.cpp / uint256.cpp / chain.h (oracle/ref_node_shim.cpp -> tests/golden/
node_glue.json), and against that code live where oracle/_ref/libnxref_node.so exists.  Host code only, no GPU."""
import os
import random
import struct

import pytest
from conftest import golden

from nodexa_chain_core_b200 import node

G = golden("node_glue.json")
ACT = G["activation_time"]
LIMITS = {"main": bytes.fromhex("00" + "ff" * 31), "regtest": bytes.fromhex("7f" + "ff" * 31)}  # big-endian


def _hdr(d):
    return node.BlockHeader(d["version"], bytes.fromhex(d["prev_block"]), bytes.fromhex(d["merkle_root"]), d["time"], d["bits"],
                            d["height"], d["nonce64"], bytes.fromhex(d["mix_hash"]))


def test_header_hash_wire_header_and_block_id_vs_reference_fixtures(oracle):
    """block.cpp:94-99 GetKAWPOWHeaderHash, block.h:61-74 wire header, block.cpp:38-56 + hash.cpp:282-291 GetHash."""
    for d in G["headers"]:
        h = _hdr(d)
        assert h.kawpow_header_hash().hex() == d["header_hash_progpow"]
        assert h.kawpow_header_hash()[::-1].hex() == d["header_hash_uint256"]
        assert h.serialize().hex() == d["wire"]
        assert h.serialize()[:80].hex() == d["kawpow_input"]          # the first 80 wire bytes ARE the CKAWPOWInput
        assert node.BlockHeader.parse(bytes.fromhex(d["wire"])) == h
        assert h.block_hash().hex() == d["block_hash"]
        # the oracle's restatement agrees with the reference too
        assert oracle.kawpow_header_hash(d["version"], h.prev_block, h.merkle_root, d["time"], d["bits"], d["height"]).hex() \
            == d["header_hash_progpow"]
        assert oracle.kawpow_hash_no_verify(d["height"], h.kawpow_header_hash(), h.mix_hash[::-1], d["nonce64"])[::-1].hex() \
            == d["block_hash"]


def test_compact_target_vs_arith_uint256(oracle):
    """arith_uint256::SetCompact (arith_uint256.cpp:209-228) and the range checks of pow.cpp:190-191."""
    for c in G["compact"]:
        ok, target = node.compact_to_boundary(c["bits"])
        usable = not c["negative"] and not c["overflow"] and int(c["value_be"], 16) != 0
        assert ok == usable, hex(c["bits"])
        if usable:
            assert target.hex() == c["value_be"], hex(c["bits"])
        o_ok, o_target = oracle.compact_to_boundary(c["bits"])
        assert o_ok == usable and (not usable or o_target.hex() == c["value_be"])


def test_check_proof_of_work_vs_reference_arithmetic():
    """pow.cpp:182-199 on 140 (hash, nBits, powLimit) triples incl. hash == target, target +- 1 and unusable nBits."""
    for c in G["check_pow"]:
        fin = bytes.fromhex(c["hash_uint256"])[::-1]  # the product takes the hash as big-endian bytes (ProgPoW order)
        assert node.check_proof_of_work(fin, c["bits"], LIMITS[c["limit"]]) == c["ok"], c


def test_varint_vs_serialize_h():
    """serialize.h:290-366; the first eleven entries are the table in that file's own comment (:309-314)."""
    for v in G["varint"]:
        n, enc = int(v["n"]), bytes.fromhex(v["bytes"])
        assert node.varint_write(n) == enc
        assert node.varint_read(enc) == (n, len(enc))
        assert node.varint_read(enc + b"\x55\xaa") == (n, len(enc))   # stops at the number's last byte
    table = {0: "00", 1: "01", 127: "7f", 128: "8000", 255: "807f", 256: "8100", 16383: "fe7f", 16384: "ff00", 16511: "ff7f",
             65535: "82fe7f", 2**32: "8efefeff00"}
    for n, enc in table.items():
        assert node.varint_write(n).hex() == enc
    for v in G["varint_bad"]:
        if v["throws"]:
            with pytest.raises(ValueError):
                node.varint_read(bytes.fromhex(v["bytes"]))
        else:
            assert node.varint_read(bytes.fromhex(v["bytes"]))[0] == int(v["value"])


def test_disk_block_index_records_vs_chain_h():
    """CDiskBlockIndex::SerializationOp (chain.h:398-431): every status combination, both eras; then the start-up check
    of LoadBlockIndexGuts (txdb.cpp:470-492): GetBlockHash() + CheckProofOfWork."""
    for d in G["disk_index"]:
        raw = bytes.fromhex(d["raw"])
        rec, used = node.DiskBlockIndex.parse(raw, ACT)
        assert used == len(raw) and rec.client_version == d["client_version"] == G["client_version"]
        want = d["record"]
        for k in ("height", "status", "tx_count", "file", "data_pos", "undo_pos", "version", "time", "bits", "nonce", "nonce64"):
            assert getattr(rec, k) == want[k], k
        for k in ("prev_block", "merkle_root", "mix_hash"):
            assert getattr(rec, k).hex() == want[k], k
        assert rec.serialize(ACT) == raw
        assert node.DiskBlockIndex.parse(raw + b"\x00" * 7, ACT)[1] == len(raw)  # records are self-delimiting
        verdict, block_id = rec.check(ACT, LIMITS["main"])
        if d["block_hash"] is None:
            assert verdict == -1                                   # X16R-era block: not this library's
        else:
            assert block_id.hex() == d["block_hash"] and verdict == int(d["pow_ok_main"])
    for d in G["disk_index_bad"]:
        raw = bytes.fromhex(d["raw"])
        if d["throws"]:
            with pytest.raises(ValueError):
                node.DiskBlockIndex.parse(raw, ACT)
        else:
            rec, used = node.DiskBlockIndex.parse(raw, ACT)
            assert (used, rec.height) == (d["consumed"], d["height"])
    with pytest.raises(ValueError):
        node.DiskBlockIndex(1, -1, 0, 0).serialize(ACT)


def test_uint256_hex_order():
    """uint256.cpp:24-60: GetHex prints the storage bytes reversed; the RPC layer's hex strings follow from that."""
    for h in G["hex"]:
        if "gethex" in h:
            assert bytes.fromhex(h["raw"])[::-1].hex() == h["gethex"]
        elif all(c in "0123456789abcdefABCDEFxX " for c in h["sethex"]) and len(h["sethex"].strip()) <= 66:
            assert node._uint256s(h["sethex"])[::-1].hex() == h["raw"]


def test_live_against_reference_code_where_built():
    """Random headers, targets and records through the reference's compiled code and the product, side by side."""
    from oracle.oracle import REF_NODE_SO, RefNode
    if not os.path.exists(REF_NODE_SO) and not os.path.isdir("/root/reference/src"):
        pytest.skip("oracle/_ref/libnxref_node.so not built here")
    rn = RefNode(ACT)
    rnd = random.Random(7)
    for _ in range(200):
        t = ACT + rnd.randrange(10**8)
        f = rn.fields(rnd.randrange(-2**31, 2**31), rnd.randbytes(32), rnd.randbytes(32), t, rnd.getrandbits(32), rnd.getrandbits(31),
                      rnd.getrandbits(64), rnd.randbytes(32))
        h = node.BlockHeader(f.version, bytes(f.prev_block), bytes(f.merkle_root), f.time, f.bits, f.height, f.nonce64, bytes(f.mix_hash))
        assert h.kawpow_header_hash() == rn.header_hash(f)[1]
        assert h.serialize() == rn.serialize_header(f)
        assert h.block_hash() == rn.block_hash(f)
        st = rnd.choice([0, 5, 8 | 5, 16 | 5, 24 | 5, 56 | 5])
        raw = rn.diskindex_serialize(f.height, st, rnd.getrandbits(20), rnd.randrange(5000), rnd.getrandbits(32), rnd.getrandbits(32), f)
        rec, used = node.DiskBlockIndex.parse(raw, ACT)
        assert used == len(raw) and rec.serialize(ACT) == raw
        verdict, block_id = rec.check(ACT, LIMITS["main"])
        _, _, bh, _ = rn.diskindex_parse(raw)
        assert block_id == bh and verdict == int(rn.check_pow(bh, f.bits, LIMITS["main"][::-1]))
        bits = rnd.getrandbits(32) if rnd.random() < 0.5 else (rnd.randrange(1, 0x23) << 24) | rnd.getrandbits(23)
        v, neg, ovf = rn.set_compact(bits)
        ok, target = node.compact_to_boundary(bits)
        assert ok == (not neg and not ovf and any(v)) and (not ok or target == v)
        hv = rnd.randbytes(32)
        assert node.check_proof_of_work(hv, bits, LIMITS["main"]) == rn.check_pow(hv[::-1], bits, LIMITS["main"][::-1])
        n = rnd.getrandbits(rnd.randrange(1, 65))
        assert node.varint_write(n) == rn.write_varint(n)
        junk = rnd.randbytes(rnd.randrange(0, 12))
        got = rn.read_varint(junk)
        if got[1] < 0:
            with pytest.raises(ValueError):
                node.varint_read(junk)
        else:
            assert node.varint_read(junk) == got
        # truncated / bit-flipped records: both sides accept or both refuse, and agree on what they read
        cut = bytearray(raw[:rnd.randrange(len(raw) + 1)])
        if cut and rnd.random() < 0.5:
            cut[rnd.randrange(min(len(cut), 12))] ^= 1 << rnd.randrange(8)
        theirs = rn.diskindex_parse(bytes(cut))
        if theirs is None:
            with pytest.raises(ValueError):
                node.DiskBlockIndex.parse(bytes(cut), ACT)
        else:
            mine, used = node.DiskBlockIndex.parse(bytes(cut), ACT)
            assert used == theirs[3] and (mine.height, mine.status, mine.tx_count, mine.file, mine.data_pos, mine.undo_pos,
                                          mine.nonce64) == tuple(theirs[0][k] for k in
                                                                 ("height", "status", "tx_count", "file", "data_pos", "undo_pos", "nonce64"))


@pytest.mark.gpu
def test_kawpow_hash_of_whole_headers_vs_reference_KAWPOWHash(nx, ctx_of):
    """hash.cpp:258-279 end to end: header fields in, (final, mix) as the node stores them (uint256) out - the reference
    ran its own KAWPOWHash on these four headers (epochs 0 and 1); here the GPU does."""
    from nodexa_chain_core_b200 import progpow
    for d in G["kawpow_hash"]:
        h = _hdr(d)
        r = progpow.hash(ctx_of(h.height // 7500), h.height, h.kawpow_header_hash(), h.nonce64)
        assert r.final_hash[::-1].hex() == d["final_uint256"] and r.mix_hash[::-1].hex() == d["mix_uint256"]
        h.mix_hash = r.mix_hash[::-1]
        assert h.block_hash().hex() == d["final_uint256"]  # GetHash() of a block carrying the right mix == GetHashFull()


def test_pre_activation_wire_header_layout():
    """block.h:61-74: before nKAWPOWActivationTime a header is the classic 80 bytes ending in the 32-bit nNonce (X16R era -
    hashing those is out of scope, but the block-index reader must know the layout to skip over such records)."""
    for d in G["pre_activation_headers"]:
        raw = struct.pack("<i32s32sIII", d["version"], bytes.fromhex(d["prev_block"]), bytes.fromhex(d["merkle_root"]), d["time"],
                          d["bits"], d["nonce"])
        assert raw.hex() == d["wire"] and len(raw) == 80 and d["time"] < ACT
        rec = node.DiskBlockIndex(G["client_version"], 1, 0, 0, version=d["version"], prev_block=raw[4:36], merkle_root=raw[36:68],
                                  time=d["time"], bits=d["bits"], nonce=d["nonce"])
        blob = rec.serialize(ACT)
        assert blob.endswith(raw) and node.DiskBlockIndex.parse(blob, ACT)[0] == rec and rec.check(ACT)[0] == -1


def _dropin(explicit=False):
    from oracle.oracle import REF_NODE_DROPIN_SO, REF_NODE_EXPLICIT_SO, RefNode
    so = REF_NODE_EXPLICIT_SO if explicit else REF_NODE_DROPIN_SO
    if not os.path.exists(so) and not os.path.isdir("/root/reference/src"):
        pytest.skip(f"{os.path.basename(so)} not built here")
    return RefNode(ACT, dropin=not explicit, explicit=explicit)


@pytest.mark.parametrize("explicit", [False, True])
def test_reference_glue_over_the_product_library_host_part(explicit):
    """The reference's own hash.cpp / block.cpp, unmodified, compiled against include/ and linked with libnxkawpow.so
    (oracle/Makefile `ref_node_dropin`): GetKAWPOWHeaderHash and GetHash() (= KAWPOWHash_OnlyMix -> the product's
    progpow::hash_no_verify) give what the pure reference build gave.  No GPU needed for these two.
    explicit=True: the same with hash.cpp replaced by examples/node_patch/hash_kawpow_nx.cpp, the binding INTEGRATION.md
    section 1 proposes (`ref_node_explicit`)."""
    rn = _dropin(explicit)
    for d in G["headers"]:
        f = rn.fields(d["version"], bytes.fromhex(d["prev_block"]), bytes.fromhex(d["merkle_root"]), d["time"], d["bits"], d["height"],
                      d["nonce64"], bytes.fromhex(d["mix_hash"]))
        assert rn.header_hash(f)[1].hex() == d["header_hash_progpow"]
        assert rn.block_hash(f).hex() == d["block_hash"]


@pytest.mark.gpu
@pytest.mark.parametrize("explicit", [False, True])
def test_reference_KAWPOWHash_glue_over_the_gpu(gpu, explicit):
    """hash.cpp:258-279 as the node would run it after swapping the library: the reference's KAWPOWHash (static context,
    GetHex/to_hash256 round trips, uint256S of the result) calling this repository's ethash_create_epoch_context and
    progpow::hash on the GPU - and returning exactly what the all-reference build returned.  explicit=True: KAWPOWHash
    from examples/node_patch/hash_kawpow_nx.cpp (nx_* calls, mutex-guarded context) under the reference's block.cpp."""
    rn = _dropin(explicit)
    for d in G["kawpow_hash"]:
        f = rn.fields(d["version"], bytes.fromhex(d["prev_block"]), bytes.fromhex(d["merkle_root"]), d["time"], d["bits"], d["height"],
                      d["nonce64"], bytes.fromhex(d["mix_hash"]))
        fin, mix = rn.kawpow_hash(f)
        assert (fin.hex(), mix.hex()) == (d["final_uint256"], d["mix_uint256"])
        assert rn.block_hash_full(f) == (fin, mix)          # CBlockHeader::GetHashFull (block.cpp:58-76)


def test_python_restatement_of_the_record_format_and_product_agree():
    """oracle.oracle.varint_* / disk_block_index_* (plain Python, pinned to the reference-made fixtures here) against the
    product on random and damaged records - the checker that also exists where oracle/_ref does not."""
    from oracle import oracle as orc
    for v in G["varint"]:
        assert orc.varint_write(int(v["n"])).hex() == v["bytes"] and orc.varint_read(bytes.fromhex(v["bytes"])) == (int(v["n"]), len(v["bytes"]) // 2)
    for v in G["varint_bad"]:
        assert (orc.varint_read(bytes.fromhex(v["bytes"])) is None) == v["throws"]
    for d in G["disk_index"]:
        raw = bytes.fromhex(d["raw"])
        rec, used = orc.disk_block_index_parse(raw, ACT)
        assert used == len(raw) and orc.disk_block_index_serialize(rec, ACT) == raw
        for k, want in d["record"].items():
            assert (rec[k].hex() if isinstance(rec[k], bytes) else rec[k]) == want, k
    for d in G["disk_index_bad"]:
        got = orc.disk_block_index_parse(bytes.fromhex(d["raw"]), ACT)
        assert (got is None) == d["throws"] and (got is None or (got[1], got[0]["height"]) == (d["consumed"], d["height"]))
    rnd = random.Random(99)
    for _ in range(3000):
        act = rnd.choice([0, ACT, 2**32 - 1])
        rec = {"client_version": rnd.getrandbits(rnd.randrange(1, 32)), "height": rnd.getrandbits(rnd.randrange(1, 32)),
               "status": rnd.getrandbits(rnd.randrange(1, 33)), "tx_count": rnd.getrandbits(rnd.randrange(1, 33)),
               "file": rnd.getrandbits(20), "data_pos": rnd.getrandbits(32), "undo_pos": rnd.getrandbits(32),
               "version": rnd.randrange(-2**31, 2**31), "prev_block": rnd.randbytes(32), "merkle_root": rnd.randbytes(32),
               "time": rnd.getrandbits(32), "bits": rnd.getrandbits(32), "nonce": rnd.getrandbits(32), "nonce64": rnd.getrandbits(64),
               "mix_hash": rnd.randbytes(32)}
        raw = orc.disk_block_index_serialize(rec, act)
        mine = node.DiskBlockIndex(rec["client_version"], rec["height"], rec["status"], rec["tx_count"], rec["file"], rec["data_pos"],
                                   rec["undo_pos"], rec["version"], rec["prev_block"], rec["merkle_root"], rec["time"], rec["bits"],
                                   rec["nonce"], rec["nonce64"], rec["mix_hash"])
        assert mine.serialize(act) == raw
        damaged = bytearray(raw[:rnd.randrange(len(raw) + 1)] if rnd.random() < 0.5 else raw)
        if damaged and rnd.random() < 0.7:
            damaged[rnd.randrange(min(len(damaged), 14))] ^= 1 << rnd.randrange(8)
        theirs = orc.disk_block_index_parse(bytes(damaged), act)
        if theirs is None:
            with pytest.raises(ValueError):
                node.DiskBlockIndex.parse(bytes(damaged), act)
        else:
            got, used = node.DiskBlockIndex.parse(bytes(damaged), act)
            assert used == theirs[1]
            t = theirs[0]
            assert (got.client_version, got.height, got.status, got.tx_count, got.file, got.data_pos, got.undo_pos, got.version,
                    got.time, got.bits, got.nonce, got.nonce64, got.prev_block, got.merkle_root, got.mix_hash) == \
                   (t["client_version"], t["height"], t["status"], t["tx_count"], t["file"], t["data_pos"], t["undo_pos"], t["version"],
                    t["time"], t["bits"], t["nonce"], t["nonce64"], t["prev_block"], t["merkle_root"], t["mix_hash"])
