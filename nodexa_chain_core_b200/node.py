"""Node-side glue for KawPow blocks: the 120-byte wire header, the header hash fed to ProgPoW, block ids and the
PoW part of CheckBlockHeader (reference src/primitives/block.{h,cpp}, src/hash.cpp:258-291, src/pow.cpp:182-199,
src/validation.cpp:11638-11665) over the nx_* C ABI."""
from __future__ import annotations

import ctypes as C
import struct
from dataclasses import dataclass

from ._lib import NxDiskBlockIndex, check, inb, lib, outb
from .ethash import EpochContext

HEADER_OK, HEADER_HIGH_HASH, HEADER_INVALID_MIX, HEADER_BAD_BITS, HEADER_WRONG_EPOCH = range(5)
KAWPOW_LIMIT = bytes.fromhex("00" + "ff" * 31)  # chainparams.cpp:116-117 (main/test)
MAINNET_KAWPOW_ACTIVATION_TIME = 1651444217      # chainparams.cpp:264
TESTNET_KAWPOW_ACTIVATION_TIME = 1653247613      # chainparams.cpp:419
REGTEST_KAWPOW_ACTIVATION_TIME = 3582830167      # chainparams.cpp:569 (year 2083: regtest never reaches KawPow by itself)
REGTEST_KAWPOW_LIMIT = bytes.fromhex("7f" + "ff" * 31)  # chainparams.cpp:438-439


@dataclass
class BlockHeader:
    """Fields of CBlockHeader for a KawPow block; uint256 fields in storage order (as on the wire)."""
    version: int
    prev_block: bytes
    merkle_root: bytes
    time: int
    bits: int
    height: int
    nonce64: int = 0
    mix_hash: bytes = bytes(32)

    def serialize(self) -> bytes:
        return struct.pack("<i32s32sIIIQ32s", self.version, self.prev_block, self.merkle_root, self.time, self.bits,
                           self.height, self.nonce64, self.mix_hash)

    @staticmethod
    def parse(raw: bytes) -> "BlockHeader":
        return BlockHeader(*struct.unpack("<i32s32sIIIQ32s", raw))

    def kawpow_header_hash(self) -> bytes:
        """What the node passes to progpow::hash (hash.cpp:269-270)."""
        o = outb(32)
        lib.nx_kawpow_header_hash(self.version, inb(self.prev_block), inb(self.merkle_root), self.time, self.bits,
                                  self.height, o)
        return bytes(o)

    def block_hash(self) -> bytes:
        """CBlockHeader::GetHash() for KawPow blocks, uint256 storage order."""
        o = outb(32)
        lib.nx_block_hash(inb(self.serialize()), o)
        return bytes(o)


BLOCK_HAVE_DATA, BLOCK_HAVE_UNDO = 8, 16  # chain.h:156-157


def varint_write(n: int) -> bytes:
    """WriteVarInt (serialize.h:330-345)."""
    o = outb(10)
    return bytes(o[:lib.nx_varint_write(n, o)])


def varint_read(data: bytes):
    """ReadVarInt<uint64_t> (serialize.h:347-366): (value, bytes consumed); raises where the reference throws."""
    v = C.c_uint64()
    used = lib.nx_varint_read(inb(data) if data else outb(1), len(data), C.byref(v))
    if not used:
        raise ValueError("ReadVarInt(): size too large or truncated input")
    return v.value, used


@dataclass
class DiskBlockIndex:
    """One block-index record as LevelDB holds it (CDiskBlockIndex, chain.h:388-431)."""
    client_version: int
    height: int
    status: int
    tx_count: int
    file: int = 0
    data_pos: int = 0
    undo_pos: int = 0
    version: int = 0
    prev_block: bytes = bytes(32)
    merkle_root: bytes = bytes(32)
    time: int = 0
    bits: int = 0
    nonce: int = 0
    nonce64: int = 0
    mix_hash: bytes = bytes(32)

    def _c(self) -> NxDiskBlockIndex:
        return NxDiskBlockIndex(self.client_version, self.height, self.status, self.tx_count, self.file, self.data_pos,
                                self.undo_pos, self.version, (C.c_uint8 * 32)(*self.prev_block),
                                (C.c_uint8 * 32)(*self.merkle_root), self.time, self.bits, self.nonce, self.nonce64,
                                (C.c_uint8 * 32)(*self.mix_hash))

    def serialize(self, kawpow_activation_time: int = MAINNET_KAWPOW_ACTIVATION_TIME) -> bytes:
        o = outb(256)
        n = lib.nx_disk_block_index_serialize(C.byref(self._c()), kawpow_activation_time, o, len(o))
        if not n:
            raise ValueError("record cannot be serialized (negative height/file/version)")
        return bytes(o[:n])

    @staticmethod
    def parse(data: bytes, kawpow_activation_time: int = MAINNET_KAWPOW_ACTIVATION_TIME):
        """Returns (record, bytes consumed); raises ValueError for a malformed record."""
        r = NxDiskBlockIndex()
        used = lib.nx_disk_block_index_parse(inb(data) if data else outb(1), len(data), kawpow_activation_time, C.byref(r))
        if not used:
            raise ValueError("malformed CDiskBlockIndex record")
        return DiskBlockIndex(r.client_version, r.height, r.status, r.tx_count, r.file, r.data_pos, r.undo_pos, r.version,
                              bytes(r.prev_block), bytes(r.merkle_root), r.time, r.bits, r.nonce, r.nonce64,
                              bytes(r.mix_hash)), used

    def check(self, kawpow_activation_time: int = MAINNET_KAWPOW_ACTIVATION_TIME, pow_limit: bytes = KAWPOW_LIMIT):
        """LoadBlockIndexGuts's start-up check (txdb.cpp:470-492): (verdict, block id); verdict 1 ok, 0 PoW fails,
        -1 pre-KawPow block (X16R, not this library's)."""
        o = outb(32)
        v = lib.nx_disk_block_index_check(C.byref(self._c()), kawpow_activation_time, inb(pow_limit), o)
        return v, bytes(o)


def compact_to_boundary(bits: int):
    o = outb(32)
    ok = lib.nx_compact_to_boundary(bits, o)
    return bool(ok), bytes(o)


def check_proof_of_work(final_hash: bytes, bits: int, pow_limit: bytes = KAWPOW_LIMIT) -> bool:
    return bool(lib.nx_check_proof_of_work(inb(final_hash), bits, inb(pow_limit)))


def check_block_headers(context: EpochContext, headers, pow_limit: bytes = KAWPOW_LIMIT, mix_only: bool = False):
    """PoW part of CheckBlockHeader for many headers of one epoch; returns a list of HEADER_* codes."""
    raws = [h.serialize() if isinstance(h, BlockHeader) else bytes(h) for h in headers]
    blob = b"".join(raws)
    st = outb(max(1, len(raws)))
    check(lib.nx_check_block_headers(context._h, inb(blob) if blob else outb(1), len(raws), inb(pow_limit), int(mix_only), st))
    return [st[i] for i in range(len(raws))]


# ---- the external-miner RPC surface (src/rpc/mining.cpp:722-737 getblocktemplate fields, :763-840 getkawpowhash,
# :842-880 pprpcsb's proof-of-work gate).  Hex strings are in the order the node prints them: `GetHex()` of a uint256
# that was itself made by `uint256S(to_hex(hash256))`, i.e. plain hex of the ProgPoW byte order used by the C ABI. ----

def pprpc_fields(header: BlockHeader) -> dict:
    """`pprpcheader` / `pprpcepoch` of getblocktemplate (mining.cpp:728-736) for a block template."""
    return {"pprpcheader": header.kawpow_header_hash().hex(), "pprpcepoch": header.height // 7500}


def _parse_hex_nonce(text: str) -> int:
    """ParseUInt64(str, &n, 16): optional 0x prefix, must fit 64 bits (mining.cpp:784-786 'Invalid nonce hex string')."""
    t = text.strip()
    try:
        n = int(t, 16)
    except ValueError:
        raise ValueError("Invalid nonce hex string") from None
    if t.startswith("-") or n >> 64:
        raise ValueError("Invalid nonce hex string")
    return n


def _uint256s(text: str) -> bytes:
    """uint256S(hex) printed back with GetHex(): leading blanks and "0x" skipped, short strings zero-extended on the left."""
    t = text.strip()
    if t[:2] in ("0x", "0X"):
        t = t[2:]
    return bytes.fromhex(t.zfill(64)[-64:])


def getkawpowhash(context: EpochContext, header_hash: str, mix_hash: str, nonce: str, height: int, target: str | None = None) -> dict:
    """The getkawpowhash RPC (mining.cpp:763-840): one KawPow hash on the GPU, compared with the miner's claimed mix
    and, when a target is given, with that target.  Same keys and "true"/"false" strings as the node's reply."""
    from . import progpow
    r = progpow.hash(context, height, bytes.fromhex(header_hash), _parse_hex_nonce(nonce))
    ret = {"result": "true" if r.mix_hash == _uint256s(mix_hash) else "false",
           "digest": r.final_hash.hex(), "mix_hash": r.mix_hash.hex(), "info": ""}
    if target is not None:
        ret["meets_target"] = "true" if r.final_hash <= _uint256s(target) else "false"
    return ret


def pprpcsb_check(context: EpochContext, template: BlockHeader, mix_hash: str, nonce: str, pow_limit: bytes = KAWPOW_LIMIT):
    """The proof-of-work gate of pprpcsb (mining.cpp:862-880): the miner's nonce and mix go into the stored template,
    the full hash is recomputed (GetHashFull) and checked against nBits.  Returns (accepted, filled-in header); the
    mix itself is compared later, by CheckBlockHeader - see check_block_headers()."""
    from . import progpow
    filled = BlockHeader(template.version, template.prev_block, template.merkle_root, template.time, template.bits,
                         template.height, _parse_hex_nonce(nonce), _uint256s(mix_hash)[::-1])
    r = progpow.hash(context, filled.height, filled.kawpow_header_hash(), filled.nonce64)
    return check_proof_of_work(r.final_hash, filled.bits, pow_limit), filled
