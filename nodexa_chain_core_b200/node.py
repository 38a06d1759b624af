"""Node-side glue for KawPow blocks: the 120-byte wire header, the header hash fed to ProgPoW, block ids and the
PoW part of CheckBlockHeader (reference src/primitives/block.{h,cpp}, src/hash.cpp:258-291, src/pow.cpp:182-199,
src/validation.cpp:11638-11665) over the nx_* C ABI."""
from __future__ import annotations

import ctypes as C
import struct
from dataclasses import dataclass

from ._lib import check, inb, lib, outb
from .ethash import EpochContext

HEADER_OK, HEADER_HIGH_HASH, HEADER_INVALID_MIX, HEADER_BAD_BITS, HEADER_WRONG_EPOCH = range(5)
KAWPOW_LIMIT = bytes.fromhex("00" + "ff" * 31)  # chainparams.cpp:116-117 (main/test)
MAINNET_KAWPOW_ACTIVATION_TIME = 1651444217      # chainparams.cpp:264


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
