// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// Second reference shim: the NODE-side glue around the KawPow path, driven through the UNMODIFIED reference sources
// compiled where they lie (oracle/Makefile, target `ref_node`, output only into oracle/_ref/):
//   primitives/block.cpp   CBlockHeader::GetKAWPOWHeaderHash / GetHash / GetHashFull        (:38-99)
//   primitives/block.h     CBlockHeader / CKAWPOWInput serialization (80- and 120-byte wire headers)  (:36-74,213-233)
//   hash.cpp               KAWPOWHash / KAWPOWHash_OnlyMix                                   (:258-291)
//   arith_uint256.cpp      arith_uint256::SetCompact / comparison (what CheckProofOfWork is made of)  (:209-228)
//   uint256.cpp            GetHex / SetHex byte order
//   chain.h + serialize.h  CDiskBlockIndex::SerializationOp, VARINT                          (chain.h:388-431, serialize.h:290-366)
// hash.cpp includes util.h (boost::filesystem, absent here) without using it on this path; the Makefile predefines that
// header's include guard so the file compiles untouched.  pow.cpp cannot be built (validation.h -> boost::variant), so
// nxrefnode_check_pow composes CheckProofOfWork (pow.cpp:182-199) from the reference's own SetCompact / UintToArith256.
// No reference source text is copied; this file only calls into it.
#include "arith_uint256.h"
#include "chain.h"
#include "clientversion.h"
#include "hash.h"
#include "primitives/block.h"
#include "primitives/transaction.h"
#include "streams.h"
#include "uint256.h"
#include "version.h"
#include "algo/sph_tiger.h"

#include <crypto/ethash/helpers.hpp>

#include <cstdlib>
#include <cstring>

// ---- link-only stubs for things block.cpp / block.h reference but this path never calls ----
std::string CTransaction::ToString() const { std::abort(); }            // CBlock::ToString (primitives/transaction.cpp)
void sph_tiger_init(void*) { std::abort(); }                             // X16RV2 only; the reference ships no tiger.c
void sph_tiger(void*, const void*, size_t) { std::abort(); }
void sph_tiger_close(void*, void*) { std::abort(); }

namespace
{
uint256 u256(const uint8_t* p)
{
    uint256 u;
    std::memcpy(u.begin(), p, 32);
    return u;
}

struct Fields  // mirrored by ctypes in oracle/oracle.py
{
    int32_t version;
    uint8_t prev_block[32];   // uint256 storage order
    uint8_t merkle_root[32];
    uint32_t time, bits, nonce, height;
    uint64_t nonce64;
    uint8_t mix_hash[32];
};

CBlockHeader header_of(const Fields& f)
{
    CBlockHeader h;
    h.nVersion = f.version;
    h.hashPrevBlock = u256(f.prev_block);
    h.hashMerkleRoot = u256(f.merkle_root);
    h.nTime = f.time;
    h.nBits = f.bits;
    h.nNonce = f.nonce;
    h.nHeight = f.height;
    h.nNonce64 = f.nonce64;
    h.mix_hash = u256(f.mix_hash);
    return h;
}

void fields_of(const CBlockHeader& h, Fields* f)
{
    f->version = h.nVersion;
    std::memcpy(f->prev_block, h.hashPrevBlock.begin(), 32);
    std::memcpy(f->merkle_root, h.hashMerkleRoot.begin(), 32);
    f->time = h.nTime;
    f->bits = h.nBits;
    f->nonce = h.nNonce;
    f->height = h.nHeight;
    f->nonce64 = h.nNonce64;
    std::memcpy(f->mix_hash, h.mix_hash.begin(), 32);
}

struct DiskIndex  // mirrored by ctypes
{
    int32_t height;
    uint32_t status, tx_count;
    int32_t file;
    uint32_t data_pos, undo_pos;
    Fields header;  // header.height is not part of the record (the VARINT above carries it)
};
}  // namespace

extern "C" {

void nxrefnode_set_activation_time(uint32_t t) { nKAWPOWActivationTime = t; }

// GetKAWPOWHeaderHash(): raw uint256 storage bytes, and what hash.cpp:269-270 feeds to progpow::hash.
void nxrefnode_header_hash(const Fields* f, uint8_t out_uint256[32], uint8_t out_progpow[32])
{
    const uint256 h = header_of(*f).GetKAWPOWHeaderHash();
    std::memcpy(out_uint256, h.begin(), 32);
    const ethash::hash256 p = to_hash256(h.GetHex());
    std::memcpy(out_progpow, p.bytes, 32);
}

// KAWPOWHash (hash.cpp:258-279): full hash + mix, both as uint256 storage bytes.  Builds a light context per epoch.
void nxrefnode_kawpow_hash(const Fields* f, uint8_t out_final[32], uint8_t out_mix[32])
{
    uint256 mix;
    const uint256 fin = KAWPOWHash(header_of(*f), mix);
    std::memcpy(out_final, fin.begin(), 32);
    std::memcpy(out_mix, mix.begin(), 32);
}

// CBlockHeader::GetHash() (block.cpp:38-56): for nTime >= activation this is KAWPOWHash_OnlyMix (hash.cpp:282-291).
// Only meaningful for KawPow-era headers here (the X16R branch needs the stubbed tiger).
void nxrefnode_block_hash(const Fields* f, uint8_t out[32])
{
    const uint256 h = header_of(*f).GetHash();
    std::memcpy(out, h.begin(), 32);
}

// CBlockHeader::GetHashFull (block.cpp:58-76)
void nxrefnode_block_hash_full(const Fields* f, uint8_t out[32], uint8_t out_mix[32])
{
    uint256 mix;
    const uint256 h = header_of(*f).GetHashFull(mix);
    std::memcpy(out, h.begin(), 32);
    std::memcpy(out_mix, mix.begin(), 32);
}

// wire header (block.h:61-74): 80 bytes before the activation time, 120 after
int nxrefnode_serialize_header(const Fields* f, uint8_t* out, int cap)
{
    CDataStream ss(SER_NETWORK, PROTOCOL_VERSION);
    ss << header_of(*f);
    if ((int)ss.size() > cap) return -1;
    std::memcpy(out, ss.data(), ss.size());
    return (int)ss.size();
}

int nxrefnode_parse_header(const uint8_t* in, int len, Fields* f)
{
    try {
        CDataStream ss(reinterpret_cast<const char*>(in), reinterpret_cast<const char*>(in) + len, SER_NETWORK, PROTOCOL_VERSION);
        CBlockHeader h;
        ss >> h;
        fields_of(h, f);
        return len - (int)ss.size();
    } catch (const std::exception&) {
        return -1;
    }
}

// the 80-byte CKAWPOWInput (block.h:213-233) whose SHA256d is the header hash
int nxrefnode_serialize_kawpow_input(const Fields* f, uint8_t out[80])
{
    const CBlockHeader h = header_of(*f);
    CKAWPOWInput input{h};
    CDataStream ss(SER_NETWORK, PROTOCOL_VERSION);
    ss << input;
    if (ss.size() != 80) return -1;
    std::memcpy(out, ss.data(), 80);
    return 80;
}

// arith_uint256::SetCompact (arith_uint256.cpp:209-228): value as 32 big-endian bytes + the two flags
void nxrefnode_set_compact(uint32_t bits, uint8_t out_be[32], int* negative, int* overflow)
{
    bool neg = false, ovf = false;
    arith_uint256 a;
    a.SetCompact(bits, &neg, &ovf);
    const uint256 u = ArithToUint256(a);
    for (int i = 0; i < 32; ++i) out_be[i] = u.begin()[31 - i];
    *negative = neg;
    *overflow = ovf;
}

uint32_t nxrefnode_get_compact(const uint8_t be[32])
{
    uint256 u;
    for (int i = 0; i < 32; ++i) u.begin()[31 - i] = be[i];
    return UintToArith256(u).GetCompact();
}

// CheckProofOfWork (pow.cpp:182-199) composed from the reference's arithmetic; hash and limit as uint256 storage bytes
int nxrefnode_check_pow(const uint8_t hash[32], uint32_t bits, const uint8_t pow_limit[32])
{
    bool neg = false, ovf = false;
    arith_uint256 target;
    target.SetCompact(bits, &neg, &ovf);
    if (neg || target == 0 || ovf || target > UintToArith256(u256(pow_limit))) return 0;
    if (UintToArith256(u256(hash)) > target) return 0;
    return 1;
}

// uint256 hex round trip (uint256.cpp:24-60): GetHex of raw bytes, SetHex into raw bytes
void nxrefnode_get_hex(const uint8_t raw[32], char out[65])
{
    const std::string s = u256(raw).GetHex();
    std::memcpy(out, s.c_str(), 65);
}
void nxrefnode_set_hex(const char* hex, uint8_t raw[32])
{
    const uint256 u = uint256S(hex);
    std::memcpy(raw, u.begin(), 32);
}

// VARINT (serialize.h:290-366)
int nxrefnode_write_varint(uint64_t n, uint8_t out[10])
{
    CDataStream ss(SER_DISK, CLIENT_VERSION);
    ss << VARINT(n);
    std::memcpy(out, ss.data(), ss.size());
    return (int)ss.size();
}
int nxrefnode_read_varint(const uint8_t* in, int len, uint64_t* n)
{
    try {
        CDataStream ss(reinterpret_cast<const char*>(in), reinterpret_cast<const char*>(in) + len, SER_DISK, CLIENT_VERSION);
        uint64_t v = 0;
        ss >> VARINT(v);
        *n = v;
        return len - (int)ss.size();
    } catch (const std::exception&) {
        return -1;
    }
}

// CDiskBlockIndex (chain.h:388-431) as LevelDB stores it (SER_DISK, leading VARINT client version)
int nxrefnode_diskindex_serialize(const DiskIndex* d, uint8_t* out, int cap)
{
    CDiskBlockIndex idx;
    idx.nHeight = d->height;
    idx.nStatus = d->status;
    idx.nTx = d->tx_count;
    idx.nFile = d->file;
    idx.nDataPos = d->data_pos;
    idx.nUndoPos = d->undo_pos;
    idx.nVersion = d->header.version;
    idx.hashPrev = u256(d->header.prev_block);
    idx.hashMerkleRoot = u256(d->header.merkle_root);
    idx.nTime = d->header.time;
    idx.nBits = d->header.bits;
    idx.nNonce = d->header.nonce;
    idx.nNonce64 = d->header.nonce64;
    idx.mix_hash = u256(d->header.mix_hash);
    CDataStream ss(SER_DISK, CLIENT_VERSION);
    ss << idx;
    if ((int)ss.size() > cap) return -1;
    std::memcpy(out, ss.data(), ss.size());
    return (int)ss.size();
}

int nxrefnode_diskindex_parse(const uint8_t* in, int len, DiskIndex* d, int32_t* client_version, uint8_t block_hash[32])
{
    try {
        CDataStream ss(reinterpret_cast<const char*>(in), reinterpret_cast<const char*>(in) + len, SER_DISK, CLIENT_VERSION);
        // the leading VARINT is the writer's client version; peek it with a second stream
        CDataStream peek(reinterpret_cast<const char*>(in), reinterpret_cast<const char*>(in) + len, SER_DISK, CLIENT_VERSION);
        int v = 0;
        peek >> VARINT(v);
        *client_version = v;
        CDiskBlockIndex idx;
        ss >> idx;
        d->height = idx.nHeight;
        d->status = idx.nStatus;
        d->tx_count = idx.nTx;
        d->file = idx.nFile;
        d->data_pos = idx.nDataPos;
        d->undo_pos = idx.nUndoPos;
        d->header.version = idx.nVersion;
        std::memcpy(d->header.prev_block, idx.hashPrev.begin(), 32);
        std::memcpy(d->header.merkle_root, idx.hashMerkleRoot.begin(), 32);
        d->header.time = idx.nTime;
        d->header.bits = idx.nBits;
        d->header.nonce = idx.nNonce;
        d->header.height = (uint32_t)idx.nHeight;
        d->header.nonce64 = idx.nNonce64;
        std::memcpy(d->header.mix_hash, idx.mix_hash.begin(), 32);
        if (block_hash && idx.nTime >= nKAWPOWActivationTime) {  // what LoadBlockIndexGuts keys the entry by (txdb.cpp:470-492)
            const uint256 h = idx.GetBlockHash();
            std::memcpy(block_hash, h.begin(), 32);
        }
        return len - (int)ss.size();
    } catch (const std::exception&) {
        return -1;
    }
}

int nxrefnode_client_version(void) { return CLIENT_VERSION; }

}  // extern "C"
