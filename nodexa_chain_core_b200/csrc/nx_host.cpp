// Host-side pieces of the path: epoch geometry, Keccak sponges, the (inherently serial) light-cache
// chain, the node glue (SHA256d header hash, compact target).  Product code; shares nothing with oracle/.
#include "nx_internal.hpp"
#include "nx_keccak.cuh"

#include <unistd.h>

#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <string>

namespace nx
{
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
void clear_error() { g_err[0] = 0; }

// Original-Keccak sponge over arbitrary input (reference lib/keccak/keccak.c:46-101).
static void sponge(uint8_t* out, size_t outlen, const uint8_t* data, size_t size)
{
    const size_t rate = 200 - 2 * outlen;
    uint64_t st[25] = {0};
    uint8_t block[200];
    while (size >= rate) {
        uint64_t w[25];
        std::memcpy(w, data, rate);
        for (size_t i = 0; i < rate / 8; ++i) st[i] ^= w[i];
        keccakf1600(st);
        data += rate;
        size -= rate;
    }
    std::memset(block, 0, sizeof block);
    if (size) std::memcpy(block, data, size);
    block[size] ^= 0x01;
    block[rate - 1] ^= 0x80;
    uint64_t w[25];
    std::memcpy(w, block, rate);
    for (size_t i = 0; i < rate / 8; ++i) st[i] ^= w[i];
    keccakf1600(st);
    std::memcpy(out, st, outlen);
}

// reference lib/ethash/ethash.cpp:101-129 — 4*n dependent keccak512 calls; no parallelism to extract, so the
// only lever is the latency of one permutation: AVX-512 version where the CPU has it, scalar otherwise.
void host_build_light_cache(uint8_t* cache, int n, const uint8_t seed[32])
{
    if (cpu_has_avx512f())
        host_build_light_cache_avx512(cache, n, seed);
    else
        host_build_light_cache_scalar(cache, n, seed);
}

// ---- on-disk light cache (epoch lifecycle): the serial chain is the slowest part of an epoch
// build (1.3 s of 1.5 s), yet its result is 16-80 MiB that never changes for an epoch.  With NX_CACHE_DIR set it
// is written once and re-read on later starts, guarded by a header (magic, epoch, item count) and a
// position-weighted 64-bit checksum of the payload; anything that does not verify is rebuilt and rewritten. ----
namespace
{
struct LightFileHeader
{
    char magic[8];
    uint32_t epoch, items;
    uint64_t checksum;
};
uint64_t light_checksum(const uint8_t* p, size_t bytes)
{
    uint64_t acc = 0;
    const uint64_t* w = reinterpret_cast<const uint64_t*>(p);
    for (size_t i = 0; i < bytes / 8; ++i) acc += w[i] * (2 * (uint64_t)i + 1);
    return acc;
}
std::string light_path(int epoch)
{
    const char* dir = std::getenv("NX_CACHE_DIR");
    if (!dir || !*dir) return "";
    return std::string(dir) + "/light_cache_epoch" + std::to_string(epoch) + ".nxlc";
}
}  // namespace

bool host_load_light_cache(uint8_t* cache, int n, int epoch)
{
    const std::string path = light_path(epoch);
    if (path.empty()) return false;
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    LightFileHeader h;
    const size_t bytes = (size_t)n * 64;
    bool ok = std::fread(&h, sizeof h, 1, f) == 1 && std::memcmp(h.magic, "NXLC0001", 8) == 0 && h.epoch == (uint32_t)epoch &&
              h.items == (uint32_t)n && std::fread(cache, 1, bytes, f) == bytes;
    std::fclose(f);
    return ok && light_checksum(cache, bytes) == h.checksum;
}

void host_store_light_cache(const uint8_t* cache, int n, int epoch)
{
    const std::string path = light_path(epoch);
    if (path.empty()) return;
    const std::string tmp = path + ".tmp." + std::to_string((long)getpid());  // one process per GPU may build the same epoch
    FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) return;
    LightFileHeader h;
    std::memcpy(h.magic, "NXLC0001", 8);
    h.epoch = (uint32_t)epoch;
    h.items = (uint32_t)n;
    h.checksum = light_checksum(cache, (size_t)n * 64);
    const bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(cache, 1, (size_t)n * 64, f) == (size_t)n * 64;
    std::fclose(f);
    if (ok)
        std::rename(tmp.c_str(), path.c_str());  // readers never see a half-written file
    else
        std::remove(tmp.c_str());
}

void host_build_light_cache_scalar(uint8_t* cache, int n, const uint8_t seed[32])
{
    uint64_t* c = reinterpret_cast<uint64_t*>(cache);
    uint64_t in[4], item[8];
    std::memcpy(in, seed, 32);
    keccak512_32(in, item);
    std::memcpy(c, item, 64);
    for (int i = 1; i < n; ++i) {
        keccak512_64(item);
        std::memcpy(c + 8 * (size_t)i, item, 64);
    }
    const uint32_t un = (uint32_t)n;
    for (int q = 0; q < 3; ++q) {
        for (int i = 0; i < n; ++i) {
            const uint32_t v = (uint32_t)c[8 * (size_t)i] % un;
            const uint32_t w = (uint32_t)(n + (i - 1)) % un;
            const uint64_t *a = c + 8 * (size_t)v, *b = c + 8 * (size_t)w;
            for (int k = 0; k < 8; ++k) item[k] = a[k] ^ b[k];
            keccak512_64(item);
            std::memcpy(c + 8 * (size_t)i, item, 64);
        }
    }
}

// ---- SHA-256 (FIPS 180-4) for the header hash; the node's own is src/crypto/sha256.cpp ----
namespace
{
const uint32_t K256[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98,
    0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786,
    0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8,
    0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
    0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819,
    0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a,
    0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7,
    0xc67178f2};
inline uint32_t rr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

void sha256_block(uint32_t h[8], const uint8_t* p)
{
    uint32_t w[64];
    for (int i = 0; i < 16; ++i)
        w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    for (int i = 16; i < 64; ++i)
        w[i] = w[i - 16] + (rr(w[i - 15], 7) ^ rr(w[i - 15], 18) ^ (w[i - 15] >> 3)) + w[i - 7] +
               (rr(w[i - 2], 17) ^ rr(w[i - 2], 19) ^ (w[i - 2] >> 10));
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 64; ++i) {
        const uint32_t t1 = hh + (rr(e, 6) ^ rr(e, 11) ^ rr(e, 25)) + ((e & f) ^ (~e & g)) + K256[i] + w[i];
        const uint32_t t2 = (rr(a, 2) ^ rr(a, 13) ^ rr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
        hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}

void sha256(const uint8_t* data, size_t n, uint8_t out[32])
{
    uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    size_t off = 0;
    for (; off + 64 <= n; off += 64) sha256_block(h, data + off);
    uint8_t tail[128] = {0};
    const size_t rem = n - off;
    if (rem) std::memcpy(tail, data + off, rem);
    tail[rem] = 0x80;
    const size_t tl = rem + 9 <= 64 ? 64 : 128;
    const uint64_t bits = (uint64_t)n * 8;
    for (int i = 0; i < 8; ++i) tail[tl - 1 - i] = (uint8_t)(bits >> (8 * i));
    sha256_block(h, tail);
    if (tl == 128) sha256_block(h, tail + 64);
    for (int i = 0; i < 8; ++i) {
        out[4 * i] = (uint8_t)(h[i] >> 24);
        out[4 * i + 1] = (uint8_t)(h[i] >> 16);
        out[4 * i + 2] = (uint8_t)(h[i] >> 8);
        out[4 * i + 3] = (uint8_t)h[i];
    }
}
}  // namespace
}  // namespace nx

using namespace nx;

extern "C" {

const char* nx_last_error(void) { return g_err; }
const char* nx_version(void) { return "nodexa-b200-pow 0.1 (KawPow/ProgPoW 0.9.4, sm_100a)"; }

int nx_find_largest_prime(int upper_bound)
{
    if (upper_bound < 2) return 0;
    if (upper_bound == 2) return 2;
    int n = (upper_bound & 1) ? upper_bound : upper_bound - 1;
    for (;; n -= 2) {
        bool prime = true;
        for (int64_t d = 3; d * d <= (int64_t)n; d += 2)
            if (n % d == 0) {
                prime = false;
                break;
            }
        if (prime) return n;
    }
}

int nx_light_cache_num_items(int epoch) { return nx_find_largest_prime((1 << 24) / 64 + epoch * ((1 << 17) / 64)); }
int nx_full_dataset_num_items(int epoch) { return nx_find_largest_prime((1 << 30) / 128 + epoch * ((1 << 23) / 128)); }

void nx_epoch_seed(int epoch, uint8_t out[32])
{
    uint64_t s[4] = {0, 0, 0, 0};
    for (int i = 0; i < epoch; ++i) keccak256_32(s);
    std::memcpy(out, s, 32);
}

int nx_find_epoch_number(const uint8_t seed[32])
{
    // The reference matches on the first 32-bit word only and gives up after 30000 epochs.
    uint32_t want;
    std::memcpy(&want, seed, 4);
    uint64_t s[4] = {0, 0, 0, 0};
    for (int i = 0; i < 30000; ++i) {
        if ((uint32_t)s[0] == want) return i;
        keccak256_32(s);
    }
    return -1;
}

void nx_keccakf1600(uint64_t state[25]) { keccakf1600(state); }
void nx_keccakf800(uint32_t state[25]) { keccakf800(state); }
void nx_keccak256(const uint8_t* data, size_t size, uint8_t out[32]) { sponge(out, 32, data, size); }
void nx_keccak512(const uint8_t* data, size_t size, uint8_t out[64]) { sponge(out, 64, data, size); }

void nx_kawpow_hash_no_verify(int block_number, const uint8_t header_hash[32], const uint8_t mix_hash[32],
                              uint64_t nonce, uint8_t out[32])
{
    static const uint32_t pad[15] = {0x72, 0x41, 0x56, 0x45, 0x4E, 0x43, 0x4F, 0x49, 0x4E, 0x4B, 0x41, 0x57, 0x50, 0x4F, 0x57};
    (void)block_number;
    uint32_t st[25], carry[8];
    std::memcpy(st, header_hash, 32);
    st[8] = (uint32_t)nonce;
    st[9] = (uint32_t)(nonce >> 32);
    for (int i = 10; i < 25; ++i) st[i] = pad[i - 10];
    keccakf800(st);
    std::memcpy(carry, st, 32);
    std::memcpy(st, carry, 32);
    std::memcpy(st + 8, mix_hash, 32);
    for (int i = 16; i < 25; ++i) st[i] = pad[i - 16];
    keccakf800(st);
    std::memcpy(out, st, 32);
}

int nx_ethash_verify_final_hash(const uint8_t header_hash[32], const uint8_t mix_hash[32], uint64_t nonce,
                                const uint8_t boundary[32])
{
    uint8_t init[40], fin[96], out[32];
    std::memcpy(init, header_hash, 32);
    std::memcpy(init + 32, &nonce, 8);
    sponge(fin, 64, init, 40);
    std::memcpy(fin + 64, mix_hash, 32);
    sponge(out, 32, fin, 96);
    return std::memcmp(out, boundary, 32) <= 0 ? 1 : 0;
}

void nx_sha256d(const uint8_t* data, size_t size, uint8_t out[32])
{
    uint8_t first[32];
    sha256(data, size, first);
    sha256(first, 32, out);
}

void nx_kawpow_header_hash(int32_t version, const uint8_t prev_block[32], const uint8_t merkle_root[32], uint32_t time,
                           uint32_t bits, uint32_t height, uint8_t out[32])
{
    uint8_t in[80], d[32];
    std::memcpy(in, &version, 4);
    std::memcpy(in + 4, prev_block, 32);
    std::memcpy(in + 36, merkle_root, 32);
    std::memcpy(in + 68, &time, 4);
    std::memcpy(in + 72, &bits, 4);
    std::memcpy(in + 76, &height, 4);
    nx_sha256d(in, 80, d);
    for (int i = 0; i < 32; ++i) out[i] = d[31 - i];
}

void nx_block_header_parse(const uint8_t raw[120], nx_block_header* out)
{
    std::memcpy(&out->version, raw, 4);
    std::memcpy(out->prev_block, raw + 4, 32);
    std::memcpy(out->merkle_root, raw + 36, 32);
    std::memcpy(&out->time, raw + 68, 4);
    std::memcpy(&out->bits, raw + 72, 4);
    std::memcpy(&out->height, raw + 76, 4);
    std::memcpy(&out->nonce64, raw + 80, 8);
    std::memcpy(out->mix_hash, raw + 88, 32);
}

void nx_block_header_serialize(const nx_block_header* in, uint8_t raw[120])
{
    std::memcpy(raw, &in->version, 4);
    std::memcpy(raw + 4, in->prev_block, 32);
    std::memcpy(raw + 36, in->merkle_root, 32);
    std::memcpy(raw + 68, &in->time, 4);
    std::memcpy(raw + 72, &in->bits, 4);
    std::memcpy(raw + 76, &in->height, 4);
    std::memcpy(raw + 80, &in->nonce64, 8);
    std::memcpy(raw + 88, in->mix_hash, 32);
}

void nx_block_hash(const uint8_t raw[120], uint8_t out_uint256[32])
{
    // the first 80 bytes of the wire header ARE the CKAWPOWInput (block.h:213-233)
    uint8_t d[32], header_hash[32], mix[32], fin[32];
    nx_sha256d(raw, 80, d);
    for (int i = 0; i < 32; ++i) header_hash[i] = d[31 - i];
    for (int i = 0; i < 32; ++i) mix[i] = raw[88 + 31 - i];
    uint64_t nonce;
    uint32_t height;
    std::memcpy(&nonce, raw + 80, 8);
    std::memcpy(&height, raw + 76, 4);
    nx_kawpow_hash_no_verify((int)height, header_hash, mix, nonce, fin);
    for (int i = 0; i < 32; ++i) out_uint256[i] = fin[31 - i];
}

// ---- block-index records (chain.h:398-431) over VARINT (serialize.h:290-366) ----
size_t nx_varint_write(uint64_t n, uint8_t out[10])
{
    // MSB base-128, one subtracted from every digit but the last so that each number has exactly one encoding
    uint8_t tmp[10];
    int len = 0;
    for (;;) {
        tmp[len] = (uint8_t)((n & 0x7f) | (len ? 0x80 : 0x00));
        if (n <= 0x7f) break;
        n = (n >> 7) - 1;
        ++len;
    }
    size_t w = 0;
    do {
        out[w++] = tmp[len];
    } while (len--);
    return w;
}

namespace
{
// ReadVarInt<I> with I's maximum passed in: the reference throws where this returns 0
size_t varint_read_max(const uint8_t* in, size_t len, uint64_t max, uint64_t* out)
{
    uint64_t n = 0;
    for (size_t i = 0; i < len; ++i) {
        const uint8_t c = in[i];
        if (n > (max >> 7)) return 0;
        n = (n << 7) | (c & 0x7f);
        if (c & 0x80) {
            if (n == max) return 0;
            ++n;
        } else {
            *out = n;
            return i + 1;
        }
    }
    return 0;  // ran out of input
}
constexpr uint64_t MAX_I32 = 0x7fffffffull, MAX_U32 = 0xffffffffull;
}  // namespace

size_t nx_varint_read(const uint8_t* in, size_t len, uint64_t* n)
{
    if (!in || !n) return 0;
    return varint_read_max(in, len, ~0ull, n);
}

size_t nx_disk_block_index_serialize(const nx_disk_block_index* r, uint32_t activation, uint8_t* out, size_t cap)
{
    if (!r || !out || r->client_version < 0 || r->height < 0 || r->file < 0) return 0;
    uint8_t buf[7 * 10 + 4 + 32 + 32 + 4 + 4 + 8 + 32];
    size_t w = 0;
    w += nx_varint_write((uint64_t)r->client_version, buf + w);
    w += nx_varint_write((uint64_t)r->height, buf + w);
    w += nx_varint_write(r->status, buf + w);
    w += nx_varint_write(r->tx_count, buf + w);
    if (r->status & (NX_BLOCK_HAVE_DATA | NX_BLOCK_HAVE_UNDO)) w += nx_varint_write((uint64_t)r->file, buf + w);
    if (r->status & NX_BLOCK_HAVE_DATA) w += nx_varint_write(r->data_pos, buf + w);
    if (r->status & NX_BLOCK_HAVE_UNDO) w += nx_varint_write(r->undo_pos, buf + w);
    std::memcpy(buf + w, &r->version, 4); w += 4;
    std::memcpy(buf + w, r->prev_block, 32); w += 32;
    std::memcpy(buf + w, r->merkle_root, 32); w += 32;
    std::memcpy(buf + w, &r->time, 4); w += 4;
    std::memcpy(buf + w, &r->bits, 4); w += 4;
    if (r->time < activation) {
        std::memcpy(buf + w, &r->nonce, 4); w += 4;
    } else {
        std::memcpy(buf + w, &r->nonce64, 8); w += 8;
        std::memcpy(buf + w, r->mix_hash, 32); w += 32;
    }
    if (w > cap) return 0;
    std::memcpy(out, buf, w);
    return w;
}

size_t nx_disk_block_index_parse(const uint8_t* in, size_t len, uint32_t activation, nx_disk_block_index* r)
{
    if (!in || !r) return 0;
    std::memset(r, 0, sizeof *r);
    size_t p = 0;
    uint64_t v = 0;
    auto take = [&](uint64_t max) -> bool {
        const size_t used = varint_read_max(in + p, len - p, max, &v);
        p += used;
        return used != 0;
    };
    if (!take(MAX_I32)) return 0;
    r->client_version = (int32_t)v;
    if (!take(MAX_I32)) return 0;
    r->height = (int32_t)v;
    if (!take(MAX_U32)) return 0;
    r->status = (uint32_t)v;
    if (!take(MAX_U32)) return 0;
    r->tx_count = (uint32_t)v;
    if (r->status & (NX_BLOCK_HAVE_DATA | NX_BLOCK_HAVE_UNDO)) {
        if (!take(MAX_I32)) return 0;
        r->file = (int32_t)v;
    }
    if (r->status & NX_BLOCK_HAVE_DATA) {
        if (!take(MAX_U32)) return 0;
        r->data_pos = (uint32_t)v;
    }
    if (r->status & NX_BLOCK_HAVE_UNDO) {
        if (!take(MAX_U32)) return 0;
        r->undo_pos = (uint32_t)v;
    }
    auto fixed = [&](void* dst, size_t n) -> bool {
        if (len - p < n) return false;
        std::memcpy(dst, in + p, n);
        p += n;
        return true;
    };
    if (!fixed(&r->version, 4) || !fixed(r->prev_block, 32) || !fixed(r->merkle_root, 32) || !fixed(&r->time, 4) ||
        !fixed(&r->bits, 4))
        return 0;
    if (r->time < activation) {
        if (!fixed(&r->nonce, 4)) return 0;
    } else {
        if (!fixed(&r->nonce64, 8) || !fixed(r->mix_hash, 32)) return 0;
    }
    return p;
}

int nx_disk_block_index_check(const nx_disk_block_index* r, uint32_t activation, const uint8_t pow_limit_be[32], uint8_t block_id[32])
{
    if (!r) return 0;
    if (r->time < activation) return -1;
    nx_block_header h;
    h.version = r->version;
    std::memcpy(h.prev_block, r->prev_block, 32);
    std::memcpy(h.merkle_root, r->merkle_root, 32);
    h.time = r->time;
    h.bits = r->bits;
    h.height = (uint32_t)r->height;
    h.nonce64 = r->nonce64;
    std::memcpy(h.mix_hash, r->mix_hash, 32);
    uint8_t raw[120], id[32], fin[32];
    nx_block_header_serialize(&h, raw);
    nx_block_hash(raw, id);
    if (block_id) std::memcpy(block_id, id, 32);
    for (int i = 0; i < 32; ++i) fin[i] = id[31 - i];
    return nx_check_proof_of_work(fin, r->bits, pow_limit_be);
}

int nx_compact_to_boundary(uint32_t bits, uint8_t out[32])
{
    const int size = (int)(bits >> 24);
    uint32_t word = bits & 0x007fffffu;
    uint8_t le[40] = {0};
    if (size <= 3) {
        word >>= 8 * (3 - size);
        le[0] = (uint8_t)word;
        le[1] = (uint8_t)(word >> 8);
        le[2] = (uint8_t)(word >> 16);
    } else if (size <= 34) {
        le[size - 3] = (uint8_t)word;
        le[size - 2] = (uint8_t)(word >> 8);
        le[size - 1] = (uint8_t)(word >> 16);
    }
    const bool negative = word != 0 && (bits & 0x00800000u) != 0;
    const bool overflow = word != 0 && ((size > 34) || (word > 0xff && size > 33) || (word > 0xffff && size > 32));
    bool nonzero = false;
    for (int i = 0; i < 32; ++i) {
        out[i] = le[31 - i];
        nonzero |= out[i] != 0;
    }
    return (!negative && !overflow && nonzero) ? 1 : 0;
}

int nx_check_proof_of_work(const uint8_t final_hash[32], uint32_t bits, const uint8_t pow_limit_be[32])
{
    uint8_t target[32];
    if (!nx_compact_to_boundary(bits, target)) return 0;
    if (pow_limit_be && std::memcmp(target, pow_limit_be, 32) > 0) return 0;
    return std::memcmp(final_hash, target, 32) <= 0 ? 1 : 0;
}

}  // extern "C"
