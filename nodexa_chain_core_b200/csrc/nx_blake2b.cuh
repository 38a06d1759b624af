// BLAKE2b (RFC 7693) compression with a personalization block and a variable digest length, usable from host
// and device code.  Needed by the Equihash engine only (initial-string generation: one digest of
// (512/n)*n/8 bytes yields 512/n strings of n bits).  The reference node contains no BLAKE2b at all
// (SURVEY.md §0 Fact 1); this follows the RFC.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define NXB_HD __host__ __device__ __forceinline__
#else
#define NXB_HD inline
#endif

namespace nx
{
struct Blake2bState
{
    uint64_t h[8];
    uint64_t t;        // bytes absorbed so far (inputs here stay far below 2^64)
    uint8_t buf[128];  // pending block
    uint32_t buflen;
    uint32_t outlen;
};

NXB_HD uint64_t b2_rotr(uint64_t x, int n) { return (x >> n) | (x << (64 - n)); }

NXB_HD uint64_t b2_iv(int i)
{
    switch (i) {
    case 0: return 0x6a09e667f3bcc908ULL;
    case 1: return 0xbb67ae8584caa73bULL;
    case 2: return 0x3c6ef372fe94f82bULL;
    case 3: return 0xa54ff53a5f1d36f1ULL;
    case 4: return 0x510e527fade682d1ULL;
    case 5: return 0x9b05688c2b3e6c1fULL;
    case 6: return 0x1f83d9abfb41bd6bULL;
    default: return 0x5be0cd19137e2179ULL;
    }
}

NXB_HD void b2_compress(uint64_t h[8], const uint64_t m[16], uint64_t t, bool last)
{
    uint64_t v[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        v[i] = h[i];
        v[i + 8] = b2_iv(i);
    }
    v[12] ^= t;  // low word of the byte counter (high word stays 0)
    if (last) v[14] = ~v[14];
#define NXB_G(a, b, c, d, x, y)            \
    v[a] = v[a] + v[b] + (x);              \
    v[d] = b2_rotr(v[d] ^ v[a], 32);       \
    v[c] = v[c] + v[d];                    \
    v[b] = b2_rotr(v[b] ^ v[c], 24);       \
    v[a] = v[a] + v[b] + (y);              \
    v[d] = b2_rotr(v[d] ^ v[a], 16);       \
    v[c] = v[c] + v[d];                    \
    v[b] = b2_rotr(v[b] ^ v[c], 63)
    // the 12 rounds written out with the message schedule (sigma) as literals, so that m[] and v[] stay in
    // registers on the GPU (a table lookup would turn them into local-memory arrays)
    // round 0
    NXB_G(0, 4, 8, 12, m[0], m[1]);
    NXB_G(1, 5, 9, 13, m[2], m[3]);
    NXB_G(2, 6, 10, 14, m[4], m[5]);
    NXB_G(3, 7, 11, 15, m[6], m[7]);
    NXB_G(0, 5, 10, 15, m[8], m[9]);
    NXB_G(1, 6, 11, 12, m[10], m[11]);
    NXB_G(2, 7, 8, 13, m[12], m[13]);
    NXB_G(3, 4, 9, 14, m[14], m[15]);
    // round 1
    NXB_G(0, 4, 8, 12, m[14], m[10]);
    NXB_G(1, 5, 9, 13, m[4], m[8]);
    NXB_G(2, 6, 10, 14, m[9], m[15]);
    NXB_G(3, 7, 11, 15, m[13], m[6]);
    NXB_G(0, 5, 10, 15, m[1], m[12]);
    NXB_G(1, 6, 11, 12, m[0], m[2]);
    NXB_G(2, 7, 8, 13, m[11], m[7]);
    NXB_G(3, 4, 9, 14, m[5], m[3]);
    // round 2
    NXB_G(0, 4, 8, 12, m[11], m[8]);
    NXB_G(1, 5, 9, 13, m[12], m[0]);
    NXB_G(2, 6, 10, 14, m[5], m[2]);
    NXB_G(3, 7, 11, 15, m[15], m[13]);
    NXB_G(0, 5, 10, 15, m[10], m[14]);
    NXB_G(1, 6, 11, 12, m[3], m[6]);
    NXB_G(2, 7, 8, 13, m[7], m[1]);
    NXB_G(3, 4, 9, 14, m[9], m[4]);
    // round 3
    NXB_G(0, 4, 8, 12, m[7], m[9]);
    NXB_G(1, 5, 9, 13, m[3], m[1]);
    NXB_G(2, 6, 10, 14, m[13], m[12]);
    NXB_G(3, 7, 11, 15, m[11], m[14]);
    NXB_G(0, 5, 10, 15, m[2], m[6]);
    NXB_G(1, 6, 11, 12, m[5], m[10]);
    NXB_G(2, 7, 8, 13, m[4], m[0]);
    NXB_G(3, 4, 9, 14, m[15], m[8]);
    // round 4
    NXB_G(0, 4, 8, 12, m[9], m[0]);
    NXB_G(1, 5, 9, 13, m[5], m[7]);
    NXB_G(2, 6, 10, 14, m[2], m[4]);
    NXB_G(3, 7, 11, 15, m[10], m[15]);
    NXB_G(0, 5, 10, 15, m[14], m[1]);
    NXB_G(1, 6, 11, 12, m[11], m[12]);
    NXB_G(2, 7, 8, 13, m[6], m[8]);
    NXB_G(3, 4, 9, 14, m[3], m[13]);
    // round 5
    NXB_G(0, 4, 8, 12, m[2], m[12]);
    NXB_G(1, 5, 9, 13, m[6], m[10]);
    NXB_G(2, 6, 10, 14, m[0], m[11]);
    NXB_G(3, 7, 11, 15, m[8], m[3]);
    NXB_G(0, 5, 10, 15, m[4], m[13]);
    NXB_G(1, 6, 11, 12, m[7], m[5]);
    NXB_G(2, 7, 8, 13, m[15], m[14]);
    NXB_G(3, 4, 9, 14, m[1], m[9]);
    // round 6
    NXB_G(0, 4, 8, 12, m[12], m[5]);
    NXB_G(1, 5, 9, 13, m[1], m[15]);
    NXB_G(2, 6, 10, 14, m[14], m[13]);
    NXB_G(3, 7, 11, 15, m[4], m[10]);
    NXB_G(0, 5, 10, 15, m[0], m[7]);
    NXB_G(1, 6, 11, 12, m[6], m[3]);
    NXB_G(2, 7, 8, 13, m[9], m[2]);
    NXB_G(3, 4, 9, 14, m[8], m[11]);
    // round 7
    NXB_G(0, 4, 8, 12, m[13], m[11]);
    NXB_G(1, 5, 9, 13, m[7], m[14]);
    NXB_G(2, 6, 10, 14, m[12], m[1]);
    NXB_G(3, 7, 11, 15, m[3], m[9]);
    NXB_G(0, 5, 10, 15, m[5], m[0]);
    NXB_G(1, 6, 11, 12, m[15], m[4]);
    NXB_G(2, 7, 8, 13, m[8], m[6]);
    NXB_G(3, 4, 9, 14, m[2], m[10]);
    // round 8
    NXB_G(0, 4, 8, 12, m[6], m[15]);
    NXB_G(1, 5, 9, 13, m[14], m[9]);
    NXB_G(2, 6, 10, 14, m[11], m[3]);
    NXB_G(3, 7, 11, 15, m[0], m[8]);
    NXB_G(0, 5, 10, 15, m[12], m[2]);
    NXB_G(1, 6, 11, 12, m[13], m[7]);
    NXB_G(2, 7, 8, 13, m[1], m[4]);
    NXB_G(3, 4, 9, 14, m[10], m[5]);
    // round 9
    NXB_G(0, 4, 8, 12, m[10], m[2]);
    NXB_G(1, 5, 9, 13, m[8], m[4]);
    NXB_G(2, 6, 10, 14, m[7], m[6]);
    NXB_G(3, 7, 11, 15, m[1], m[5]);
    NXB_G(0, 5, 10, 15, m[15], m[11]);
    NXB_G(1, 6, 11, 12, m[9], m[14]);
    NXB_G(2, 7, 8, 13, m[3], m[12]);
    NXB_G(3, 4, 9, 14, m[13], m[0]);
    // round 10
    NXB_G(0, 4, 8, 12, m[0], m[1]);
    NXB_G(1, 5, 9, 13, m[2], m[3]);
    NXB_G(2, 6, 10, 14, m[4], m[5]);
    NXB_G(3, 7, 11, 15, m[6], m[7]);
    NXB_G(0, 5, 10, 15, m[8], m[9]);
    NXB_G(1, 6, 11, 12, m[10], m[11]);
    NXB_G(2, 7, 8, 13, m[12], m[13]);
    NXB_G(3, 4, 9, 14, m[14], m[15]);
    // round 11
    NXB_G(0, 4, 8, 12, m[14], m[10]);
    NXB_G(1, 5, 9, 13, m[4], m[8]);
    NXB_G(2, 6, 10, 14, m[9], m[15]);
    NXB_G(3, 7, 11, 15, m[13], m[6]);
    NXB_G(0, 5, 10, 15, m[1], m[12]);
    NXB_G(1, 6, 11, 12, m[0], m[2]);
    NXB_G(2, 7, 8, 13, m[11], m[7]);
    NXB_G(3, 4, 9, 14, m[5], m[3]);
#undef NXB_G
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] ^= v[i] ^ v[i + 8];
}

// outlen 1..64, no key, 16-byte personalization (parameter block bytes 48..63).
NXB_HD void b2_init(Blake2bState& s, uint32_t outlen, const uint8_t personal[16])
{
#pragma unroll
    for (int i = 0; i < 8; ++i) s.h[i] = b2_iv(i);
    s.h[0] ^= 0x01010000ULL ^ (uint64_t)outlen;  // digest length, key length 0, fanout 1, depth 1
    uint64_t p0 = 0, p1 = 0;
    for (int i = 0; i < 8; ++i) {
        p0 |= (uint64_t)personal[i] << (8 * i);
        p1 |= (uint64_t)personal[8 + i] << (8 * i);
    }
    s.h[6] ^= p0;
    s.h[7] ^= p1;
    s.t = 0;
    s.buflen = 0;
    s.outlen = outlen;
    for (int i = 0; i < 128; ++i) s.buf[i] = 0;
}

NXB_HD void b2_load_block(const uint8_t* b, uint64_t m[16])
{
    for (int i = 0; i < 16; ++i) {
        uint64_t w = 0;
        for (int j = 0; j < 8; ++j) w |= (uint64_t)b[8 * i + j] << (8 * j);
        m[i] = w;
    }
}

NXB_HD void b2_update(Blake2bState& s, const uint8_t* in, size_t len)
{
    for (size_t i = 0; i < len; ++i) {
        if (s.buflen == 128) {  // a full block is only compressed once more input follows it
            uint64_t m[16];
            b2_load_block(s.buf, m);
            s.t += 128;
            b2_compress(s.h, m, s.t, false);
            s.buflen = 0;
        }
        s.buf[s.buflen++] = in[i];
    }
}

NXB_HD void b2_final(Blake2bState s, uint8_t* out)  // by value: the midstate can be finalised many times
{
    for (uint32_t i = s.buflen; i < 128; ++i) s.buf[i] = 0;
    uint64_t m[16];
    b2_load_block(s.buf, m);
    s.t += s.buflen;
    b2_compress(s.h, m, s.t, true);
    for (uint32_t i = 0; i < s.outlen; ++i) out[i] = (uint8_t)(s.h[i >> 3] >> (8 * (i & 7)));
}

}  // namespace nx
