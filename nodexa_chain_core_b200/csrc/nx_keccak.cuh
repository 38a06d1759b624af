// Keccak-f[1600] / Keccak-f[800] and the two fixed-size sponges the Ethash/KawPow path needs,
// usable from host and device code.  Product code (not the oracle).
//
// Replaces (behaviourally): reference src/crypto/ethash/lib/keccak/keccakf1600.c:40-255,
// keccakf800.c:38-253 and the 32/64-byte cases of keccak.c:46-127 (original Keccak padding:
// 0x01 after the message, 0x80 in the last byte of the rate).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define NX_HD __host__ __device__ __forceinline__
#else
#define NX_HD inline
#endif

namespace nx
{
#define NX_KECCAK_RC_VALUES                                                                           \
    0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,      \
    0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,      \
    0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,      \
    0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,      \
    0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,      \
    0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL

static const uint64_t KECCAK_RC_HOST[24] = {NX_KECCAK_RC_VALUES};
#ifdef __CUDACC__
static __device__ __constant__ uint64_t KECCAK_RC_DEV[24] = {NX_KECCAK_RC_VALUES};
#endif

NX_HD uint64_t keccak_rc(int r)
{
#ifdef __CUDA_ARCH__
    return KECCAK_RC_DEV[r];
#else
    return KECCAK_RC_HOST[r];
#endif
}

template <typename T>
NX_HD T rol(T v, unsigned n)
{
    constexpr unsigned bits = sizeof(T) * 8;
    n %= bits;
    return n ? (T)((v << n) | (v >> (bits - n))) : v;
}

// One round without iota.  All indices are literals so the 25 lanes stay in registers on the GPU.
// rho offsets are for 64-bit lanes; 32-bit lanes take them modulo 32 (done inside rol).
template <typename T>
NX_HD void keccak_round(T a[25])
{
    const T c0 = a[0] ^ a[5] ^ a[10] ^ a[15] ^ a[20];
    const T c1 = a[1] ^ a[6] ^ a[11] ^ a[16] ^ a[21];
    const T c2 = a[2] ^ a[7] ^ a[12] ^ a[17] ^ a[22];
    const T c3 = a[3] ^ a[8] ^ a[13] ^ a[18] ^ a[23];
    const T c4 = a[4] ^ a[9] ^ a[14] ^ a[19] ^ a[24];
    const T d0 = c4 ^ rol(c1, 1), d1 = c0 ^ rol(c2, 1), d2 = c1 ^ rol(c3, 1), d3 = c2 ^ rol(c4, 1),
            d4 = c3 ^ rol(c0, 1);

    // theta + rho + pi: b[y + 5*((2x+3y)%5)] = rol(a[x+5y]^d[x], rho[x+5y])
    const T b0 = a[0] ^ d0;
    const T b10 = rol(a[1] ^ d1, 1);
    const T b20 = rol(a[2] ^ d2, 62);
    const T b5 = rol(a[3] ^ d3, 28);
    const T b15 = rol(a[4] ^ d4, 27);
    const T b16 = rol(a[5] ^ d0, 36);
    const T b1 = rol(a[6] ^ d1, 44);
    const T b11 = rol(a[7] ^ d2, 6);
    const T b21 = rol(a[8] ^ d3, 55);
    const T b6 = rol(a[9] ^ d4, 20);
    const T b7 = rol(a[10] ^ d0, 3);
    const T b17 = rol(a[11] ^ d1, 10);
    const T b2 = rol(a[12] ^ d2, 43);
    const T b12 = rol(a[13] ^ d3, 25);
    const T b22 = rol(a[14] ^ d4, 39);
    const T b23 = rol(a[15] ^ d0, 41);
    const T b8 = rol(a[16] ^ d1, 45);
    const T b18 = rol(a[17] ^ d2, 15);
    const T b3 = rol(a[18] ^ d3, 21);
    const T b13 = rol(a[19] ^ d4, 8);
    const T b14 = rol(a[20] ^ d0, 18);
    const T b24 = rol(a[21] ^ d1, 2);
    const T b9 = rol(a[22] ^ d2, 61);
    const T b19 = rol(a[23] ^ d3, 56);
    const T b4 = rol(a[24] ^ d4, 14);

    // chi
    a[0] = b0 ^ (~b1 & b2);
    a[1] = b1 ^ (~b2 & b3);
    a[2] = b2 ^ (~b3 & b4);
    a[3] = b3 ^ (~b4 & b0);
    a[4] = b4 ^ (~b0 & b1);
    a[5] = b5 ^ (~b6 & b7);
    a[6] = b6 ^ (~b7 & b8);
    a[7] = b7 ^ (~b8 & b9);
    a[8] = b8 ^ (~b9 & b5);
    a[9] = b9 ^ (~b5 & b6);
    a[10] = b10 ^ (~b11 & b12);
    a[11] = b11 ^ (~b12 & b13);
    a[12] = b12 ^ (~b13 & b14);
    a[13] = b13 ^ (~b14 & b10);
    a[14] = b14 ^ (~b10 & b11);
    a[15] = b15 ^ (~b16 & b17);
    a[16] = b16 ^ (~b17 & b18);
    a[17] = b17 ^ (~b18 & b19);
    a[18] = b18 ^ (~b19 & b15);
    a[19] = b19 ^ (~b15 & b16);
    a[20] = b20 ^ (~b21 & b22);
    a[21] = b21 ^ (~b22 & b23);
    a[22] = b22 ^ (~b23 & b24);
    a[23] = b23 ^ (~b24 & b20);
    a[24] = b24 ^ (~b20 & b21);
}

NX_HD void keccakf1600(uint64_t a[25])
{
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int r = 0; r < 24; ++r) {
        keccak_round<uint64_t>(a);
        a[0] ^= keccak_rc(r);
    }
}

// 22 rounds; round constants are the low halves of the f1600 ones.
NX_HD void keccakf800(uint32_t a[25])
{
#ifdef __CUDA_ARCH__
#pragma unroll 1
#endif
    for (int r = 0; r < 22; ++r) {
        keccak_round<uint32_t>(a);
        a[0] ^= (uint32_t)keccak_rc(r);
    }
}

// keccak512 of exactly 64 bytes (8 lanes in, 8 lanes out), rate 72 bytes = 9 lanes.
NX_HD void keccak512_64(uint64_t w[8])
{
    uint64_t a[25];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = w[i];
    a[8] = 0x8000000000000001ULL;  // 0x01 at byte 64, 0x80 at byte 71: both inside lane 8
#pragma unroll
    for (int i = 9; i < 25; ++i) a[i] = 0;
    keccakf1600(a);
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = a[i];
}

// keccak512 of exactly 32 bytes -> 64 bytes.
NX_HD void keccak512_32(const uint64_t in[4], uint64_t out[8])
{
    uint64_t a[25];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = in[i];
    a[4] = 0x01;
#pragma unroll
    for (int i = 5; i < 25; ++i) a[i] = 0;
    a[8] = 0x8000000000000000ULL;
    keccakf1600(a);
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = a[i];
}

// keccak256 of exactly 32 bytes, rate 136 bytes = 17 lanes.
NX_HD void keccak256_32(uint64_t w[4])
{
    uint64_t a[25];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = w[i];
    a[4] = 0x01;
#pragma unroll
    for (int i = 5; i < 25; ++i) a[i] = 0;
    a[16] = 0x8000000000000000ULL;
    keccakf1600(a);
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = a[i];
}

}  // namespace nx
