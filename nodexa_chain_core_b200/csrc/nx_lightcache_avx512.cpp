// AVX-512 version of the Ethash light-cache chain (reference lib/ethash/ethash.cpp:101-129).
//
// The chain is 4*n dependent keccak512 calls, so the only lever is the latency of ONE Keccak-f[1600].
// Scalar x86-64 code (16 GPRs for 25 lanes) needs ~1000 cycles per permutation; with the five planes of the
// state in five zmm registers a round is ~37 vector ops, 19 of them permutes:
//   theta  column parities as a vertical XOR of the planes, neighbours fetched with two permutes;
//   rho    one variable rotate per plane (VPROLVQ) on (plane ^ D);
//   pi     one permute per plane puts lane (x,y) where column register x' = y needs it;
//   chi    with the data now held as COLUMN registers, a ^ (~b & c) is one VPTERNLOGQ per column, no shuffles;
//   back   a 5x5 transpose (unpacks + two-source permutes) returns to plane registers for the next theta.
// Compiled with -mavx512f by g++ (not nvcc) and only called when the CPU reports AVX-512F.
#include <immintrin.h>
#include <stdint.h>
#include <string.h>

namespace nx
{
namespace
{
alignas(64) const uint64_t RC[24] = {
    0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL, 0x000000000000808bULL,
    0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL, 0x000000000000008aULL, 0x0000000000000088ULL,
    0x0000000080008009ULL, 0x000000008000000aULL, 0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL,
    0x8000000000008003ULL, 0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
    0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};

struct Consts
{
    __m512i prev, next;        // theta: element x <- C[x-1], C[x+1]
    __m512i rho[5];            // rotation counts of plane y
    __m512i pi1[5];            // F_y[j] = E_y[(3j + y) % 5]
    __m512i wl, wh;            // merge column 4 into the unpacked columns 2,3
    __m512i p0, p2, p1, p3;    // final two-source selections of the transpose
    __m512i p4a, p4b;
};

inline Consts make_consts()
{
    Consts c;
    c.prev = _mm512_setr_epi64(4, 0, 1, 2, 3, 5, 6, 7);
    c.next = _mm512_setr_epi64(1, 2, 3, 4, 0, 5, 6, 7);
    c.rho[0] = _mm512_setr_epi64(0, 1, 62, 28, 27, 0, 0, 0);
    c.rho[1] = _mm512_setr_epi64(36, 44, 6, 55, 20, 0, 0, 0);
    c.rho[2] = _mm512_setr_epi64(3, 10, 43, 25, 39, 0, 0, 0);
    c.rho[3] = _mm512_setr_epi64(41, 45, 15, 21, 8, 0, 0, 0);
    c.rho[4] = _mm512_setr_epi64(18, 2, 61, 56, 14, 0, 0, 0);
    c.pi1[0] = _mm512_setr_epi64(0, 3, 1, 4, 2, 5, 6, 7);
    c.pi1[1] = _mm512_setr_epi64(1, 4, 2, 0, 3, 5, 6, 7);
    c.pi1[2] = _mm512_setr_epi64(2, 0, 3, 1, 4, 5, 6, 7);
    c.pi1[3] = _mm512_setr_epi64(3, 1, 4, 2, 0, 5, 6, 7);
    c.pi1[4] = _mm512_setr_epi64(4, 2, 0, 3, 1, 5, 6, 7);
    // two-source permutes: index i < 8 takes element i of the first operand, 8+i element i of the second
    // WL = [L23[0],L23[1],G4[0], L23[2],L23[3],G4[2], L23[4],L23[5]]
    c.wl = _mm512_setr_epi64(0, 1, 8 + 0, 2, 3, 8 + 2, 4, 5);
    // WH = [H23[0],H23[1],G4[1], H23[2],H23[3],G4[3], G4[4], -]
    c.wh = _mm512_setr_epi64(0, 1, 8 + 1, 2, 3, 8 + 3, 8 + 4, 7);
    c.p0 = _mm512_setr_epi64(0, 1, 8 + 0, 8 + 1, 8 + 2, 5, 6, 7);  // (L01, WL)
    c.p2 = _mm512_setr_epi64(2, 3, 8 + 3, 8 + 4, 8 + 5, 5, 6, 7);  // (L01, WL)
    c.p1 = _mm512_setr_epi64(0, 1, 8 + 0, 8 + 1, 8 + 2, 5, 6, 7);  // (H01, WH)
    c.p3 = _mm512_setr_epi64(2, 3, 8 + 3, 8 + 4, 8 + 5, 5, 6, 7);  // (H01, WH)
    c.p4a = _mm512_setr_epi64(4, 5, 8 + 6, 8 + 7, 4, 5, 6, 7);     // (L01, WL) -> [G0[4],G1[4],G2[4],G3[4],-]
    c.p4b = _mm512_setr_epi64(0, 1, 2, 3, 8 + 6, 5, 6, 7);         // (that, WH) -> element 4 = G4[4]
    return c;
}

// state as five plane registers, P[y] element x = lane x + 5y (elements 5..7 are don't-care)
inline void permute(__m512i& P0, __m512i& P1, __m512i& P2, __m512i& P3, __m512i& P4, const Consts& k)
{
    for (int r = 0; r < 24; ++r) {
        // theta
        __m512i C = _mm512_ternarylogic_epi64(P0, P1, P2, 0x96);
        C = _mm512_ternarylogic_epi64(C, P3, P4, 0x96);
        const __m512i Cm = _mm512_permutexvar_epi64(k.prev, C);
        const __m512i Cp1 = _mm512_rol_epi64(_mm512_permutexvar_epi64(k.next, C), 1);
        // theta applied + rho
        const __m512i E0 = _mm512_rolv_epi64(_mm512_ternarylogic_epi64(P0, Cm, Cp1, 0x96), k.rho[0]);
        const __m512i E1 = _mm512_rolv_epi64(_mm512_ternarylogic_epi64(P1, Cm, Cp1, 0x96), k.rho[1]);
        const __m512i E2 = _mm512_rolv_epi64(_mm512_ternarylogic_epi64(P2, Cm, Cp1, 0x96), k.rho[2]);
        const __m512i E3 = _mm512_rolv_epi64(_mm512_ternarylogic_epi64(P3, Cm, Cp1, 0x96), k.rho[3]);
        const __m512i E4 = _mm512_rolv_epi64(_mm512_ternarylogic_epi64(P4, Cm, Cp1, 0x96), k.rho[4]);
        // pi, first half: F_s element j = B[x'=s][y'=j]
        const __m512i F0 = _mm512_permutexvar_epi64(k.pi1[0], E0);
        const __m512i F1 = _mm512_permutexvar_epi64(k.pi1[1], E1);
        const __m512i F2 = _mm512_permutexvar_epi64(k.pi1[2], E2);
        const __m512i F3 = _mm512_permutexvar_epi64(k.pi1[3], E3);
        const __m512i F4 = _mm512_permutexvar_epi64(k.pi1[4], E4);
        // chi across the column registers: a ^ (~b & c)
        __m512i G0 = _mm512_ternarylogic_epi64(F0, F1, F2, 0xD2);
        const __m512i G1 = _mm512_ternarylogic_epi64(F1, F2, F3, 0xD2);
        const __m512i G2 = _mm512_ternarylogic_epi64(F2, F3, F4, 0xD2);
        const __m512i G3 = _mm512_ternarylogic_epi64(F3, F4, F0, 0xD2);
        const __m512i G4 = _mm512_ternarylogic_epi64(F4, F0, F1, 0xD2);
        // iota: lane (0,0) = column 0, element 0
        G0 = _mm512_xor_si512(G0, _mm512_maskz_set1_epi64(1, (long long)RC[r]));
        // pi, second half: transpose column registers G_s[j] into plane registers P_j[s]
        const __m512i L01 = _mm512_unpacklo_epi64(G0, G1);
        const __m512i H01 = _mm512_unpackhi_epi64(G0, G1);
        const __m512i L23 = _mm512_unpacklo_epi64(G2, G3);
        const __m512i H23 = _mm512_unpackhi_epi64(G2, G3);
        const __m512i WL = _mm512_permutex2var_epi64(L23, k.wl, G4);
        const __m512i WH = _mm512_permutex2var_epi64(H23, k.wh, G4);
        P0 = _mm512_permutex2var_epi64(L01, k.p0, WL);
        P2 = _mm512_permutex2var_epi64(L01, k.p2, WL);
        P1 = _mm512_permutex2var_epi64(H01, k.p1, WH);
        P3 = _mm512_permutex2var_epi64(H01, k.p3, WH);
        P4 = _mm512_permutex2var_epi64(_mm512_permutex2var_epi64(L01, k.p4a, WL), k.p4b, WH);
    }
}

// keccak512 of 64 bytes: rate 72 bytes = lanes 0..8, pad 0x01 at byte 64 and 0x80 at byte 71 (both in lane 8)
inline void keccak512_64(const uint64_t in[8], uint64_t out[8], const Consts& k)
{
    __m512i P0 = _mm512_maskz_loadu_epi64(0x1f, in);
    __m512i P1 = _mm512_maskz_loadu_epi64(0x07, in + 5);
    P1 = _mm512_or_si512(P1, _mm512_maskz_set1_epi64(0x08, (long long)0x8000000000000001ULL));
    __m512i P2 = _mm512_setzero_si512(), P3 = P2, P4 = P2;
    permute(P0, P1, P2, P3, P4, k);
    _mm512_mask_storeu_epi64(out, 0x1f, P0);
    _mm512_mask_storeu_epi64(out + 5, 0x07, P1);
}
}  // namespace

// Raw permutation on 25 lanes (used by the self-test against the scalar code).
void keccakf1600_avx512(uint64_t a[25])
{
    const Consts k = make_consts();
    __m512i P0 = _mm512_maskz_loadu_epi64(0x1f, a), P1 = _mm512_maskz_loadu_epi64(0x1f, a + 5),
            P2 = _mm512_maskz_loadu_epi64(0x1f, a + 10), P3 = _mm512_maskz_loadu_epi64(0x1f, a + 15),
            P4 = _mm512_maskz_loadu_epi64(0x1f, a + 20);
    permute(P0, P1, P2, P3, P4, k);
    _mm512_mask_storeu_epi64(a, 0x1f, P0);
    _mm512_mask_storeu_epi64(a + 5, 0x1f, P1);
    _mm512_mask_storeu_epi64(a + 10, 0x1f, P2);
    _mm512_mask_storeu_epi64(a + 15, 0x1f, P3);
    _mm512_mask_storeu_epi64(a + 20, 0x1f, P4);
}

void host_build_light_cache_avx512(uint8_t* cache, int n, const uint8_t seed[32])
{
    const Consts k = make_consts();
    uint64_t* c = reinterpret_cast<uint64_t*>(cache);
    uint64_t item[8];
    {
        // keccak512 of the 32-byte seed: lanes 0..3, 0x01 at byte 32 (lane 4), 0x80 at byte 71 (lane 8)
        uint64_t s[4];
        memcpy(s, seed, 32);
        __m512i P0 = _mm512_maskz_loadu_epi64(0x0f, s);
        P0 = _mm512_or_si512(P0, _mm512_maskz_set1_epi64(0x10, 1));
        __m512i P1 = _mm512_maskz_set1_epi64(0x08, (long long)0x8000000000000000ULL);
        __m512i P2 = _mm512_setzero_si512(), P3 = P2, P4 = P2;
        permute(P0, P1, P2, P3, P4, k);
        _mm512_mask_storeu_epi64(item, 0x1f, P0);
        _mm512_mask_storeu_epi64(item + 5, 0x07, P1);
    }
    memcpy(c, item, 64);
    for (int i = 1; i < n; ++i) {
        keccak512_64(item, item, k);
        memcpy(c + 8 * (size_t)i, item, 64);
    }
    const uint32_t un = (uint32_t)n;
    for (int q = 0; q < 3; ++q) {
        for (int i = 0; i < n; ++i) {
            const uint32_t v = (uint32_t)c[8 * (size_t)i] % un;
            const uint32_t w = (uint32_t)(n + (i - 1)) % un;
            const __m512i x = _mm512_xor_si512(_mm512_loadu_si512(c + 8 * (size_t)v), _mm512_loadu_si512(c + 8 * (size_t)w));
            _mm512_storeu_si512(item, x);
            keccak512_64(item, item, k);
            memcpy(c + 8 * (size_t)i, item, 64);
        }
    }
}

}  // namespace nx
