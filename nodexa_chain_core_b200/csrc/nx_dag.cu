// Ethash DAG generation (K3), DAG checksum and light-cache kernels for sm_100a.
//
// Replaces: item_state / calculate_dataset_item_512/1024/2048
// (reference src/crypto/ethash/lib/ethash/ethash.cpp:180-251, 512 parents per 64-byte item, :27) and,
// for the L1 cache, the tail of create_epoch_context (:168-170): the L1 cache is simply the first
// 16 KiB of the DAG, so it falls out of K3.
//
// Layout in HBM: the DAG is an array of 256-byte items ("hash2048"); item i is made of the four
// 64-byte Ethash items 4i..4i+3.  The light cache is an array of 64-byte items and is read with
// 16-byte loads, four lanes per parent, so a warp-wide load touches 8 parents x 64 contiguous bytes
// (2 full sectors each) instead of 32 half-used sectors.
//
// Bound: L2 gather bandwidth.  Every 64 output bytes need 512 random 64-byte reads from the
// 16-80 MiB light cache (32 KiB of L2 traffic per 64 B written); HBM only sees the
// DAG being written once.
#include "nx_internal.hpp"
#include "nx_keccak.cuh"

#include <cstdlib>
#include <cstring>
#include <mutex>

namespace nx
{
namespace
{
constexpr uint32_t FNV_PRIME = 0x01000193u;
__device__ __forceinline__ uint32_t fnv1(uint32_t u, uint32_t v) { return (u * FNV_PRIME) ^ v; }

// x % d for a run-time d (the prime light-cache item count), without an integer divide:
// q = floor(x * floor(2^32/d) / 2^32) is floor(x/d) or one less, so one conditional subtract fixes it.
struct FastMod
{
    uint32_t d, m;
};
__device__ __forceinline__ uint32_t fast_mod(uint32_t x, FastMod f)
{
    const uint32_t r = x - __umulhi(x, f.m) * f.d;
    return min(r, r - f.d);  // r - d wraps to a huge value when r < d
}

constexpr int DAG_THREADS = 128;

// 4x4 exchange inside a lane quad.  Before: lane q holds the 16 mix words of ITS item in w[0..15].
// After: lane s holds words 4s..4s+3 of all four items of the quad, a[q][c] = word 4s+c of lane q's item.
// (And back again with the same pattern.)  Register indices must be literals, so the word group a lane
// sends in step d - the one its partner s^d needs - is picked with selects.
__device__ __forceinline__ void quad_scatter(const uint32_t w[16], uint32_t a[4][4], uint32_t sub)
{
#pragma unroll
    for (uint32_t d = 0; d < 4; ++d) {
        const uint32_t want = sub ^ d;  // my partner in this step, and the word group it needs from me
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t send = want == 0 ? w[c] : want == 1 ? w[4 + c] : want == 2 ? w[8 + c] : w[12 + c];
            const uint32_t got = __shfl_xor_sync(0xffffffffu, send, d, 4);
            // `got` is word 4*sub+c of lane (sub^d)'s item
            if (want == 0) a[0][c] = got;
            if (want == 1) a[1][c] = got;
            if (want == 2) a[2][c] = got;
            if (want == 3) a[3][c] = got;
        }
    }
}

__device__ __forceinline__ void quad_gather(const uint32_t a[4][4], uint32_t w[16], uint32_t sub)
{
#pragma unroll
    for (uint32_t d = 0; d < 4; ++d) {
        const uint32_t partner = sub ^ d;  // I hold words 4*sub.. of partner's item in a[partner][*]
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t send = partner == 0 ? a[0][c] : partner == 1 ? a[1][c] : partner == 2 ? a[2][c] : a[3][c];
            const uint32_t got = __shfl_xor_sync(0xffffffffu, send, d, 4);
            // `got` is word 4*partner+c of MY item
            if (partner == 0) w[c] = got;
            if (partner == 1) w[4 + c] = got;
            if (partner == 2) w[8 + c] = got;
            if (partner == 3) w[12 + c] = got;
        }
    }
}

// One thread owns one 64-byte item for the two Keccak calls.  The 512-parent FNV walk is done by the four
// lanes of a quad for their four items AT THE SAME TIME: lane s holds words 4s..4s+3 of each item, so every
// step issues four independent 16-byte loads per lane (64 contiguous bytes per parent across the quad).
// Four parents in flight per lane is what hides the L2 latency of the gather.
// Fused generate + all-gather: the slice is stored into the DAG buffer of EVERY GPU of the job (outs.ptr[0] is
// the local one, the others are peer-mapped pointers reached over NVLink), so the exchange overlaps the L2
// gather instead of following it as a separate collective.
struct DagOuts
{
    uint4* ptr[NX_MAX_PEERS];  // each already offset to the first item of this launch
    int n;
};

__global__ void __launch_bounds__(DAG_THREADS)
dag_items_kernel(const uint4* __restrict__ light, FastMod fm, uint32_t first512, uint32_t count512, DagOuts outs)
{
    const uint32_t gid = blockIdx.x * DAG_THREADS + threadIdx.x;
    const uint32_t sub = threadIdx.x & 3u;
    const bool live = gid < count512;
    const uint32_t k = first512 + (live ? gid : 0u);  // dead lanes recompute item `first512`; never stored

    uint32_t mix[16];
    {
        const uint4* src = light + (size_t)fast_mod(k, fm) * 4;
        uint64_t w[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint4 v = __ldg(src + i);
            w[2 * i] = (uint64_t)v.x | ((uint64_t)v.y << 32);
            w[2 * i + 1] = (uint64_t)v.z | ((uint64_t)v.w << 32);
        }
        w[0] ^= (uint64_t)k;  // mix.word32s[0] ^= seed
        keccak512_64(w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            mix[2 * i] = (uint32_t)w[i];
            mix[2 * i + 1] = (uint32_t)(w[i] >> 32);
        }
    }

    uint32_t a[4][4];
    quad_scatter(mix, a, sub);
    uint32_t seed[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) seed[q] = __shfl_sync(0xffffffffu, k, q, 4);

#pragma unroll 1
    for (uint32_t j = 0; j < 512; j += 16) {
#pragma unroll
        for (uint32_t jj = 0; jj < 16; ++jj) {
            // mix word jj of every item lives in quad-lane jj/4, component jj%4
            uint4 pv[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint32_t t = fnv1(seed[q] ^ (j + jj), __shfl_sync(0xffffffffu, a[q][jj & 3], jj >> 2, 4));
                pv[q] = __ldg(light + (size_t)fast_mod(t, fm) * 4 + sub);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                a[q][0] = fnv1(a[q][0], pv[q].x);
                a[q][1] = fnv1(a[q][1], pv[q].y);
                a[q][2] = fnv1(a[q][2], pv[q].z);
                a[q][3] = fnv1(a[q][3], pv[q].w);
            }
        }
    }
    quad_gather(a, mix, sub);

    {
        uint64_t w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = (uint64_t)mix[2 * i] | ((uint64_t)mix[2 * i + 1] << 32);
        keccak512_64(w);
        if (live) {
            // streaming stores: the DAG is written once and must not push the light cache out of L2
            uint4 v[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                v[i] = make_uint4((uint32_t)w[2 * i], (uint32_t)(w[2 * i] >> 32), (uint32_t)w[2 * i + 1],
                                  (uint32_t)(w[2 * i + 1] >> 32));
            for (int t = 0; t < outs.n; ++t) {
                uint4* dst = outs.ptr[t] + (size_t)gid * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) __stcs(dst + i, v[i]);
            }
        }
    }
}

// ---- position-weighted checksum: sum_i w_i * (2i+1) mod 2^64 ----
constexpr int CSUM_THREADS = 256;

__global__ void __launch_bounds__(CSUM_THREADS)
checksum_partial_kernel(const uint64_t* __restrict__ words, size_t n, uint64_t* __restrict__ partials)
{
    __shared__ uint64_t red[CSUM_THREADS / 32];
    uint64_t acc = 0;
    const size_t stride = (size_t)gridDim.x * CSUM_THREADS;
    for (size_t i = (size_t)blockIdx.x * CSUM_THREADS + threadIdx.x; i < n; i += stride)
        acc += __ldg(words + i) * (2 * (uint64_t)i + 1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t s = 0;
        for (int i = 0; i < CSUM_THREADS / 32; ++i) s += red[i];
        partials[blockIdx.x] = s;
    }
}

__global__ void checksum_final_kernel(const uint64_t* __restrict__ partials, int n, uint64_t* __restrict__ out)
{
    uint64_t acc = 0;
    for (int i = threadIdx.x; i < n; i += 32) acc += partials[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) *out = acc;
}

// ---- light cache on the GPU (K4) ----
// The chain is strictly serial (reference ethash.cpp:104-127): cache[i] needs cache[i-1].  One warp,
// lane 0 does the permutation; it exists so that a host without spare cores can still build an epoch
// entirely on the device.  See DESIGN.md for why the host build is the default.
__global__ void light_cache_kernel(uint64_t* __restrict__ cache, uint32_t n, uint64_t s0, uint64_t s1, uint64_t s2,
                                   uint64_t s3)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint64_t item[8];
    {
        const uint64_t in[4] = {s0, s1, s2, s3};
        keccak512_32(in, item);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) cache[k] = item[k];
    for (uint32_t i = 1; i < n; ++i) {
        keccak512_64(item);
#pragma unroll
        for (int k = 0; k < 8; ++k) cache[8 * (size_t)i + k] = item[k];
    }
    __threadfence();
    for (int q = 0; q < 3; ++q) {
        for (uint32_t i = 0; i < n; ++i) {
            // cache[(i-1) mod n] is `item` from the previous step except at i == 0
            const uint32_t v = (uint32_t)__ldcg(cache + 8 * (size_t)i) % n;
            const uint32_t w = (n + i - 1) % n;
            const uint64_t* a = cache + 8 * (size_t)v;
            uint64_t x[8];
            if (i == 0) {
                const uint64_t* b = cache + 8 * (size_t)w;
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = __ldcg(a + k) ^ __ldcg(b + k);
            } else if (v == w) {
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = 0;
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = __ldcg(a + k) ^ item[k];
            }
            keccak512_64(x);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                item[k] = x[k];
                __stcg(cache + 8 * (size_t)i + k, x[k]);
            }
        }
    }
}
// ---- light cache on the GPU, one WARP per permutation (K4w) ----
// Same serial chain, but Keccak-f[1600] is spread over 25 lanes: lane l = x + 5y holds state word A[x][y].  What counts
// is the latency of ONE permutation, so the round is laid out for the shortest chain of dependent shuffles:
//   theta   C[x] by a 2-step butterfly over the five lanes of a column (+1 independent shuffle), D from two shuffles;
//   rho     a per-lane rotation amount (two funnel shifts after a conditional swap of the halves);
//   pi+chi  ONE step: the three operands b[x], b[x+1], b[x+2] of chi are fetched straight from the pre-pi lanes
//           (source lane = pi^-1 of the operand's position), so pi costs no shuffle of its own;
//   iota    lane 0.
// Four dependent 64-bit shuffles per round.  The item index of the second phase (first word of the OLD item i) is
// prefetched 32 items at a time, and cache[v] of item i+1 is loaded before the permutation of item i starts (the only
// hazard, v == i, is the case the chain defines as x = 0 anyway).  Bit-identical to the host chain (tests/test_gpu_dag.py).
struct WarpKeccak
{
    uint32_t rot;                 // rho offset of this lane's word
    int up5, up10, up20;          // lanes of the same column: y+1, y+2, y+4
    int xm1, xp1;                 // same row, x-1 and x+1
    int src0, src1, src2;         // pre-pi lanes of b[x][y], b[x+1][y], b[x+2][y]
};

__device__ __forceinline__ uint64_t shfl64(uint64_t v, int src) { return __shfl_sync(0xffffffffu, v, src); }

__device__ __forceinline__ uint64_t rol64_var(uint64_t v, uint32_t n)
{
    uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
    if (n & 32u) {
        const uint32_t t = lo;
        lo = hi;
        hi = t;
    }
    const uint32_t s = n & 31u;
    const uint32_t nhi = __funnelshift_l(lo, hi, s), nlo = __funnelshift_l(hi, lo, s);
    return ((uint64_t)nhi << 32) | nlo;
}

__device__ __forceinline__ uint64_t warp_keccakf1600(uint64_t a, const WarpKeccak& k, uint32_t lane)
{
#pragma unroll 1
    for (int r = 0; r < 24; ++r) {
        const uint64_t far = shfl64(a, k.up20);
        const uint64_t t = a ^ shfl64(a, k.up5);
        const uint64_t c = t ^ shfl64(t, k.up10) ^ far;
        const uint64_t cp = shfl64(c, k.xp1), cm = shfl64(c, k.xm1);
        a ^= cm ^ ((cp << 1) | (cp >> 63));
        const uint64_t b = rol64_var(a, k.rot);
        const uint64_t b0 = shfl64(b, k.src0), b1 = shfl64(b, k.src1), b2 = shfl64(b, k.src2);
        a = b0 ^ (~b1 & b2);
        if (lane == 0) a ^= KECCAK_RC_DEV[r];
    }
    return a;
}

__global__ void __launch_bounds__(32) light_cache_warp_kernel(uint64_t* __restrict__ cache, uint32_t n, uint64_t s0,
                                                              uint64_t s1, uint64_t s2, uint64_t s3)
{
    if (blockIdx.x != 0) return;
    const uint32_t lane = threadIdx.x;
    WarpKeccak k;
    {
        const uint32_t RHO[25] = {0, 1, 62, 28, 27, 36, 44, 6, 55, 20, 3, 10, 43, 25, 39, 41, 45, 15, 21, 8, 18, 2, 61, 56, 14};
        const uint32_t l = lane < 25u ? lane : 0u;  // lanes 25..31 mirror lane 0's wiring and carry zeros
        const uint32_t x = l % 5u, y = l / 5u;
        k.rot = RHO[l];
        k.up5 = (int)((l + 5u) % 25u);
        k.up10 = (int)((l + 10u) % 25u);
        k.up20 = (int)((l + 20u) % 25u);
        k.xm1 = (int)((x + 4u) % 5u + 5u * y);
        k.xp1 = (int)((x + 1u) % 5u + 5u * y);
        // pi: the word at (x', y') moves to (y', 2x' + 3y'); so position (X, Y) is fed from x' = (X + 3Y) % 5, y' = X
        auto pre_pi = [](uint32_t X, uint32_t Y) { return (int)((X + 3u * Y) % 5u + 5u * X); };
        k.src0 = pre_pi(x, y);
        k.src1 = pre_pi((x + 1u) % 5u, y);
        k.src2 = pre_pi((x + 2u) % 5u, y);
    }
    const uint64_t PAD64 = 0x8000000000000001ULL;  // keccak512 of 64 bytes: 0x01 and 0x80 both land in word 8
    // item 0 = keccak512(seed)
    uint64_t a = lane == 0 ? s0 : lane == 1 ? s1 : lane == 2 ? s2 : lane == 3 ? s3 : lane == 4 ? 0x01ULL
                                                                                 : lane == 8 ? 0x8000000000000000ULL : 0ULL;
    a = warp_keccakf1600(a, k, lane);
    if (lane < 8) cache[lane] = a;
    for (uint32_t i = 1; i < n; ++i) {
        a = lane < 8 ? a : lane == 8 ? PAD64 : 0ULL;
        a = warp_keccakf1600(a, k, lane);
        if (lane < 8) cache[8 * (size_t)i + lane] = a;
    }
    __threadfence();
    for (int q = 0; q < 3; ++q) {
        // lanes 0..7 of `a` hold item (i-1) mod n; before i = 0 that is the last item of the previous pass (still in `a`)
        uint32_t idx_words = 0, idx_next = 0;
        uint64_t operand = 0;  // cache[v] of the item about to be processed, lanes 0..7
        uint32_t v_cur = 0;
        for (uint32_t i = 0; i < n; ++i) {
            const uint32_t j = i & 31u;
            if (j == 0) {
                // old first words of items i..i+31 (none of them rewritten yet in this pass)
                if (i == 0) {
                    const uint32_t it = i + lane;
                    idx_words = it < n ? (uint32_t)__ldcg(cache + 8 * (size_t)it) : 0u;
                    v_cur = __shfl_sync(0xffffffffu, idx_words, 0) % n;
                    operand = lane < 8 ? __ldcg(cache + 8 * (size_t)v_cur + lane) : 0ULL;
                } else {
                    idx_words = idx_next;
                }
            }
            if (j == 8) {  // next group's index words, 24 items ahead of their first use
                const uint32_t it = i - 8u + 32u + lane;
                idx_next = it < n ? (uint32_t)__ldcg(cache + 8 * (size_t)it) : 0u;
            }
            // v of the NEXT item and its operand, issued before this item's permutation
            const uint32_t w_next = i;  // item i is (i+1)-1
            uint32_t v_next = 0;
            uint64_t operand_next = 0;
            const bool has_next = i + 1u < n;
            if (has_next) {
                const uint32_t word = j == 31u ? __shfl_sync(0xffffffffu, idx_next, 0) : __shfl_sync(0xffffffffu, idx_words, (int)(j + 1u));
                v_next = word % n;
                if (v_next != w_next && lane < 8) operand_next = __ldcg(cache + 8 * (size_t)v_next + lane);
            }
            const uint32_t w = i == 0 ? n - 1u : i - 1u;
            uint64_t x = lane < 8 ? (v_cur == w ? 0ULL : (a ^ operand)) : lane == 8 ? PAD64 : 0ULL;
            a = warp_keccakf1600(x, k, lane);
            if (lane < 8) __stcg(cache + 8 * (size_t)i + lane, a);
            v_cur = v_next;
            operand = operand_next;
        }
    }
}

// ---- classic Ethash hash (K7) ----
// Replaces hash_seed / hash_kernel / hash_final (reference ethash.cpp:257-303) applied to a batch of
// nonces.  The DAG is addressed as 128-byte items here (index_limit = full_dataset_num_items, :279).
// Every thread owns one nonce for the two Keccak calls; the 64-access FNV walk of each nonce is done by its
// group of 8 lanes together, four nonces at a time: lane s holds mix words 4s..4s+3 of each, so an access is one
// 16-byte load per lane = one contiguous 128-byte item per group (4 items per warp-wide load).
// Algorithmic bytes: 64 x 128 B = 8 192 B of DAG per hash.
constexpr int ETHASH_THREADS = 128;
// Q = nonces of a group walked at the same time (4 or 8): Q independent 128-byte reads in flight per group.
template <int Q>
__global__ void __launch_bounds__(ETHASH_THREADS)
ethash_hash_kernel(const uint4* __restrict__ dag, FastMod fm, uint64_t h0, uint64_t h1, uint64_t h2, uint64_t h3,
                   uint64_t start_nonce, uint64_t count, uint4* __restrict__ out)
{
    const uint64_t gid = (uint64_t)blockIdx.x * ETHASH_THREADS + threadIdx.x;
    const uint32_t sub = threadIdx.x & 7u;
    const bool live = gid < count;  // dead lanes still take part in the shuffles
    uint64_t a[25];
    a[0] = h0; a[1] = h1; a[2] = h2; a[3] = h3;
    a[4] = start_nonce + gid;
    a[5] = 0x01;  // 40-byte message: pad byte right after it
#pragma unroll
    for (int i = 6; i < 25; ++i) a[i] = 0;
    a[8] = 0x8000000000000000ULL;  // rate 72 bytes
    keccakf1600(a);
    uint32_t seed[16];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        seed[2 * i] = (uint32_t)a[i];
        seed[2 * i + 1] = (uint32_t)(a[i] >> 32);
    }
    uint32_t mh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) mh[i] = 0;

#pragma unroll 1
    for (uint32_t p0 = 0; p0 < 8; p0 += Q) {
        // the 32-word mix of a nonce starts as its seed twice over: lane s takes words 4s..4s+3,
        // i.e. seed words (4s+c) mod 16 = 4*(s&3)+c
        uint32_t m[Q][4];
        uint32_t seed0[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) {
#pragma unroll
            for (int c = 0; c < 4; ++c) m[q][c] = 0;
#pragma unroll
            for (int w = 0; w < 16; ++w) {
                const uint32_t v = __shfl_sync(0xffffffffu, seed[w], p0 + q, 8);
                if ((uint32_t)(w >> 2) == (sub & 3u)) m[q][w & 3] = v;
            }
            seed0[q] = __shfl_sync(0xffffffffu, seed[0], p0 + q, 8);
        }
#pragma unroll 1
        for (uint32_t i = 0; i < 64; i += 32) {
#pragma unroll
            for (uint32_t ii = 0; ii < 32; ++ii) {
                // mix word ii lives in lane ii/4, component ii%4
                uint4 v[Q];
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    const uint32_t t = fnv1((i + ii) ^ seed0[q], __shfl_sync(0xffffffffu, m[q][ii & 3], ii >> 2, 8));
                    v[q] = __ldg(dag + (size_t)fast_mod(t, fm) * 8 + sub);
                }
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    m[q][0] = fnv1(m[q][0], v[q].x);
                    m[q][1] = fnv1(m[q][1], v[q].y);
                    m[q][2] = fnv1(m[q][2], v[q].z);
                    m[q][3] = fnv1(m[q][3], v[q].w);
                }
            }
        }
        // lane s holds words 4s..4s+3 -> compressed word s of each mix hash; hand all 8 words to the owner lane
#pragma unroll
        for (int q = 0; q < Q; ++q) {
            const uint32_t h = fnv1(fnv1(fnv1(m[q][0], m[q][1]), m[q][2]), m[q][3]);
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                const uint32_t v = __shfl_sync(0xffffffffu, h, w, 8);
                if (sub == p0 + q) mh[w] = v;
            }
        }
    }
    // final = keccak256(seed[64] || mix_hash[32]): 96 bytes, rate 136
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = (uint64_t)seed[2 * i] | ((uint64_t)seed[2 * i + 1] << 32);
#pragma unroll
    for (int i = 0; i < 4; ++i) a[8 + i] = (uint64_t)mh[2 * i] | ((uint64_t)mh[2 * i + 1] << 32);
    a[12] = 0x01;
#pragma unroll
    for (int i = 13; i < 25; ++i) a[i] = 0;
    a[16] = 0x8000000000000000ULL;
    keccakf1600(a);
    if (!live) return;
    uint4* o = out + gid * 4;
    o[0] = make_uint4((uint32_t)a[0], (uint32_t)(a[0] >> 32), (uint32_t)a[1], (uint32_t)(a[1] >> 32));
    o[1] = make_uint4((uint32_t)a[2], (uint32_t)(a[2] >> 32), (uint32_t)a[3], (uint32_t)(a[3] >> 32));
    o[2] = make_uint4(mh[0], mh[1], mh[2], mh[3]);
    o[3] = make_uint4(mh[4], mh[5], mh[6], mh[7]);
}

// ---- allocation-free DAG checksum (see nx_dag_checksum_noalloc in include/nx_kawpow.h) ----
// Same weighted sum as checksum_partial_kernel, but the per-CTA partial sums live in a fixed array
// inside the library image instead of a buffer the context cudaMalloc'ed.  One CTA reduces one
// contiguous 4 MiB chunk, so the number of partials - and with it the largest input the fixed array
// can serve - is known up front.  Every GPU that loads the library keeps these 16.25 KiB until exit.
constexpr size_t NOALLOC_CHUNK_WORDS = (size_t)1 << 19;  // 2^19 u64 words = 4 MiB per CTA
constexpr int NOALLOC_MAX_PARTIALS = 2048;                // x 4 MiB = 8 GiB of input at most
constexpr int NOALLOC_RESULT_SLOT = NOALLOC_MAX_PARTIALS; // one more u64, 256-byte aligned, for the total
static __device__ __align__(256) uint64_t g_noalloc_scratch[NOALLOC_MAX_PARTIALS + 32];

__global__ void __launch_bounds__(CSUM_THREADS)
checksum_chunk_kernel(const uint64_t* __restrict__ words, size_t n, uint64_t* __restrict__ partials)
{
    __shared__ uint64_t red[CSUM_THREADS / 32];
    const size_t begin = (size_t)blockIdx.x * NOALLOC_CHUNK_WORDS;
    const size_t end = begin + NOALLOC_CHUNK_WORDS < n ? begin + NOALLOC_CHUNK_WORDS : n;
    uint64_t acc = 0;
    for (size_t i = begin + threadIdx.x; i < end; i += CSUM_THREADS) acc += __ldg(words + i) * (2 * (uint64_t)i + 1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t s = 0;
        for (int i = 0; i < CSUM_THREADS / 32; ++i) s += red[i];
        partials[blockIdx.x] = s;  // written unconditionally: no slot that is read later keeps an old value
    }
}

// Host code may not use g_noalloc_scratch as a pointer: on the host that name is only a handle for the
// device symbol.  The real device address is looked up once per device and kept here.
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
uint64_t* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};
}  // namespace

}  // namespace nx

extern "C" int nx_dag_checksum_noalloc(const void* data_device, size_t n_bytes, void* stream, uint64_t* out_host)
{
    using namespace nx;
    if (!data_device || !out_host) {
        set_error("nx_dag_checksum_noalloc: null argument");
        return NX_ERR_INVALID;
    }
    if (n_bytes % 8 != 0 || ((uintptr_t)data_device & 7) != 0) {
        set_error("nx_dag_checksum_noalloc: data must be 8-byte aligned and a whole number of 64-bit words");
        return NX_ERR_INVALID;
    }
    const size_t n_words = n_bytes / 8;
    const size_t chunks = (n_words + NOALLOC_CHUNK_WORDS - 1) / NOALLOC_CHUNK_WORDS;
    if (chunks > (size_t)NOALLOC_MAX_PARTIALS) {
        set_error("nx_dag_checksum_noalloc: %zu bytes exceed the fixed scratch (limit %llu bytes = %d chunks of 4 MiB); "
                  "use nx_dag_checksum for larger inputs",
                  n_bytes, (unsigned long long)NX_NOALLOC_MAX_BYTES, NOALLOC_MAX_PARTIALS);
        return NX_ERR_INVALID;
    }
    int dev = -1;
    NX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        set_error("nx_dag_checksum_noalloc: device index %d not supported", dev);
        return NX_ERR_INVALID;
    }
    cudaPointerAttributes attr;
    NX_CUDA(cudaPointerGetAttributes(&attr, data_device));
    if (attr.type != cudaMemoryTypeDevice || attr.device != dev) {
        set_error("nx_dag_checksum_noalloc: data is not device memory of the current device %d "
                  "(make the device that owns it current first)", dev);
        return NX_ERR_INVALID;
    }
    {
        // The attributes above describe the base pointer only.  Ask the driver for the extent of the allocation
        // around it (a query, nothing is allocated) and refuse a range that runs past its end: the kernel would
        // otherwise fault and leave the context with a sticky error.  Memory the driver cannot describe this way
        // (some mapped ranges) is let through unchecked.
        typedef int (*get_range_fn)(unsigned long long*, size_t*, unsigned long long);
        static get_range_fn get_range = [] {
            void* fn = nullptr;
            cudaDriverEntryPointQueryResult st;
            if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &st) != cudaSuccess ||
                st != cudaDriverEntryPointSuccess)
                fn = nullptr;
            cudaGetLastError();
            return reinterpret_cast<get_range_fn>(fn);
        }();
        unsigned long long base = 0;
        size_t size = 0;
        if (get_range && get_range(&base, &size, (unsigned long long)(uintptr_t)data_device) == 0 && size != 0) {
            const unsigned long long p = (unsigned long long)(uintptr_t)data_device;
            if (p < base || n_bytes > size || p - base > size - n_bytes) {
                set_error("nx_dag_checksum_noalloc: %zu bytes from this pointer run past the end of its allocation "
                          "(%zu bytes, the pointer sits at offset %llu)", n_bytes, size, p - base);
                return NX_ERR_INVALID;
            }
        }
    }
    uint64_t* scratch = nullptr;
    {
        // One-time set-up per device, with that device current: resolve the symbol to its device address.  The
        // address dies with the context (cudaDeviceReset): a cached one that the driver no longer knows as memory
        // of this device is looked up again.
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (g_noalloc_ptr[dev]) {
            cudaPointerAttributes sa;
            if (cudaPointerGetAttributes(&sa, g_noalloc_ptr[dev]) != cudaSuccess || sa.type != cudaMemoryTypeDevice ||
                sa.device != dev)
                g_noalloc_ptr[dev] = nullptr;
            cudaGetLastError();
        }
        if (!g_noalloc_ptr[dev]) {
            void* p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                set_error("cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
                cudaGetLastError();
                return NX_ERR_CUDA;
            }
            g_noalloc_ptr[dev] = static_cast<uint64_t*>(p);
        }
        scratch = g_noalloc_ptr[dev];
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (chunks == 0) {
        *out_host = 0;
        return NX_OK;
    }
    // Nothing in the array has to start at zero: partial[0..chunks) and the result slot are all
    // overwritten by this call before they are read, whatever the previous call left behind.
    checksum_chunk_kernel<<<(unsigned)chunks, CSUM_THREADS, 0, s>>>(static_cast<const uint64_t*>(data_device), n_words,
                                                                   scratch);
    NX_CUDA(cudaGetLastError());
    checksum_final_kernel<<<1, 32, 0, s>>>(scratch, (int)chunks, scratch + NOALLOC_RESULT_SLOT);
    NX_CUDA(cudaGetLastError());
    NX_CUDA(cudaMemcpyAsync(out_host, scratch + NOALLOC_RESULT_SLOT, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    NX_CUDA(cudaStreamSynchronize(s));
    return NX_OK;
}

namespace nx
{
int dag_generate(const void* light_dev, int light_items, uint32_t first_item256, uint32_t count_item256, void* out_dev,
                 cudaStream_t s)
{
    void* one[1] = {out_dev};
    return dag_generate_multi(light_dev, light_items, first_item256, count_item256, one, 1, s);
}

int dag_generate_multi(const void* light_dev, int light_items, uint32_t first_item256, uint32_t count_item256,
                       void* const* outs_dev, int n_outs, cudaStream_t s)
{
    if (count_item256 == 0) return NX_OK;
    if (n_outs < 1 || n_outs > NX_MAX_PEERS) {
        set_error("dag_generate: %d output buffers (1..%d supported)", n_outs, NX_MAX_PEERS);
        return NX_ERR_INVALID;
    }
    const uint64_t count512 = (uint64_t)count_item256 * 4;
    if ((uint64_t)first_item256 * 4 + count512 > 0xffffffffull) {
        set_error("dag_generate: item range exceeds 32-bit index space");
        return NX_ERR_INVALID;
    }
    FastMod fm;
    fm.d = (uint32_t)light_items;
    fm.m = (uint32_t)((1ull << 32) / fm.d);
    const uint32_t blocks = (uint32_t)((count512 + DAG_THREADS - 1) / DAG_THREADS);
    DagOuts outs;
    outs.n = n_outs;
    for (int i = 0; i < NX_MAX_PEERS; ++i) outs.ptr[i] = i < n_outs ? static_cast<uint4*>(outs_dev[i]) : nullptr;
    dag_items_kernel<<<blocks, DAG_THREADS, 0, s>>>(static_cast<const uint4*>(light_dev), fm, first_item256 * 4u,
                                                    (uint32_t)count512, outs);
    NX_CUDA(cudaGetLastError());
    return NX_OK;
}

int ethash_hash_batch(const void* dag_dev, uint32_t num_items_128, const uint8_t header[32], uint64_t start_nonce,
                      uint64_t count, void* out_dev, cudaStream_t s)
{
    if (count == 0) return NX_OK;
    uint64_t h[4];
    memcpy(h, header, 32);
    const uint64_t blocks = (count + ETHASH_THREADS - 1) / ETHASH_THREADS;
    if (blocks > 0x7fffffffull) {
        set_error("ethash batch too large");
        return NX_ERR_INVALID;
    }
    FastMod fm;
    fm.d = num_items_128;
    fm.m = (uint32_t)((1ull << 32) / fm.d);
    static const int q = [] {  // NX_ETHASH_INFLIGHT=4|8 (measurement knob)
        const char* e = std::getenv("NX_ETHASH_INFLIGHT");
        return e && std::atoi(e) == 8 ? 8 : 4;
    }();
    if (q == 8)
        ethash_hash_kernel<8><<<(unsigned)blocks, ETHASH_THREADS, 0, s>>>(static_cast<const uint4*>(dag_dev), fm, h[0], h[1],
                                                                         h[2], h[3], start_nonce, count, static_cast<uint4*>(out_dev));
    else
        ethash_hash_kernel<4><<<(unsigned)blocks, ETHASH_THREADS, 0, s>>>(static_cast<const uint4*>(dag_dev), fm, h[0], h[1],
                                                                         h[2], h[3], start_nonce, count, static_cast<uint4*>(out_dev));
    NX_CUDA(cudaGetLastError());
    return NX_OK;
}

int light_cache_build_gpu(void* light_dev, int light_items, const uint8_t seed[32], cudaStream_t s)
{
    uint64_t w[4];
    memcpy(w, seed, 32);
    // NX_LIGHT_CACHE_KERNEL=thread selects the first version (one thread does the whole permutation)
    const char* which = std::getenv("NX_LIGHT_CACHE_KERNEL");
    if (which && std::strcmp(which, "thread") == 0)
        light_cache_kernel<<<1, 32, 0, s>>>(static_cast<uint64_t*>(light_dev), (uint32_t)light_items, w[0], w[1], w[2], w[3]);
    else
        light_cache_warp_kernel<<<1, 32, 0, s>>>(static_cast<uint64_t*>(light_dev), (uint32_t)light_items, w[0], w[1], w[2], w[3]);
    NX_CUDA(cudaGetLastError());
    return NX_OK;
}

int dag_checksum(const void* dag_dev, size_t n_words64, uint64_t* partials_dev, int n_partials, uint64_t* out_dev,
                 cudaStream_t s)
{
    checksum_partial_kernel<<<n_partials, CSUM_THREADS, 0, s>>>(static_cast<const uint64_t*>(dag_dev), n_words64,
                                                                 partials_dev);
    NX_CUDA(cudaGetLastError());
    checksum_final_kernel<<<1, 32, 0, s>>>(partials_dev, n_partials, out_dev);
    NX_CUDA(cudaGetLastError());
    return NX_OK;
}

}  // namespace nx
