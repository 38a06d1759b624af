// KawPow (ProgPoW 0.9.4, Ravencoin parameters) search / batch-hash kernel for sm_100a.
//
// This text is NOT compiled into the library.  It is embedded as a string and compiled at run time
// by NVRTC, once per ProgPoW period, after the generator (nx_codegen.cpp) has prepended:
//     #define NX_DAG_ITEMS   <number of 256-byte DAG items of the epoch>
//     #define NX_ROUND_CACHE_MATH(m, L1)        the period's 11 cache + 18 math ops, straight-line
//     #define NX_ROUND_DAG(m, d0, d1, d2, d3)   the period's 4 DAG merges
// so that every src/dst/selector of the random program is a literal and the 32 mix words of a
// lane live in registers.  It also compiles as plain C++ (no __CUDACC__) so that the generated
// program can be checked on the CPU against the oracle (tests/test_codegen.py).
//
// Replaces: progpow::hash / search / verify, hash_mix, init_mix, round, random_math, random_merge
// (reference src/crypto/ethash/lib/ethash/progpow.cpp:92-144,179-355,431-495,553-579).
//
// Work decomposition (persistent CTAs: the grid is sized to the SM count and every WARP walks the job on its own):
//   * the CTA stages the L1 cache in shared memory once (the only CTA-wide barrier), then each warp takes
//     "units" of 2 x job.per_hw nonces - unit u, u + (warps in the grid), ... in ascending nonce order;
//   * per unit every thread absorbs its own nonce with keccak-f800 (25 words in registers);
//   * the 16 threads of a half-warp then act as the 16 ProgPoW lanes and walk through the half-warp's
//     per_hw (normally 16) nonces one after the other: the seed is broadcast by shuffle, each lane keeps its
//     32 mix words in registers, the 256-byte DAG item of a round is fetched as sixteen coalesced 16-byte
//     loads (one per lane), the 16 KiB L1 cache sits in shared memory;
//   * every thread finishes its own nonce with the second keccak-f800 and either stores {final, mix}
//     (batch mode) or compares against the boundary (search mode).
// Because warps are independent after the first barrier, the keccak phases (integer pipe only) of some warps
// overlap the round loops (shared-memory pipe) of others, a search stops at the first unit that lies above an
// already reported nonce (the reference returns at its first hit, progpow.cpp:572-577), and a small call
// (per_hw < 16) is spread over all SMs instead of costing a full 16-nonce walk.

typedef unsigned int u32;
typedef unsigned long long u64;

#ifdef __CUDACC__
#define NX_DEV __device__ __forceinline__
#else
#define NX_DEV static inline
#endif

#ifndef NX_THREADS
#define NX_THREADS 256
#endif
#ifndef NX_MIN_BLOCKS
#define NX_MIN_BLOCKS 4
#endif
// Tunables set by the run-time (nx_kawpow_rt.cpp) through -D:
//   NX_L1_COPIES  1|2|4|8  replicas of the 16 KiB L1 cache in shared memory.  Copy g is interleaved at a
//                 word stride of NX_L1_COPIES, so it only occupies banks {g, g+C, g+2C, ...}; the lanes of a
//                 warp are spread over the copies, which lowers the expected bank-conflict degree of the
//                 random 4-byte reads (3.53 -> 2.96 for 8 copies) at the same instruction count.
//   NX_ROT_IMAD   0..16    how many of every 16 constant rotates of the merge steps are done as a
//                 32x32->64 multiply (FMA pipe) instead of a funnel shift (ALU pipe).
//   NX_UNROLL     1|2|4    unroll factor of the 64-round loop.
#ifndef NX_L1_COPIES
#define NX_L1_COPIES 1
#endif
#ifndef NX_ROT_IMAD
#define NX_ROT_IMAD 0
#endif
#ifndef NX_DAG_HINT
#define NX_DAG_HINT 0  // cache hints on the DAG load (see nx_ld_dag)
#endif
#ifndef NX_L1_HI
#define NX_L1_HI 0
#endif
#ifndef NX_UNROLL
#define NX_UNROLL 1
#endif
#if NX_L1_COPIES == 1
#define NX_L1_LOG 0
#elif NX_L1_COPIES == 2
#define NX_L1_LOG 1
#elif NX_L1_COPIES == 4
#define NX_L1_LOG 2
#elif NX_L1_COPIES == 8
#define NX_L1_LOG 3
#else
#error "NX_L1_COPIES must be 1, 2, 4 or 8"
#endif

#define NX_FNV_PRIME 0x01000193u
#define NX_FNV_BASIS 0x811c9dc5u

// ---- scalar helpers used by the generated program (reference bit_manipulation.h:17-77) ----
#ifdef __CUDACC__
NX_DEV u32 nx_rotl(u32 x, u32 n) { return __funnelshift_l(x, x, n); }
NX_DEV u32 nx_rotr(u32 x, u32 n) { return __funnelshift_r(x, x, n); }
NX_DEV u32 nx_mulhi(u32 a, u32 b) { return __umulhi(a, b); }
NX_DEV u32 nx_min(u32 a, u32 b) { return min(a, b); }
NX_DEV u32 nx_clz(u32 a) { return (u32)__clz((int)a); }  // 32 for 0, as the reference's clz32
NX_DEV u32 nx_popc(u32 a) { return (u32)__popc(a); }
#else
NX_DEV u32 nx_rotl(u32 x, u32 n) { n &= 31; return n ? (x << n) | (x >> (32 - n)) : x; }
NX_DEV u32 nx_rotr(u32 x, u32 n) { n &= 31; return n ? (x >> n) | (x << (32 - n)) : x; }
NX_DEV u32 nx_mulhi(u32 a, u32 b) { return (u32)(((u64)a * b) >> 32); }
NX_DEV u32 nx_min(u32 a, u32 b) { return a < b ? a : b; }
NX_DEV u32 nx_clz(u32 a) { return a ? (u32)__builtin_clz(a) : 32u; }
NX_DEV u32 nx_popc(u32 a) { return (u32)__builtin_popcount(a); }
#endif
NX_DEV u32 nx_fnv1a(u32 h, u32 d) { return (h ^ d) * NX_FNV_PRIME; }

// rotate by a literal 1..31 (merge steps); k is the ordinal of the rotate within the period program.
// NX_ROT_IMAD = N (0..16) turns N out of every 16 of them, evenly spread, into a 32x32->64 multiply whose
// halves are OR-ed (x*2^n = (x<<n) low, (x>>(32-n)) high): that is FMA-pipe work (at two issue slots per
// wide multiply) taken off the busier ALU pipe.
#if defined(__CUDACC__) && NX_ROT_IMAD
NX_DEV u32 nx_rotl_c(u32 x, u32 n, u32 k)
{
    if (((k * NX_ROT_IMAD) >> 4) != (((k + 1u) * NX_ROT_IMAD) >> 4)) {
        u32 lo, hi;
        asm("{\n\t.reg .u64 p;\n\tmul.wide.u32 p, %2, %3;\n\tmov.b64 {%0, %1}, p;\n\t}" : "=r"(lo), "=r"(hi) : "r"(x), "r"(1u << n));
        return lo | hi;
    }
    return nx_rotl(x, n);
}
#else
NX_DEV u32 nx_rotl_c(u32 x, u32 n, u32 k) { (void)k; return nx_rotl(x, n); }
#endif
NX_DEV u32 nx_rotr_c(u32 x, u32 n, u32 k) { return nx_rotl_c(x, 32u - n, k); }

// ---- KISS99 (reference lib/ethash/kiss99.hpp:31-62) ----
struct NxKiss99
{
    u32 z, w, jsr, jcong;
};
NX_DEV u32 nx_kiss99(NxKiss99& s)
{
    s.z = 36969u * (s.z & 0xffffu) + (s.z >> 16);
    s.w = 18000u * (s.w & 0xffffu) + (s.w >> 16);
    s.jcong = 69069u * s.jcong + 1234567u;
    s.jsr ^= s.jsr << 17;
    s.jsr ^= s.jsr >> 13;
    s.jsr ^= s.jsr << 5;
    return (((s.z << 16) + s.w) ^ s.jcong) + s.jsr;
}

// init_mix for one lane (reference progpow.cpp:247-263)
NX_DEV void nx_fill_mix(u32 seed0, u32 seed1, u32 lane, u32 m[32])
{
    NxKiss99 g;
    g.z = nx_fnv1a(NX_FNV_BASIS, seed0);
    g.w = nx_fnv1a(g.z, seed1);
    g.jsr = nx_fnv1a(g.w, lane);
    g.jcong = nx_fnv1a(g.jsr, lane);
#pragma unroll
    for (int i = 0; i < 32; ++i) m[i] = nx_kiss99(g);
}

// per-lane FNV1a reduction of the 32 mix words (reference progpow.cpp:279-285)
NX_DEV u32 nx_lane_reduce(const u32 m[32])
{
    u32 h = NX_FNV_BASIS;
#pragma unroll
    for (int i = 0; i < 32; ++i) h = nx_fnv1a(h, m[i]);
    return h;
}

// ---- Keccak-f[800] (reference lib/keccak/keccakf800.c:38-253); lanes stay in registers ----
#ifdef __CUDACC__
__constant__ u32 NX_RC800[22] = {
#else
static const u32 NX_RC800[22] = {
#endif
    0x00000001u, 0x00008082u, 0x0000808Au, 0x80008000u, 0x0000808Bu, 0x80000001u, 0x80008081u, 0x00008009u,
    0x0000008Au, 0x00000088u, 0x80008009u, 0x8000000Au, 0x8000808Bu, 0x0000008Bu, 0x00008089u, 0x00008003u,
    0x00008002u, 0x00000080u, 0x0000800Au, 0x8000000Au, 0x80008081u, 0x00008080u};

#define NX_CHI(a, b, c) ((a) ^ (~(b) & (c)))
NX_DEV void nxk_f800(u32 a[25])
{
#ifdef __CUDACC__
#pragma unroll 1
#endif
    for (int r = 0; r < 22; ++r) {
        const u32 c0 = a[0] ^ a[5] ^ a[10] ^ a[15] ^ a[20], c1 = a[1] ^ a[6] ^ a[11] ^ a[16] ^ a[21],
                  c2 = a[2] ^ a[7] ^ a[12] ^ a[17] ^ a[22], c3 = a[3] ^ a[8] ^ a[13] ^ a[18] ^ a[23],
                  c4 = a[4] ^ a[9] ^ a[14] ^ a[19] ^ a[24];
        const u32 d0 = c4 ^ nx_rotl(c1, 1), d1 = c0 ^ nx_rotl(c2, 1), d2 = c1 ^ nx_rotl(c3, 1),
                  d3 = c2 ^ nx_rotl(c4, 1), d4 = c3 ^ nx_rotl(c0, 1);
        // theta+rho+pi; rho offsets are the Keccak ones modulo 32
        const u32 b0 = a[0] ^ d0, b10 = nx_rotl(a[1] ^ d1, 1), b20 = nx_rotl(a[2] ^ d2, 30),
                  b5 = nx_rotl(a[3] ^ d3, 28), b15 = nx_rotl(a[4] ^ d4, 27), b16 = nx_rotl(a[5] ^ d0, 4),
                  b1 = nx_rotl(a[6] ^ d1, 12), b11 = nx_rotl(a[7] ^ d2, 6), b21 = nx_rotl(a[8] ^ d3, 23),
                  b6 = nx_rotl(a[9] ^ d4, 20), b7 = nx_rotl(a[10] ^ d0, 3), b17 = nx_rotl(a[11] ^ d1, 10),
                  b2 = nx_rotl(a[12] ^ d2, 11), b12 = nx_rotl(a[13] ^ d3, 25), b22 = nx_rotl(a[14] ^ d4, 7),
                  b23 = nx_rotl(a[15] ^ d0, 9), b8 = nx_rotl(a[16] ^ d1, 13), b18 = nx_rotl(a[17] ^ d2, 15),
                  b3 = nx_rotl(a[18] ^ d3, 21), b13 = nx_rotl(a[19] ^ d4, 8), b14 = nx_rotl(a[20] ^ d0, 18),
                  b24 = nx_rotl(a[21] ^ d1, 2), b9 = nx_rotl(a[22] ^ d2, 29), b19 = nx_rotl(a[23] ^ d3, 24),
                  b4 = nx_rotl(a[24] ^ d4, 14);
        a[0] = NX_CHI(b0, b1, b2) ^ NX_RC800[r];
        a[1] = NX_CHI(b1, b2, b3);
        a[2] = NX_CHI(b2, b3, b4);
        a[3] = NX_CHI(b3, b4, b0);
        a[4] = NX_CHI(b4, b0, b1);
        a[5] = NX_CHI(b5, b6, b7);
        a[6] = NX_CHI(b6, b7, b8);
        a[7] = NX_CHI(b7, b8, b9);
        a[8] = NX_CHI(b8, b9, b5);
        a[9] = NX_CHI(b9, b5, b6);
        a[10] = NX_CHI(b10, b11, b12);
        a[11] = NX_CHI(b11, b12, b13);
        a[12] = NX_CHI(b12, b13, b14);
        a[13] = NX_CHI(b13, b14, b10);
        a[14] = NX_CHI(b14, b10, b11);
        a[15] = NX_CHI(b15, b16, b17);
        a[16] = NX_CHI(b16, b17, b18);
        a[17] = NX_CHI(b17, b18, b19);
        a[18] = NX_CHI(b18, b19, b15);
        a[19] = NX_CHI(b19, b15, b16);
        a[20] = NX_CHI(b20, b21, b22);
        a[21] = NX_CHI(b21, b22, b23);
        a[22] = NX_CHI(b22, b23, b24);
        a[23] = NX_CHI(b23, b24, b20);
        a[24] = NX_CHI(b24, b20, b21);
    }
}

// "rAVENCOINKAWPOW": words 10..24 of the first absorb, words 16..24 of the second
// (reference progpow.cpp:157-173,317-319,343-345)
#define NX_PAD_FIRST(st)                                                                                    \
    st[10] = 0x72u; st[11] = 0x41u; st[12] = 0x56u; st[13] = 0x45u; st[14] = 0x4Eu; st[15] = 0x43u;       \
    st[16] = 0x4Fu; st[17] = 0x49u; st[18] = 0x4Eu; st[19] = 0x4Bu; st[20] = 0x41u; st[21] = 0x57u;       \
    st[22] = 0x50u; st[23] = 0x4Fu; st[24] = 0x57u
#define NX_PAD_LAST(st)                                                                                     \
    st[16] = 0x72u; st[17] = 0x41u; st[18] = 0x56u; st[19] = 0x45u; st[20] = 0x4Eu; st[21] = 0x43u;       \
    st[22] = 0x4Fu; st[23] = 0x49u; st[24] = 0x4Eu

#if defined(__CUDACC__) && !defined(NX_HELPERS_ONLY)
// ------------------------------------------------------------------------------------------ device

struct NxJob
{
    u32 header[8];       // header_hash as 8 little-endian words
    u32 boundary_be[8];  // boundary words, byte-swapped: compared most-significant first
    u64 start_nonce;
    u64 count;           // nonces that may be reported; the grid is rounded up to whole CTAs
    u64 report_base;     // search mode reports report_base + (nonce - start_nonce)
    u32 mode;            // 0: store {final, mix} per nonce;  1: search (compare on device, atomicMin the offset)
    u32 l1_mul_lo;       // 2^20                      } NX_L1_HI == 2: L1 addressing multipliers handed over as kernel
    u32 l1_mul_hi;       // 2^(14 + log2 L1 replicas) } parameters, so that ptxas cannot fold the multiplies into ALU shifts
    u32 per_hw;          // nonces a half-warp walks per unit, 1..16 (16 unless the call is too small to fill the GPU)
    u32 n_units;         // ceil(count / (2 * per_hw)); unit u covers offsets [u * 2 * per_hw, (u + 1) * 2 * per_hw)
    u32 pad;
};

NX_DEV u32 nx_bswap(u32 x) { return __byte_perm(x, 0, 0x0123); }

// one L1-cache word.  Address = lane_base + (x & 4095) * 4 * NX_L1_COPIES.
//   default:    AND (ALU pipe) + multiply-add (FMA pipe) + LDS
//   NX_L1_HI:   x*2^20 keeps only the low 12 bits of x; the high half of (that * 2^(14+log2 copies)) is the
//               scaled index, so both steps run on the FMA pipe and the ALU pipe (the busier one) is spared.
NX_DEV u32 nx_l1_load(u32 x, u32 lane_base, u32 mul_lo, u32 mul_hi)
{
    u32 v;
#if NX_L1_HI == 2
    // both multipliers are kernel parameters (constant-bank operands): two FMA-pipe multiplies, no ALU-pipe work
    asm("{\n\t.reg .u32 t;\n\tmul.lo.u32 t, %1, %2;\n\tmad.hi.u32 t, t, %3, %4;\n\tld.shared.u32 %0, [t];\n\t}"
        : "=r"(v)
        : "r"(x), "r"(mul_lo), "r"(mul_hi), "r"(lane_base));
#elif NX_L1_HI
    asm("{\n\t.reg .u32 t;\n\tmul.lo.u32 t, %1, 1048576;\n\tmad.hi.u32 t, t, %2, %3;\n\tld.shared.u32 %0, [t];\n\t}"
        : "=r"(v)
        : "r"(x), "n"(1 << (14 + NX_L1_LOG)), "r"(lane_base));
#else
    asm("{\n\t.reg .u32 t;\n\tand.b32 t, %1, 4095;\n\tmad.lo.u32 t, t, %2, %3;\n\tld.shared.u32 %0, [t];\n\t}"
        : "=r"(v)
        : "r"(x), "n"(4 * NX_L1_COPIES), "r"(lane_base));
#endif
    return v;
}

NX_DEV uint4 nx_ld_dag(const uint4* p)
{
    uint4 v;
#if NX_DAG_HINT == 1  // ask L2 to bring in the whole 256-byte item on the first sector that misses
    asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
#elif NX_DAG_HINT == 2  // plain cached load
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
#elif NX_DAG_HINT == 3  // evict-first in L1, 128-byte L2 granularity
    asm volatile("ld.global.nc.L1::evict_first.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
#else
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
#endif
    return v;
}

// One kernel for both uses, so that a period costs ONE NVRTC compilation: job.mode 0 stores {final[8], mix[8]} per
// nonce into `result` (u32[16] each), job.mode 1 atomicMins the offset of qualifying nonces into `result` (one u64).
// The mode only decides what happens to the finished hash - a uniform branch after the last keccak.
//
// Measurement-only shapes (NX_KAWPOW_TUNE nodag=1 / nomath=1; results are NOT KawPow hashes):
//   NX_NODAG   the round's DAG words are derived from the item index in registers instead of being loaded: index
//              chain, L1 reads, math, merges, shuffles and both keccaks unchanged -> the on-chip ceiling of the kernel;
//   NX_NOMATH  the 11 cache reads and 18 math ops are dropped, the index chain, the sixteen 16-byte loads per round and
//              the four DAG merges stay -> what the dependent 256-byte reads alone achieve at this occupancy.
extern "C" __global__ void __launch_bounds__(NX_THREADS, NX_MIN_BLOCKS)
nx_kawpow_kernel(const uint4* __restrict__ dag, const NxJob job, void* __restrict__ result)
{
    // dynamic shared memory: [L1 cache, NX_L1_COPIES interleaved replicas][carry 6 x T][digest 8 x T]
    extern __shared__ __align__(16) u32 nx_smem[];
    u32* const l1s = nx_smem;
    u32(*const carry_s)[NX_THREADS] = reinterpret_cast<u32(*)[NX_THREADS]>(nx_smem + 4096 * NX_L1_COPIES);
    u32(*const digest_s)[NX_THREADS] = reinterpret_cast<u32(*)[NX_THREADS]>(nx_smem + 4096 * NX_L1_COPIES + 6 * NX_THREADS);

    const u32 tid = threadIdx.x;
    // this lane's replica: word g of every interleaved group of NX_L1_COPIES words
    u32 l1_rep = (tid & 31u) >> (5 - NX_L1_LOG);
#if NX_L1_COPIES > 1
    asm volatile("" : "+r"(l1_rep));  // keep it a plain register instead of re-deriving it from tid at each use
#endif
    // 32-bit shared-window address of word 0 of this lane's replica
    const u32 l1_base = (u32)__cvta_generic_to_shared(l1s) + l1_rep * 4u;
    const u32 lane = tid & 15u;

    // L1 cache = first 16 KiB of the DAG (reference ethash.cpp:168-170)
#if NX_L1_COPIES == 1
    for (u32 i = tid; i < 1024u; i += NX_THREADS) reinterpret_cast<uint4*>(l1s)[i] = __ldg(dag + i);
#else
    for (u32 i = tid; i < 4096u; i += NX_THREADS) {
        const u32 v = __ldg(reinterpret_cast<const u32*>(dag) + i);
#pragma unroll
        for (int g = 0; g < NX_L1_COPIES; ++g) l1s[i * NX_L1_COPIES + g] = v;
    }
#endif
    __syncthreads();  // the only CTA-wide barrier: carry_s / digest_s slots are private to their thread

    const u32 per_hw = job.per_hw;
    const u32 unit_nonces = 2u * per_hw;
    // The loop counter is built from launch-wide values only, so that the compiler can see that a warp never
    // diverges here (a counter derived from threadIdx would cost a convergence check before every shuffle of the round
    // loop).  A warp whose unit lies past the end still walks it - its results are dropped - and only in the last stripe.
#pragma unroll 1
    for (u32 stripe = blockIdx.x * (NX_THREADS / 32u); stripe < job.n_units; stripe += gridDim.x * (NX_THREADS / 32u)) {
        const u32 unit = stripe + (tid >> 5);
        if (job.mode != 0u) {
            // units ascend in nonce order within a warp: once a nonce below this unit has been reported, neither this
            // unit nor any later one of this warp can lower the minimum
            const u64 best = *static_cast<volatile const u64*>(result);
            if (__any_sync(0xffffffffu, best < job.report_base + (u64)unit * unit_nonces)) break;
        }
        u32 seed0, seed1;
        {
            // offset of this thread's nonce; lanes >= per_hw of a half-warp hash a neighbour's nonce and drop the result
            const u64 nonce = job.start_nonce + (u64)unit * unit_nonces + ((tid >> 4) & 1u) * per_hw + lane;
            u32 st[25];
#pragma unroll
            for (int i = 0; i < 8; ++i) st[i] = job.header[i];
            st[8] = (u32)nonce;
            st[9] = (u32)(nonce >> 32);
            NX_PAD_FIRST(st);
            nxk_f800(st);
            seed0 = st[0];
            seed1 = st[1];
#pragma unroll
            for (int i = 0; i < 6; ++i) carry_s[i][tid] = st[2 + i];
        }

#pragma unroll 1
        for (u32 h = 0; h < per_hw; ++h) {
            u32 m[32];
            nx_fill_mix(__shfl_sync(0xffffffffu, seed0, h, 16), __shfl_sync(0xffffffffu, seed1, h, 16), lane, m);

#pragma unroll NX_UNROLL
            for (u32 r = 0; r < 64u; ++r) {
                // item index comes from lane r%16 (reference progpow.cpp:182-183)
                const u32 item = __shfl_sync(0xffffffffu, m[0], r & 15u, 16) % NX_DAG_ITEMS;
#ifdef NX_NODAG
                const u32 di = item * 16u + ((lane ^ r) & 15u);
                const uint4 d = make_uint4(di, di * 0x9E3779B1u, ~di, di + 0x7F4A7C15u);
#else
                const uint4 d = nx_ld_dag(dag + ((u64)item * 16u + ((lane ^ r) & 15u)));
#endif
#ifndef NX_NOMATH
#define NX_L1(x) nx_l1_load((x), l1_base, job.l1_mul_lo, job.l1_mul_hi)
                NX_ROUND_CACHE_MATH(m, NX_L1);
#undef NX_L1
#endif
                NX_ROUND_DAG(m, d.x, d.y, d.z, d.w);
            }

            // 16 lane hashes -> 8 words (reference progpow.cpp:288-293)
            const u32 lh = nx_lane_reduce(m);
            u32 dg[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                dg[i] = nx_fnv1a(nx_fnv1a(NX_FNV_BASIS, __shfl_sync(0xffffffffu, lh, i, 16)),
                                 __shfl_sync(0xffffffffu, lh, i + 8, 16));
            if (h == lane) {
#pragma unroll
                for (int i = 0; i < 8; ++i) digest_s[i][tid] = dg[i];
            }
        }
        // digest_s[.][tid] was written by this very thread: no barrier needed

        u32 st[25];
        st[0] = seed0;
        st[1] = seed1;
#pragma unroll
        for (int i = 0; i < 6; ++i) st[2 + i] = carry_s[i][tid];
#pragma unroll
        for (int i = 0; i < 8; ++i) st[8 + i] = digest_s[i][tid];
        NX_PAD_LAST(st);
        nxk_f800(st);

        const u64 gid = (u64)unit * unit_nonces + ((tid >> 4) & 1u) * per_hw + lane;
        if (lane < per_hw && gid < job.count) {  // also false for a unit past the end (gid >= count)
            if (job.mode == 0u) {
                uint4* o = reinterpret_cast<uint4*>(static_cast<u32*>(result) + gid * 16u);
                o[0] = make_uint4(st[0], st[1], st[2], st[3]);
                o[1] = make_uint4(st[4], st[5], st[6], st[7]);
                o[2] = make_uint4(digest_s[0][tid], digest_s[1][tid], digest_s[2][tid], digest_s[3][tid]);
                o[3] = make_uint4(digest_s[4][tid], digest_s[5][tid], digest_s[6][tid], digest_s[7][tid]);
            } else {
                // final <= boundary as big-endian byte strings (reference ethash-internal.hpp:32-43)
                bool ok = true;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const u32 a = nx_bswap(st[i]), b = job.boundary_be[i];
                    if (a != b) {
                        ok = a < b;
                        break;
                    }
                }
                if (ok) atomicMin(static_cast<u64*>(result), job.report_base + gid);
            }
        }
    }
}
#endif  // __CUDACC__ && !NX_HELPERS_ONLY
