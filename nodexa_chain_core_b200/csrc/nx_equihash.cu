// GATED "next" row: a parameterised Equihash engine for sm_100a.
//
// *** The reference node contains NO Equihash: no solver, no validator, no (n,k), no personalization, no header
// *** field, no test vector (SURVEY.md §0 Fact 1).  Nothing here can be checked against it; (n,k) and the 8-byte
// *** personalization tag are therefore INPUTS, and correctness is established against an independent CPU
// *** implementation of the published algorithm (oracle/nx_equihash_oracle.c) - self-consistency, parity unpinned.
//
// Algorithm (Biryukov-Khovratovich Equihash, Wagner's generalised birthday): N = 2^(n/(k+1)+1) strings of n bits
// from BLAKE2b(personal = tag||LE32(n)||LE32(k); header||LE32(i / (512/n))), k collision rounds of cbl = n/(k+1)
// bits, the last one cancelling two chunks at once; a solution is 2^k distinct indices whose tree satisfies
// "left subtree's first index < right subtree's first index".
//
// GPU mapping: rows LIVE in buckets.  Bucket b of round r holds the rows whose chunk r has top bits b
// (cbl-8 of them, so a bucket's mean load is N / 2^(cbl-8) = 512 rows whatever (n,k) is); a row id is
// bucket*slots + slot.  A round is ONE kernel, one CTA per bucket: the bucket's rows are copied into shared
// memory (contiguous, 16-byte loads) and chained by the remaining 8 bits of the chunk on the way in, and every
// colliding pair is XOR-ed and appended straight to the bucket its NEXT chunk selects (one atomic slot counter per
// bucket, each in a 32-byte sector of its own because the L2 serialises atomics per line) as whole 16-byte stores,
// together with a two-id back-reference.  The generator does the same with the BLAKE2b strings, so there is no
// separate scatter pass.  Candidates of the last round are
// expanded, canonically ordered and checked for distinct indices by one CTA each; an independent validator
// kernel recomputes the BLAKE2b strings of a solution and re-checks every rule.
#include "../../include/nx_equihash.h"
#include "nx_blake2b.cuh"
#include "nx_internal.hpp"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>

namespace nx
{
namespace
{
constexpr int EQ_MAX_CHUNKS = 12;      // k+1
constexpr int EQ_COLLIDE_THREADS = 256;
constexpr int EQ_MAX_CANDIDATES = 8192;
constexpr uint32_t EQ_RECOVER_FIRST = 64;  // candidates expanded (and copied to the host) before their count is known
// Bucket slot counters sit EQ_CNT_STRIDE words apart.  Every new row reserves its slot with an atomicAdd on its bucket's
// counter, 2^21 of them per round; the L2 serialises atomics that hit one line, so the stride decides how many of them run
// side by side.  NX_EQ_CNT_STRIDE (1, 8 or 32 words; read at create time) is kept as a measurement knob.
#ifndef EQ_CNT_STRIDE_DEFAULT
#define EQ_CNT_STRIDE_DEFAULT 8
#endif

// Rows are stored at a stride rounded up to a multiple of 4 words: a row is then read and written as whole 16-byte
// accesses (a new row goes to a random bucket, so every store of it opens a new 32-byte sector: 3 x 16 B instead of
// 9 x 4 B is a third of the L2 write operations).
__host__ __device__ __forceinline__ uint32_t eq_stride(uint32_t live_words) { return (live_words + 3u) & ~3u; }
// Shared-memory stride of a staged row (its S-1 words behind the sort key): rounded up to an even number of words so
// that a row is read as 8- or 16-byte accesses - except for the widest rows, where the extra word per row would cost a
// resident CTA (round 0 of (200,9): 9 words x 1024 slots is what lets five CTAs share an SM).
__host__ __device__ constexpr uint32_t eq_smem_stride(int live_words)
{
    return live_words >= 10 ? (uint32_t)(live_words - 1) : (uint32_t)(live_words & ~1);  // S-1, or S-1 rounded up to even
}

struct EqParams
{
    int n, k, cbl, chunks, ipp /* strings per digest */, hash_bytes, digest_bytes;
    uint32_t N;
    int bucket_bits;      // cbl - 8 (0 when cbl == 8)
    uint32_t num_buckets;
    uint32_t slots;       // row slots per bucket: 1024 (mean load 512); 2048 for tiny N, whose single bucket swings widely
    uint32_t cnt_stride;  // words between the slot counters of neighbouring buckets
};

// Word-level BLAKE2b tail: the host has absorbed the header up to the last (partial) block; a thread appends its
// 32-bit string counter at byte offset `fill` of that block and finalises.  No byte arrays, no local memory.
struct EqMid
{
    uint64_t h[8];
    uint64_t m[16];   // pending block as little-endian words, zero padded
    uint64_t t;       // bytes absorbed before the pending block
    uint32_t fill;    // bytes used in the pending block (multiple of 4, <= 124)
};

__device__ __forceinline__ void eq_digest_words(const EqMid& mid, uint32_t g, uint64_t out[8])
{
    uint64_t m[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) m[i] = mid.m[i];
    const uint32_t w = mid.fill >> 3, sh = (mid.fill & 4u) * 8u;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if ((uint32_t)i == w) m[i] |= (uint64_t)g << sh;
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = mid.h[i];
    b2_compress(out, m, mid.t + mid.fill + 4, true);
}

// chunk c of the string that starts at byte `byte0` of a digest held as little-endian words.  The string is a
// big-endian bit string: build the 64-bit big-endian window that contains the chunk (cbl <= 24 bits, so 5 bytes
// always cover it) and shift.
__device__ __forceinline__ uint32_t eq_chunk_words(const uint64_t d[8], int byte0, int cbl, int c)
{
    const int bit0 = byte0 * 8 + c * cbl;
    const int first = bit0 >> 3;  // first byte of the window
    uint64_t win = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const int by = first + i;
        const uint64_t byte = by < 64 ? (d[by >> 3] >> (8 * (by & 7))) & 0xffull : 0ull;
        win = (win << 8) | byte;
    }
    // window = 40 bits, chunk starts (bit0 & 7) bits into it
    return (uint32_t)((win >> (40 - (bit0 & 7) - cbl)) & ((1ull << cbl) - 1ull));
}

// chunks of initial string i (used by the validator)
__device__ void eq_string_chunks(const EqParams& p, const EqMid& mid, uint32_t i, uint32_t* out)
{
    uint64_t d[8];
    eq_digest_words(mid, i / (uint32_t)p.ipp, d);
    const int byte0 = (int)(i % (uint32_t)p.ipp) * p.hash_bytes;
    for (int c = 0; c < p.chunks; ++c) out[c] = eq_chunk_words(d, byte0, p.cbl, c);
}

// ---- round 0: N strings, each appended to the bucket its chunk 0 selects ----
// NN, KK > 0: the parameter set is known at compile time, so every byte position of every chunk is a literal and the
// digest and the chunks stay in registers (the generic form indexes the digest dynamically: 112 bytes of local memory
// per thread and twice the time).  NN = 0: any (n, k) the engine accepts.
template <int NN, int KK>
__global__ void eq_generate_kernel(EqParams p, EqMid mid, uint32_t* __restrict__ rows, uint32_t* __restrict__ leaf,
                                   uint32_t* __restrict__ bucket_cnt, uint32_t* __restrict__ dropped)
{
    const int chunks = NN ? KK + 1 : p.chunks;
    const int cbl = NN ? NN / (KK + 1) : p.cbl;
    const int ipp = NN ? 512 / NN : p.ipp;
    const int hash_bytes = NN ? NN / 8 : p.hash_bytes;
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;  // one digest per thread
    if (g >= (p.N + (uint32_t)ipp - 1) / (uint32_t)ipp) return;  // N need not be a multiple of 512/n
    uint64_t d[8];
    eq_digest_words(mid, g, d);
    const uint32_t st = eq_stride((uint32_t)chunks);  // live words at a stride of whole 16-byte stores
    // specialised: 512/NN iterations, unrolled (every position a literal); generic: ipp iterations (up to 32 for n = 16)
#pragma unroll(NN ? 512 / NN : 1)
    for (int j = 0; j < ipp; ++j) {
        const uint32_t i = g * (uint32_t)ipp + (uint32_t)j;
        if (i >= p.N) break;
        uint32_t ch[EQ_MAX_CHUNKS];
#pragma unroll
        for (int c = 0; c < EQ_MAX_CHUNKS; ++c) ch[c] = c < chunks ? eq_chunk_words(d, j * hash_bytes, cbl, c) : 0u;
        const uint32_t b = p.bucket_bits ? (ch[0] >> 8) : 0u;
        const uint32_t slot = atomicAdd(&bucket_cnt[b * p.cnt_stride], 1u);
        if (slot >= p.slots) {
            atomicAdd(dropped, 1u);
            continue;
        }
        const uint32_t id = b * p.slots + slot;
        uint4* dst = reinterpret_cast<uint4*>(rows + (size_t)id * st);
#pragma unroll
        for (int q = 0; q < EQ_MAX_CHUNKS / 4; ++q)
            if ((uint32_t)q * 4u < st) dst[q] = make_uint4(ch[4 * q], ch[4 * q + 1], ch[4 * q + 2], ch[4 * q + 3]);
        leaf[id] = i;
    }
}

typedef void (*EqGenerateFn)(EqParams, EqMid, uint32_t*, uint32_t*, uint32_t*, uint32_t*);
EqGenerateFn eq_generate_fn(int n, int k)
{
    if (n == 200 && k == 9) return eq_generate_kernel<200, 9>;
    if (n == 144 && k == 5) return eq_generate_kernel<144, 5>;
    if (n == 192 && k == 7) return eq_generate_kernel<192, 7>;
    if (n == 96 && k == 5) return eq_generate_kernel<96, 5>;
    return eq_generate_kernel<0, 0>;
}

// one staged row out of shared memory, in the widest accesses its stride allows
template <uint32_t SP>
__device__ __forceinline__ void eq_load_row(const uint32_t* r, uint32_t out[SP])
{
    if constexpr (SP % 4 == 0) {
#pragma unroll
        for (uint32_t q = 0; q < SP / 4; ++q) {
            const uint4 v = reinterpret_cast<const uint4*>(r)[q];
            out[4 * q] = v.x, out[4 * q + 1] = v.y, out[4 * q + 2] = v.z, out[4 * q + 3] = v.w;
        }
    } else if constexpr (SP % 2 == 0) {
#pragma unroll
        for (uint32_t q = 0; q < SP / 2; ++q) {
            const uint2 v = reinterpret_cast<const uint2*>(r)[q];
            out[2 * q] = v.x, out[2 * q + 1] = v.y;
        }
    } else {
#pragma unroll
        for (uint32_t q = 0; q < SP; ++q) out[q] = r[q];
    }
}

// Resident CTAs per SM the collide kernel is compiled for: 5 is what the shared memory of the widest rows allows
// (9 words x 1024 slots); rows of 7 words and fewer leave room for a sixth if the kernel keeps to 40 registers
// (measured at (200,9): 78 -> 70, 76 -> 68, 66 -> 62, 62 -> 59 us for 7, 6, 5, 4 words).
#ifndef EQ_COLLIDE_BLOCKS_NARROW
#define EQ_COLLIDE_BLOCKS_NARROW 6
#endif
// Measured alternatives: six for 8-word rows is slower (89 vs 76 us), eight for rows of 5 words and fewer (32 registers)
// gains nothing (74 vs 62 us at 5 words).
#define EQ_COLLIDE_BLOCKS(S) ((S) >= 8 ? 5 : EQ_COLLIDE_BLOCKS_NARROW)

// ---- one round: one CTA per bucket ----
// Rows of round r carry only the chunks that are still alive, r..k: S = chunks - r live words (template parameter, so
// that a row lives in registers while it is built), stored in global memory at stride eq_stride(S).
// Rows that collide on chunk r share its low 8 bits inside a bucket, so while the bucket is staged every row is pushed
// onto the chain of its key (one shared-memory atomicExch per row: head[key] <- row, next[row] <- old head); a row
// then pairs with everything behind it in its chain.  (Round 1 counting-sorted the bucket instead: two shared-memory
// atomics per row, a 256-bin scan and three more barriers - the atomics alone were a third of a round.)
// dynamic shared memory: [the bucket's rows without their sort key: slots x eq_smem_stride(S) words][next: slots x u16]
template <int S>
__global__ void __launch_bounds__(EQ_COLLIDE_THREADS, EQ_COLLIDE_BLOCKS(S))
eq_collide_kernel(EqParams p, int r, const uint32_t* __restrict__ rows, uint32_t* __restrict__ bucket_cnt,
                  uint32_t* __restrict__ next_rows, uint32_t* __restrict__ next_cnt, uint2* __restrict__ tree,
                  const uint2* __restrict__ prev_tree, uint32_t* __restrict__ cand_count)
{
    extern __shared__ __align__(16) uint32_t eq_smem[];
    __shared__ uint32_t head[256];
    constexpr uint32_t NIL = 0xffffu;
    constexpr uint32_t ST = (S + 3) & ~3;         // stride of the rows read
    constexpr uint32_t SO = S - 1;                // live words of the rows written (word 0 = chunk r + 1)
    constexpr uint32_t STO = (SO + 3) & ~3;       // their stride
    constexpr uint32_t SP = eq_smem_stride(S);    // stride of a row in shared memory (even where that is free: 8/16-byte reads)
    // shared memory keeps only what a pair needs: the SO words behind the sort key (word 0 picks the chain and is done)
    uint32_t* const rs = eq_smem;
    uint16_t* const next = reinterpret_cast<uint16_t*>(eq_smem + (size_t)p.slots * SP);
    const uint32_t b = blockIdx.x, t = threadIdx.x;
    const uint32_t base = b * p.slots;
    constexpr uint32_t QPR = ST / 4;  // 16-byte words per row
    constexpr uint32_t U = 4;         // row loads in flight per thread
    const uint4* const src = reinterpret_cast<const uint4*>(rows + (size_t)base * ST);
    // The first round of row loads is issued together with the load of the bucket's fill count instead of behind it:
    // the slots of a bucket are all allocated, so reading past the fill is harmless, and what lies there is not used.
    // Only for the wide rows: the narrow ones are compiled for 40 registers, where four more live 16-byte values spill
    // (measured: 108 -> 104 and 84 -> 80 us for 10 and 9 words, but 68 -> 74 us for 7).
    constexpr bool EARLY = S >= 8;
    uint4 first[U];
#pragma unroll
    for (uint32_t u = 0; u < U; ++u) {
        const uint32_t i = t + u * EQ_COLLIDE_THREADS;
        first[u] = EARLY && i < p.slots * QPR ? src[i] : make_uint4(0u, 0u, 0u, 0u);
    }
    const uint32_t cnt = min(bucket_cnt[b * p.cnt_stride], p.slots);
    head[t] = NIL;
    // This CTA is the only reader of its bucket's counter, and the array it sits in is the one the NEXT round appends
    // to: leave it at zero, so that no memset is needed between rounds (or between solves).
    __syncthreads();
    if (t == 0) bucket_cnt[b * p.cnt_stride] = 0u;
    const bool last = r == p.k - 1;
    if (cnt == 0) return;

    {   // the bucket's rows: one contiguous region, read as 16-byte words, four loads in flight per thread before the
        // first is used (the copy is a chain of global-memory latencies)
        for (uint32_t i0 = t; i0 < cnt * QPR; i0 += EQ_COLLIDE_THREADS * U) {
            uint4 v[U];
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                const uint32_t i = i0 + u * EQ_COLLIDE_THREADS;
                v[u] = EARLY && i0 == t ? first[u] : (i < cnt * QPR ? src[i] : make_uint4(0u, 0u, 0u, 0u));
            }
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                const uint32_t i = i0 + u * EQ_COLLIDE_THREADS;
                if (i >= cnt * QPR) break;
                const uint32_t row = i / QPR, w0 = (i % QPR) * 4u;
                const uint32_t vv[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                for (uint32_t q = 0; q < 4; ++q) {
                    const uint32_t w = w0 + q;
                    if (w == 0) {
                        // the low 8 bits of chunk r: push the row onto that key's chain
                        next[row] = (uint16_t)atomicExch(&head[vv[q] & 0xffu], row);
                    } else if (w < (uint32_t)S) {
                        rs[row * SP + (w - 1)] = vv[q];
                    }
                }
            }
        }
    }
    __syncthreads();
    // Row i pairs with every row behind it in its chain.  A thread owns rows t, t+256, ... (EQ_CURSORS of them at a
    // time; a bucket holds about 512 rows, so two cover it) and advances them in lock step: per step it XORs the pair of
    // every live cursor into registers and issues the slot-reserving atomic, and only then uses the returned slots - so
    // the atomics of a thread are in flight together (their latency is what this kernel waits on) and a pair's rows are
    // read from shared memory once.
    constexpr int EQ_CURSORS = 2;
    for (uint32_t i0 = t; i0 < cnt; i0 += EQ_COLLIDE_THREADS * EQ_CURSORS) {
        uint32_t sa[EQ_CURSORS], j[EQ_CURSORS];
        bool live = false;
#pragma unroll
        for (int c = 0; c < EQ_CURSORS; ++c) {
            const uint32_t i = i0 + (uint32_t)c * EQ_COLLIDE_THREADS;
            sa[c] = i < cnt ? i : 0u;
            j[c] = i < cnt ? (uint32_t)next[i] : NIL;
            live |= j[c] != NIL;
        }
        while (live) {
            uint32_t slot[EQ_CURSORS], nb[EQ_CURSORS], sb[EQ_CURSORS];
            uint32_t o[EQ_CURSORS][STO];
            bool emit[EQ_CURSORS];
            // phase A: XOR, decide and reserve
#pragma unroll
            for (int c = 0; c < EQ_CURSORS; ++c) {
                emit[c] = false;
                sb[c] = j[c];
                if (j[c] == NIL) continue;
                uint32_t xa[SP], xb[SP];
                eq_load_row<SP>(rs + sa[c] * SP, xa);
                eq_load_row<SP>(rs + sb[c] * SP, xb);
                uint32_t any = 0;
#pragma unroll
                for (uint32_t w = 0; w < STO; ++w) {  // word w = chunk r + 1 + w
                    o[c][w] = w < SO ? (xa[w] ^ xb[w]) : 0u;
                    any |= o[c][w];
                }
                if (last) {
                    if (any != 0) continue;  // the last round must cancel both remaining chunks
                    if (prev_tree) {         // cheap early filter: two rows that share a child share indices
                        const uint2 pa = prev_tree[base + sa[c]], pb = prev_tree[base + sb[c]];
                        if (pa.x == pb.x || pa.x == pb.y || pa.y == pb.x || pa.y == pb.y) continue;
                    }
                    slot[c] = atomicAdd(cand_count, 1u);
                } else {
                    if (any == 0) continue;  // identical remainders: the two subtrees share their index set
                    nb[c] = p.bucket_bits ? (o[c][0] >> 8) : 0u;
                    slot[c] = atomicAdd(&next_cnt[nb[c] * p.cnt_stride], 1u);
                }
                emit[c] = true;
            }
            // phase B: use the reserved slots, step along the chains
            live = false;
#pragma unroll
            for (int c = 0; c < EQ_CURSORS; ++c) {
                if (emit[c]) {
                    if (last) {
                        if (slot[c] < EQ_MAX_CANDIDATES) tree[slot[c]] = make_uint2(base + sa[c], base + sb[c]);
                    } else if (slot[c] >= p.slots) {
                        atomicAdd(cand_count + 1, 1u);  // a row that did not fit its bucket is lost: counted, reported
                    } else {
                        const uint32_t id = nb[c] * p.slots + slot[c];
                        uint4* dst = reinterpret_cast<uint4*>(next_rows + (size_t)id * STO);
#pragma unroll
                        for (uint32_t q = 0; q < STO / 4; ++q)
                            dst[q] = make_uint4(o[c][4 * q], o[c][4 * q + 1], o[c][4 * q + 2], o[c][4 * q + 3]);
                        tree[id] = make_uint2(base + sa[c], base + sb[c]);
                    }
                }
                if (j[c] != NIL) j[c] = next[j[c]];
                live |= j[c] != NIL;
            }
        }
    }
}

// one instantiation per number of live words (2 .. EQ_MAX_CHUNKS)
typedef void (*EqCollideFn)(EqParams, int, const uint32_t*, uint32_t*, uint32_t*, uint32_t*, uint2*, const uint2*, uint32_t*);
EqCollideFn eq_collide_fn(int live_words)
{
    switch (live_words) {
    case 2: return eq_collide_kernel<2>;
    case 3: return eq_collide_kernel<3>;
    case 4: return eq_collide_kernel<4>;
    case 5: return eq_collide_kernel<5>;
    case 6: return eq_collide_kernel<6>;
    case 7: return eq_collide_kernel<7>;
    case 8: return eq_collide_kernel<8>;
    case 9: return eq_collide_kernel<9>;
    case 10: return eq_collide_kernel<10>;
    case 11: return eq_collide_kernel<11>;
    case 12: return eq_collide_kernel<12>;
    default: return nullptr;
    }
}

// ---- candidates -> index lists: expand the tree, order canonically, require distinct indices ----
// grid = candidates, block = 2^k threads (<= 512); trees[r] is the pair table written by round r-1 (r = 1..k).
struct EqTrees
{
    const uint2* t[EQ_MAX_CHUNKS];
    const uint32_t* leaf;  // row id of round 0 -> initial string index
};

__global__ void eq_recover_kernel(EqParams p, EqTrees trees, const uint32_t* __restrict__ cand_count, uint32_t cand_cap,
                                  uint32_t cand_base, uint32_t* __restrict__ out_indices, uint32_t* __restrict__ out_valid)
{
    __shared__ uint32_t idx[512];
    __shared__ uint32_t tmp[1024];
    __shared__ int ok;
    const uint32_t cand = cand_base + blockIdx.x;
    const uint32_t leaves = 1u << p.k;
    const uint32_t t = threadIdx.x;
    if (cand >= min(*cand_count, cand_cap)) return;
    if (t == 0) ok = 1;
    // walk from the root to leaf t: bit (k-1-d) of t picks the child at depth d
    uint32_t slot = cand;
    for (int round = p.k; round >= 1; --round) {
        const uint2 pr = trees.t[round][slot];
        slot = ((t >> (round - 1)) & 1u) ? pr.y : pr.x;
    }
    idx[t] = trees.leaf[slot];  // round-0 row id -> initial string index
    __syncthreads();
    // ordering rule, bottom-up: at every node the subtree with the smaller first index goes left
    for (int level = 0; level < p.k; ++level) {
        const uint32_t width = 1u << level;
        const uint32_t node = t / (2 * width);
        const uint32_t left = node * 2 * width;
        const bool swap = idx[left] > idx[left + width];
        const uint32_t v = idx[t];
        __syncthreads();
        if (swap) idx[t ^ width] = v;  // exchange the two halves of the node
        __syncthreads();
    }
    // distinct indices: every thread puts its index into a small open-addressing table; meeting itself means a duplicate
    const uint32_t mine = idx[t];
    tmp[t] = 0xffffffffu;
    tmp[t + leaves] = 0xffffffffu;  // table of 2 x leaves entries (indices are < 2^25, so all-ones is free as "empty")
    __syncthreads();
    {
        const uint32_t mask = 2u * leaves - 1u;
        uint32_t slot = (mine * 2654435761u) >> 7 & mask;
        for (;;) {
            const uint32_t old = atomicCAS(&tmp[slot], 0xffffffffu, mine);
            if (old == 0xffffffffu) break;
            if (old == mine) {
                ok = 0;
                break;
            }
            slot = (slot + 1u) & mask;
        }
    }
    __syncthreads();
    out_indices[(size_t)cand * leaves + t] = idx[t];
    if (t == 0) out_valid[cand] = (uint32_t)ok;
}

// ---- GPU-side validator: recompute the strings of a claimed solution and re-check every rule ----
// grid = solutions, block = 2^k threads.  rows live in dynamic shared memory: leaves * chunks words.
__global__ void eq_validate_kernel(EqParams p, EqMid mid, const uint32_t* __restrict__ indices, uint32_t* __restrict__ verdict)
{
    extern __shared__ uint32_t vrows[];
    __shared__ uint32_t vidx[512];
    __shared__ int ok;
    const uint32_t leaves = 1u << p.k;
    const uint32_t t = threadIdx.x;
    const uint32_t* sol = indices + (size_t)blockIdx.x * leaves;
    if (t == 0) ok = 1;
    __syncthreads();
    const uint32_t mine = sol[t];
    vidx[t] = mine;
    if (mine >= p.N) {
        ok = 0;
    } else {
        uint32_t ch[EQ_MAX_CHUNKS];
        eq_string_chunks(p, mid, mine, ch);
        for (int c = 0; c < p.chunks; ++c) vrows[t * p.chunks + c] = ch[c];
    }
    __syncthreads();
    for (uint32_t j = t + 1; j < leaves; ++j)
        if (vidx[j] == mine) ok = 0;
    __syncthreads();
    if (ok) {
        for (int level = 0; level < p.k; ++level) {
            const uint32_t width = 1u << level;
            if (t % (2 * width) == 0) {  // t is the first leaf of a left subtree at this level
                uint32_t* a = vrows + (size_t)t * p.chunks;
                const uint32_t* b = vrows + (size_t)(t + width) * p.chunks;
                if (a[level] != b[level]) ok = 0;            // collision on this round's chunk
                if (vidx[t] >= vidx[t + width]) ok = 0;      // ordering rule
                for (int c = 0; c < p.chunks; ++c) a[c] ^= b[c];
            }
            __syncthreads();
        }
        if (t == 0)
            for (int c = 0; c < p.chunks; ++c)
                if (vrows[c]) ok = 0;
    }
    __syncthreads();
    if (t == 0) verdict[blockIdx.x] = (uint32_t)ok;
}

bool eq_make_params(EqParams& p, int n, int k)
{
    if (n <= 0 || k <= 0 || n % 8 || n % (k + 1) || n > 512 || k + 1 > EQ_MAX_CHUNKS || k > 9) return false;
    const int cbl = n / (k + 1);
    if (cbl < 8 || cbl > 24) return false;
    p.n = n;
    p.k = k;
    p.cbl = cbl;
    p.chunks = k + 1;
    p.ipp = 512 / n;
    p.hash_bytes = n / 8;
    p.digest_bytes = p.ipp * p.hash_bytes;
    p.N = 1u << (cbl + 1);
    p.bucket_bits = cbl - 8;
    p.num_buckets = 1u << p.bucket_bits;
    p.slots = p.N <= (1u << 12) ? 2048u : 1024u;
    p.cnt_stride = EQ_CNT_STRIDE_DEFAULT;
    return true;
}

// Absorb the header on the host up to its last partial block.  The string counter is appended on the device as one
// aligned 32-bit store, so the header length must be a multiple of 4 (every real header is).
bool eq_midstate(const EqParams& p, const uint8_t tag[8], const uint8_t* header, size_t len, EqMid& mid)
{
    if (len % 4 != 0) return false;
    uint8_t personal[16];
    std::memcpy(personal, tag, 8);
    const uint32_t le_n = (uint32_t)p.n, le_k = (uint32_t)p.k;
    std::memcpy(personal + 8, &le_n, 4);
    std::memcpy(personal + 12, &le_k, 4);
    Blake2bState s;
    b2_init(s, (uint32_t)p.digest_bytes, personal);
    b2_update(s, header, len);
    // b2_update leaves the last block pending even when it is full; 4 more bytes will follow, so a full one is
    // compressed here and the counter starts a fresh block (len % 4 == 0 means a partial block has >= 4 free bytes)
    if (s.buflen == 128) {
        uint64_t m[16];
        b2_load_block(s.buf, m);
        s.t += 128;
        b2_compress(s.h, m, s.t, false);
        s.buflen = 0;
        std::memset(s.buf, 0, 128);
    }
    for (int i = 0; i < 8; ++i) mid.h[i] = s.h[i];
    uint8_t block[128] = {0};
    std::memcpy(block, s.buf, s.buflen);
    b2_load_block(block, mid.m);
    mid.t = s.t;
    mid.fill = s.buflen;
    return true;
}
}  // namespace
}  // namespace nx

using namespace nx;

struct nx_equihash
{
    EqParams p;
    int device;
    uint8_t tag[8];
    size_t smem_bytes = 0;            // dynamic shared memory of eq_collide_kernel
    uint32_t* rows[2] = {nullptr, nullptr};     // num_buckets x slots x chunks words, ping-pong
    uint32_t* bucket_cnt[2] = {nullptr, nullptr};
    uint32_t* leaf = nullptr;                   // round-0 row id -> initial string index
    uint2* tree[EQ_MAX_CHUNKS] = {nullptr};     // tree[r], r = 1..k: pairs that made the rows of round r
    uint32_t* cand_count = nullptr;
    uint32_t* cand_indices = nullptr;           // EQ_MAX_CANDIDATES x 2^k
    uint32_t* cand_valid = nullptr;
    uint32_t* verdict = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0;
    int last_launches = 0;
    uint32_t last_candidates = 0;
    uint32_t last_dropped = 0;  // rows/candidates lost to a full bucket in the last solve (0 in normal operation)
    bool counters_clean = false;  // both slot-counter arrays are known to be all zero
    nx_equihash* twin = nullptr;     // second instance of the same solver, created by the first nx_equihash_solve_batch
    uint32_t* host_first = nullptr;  // pinned: {valid[F], indices[F x 2^k], stats[2]} of the first F candidates, one async copy each
    std::mutex mu;
};

extern "C" {

size_t nx_equihash_solution_bytes(int n, int k)
{
    EqParams p;
    if (!eq_make_params(p, n, k)) return 0;
    return ((size_t)1 << k) * (size_t)(p.cbl + 1) / 8;
}

void nx_equihash_destroy(nx_equihash* s)
{
    if (!s) return;
    if (s->twin) nx_equihash_destroy(s->twin);
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    for (int i = 0; i < 2; ++i) {
        if (s->rows[i]) cudaFree(s->rows[i]);
        if (s->bucket_cnt[i]) cudaFree(s->bucket_cnt[i]);
    }
    if (s->leaf) cudaFree(s->leaf);
    for (int i = 0; i < EQ_MAX_CHUNKS; ++i)
        if (s->tree[i]) cudaFree(s->tree[i]);
    if (s->cand_count) cudaFree(s->cand_count);
    if (s->cand_indices) cudaFree(s->cand_indices);
    if (s->cand_valid) cudaFree(s->cand_valid);
    if (s->verdict) cudaFree(s->verdict);
    if (s->host_first) cudaFreeHost(s->host_first);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

int nx_equihash_create(int device, int n, int k, const uint8_t tag[8], nx_equihash** out)
{
    if (!out || !tag) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    *out = nullptr;
    EqParams p;
    if (!eq_make_params(p, n, k)) {
        set_error("unsupported Equihash parameters (n=%d, k=%d): need n %% 8 == 0, n %% (k+1) == 0, 8 <= n/(k+1) <= 24, k <= 9", n, k);
        return NX_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: this library has no CPU fallback");
        return NX_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
        set_error("device %d out of range (have %d)", device, ndev);
        return NX_ERR_INVALID;
    }
    NX_CUDA(cudaSetDevice(device));
    nx_equihash* s = new (std::nothrow) nx_equihash;
    if (!s) return NX_ERR_OOM;
    if (const char* e = std::getenv("NX_EQ_CNT_STRIDE")) {
        const int v = std::atoi(e);
        if (v == 1 || v == 8 || v == 32) p.cnt_stride = (uint32_t)v;
    }
    s->p = p;
    s->device = device;
    std::memcpy(s->tag, tag, 8);
    const size_t ids = (size_t)p.num_buckets * p.slots;  // row ids per round
    const size_t row_bytes = ids * eq_stride((uint32_t)p.chunks) * 4;
    const uint32_t leaves = 1u << k;
    s->smem_bytes = (size_t)p.slots * p.chunks * 4 + (size_t)p.slots * 2 * 2;
    bool ok = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) == cudaSuccess && cudaEventCreate(&s->ev0) == cudaSuccess &&
              cudaEventCreate(&s->ev1) == cudaSuccess && cudaMalloc((void**)&s->rows[0], row_bytes) == cudaSuccess &&
              cudaMalloc((void**)&s->rows[1], row_bytes) == cudaSuccess &&
              cudaMalloc((void**)&s->bucket_cnt[0], (size_t)p.num_buckets * p.cnt_stride * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->bucket_cnt[1], (size_t)p.num_buckets * p.cnt_stride * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->leaf, ids * 4) == cudaSuccess && cudaMalloc((void**)&s->cand_count, 8) == cudaSuccess &&
              cudaMalloc((void**)&s->cand_indices, (size_t)EQ_MAX_CANDIDATES * leaves * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->cand_valid, (size_t)EQ_MAX_CANDIDATES * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->verdict, (size_t)EQ_MAX_CANDIDATES * 4) == cudaSuccess &&
              cudaMallocHost((void**)&s->host_first, ((size_t)EQ_RECOVER_FIRST * (leaves + 1) + 2) * 4) == cudaSuccess;
    for (int r = 1; ok && r <= k; ++r)
        ok = cudaMalloc((void**)&s->tree[r], (r == k ? (size_t)EQ_MAX_CANDIDATES : ids) * sizeof(uint2)) == cudaSuccess;
    for (int live = 2; ok && live <= p.chunks; ++live) {
        // every instantiation that will run gets its own opt-in (the 48 KiB default counts the 2 KiB of static shared
        // memory too, so sizes just below it need the attribute as well)
        const size_t smem = (size_t)p.slots * eq_smem_stride(live) * 4 + (size_t)p.slots * 2;
        if (smem > 40 * 1024)
            ok = cudaFuncSetAttribute(eq_collide_fn(live), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess;
    }
    if (!ok) {
        set_error("Equihash(%d,%d): device set-up failed: %s", n, k, cudaGetErrorString(cudaGetLastError()));
        nx_equihash_destroy(s);
        return NX_ERR_OOM;
    }
    *out = s;
    return NX_OK;
}

// One solve in two halves, so that two solver instances can be kept in flight (nx_equihash_solve_batch): eq_begin queues
// everything on the instance's stream - generator, k rounds, the first candidate expansion, the copies of its results -
// and returns; eq_end waits for it and turns the candidates into encoded solutions.
static int eq_begin(nx_equihash* s, const uint8_t* header, size_t header_len)
{
    NX_CUDA(cudaSetDevice(s->device));
    const EqParams& p = s->p;
    EqMid mid;
    if (!eq_midstate(p, s->tag, header, header_len, mid)) {
        set_error("Equihash header length %zu is not a multiple of 4", header_len);
        return NX_ERR_INVALID;
    }
    const int k = p.k;
    int launches = 0;
    const size_t cnt_bytes = (size_t)p.num_buckets * p.cnt_stride * 4;

    NX_CUDA(cudaEventRecord(s->ev0, s->stream));
    NX_CUDA(cudaMemsetAsync(s->cand_count, 0, 8, s->stream));
    // The slot counters are left at zero by the kernels that read them (eq_collide_kernel), so a solve normally starts
    // on clean arrays; only after a solve that did not run to its end are they cleared here.
    if (!s->counters_clean) {
        NX_CUDA(cudaMemsetAsync(s->bucket_cnt[0], 0, cnt_bytes, s->stream));
        NX_CUDA(cudaMemsetAsync(s->bucket_cnt[1], 0, cnt_bytes, s->stream));
    }
    s->counters_clean = false;
    const uint32_t digests = (p.N + (uint32_t)p.ipp - 1) / (uint32_t)p.ipp;
    eq_generate_fn(p.n, p.k)<<<(digests + 127) / 128, 128, 0, s->stream>>>(p, mid, s->rows[0], s->leaf, s->bucket_cnt[0], s->cand_count + 1);
    NX_CUDA(cudaGetLastError());
    ++launches;
    for (int r = 0; r < k; ++r) {
        const size_t smem = (size_t)p.slots * eq_smem_stride(p.chunks - r) * 4 + (size_t)p.slots * 2;
        eq_collide_fn(p.chunks - r)<<<p.num_buckets, EQ_COLLIDE_THREADS, smem, s->stream>>>(
            p, r, s->rows[r & 1], s->bucket_cnt[r & 1], s->rows[(r + 1) & 1], s->bucket_cnt[(r + 1) & 1], s->tree[r + 1],
            r >= 1 ? s->tree[r] : nullptr, s->cand_count);
        NX_CUDA(cudaGetLastError());
        ++launches;
    }
    // Candidates are expanded without waiting for their count: a first launch covers the usual handful (its CTAs look
    // at the count themselves), and only a header with more than EQ_RECOVER_FIRST candidates costs a second launch.
    const uint32_t leaves = 1u << k;
    EqTrees trees;
    for (int i = 0; i < EQ_MAX_CHUNKS; ++i) trees.t[i] = s->tree[i];
    trees.leaf = s->leaf;
    eq_recover_kernel<<<EQ_RECOVER_FIRST, leaves, 0, s->stream>>>(p, trees, s->cand_count, EQ_MAX_CANDIDATES, 0u, s->cand_indices,
                                                                 s->cand_valid);
    NX_CUDA(cudaGetLastError());
    ++launches;
    NX_CUDA(cudaEventRecord(s->ev1, s->stream));
    // results of the first launch and the counts travel in the same queue: one wait for the whole solve
    uint32_t* const h_valid = s->host_first;
    uint32_t* const h_idx = s->host_first + EQ_RECOVER_FIRST;
    uint32_t* const stats = s->host_first + (size_t)EQ_RECOVER_FIRST * (leaves + 1);
    NX_CUDA(cudaMemcpyAsync(h_valid, s->cand_valid, (size_t)EQ_RECOVER_FIRST * 4, cudaMemcpyDeviceToHost, s->stream));
    NX_CUDA(cudaMemcpyAsync(h_idx, s->cand_indices, (size_t)EQ_RECOVER_FIRST * leaves * 4, cudaMemcpyDeviceToHost, s->stream));
    NX_CUDA(cudaMemcpyAsync(stats, s->cand_count, 8, cudaMemcpyDeviceToHost, s->stream));
    s->last_launches = launches;
    return NX_OK;
}

static int eq_end(nx_equihash* s, std::vector<std::vector<uint8_t>>& sols)
{
    sols.clear();
    NX_CUDA(cudaSetDevice(s->device));
    const EqParams& p = s->p;
    const int k = p.k;
    const uint32_t leaves = 1u << k;
    uint32_t* const h_valid = s->host_first;
    uint32_t* const h_idx = s->host_first + EQ_RECOVER_FIRST;
    uint32_t* const stats = s->host_first + (size_t)EQ_RECOVER_FIRST * (leaves + 1);
    NX_CUDA(cudaStreamSynchronize(s->stream));
    s->counters_clean = true;
    s->last_dropped = stats[1] + (stats[0] > EQ_MAX_CANDIDATES ? stats[0] - EQ_MAX_CANDIDATES : 0);
    const uint32_t ncand = std::min<uint32_t>(stats[0], EQ_MAX_CANDIDATES);
    NX_CUDA(cudaEventElapsedTime(&s->last_ms, s->ev0, s->ev1));
    if (ncand > EQ_RECOVER_FIRST) {
        EqTrees trees;
        for (int i = 0; i < EQ_MAX_CHUNKS; ++i) trees.t[i] = s->tree[i];
        trees.leaf = s->leaf;
        NX_CUDA(cudaEventRecord(s->ev0, s->stream));
        eq_recover_kernel<<<ncand - EQ_RECOVER_FIRST, leaves, 0, s->stream>>>(p, trees, s->cand_count, EQ_MAX_CANDIDATES,
                                                                             EQ_RECOVER_FIRST, s->cand_indices, s->cand_valid);
        NX_CUDA(cudaGetLastError());
        ++s->last_launches;
        NX_CUDA(cudaEventRecord(s->ev1, s->stream));
        NX_CUDA(cudaStreamSynchronize(s->stream));
        float more = 0;
        NX_CUDA(cudaEventElapsedTime(&more, s->ev0, s->ev1));
        s->last_ms += more;
    }
    s->last_candidates = ncand;
    if (ncand == 0) return NX_OK;
    std::vector<uint32_t> valid(ncand), idx((size_t)ncand * leaves);
    const uint32_t have = std::min<uint32_t>(ncand, EQ_RECOVER_FIRST);  // already on the host
    std::memcpy(valid.data(), h_valid, (size_t)have * 4);
    std::memcpy(idx.data(), h_idx, (size_t)have * leaves * 4);
    if (ncand > have) {
        NX_CUDA(cudaMemcpy(valid.data() + have, s->cand_valid + have, (size_t)(ncand - have) * 4, cudaMemcpyDeviceToHost));
        NX_CUDA(cudaMemcpy(idx.data() + (size_t)have * leaves, s->cand_indices + (size_t)have * leaves,
                           (size_t)(ncand - have) * leaves * 4, cudaMemcpyDeviceToHost));
    }
    const size_t sb = ((size_t)1 << k) * (size_t)(p.cbl + 1) / 8;
    for (uint32_t c = 0; c < ncand; ++c) {
        if (!valid[c]) continue;
        std::vector<uint8_t> enc(sb, 0);
        size_t pos = 0;
        for (uint32_t i = 0; i < leaves; ++i)
            for (int b = p.cbl; b >= 0; --b, ++pos)
                if ((idx[(size_t)c * leaves + i] >> b) & 1u) enc[pos >> 3] |= (uint8_t)(0x80u >> (pos & 7));
        sols.push_back(std::move(enc));
    }
    std::sort(sols.begin(), sols.end());
    sols.erase(std::unique(sols.begin(), sols.end()), sols.end());
    return NX_OK;
}

// All solutions of one header (header = everything that is hashed before the LE32 string counter, nonce included;
// its length must be a multiple of 4).  solutions_out receives up to max_solutions minimal encodings, sorted
// bytewise; *n_solutions the number found.
int nx_equihash_solve(nx_equihash* s, const uint8_t* header, size_t header_len, uint8_t* solutions_out, int max_solutions,
                      int* n_solutions)
{
    if (!s || !header || !n_solutions || (!solutions_out && max_solutions > 0)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    *n_solutions = 0;
    return guarded([&]() -> int {
        std::lock_guard<std::mutex> lk(s->mu);
        int rc = eq_begin(s, header, header_len);
        if (rc != NX_OK) return rc;
        std::vector<std::vector<uint8_t>> sols;
        if ((rc = eq_end(s, sols)) != NX_OK) return rc;
        const size_t sb = nx_equihash_solution_bytes(s->p.n, s->p.k);
        *n_solutions = (int)sols.size();
        for (size_t i = 0; i < sols.size() && (int)i < max_solutions; ++i) std::memcpy(solutions_out + i * sb, sols[i].data(), sb);
        return NX_OK;
    });
}

// Solutions of n headers of equal length, two solves in flight: while one instance's kernels run, the other's results
// are decoded on the host and its next header is queued, so the GPU never waits for the host between nonces and the
// thin tail of one solve (candidate expansion, the last narrow rounds) overlaps the wide start of the next.
// counts[i] = solutions of header i; their encodings follow each other in solutions_out (header order, each header's
// sorted), as many as max_solutions_total allows.  The second instance (another set of row buffers) is created on the
// first call and lives as long as the solver.
int nx_equihash_solve_batch(nx_equihash* s, const uint8_t* headers, size_t header_len, int n, uint8_t* solutions_out,
                            int max_solutions_total, int* counts, float* wall_ms)
{
    if (!s || (n > 0 && (!headers || !counts)) || (!solutions_out && max_solutions_total > 0) || n < 0) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    return guarded([&]() -> int {
        std::lock_guard<std::mutex> lk(s->mu);
        if (!s->twin) {
            const int rc = nx_equihash_create(s->device, s->p.n, s->p.k, s->tag, &s->twin);
            if (rc != NX_OK) return rc;
            s->twin->p.cnt_stride = s->p.cnt_stride;
        }
        nx_equihash* lane[2] = {s, s->twin};
        const size_t sb = nx_equihash_solution_bytes(s->p.n, s->p.k);
        const auto t0 = std::chrono::steady_clock::now();
        int written = 0;
        uint32_t dropped = 0, cands = 0;
        std::vector<std::vector<uint8_t>> sols;
        auto finish = [&](int i) -> int {  // header i ran on lane i & 1
            const int rc = eq_end(lane[i & 1], sols);
            if (rc != NX_OK) return rc;
            counts[i] = (int)sols.size();
            dropped += lane[i & 1]->last_dropped;
            cands += lane[i & 1]->last_candidates;
            for (size_t q = 0; q < sols.size() && written < max_solutions_total; ++q, ++written)
                std::memcpy(solutions_out + (size_t)written * sb, sols[q].data(), sb);
            return NX_OK;
        };
        for (int i = 0; i < n; ++i) {
            int rc;
            if (i >= 2 && (rc = finish(i - 2)) != NX_OK) return rc;
            if ((rc = eq_begin(lane[i & 1], headers + (size_t)i * header_len, header_len)) != NX_OK) return rc;
        }
        for (int i = n >= 2 ? n - 2 : 0; i < n; ++i) {
            const int rc = finish(i);
            if (rc != NX_OK) return rc;
        }
        const float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (wall_ms) *wall_ms = ms;
        s->last_ms = ms;
        s->last_dropped = dropped;
        s->last_candidates = cands;
        return NX_OK;
    });
}

// GPU-side validator for `count` encoded solutions of one header: verdict[i] = 1 when solution i is valid.
int nx_equihash_validate(nx_equihash* s, const uint8_t* header, size_t header_len, const uint8_t* solutions, int count,
                         uint8_t* verdict)
{
    if (!s || !header || ((!solutions || !verdict) && count > 0)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (count <= 0) return NX_OK;
    if (count > EQ_MAX_CANDIDATES) {
        set_error("at most %d solutions per call", EQ_MAX_CANDIDATES);
        return NX_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(s->mu);
    NX_CUDA(cudaSetDevice(s->device));
    const EqParams& p = s->p;
    const uint32_t leaves = 1u << p.k;
    const size_t sb = (size_t)leaves * (size_t)(p.cbl + 1) / 8;
    std::vector<uint32_t> idx((size_t)count * leaves);
    for (int c = 0; c < count; ++c) {
        size_t pos = 0;
        const uint8_t* in = solutions + (size_t)c * sb;
        for (uint32_t i = 0; i < leaves; ++i) {
            uint32_t v = 0;
            for (int b = 0; b <= p.cbl; ++b, ++pos) v = (v << 1) | ((in[pos >> 3] >> (7 - (pos & 7))) & 1u);
            idx[(size_t)c * leaves + i] = v;
        }
    }
    EqMid mid;
    if (!eq_midstate(p, s->tag, header, header_len, mid)) {
        set_error("Equihash header length %zu is not a multiple of 4", header_len);
        return NX_ERR_INVALID;
    }
    NX_CUDA(cudaMemcpyAsync(s->cand_indices, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, s->stream));
    eq_validate_kernel<<<count, leaves, (size_t)leaves * p.chunks * 4, s->stream>>>(p, mid, s->cand_indices, s->verdict);
    NX_CUDA(cudaGetLastError());
    std::vector<uint32_t> v((size_t)count);
    NX_CUDA(cudaMemcpyAsync(v.data(), s->verdict, (size_t)count * 4, cudaMemcpyDeviceToHost, s->stream));
    NX_CUDA(cudaStreamSynchronize(s->stream));
    for (int c = 0; c < count; ++c) verdict[c] = (uint8_t)(v[(size_t)c] ? 1 : 0);
    return NX_OK;
}

int nx_equihash_last_dropped(const nx_equihash* s) { return s ? (int)s->last_dropped : 0; }
int nx_equihash_last_candidates(const nx_equihash* s) { return s ? (int)s->last_candidates : 0; }

int nx_equihash_last_timing(const nx_equihash* s, float* kernel_ms, int* launches)
{
    if (!s) return NX_ERR_INVALID;
    if (kernel_ms) *kernel_ms = s->last_ms;
    if (launches) *launches = s->last_launches;
    return NX_OK;
}

}  // extern "C"
