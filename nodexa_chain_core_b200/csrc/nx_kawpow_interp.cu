// KawPow for MANY INDEPENDENT (header, nonce, height) records — kernel K6, the verify direction.
//
// Header validation hashes each block once, at its own height, so nearly every record belongs to a
// different ProgPoW period (period = height / 3).  Compiling a program per period (K1/K2, ~0.8 s of NVRTC
// each) would be absurd there, so this kernel INTERPRETS the period program like the reference does
// (progpow.cpp:92-144,179-245), with the program a small table in shared memory: one half-warp = the 16
// lanes of one record, the 32 mix words of each lane in shared memory (the register index is data here),
// the 16 KiB L1 cache in shared memory, the DAG item of a round fetched as 16 coalesced 16-byte loads.
// It runs at roughly a quarter of the compiled kernel's speed and needs no compilation at all.
//
// Replaces: progpow::hash / verify called once per header (reference src/validation.cpp:11652-11662,
// src/rpc/mining.cpp:811,880, src/net_processing.cpp:2596-2597) and the 1000-height loop of
// kawpow_tests.cpp:41-56.
#include "nx_internal.hpp"

#define NX_HELPERS_ONLY 1
#include "kawpow_kernel.cuh"

namespace nx
{
namespace
{
constexpr int IT_THREADS = 128;                 // 8 records per CTA
constexpr int IT_RECORDS = IT_THREADS / 16;
constexpr int IT_MIX_STRIDE = 32 * 16 + 16;     // words per record: [reg][lane], +16 so neighbours differ in bank half

struct FastMod32
{
    u32 d, m;
};
__device__ __forceinline__ u32 fast_mod32(u32 x, FastMod32 f)
{
    const u32 r = x - __umulhi(x, f.m) * f.d;
    return min(r, r - f.d);
}

__device__ __forceinline__ u32 it_merge(u32 a, u32 b, u32 kind, u32 x)
{
    switch (kind) {
    case 0: return a * 33u + b;
    case 1: return (a ^ b) * 33u;
    case 2: return nx_rotl(a, x) ^ b;
    default: return nx_rotr(a, x) ^ b;
    }
}

__device__ __forceinline__ u32 it_math(u32 a, u32 b, u32 kind)
{
    switch (kind) {
    default:
    case 0: return a + b;
    case 1: return a * b;
    case 2: return nx_mulhi(a, b);
    case 3: return nx_min(a, b);
    case 4: return nx_rotl(a, b);
    case 5: return nx_rotr(a, b);
    case 6: return a & b;
    case 7: return a | b;
    case 8: return a ^ b;
    case 9: return nx_clz(a) + nx_clz(b);
    case 10: return nx_popc(a) + nx_popc(b);
    }
}

// ops: per program 40 words-of-4: {kind, src1, src2|(unused), dst} packed as x = kind | src1<<8 | src2<<16 | dst<<24,
//      y = math selector (0..10), z = merge kind (0..3), w = merge rotate amount.
__global__ void __launch_bounds__(IT_THREADS)
kawpow_interp_kernel(const uint4* __restrict__ dag, FastMod32 fm, const KawpowRecord* __restrict__ recs, u32 n_recs,
                     const uint4* __restrict__ programs, u32* __restrict__ out)
{
    __shared__ __align__(16) u32 l1s[4096];
    __shared__ u32 mixs[IT_RECORDS * IT_MIX_STRIDE];
    __shared__ uint4 progs[IT_RECORDS][36];

    const u32 tid = threadIdx.x;
    const u32 lane = tid & 15u;
    const u32 slot = tid >> 4;
    const u32 rec_idx = blockIdx.x * IT_RECORDS + slot;
    const bool live = rec_idx < n_recs;
    const KawpowRecord rec = recs[live ? rec_idx : n_recs - 1];

    for (u32 i = tid; i < 1024u; i += IT_THREADS) reinterpret_cast<uint4*>(l1s)[i] = __ldg(dag + i);
    for (u32 i = lane; i < 33u; i += 16u) progs[slot][i] = __ldg(programs + (size_t)rec.program_slot * 36 + i);

    // every lane of the record repeats the first absorb (no divergence; 16x a cheap step)
    u32 st[25];
#pragma unroll
    for (int i = 0; i < 8; ++i) st[i] = rec.header[i];
    st[8] = (u32)rec.nonce;
    st[9] = (u32)(rec.nonce >> 32);
    NX_PAD_FIRST(st);
    nxk_f800(st);
    u32 carry[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) carry[i] = st[i];

    u32* const mix = mixs + slot * IT_MIX_STRIDE + lane;  // word r of this lane at mix[r * 16]
    {
        u32 m[32];
        nx_fill_mix(carry[0], carry[1], lane, m);
#pragma unroll
        for (int i = 0; i < 32; ++i) mix[i * 16] = m[i];
    }
    __syncthreads();

#pragma unroll 1
    for (u32 r = 0; r < 64u; ++r) {
        const u32 item = fast_mod32(mixs[slot * IT_MIX_STRIDE + (r & 15u)], fm);  // mix[r%16][0]
        __syncwarp();  // everybody has read it before lane r%16 may overwrite its word 0 below
        const uint4 d = __ldg(dag + ((u64)item * 16u + ((lane ^ r) & 15u)));
#pragma unroll 1
        for (u32 i = 0; i < 29u; ++i) {
            const uint4 op = progs[slot][i];
            const u32 kind = op.x & 0xffu, s1 = (op.x >> 8) & 0xffu, s2 = (op.x >> 16) & 0xffu, dst = op.x >> 24;
            u32 data;
            if (kind == 0)
                data = l1s[mix[s1 * 16] & 4095u];
            else
                data = it_math(mix[s1 * 16], mix[s2 * 16], op.y);
            mix[dst * 16] = it_merge(mix[dst * 16], data, op.z, op.w);
        }
        const u32 dw[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (u32 i = 0; i < 4u; ++i) {
            const uint4 op = progs[slot][29 + i];
            const u32 dst = op.x >> 24;
            mix[dst * 16] = it_merge(mix[dst * 16], dw[i], op.z, op.w);
        }
        __syncwarp();  // next round's item index is read from another lane's word 0
    }

    u32 lh = NX_FNV_BASIS;
#pragma unroll 1
    for (int i = 0; i < 32; ++i) lh = nx_fnv1a(lh, mix[i * 16]);
    u32 dg[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
        dg[i] = nx_fnv1a(nx_fnv1a(NX_FNV_BASIS, __shfl_sync(0xffffffffu, lh, i, 16)), __shfl_sync(0xffffffffu, lh, i + 8, 16));

#pragma unroll
    for (int i = 0; i < 8; ++i) st[i] = carry[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) st[8 + i] = dg[i];
    NX_PAD_LAST(st);
    nxk_f800(st);
    if (live && lane == 0) {
        uint4* o = reinterpret_cast<uint4*>(out + (size_t)rec.out_index * 16u);  // records arrive grouped by period
        o[0] = make_uint4(st[0], st[1], st[2], st[3]);
        o[1] = make_uint4(st[4], st[5], st[6], st[7]);
        o[2] = make_uint4(dg[0], dg[1], dg[2], dg[3]);
        o[3] = make_uint4(dg[4], dg[5], dg[6], dg[7]);
    }
}
}  // namespace

// Pack the period program of prog_seed into 36 uint4 (33 used).
void kawpow_pack_program(uint64_t prog_seed, uint32_t out[36 * 4])
{
    const std::vector<ProgramOp> ops = kawpow_program(prog_seed);
    memset(out, 0, 36 * 16);
    for (size_t i = 0; i < ops.size() && i < 36; ++i) {
        const ProgramOp& op = ops[i];
        const uint32_t kind = op.kind == ProgramOp::CACHE ? 0u : op.kind == ProgramOp::MATH ? 1u : 2u;
        out[4 * i + 0] = kind | (op.src1 << 8) | (op.src2 << 16) | (op.dst << 24);
        out[4 * i + 1] = op.sel_math % 11u;
        out[4 * i + 2] = op.sel_merge % 4u;
        out[4 * i + 3] = ((op.sel_merge >> 16) % 31u) + 1u;
    }
}

int kawpow_interp_launch(const void* dag, uint32_t dag_items, const KawpowRecord* recs_dev, uint32_t n_recs,
                         const void* programs_dev, void* out_dev, cudaStream_t s)
{
    if (n_recs == 0) return NX_OK;
    FastMod32 fm;
    fm.d = dag_items;
    fm.m = (uint32_t)((1ull << 32) / dag_items);
    const uint32_t blocks = (n_recs + IT_RECORDS - 1) / IT_RECORDS;
    kawpow_interp_kernel<<<blocks, IT_THREADS, 0, s>>>(static_cast<const uint4*>(dag), fm, recs_dev, n_recs,
                                                      static_cast<const uint4*>(programs_dev), static_cast<u32*>(out_dev));
    NX_CUDA(cudaGetLastError());
    return NX_OK;
}

}  // namespace nx
