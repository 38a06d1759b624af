// Micro-benchmarks behind the roofline statements of DESIGN.md section 4 ("verify with a
// microbenchmark ... declare the binder").  Stand-alone program, not part of libnxkawpow.so:
//
//   int     INT32 issue rate of one SM: dependent-free LOP3 / IMAD / SHF streams and the 64:38 mix of a KawPow round
//   lds     random 4-byte shared-memory reads from a 16 KiB table, 1 replica and 8 interleaved replicas (K1's L1 cache)
//   gather  random 64-byte gathers from a 72 MiB table that lives in L2 (K3's light-cache walk), 1 and 4 in flight
//   dram    independent random 256-byte reads from a 4.5 GiB buffer (K1's DAG item), against the copy peak
//
// Prints one JSON object.  Every kernel is timed with CUDA events after a warm-up launch; SM cycles come from clock64().
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            fprintf(stderr, "%s failed: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

typedef unsigned int u32;
typedef unsigned long long u64;

__device__ __forceinline__ u32 xorshift(u32& s)
{
    s ^= s << 13;
    s ^= s >> 17;
    s ^= s << 5;
    return s;
}

// ------------------------------------------------------------------------------------------------ int
// 8 independent chains per thread, ITER x 8 instructions; 1024 threads per SM like K1.  kind 0: LOP3, 1: IMAD,
// 2: SHF (funnel shift), 3: the 64 ALU : 38 FMA mix of one KawPow round (profiles: 64 LOP3/SHF/IADD3 + 38 IMAD of 125).
template <int KIND>
__global__ void __launch_bounds__(1024, 1) int_kernel(u32* out, int iters, u32 seed, u64* cycles)
{
    u32 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    const u32 k1 = seed | 1u, k2 = seed ^ 0x5bd1e995u;
    const u64 t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (KIND == 0) a[i] = (a[i] & k1) ^ k2;                       // LOP3
                if (KIND == 1) a[i] = a[i] * k1 + k2;                         // IMAD
                if (KIND == 2) a[i] = __funnelshift_l(a[i], k2, 7);           // SHF
                if (KIND == 3) {                                              // 5 ALU : 3 FMA per 8 instructions
                    if (i < 3) a[i] = a[i] * 33u + a[(i + 1) & 7];            // IMAD
                    else if (i == 3) a[i] = __funnelshift_l(a[i], a[i], 11) ^ a[7];  // SHF + LOP3
                    else if (i < 7) a[i] = (a[i] & k1) ^ a[(i + 5) & 7];      // LOP3
                }
            }
        }
    }
    const u64 t1 = clock64();
    u32 x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// ------------------------------------------------------------------------------------------------ lds
// Each lane does ITER x 8 random reads of a 4096-word table, the index depending on the values read before (like K1's
// index = mix word & 4095).  COPIES interleaved replicas as in kawpow_kernel.cuh.  Five integer instructions ride along
// with each read (LCG, shift, mask, address, merge) - fewer than the ~10 a KawPow round has per read.
template <int COPIES>
__global__ void __launch_bounds__(1024, 1) lds_kernel(u32* out, int iters, u32 seed, u64* cycles)
{
    extern __shared__ u32 tab[];
    for (u32 i = threadIdx.x; i < 4096u * COPIES; i += blockDim.x) tab[i] = (i / COPIES) * 2654435761u + seed;
    __syncthreads();
    const u32 rep = (threadIdx.x & 31u) / (32 / COPIES);
    // The index of each read mixes a per-lane LCG stream with the values read so far: it depends on the data (as
    // KawPow's `mix[src] % 4096` does) but cannot fall into a short cycle the way an index fed by the table alone would.
    u32 a[8], s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = seed * (threadIdx.x * 8 + i + 1);
        s[i] = (seed ^ ((blockIdx.x * blockDim.x + threadIdx.x) * 8u + i) * 2654435761u) | 1u;
    }
    const u64 t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            s[i] = s[i] * 1664525u + 1013904223u;  // one multiply-add; the high half of an LCG is a long stream
            a[i] = a[i] * 33u + tab[((a[i] ^ (s[i] >> 16)) & 4095u) * COPIES + rep];
        }
    }
    const u64 t1 = clock64();
    u32 x = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) x ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = x;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// ------------------------------------------------------------------------------------------------ gather
// A lane quad reads one random 64-byte item (4 x 16 B) per step, INFLIGHT independent items at a time.
template <int INFLIGHT>
__global__ void __launch_bounds__(256) gather_kernel(const uint4* __restrict__ table, u32 items, u32* out, int iters, u32 seed)
{
    const u32 gid = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 sub = threadIdx.x & 3u;
    u32 s[INFLIGHT];
#pragma unroll
    for (int j = 0; j < INFLIGHT; ++j) s[j] = (seed + (gid >> 2) * 2654435761u + j * 0x85ebca6bu) | 1u;
    u32 acc = 0;
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        uint4 v[INFLIGHT];
#pragma unroll
        for (int j = 0; j < INFLIGHT; ++j) {
            const u32 idx = (u32)(((u64)xorshift(s[j]) * items) >> 32);
            v[j] = __ldg(table + (size_t)idx * 4 + sub);
        }
#pragma unroll
        for (int j = 0; j < INFLIGHT; ++j) {
            acc ^= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
            s[j] ^= __shfl_sync(0xffffffffu, v[j].x, 0, 4) & 0xffu;  // the next index depends on the data, as in the FNV walk
            s[j] |= 1u;
        }
    }
    out[gid] = acc;
}

// ------------------------------------------------------------------------------------------------ dram
// A group of LANES threads reads one random item of LANES x 16 bytes per step (16 lanes = KawPow's 256-byte DAG item,
// 8 lanes = classic Ethash's 128-byte item), INFLIGHT independent items at a time.
template <int INFLIGHT, int LANES>
__global__ void __launch_bounds__(256) dram_kernel(const uint4* __restrict__ buf, u32 items, u32* out, int iters, u32 seed)
{
    const u32 gid = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 lane = threadIdx.x & (LANES - 1);
    u32 s[INFLIGHT];
#pragma unroll
    for (int j = 0; j < INFLIGHT; ++j) s[j] = (seed + (gid / LANES) * 2654435761u + j * 0x85ebca6bu) | 1u;
    u32 acc = 0;
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        uint4 v[INFLIGHT];
#pragma unroll
        for (int j = 0; j < INFLIGHT; ++j) {
            const u32 idx = (u32)(((u64)xorshift(s[j]) * items) >> 32);
            const uint4* p = buf + (size_t)idx * LANES + lane;
            asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w)
                         : "l"(p));
        }
#pragma unroll
        for (int j = 0; j < INFLIGHT; ++j) acc ^= v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
    }
    out[gid] = acc;
}

__global__ void fill_kernel(uint4* p, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_uint4((u32)i, (u32)(i >> 7), (u32)i * 2654435761u, 0x01000193u);
}

template <class F>
static float time_ms(F&& launch, int reps = 3)
{
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    launch();  // warm-up
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    CK(cudaEventDestroy(e0));
    CK(cudaEventDestroy(e1));
    return best;
}

int main(int argc, char** argv)
{
    std::string which = argc > 1 ? argv[1] : "all";
    int dev = 0;
    CK(cudaSetDevice(dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    const int sms = prop.multiProcessorCount;
    u32* out = nullptr;
    u64* cyc = nullptr;
    CK(cudaMalloc(&out, (size_t)sms * 64 * 1024 * 4));
    CK(cudaMalloc(&cyc, (size_t)sms * 8));
    std::vector<u64> hc(sms);
    auto mean_cycles = [&] {
        CK(cudaMemcpy(hc.data(), cyc, (size_t)sms * 8, cudaMemcpyDeviceToHost));
        double s = 0;
        for (u64 c : hc) s += (double)c;
        return s / sms;
    };
    printf("{\"device\": \"%s\", \"sms\": %d", prop.name, sms);

    if (which == "all" || which == "int") {
        const int iters = 4096;
        const char* names[4] = {"lop3", "imad", "shf", "kawpow_mix_5alu_3fma"};
        const double instr_per_iter[4] = {32, 32, 32, 32};
        printf(", \"int\": {\"threads_per_sm\": 1024");
        for (int k = 0; k < 4; ++k) {
            float ms = time_ms([&] {
                if (k == 0) int_kernel<0><<<sms, 1024>>>(out, iters, 12345u, cyc);
                if (k == 1) int_kernel<1><<<sms, 1024>>>(out, iters, 12345u, cyc);
                if (k == 2) int_kernel<2><<<sms, 1024>>>(out, iters, 12345u, cyc);
                if (k == 3) int_kernel<3><<<sms, 1024>>>(out, iters, 12345u, cyc);
            });
            const double cycles = mean_cycles();
            const double warp_instr = 32.0 * iters * instr_per_iter[k];  // per SM: 32 warps
            printf(", \"%s\": {\"warp_instr_per_clk_per_sm\": %.3f, \"ms\": %.3f, \"sm_mhz\": %.0f}", names[k],
                   warp_instr / cycles, ms, cycles / (ms * 1e3));
        }
        printf("}");
    }
    if (which == "all" || which == "lds") {
        const int iters = 2048;
        printf(", \"lds\": {\"threads_per_sm\": 1024, \"table_bytes\": 16384");
        for (int copies : {1, 8}) {
            const size_t smem = (size_t)copies * 16384;
            if (copies == 8) CK(cudaFuncSetAttribute(lds_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            float ms = time_ms([&] {
                if (copies == 1) lds_kernel<1><<<sms, 1024, smem>>>(out, iters, 777u, cyc);
                else lds_kernel<8><<<sms, 1024, smem>>>(out, iters, 777u, cyc);
            });
            const double cycles = mean_cycles();
            const double warp_lds = 32.0 * iters * 8;
            printf(", \"copies_%d\": {\"warp_lds_per_clk_per_sm\": %.4f, \"clk_per_warp_lds\": %.3f, \"lane_reads_per_s_per_gpu\": %.4g, "
                   "\"ms\": %.3f, \"sm_mhz\": %.0f}",
                   copies, warp_lds / cycles, cycles / warp_lds, warp_lds * 32 * sms / (ms * 1e-3), ms, cycles / (ms * 1e3));
        }
        printf("}");
    }
    if (which == "all" || which == "gather") {
        const u32 items = 1179641;  // light cache of epoch 448: 72 MiB of 64-byte items
        uint4* table = nullptr;
        CK(cudaMalloc(&table, (size_t)items * 64));
        fill_kernel<<<sms * 8, 256>>>(table, (size_t)items * 4);
        const int iters = 2048;
        const int blocks = sms * 8;
        printf(", \"gather\": {\"table_bytes\": %zu, \"item_bytes\": 64", (size_t)items * 64);
        for (int inflight : {1, 4, 8}) {
            float ms = time_ms([&] {
                if (inflight == 1) gather_kernel<1><<<blocks, 256>>>(table, items, out, iters * 4, 99u);
                if (inflight == 4) gather_kernel<4><<<blocks, 256>>>(table, items, out, iters, 99u);
                if (inflight == 8) gather_kernel<8><<<blocks, 256>>>(table, items, out, iters / 2, 99u);
            });
            const double bytes = (double)blocks * 256 / 4 * (double)iters * 4 * 64;
            printf(", \"inflight_%d\": {\"tb_per_s\": %.3f, \"ms\": %.3f}", inflight, bytes / (ms * 1e-3) / 1e12, ms);
        }
        printf("}");
        CK(cudaFree(table));
    }
    if (which == "all" || which == "dram") {
        const u32 items = 18874358;  // DAG of epoch 448: 4.5 GiB of 256-byte items
        uint4* buf = nullptr;
        CK(cudaMalloc(&buf, (size_t)items * 256));
        fill_kernel<<<sms * 8, 256>>>(buf, (size_t)items * 16);
        const int iters = 1024;
        const int blocks = sms * 8;
        printf(", \"dram\": {\"buffer_bytes\": %zu", (size_t)items * 256);
        for (int lanes : {16, 8}) {
            printf(", \"item_%d_bytes\": {\"threads_per_sm\": 2048", lanes * 16);
            for (int inflight : {1, 2, 4, 8}) {
                float ms = time_ms([&] {
                    if (lanes == 16) {
                        if (inflight == 1) dram_kernel<1, 16><<<blocks, 256>>>(buf, items, out, iters * 8, 5u);
                        if (inflight == 2) dram_kernel<2, 16><<<blocks, 256>>>(buf, items, out, iters * 4, 5u);
                        if (inflight == 4) dram_kernel<4, 16><<<blocks, 256>>>(buf, items, out, iters * 2, 5u);
                        if (inflight == 8) dram_kernel<8, 16><<<blocks, 256>>>(buf, items, out, iters, 5u);
                    } else {
                        if (inflight == 1) dram_kernel<1, 8><<<blocks, 256>>>(buf, items * 2, out, iters * 8, 5u);
                        if (inflight == 2) dram_kernel<2, 8><<<blocks, 256>>>(buf, items * 2, out, iters * 4, 5u);
                        if (inflight == 4) dram_kernel<4, 8><<<blocks, 256>>>(buf, items * 2, out, iters * 2, 5u);
                        if (inflight == 8) dram_kernel<8, 8><<<blocks, 256>>>(buf, items * 2, out, iters, 5u);
                    }
                });
                const double bytes = (double)blocks * 256 / lanes * (double)iters * 8 * lanes * 16;
                printf(", \"inflight_%d\": {\"gb_per_s\": %.1f, \"ms\": %.3f}", inflight, bytes / (ms * 1e-3) / 1e9, ms);
            }
            printf("}");
        }
        printf("}");
        CK(cudaFree(buf));
    }
    printf("}\n");
    return 0;
}
