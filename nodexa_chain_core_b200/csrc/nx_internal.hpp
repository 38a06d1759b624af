// Internal declarations shared by the translation units of libnxkawpow.so.  Not installed.
#pragma once
#include "../../include/nx_kawpow.h"

#include <cstdint>
#include <cstdio>
#include <memory>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>

#include <cuda_runtime.h>

namespace nx
{
// ---- error plumbing (thread-local message behind nx_last_error) ----
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
void clear_error();

#define NX_CUDA(call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            nx::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return NX_ERR_CUDA;                                                                    \
        }                                                                                          \
    } while (0)

// "None throws" (include/nx_kawpow.h): entry points that allocate on the host run their body through this.
template <class F>
int guarded(F&& body) noexcept
{
    try {
        return body();
    } catch (const std::bad_alloc&) {
        set_error("out of host memory");
        return NX_ERR_OOM;
    } catch (const std::exception& e) {
        set_error("unexpected host error: %s", e.what());
        return NX_ERR_STATE;
    } catch (...) {
        set_error("unexpected host error");
        return NX_ERR_STATE;
    }
}

// ---- host algorithms (nx_host.cpp) ----
void host_build_light_cache(uint8_t* cache, int num_items, const uint8_t seed[32]);  // dispatches on the CPU
void host_build_light_cache_scalar(uint8_t* cache, int num_items, const uint8_t seed[32]);
void host_build_light_cache_avx512(uint8_t* cache, int num_items, const uint8_t seed[32]);  // nx_lightcache_avx512.cpp
void keccakf1600_avx512(uint64_t a[25]);
bool cpu_has_avx512f();  // nx_cpu_dispatch.cpp
bool host_load_light_cache(uint8_t* cache, int num_items, int epoch);        // NX_CACHE_DIR, verified; false = rebuild
void host_store_light_cache(const uint8_t* cache, int num_items, int epoch);

// ---- program generator (nx_codegen.cpp) ----
struct ProgramOp
{
    enum Kind { CACHE = 0, MATH = 1, DAG = 2 } kind;
    uint32_t src1, src2, dst, sel_math, sel_merge;
};
std::vector<ProgramOp> kawpow_program(uint64_t prog_seed);
std::string kawpow_program_text(uint64_t prog_seed);
std::string kawpow_source(uint64_t prog_seed, uint32_t dag_items_256);
const char* kawpow_kernel_skeleton();  // the embedded text of kawpow_kernel.cuh

// ---- NVRTC program cache + launches (nx_kawpow_rt.cpp) ----
struct KawpowJob
{
    uint32_t header[8];
    uint32_t boundary_be[8];
    uint64_t start_nonce;
    uint64_t count;
    uint64_t report_base;
    uint32_t mode;  // 0 batch hash, 1 search (set by the launch helpers)
    uint32_t l1_mul_lo, l1_mul_hi;  // L1 addressing multipliers for the NX_L1_HI == 2 shape (set by the launch helpers)
    uint32_t per_hw, n_units;       // launch shape (set by the launch helpers, see kawpow_launch_shape)
    uint32_t pad;
};
static_assert(sizeof(KawpowJob) == 112, "KawpowJob must mirror NxJob in kawpow_kernel.cuh");
void kawpow_launch_shape(uint64_t count, int threads, int sms, int minblocks, bool persist, uint32_t* per_hw,
                         uint32_t* n_units, uint32_t* grid);
struct KawpowProgram;  // one loaded module: (device, prog_seed, dag_items)
int kawpow_get_program(int device, uint64_t prog_seed, uint32_t dag_items_256, std::shared_ptr<KawpowProgram>* out);
int kawpow_compile_only(uint64_t prog_seed, uint32_t dag_items_256, std::vector<char>& cubin, std::string& log);
int kawpow_launch_hash(const KawpowProgram* p, const void* dag, const KawpowJob& job, void* out_dev, cudaStream_t s);
int kawpow_launch_search(const KawpowProgram* p, const void* dag, const KawpowJob& job, void* found_min_dev,
                         cudaStream_t s);

// ---- interpreted KawPow for independent records (nx_kawpow_interp.cu, kernel K6) ----
struct KawpowRecord
{
    uint32_t header[8];
    uint64_t nonce;
    uint32_t program_slot;  // index into the packed program table of this launch
    uint32_t out_index;     // where this record's {final, mix} go in the output array (the caller's record order)
};
void kawpow_pack_program(uint64_t prog_seed, uint32_t out[36 * 4]);
int kawpow_interp_launch(const void* dag, uint32_t dag_items, const KawpowRecord* recs_dev, uint32_t n_recs,
                         const void* programs_dev, void* out_dev, cudaStream_t s);
bool kawpow_program_is_cached(int device, uint64_t prog_seed, uint32_t dag_items_256);

// ---- CUDA kernels (nx_dag.cu) ----
constexpr int NX_MAX_PEERS = 16;
int dag_generate(const void* light_dev, int light_items, uint32_t first_item256, uint32_t count_item256,
                 void* out_dev, cudaStream_t s);
// same slice stored into n_outs buffers (local + peer-mapped); each pointer already offset to the slice start
int dag_generate_multi(const void* light_dev, int light_items, uint32_t first_item256, uint32_t count_item256,
                       void* const* outs_dev, int n_outs, cudaStream_t s);
int ethash_hash_batch(const void* dag_dev, uint32_t num_items_128, const uint8_t header[32], uint64_t start_nonce,
                      uint64_t count, void* out_dev, cudaStream_t s);
int light_cache_build_gpu(void* light_dev, int light_items, const uint8_t seed[32], cudaStream_t s);
int dag_checksum(const void* dag_dev, size_t n_words64, uint64_t* partials_dev, int n_partials, uint64_t* out_dev,
                 cudaStream_t s);
constexpr int CHECKSUM_PARTIALS = 148 * 8;

}  // namespace nx

// The epoch context behind the opaque nx_epoch handle.
struct nx_epoch
{
    int epoch = 0;
    int device = 0;
    int light_items = 0;
    int dataset_items = 0;      // 128-byte items (reference convention)
    uint32_t dag_items = 0;     // 256-byte items = dataset_items / 2
    int rank = 0, world = 1;
    uint32_t share_items = 0;   // items per rank in the padded layout
    bool ready = false;
    uint8_t* light_host = nullptr;
    uint32_t l1_host[4096];
    void* light_dev = nullptr;
    void* dag_dev = nullptr;
    size_t dag_padded_bytes = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // search/batch scratch owned by the context
    uint64_t* found_min_dev = nullptr;
    uint64_t* found_min_host = nullptr;  // pinned
    uint64_t* checksum_partials_dev = nullptr;
    void* batch_out_dev = nullptr;
    size_t batch_out_bytes = 0;
    bool deferred = false;     // slice not generated yet (waiting for nx_epoch_generate_fused*)
    bool light_pending = false;  // created by nx_epoch_create_shell: the light cache has not arrived yet
    std::vector<void*> ipc_opened;  // peer DAG buffers opened with cudaIpcOpenMemHandle
    std::vector<void*> ipc_bases;   // the same, indexed by rank (own entry null), once nx_epoch_open_ipc_peers has run
    void* many_dev = nullptr;  // records + packed programs of nx_kawpow_hash_many
    size_t many_bytes = 0;
    // bookkeeping for enqueue/sync
    uint64_t pending_base = 0;  // start nonce that offset 0 of found_min refers to
    bool pending = false;
    int pending_block = 0;
    uint8_t pending_header[32];
    int launches_since_sync = 0;
    float last_kernel_ms = 0.f;
    int last_launches = 0;
    double light_ms = 0, dag_ms = 0;
    bool l2_pinned = false;       // light cache currently marked persisting in L2 (DAG generation only)
    size_t l2_prev_limit = 0;     // device's persisting-L2 carve-out before we raised it; restored on unpin
    std::recursive_mutex mu;  // serialises use of the stream + scratch; recursive so that a whole search
                              // (enqueue, sync, re-hash of the hit) is one critical section
};
