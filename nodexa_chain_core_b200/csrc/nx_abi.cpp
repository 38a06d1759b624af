// extern "C" entry points that own device state: epoch contexts and the KawPow calls.
// (Pure host entry points live in nx_host.cpp.)
#include "nx_internal.hpp"

#include <chrono>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <map>
#include <vector>
#include <new>

using namespace nx;

namespace
{
constexpr uint64_t NOT_FOUND = ~0ull;
constexpr uint64_t SEARCH_BATCH = 1ull << 26;  // nonces per launch inside nx_kawpow_search (0.34 s of work on a B200)
constexpr int SEARCH_QUEUE = 8;                  // launches queued between two reads of the result

inline uint32_t bswap32(uint32_t x) { return __builtin_bswap32(x); }

void fill_job(KawpowJob& job, const uint8_t header[32], const uint8_t* boundary, uint64_t start, uint64_t count,
              uint64_t report_base)
{
    std::memcpy(job.header, header, 32);
    if (boundary) {
        uint32_t w[8];
        std::memcpy(w, boundary, 32);
        for (int i = 0; i < 8; ++i) job.boundary_be[i] = bswap32(w[i]);
    } else {
        std::memset(job.boundary_be, 0, 32);
    }
    job.start_nonce = start;
    job.count = count;
    job.report_base = report_base;
}

int ensure_batch_out(nx_epoch* c, size_t bytes)
{
    if (c->batch_out_bytes >= bytes) return NX_OK;
    if (c->batch_out_dev) cudaFree(c->batch_out_dev);
    c->batch_out_dev = nullptr;
    c->batch_out_bytes = 0;
    if (cudaMalloc(&c->batch_out_dev, bytes) != cudaSuccess) {
        cudaGetLastError();
        set_error("out of device memory for %zu result bytes", bytes);
        return NX_ERR_OOM;
    }
    c->batch_out_bytes = bytes;
    return NX_OK;
}

int check_ctx(const nx_epoch* c, int block_number)
{
    if (!c) {
        set_error("null epoch context");
        return NX_ERR_INVALID;
    }
    if (!c->ready) {
        set_error("epoch context not finalized (sharded DAG still waiting for its all-gather)");
        return NX_ERR_STATE;
    }
    if (block_number < 0) {
        set_error("negative block number");
        return NX_ERR_INVALID;
    }
    return NX_OK;
}

// K3 gathers 512 random 64-byte parents per item from the light cache (16-80 MiB).  It only just fits the 126 MB
// L2, and on a two-die part a plain LRU keeps ~76 % of it resident; marking the window as persisting while the DAG
// is generated keeps the gather in L2.  Best effort: unsupported or refused settings are ignored.
void l2_pin_light_cache(nx_epoch* c, bool on)
{
    if (const char* e = std::getenv("NX_NO_L2_PIN"))
        if (*e && *e != '0') return;
    cudaStreamAttrValue attr;
    std::memset(&attr, 0, sizeof attr);
    if (on) {
        int max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, c->device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, c->device);
        if (max_persist <= 0 || max_window <= 0) return;
        const size_t light_bytes = (size_t)c->light_items * 64;
        if (!c->l2_pinned) cudaDeviceGetLimit(&c->l2_prev_limit, cudaLimitPersistingL2CacheSize);
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist);
        c->l2_pinned = true;
        const size_t window = light_bytes < (size_t)max_window ? light_bytes : (size_t)max_window;
        attr.accessPolicyWindow.base_ptr = c->light_dev;
        attr.accessPolicyWindow.num_bytes = window;
        attr.accessPolicyWindow.hitRatio = (float)((size_t)max_persist >= window ? 1.0 : (double)max_persist / (double)window);
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    } else {
        if (!c->l2_pinned) return;
        attr.accessPolicyWindow.num_bytes = 0;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
    }
    cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &attr);
    if (!on) {  // hand the L2 back exactly as the host process had it
        cudaCtxResetPersistingL2Cache();
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, c->l2_prev_limit);
        c->l2_pinned = false;
    }
    cudaGetLastError();
}

struct L2Pin  // pinned for the scope, released on every exit path
{
    nx_epoch* c;
    explicit L2Pin(nx_epoch* ctx) : c(ctx) { l2_pin_light_cache(c, true); }
    ~L2Pin() { l2_pin_light_cache(c, false); }
};

int program_for(const nx_epoch* c, int block_number, std::shared_ptr<KawpowProgram>* p)
{
    return kawpow_get_program(c->device, (uint64_t)(block_number / NX_PERIOD_LENGTH), c->dag_items, p);
}

void destroy(nx_epoch* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->l2_pinned) l2_pin_light_cache(c, false);  // a create() that failed between pin and unpin
    for (void* p : c->ipc_opened) cudaIpcCloseMemHandle(p);
    if (c->dag_dev) cudaFree(c->dag_dev);
    if (c->light_dev) cudaFree(c->light_dev);
    if (c->found_min_dev) cudaFree(c->found_min_dev);
    if (c->checksum_partials_dev) cudaFree(c->checksum_partials_dev);
    if (c->batch_out_dev) cudaFree(c->batch_out_dev);
    if (c->many_dev) cudaFree(c->many_dev);
    if (c->found_min_host) cudaFreeHost(c->found_min_host);
    if (c->light_host) cudaFreeHost(c->light_host);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

// light_later: allocate only; the light cache arrives through nx_epoch_build_light_cache (the rank that runs the
// serial chain) or nx_epoch_adopt_light_cache (the ranks it is broadcast to).  Implies deferred.
int create(int epoch, int device, int rank, int world, const uint8_t* light_in, bool deferred, nx_epoch** out,
           bool light_later = false)
{
    clear_error();
    if (!out) {
        set_error("null out pointer");
        return NX_ERR_INVALID;
    }
    *out = nullptr;
    if (epoch < 0 || epoch > 30000 || world < 1 || rank < 0 || rank >= world) {
        set_error("bad epoch/rank/world (%d, %d, %d)", epoch, rank, world);
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

    nx_epoch* c = new (std::nothrow) nx_epoch;
    if (!c) return NX_ERR_OOM;
    c->epoch = epoch;
    c->device = device;
    c->rank = rank;
    c->world = world;
    c->light_items = nx_light_cache_num_items(epoch);
    c->dataset_items = nx_full_dataset_num_items(epoch);
    c->dag_items = (uint32_t)(c->dataset_items / 2);
    c->share_items = (c->dag_items + (uint32_t)world - 1) / (uint32_t)world;
    c->dag_padded_bytes = (size_t)c->share_items * (size_t)world * 256;
    const size_t light_bytes = (size_t)c->light_items * 64;

    auto fail = [&](int rc) {
        destroy(c);
        return rc;
    };
#define NX_TRY(call)                                                                               \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            set_error("%s failed: %s", #call, cudaGetErrorString(e_));                             \
            cudaGetLastError();                                                                    \
            return fail(e_ == cudaErrorMemoryAllocation ? NX_ERR_OOM : NX_ERR_CUDA);               \
        }                                                                                          \
    } while (0)

    NX_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    NX_TRY(cudaEventCreate(&c->ev0));
    NX_TRY(cudaEventCreate(&c->ev1));
    // the DAG first: it is the allocation that fails when an epoch does not fit the device, and failing here costs nothing
    NX_TRY(cudaMalloc(&c->dag_dev, c->dag_padded_bytes + 256));  // +1 item: classic Ethash reaches 128-byte item n-1
    NX_TRY(cudaMalloc(&c->light_dev, light_bytes));
    NX_TRY(cudaMallocHost((void**)&c->light_host, light_bytes));
    NX_TRY(cudaMalloc((void**)&c->found_min_dev, 2 * sizeof(uint64_t)));
    NX_TRY(cudaMallocHost((void**)&c->found_min_host, 2 * sizeof(uint64_t)));
    NX_TRY(cudaMalloc((void**)&c->checksum_partials_dev, CHECKSUM_PARTIALS * sizeof(uint64_t)));

    // ---- light cache: serial chain.  Host by default; NX_LIGHT_CACHE=gpu runs kernel K4. ----
    uint8_t seed[32];
    nx_epoch_seed(epoch, seed);
    const char* mode = std::getenv("NX_LIGHT_CACHE");
    const auto t0 = std::chrono::steady_clock::now();
    if (light_later) {
        c->light_pending = true;
    } else if (light_in) {  // caller already ran the serial chain (one host build shared by several devices)
        std::memcpy(c->light_host, light_in, light_bytes);
        NX_TRY(cudaMemcpyAsync(c->light_dev, c->light_host, light_bytes, cudaMemcpyHostToDevice, c->stream));
    } else if (mode && std::strcmp(mode, "gpu") == 0) {
        int rc = light_cache_build_gpu(c->light_dev, c->light_items, seed, c->stream);
        if (rc != NX_OK) return fail(rc);
        NX_TRY(cudaMemcpyAsync(c->light_host, c->light_dev, light_bytes, cudaMemcpyDeviceToHost, c->stream));
        NX_TRY(cudaStreamSynchronize(c->stream));
    } else {
        if (!host_load_light_cache(c->light_host, c->light_items, epoch)) {  // NX_CACHE_DIR: reuse an earlier build
            host_build_light_cache(c->light_host, c->light_items, seed);
            host_store_light_cache(c->light_host, c->light_items, epoch);
        }
        NX_TRY(cudaMemcpyAsync(c->light_dev, c->light_host, light_bytes, cudaMemcpyHostToDevice, c->stream));
    }
    c->light_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();

    // ---- DAG: this rank's slice, written at its final offset ----
    const uint32_t first = (uint32_t)rank * c->share_items;
    const uint32_t last = first + c->share_items < c->dag_items ? first + c->share_items : c->dag_items;
    const uint32_t count = last > first ? last - first : 0;
    if (c->dag_padded_bytes > (size_t)c->dag_items * 256)  // padding tail is never addressed; keep it defined
        NX_TRY(cudaMemsetAsync((char*)c->dag_dev + (size_t)c->dag_items * 256, 0,
                               c->dag_padded_bytes - (size_t)c->dag_items * 256, c->stream));
    c->deferred = deferred;
    if (!deferred) {
        l2_pin_light_cache(c, true);
        NX_TRY(cudaEventRecord(c->ev0, c->stream));
        int rc = dag_generate(c->light_dev, c->light_items, first, count, (char*)c->dag_dev + (size_t)first * 256,
                              c->stream);
        if (rc != NX_OK) return fail(rc);
        NX_TRY(cudaEventRecord(c->ev1, c->stream));
    }
    NX_TRY(cudaStreamSynchronize(c->stream));
    if (!deferred) l2_pin_light_cache(c, false);
    if (!deferred) {
        float ms = 0;
        NX_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        c->dag_ms = ms;
    }
#undef NX_TRY

    *out = c;
    if (world == 1) return nx_epoch_finalize(c);
    return NX_OK;
}
}  // namespace

extern "C" {

int nx_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int nx_epoch_create(int epoch, int device, nx_epoch** out) { return create(epoch, device, 0, 1, nullptr, false, out); }

int nx_epoch_create_from_light(int epoch, int device, const uint8_t* light_cache, nx_epoch** out)
{
    if (!light_cache) {
        set_error("null light cache");
        return NX_ERR_INVALID;
    }
    return create(epoch, device, 0, 1, light_cache, false, out);
}

int nx_epoch_create_sharded(int epoch, int device, int rank, int world, nx_epoch** out)
{
    return create(epoch, device, rank, world, nullptr, false, out);
}

int nx_epoch_create_deferred(int epoch, int device, int rank, int world, const uint8_t* light_cache, nx_epoch** out)
{
    return create(epoch, device, rank, world, light_cache, true, out);
}

int nx_epoch_create_shell(int epoch, int device, int rank, int world, nx_epoch** out)
{
    return create(epoch, device, rank, world, nullptr, true, out, true);
}

int nx_epoch_build_light_cache(nx_epoch* c)
{
    if (!c) {
        set_error("null epoch context");
        return NX_ERR_INVALID;
    }
    if (!c->light_pending) {
        set_error("this context already has its light cache");
        return NX_ERR_STATE;
    }
    // the serial chain touches only light_host: the context stays usable from other threads (IPC set-up) meanwhile
    uint8_t seed[32];
    nx_epoch_seed(c->epoch, seed);
    const auto t0 = std::chrono::steady_clock::now();
    const char* mode = std::getenv("NX_LIGHT_CACHE");
    const size_t light_bytes = (size_t)c->light_items * 64;
    const bool on_gpu = mode && std::strcmp(mode, "gpu") == 0;
    if (!on_gpu && !host_load_light_cache(c->light_host, c->light_items, c->epoch)) {
        host_build_light_cache(c->light_host, c->light_items, seed);
        host_store_light_cache(c->light_host, c->light_items, c->epoch);
    }
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    if (on_gpu) {
        const int rc = light_cache_build_gpu(c->light_dev, c->light_items, seed, c->stream);
        if (rc != NX_OK) return rc;
        NX_CUDA(cudaMemcpyAsync(c->light_host, c->light_dev, light_bytes, cudaMemcpyDeviceToHost, c->stream));
    } else {
        NX_CUDA(cudaMemcpyAsync(c->light_dev, c->light_host, light_bytes, cudaMemcpyHostToDevice, c->stream));
    }
    NX_CUDA(cudaStreamSynchronize(c->stream));
    c->light_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->light_pending = false;
    return NX_OK;
}

void* nx_epoch_light_ptr(const nx_epoch* c) { return c ? c->light_dev : nullptr; }

int nx_epoch_adopt_light_cache(nx_epoch* c)
{
    if (!c) {
        set_error("null epoch context");
        return NX_ERR_INVALID;
    }
    if (!c->light_pending) {
        set_error("this context already has its light cache");
        return NX_ERR_STATE;
    }
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    const auto t0 = std::chrono::steady_clock::now();
    // the device buffer was filled by the caller (a broadcast from the rank that ran the chain): mirror it for the
    // reference-visible light_cache field
    NX_CUDA(cudaMemcpyAsync(c->light_host, c->light_dev, (size_t)c->light_items * 64, cudaMemcpyDeviceToHost, c->stream));
    NX_CUDA(cudaStreamSynchronize(c->stream));
    c->light_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->light_pending = false;
    return NX_OK;
}

int nx_epoch_ipc_handle(const nx_epoch* c, uint8_t handle[64])
{
    if (!c || !handle) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    NX_CUDA(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    NX_CUDA(cudaIpcGetMemHandle(&h, c->dag_dev));
    std::memcpy(handle, &h, 64);
    return NX_OK;
}

// Generate this rank's slice and store it into every rank's DAG buffer in the same kernel.
// peer_bases == nullptr: the slice goes into the own buffer only (the caller all-gathers afterwards).
static int generate_fused(nx_epoch* c, void* const* peer_bases /* world entries; [rank] may be null */)
{
    if (!c->deferred) {
        set_error("context was not created with nx_epoch_create_deferred (its slice is already generated)");
        return NX_ERR_STATE;
    }
    if (c->light_pending) {
        set_error("this context has no light cache yet (nx_epoch_build_light_cache / nx_epoch_adopt_light_cache come first)");
        return NX_ERR_STATE;
    }
    if (c->world > NX_MAX_PEERS) {
        set_error("world %d exceeds the %d peers one launch can address", c->world, NX_MAX_PEERS);
        return NX_ERR_INVALID;
    }
    const uint32_t first = (uint32_t)c->rank * c->share_items;
    const uint32_t last = first + c->share_items < c->dag_items ? first + c->share_items : c->dag_items;
    const uint32_t count = last > first ? last - first : 0;
    void* outs[NX_MAX_PEERS];
    int n = 0;
    outs[n++] = (char*)c->dag_dev + (size_t)first * 256;  // local copy first
    for (int r = 0; peer_bases && r < c->world; ++r) {
        if (r == c->rank) continue;
        if (!peer_bases[r]) {  // e.g. a rank whose context could not be created: never offset a null pointer
            set_error("no DAG buffer given for rank %d", r);
            return NX_ERR_INVALID;
        }
        outs[n++] = (char*)peer_bases[r] + (size_t)first * 256;
    }
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    {
        L2Pin pin(c);
        NX_CUDA(cudaEventRecord(c->ev0, c->stream));
        const int rc = dag_generate_multi(c->light_dev, c->light_items, first, count, outs, n, c->stream);
        if (rc != NX_OK) return rc;
        NX_CUDA(cudaEventRecord(c->ev1, c->stream));
        NX_CUDA(cudaStreamSynchronize(c->stream));
    }
    float ms = 0;
    NX_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->dag_ms = ms;
    c->deferred = false;
    return NX_OK;
}

// Map every other rank's DAG buffer into this process (cudaIpcOpenMemHandle, tens of ms each for a multi-GiB buffer).
// Needs no light cache, so a multi-process build does it while the one serial chain is still running.
static int open_ipc_peers(nx_epoch* c, const uint8_t* handles /* world x 64 bytes, indexed by rank */)
{
    if (!c->ipc_bases.empty()) return NX_OK;  // already open
    NX_CUDA(cudaSetDevice(c->device));
    std::vector<void*> bases((size_t)c->world, nullptr);
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + 64 * (size_t)r, 64);
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            set_error("cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e));
            cudaGetLastError();
            return NX_ERR_CUDA;
        }
        c->ipc_opened.push_back(p);
        bases[(size_t)r] = p;
    }
    c->ipc_bases = bases;
    return NX_OK;
}

static int nx_epoch_generate_fused_ipc_body(nx_epoch* c, const uint8_t* handles /* may be null after nx_epoch_open_ipc_peers */)
{
    if (!c || (!handles && c->ipc_bases.empty())) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (c->ipc_bases.empty()) {
        const int rc = open_ipc_peers(c, handles);
        if (rc != NX_OK) return rc;
    }
    return generate_fused(c, c->ipc_bases.data());
}

int nx_epoch_open_ipc_peers(nx_epoch* c, const uint8_t* handles)
{
    if (!c || !handles) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    return guarded([&] { return open_ipc_peers(c, handles); });
}

int nx_epoch_generate_slice(nx_epoch* c)
{
    if (!c) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    return generate_fused(c, nullptr);  // own buffer only: the exchange is the caller's in-place all-gather
}

int nx_epoch_generate_fused_peers(nx_epoch* c, void* const* peer_dag_ptrs /* world entries, same process */)
{
    if (!c || !peer_dag_ptrs) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    return generate_fused(c, peer_dag_ptrs);
}

int nx_epoch_finalize(nx_epoch* c)
{
    if (!c) {
        set_error("null epoch context");
        return NX_ERR_INVALID;
    }
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    if (c->deferred) {
        set_error("this context's DAG slice has not been generated yet (nx_epoch_generate_fused_* comes first)");
        return NX_ERR_STATE;
    }
    NX_CUDA(cudaSetDevice(c->device));
    // KawPow addresses floor(n/2) 256-byte items; classic Ethash also reaches the odd last 128-byte item
    // n-1, i.e. the first half of 256-byte item floor(n/2): generate that one item behind the rest.
    {
        const int rc = dag_generate(c->light_dev, c->light_items, c->dag_items, 1,
                                    (char*)c->dag_dev + (size_t)c->dag_items * 256, c->stream);
        if (rc != NX_OK) return rc;
        NX_CUDA(cudaStreamSynchronize(c->stream));
    }
    // L1 cache mirror = first 16 KiB of the DAG (reference ethash.cpp:168-170)
    NX_CUDA(cudaMemcpy(c->l1_host, c->dag_dev, sizeof c->l1_host, cudaMemcpyDeviceToHost));
    c->ready = true;
    return NX_OK;
}

void nx_epoch_destroy(nx_epoch* c) { destroy(c); }

int nx_epoch_number(const nx_epoch* c) { return c ? c->epoch : -1; }
int nx_epoch_light_items(const nx_epoch* c) { return c ? c->light_items : 0; }
int nx_epoch_dataset_items(const nx_epoch* c) { return c ? c->dataset_items : 0; }
const uint8_t* nx_epoch_light_cache(const nx_epoch* c) { return c ? c->light_host : nullptr; }
const uint32_t* nx_epoch_l1_cache(const nx_epoch* c) { return c ? c->l1_host : nullptr; }
int nx_epoch_device(const nx_epoch* c) { return c ? c->device : -1; }
void* nx_epoch_dag_ptr(const nx_epoch* c) { return c ? c->dag_dev : nullptr; }
size_t nx_epoch_dag_bytes(const nx_epoch* c) { return c ? (size_t)c->dag_items * 256 : 0; }
size_t nx_epoch_dag_padded_bytes(const nx_epoch* c) { return c ? c->dag_padded_bytes : 0; }
size_t nx_epoch_dag_share_bytes(const nx_epoch* c) { return c ? (size_t)c->share_items * 256 : 0; }
void nx_epoch_build_times(const nx_epoch* c, double* light_ms, double* dag_ms)
{
    if (light_ms) *light_ms = c ? c->light_ms : 0;
    if (dag_ms) *dag_ms = c ? c->dag_ms : 0;
}

int nx_dag_generate_range(const nx_epoch* cc, uint32_t first, uint32_t count, void* out_device)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    if (!c || !out_device) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if ((uint64_t)first + count > c->dag_items) {
        set_error("item range [%u,+%u) beyond the DAG (%u items)", first, count, c->dag_items);
        return NX_ERR_INVALID;
    }
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    {
        L2Pin pin(c);
        NX_CUDA(cudaEventRecord(c->ev0, c->stream));
        int rc = dag_generate(c->light_dev, c->light_items, first, count, out_device, c->stream);
        if (rc != NX_OK) return rc;
        NX_CUDA(cudaEventRecord(c->ev1, c->stream));
        NX_CUDA(cudaStreamSynchronize(c->stream));
    }
    NX_CUDA(cudaEventElapsedTime(&c->last_kernel_ms, c->ev0, c->ev1));
    c->last_launches = 1;
    return NX_OK;
}

int nx_dag_read_items(const nx_epoch* c, uint32_t first, uint32_t count, uint8_t* out_host)
{
    if (!c || !out_host) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if ((uint64_t)first + count > (uint64_t)c->share_items * c->world + 1) {  // +1: classic Ethash's extra item
        set_error("item range beyond the DAG buffer");
        return NX_ERR_INVALID;
    }
    NX_CUDA(cudaSetDevice(c->device));
    NX_CUDA(cudaMemcpy(out_host, (const char*)c->dag_dev + (size_t)first * 256, (size_t)count * 256,
                       cudaMemcpyDeviceToHost));
    return NX_OK;
}

int nx_dag_checksum(const nx_epoch* cc, uint64_t* out)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    if (!c || !out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    int rc = dag_checksum(c->dag_dev, (size_t)c->dag_items * 32, c->checksum_partials_dev, CHECKSUM_PARTIALS,
                          c->found_min_dev + 1, c->stream);
    if (rc != NX_OK) return rc;
    NX_CUDA(cudaMemcpyAsync(c->found_min_host + 1, c->found_min_dev + 1, 8, cudaMemcpyDeviceToHost, c->stream));
    NX_CUDA(cudaStreamSynchronize(c->stream));
    *out = c->found_min_host[1];
    return NX_OK;
}

int nx_kawpow_prepare(const nx_epoch* c, int block_number)
{
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    std::shared_ptr<KawpowProgram> p;
    return program_for(c, block_number, &p);
}

int nx_kawpow_hash_batch_device(const nx_epoch* cc, int block_number, const uint8_t header_hash[32],
                                uint64_t start_nonce, uint64_t count, void* out_device)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    if (!header_hash || (!out_device && count)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (count == 0) return NX_OK;
    std::shared_ptr<KawpowProgram> p;
    if ((rc = program_for(c, block_number, &p)) != NX_OK) return rc;
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    KawpowJob job;
    fill_job(job, header_hash, nullptr, start_nonce, count, 0);
    NX_CUDA(cudaEventRecord(c->ev0, c->stream));
    if ((rc = kawpow_launch_hash(p.get(), c->dag_dev, job, out_device, c->stream)) != NX_OK) return rc;
    NX_CUDA(cudaEventRecord(c->ev1, c->stream));
    NX_CUDA(cudaStreamSynchronize(c->stream));
    NX_CUDA(cudaEventElapsedTime(&c->last_kernel_ms, c->ev0, c->ev1));
    c->last_launches = 1;
    return NX_OK;
}

int nx_kawpow_hash_batch(const nx_epoch* cc, int block_number, const uint8_t header_hash[32], uint64_t start_nonce,
                         uint64_t count, uint8_t* out_host)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    if (!header_hash || (!out_host && count)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (count == 0) return NX_OK;
    std::shared_ptr<KawpowProgram> p;
    if ((rc = program_for(c, block_number, &p)) != NX_OK) return rc;
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    constexpr uint64_t CHUNK = 1ull << 22;  // 256 MiB of results per launch
    if ((rc = ensure_batch_out(c, (size_t)(count < CHUNK ? count : CHUNK) * 64)) != NX_OK) return rc;
    float total_ms = 0;
    int launches = 0;
    for (uint64_t done = 0; done < count; done += CHUNK) {
        const uint64_t n = count - done < CHUNK ? count - done : CHUNK;
        KawpowJob job;
        fill_job(job, header_hash, nullptr, start_nonce + done, n, 0);
        NX_CUDA(cudaEventRecord(c->ev0, c->stream));
        if ((rc = kawpow_launch_hash(p.get(), c->dag_dev, job, c->batch_out_dev, c->stream)) != NX_OK) return rc;
        NX_CUDA(cudaEventRecord(c->ev1, c->stream));
        NX_CUDA(cudaMemcpyAsync(out_host + done * 64, c->batch_out_dev, (size_t)n * 64, cudaMemcpyDeviceToHost, c->stream));
        NX_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0;
        NX_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        total_ms += ms;
        ++launches;
    }
    c->last_kernel_ms = total_ms;
    c->last_launches = launches;
    return NX_OK;
}

int nx_kawpow_hash(const nx_epoch* c, int block_number, const uint8_t header_hash[32], uint64_t nonce, nx_result* out)
{
    if (!out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    uint8_t buf[64];
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    if (!header_hash) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    // A one-off hash (the node's verify path, the re-hash of a search hit) goes through the interpreted kernel K6:
    // no NVRTC wait, and one half-warp finishes one nonce in 0.3 ms where the compiled kernel - whose threads each
    // walk 16 nonce slots - needs a whole 0.8 ms wave.  Only a block outside this context's epoch (which K6's
    // record check refuses) takes the compiled path, as the reference lets context and height disagree.
    if (block_number / NX_EPOCH_LENGTH != c->epoch) {
        rc = nx_kawpow_hash_batch(c, block_number, header_hash, nonce, 1, buf);
    } else {
        nx_hash_record r;
        r.block_number = block_number;
        r.pad = 0;
        r.nonce = nonce;
        std::memcpy(r.header_hash, header_hash, 32);
        rc = nx_kawpow_hash_many(c, &r, 1, buf);
    }
    if (rc != NX_OK) return rc;
    std::memcpy(out->final_hash, buf, 32);
    std::memcpy(out->mix_hash, buf + 32, 32);
    return NX_OK;
}

int nx_ethash_hash_batch(const nx_epoch* cc, const uint8_t header_hash[32], uint64_t start_nonce, uint64_t count,
                         uint8_t* out_host)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    int rc = check_ctx(c, 0);
    if (rc != NX_OK) return rc;
    if (!header_hash || (!out_host && count)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (count == 0) return NX_OK;
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    constexpr uint64_t CHUNK = 1ull << 22;
    if ((rc = ensure_batch_out(c, (size_t)(count < CHUNK ? count : CHUNK) * 64)) != NX_OK) return rc;
    float total_ms = 0;
    int launches = 0;
    for (uint64_t done = 0; done < count; done += CHUNK) {
        const uint64_t n = count - done < CHUNK ? count - done : CHUNK;
        NX_CUDA(cudaEventRecord(c->ev0, c->stream));
        if ((rc = ethash_hash_batch(c->dag_dev, (uint32_t)c->dataset_items, header_hash, start_nonce + done, n,
                                    c->batch_out_dev, c->stream)) != NX_OK)
            return rc;
        NX_CUDA(cudaEventRecord(c->ev1, c->stream));
        NX_CUDA(cudaMemcpyAsync(out_host + done * 64, c->batch_out_dev, (size_t)n * 64, cudaMemcpyDeviceToHost, c->stream));
        NX_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0;
        NX_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
        total_ms += ms;
        ++launches;
    }
    c->last_kernel_ms = total_ms;
    c->last_launches = launches;
    return NX_OK;
}

int nx_ethash_hash(const nx_epoch* c, const uint8_t header_hash[32], uint64_t nonce, nx_result* out)
{
    if (!out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    uint8_t buf[64];
    const int rc = nx_ethash_hash_batch(c, header_hash, nonce, 1, buf);
    if (rc != NX_OK) return rc;
    std::memcpy(out->final_hash, buf, 32);
    std::memcpy(out->mix_hash, buf + 32, 32);
    return NX_OK;
}

int nx_kawpow_search_enqueue(const nx_epoch* cc, int block_number, const uint8_t header_hash[32],
                             const uint8_t boundary[32], uint64_t start_nonce, uint64_t count)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    if (!header_hash || !boundary) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (count == 0) return NX_OK;
    std::shared_ptr<KawpowProgram> p;
    if ((rc = program_for(c, block_number, &p)) != NX_OK) return rc;
    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    if (!c->pending) {
        c->found_min_host[0] = NOT_FOUND;
        NX_CUDA(cudaMemcpyAsync(c->found_min_dev, c->found_min_host, 8, cudaMemcpyHostToDevice, c->stream));
        NX_CUDA(cudaEventRecord(c->ev0, c->stream));
        c->pending = true;
        c->pending_base = start_nonce;
        c->pending_block = block_number;
        std::memcpy(c->pending_header, header_hash, 32);
        c->launches_since_sync = 0;
    } else if (block_number != c->pending_block || std::memcmp(header_hash, c->pending_header, 32) != 0 ||
               start_nonce < c->pending_base) {
        set_error("batches enqueued between two syncs must share header and block number and ascend in nonce");
        return NX_ERR_STATE;
    }
    KawpowJob job;
    fill_job(job, header_hash, boundary, start_nonce, count, start_nonce - c->pending_base);
    if ((rc = kawpow_launch_search(p.get(), c->dag_dev, job, c->found_min_dev, c->stream)) != NX_OK) return rc;
    ++c->launches_since_sync;
    return NX_OK;
}

int nx_kawpow_search_sync(const nx_epoch* cc, nx_search_result* out)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    if (!c || !out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    std::memset(out, 0, sizeof *out);
    uint64_t hit = NOT_FOUND;
    int block = 0;
    uint8_t header[32];
    {
        std::lock_guard<std::recursive_mutex> lk(c->mu);
        if (!c->pending) return NX_OK;
        NX_CUDA(cudaSetDevice(c->device));
        NX_CUDA(cudaEventRecord(c->ev1, c->stream));
        NX_CUDA(cudaMemcpyAsync(c->found_min_host, c->found_min_dev, 8, cudaMemcpyDeviceToHost, c->stream));
        NX_CUDA(cudaStreamSynchronize(c->stream));
        NX_CUDA(cudaEventElapsedTime(&c->last_kernel_ms, c->ev0, c->ev1));
        c->last_launches = c->launches_since_sync;
        c->pending = false;
        if (c->found_min_host[0] != NOT_FOUND) hit = c->pending_base + c->found_min_host[0];
        block = c->pending_block;
        std::memcpy(header, c->pending_header, 32);
        if (c->found_min_host[0] == NOT_FOUND) return NX_OK;
    }
    // a hit: re-hash that one nonce to obtain {final, mix} (keeps the hot kernel free of result traffic)
    nx_result r;
    const float keep_ms = c->last_kernel_ms;
    const int keep_n = c->last_launches;
    const int rc = nx_kawpow_hash(c, block, header, hit, &r);
    c->last_kernel_ms = keep_ms;
    c->last_launches = keep_n + 1;
    if (rc != NX_OK) return rc;
    out->found = 1;
    out->nonce = hit;
    std::memcpy(out->final_hash, r.final_hash, 32);
    std::memcpy(out->mix_hash, r.mix_hash, 32);
    return NX_OK;
}

int nx_kawpow_last_timing(const nx_epoch* c, float* kernel_ms, int* launches)
{
    if (!c) return NX_ERR_INVALID;
    if (kernel_ms) *kernel_ms = c->last_kernel_ms;
    if (launches) *launches = c->last_launches;
    return NX_OK;
}

int nx_kawpow_search(const nx_epoch* c, int block_number, const uint8_t header_hash[32], const uint8_t boundary[32],
                     uint64_t start_nonce, uint64_t iterations, nx_search_result* out)
{
    if (!out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    std::memset(out, 0, sizeof *out);
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    // ascending batches; the first batch with a hit holds the lowest qualifying nonce
    // (reference progpow.cpp:572-577 returns the first hit of an ascending scan).  The context is held for the
    // whole scan so that concurrent searches from other host threads cannot interleave their batches with ours.
    // The reference's loop is `for (nonce = start; nonce < start + iterations; ++nonce)` in uint64_t: when the
    // end wraps past 2^64 it is not above the start and nothing is scanned.  Same here.
    if (start_nonce + iterations <= start_nonce) return NX_OK;
    std::lock_guard<std::recursive_mutex> hold(const_cast<nx_epoch*>(c)->mu);
    // Launches are queued SEARCH_QUEUE at a time and the result is read once per group: the kernel itself stops
    // scanning above a reported nonce (every warp looks at the running minimum before each unit), so after a hit the
    // launches still in the queue return at once and nothing is gained by a host round trip per launch.
    float total_ms = 0;
    int launches = 0;
    uint64_t done = 0;
    while (done < iterations) {
        for (int q = 0; q < SEARCH_QUEUE && done < iterations; ++q) {
            const uint64_t n = iterations - done < SEARCH_BATCH ? iterations - done : SEARCH_BATCH;
            if ((rc = nx_kawpow_search_enqueue(c, block_number, header_hash, boundary, start_nonce + done, n)) != NX_OK) {
                nx_search_result drop;
                nx_kawpow_search_sync(c, &drop);  // leave no launch pending behind an error
                return rc;
            }
            done += n;  // ends exactly at `iterations`: no wrap for ranges that reach up to 2^64 - 1
        }
        if ((rc = nx_kawpow_search_sync(c, out)) != NX_OK) return rc;
        total_ms += c->last_kernel_ms;
        launches += c->last_launches;
        if (out->found) break;
    }
    nx_epoch* m = const_cast<nx_epoch*>(c);
    m->last_kernel_ms = total_ms;
    m->last_launches = launches;
    return NX_OK;
}

int nx_kawpow_verify(const nx_epoch* c, int block_number, const uint8_t header_hash[32], const uint8_t mix_hash[32],
                     uint64_t nonce, const uint8_t boundary[32])
{
    if (!header_hash || !mix_hash || !boundary) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    int rc = check_ctx(c, block_number);
    if (rc != NX_OK) return rc;
    // cheap test first, exactly as the reference orders it (progpow.cpp:487-494)
    uint8_t fin[32];
    nx_kawpow_hash_no_verify(block_number, header_hash, mix_hash, nonce, fin);
    if (std::memcmp(fin, boundary, 32) > 0) return 0;
    nx_result r;
    if ((rc = nx_kawpow_hash(c, block_number, header_hash, nonce, &r)) != NX_OK) return rc;
    return std::memcmp(r.mix_hash, mix_hash, 32) == 0 ? 1 : 0;
}

// Device-side records and packed programs for n caller records (all inside the context's epoch): one packed
// program per distinct period, recs[i].out_index = i.  Shared by nx_kawpow_hash_many and its guarded test twin.
static int build_interp_records(const nx_epoch* c, const nx_hash_record* records, size_t n, std::vector<KawpowRecord>& recs,
                                std::vector<uint32_t>& progs, size_t* n_periods)
{
    std::map<uint64_t, uint32_t> slot_of;
    recs.resize(n);
    for (size_t i = 0; i < n; ++i) {
        const int bn = records[i].block_number;
        if (bn < 0 || bn / NX_EPOCH_LENGTH != c->epoch) {
            set_error("record %zu: block %d is not in this context's epoch %d", i, bn, c->epoch);
            return NX_ERR_INVALID;
        }
        const uint64_t seed = (uint64_t)(bn / NX_PERIOD_LENGTH);
        auto it = slot_of.find(seed);
        if (it == slot_of.end()) it = slot_of.emplace(seed, (uint32_t)slot_of.size()).first;
        std::memcpy(recs[i].header, records[i].header_hash, 32);
        recs[i].nonce = records[i].nonce;
        recs[i].program_slot = it->second;
        recs[i].out_index = (uint32_t)i;
    }
    progs.assign(slot_of.size() * 36 * 4, 0u);
    for (const auto& kv : slot_of) kawpow_pack_program(kv.first, progs.data() + (size_t)kv.second * 36 * 4);
    *n_periods = slot_of.size();
    return NX_OK;
}

static int nx_kawpow_hash_many_body(const nx_epoch* cc, const nx_hash_record* records, size_t n, uint8_t* out_host)
{
    nx_epoch* c = const_cast<nx_epoch*>(cc);
    int rc = check_ctx(c, 0);
    if (rc != NX_OK) return rc;
    if ((!records || !out_host) && n) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (n == 0) return NX_OK;
    if (n > 0x7fffffffu / 16) {
        set_error("too many records in one call");
        return NX_ERR_INVALID;
    }
    std::vector<KawpowRecord> recs;
    std::vector<uint32_t> progs;
    size_t n_periods = 0;
    if ((rc = build_interp_records(c, records, n, recs, progs, &n_periods)) != NX_OK) return rc;
    // The two half-warps of a warp interpret one record each: when both belong to the same period they take the same
    // branches of the op dispatch.  Group the records by period (results still land in the caller's order).
    if (n_periods < n)
        std::stable_sort(recs.begin(), recs.end(),
                         [](const KawpowRecord& a, const KawpowRecord& b) { return a.program_slot < b.program_slot; });

    std::lock_guard<std::recursive_mutex> lk(c->mu);
    NX_CUDA(cudaSetDevice(c->device));
    const size_t rec_bytes = (n * sizeof(KawpowRecord) + 255) & ~(size_t)255;
    const size_t prog_bytes = (progs.size() * 4 + 255) & ~(size_t)255;
    const size_t need = rec_bytes + prog_bytes;
    if (c->many_bytes < need) {
        if (c->many_dev) cudaFree(c->many_dev);
        c->many_dev = nullptr;
        c->many_bytes = 0;
        if (cudaMalloc(&c->many_dev, need) != cudaSuccess) {
            cudaGetLastError();
            set_error("out of device memory for %zu record bytes", need);
            return NX_ERR_OOM;
        }
        c->many_bytes = need;
    }
    if ((rc = ensure_batch_out(c, n * 64)) != NX_OK) return rc;
    char* base = static_cast<char*>(c->many_dev);
    NX_CUDA(cudaMemcpyAsync(base, recs.data(), n * sizeof(KawpowRecord), cudaMemcpyHostToDevice, c->stream));
    NX_CUDA(cudaMemcpyAsync(base + rec_bytes, progs.data(), progs.size() * 4, cudaMemcpyHostToDevice, c->stream));
    NX_CUDA(cudaEventRecord(c->ev0, c->stream));
    if ((rc = kawpow_interp_launch(c->dag_dev, c->dag_items, reinterpret_cast<const KawpowRecord*>(base), (uint32_t)n,
                                   base + rec_bytes, c->batch_out_dev, c->stream)) != NX_OK)
        return rc;
    NX_CUDA(cudaEventRecord(c->ev1, c->stream));
    NX_CUDA(cudaMemcpyAsync(out_host, c->batch_out_dev, n * 64, cudaMemcpyDeviceToHost, c->stream));
    NX_CUDA(cudaStreamSynchronize(c->stream));  // recs/progs are pageable host vectors: keep them alive until here
    NX_CUDA(cudaEventElapsedTime(&c->last_kernel_ms, c->ev0, c->ev1));
    c->last_launches = 1;
    return NX_OK;
}

static int nx_check_block_headers_body(const nx_epoch* c, const uint8_t* raw, size_t n, const uint8_t pow_limit_be[32], int mix_only,
                           uint8_t* status)
{
    if ((!raw || !status) && n) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    int rc = check_ctx(c, 0);
    if (rc != NX_OK) return rc;
    std::vector<nx_hash_record> todo;
    std::vector<size_t> index;
    std::vector<uint8_t> targets;  // 32 bytes per surviving header
    for (size_t i = 0; i < n; ++i) {
        const uint8_t* h = raw + 120 * i;
        nx_block_header f;
        nx_block_header_parse(h, &f);
        uint8_t target[32];
        if (!nx_compact_to_boundary(f.bits, target) || (pow_limit_be && std::memcmp(target, pow_limit_be, 32) > 0)) {
            status[i] = NX_HEADER_BAD_BITS;
            continue;
        }
        if (mix_only) {  // below the last checkpoint: the final hash from the header's own mix is enough
            uint8_t id[32], fin[32];
            nx_block_hash(h, id);
            for (int k = 0; k < 32; ++k) fin[k] = id[31 - k];
            status[i] = std::memcmp(fin, target, 32) <= 0 ? NX_HEADER_OK : NX_HEADER_HIGH_HASH;
            continue;
        }
        if ((int)(f.height / NX_EPOCH_LENGTH) != c->epoch || (int32_t)f.height < 0) {
            status[i] = NX_HEADER_WRONG_EPOCH;
            continue;
        }
        nx_hash_record r;
        r.block_number = (int32_t)f.height;
        r.pad = 0;
        r.nonce = f.nonce64;
        uint8_t d[32];
        nx_sha256d(h, 80, d);
        for (int k = 0; k < 32; ++k) r.header_hash[k] = d[31 - k];
        todo.push_back(r);
        index.push_back(i);
        targets.insert(targets.end(), target, target + 32);
    }
    if (todo.empty()) return NX_OK;
    std::vector<uint8_t> out(todo.size() * 64);
    if ((rc = nx_kawpow_hash_many(c, todo.data(), todo.size(), out.data())) != NX_OK) return rc;
    for (size_t k = 0; k < todo.size(); ++k) {
        const uint8_t* fin = out.data() + 64 * k;
        const uint8_t* mix = fin + 32;
        const uint8_t* claimed = raw + 120 * index[k] + 88;  // uint256 storage order = reversed
        if (std::memcmp(fin, targets.data() + 32 * k, 32) > 0) {
            status[index[k]] = NX_HEADER_HIGH_HASH;
            continue;
        }
        bool same = true;
        for (int b = 0; b < 32; ++b) same &= mix[b] == claimed[31 - b];
        status[index[k]] = same ? NX_HEADER_OK : NX_HEADER_INVALID_MIX;
    }
    return NX_OK;
}

static int nx_kawpow_verify_batch_body(const nx_epoch* c, const nx_verify_record* records, size_t n, uint8_t* ok)
{
    if ((!records || !ok) && n) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    // progpow::verify's order (progpow.cpp:487-494): the cheap final-hash-vs-boundary test on the host first;
    // only records that pass it have their mix recomputed, all in one interpreted-kernel launch (K6).
    std::vector<nx_hash_record> todo;
    std::vector<size_t> index;
    for (size_t i = 0; i < n; ++i) {
        uint8_t fin[32];
        nx_kawpow_hash_no_verify(records[i].block_number, records[i].header_hash, records[i].mix_hash, records[i].nonce, fin);
        ok[i] = 0;
        if (std::memcmp(fin, records[i].boundary, 32) > 0) continue;
        nx_hash_record r;
        r.block_number = records[i].block_number;
        r.pad = 0;
        r.nonce = records[i].nonce;
        std::memcpy(r.header_hash, records[i].header_hash, 32);
        todo.push_back(r);
        index.push_back(i);
    }
    if (todo.empty()) return check_ctx(c, 0);
    std::vector<uint8_t> out(todo.size() * 64);
    const int rc = nx_kawpow_hash_many(c, todo.data(), todo.size(), out.data());
    if (rc != NX_OK) return rc;
    for (size_t k = 0; k < todo.size(); ++k)
        ok[index[k]] = std::memcmp(out.data() + 64 * k + 32, records[index[k]].mix_hash, 32) == 0 ? 1 : 0;
    return NX_OK;
}

int nx_epoch_generate_fused_ipc(nx_epoch* c, const uint8_t* handles /* world x 64 bytes, indexed by rank */)
{
    return guarded([&] { return nx_epoch_generate_fused_ipc_body(c, handles); });
}

// Test hook (not declared in include/: the tests bind it through ctypes).  compute-sanitizer is not available on every
// pool, so this runs the per-record kernel K6 on buffers of EXACTLY the bytes it may touch, each between two 4 KiB
// guard zones filled with a pattern, and reports how many guard bytes the kernel changed (0 = no out-of-bounds
// store).  Records and programs are laid out as nx_kawpow_hash_many lays them out; out_host receives n x 64 bytes.
int nx_test_hash_many_guarded(const nx_epoch* cc, const nx_hash_record* records, size_t n, uint8_t* out_host,
                              uint64_t* guard_bytes_changed)
{
    return guarded([&]() -> int {
        nx_epoch* c = const_cast<nx_epoch*>(cc);
        int rc = check_ctx(c, 0);
        if (rc != NX_OK) return rc;
        if (!records || !out_host || !guard_bytes_changed || n == 0 || n > (1u << 20)) {
            set_error("bad argument");
            return NX_ERR_INVALID;
        }
        std::vector<KawpowRecord> recs;
        std::vector<uint32_t> progs;
        size_t n_periods = 0;
        if ((rc = build_interp_records(c, records, n, recs, progs, &n_periods)) != NX_OK) return rc;

        constexpr size_t G = 4096;
        const size_t sizes[3] = {n * sizeof(KawpowRecord), progs.size() * 4, n * 64};
        size_t off[3], total = 0;
        for (int k = 0; k < 3; ++k) {
            total += G;
            off[k] = total;
            total += (sizes[k] + 255) & ~(size_t)255;  // the next guard starts right behind the rounded payload ...
        }
        total += G;
        std::lock_guard<std::recursive_mutex> lk(c->mu);
        NX_CUDA(cudaSetDevice(c->device));
        char* buf = nullptr;
        if (cudaMalloc((void**)&buf, total) != cudaSuccess) {
            cudaGetLastError();
            set_error("out of device memory");
            return NX_ERR_OOM;
        }
        std::vector<uint8_t> before(total, 0xA5), after(total);
        auto done = [&](int code) {
            cudaFree(buf);
            return code;
        };
        if (cudaMemcpy(buf, before.data(), total, cudaMemcpyHostToDevice) != cudaSuccess) return done(NX_ERR_CUDA);
        if (cudaMemcpy(buf + off[0], recs.data(), sizes[0], cudaMemcpyHostToDevice) != cudaSuccess) return done(NX_ERR_CUDA);
        if (cudaMemcpy(buf + off[1], progs.data(), sizes[1], cudaMemcpyHostToDevice) != cudaSuccess) return done(NX_ERR_CUDA);
        rc = kawpow_interp_launch(c->dag_dev, c->dag_items, reinterpret_cast<const KawpowRecord*>(buf + off[0]), (uint32_t)n,
                                  buf + off[1], buf + off[2], c->stream);
        if (rc != NX_OK) return done(rc);
        if (cudaStreamSynchronize(c->stream) != cudaSuccess || cudaMemcpy(after.data(), buf, total, cudaMemcpyDeviceToHost) != cudaSuccess) {
            set_error("guarded run failed: %s", cudaGetErrorString(cudaGetLastError()));
            return done(NX_ERR_CUDA);
        }
        std::memcpy(out_host, after.data() + off[2], sizes[2]);
        // ... so everything outside the three exact payloads - guards AND the rounding slack - must still hold the pattern
        uint64_t changed = 0;
        size_t pos = 0;
        for (int k = 0; k < 3; ++k) {
            for (; pos < off[k]; ++pos) changed += after[pos] != 0xA5;
            pos = off[k] + sizes[k];
            if (k < 2)  // inputs: the kernel never writes them
                for (size_t q = off[k]; q < off[k] + sizes[k]; ++q)
                    changed += after[q] != (k == 0 ? reinterpret_cast<const uint8_t*>(recs.data())[q - off[k]]
                                                   : reinterpret_cast<const uint8_t*>(progs.data())[q - off[k]]);
        }
        for (; pos < total; ++pos) changed += after[pos] != 0xA5;
        *guard_bytes_changed = changed;
        return done(NX_OK);
    });
}

int nx_kawpow_hash_many(const nx_epoch* cc, const nx_hash_record* records, size_t n, uint8_t* out_host)
{
    return guarded([&] { return nx_kawpow_hash_many_body(cc, records, n, out_host); });
}

int nx_check_block_headers(const nx_epoch* c, const uint8_t* raw, size_t n, const uint8_t pow_limit_be[32], int mix_only,
                           uint8_t* status)
{
    return guarded([&] { return nx_check_block_headers_body(c, raw, n, pow_limit_be, mix_only, status); });
}

int nx_kawpow_verify_batch(const nx_epoch* c, const nx_verify_record* records, size_t n, uint8_t* ok)
{
    return guarded([&] { return nx_kawpow_verify_batch_body(c, records, n, ok); });
}

}  // extern "C"
