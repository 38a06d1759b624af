// Run-time side of the KawPow kernels: NVRTC compilation of the per-period program for sm_100a,
// an in-process cache keyed by (device, prog_seed, DAG item count), and the launches.
//
// The period changes every 3 blocks (reference include/ethash/progpow.hpp:21), so compilation
// (a few hundred ms) is kept off the hashing path by nx_kawpow_prepare() and by the cache; the
// cubin can also be persisted in NX_PROGRAM_CACHE_DIR.
#include "nx_internal.hpp"

#include <dlfcn.h>
#include <unistd.h>
#include <nvrtc.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <future>
#include <map>
#include <memory>
#include <sstream>
#include <tuple>

namespace nx
{
namespace
{
// ---- the handful of driver-API entry points we need, resolved lazily from libcuda.so.1 so that
//      the library itself loads (and its symbols can be inspected) on a machine without a driver ----
typedef int CUresult_;
typedef struct CUmod_st* CUmodule_;
typedef struct CUfunc_st* CUfunction_;
struct Driver
{
    void* lib = nullptr;
    CUresult_ (*ModuleLoadData)(CUmodule_*, const void*) = nullptr;
    CUresult_ (*ModuleUnload)(CUmodule_) = nullptr;
    CUresult_ (*ModuleGetFunction)(CUfunction_*, CUmodule_, const char*) = nullptr;
    CUresult_ (*LaunchKernel)(CUfunction_, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned,
                              void* /*CUstream*/, void**, void**) = nullptr;
    CUresult_ (*GetErrorString)(CUresult_, const char**) = nullptr;
    CUresult_ (*FuncSetAttribute)(CUfunction_, int, int) = nullptr;
    CUresult_ (*FuncGetAttribute)(int*, int, CUfunction_) = nullptr;
    bool ok = false;
};

Driver& driver()
{
    static Driver d;
    static std::once_flag once;
    std::call_once(once, [] {
        d.lib = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
        if (!d.lib) return;
#define NX_SYM(field, name) *(void**)(&d.field) = dlsym(d.lib, name)
        NX_SYM(ModuleLoadData, "cuModuleLoadData");
        NX_SYM(ModuleUnload, "cuModuleUnload");
        NX_SYM(ModuleGetFunction, "cuModuleGetFunction");
        NX_SYM(LaunchKernel, "cuLaunchKernel");
        NX_SYM(GetErrorString, "cuGetErrorString");
        NX_SYM(FuncSetAttribute, "cuFuncSetAttribute");
        NX_SYM(FuncGetAttribute, "cuFuncGetAttribute");
#undef NX_SYM
        d.ok = d.ModuleLoadData && d.ModuleUnload && d.ModuleGetFunction && d.LaunchKernel && d.GetErrorString;
    });
    return d;
}

const char* cu_err(CUresult_ r)
{
    const char* s = nullptr;
    if (driver().GetErrorString) driver().GetErrorString(r, &s);
    return s ? s : "unknown driver error";
}

// Kernel shape.  Defaults are the measured best on B200 (profiles/r1_kawpow_tune_sweep3_sustained.txt: 8 L1
// replicas in one 1024-thread CTA per SM, 192.2 MH/s sustained vs 186.3 for 1 replica x 4 CTAs of 256);
// NX_KAWPOW_TUNE overrides them, e.g. NX_KAWPOW_TUNE="copies=1,threads=256,minblocks=4".  Read once per process.
struct Tune
{
    int copies = 8, threads = 1024, minblocks = 1, rotimad = 0, l1hi = 0, unroll = 1, daghint = 0;
    int persist = 1;             // 1: grid = SMs x minblocks CTAs whose warps stride over the job; 0: one unit per warp
    int nodag = 0, nomath = 0;   // measurement-only shapes (wrong hashes): see kawpow_kernel.cuh
    size_t smem_bytes() const { return (size_t)copies * 16384 + (size_t)14 * 4 * threads; }
    std::string tag() const
    {
        std::ostringstream s;
        s << "c" << copies << "t" << threads << "b" << minblocks << "r" << rotimad << "h" << l1hi << "u" << unroll << "d" << daghint;
        if (nodag) s << "ND";
        if (nomath) s << "NM";
        return s.str();
    }
};

const Tune& tune()
{
    static Tune t;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* e = std::getenv("NX_KAWPOW_TUNE");
        if (!e) return;
        std::string str(e);
        size_t pos = 0;
        while (pos < str.size()) {
            size_t end = str.find(',', pos);
            if (end == std::string::npos) end = str.size();
            const std::string kv = str.substr(pos, end - pos);
            const size_t eq = kv.find('=');
            if (eq != std::string::npos) {
                const std::string k = kv.substr(0, eq);
                const int v = std::atoi(kv.c_str() + eq + 1);
                if (k == "copies" && (v == 1 || v == 2 || v == 4 || v == 8)) t.copies = v;
                if (k == "threads" && v >= 32 && v <= 1024 && v % 32 == 0) t.threads = v;
                if (k == "minblocks" && v >= 1 && v <= 8) t.minblocks = v;
                if (k == "rotimad" && v >= 0 && v <= 16) t.rotimad = v;
                if (k == "l1hi" && v >= 0 && v <= 2) t.l1hi = v;
                if (k == "unroll" && (v == 1 || v == 2 || v == 4 || v == 8)) t.unroll = v;
                if (k == "persist" && (v == 0 || v == 1)) t.persist = v;
                if (k == "nodag" && (v == 0 || v == 1)) t.nodag = v;
                if (k == "nomath" && (v == 0 || v == 1)) t.nomath = v;
                if (k == "daghint" && v >= 0 && v <= 3) t.daghint = v;
            }
            pos = end + 1;
        }
    });
    return t;
}

std::string cache_path(uint64_t prog_seed, uint32_t dag_items, const char* arch)
{
    const char* dir = std::getenv("NX_PROGRAM_CACHE_DIR");
    if (!dir || !*dir) return "";
    std::ostringstream s;
    s << dir << "/kawpow_" << arch << "_" << prog_seed << "_" << dag_items << "_" << tune().tag() << "_v4.cubin";
    return s.str();
}
}  // namespace

struct KawpowProgram
{
    int device = 0;
    uint64_t prog_seed = 0;
    uint32_t dag_items = 0;
    CUmodule_ module = nullptr;
    CUfunction_ fn = nullptr;  // nx_kawpow_kernel: batch hash and search are two modes of one kernel
    double compile_ms = 0;
    int regs = 0;
    int threads = 256;
    int sms = 1;  // multiprocessors of the device: the persistent grid is sms x minblocks CTAs
    size_t smem_bytes = 0;
    ~KawpowProgram();
};

KawpowProgram::~KawpowProgram()
{
    if (module && driver().ModuleUnload) driver().ModuleUnload(module);
}

namespace
{
std::mutex g_cache_mu;
std::map<std::tuple<int, uint64_t, uint32_t>, std::shared_ptr<KawpowProgram>> g_cache;
std::mutex g_compile_mu;
std::map<std::pair<uint64_t, uint32_t>, std::shared_ptr<const std::vector<char>>> g_cubins;  // device independent

int compile_cubin(uint64_t prog_seed, uint32_t dag_items, std::vector<char>& cubin, std::string& log)
{
    const std::string src = kawpow_source(prog_seed, dag_items);
    nvrtcProgram prog = nullptr;
    if (nvrtcCreateProgram(&prog, src.c_str(), "nx_kawpow_period.cu", 0, nullptr, nullptr) != NVRTC_SUCCESS) {
        set_error("nvrtcCreateProgram failed");
        return NX_ERR_NVRTC;
    }
    const char* arch_env = std::getenv("NX_KAWPOW_ARCH");
    const std::string arch = std::string("--gpu-architecture=") + (arch_env && *arch_env ? arch_env : "sm_100a");
    const Tune& tn = tune();
    const std::string d_threads = "-DNX_THREADS=" + std::to_string(tn.threads);
    const std::string d_blocks = "-DNX_MIN_BLOCKS=" + std::to_string(tn.minblocks);
    const std::string d_copies = "-DNX_L1_COPIES=" + std::to_string(tn.copies);
    const std::string d_rot = "-DNX_ROT_IMAD=" + std::to_string(tn.rotimad);
    const std::string d_hi = "-DNX_L1_HI=" + std::to_string(tn.l1hi);
    const std::string d_unroll = "-DNX_UNROLL=" + std::to_string(tn.unroll);
    const std::string d_hint = "-DNX_DAG_HINT=" + std::to_string(tn.daghint);
    std::vector<const char*> opts = {arch.c_str(), "--std=c++17", "-lineinfo", "--extra-device-vectorization",
                                     d_threads.c_str(), d_blocks.c_str(), d_copies.c_str(), d_rot.c_str(),
                                     d_hi.c_str(), d_unroll.c_str(), d_hint.c_str()};
    if (tn.nodag) opts.push_back("-DNX_NODAG=1");
    if (tn.nomath) opts.push_back("-DNX_NOMATH=1");
    std::vector<std::string> extra;
    if (const char* e = std::getenv("NX_KAWPOW_NVRTC_FLAGS")) {  // space separated, for experiments
        std::istringstream is(e);
        std::string tok;
        while (is >> tok) extra.push_back(tok);
        for (auto& t : extra) opts.push_back(t.c_str());
    }
    const nvrtcResult rc = nvrtcCompileProgram(prog, (int)opts.size(), opts.data());
    size_t log_size = 0;
    nvrtcGetProgramLogSize(prog, &log_size);
    if (log_size > 1) {
        log.resize(log_size);
        nvrtcGetProgramLog(prog, &log[0]);
    }
    if (rc != NVRTC_SUCCESS) {
        set_error("NVRTC: %s: %.400s", nvrtcGetErrorString(rc), log.c_str());
        nvrtcDestroyProgram(&prog);
        return NX_ERR_NVRTC;
    }
    size_t n = 0;
    if (nvrtcGetCUBINSize(prog, &n) != NVRTC_SUCCESS || n == 0) {
        set_error("NVRTC produced no cubin (arch %s)", arch.c_str());
        nvrtcDestroyProgram(&prog);
        return NX_ERR_NVRTC;
    }
    cubin.resize(n);
    nvrtcGetCUBIN(prog, cubin.data());
    nvrtcDestroyProgram(&prog);
    return NX_OK;
}
}  // namespace

// Exposed for tests and tools: compile only (no device needed) and hand back the cubin bytes.
int kawpow_compile_only(uint64_t prog_seed, uint32_t dag_items, std::vector<char>& cubin, std::string& log)
{
    return compile_cubin(prog_seed, dag_items, cubin, log);
}

bool kawpow_program_is_cached(int device, uint64_t prog_seed, uint32_t dag_items)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    return g_cache.count(std::make_tuple(device, prog_seed, dag_items)) != 0;
}

// The cubin of (prog_seed, dag_items): from this process's map, else from NX_PROGRAM_CACHE_DIR, else compiled now.
// One NVRTC run serves every device of the process.  g_compile_mu guards the maps only: the compilation itself runs
// unlocked, so different periods compile side by side (a background prefetch never delays a foreground request for
// another period), while a second request for a period that is being compiled waits for that run instead of repeating it.
namespace
{
std::map<std::pair<uint64_t, uint32_t>, std::shared_future<int>> g_inflight;
std::atomic<bool> g_nvrtc_used{false};  // NVRTC has compiled something in this process (its lazy state exists)
}

static int obtain_cubin(uint64_t prog_seed, uint32_t dag_items, std::shared_ptr<const std::vector<char>>* out)
{
    const auto ckey = std::make_pair(prog_seed, dag_items);
    std::promise<int> mine;
    {
        std::unique_lock<std::mutex> lk(g_compile_mu);
        for (;;) {
            auto it = g_cubins.find(ckey);
            if (it != g_cubins.end()) {
                *out = it->second;
                return NX_OK;
            }
            auto fl = g_inflight.find(ckey);
            if (fl == g_inflight.end()) break;
            std::shared_future<int> other = fl->second;
            lk.unlock();
            const int rc = other.get();  // somebody else is compiling exactly this
            lk.lock();
            if (rc != NX_OK && !g_cubins.count(ckey)) {
                set_error("compilation of period %llu failed in another thread", (unsigned long long)prog_seed);
                return rc;
            }
        }
        g_inflight[ckey] = mine.get_future().share();
    }
    auto fresh = std::make_shared<std::vector<char>>();
    int rc = NX_OK;
    const char* arch_env = std::getenv("NX_KAWPOW_ARCH");
    const std::string path = cache_path(prog_seed, dag_items, arch_env && *arch_env ? arch_env : "sm_100a");
    if (!path.empty()) {
        std::ifstream f(path, std::ios::binary);
        if (f) fresh->assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
        // a file that is not a whole ELF image (foreign, truncated by a crash) is ignored and rewritten
        if (fresh->size() < 64 || std::memcmp(fresh->data(), "\x7f" "ELF", 4) != 0) fresh->clear();
    }
    if (fresh->empty()) {
        std::string log;
        rc = compile_cubin(prog_seed, dag_items, *fresh, log);
        if (rc == NX_OK) g_nvrtc_used.store(true);
        if (rc == NX_OK && !path.empty()) {
            // several processes (one per GPU) may compile the same period at the same time: write beside the
            // final name and rename, so that no reader ever sees a half-written cubin
            const std::string tmp = path + ".tmp." + std::to_string((long)getpid());
            bool ok = false;
            {
                std::ofstream f(tmp, std::ios::binary);
                f.write(fresh->data(), (std::streamsize)fresh->size());
                ok = f.good();
            }
            if (!ok || std::rename(tmp.c_str(), path.c_str()) != 0) std::remove(tmp.c_str());
        }
    }
    {
        std::lock_guard<std::mutex> lk(g_compile_mu);
        if (rc == NX_OK) {
            while (g_cubins.size() >= 8) g_cubins.erase(g_cubins.begin());  // lowest period first: heights ascend
            g_cubins[ckey] = fresh;
            *out = fresh;
        }
        g_inflight.erase(ckey);
    }
    mine.set_value(rc);
    return rc;
}

// The period changes every 3 blocks.  A caller that searches period p will want p+1 within minutes, and compiling it
// then stalls its first call by the whole NVRTC run.  With NX_PREFETCH=1 the use of a period starts the compilation of
// the next one on a background thread (one at a time; only the device-independent cubin: loading it into a device
// later takes about 0.1 ms).  Opt-in, because a thread that is inside NVRTC when the process exits races the
// library's own teardown: the wait for it is registered with atexit() at the first prefetch, which runs before the
// destructors of everything constructed earlier - so a prefetch only starts once NVRTC has compiled something in the
// foreground (its lazily built state then predates the registration).  Callers that plan ahead themselves
// (nx_kawpow_prepare, nx_miner_prepare_async) need none of this.
namespace
{
struct Prefetch
{
    std::mutex mu;
    std::future<void> fut;
    void wait()
    {
        std::lock_guard<std::mutex> lk(mu);
        if (fut.valid()) fut.wait();
    }
    ~Prefetch() { wait(); }
} g_prefetch;
}  // namespace

static void prefetch_next_period(int device, uint64_t prog_seed, uint32_t dag_items)
{
    static const bool on = [] {
        const char* e = std::getenv("NX_PREFETCH");
        return e && *e && *e != '0';
    }();
    if (!on || !g_nvrtc_used.load()) return;
    const uint64_t next = prog_seed + 1;
    if ((next * NX_PERIOD_LENGTH) / NX_EPOCH_LENGTH != (prog_seed * NX_PERIOD_LENGTH) / NX_EPOCH_LENGTH) return;  // next epoch: other DAG size
    // only where periods are used in ascending order (a chain being mined or followed), or the same one again: a caller
    // that walks heights downwards or at random would get nothing from it
    static std::atomic<uint64_t> last_seed{~0ull};
    const uint64_t prev = last_seed.exchange(prog_seed);
    if (!(prev == ~0ull || prev == prog_seed || prev + 1 == prog_seed)) return;
    if (kawpow_program_is_cached(device, next, dag_items)) return;
    {
        std::lock_guard<std::mutex> lk(g_compile_mu);
        if (g_cubins.count(std::make_pair(next, dag_items)) || g_inflight.count(std::make_pair(next, dag_items))) return;
    }
    static std::once_flag at_exit;
    std::call_once(at_exit, [] { std::atexit([] { g_prefetch.wait(); }); });  // before any teardown, see above
    std::lock_guard<std::mutex> lk(g_prefetch.mu);
    if (g_prefetch.fut.valid() && g_prefetch.fut.wait_for(std::chrono::seconds(0)) != std::future_status::ready) return;
    g_prefetch.fut = std::async(std::launch::async, [next, dag_items] {
        std::shared_ptr<const std::vector<char>> drop;
        obtain_cubin(next, dag_items, &drop);  // a failure here is the foreground call's to report when it gets there
        clear_error();
    });
}

// Three steps with separate locks, so that a cached lookup (every search call makes one) never waits behind another
// thread's compilation of the next period: (1) map lookup under g_cache_mu; (2) cubin for (prog_seed, dag_items) under
// g_compile_mu - one NVRTC run serves every device of the process; (3) per-device module load, then publish.
static int get_program(int device, uint64_t prog_seed, uint32_t dag_items, std::shared_ptr<KawpowProgram>* out)
{
    const auto key = std::make_tuple(device, prog_seed, dag_items);
    {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        auto it = g_cache.find(key);
        if (it != g_cache.end()) {
            *out = it->second;
            return NX_OK;
        }
    }
    Driver& d = driver();
    if (!d.ok) {
        set_error("libcuda.so.1 not available: a CUDA driver is required (no CPU fallback)");
        return NX_ERR_CUDA;
    }
    NX_CUDA(cudaSetDevice(device));
    NX_CUDA(cudaFree(nullptr));  // make sure the primary context exists and is current

    const auto t0 = std::chrono::steady_clock::now();
    std::shared_ptr<const std::vector<char>> cubin;
    {
        const int rc = obtain_cubin(prog_seed, dag_items, &cubin);
        if (rc != NX_OK) return rc;
    }
    std::shared_ptr<KawpowProgram> p(new KawpowProgram);
    p->device = device;
    p->prog_seed = prog_seed;
    p->dag_items = dag_items;
    CUresult_ r = d.ModuleLoadData(&p->module, cubin->data());
    if (r != 0) {
        set_error("cuModuleLoadData failed: %s", cu_err(r));
        return NX_ERR_CUDA;
    }
    if ((r = d.ModuleGetFunction(&p->fn, p->module, "nx_kawpow_kernel")) != 0) {
        set_error("cuModuleGetFunction failed: %s", cu_err(r));
        return NX_ERR_CUDA;
    }
    p->threads = tune().threads;
    p->smem_bytes = tune().smem_bytes();
    if (p->smem_bytes > 48 * 1024) {
        if (!d.FuncSetAttribute ||
            (r = d.FuncSetAttribute(p->fn, 8 /*CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES*/, (int)p->smem_bytes)) != 0) {
            set_error("cuFuncSetAttribute(max dynamic shared memory = %zu) failed: %s", p->smem_bytes, cu_err(r));
            return NX_ERR_CUDA;
        }
    }
    if (d.FuncGetAttribute) d.FuncGetAttribute(&p->regs, 4 /*CU_FUNC_ATTRIBUTE_NUM_REGS*/, p->fn);
    NX_CUDA(cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, device));
    p->compile_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (std::getenv("NX_VERBOSE"))
        fprintf(stderr, "[nx] kawpow program seed=%llu items=%u tune=%s: %.1f ms, regs=%d smem=%zu\n",
                (unsigned long long)prog_seed, dag_items, tune().tag().c_str(), p->compile_ms, p->regs, p->smem_bytes);
    std::lock_guard<std::mutex> lk(g_cache_mu);
    auto it = g_cache.find(key);
    if (it != g_cache.end()) {  // another thread loaded the same program meanwhile: keep one
        *out = it->second;
        return NX_OK;
    }
    // keep the cache bounded: a node needs the current and the next period per device.  The lowest period goes first;
    // an evicted program is unloaded when its last user lets go of the shared_ptr.
    while (g_cache.size() >= 64) {
        auto victim = g_cache.begin();
        for (auto jt = g_cache.begin(); jt != g_cache.end(); ++jt)
            if (std::get<1>(jt->first) < std::get<1>(victim->first)) victim = jt;
        g_cache.erase(victim);
    }
    g_cache[key] = p;
    *out = p;
    return NX_OK;
}

int kawpow_get_program(int device, uint64_t prog_seed, uint32_t dag_items, std::shared_ptr<KawpowProgram>* out)
{
    return guarded([&] {
        const int rc = get_program(device, prog_seed, dag_items, out);
        if (rc == NX_OK) prefetch_next_period(device, prog_seed, dag_items);
        return rc;
    });
}

// Shape of one launch.  A unit is the work one warp does between two looks at the job: 2 x per_hw nonces (one
// half-warp walks per_hw of them).  per_hw is 16 unless the call is too small to give every resident warp a full unit:
// then the nonces are spread thinner so that the call costs per_hw / 16 of a full walk instead of a whole one (a
// 65 536-nonce call - the reference's nInnerLoopCount, src/rpc/mining.cpp:119 - gets per_hw = 7 on 148 SMs).
void kawpow_launch_shape(uint64_t count, int threads, int sms, int minblocks, bool persist, uint32_t* per_hw,
                         uint32_t* n_units, uint32_t* grid)
{
    const uint64_t warps_per_cta = (uint64_t)threads / 32;
    const uint64_t resident_warps = (uint64_t)sms * minblocks * warps_per_cta;
    uint64_t k = (count + 2 * resident_warps - 1) / (2 * resident_warps);
    if (k < 1) k = 1;
    if (k > 16) k = 16;
    const uint64_t units = (count + 2 * k - 1) / (2 * k);
    uint64_t ctas = (units + warps_per_cta - 1) / warps_per_cta;
    if (persist && ctas > (uint64_t)sms * minblocks) ctas = (uint64_t)sms * minblocks;
    *per_hw = (uint32_t)k;
    *n_units = (uint32_t)units;
    *grid = (uint32_t)ctas;
}

static int launch(const KawpowProgram* p, uint32_t mode, const void* dag, const KawpowJob& job_in, void* third,
                  cudaStream_t s)
{
    if (job_in.count == 0) return NX_OK;
    if (job_in.count > (1ull << 32)) {
        set_error("kawpow launch: batch of %llu nonces is too large for one launch", (unsigned long long)job_in.count);
        return NX_ERR_INVALID;
    }
    KawpowJob job = job_in;
    job.mode = mode;
    int log_copies = 0;
    while ((1 << log_copies) < tune().copies) ++log_copies;
    job.l1_mul_lo = 1u << 20;
    job.l1_mul_hi = 1u << (14 + log_copies);
    job.pad = 0;
    uint32_t grid = 0;
    kawpow_launch_shape(job.count, p->threads, p->sms, tune().minblocks, tune().persist != 0, &job.per_hw, &job.n_units, &grid);
    void* args[3] = {(void*)&dag, (void*)&job, (void*)&third};
    const CUresult_ r = driver().LaunchKernel(p->fn, grid, 1, 1, (unsigned)p->threads, 1, 1,
                                              (unsigned)p->smem_bytes, (void*)s, args, nullptr);
    if (r != 0) {
        set_error("cuLaunchKernel failed: %s", cu_err(r));
        return NX_ERR_CUDA;
    }
    return NX_OK;
}

int kawpow_launch_hash(const KawpowProgram* p, const void* dag, const KawpowJob& job, void* out_dev, cudaStream_t s)
{
    return launch(p, 0u, dag, job, out_dev, s);
}

int kawpow_launch_search(const KawpowProgram* p, const void* dag, const KawpowJob& job, void* found_min_dev, cudaStream_t s)
{
    return launch(p, 1u, dag, job, found_min_dev, s);
}

}  // namespace nx
