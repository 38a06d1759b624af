// Single-process, multi-GPU mining front end: the replacement for the node's nonce loops
// (reference src/miner.cpp:656-713 CloreMiner, src/rpc/mining.cpp:117-173 generateBlocks) and the
// epoch/period lifecycle around them (src/hash.cpp:260-266 static context, lib/ethash/managed.cpp).
//
//  * one epoch context per device; the (serial) light cache is built ONCE on the host and shared, every device
//    then generates its own DAG copy in parallel (0.2 s on a B200 - cheaper than moving 4.5 GiB around);
//  * nonce space cut into super-batches of `batch` nonces that the devices claim in ascending order as they
//    become free; the answer is the LOWEST qualifying nonce (progpow.cpp:572-577): windows above the best hit so
//    far are dropped, windows below it finish - a host-side min reduction, no device collective.  A device whose
//    launch fails drops out and its window returns to the pool for the surviving devices;
//  * nx_miner_prepare_async() builds the next epoch's contexts and NVRTC-compiles the next period's program
//    in a background thread, so neither ever sits on the mining path.
#include "nx_internal.hpp"

#include <atomic>
#include <chrono>
#include <cstring>
#include <cstdlib>
#include <functional>
#include <future>
#include <thread>

using namespace nx;

struct nx_miner
{
    std::vector<int> devices;
    std::mutex mu;                      // guards current/next
    int epoch = -1;
    std::vector<nx_epoch*> current;     // one per device
    int next_epoch = -1;
    std::vector<nx_epoch*> next;        // prebuilt contexts for next_epoch (may be empty)
    std::mutex bg_mu;                   // guards `background` itself (prepare_async may be called beside a search)
    std::shared_future<int> background;
    uint64_t batch = 1ull << 22;
    // statistics of the last search
    std::atomic<uint64_t> hashes{0};
    double last_seconds = 0;
    int failed_devices = 0;             // devices that dropped out during the last search
};

namespace
{
std::shared_future<int> background_of(nx_miner* m)
{
    std::lock_guard<std::mutex> lk(m->bg_mu);
    return m->background;
}

void destroy_all(std::vector<nx_epoch*>& v)
{
    for (nx_epoch* c : v) nx_epoch_destroy(c);
    v.clear();
}

// Light cache once on the host.  With peer access between all devices each device then generates 1/N of the
// DAG and its kernel stores the slice into every device's buffer (fused generate + all-gather over NVLink);
// otherwise every device generates the whole DAG for itself.  Either way the devices work in parallel.
bool all_peers_reachable(const std::vector<int>& devices)
{
    if (devices.size() < 2 || devices.size() > (size_t)NX_MAX_PEERS) return false;
    if (const char* e = std::getenv("NX_MINER_NO_P2P"))
        if (*e && *e != '0') return false;
    for (int a : devices)
        for (int b : devices) {
            if (a == b) continue;
            int ok = 0;
            if (cudaDeviceCanAccessPeer(&ok, a, b) != cudaSuccess || !ok) {
                cudaGetLastError();
                return false;
            }
        }
    for (int a : devices) {
        if (cudaSetDevice(a) != cudaSuccess) return false;
        for (int b : devices) {
            if (a == b) continue;
            const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                return false;
            }
            cudaGetLastError();
        }
    }
    return true;
}

int build_contexts(const std::vector<int>& devices, int epoch, std::vector<nx_epoch*>& out)
{
    const int n_light = nx_light_cache_num_items(epoch);
    std::vector<uint8_t> light((size_t)n_light * 64);
    uint8_t seed[32];
    nx_epoch_seed(epoch, seed);
    if (!host_load_light_cache(light.data(), n_light, epoch)) {
        host_build_light_cache(light.data(), n_light, seed);
        host_store_light_cache(light.data(), n_light, epoch);
    }

    const size_t n = devices.size();
    const bool fused = all_peers_reachable(devices);
    out.assign(n, nullptr);
    std::vector<int> rcs(n, NX_OK);
    std::vector<std::string> errs(n);
    auto run_all = [&](const std::function<int(size_t)>& fn) {
        std::vector<std::thread> th;
        for (size_t i = 0; i < n; ++i)
            th.emplace_back([&, i] {
                if (rcs[i] != NX_OK) return;
                rcs[i] = fn(i);
                if (rcs[i] != NX_OK) errs[i] = nx_last_error();
            });
        for (auto& t : th) t.join();  // doubles as the barrier between the phases
    };
    if (fused) {
        // Every phase needs every rank: a slice that was not generated, or a buffer that does not exist, would leave
        // holes in (or stray stores into) the DAGs of the healthy devices.  So one failure ends the build there.
        auto any_failed = [&] {
            for (size_t i = 0; i < n; ++i)
                if (rcs[i] != NX_OK) return true;
            return false;
        };
        // test hook: NX_MINER_INJECT_CREATE_FAIL=<index> makes that entry's context creation fail (stands for an
        // allocation that does not fit): the build must stop before any fused kernel runs
        long fail_create = -1;
        if (const char* e = std::getenv("NX_MINER_INJECT_CREATE_FAIL")) fail_create = std::strtol(e, nullptr, 10);
        run_all([&](size_t i) {
            if ((long)i == fail_create) {
                set_error("injected failure (NX_MINER_INJECT_CREATE_FAIL)");
                return (int)NX_ERR_OOM;
            }
            return nx_epoch_create_deferred(epoch, devices[i], (int)i, (int)n, light.data(), &out[i]);
        });
        std::vector<void*> bases(n, nullptr);
        for (size_t i = 0; i < n; ++i) bases[i] = out[i] ? nx_epoch_dag_ptr(out[i]) : nullptr;
        if (!any_failed()) run_all([&](size_t i) { return nx_epoch_generate_fused_peers(out[i], bases.data()); });
        if (!any_failed()) run_all([&](size_t i) { return nx_epoch_finalize(out[i]); });
    } else {
        run_all([&](size_t i) { return nx_epoch_create_from_light(epoch, devices[i], light.data(), &out[i]); });
    }
    for (size_t i = 0; i < n; ++i)
        if (rcs[i] != NX_OK) {
            set_error("device %d: %s", devices[i], errs[i].c_str());
            destroy_all(out);
            return rcs[i];
        }
    return NX_OK;
}
}  // namespace

extern "C" {

static int nx_miner_create_body(const int* devices, int n_devices, nx_miner** out)
{
    if (!out) {
        set_error("null out pointer");
        return NX_ERR_INVALID;
    }
    *out = nullptr;
    const int have = nx_device_count();
    if (have < 1) {
        set_error("no CUDA device: this library has no CPU fallback");
        return NX_ERR_CUDA;
    }
    std::unique_ptr<nx_miner> m(new nx_miner);
    if (!devices || n_devices <= 0) {
        for (int d = 0; d < have; ++d) m->devices.push_back(d);  // all visible devices
    } else {
        for (int i = 0; i < n_devices; ++i) {
            if (devices[i] < 0 || devices[i] >= have) {
                set_error("device %d out of range (have %d)", devices[i], have);
                return NX_ERR_INVALID;
            }
            m->devices.push_back(devices[i]);
        }
    }
    *out = m.release();
    return NX_OK;
}

void nx_miner_destroy(nx_miner* m)
{
    if (!m) return;
    if (auto bg = background_of(m); bg.valid()) bg.wait();
    destroy_all(m->current);
    destroy_all(m->next);
    delete m;
}

int nx_miner_device_count(const nx_miner* m) { return m ? (int)m->devices.size() : 0; }

int nx_miner_set_batch(nx_miner* m, uint64_t nonces_per_batch)
{
    if (!m || nonces_per_batch == 0 || nonces_per_batch > (1ull << 30)) {
        set_error("batch must be in [1, 2^30]");
        return NX_ERR_INVALID;
    }
    m->batch = nonces_per_batch;
    return NX_OK;
}

// Make `epoch` current on every device: adopts the contexts prepared in the background when they match.
static int nx_miner_set_epoch_body(nx_miner* m, int epoch)
{
    if (!m || epoch < 0) {
        set_error("bad argument");
        return NX_ERR_INVALID;
    }
    {
        // the common case - same epoch as last time - must not wait for a background task: that task may be the
        // compilation of the NEXT period, which is meant to overlap the mining of this one
        std::lock_guard<std::mutex> lk(m->mu);
        if (m->epoch == epoch && !m->current.empty()) return NX_OK;
    }
    // an epoch switch replaces (and frees) the current contexts: let a background task that may be using them, or
    // that may be building this very epoch, finish first
    if (auto bg = background_of(m); bg.valid()) bg.wait();
    std::lock_guard<std::mutex> lk(m->mu);
    if (m->epoch == epoch && !m->current.empty()) return NX_OK;
    if (m->next_epoch == epoch && !m->next.empty()) {
        destroy_all(m->current);
        m->current.swap(m->next);
        m->epoch = epoch;
        m->next_epoch = -1;
        return NX_OK;
    }
    destroy_all(m->current);  // free the old DAGs before allocating new ones
    m->epoch = -1;
    std::vector<nx_epoch*> fresh;
    const int rc = build_contexts(m->devices, epoch, fresh);
    if (rc != NX_OK) return rc;
    m->current.swap(fresh);
    m->epoch = epoch;
    return NX_OK;
}

// Background preparation for the blocks after `block_number`: the next period's program on every device,
// and - when the next epoch starts within `lookahead_blocks` - that epoch's contexts.  Returns at once.
static int nx_miner_prepare_async_body(nx_miner* m, int block_number, int lookahead_blocks)
{
    if (!m || block_number < 0) {
        set_error("bad argument");
        return NX_ERR_INVALID;
    }
    std::lock_guard<std::mutex> bg_lk(m->bg_mu);
    if (m->background.valid() &&
        m->background.wait_for(std::chrono::seconds(0)) != std::future_status::ready)
        return NX_OK;  // still busy with the previous request
    m->background = std::async(std::launch::async, [m, block_number, lookahead_blocks]() -> int {
        int rc = NX_OK;
        std::vector<nx_epoch*> cur;
        int cur_epoch;
        {
            std::lock_guard<std::mutex> lk(m->mu);
            cur = m->current;
            cur_epoch = m->epoch;
        }
        const int next_period_block = (block_number / NX_PERIOD_LENGTH + 1) * NX_PERIOD_LENGTH;
        if (next_period_block / NX_EPOCH_LENGTH == cur_epoch)
            for (nx_epoch* c : cur) {
                const int r = nx_kawpow_prepare(c, next_period_block);
                if (r != NX_OK) rc = r;
            }
        const int ahead_epoch = (block_number + (lookahead_blocks > 0 ? lookahead_blocks : 0)) / NX_EPOCH_LENGTH;
        bool need;
        {
            std::lock_guard<std::mutex> lk(m->mu);
            need = ahead_epoch != cur_epoch && m->next_epoch != ahead_epoch;
        }
        if (need) {
            std::vector<nx_epoch*> fresh;
            const int r = build_contexts(m->devices, ahead_epoch, fresh);
            if (r != NX_OK) return r;
            for (nx_epoch* c : fresh) nx_kawpow_prepare(c, ahead_epoch * NX_EPOCH_LENGTH);
            std::lock_guard<std::mutex> lk(m->mu);
            destroy_all(m->next);
            m->next.swap(fresh);
            m->next_epoch = ahead_epoch;
        }
        return rc;
    });
    return NX_OK;
}

int nx_miner_wait_prepared(nx_miner* m)
{
    if (!m) return NX_ERR_INVALID;
    const std::shared_future<int> bg = background_of(m);
    if (!bg.valid()) return NX_OK;
    return guarded([&] { return bg.get(); });  // get() rethrows whatever the background task threw
}

int nx_miner_epoch(const nx_miner* m) { return m ? m->epoch : -1; }
int nx_miner_prepared_epoch(const nx_miner* m) { return m ? m->next_epoch : -1; }

const nx_epoch* nx_miner_context(const nx_miner* m, int index)
{
    if (!m || index < 0 || index >= (int)m->current.size()) return nullptr;
    return m->current[(size_t)index];
}

// progpow::search over all devices.  `cancel` (may be NULL) is polled between batches: a non-zero value makes
// the call return early with found = 0 unless a hit was already confirmed as the lowest.
static int nx_miner_search_body(nx_miner* m, int block_number, const uint8_t header_hash[32], const uint8_t boundary[32],
                    uint64_t start_nonce, uint64_t iterations, const volatile int* cancel, nx_search_result* out)
{
    if (!m || !header_hash || !boundary || !out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    std::memset(out, 0, sizeof *out);
    if (block_number < 0) {
        set_error("negative block number");
        return NX_ERR_INVALID;
    }
    int rc = nx_miner_set_epoch(m, block_number / NX_EPOCH_LENGTH);
    if (rc != NX_OK) return rc;
    if (start_nonce + iterations <= start_nonce) return NX_OK;  // uint64_t wrap: the reference's loop scans nothing
    std::vector<nx_epoch*> ctxs;
    {
        std::lock_guard<std::mutex> lk(m->mu);
        ctxs = m->current;
    }
    const size_t n = ctxs.size();
    for (nx_epoch* c : ctxs)  // compile once per device, before the clocks of the workers start
        if ((rc = nx_kawpow_prepare(c, block_number)) != NX_OK) return rc;

    const uint64_t B = m->batch;
    const uint64_t total_batches = iterations / B + (iterations % B != 0);  // no wrap for iterations near 2^64
    std::atomic<uint64_t> best_offset{~0ull};  // lowest hit so far, as an offset from start_nonce
    std::atomic<uint64_t> scanned{0};
    std::atomic<uint64_t> next_batch{0};       // windows are claimed in ascending order by whichever device is free
    std::mutex retry_mu;
    std::vector<uint64_t> retry;               // windows given back by a device that failed
    std::vector<int> rcs(n, NX_OK);
    std::vector<std::string> errs(n);
    std::vector<nx_search_result> hits(n);
    std::atomic<int> alive{(int)n};
    const auto t0 = std::chrono::steady_clock::now();

    // test hook: NX_MINER_INJECT_FAIL="<device index>:<windows it may complete first>"
    int fail_dev = -1;
    long fail_after = 0;
    if (const char* e = std::getenv("NX_MINER_INJECT_FAIL")) {
        char* end = nullptr;
        fail_dev = (int)std::strtol(e, &end, 10);
        if (end && *end == ':') fail_after = std::strtol(end + 1, nullptr, 10);
    }

    auto worker = [&](size_t g) {
        std::memset(&hits[g], 0, sizeof hits[g]);
        long done_here = 0;
        for (;;) {
            uint64_t k;
            bool reissued = false;
            {
                std::lock_guard<std::mutex> lk(retry_mu);
                if (!retry.empty()) {
                    k = retry.back();
                    retry.pop_back();
                    reissued = true;
                } else {
                    k = next_batch.fetch_add(1, std::memory_order_relaxed);
                }
            }
            if (k >= total_batches) break;
            const uint64_t off = k * B;
            // Windows are handed out in ascending order, so once a hit exists every window starting above it can
            // be dropped; windows below it that are still running on other devices finish and may lower it.
            // A fresh window above the hit ends this worker (every later one lies higher still: with an "unlimited"
            // iteration count there are 2^40 and more of them, which must not be walked one by one); a re-issued
            // window above the hit is just dropped, since lower ones may still wait in the retry list.
            if (off >= best_offset.load(std::memory_order_relaxed)) {
                if (reissued) continue;
                break;
            }
            if (cancel && *cancel) break;
            const uint64_t count = iterations - off < B ? iterations - off : B;
            nx_search_result r;
            int e = NX_OK;
            if ((int)g == fail_dev && done_here >= fail_after) {
                set_error("injected failure (NX_MINER_INJECT_FAIL)");
                e = NX_ERR_CUDA;
            }
            if (e == NX_OK) e = nx_kawpow_search_enqueue(ctxs[g], block_number, header_hash, boundary, start_nonce + off, count);
            if (e == NX_OK) e = nx_kawpow_search_sync(ctxs[g], &r);
            if (e != NX_OK) {
                // this device drops out; its window goes back to the pool for the surviving devices
                rcs[g] = e;
                errs[g] = nx_last_error();
                std::lock_guard<std::mutex> lk(retry_mu);
                retry.push_back(k);
                alive.fetch_sub(1);
                return;
            }
            ++done_here;
            scanned.fetch_add(count, std::memory_order_relaxed);
            if (r.found) {
                const uint64_t hit_off = r.nonce - start_nonce;
                uint64_t cur = best_offset.load();
                while (hit_off < cur && !best_offset.compare_exchange_weak(cur, hit_off)) {
                }
                if (!hits[g].found || r.nonce - start_nonce < hits[g].nonce - start_nonce) hits[g] = r;
            }
        }
    };
    std::vector<std::thread> th;
    for (size_t g = 1; g < n; ++g) th.emplace_back(worker, g);
    worker(0);
    for (auto& t : th) t.join();
    // A device may have failed after the others had already run out of windows: finish its window(s) on a survivor.
    for (size_t g = 0; g < n && !retry.empty(); ++g)
        if (rcs[g] == NX_OK) {
            const nx_search_result keep = hits[g];
            worker(g);
            if (keep.found && (!hits[g].found || keep.nonce - start_nonce < hits[g].nonce - start_nonce)) hits[g] = keep;
        }
    m->last_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    m->hashes = scanned.load();
    m->failed_devices = (int)n - alive.load();
    if (!retry.empty() || alive.load() == 0) {  // nobody left to take the work over
        for (size_t g = 0; g < n; ++g)
            if (rcs[g] != NX_OK) {
                set_error("device %d: %s", m->devices[g], errs[g].c_str());
                return rcs[g];
            }
    }
    const uint64_t best = best_offset.load();
    if (best == ~0ull) return NX_OK;
    for (size_t g = 0; g < n; ++g)
        if (hits[g].found && hits[g].nonce - start_nonce == best) *out = hits[g];
    return NX_OK;
}

int nx_miner_last_failed_devices(const nx_miner* m) { return m ? m->failed_devices : 0; }

int nx_miner_last_stats(const nx_miner* m, uint64_t* hashes, double* seconds)
{
    if (!m) return NX_ERR_INVALID;
    if (hashes) *hashes = m->hashes.load();
    if (seconds) *seconds = m->last_seconds;
    return NX_OK;
}

// The node's inner loop in one call (miner.cpp:656-690 / rpc/mining.cpp:141-158): header fields in, nBits
// decoded to the boundary, nNonce64 scanned from start_nonce for at most max_tries nonces.
// Returns NX_OK with out->found = 1 and the block's nNonce64 / mix_hash / final hash when a nonce qualifies.
int nx_miner_mine_header(nx_miner* m, int32_t version, const uint8_t prev_block[32], const uint8_t merkle_root[32],
                         uint32_t time, uint32_t bits, uint32_t height, uint64_t start_nonce, uint64_t max_tries,
                         const volatile int* cancel, nx_search_result* out)
{
    if (!out) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    std::memset(out, 0, sizeof *out);
    uint8_t header_hash[32], boundary[32];
    if (!nx_compact_to_boundary(bits, boundary)) {
        set_error("nBits 0x%08x does not decode to a usable target (negative, zero or overflow)", bits);
        return NX_ERR_INVALID;
    }
    nx_kawpow_header_hash(version, prev_block, merkle_root, time, bits, height, header_hash);
    return nx_miner_search(m, (int)height, header_hash, boundary, start_nonce, max_tries, cancel, out);
}

int nx_miner_create(const int* devices, int n_devices, nx_miner** out)
{
    return guarded([&] { return nx_miner_create_body(devices, n_devices, out); });
}

int nx_miner_set_epoch(nx_miner* m, int epoch)
{
    return guarded([&] { return nx_miner_set_epoch_body(m, epoch); });
}

int nx_miner_prepare_async(nx_miner* m, int block_number, int lookahead_blocks)
{
    return guarded([&] { return nx_miner_prepare_async_body(m, block_number, lookahead_blocks); });
}

int nx_miner_search(nx_miner* m, int block_number, const uint8_t header_hash[32], const uint8_t boundary[32],
                    uint64_t start_nonce, uint64_t iterations, const volatile int* cancel, nx_search_result* out)
{
    return guarded([&] { return nx_miner_search_body(m, block_number, header_hash, boundary, start_nonce, iterations, cancel, out); });
}

}  // extern "C"
