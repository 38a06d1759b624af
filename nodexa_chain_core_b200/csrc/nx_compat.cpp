// The reference's own entry points — ethash_* (C) and ethash:: / progpow:: (C++) — implemented on top of
// the nx_* C ABI, so that src/hash.cpp:258-291, src/rpc/mining.cpp:811 and src/test/kawpow_tests.cpp link
// against libnxkawpow.so without source changes.  Declarations: include/nx/ethash_compat.{h,hpp}.
//
// Error behaviour follows the reference: no exceptions, no errno; create returns nullptr on failure
// (ethash.cpp:147-148), verify returns false, search returns the all-zero result, find_epoch_number -1.
// A failed device call inside hash() FAILS CLOSED: the reference's signature has no error channel, so the result is
// {final_hash = ff..ff, mix_hash = ff..ff} - a final hash no boundary accepts and a mix no honest header carries -
// and the reason stays in nx_last_error().  (An all-zero result would pass CheckProofOfWork for every nBits.)
// With NX_COMPAT_ON_ERROR=abort the process stops instead, which is what the reference's CPU code amounts to when
// it cannot continue.
#include "../../include/nx/ethash_compat.hpp"
#include "nx_internal.hpp"

#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>

// The object behind both context types: the reference-visible prefix, then our handle.
struct ethash_epoch_context_full : ethash_epoch_context
{
    nx_epoch* nx;
    ethash_epoch_context_full(nx_epoch* h) noexcept
      : ethash_epoch_context{nx_epoch_number(h), nx_epoch_light_items(h),
                             reinterpret_cast<const ethash_hash512*>(nx_epoch_light_cache(h)), nx_epoch_l1_cache(h),
                             nx_epoch_dataset_items(h)},
        nx(h)
    {}
};

namespace
{
inline const nx_epoch* handle(const ethash_epoch_context& c) noexcept
{
    return static_cast<const ethash_epoch_context_full&>(c).nx;  // every context this library hands out is a _full
}

ethash_epoch_context_full* create(int epoch) noexcept
{
    const char* dev = std::getenv("NX_DEVICE");
    nx_epoch* h = nullptr;
    if (nx_epoch_create(epoch, dev ? std::atoi(dev) : 0, &h) != NX_OK) return nullptr;
    auto* c = new (std::nothrow) ethash_epoch_context_full(h);
    if (!c) nx_epoch_destroy(h);
    return c;
}

// What hash() hands back when the device call did not succeed (see the header comment).
ethash::result failed_result() noexcept
{
    if (const char* e = std::getenv("NX_COMPAT_ON_ERROR"))
        if (std::strcmp(e, "abort") == 0) {
            std::fprintf(stderr, "libnxkawpow: hash failed: %s\n", nx_last_error());
            std::abort();
        }
    ethash::result out;
    std::memset(out.final_hash.bytes, 0xff, 32);
    std::memset(out.mix_hash.bytes, 0xff, 32);
    return out;
}

// test hook: NX_COMPAT_INJECT_FAIL=1 makes every compat hash() behave as if its device call had failed
bool injected_failure() noexcept
{
    const char* e = std::getenv("NX_COMPAT_INJECT_FAIL");
    if (!e || !*e || *e == '0') return false;
    nx::set_error("injected failure (NX_COMPAT_INJECT_FAIL)");
    return true;
}

ethash::result to_result(const nx_result& r) noexcept
{
    ethash::result out;
    std::memcpy(out.final_hash.bytes, r.final_hash, 32);
    std::memcpy(out.mix_hash.bytes, r.mix_hash, 32);
    return out;
}

// managed.cpp:23-81 re-done with std::mutex: one shared context per kind, a thread-local reference to it.
std::mutex g_shared_mu;
std::shared_ptr<ethash_epoch_context_full> g_shared;
thread_local std::shared_ptr<ethash_epoch_context_full> t_local;

const ethash_epoch_context_full* global_context(int epoch) noexcept
{
    if (t_local && t_local->epoch_number == epoch) return t_local.get();
    t_local.reset();
    std::lock_guard<std::mutex> lk(g_shared_mu);
    if (!g_shared || g_shared->epoch_number != epoch) {
        g_shared.reset();  // release the old DAG before building the next one
        ethash_epoch_context_full* c = create(epoch);
        if (!c) return nullptr;
        g_shared.reset(c, [](ethash_epoch_context_full* p) { ethash_destroy_epoch_context_full(p); });
    }
    t_local = g_shared;
    return t_local.get();
}
}  // namespace

extern "C" {

int ethash_calculate_light_cache_num_items(int e) noexcept { return nx_light_cache_num_items(e); }
int ethash_calculate_full_dataset_num_items(int e) noexcept { return nx_full_dataset_num_items(e); }
ethash_hash256 ethash_calculate_epoch_seed(int e) noexcept
{
    ethash_hash256 s;
    nx_epoch_seed(e, s.bytes);
    return s;
}

ethash_epoch_context* ethash_create_epoch_context(int e) noexcept { return create(e); }
ethash_epoch_context_full* ethash_create_epoch_context_full(int e) noexcept { return create(e); }
void ethash_destroy_epoch_context_full(ethash_epoch_context_full* c) noexcept
{
    if (!c) return;
    nx_epoch_destroy(c->nx);
    delete c;
}
void ethash_destroy_epoch_context(ethash_epoch_context* c) noexcept
{
    ethash_destroy_epoch_context_full(static_cast<ethash_epoch_context_full*>(c));
}

const ethash_epoch_context* ethash_get_global_epoch_context(int e) noexcept { return global_context(e); }
const ethash_epoch_context_full* ethash_get_global_epoch_context_full(int e) noexcept { return global_context(e); }

ethash_result ethash_hash(const ethash_epoch_context* c, const ethash_hash256* header, uint64_t nonce) noexcept
{
    nx_result r;
    if (injected_failure() || nx_ethash_hash(handle(*c), header->bytes, nonce, &r) != NX_OK) return failed_result();
    return to_result(r);
}

bool ethash_verify_final_hash(const ethash_hash256* header, const ethash_hash256* mix, uint64_t nonce,
                              const ethash_hash256* boundary) noexcept
{
    return nx_ethash_verify_final_hash(header->bytes, mix->bytes, nonce, boundary->bytes) == 1;
}

bool ethash_verify(const ethash_epoch_context* c, const ethash_hash256* header, const ethash_hash256* mix, uint64_t nonce,
                   const ethash_hash256* boundary) noexcept
{
    if (!ethash_verify_final_hash(header, mix, nonce, boundary)) return false;
    const ethash_result r = ethash_hash(c, header, nonce);
    return std::memcmp(r.mix_hash.bytes, mix->bytes, 32) == 0;
}

void ethash_keccakf1600(uint64_t st[25]) noexcept { nx_keccakf1600(st); }
void ethash_keccakf800(uint32_t st[25]) noexcept { nx_keccakf800(st); }
ethash_hash256 ethash_keccak256(const uint8_t* d, size_t n) noexcept
{
    ethash_hash256 h;
    nx_keccak256(d, n, h.bytes);
    return h;
}
ethash_hash256 ethash_keccak256_32(const uint8_t d[32]) noexcept { return ethash_keccak256(d, 32); }
ethash_hash512 ethash_keccak512(const uint8_t* d, size_t n) noexcept
{
    ethash_hash512 h;
    nx_keccak512(d, n, h.bytes);
    return h;
}
ethash_hash512 ethash_keccak512_64(const uint8_t d[64]) noexcept { return ethash_keccak512(d, 64); }
int ethash_find_largest_prime(int ub) noexcept { return nx_find_largest_prime(ub); }

}  // extern "C"

namespace ethash
{
result hash(const epoch_context_full& c, const hash256& header, uint64_t nonce) noexcept
{
    return ethash_hash(&c, &header, nonce);
}

// Classic Ethash search (ethash.cpp:326-350): ascending scan in batches, first hit wins.
static search_result classic_search(const epoch_context& c, const hash256& header, const hash256& boundary,
                                    uint64_t start, size_t iterations) noexcept
{
    constexpr size_t BATCH = 1 << 16;
    std::unique_ptr<uint8_t[]> buf(new (std::nothrow) uint8_t[BATCH * 64]);
    if (!buf) return {};
    for (size_t done = 0; done < iterations; done += BATCH) {
        const size_t n = iterations - done < BATCH ? iterations - done : BATCH;
        if (nx_ethash_hash_batch(handle(c), header.bytes, start + done, n, buf.get()) != NX_OK) return {};
        for (size_t i = 0; i < n; ++i)
            if (std::memcmp(buf.get() + 64 * i, boundary.bytes, 32) <= 0) {
                result r;
                std::memcpy(r.final_hash.bytes, buf.get() + 64 * i, 32);
                std::memcpy(r.mix_hash.bytes, buf.get() + 64 * i + 32, 32);
                return {r, start + done + i};
            }
    }
    return {};
}
search_result search_light(const epoch_context& c, const hash256& h, const hash256& b, uint64_t s, size_t n) noexcept
{
    return classic_search(c, h, b, s, n);
}
search_result search(const epoch_context_full& c, const hash256& h, const hash256& b, uint64_t s, size_t n) noexcept
{
    return classic_search(c, h, b, s, n);
}
int find_epoch_number(const hash256& seed) noexcept { return nx_find_epoch_number(seed.bytes); }
}  // namespace ethash

namespace progpow
{
result hash(const epoch_context& c, int block_number, const hash256& header, uint64_t nonce) noexcept
{
    nx_result r;
    if (injected_failure() || nx_kawpow_hash(handle(c), block_number, header.bytes, nonce, &r) != NX_OK) return failed_result();
    return to_result(r);
}
result hash(const epoch_context_full& c, int block_number, const hash256& header, uint64_t nonce) noexcept
{
    return hash(static_cast<const epoch_context&>(c), block_number, header, nonce);
}
bool verify(const epoch_context& c, int block_number, const hash256& header, const hash256& mix, uint64_t nonce,
            const hash256& boundary) noexcept
{
    return nx_kawpow_verify(handle(c), block_number, header.bytes, mix.bytes, nonce, boundary.bytes) == 1;
}
hash256 hash_no_verify(const int& block_number, const hash256& header, const hash256& mix, const uint64_t& nonce) noexcept
{
    hash256 out;
    nx_kawpow_hash_no_verify(block_number, header.bytes, mix.bytes, nonce, out.bytes);
    return out;
}
static search_result do_search(const epoch_context& c, int block_number, const hash256& header, const hash256& boundary,
                               uint64_t start, size_t iterations) noexcept
{
    nx_search_result sr;
    if (nx_kawpow_search(handle(c), block_number, header.bytes, boundary.bytes, start, iterations, &sr) != NX_OK || !sr.found)
        return {};
    result r;
    std::memcpy(r.final_hash.bytes, sr.final_hash, 32);
    std::memcpy(r.mix_hash.bytes, sr.mix_hash, 32);
    return {r, sr.nonce};
}
search_result search_light(const epoch_context& c, int bn, const hash256& h, const hash256& b, uint64_t s, size_t n) noexcept
{
    return do_search(c, bn, h, b, s, n);
}
search_result search(const epoch_context_full& c, int bn, const hash256& h, const hash256& b, uint64_t s, size_t n) noexcept
{
    return do_search(c, bn, h, b, s, n);
}
}  // namespace progpow
