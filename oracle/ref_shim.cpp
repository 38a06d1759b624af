// TEST INFRASTRUCTURE ONLY — never linked into the product library.
//
// Thin extern "C" shim over the UNMODIFIED reference sources under
// /root/reference/src/crypto/ethash (compiled in place by oracle/Makefile,
// output only into oracle/_ref/).  It lets Python (ctypes) drive the
// reference's own C++ API:
//   progpow::hash / verify / hash_no_verify / search_light   (lib/ethash/progpow.cpp:298-579)
//   ethash::calculate_dataset_item_512 / _2048               (lib/ethash/ethash.cpp:209-251)
//   ethash_create_epoch_context / ethash_hash / ...          (lib/ethash/ethash.cpp:355-442)
// No reference source text is copied; this file only calls into it.
#include <crypto/ethash/include/ethash/progpow.hpp>
#include <crypto/ethash/include/ethash/keccak.hpp>
#include <crypto/ethash/lib/ethash/ethash-internal.hpp>

#include <atomic>
#include <cstring>
#include <thread>
#include <vector>

using namespace ethash;

extern "C" {

void* nxref_ctx_create(int epoch) { return ethash_create_epoch_context(epoch); }
void nxref_ctx_destroy(void* c) { ethash_destroy_epoch_context(static_cast<epoch_context*>(c)); }

int nxref_ctx_light_items(const void* c) { return static_cast<const epoch_context*>(c)->light_cache_num_items; }
int nxref_ctx_dataset_items(const void* c) { return static_cast<const epoch_context*>(c)->full_dataset_num_items; }
const void* nxref_ctx_light_cache(const void* c) { return static_cast<const epoch_context*>(c)->light_cache; }
const uint32_t* nxref_ctx_l1(const void* c) { return static_cast<const epoch_context*>(c)->l1_cache; }

void nxref_kawpow_hash(const void* c, int height, const uint8_t header[32], uint64_t nonce,
                       uint8_t final_out[32], uint8_t mix_out[32])
{
    hash256 h;
    std::memcpy(h.bytes, header, 32);
    const result r = progpow::hash(*static_cast<const epoch_context*>(c), height, h, nonce);
    std::memcpy(final_out, r.final_hash.bytes, 32);
    std::memcpy(mix_out, r.mix_hash.bytes, 32);
}

// count nonces starting at start, light path, fanned over `threads` host threads.
// out = count x {final[32], mix[32]}.
void nxref_kawpow_hash_range(const void* c, int height, const uint8_t header[32], uint64_t start,
                             uint64_t count, uint8_t* out, int threads)
{
    if (threads < 1) threads = 1;
    std::atomic<uint64_t> next{0};
    auto work = [&] {
        for (;;) {
            const uint64_t i = next.fetch_add(1);
            if (i >= count) return;
            nxref_kawpow_hash(c, height, header, start + i, out + 64 * i, out + 64 * i + 32);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

int nxref_kawpow_verify(const void* c, int height, const uint8_t header[32], const uint8_t mix[32],
                        uint64_t nonce, const uint8_t boundary[32])
{
    hash256 h, m, b;
    std::memcpy(h.bytes, header, 32);
    std::memcpy(m.bytes, mix, 32);
    std::memcpy(b.bytes, boundary, 32);
    return progpow::verify(*static_cast<const epoch_context*>(c), height, h, m, nonce, b) ? 1 : 0;
}

void nxref_kawpow_hash_no_verify(int height, const uint8_t header[32], const uint8_t mix[32],
                                 uint64_t nonce, uint8_t out[32])
{
    hash256 h, m;
    std::memcpy(h.bytes, header, 32);
    std::memcpy(m.bytes, mix, 32);
    const hash256 r = progpow::hash_no_verify(height, h, m, nonce);
    std::memcpy(out, r.bytes, 32);
}

// returns 1 when found; nonce/final/mix filled (zeroed otherwise, as the reference does).
int nxref_kawpow_search_light(const void* c, int height, const uint8_t header[32],
                              const uint8_t boundary[32], uint64_t start, uint64_t iterations,
                              uint64_t* nonce, uint8_t final_out[32], uint8_t mix_out[32])
{
    hash256 h, b;
    std::memcpy(h.bytes, header, 32);
    std::memcpy(b.bytes, boundary, 32);
    const search_result r =
        progpow::search_light(*static_cast<const epoch_context*>(c), height, h, b, start, iterations);
    *nonce = r.nonce;
    std::memcpy(final_out, r.final_hash.bytes, 32);
    std::memcpy(mix_out, r.mix_hash.bytes, 32);
    return r.solution_found ? 1 : 0;
}

// ---- full-DAG context: the reference's best-case CPU path (progpow.cpp:357-429, lazy DAG) ----
// ethash_create_epoch_context_full callocs the whole dataset (zero = "not computed yet"); the caller may
// pre-fill it (e.g. with a DAG that was already verified) so that no lazy item generation happens while timing.
void* nxref_ctx_create_full(int epoch) { return ethash_create_epoch_context_full(epoch); }
void nxref_ctx_destroy_full(void* c) { ethash_destroy_epoch_context_full(static_cast<epoch_context_full*>(c)); }
void* nxref_ctx_full_dataset(void* c) { return static_cast<epoch_context_full*>(c)->full_dataset; }

void nxref_kawpow_hash_range_full(const void* c, int height, const uint8_t header[32], uint64_t start, uint64_t count,
                                  uint8_t* out, int threads)
{
    hash256 h;
    std::memcpy(h.bytes, header, 32);
    const epoch_context_full& ctx = *static_cast<const epoch_context_full*>(c);
    if (threads < 1) threads = 1;
    std::atomic<uint64_t> next{0};
    auto work = [&] {
        for (;;) {
            const uint64_t i0 = next.fetch_add(64);
            if (i0 >= count) return;
            for (uint64_t i = i0; i < i0 + 64 && i < count; ++i) {
                const result r = progpow::hash(ctx, height, h, start + i);
                if (out) {
                    std::memcpy(out + 64 * i, r.final_hash.bytes, 32);
                    std::memcpy(out + 64 * i + 32, r.mix_hash.bytes, 32);
                }
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

void nxref_dataset_item_512(const void* c, int64_t index, uint8_t out[64])
{
    const hash512 r = calculate_dataset_item_512(*static_cast<const epoch_context*>(c), index);
    std::memcpy(out, r.bytes, 64);
}

void nxref_dataset_item_2048(const void* c, uint32_t index, uint8_t out[256])
{
    const hash2048 r = calculate_dataset_item_2048(*static_cast<const epoch_context*>(c), index);
    std::memcpy(out, r.bytes, 256);
}

// count 2048-bit items starting at first, multi-threaded.
void nxref_dataset_items_2048(const void* c, uint32_t first, uint32_t count, uint8_t* out, int threads)
{
    if (threads < 1) threads = 1;
    std::atomic<uint32_t> next{0};
    auto work = [&] {
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= count) return;
            nxref_dataset_item_2048(c, first + i, out + 256 * size_t(i));
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

void nxref_ethash_hash(const void* c, const uint8_t header[32], uint64_t nonce, uint8_t final_out[32],
                       uint8_t mix_out[32])
{
    hash256 h;
    std::memcpy(h.bytes, header, 32);
    const ethash_result r = ethash_hash(static_cast<const epoch_context*>(c), &h, nonce);
    std::memcpy(final_out, r.final_hash.bytes, 32);
    std::memcpy(mix_out, r.mix_hash.bytes, 32);
}

void nxref_epoch_seed(int epoch, uint8_t out[32])
{
    const ethash_hash256 s = ethash_calculate_epoch_seed(epoch);
    std::memcpy(out, s.bytes, 32);
}

int nxref_find_epoch_number(const uint8_t seed[32])
{
    hash256 s;
    std::memcpy(s.bytes, seed, 32);
    return find_epoch_number(s);
}

int nxref_light_cache_num_items(int epoch) { return ethash_calculate_light_cache_num_items(epoch); }
int nxref_full_dataset_num_items(int epoch) { return ethash_calculate_full_dataset_num_items(epoch); }

void nxref_keccak256(const uint8_t* d, size_t n, uint8_t out[32])
{
    const ethash_hash256 r = ethash_keccak256(d, n);
    std::memcpy(out, r.bytes, 32);
}
void nxref_keccak512(const uint8_t* d, size_t n, uint8_t out[64])
{
    const ethash_hash512 r = ethash_keccak512(d, n);
    std::memcpy(out, r.bytes, 64);
}
void nxref_keccakf800(uint32_t st[25]) { ethash_keccakf800(st); }
void nxref_keccakf1600(uint64_t st[25]) { ethash_keccakf1600(st); }

}  // extern "C"
