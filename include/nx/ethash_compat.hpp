// C++ view of the reference's ethash:: and progpow:: namespaces, served by libnxkawpow.so.
// Mirrors src/crypto/ethash/include/ethash/{hash_types.hpp, ethash.hpp:24-172, keccak.hpp, progpow.hpp:13-49}:
// same names, signatures and noexcept contract, so src/hash.cpp, src/rpc/mining.cpp and
// src/test/kawpow_tests.cpp build against it unchanged.  The out-of-line functions are defined in
// nodexa_chain_core_b200/csrc/nx_compat.cpp on top of the nx_* C ABI.
#pragma once
#include "ethash_compat.h"

#include <cstring>
#include <memory>

namespace ethash
{
using hash256 = ethash_hash256;
using hash512 = ethash_hash512;
using hash1024 = ethash_hash1024;
using hash2048 = ethash_hash2048;
using epoch_context = ethash_epoch_context;
using epoch_context_full = ethash_epoch_context_full;
using result = ethash_result;

constexpr auto revision = ETHASH_REVISION;
static constexpr int epoch_length = ETHASH_EPOCH_LENGTH;
static constexpr int light_cache_item_size = ETHASH_LIGHT_CACHE_ITEM_SIZE;
static constexpr int full_dataset_item_size = ETHASH_FULL_DATASET_ITEM_SIZE;
static constexpr int num_dataset_accesses = ETHASH_NUM_DATASET_ACCESSES;

struct search_result  // ethash.hpp:49-61
{
    bool solution_found = false;
    uint64_t nonce = 0;
    hash256 final_hash = {};
    hash256 mix_hash = {};
    search_result() noexcept = default;
    search_result(result r, uint64_t n) noexcept : solution_found(true), nonce(n), final_hash(r.final_hash), mix_hash(r.mix_hash) {}
};

inline hash256 hash256_from_bytes(const uint8_t bytes[32]) noexcept
{
    hash256 h;
    std::memcpy(&h, bytes, sizeof h);
    return h;
}

static constexpr auto calculate_light_cache_num_items = ethash_calculate_light_cache_num_items;
static constexpr auto calculate_full_dataset_num_items = ethash_calculate_full_dataset_num_items;
static constexpr auto calculate_epoch_seed = ethash_calculate_epoch_seed;

inline constexpr int get_epoch_number(int block_number) noexcept { return block_number ? block_number / epoch_length : 0; }
inline constexpr size_t get_light_cache_size(int num_items) noexcept { return static_cast<size_t>(num_items) * light_cache_item_size; }
inline constexpr uint64_t get_full_dataset_size(int num_items) noexcept { return static_cast<uint64_t>(num_items) * full_dataset_item_size; }

using epoch_context_ptr = std::unique_ptr<epoch_context, decltype(&ethash_destroy_epoch_context)>;
using epoch_context_full_ptr = std::unique_ptr<epoch_context_full, decltype(&ethash_destroy_epoch_context_full)>;
inline epoch_context_ptr create_epoch_context(int epoch_number) noexcept { return {ethash_create_epoch_context(epoch_number), ethash_destroy_epoch_context}; }
inline epoch_context_full_ptr create_epoch_context_full(int epoch_number) noexcept { return {ethash_create_epoch_context_full(epoch_number), ethash_destroy_epoch_context_full}; }

inline result hash(const epoch_context& context, const hash256& header_hash, uint64_t nonce) noexcept { return ethash_hash(&context, &header_hash, nonce); }
result hash(const epoch_context_full& context, const hash256& header_hash, uint64_t nonce) noexcept;
inline bool verify_final_hash(const hash256& header_hash, const hash256& mix_hash, uint64_t nonce, const hash256& boundary) noexcept { return ethash_verify_final_hash(&header_hash, &mix_hash, nonce, &boundary); }
inline bool verify(const epoch_context& context, const hash256& header_hash, const hash256& mix_hash, uint64_t nonce, const hash256& boundary) noexcept { return ethash_verify(&context, &header_hash, &mix_hash, nonce, &boundary); }
search_result search_light(const epoch_context& context, const hash256& header_hash, const hash256& boundary, uint64_t start_nonce, size_t iterations) noexcept;
search_result search(const epoch_context_full& context, const hash256& header_hash, const hash256& boundary, uint64_t start_nonce, size_t iterations) noexcept;
int find_epoch_number(const hash256& seed) noexcept;

inline const epoch_context& get_global_epoch_context(int epoch_number) noexcept { return *ethash_get_global_epoch_context(epoch_number); }
inline const epoch_context_full& get_global_epoch_context_full(int epoch_number) noexcept { return *ethash_get_global_epoch_context_full(epoch_number); }

// keccak.hpp
inline hash256 keccak256(const uint8_t* data, size_t size) noexcept { return ethash_keccak256(data, size); }
inline hash256 keccak256(const hash256& input) noexcept { return ethash_keccak256_32(input.bytes); }
inline hash512 keccak512(const uint8_t* data, size_t size) noexcept { return ethash_keccak512(data, size); }
inline hash512 keccak512(const hash512& input) noexcept { return ethash_keccak512_64(input.bytes); }
static constexpr auto keccak256_32 = ethash_keccak256_32;
static constexpr auto keccak512_64 = ethash_keccak512_64;

// lib/ethash/endianness.hpp: the tests call ethash::le::uint32(); this library targets little-endian hosts only.
namespace le
{
inline uint32_t uint32(uint32_t x) noexcept { return x; }
inline uint64_t uint64(uint64_t x) noexcept { return x; }
}  // namespace le
}  // namespace ethash

namespace progpow
{
using namespace ethash;

constexpr auto revision = "0.9.4";
constexpr int period_length = 3;
constexpr uint32_t num_regs = 32;
constexpr size_t num_lanes = 16;
constexpr int num_cache_accesses = 11;
constexpr int num_math_operations = 18;
constexpr size_t l1_cache_size = 16 * 1024;
constexpr size_t l1_cache_num_items = l1_cache_size / sizeof(uint32_t);

result hash(const epoch_context& context, int block_number, const hash256& header_hash, uint64_t nonce) noexcept;
result hash(const epoch_context_full& context, int block_number, const hash256& header_hash, uint64_t nonce) noexcept;
bool verify(const epoch_context& context, int block_number, const hash256& header_hash, const hash256& mix_hash, uint64_t nonce, const hash256& boundary) noexcept;
hash256 hash_no_verify(const int& block_number, const hash256& header_hash, const hash256& mix_hash, const uint64_t& nonce) noexcept;
search_result search_light(const epoch_context& context, int block_number, const hash256& header_hash, const hash256& boundary, uint64_t start_nonce, size_t iterations) noexcept;
search_result search(const epoch_context_full& context, int block_number, const hash256& header_hash, const hash256& boundary, uint64_t start_nonce, size_t iterations) noexcept;
}  // namespace progpow
