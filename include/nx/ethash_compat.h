/* C view of the reference's Ethash library interface, served by libnxkawpow.so.
 *
 * Same type names, field order and function signatures as the reference's
 * src/crypto/ethash/include/ethash/{hash_types.h:14-46, ethash.h:20-131, keccak.h:30-45} and
 * lib/ethash/primes.h:21, so that node code written against those headers compiles and links
 * unchanged (the reference paths are provided as one-line forwarders under include/crypto/ethash/).
 * Every function here is a thin wrapper over the nx_* C ABI (include/nx_kawpow.h).
 */
#ifndef NX_ETHASH_COMPAT_H
#define NX_ETHASH_COMPAT_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
#define NX_NOEXCEPT noexcept
extern "C" {
#else
#define NX_NOEXCEPT
#endif

/* hash_types.h:14-46 */
union ethash_hash256  { uint64_t word64s[4];  uint32_t word32s[8];  uint8_t bytes[32];  char str[32]; };
union ethash_hash512  { uint64_t word64s[8];  uint32_t word32s[16]; uint8_t bytes[64];  char str[64]; };
union ethash_hash1024 { union ethash_hash512 hash512s[2]; uint64_t word64s[16]; uint32_t word32s[32]; uint8_t bytes[128]; char str[128]; };
union ethash_hash2048 { union ethash_hash512 hash512s[4]; uint64_t word64s[32]; uint32_t word32s[64]; uint8_t bytes[256]; char str[256]; };

/* ethash.h:26-32 */
#define ETHASH_REVISION "23"
#define ETHASH_EPOCH_LENGTH 7500
#define ETHASH_LIGHT_CACHE_ITEM_SIZE 64
#define ETHASH_FULL_DATASET_ITEM_SIZE 128
#define ETHASH_NUM_DATASET_ACCESSES 64

/* ethash.h:35-42 — callers read these fields directly.  light_cache and l1_cache point at HOST mirrors;
 * the DAG itself lives in HBM behind the opaque tail of the object. */
struct ethash_epoch_context
{
    const int epoch_number;
    const int light_cache_num_items;
    const union ethash_hash512* const light_cache;
    const uint32_t* const l1_cache;
    const int full_dataset_num_items;
};
struct ethash_epoch_context_full; /* opaque, as in ethash.h:45 */

struct ethash_result /* ethash.h:48-52 */
{
    union ethash_hash256 final_hash;
    union ethash_hash256 mix_hash;
};

int ethash_calculate_light_cache_num_items(int epoch_number) NX_NOEXCEPT;
int ethash_calculate_full_dataset_num_items(int epoch_number) NX_NOEXCEPT;
union ethash_hash256 ethash_calculate_epoch_seed(int epoch_number) NX_NOEXCEPT;

/* NULL on failure (the reference returns NULL on out-of-memory, ethash.cpp:147-148); nx_last_error() says why.
 * The GPU is chosen by the environment variable NX_DEVICE (default 0). */
struct ethash_epoch_context* ethash_create_epoch_context(int epoch_number) NX_NOEXCEPT;
struct ethash_epoch_context_full* ethash_create_epoch_context_full(int epoch_number) NX_NOEXCEPT;
void ethash_destroy_epoch_context(struct ethash_epoch_context* context) NX_NOEXCEPT;
void ethash_destroy_epoch_context_full(struct ethash_epoch_context_full* context) NX_NOEXCEPT;

/* managed.cpp:84-100 — borrowed pointers, valid until the calling thread asks for another epoch. */
const struct ethash_epoch_context* ethash_get_global_epoch_context(int epoch_number) NX_NOEXCEPT;
const struct ethash_epoch_context_full* ethash_get_global_epoch_context_full(int epoch_number) NX_NOEXCEPT;

struct ethash_result ethash_hash(const struct ethash_epoch_context* context, const union ethash_hash256* header_hash,
                                 uint64_t nonce) NX_NOEXCEPT;
bool ethash_verify(const struct ethash_epoch_context* context, const union ethash_hash256* header_hash,
                   const union ethash_hash256* mix_hash, uint64_t nonce, const union ethash_hash256* boundary) NX_NOEXCEPT;
bool ethash_verify_final_hash(const union ethash_hash256* header_hash, const union ethash_hash256* mix_hash,
                              uint64_t nonce, const union ethash_hash256* boundary) NX_NOEXCEPT;

/* keccak.h:30-45 */
void ethash_keccakf1600(uint64_t state[25]) NX_NOEXCEPT;
void ethash_keccakf800(uint32_t state[25]) NX_NOEXCEPT;
union ethash_hash256 ethash_keccak256(const uint8_t* data, size_t size) NX_NOEXCEPT;
union ethash_hash256 ethash_keccak256_32(const uint8_t data[32]) NX_NOEXCEPT;
union ethash_hash512 ethash_keccak512(const uint8_t* data, size_t size) NX_NOEXCEPT;
union ethash_hash512 ethash_keccak512_64(const uint8_t data[64]) NX_NOEXCEPT;

/* primes.h:21 */
int ethash_find_largest_prime(int upper_bound) NX_NOEXCEPT;

#ifdef __cplusplus
}
#endif
#endif
