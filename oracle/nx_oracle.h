/* TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's KawPow/Ethash path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this.  The product library (nodexa_chain_core_b200/csrc) never links,
 * loads or calls it.
 *
 * Parity status: PINNED — checked against every KAT the reference holds for this path
 * (progpow_test_vectors.hpp:22-62, kawpow_tests.cpp:29-33,52-53,60-69,115-139) and
 * against the reference's own objects compiled in place (oracle/_ref/libnxref.so);
 * see tests/test_oracle.py.
 *
 * Each function cites the reference file:line (under src/crypto/ethash/ unless noted)
 * whose behaviour it restates.
 */
#ifndef NX_ORACLE_H
#define NX_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- Keccak (lib/keccak/keccak.c:46-127, keccakf1600.c:40-255, keccakf800.c:38-253) ---- */
void nxo_keccakf1600(uint64_t st[25]);
void nxo_keccakf800(uint32_t st[25]);
void nxo_keccak256(const uint8_t* data, size_t size, uint8_t out[32]);
void nxo_keccak512(const uint8_t* data, size_t size, uint8_t out[64]);

/* ---- epoch geometry (lib/ethash/ethash.cpp:357-393, primes.c:24-43, ethash.cpp:57-97) ---- */
int nxo_find_largest_prime(int upper_bound);
int nxo_light_cache_num_items(int epoch);
int nxo_full_dataset_num_items(int epoch);
void nxo_epoch_seed(int epoch, uint8_t out[32]);
int nxo_find_epoch_number(const uint8_t seed[32]);

/* ---- epoch context (lib/ethash/ethash.cpp:101-172) ---- */
typedef struct nxo_ctx
{
    int epoch;
    int light_items;     /* 64-byte items */
    int dataset_items;   /* 128-byte items (ProgPoW uses dataset_items/2 items of 256 bytes) */
    uint8_t* light;      /* light_items * 64 bytes */
    uint32_t l1[4096];   /* first 16 KiB of the DAG */
    const uint8_t* dag;  /* optional: full DAG as 256-byte items, borrowed (NULL = light path) */
} nxo_ctx;

nxo_ctx* nxo_ctx_create(int epoch);
void nxo_ctx_destroy(nxo_ctx* c);
void nxo_ctx_set_dag(nxo_ctx* c, const uint8_t* dag256);
void nxo_build_light_cache(uint8_t* cache, int num_items, const uint8_t seed[32]);

/* ---- DAG items (lib/ethash/ethash.cpp:180-251) ---- */
void nxo_dataset_item_512(const nxo_ctx* c, int64_t index, uint8_t out[64]);
void nxo_dataset_item_2048(const nxo_ctx* c, uint32_t index, uint8_t out[256]);
void nxo_dataset_items_2048(const nxo_ctx* c, uint32_t first, uint32_t count, uint8_t* out, int threads);

/* ---- KawPow (lib/ethash/progpow.cpp:46-579) ---- */
void nxo_kawpow_hash(const nxo_ctx* c, int height, const uint8_t header[32], uint64_t nonce,
                     uint8_t final_out[32], uint8_t mix_out[32]);
void nxo_kawpow_hash_range(const nxo_ctx* c, int height, const uint8_t header[32], uint64_t start,
                           uint64_t count, uint8_t* out /* count x {final,mix} */, int threads);
int nxo_kawpow_verify(const nxo_ctx* c, int height, const uint8_t header[32], const uint8_t mix[32],
                      uint64_t nonce, const uint8_t boundary[32]);
void nxo_kawpow_hash_no_verify(int height, const uint8_t header[32], const uint8_t mix[32],
                               uint64_t nonce, uint8_t out[32]);
int nxo_kawpow_search(const nxo_ctx* c, int height, const uint8_t header[32], const uint8_t boundary[32],
                      uint64_t start, uint64_t iterations, uint64_t* nonce, uint8_t final_out[32],
                      uint8_t mix_out[32]);
/* multi-threaded: lowest qualifying nonce in [start,start+iterations) — same answer as the serial scan */
int nxo_kawpow_search_mt(const nxo_ctx* c, int height, const uint8_t header[32], const uint8_t boundary[32],
                         uint64_t start, uint64_t iterations, uint64_t* nonce, uint8_t final_out[32],
                         uint8_t mix_out[32], int threads);

/* per-period program as a flat op list, for checking the product's code generator
 * (progpow.cpp:46-88 mix_rng_state, :179-233 round's draw order).
 * ops[i] = {kind, a, b, c, d}: kind 0 cache {src,dst,sel,0}; kind 1 math {src1,src2,sel1,dst,sel2 in ops[i][4]};
 * kind 2 dag {dst,sel}.  Returns number of ops written (11+18+4 = 33). */
int nxo_kawpow_program(uint64_t prog_seed, uint32_t ops[33][6]);

/* ---- classic Ethash (lib/ethash/ethash.cpp:257-303,416-440) ---- */
void nxo_ethash_hash(const nxo_ctx* c, const uint8_t header[32], uint64_t nonce, uint8_t final_out[32],
                     uint8_t mix_out[32]);
int nxo_ethash_verify_final(const uint8_t header[32], const uint8_t mix[32], uint64_t nonce,
                            const uint8_t boundary[32]);

/* ---- 256-bit compare as big-endian bytes (lib/ethash/ethash-internal.hpp:32-47) ---- */
int nxo_less_or_equal(const uint8_t a[32], const uint8_t b[32]);

/* ---- node glue (src/primitives/block.cpp:94-99, block.h:213-233, src/hash.cpp:258-291,
 *      src/arith_uint256.cpp:209-228, src/pow.cpp:182-199) ---- */
void nxo_sha256(const uint8_t* data, size_t n, uint8_t out[32]);
/* header_hash fed to progpow = byte-reversed SHA256d of the 80-byte CKAWPOWInput */
void nxo_kawpow_header_hash(int32_t version, const uint8_t prev[32], const uint8_t merkle[32], uint32_t time,
                            uint32_t bits, uint32_t height, uint8_t out[32]);
/* compact nBits -> 32-byte big-endian boundary; returns 0 if negative/zero/overflow */
int nxo_compact_to_boundary(uint32_t bits, uint8_t out[32]);

#ifdef __cplusplus
}
#endif
#endif
