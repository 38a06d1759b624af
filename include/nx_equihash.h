/* nx_equihash.h — GATED "next" row: a parameterised Equihash solver + GPU-side validator.
 *
 * There is NO reference interface to cite: DeonDavisV/Nodexa-Chain-Core contains no Equihash solver, validator,
 * (n,k), personalization, header field or test (SURVEY.md §0 Fact 1; the only mention is a sentence in README.md:7).
 * Parity with the reference is therefore UNPINNED by construction.  (n,k) and the 8-byte personalization tag are
 * inputs; results are checked against an independent CPU implementation of the published algorithm
 * (oracle/nx_equihash_oracle.c) — self-consistency only.  Whoever defines the chain's parameters can plug them in.
 *
 * Conventions (the ones popularised by Zcash, stated so that they can be changed in one place):
 *   string i = bytes [(i mod m)*n/8, +n/8) of BLAKE2b-(m*n/8)(personal = tag || LE32(n) || LE32(k);
 *              header || LE32(i div m)),  m = 512 div n;   header = everything hashed before the counter, nonce included;
 *   a solution = 2^k distinct indices < 2^(n/(k+1)+1), tree rule "left subtree's first index < right's";
 *   encoding   = the indices as big-endian (n/(k+1)+1)-bit fields, packed MSB first (2^k*(n/(k+1)+1)/8 bytes).
 * Supported: n % 8 == 0, n % (k+1) == 0, 8 <= n/(k+1) <= 24, k <= 9  — e.g. (200,9), (144,5), (192,7), (96,5), (48,5).
 */
#ifndef NX_EQUIHASH_H
#define NX_EQUIHASH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nx_equihash nx_equihash; /* device buffers of one (device, n, k, tag) */

size_t nx_equihash_solution_bytes(int n, int k); /* 0 for unsupported parameters */
int nx_equihash_create(int device, int n, int k, const uint8_t tag[8], nx_equihash** out);
void nx_equihash_destroy(nx_equihash* solver);

/* Every solution of one header.  solutions_out receives up to max_solutions encodings, sorted bytewise and
 * without duplicates; *n_solutions is the number found (it may exceed max_solutions). */
int nx_equihash_solve(nx_equihash* solver, const uint8_t* header, size_t header_len, uint8_t* solutions_out,
                      int max_solutions, int* n_solutions);

/* The same for n headers of header_len bytes each, laid end to end, with two solves kept in flight (a second set of device
 * buffers is created on the first call): counts[i] receives the number of solutions of header i, solutions_out their
 * encodings in header order (each header's sorted, without duplicates), at most max_solutions_total of them; *wall_ms (may
 * be NULL) the host time of the whole call.  nx_equihash_last_dropped / _candidates then refer to the whole batch. */
int nx_equihash_solve_batch(nx_equihash* solver, const uint8_t* headers, size_t header_len, int n, uint8_t* solutions_out,
                            int max_solutions_total, int* counts, float* wall_ms);

/* GPU-side validator: recomputes the BLAKE2b strings of `count` claimed solutions of one header and re-checks the
 * collisions of every round, the ordering rule, distinctness and the final zero XOR.  verdict[i] = 1 / 0. */
int nx_equihash_validate(nx_equihash* solver, const uint8_t* header, size_t header_len, const uint8_t* solutions,
                         int count, uint8_t* verdict);

/* Rows or candidates that did not fit their bucket during the last solve and were lost (each bucket has twice its
 * mean load in slots, so this is 0 in normal operation; a non-zero value means solutions may have been missed). */
int nx_equihash_last_dropped(const nx_equihash* solver);
/* Final-round candidates of the last solve (valid solutions plus trees that repeat an index). */
int nx_equihash_last_candidates(const nx_equihash* solver);

/* Device time (CUDA events on the solver's stream) and kernel launches of the last nx_equihash_solve. */
int nx_equihash_last_timing(const nx_equihash* solver, float* kernel_ms, int* launches);

#ifdef __cplusplus
}
#endif
#endif
