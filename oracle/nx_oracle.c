/* TEST INFRASTRUCTURE ONLY — see nx_oracle.h.  Plain-C CPU restatement of the reference's
 * KawPow / Ethash path, written from the reference's behaviour (file:line cited per function,
 * paths relative to /root/reference/src/crypto/ethash/ unless noted).  Little-endian hosts only
 * (the reference's le:: helpers are identities there, lib/ethash/endianness.hpp). */
#define _GNU_SOURCE
#include "nx_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#if defined(__BYTE_ORDER__) && __BYTE_ORDER__ != __ORDER_LITTLE_ENDIAN__
#error "oracle assumes a little-endian host"
#endif

/* ------------------------------------------------------------------ Keccak */

static const uint64_t RC64[24] = {
    0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
    0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
    0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
    0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
    0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
    0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};

/* rho offsets indexed [x + 5*y] (Keccak reference, section 2.3.4) */
static const unsigned RHO[25] = {0,  1,  62, 28, 27, 36, 44, 6,  55, 20, 3,  10, 43,
                                 25, 39, 41, 45, 15, 21, 8,  18, 2,  61, 56, 14};

static inline uint64_t rol64(uint64_t v, unsigned n) { return n ? (v << n) | (v >> (64 - n)) : v; }
static inline uint32_t rol32(uint32_t v, unsigned n)
{
    n &= 31;
    return n ? (v << n) | (v >> (32 - n)) : v;
}
static inline uint32_t ror32(uint32_t v, unsigned n)
{
    n &= 31;
    return n ? (v >> n) | (v << (32 - n)) : v;
}

/* One Keccak round on lane type T (theta, rho+pi, chi); iota is applied by the caller.
 * Written as a macro so both widths share it and the compiler sees constant indices. */
#define PI_DST(x, y) ((y) + 5 * ((2 * (x) + 3 * (y)) % 5))
#define KECCAK_ROUND(T, ROL, BITS, a)                                                              \
    do {                                                                                           \
        T c0 = a[0] ^ a[5] ^ a[10] ^ a[15] ^ a[20], c1 = a[1] ^ a[6] ^ a[11] ^ a[16] ^ a[21],      \
          c2 = a[2] ^ a[7] ^ a[12] ^ a[17] ^ a[22], c3 = a[3] ^ a[8] ^ a[13] ^ a[18] ^ a[23],      \
          c4 = a[4] ^ a[9] ^ a[14] ^ a[19] ^ a[24];                                                \
        const T d0 = c4 ^ ROL(c1, 1), d1 = c0 ^ ROL(c2, 1), d2 = c1 ^ ROL(c3, 1),                  \
                d3 = c2 ^ ROL(c4, 1), d4 = c3 ^ ROL(c0, 1);                                        \
        T b[25];                                                                                   \
        _Pragma("GCC unroll 5") for (int y = 0; y < 5; ++y)                                        \
        {                                                                                          \
            b[PI_DST(0, y)] = ROL(a[0 + 5 * y] ^ d0, RHO[0 + 5 * y] % BITS);                       \
            b[PI_DST(1, y)] = ROL(a[1 + 5 * y] ^ d1, RHO[1 + 5 * y] % BITS);                       \
            b[PI_DST(2, y)] = ROL(a[2 + 5 * y] ^ d2, RHO[2 + 5 * y] % BITS);                       \
            b[PI_DST(3, y)] = ROL(a[3 + 5 * y] ^ d3, RHO[3 + 5 * y] % BITS);                       \
            b[PI_DST(4, y)] = ROL(a[4 + 5 * y] ^ d4, RHO[4 + 5 * y] % BITS);                       \
        }                                                                                          \
        _Pragma("GCC unroll 5") for (int y = 0; y < 25; y += 5)                                    \
        {                                                                                          \
            a[y + 0] = b[y + 0] ^ (~b[y + 1] & b[y + 2]);                                          \
            a[y + 1] = b[y + 1] ^ (~b[y + 2] & b[y + 3]);                                          \
            a[y + 2] = b[y + 2] ^ (~b[y + 3] & b[y + 4]);                                          \
            a[y + 3] = b[y + 3] ^ (~b[y + 4] & b[y + 0]);                                          \
            a[y + 4] = b[y + 4] ^ (~b[y + 0] & b[y + 1]);                                          \
        }                                                                                          \
    } while (0)

/* Keccak-f[1600], 24 rounds: restates lib/keccak/keccakf1600.c:40-255 (the same permutation, unrolled there). */
void nxo_keccakf1600(uint64_t st[25])
{
    uint64_t a[25];
    memcpy(a, st, sizeof a);
    for (int round = 0; round < 24; ++round) {
        KECCAK_ROUND(uint64_t, rol64, 64, a);
        a[0] ^= RC64[round];
    }
    memcpy(st, a, sizeof a);
}

/* Keccak-f[800], 22 rounds, 32-bit lanes: restates lib/keccak/keccakf800.c:38-253.
 * Round constants are the low 32 bits of the f1600 constants; rho offsets are taken mod 32. */
void nxo_keccakf800(uint32_t st[25])
{
    uint32_t a[25];
    memcpy(a, st, sizeof a);
    for (int round = 0; round < 22; ++round) {
        KECCAK_ROUND(uint32_t, rol32, 32, a);
        a[0] ^= (uint32_t)RC64[round];
    }
    memcpy(st, a, sizeof a);
}

/* Original-Keccak sponge (pad 0x01 ... 0x80), rate = 200 - 2*outlen: lib/keccak/keccak.c:46-101. */
static void keccak_sponge(uint8_t* out, size_t outlen, const uint8_t* data, size_t size)
{
    const size_t rate = 200 - 2 * outlen;
    uint64_t st[25];
    uint8_t block[200];
    memset(st, 0, sizeof st);
    while (size >= rate) {
        uint64_t w[25];
        memcpy(w, data, rate);
        for (size_t i = 0; i < rate / 8; ++i) st[i] ^= w[i];
        nxo_keccakf1600(st);
        data += rate;
        size -= rate;
    }
    memset(block, 0, sizeof block);
    memcpy(block, data, size);
    block[size] ^= 0x01;
    block[rate - 1] ^= 0x80;
    {
        uint64_t w[25];
        memcpy(w, block, rate);
        for (size_t i = 0; i < rate / 8; ++i) st[i] ^= w[i];
    }
    nxo_keccakf1600(st);
    memcpy(out, st, outlen);
}

void nxo_keccak256(const uint8_t* data, size_t size, uint8_t out[32]) { keccak_sponge(out, 32, data, size); }
void nxo_keccak512(const uint8_t* data, size_t size, uint8_t out[64]) { keccak_sponge(out, 64, data, size); }

/* ------------------------------------------------------------------ epoch geometry */

/* lib/ethash/primes.c:24-43 */
int nxo_find_largest_prime(int upper_bound)
{
    if (upper_bound < 2) return 0;
    if (upper_bound == 2) return 2;
    for (int n = (upper_bound % 2) ? upper_bound : upper_bound - 1;; n -= 2) {
        int prime = 1;
        for (int64_t d = 3; d * d <= n; d += 2)
            if (n % d == 0) {
                prime = 0;
                break;
            }
        if (prime) return n;
    }
}

/* lib/ethash/ethash.cpp:22-26,365-393: 2^24 B + epoch*2^17 B of 64-byte items; 2^30 B + epoch*2^23 B of 128-byte items */
int nxo_light_cache_num_items(int epoch) { return nxo_find_largest_prime((1 << 24) / 64 + epoch * ((1 << 17) / 64)); }
int nxo_full_dataset_num_items(int epoch) { return nxo_find_largest_prime((1 << 30) / 128 + epoch * ((1 << 23) / 128)); }

/* lib/ethash/ethash.cpp:357-363 */
void nxo_epoch_seed(int epoch, uint8_t out[32])
{
    memset(out, 0, 32);
    for (int i = 0; i < epoch; ++i) nxo_keccak256(out, 32, out);
}

/* lib/ethash/ethash.cpp:57-97 (without the thread-local memo, which only changes speed);
 * the reference compares the first 32-bit word only. */
int nxo_find_epoch_number(const uint8_t seed[32])
{
    uint8_t s[32];
    memset(s, 0, 32);
    for (int i = 0; i < 30000; ++i) {
        if (memcmp(s, seed, 4) == 0) return i;
        nxo_keccak256(s, 32, s);
    }
    return -1;
}

/* ------------------------------------------------------------------ FNV, KISS99 */

#define FNV_PRIME 0x01000193u
#define FNV_BASIS 0x811c9dc5u
static inline uint32_t fnv1(uint32_t u, uint32_t v) { return (u * FNV_PRIME) ^ v; }  /* bit_manipulation.h:62-65 */
static inline uint32_t fnv1a(uint32_t u, uint32_t v) { return (u ^ v) * FNV_PRIME; } /* bit_manipulation.h:73-76 */

typedef struct
{
    uint32_t z, w, jsr, jcong;
} kiss99_t;

/* lib/ethash/kiss99.hpp:48-61 */
static inline uint32_t kiss99(kiss99_t* s)
{
    s->z = 36969u * (s->z & 0xffff) + (s->z >> 16);
    s->w = 18000u * (s->w & 0xffff) + (s->w >> 16);
    s->jcong = 69069u * s->jcong + 1234567u;
    s->jsr ^= s->jsr << 17;
    s->jsr ^= s->jsr >> 13;
    s->jsr ^= s->jsr << 5;
    return (((s->z << 16) + s->w) ^ s->jcong) + s->jsr;
}

/* ------------------------------------------------------------------ light cache, DAG */

/* lib/ethash/ethash.cpp:101-129 */
void nxo_build_light_cache(uint8_t* cache, int n, const uint8_t seed[32])
{
    nxo_keccak512(seed, 32, cache);
    for (int i = 1; i < n; ++i) nxo_keccak512(cache + 64 * (size_t)(i - 1), 64, cache + 64 * (size_t)i);
    for (int q = 0; q < 3; ++q) {
        for (int i = 0; i < n; ++i) {
            uint32_t first;
            uint64_t x[8], a[8], b[8];
            memcpy(&first, cache + 64 * (size_t)i, 4);
            const uint32_t v = first % (uint32_t)n;
            const uint32_t w = (uint32_t)(n + (i - 1)) % (uint32_t)n;
            memcpy(a, cache + 64 * (size_t)v, 64);
            memcpy(b, cache + 64 * (size_t)w, 64);
            for (int k = 0; k < 8; ++k) x[k] = a[k] ^ b[k];
            nxo_keccak512((const uint8_t*)x, 64, cache + 64 * (size_t)i);
        }
    }
}

/* lib/ethash/ethash.cpp:180-215: one 64-byte DAG item, 512 parents */
void nxo_dataset_item_512(const nxo_ctx* c, int64_t index, uint8_t out[64])
{
    const uint32_t n = (uint32_t)c->light_items;
    const uint32_t seed = (uint32_t)index;
    uint32_t mix[16];
    memcpy(mix, c->light + 64 * (size_t)(index % c->light_items), 64);
    mix[0] ^= seed;
    nxo_keccak512((const uint8_t*)mix, 64, (uint8_t*)mix);
    for (uint32_t j = 0; j < 512; ++j) {
        const uint32_t t = fnv1(seed ^ j, mix[j % 16]);
        uint32_t parent[16];
        memcpy(parent, c->light + 64 * (size_t)(t % n), 64);
        for (int k = 0; k < 16; ++k) mix[k] = fnv1(mix[k], parent[k]);
    }
    nxo_keccak512((const uint8_t*)mix, 64, out);
}

/* lib/ethash/ethash.cpp:235-251: 256-byte item i = 64-byte items 4i..4i+3, the four parent walks
 * advanced in lockstep (as the reference does) so their cache misses overlap. */
void nxo_dataset_item_2048(const nxo_ctx* c, uint32_t index, uint8_t out[256])
{
    const uint32_t n = (uint32_t)c->light_items;
    uint32_t mix[4][16], seed[4];
    for (int k = 0; k < 4; ++k) {
        const int64_t idx = (int64_t)index * 4 + k;
        seed[k] = (uint32_t)idx;
        memcpy(mix[k], c->light + 64 * (size_t)(idx % c->light_items), 64);
        mix[k][0] ^= seed[k];
        nxo_keccak512((const uint8_t*)mix[k], 64, (uint8_t*)mix[k]);
    }
    for (uint32_t j = 0; j < 512; ++j) {
        const uint32_t* parent[4];
        for (int k = 0; k < 4; ++k)
            parent[k] = (const uint32_t*)(c->light + 64 * (size_t)(fnv1(seed[k] ^ j, mix[k][j % 16]) % n));
        for (int k = 0; k < 4; ++k)
            for (int w = 0; w < 16; ++w) mix[k][w] = fnv1(mix[k][w], parent[k][w]);
    }
    for (int k = 0; k < 4; ++k) nxo_keccak512((const uint8_t*)mix[k], 64, out + 64 * k);
}

typedef struct
{
    const nxo_ctx* c;
    uint32_t first, count;
    uint8_t* out;
    volatile uint32_t* next;
} items_job;

static void* items_worker(void* p)
{
    items_job* j = (items_job*)p;
    for (;;) {
        const uint32_t i = __atomic_fetch_add(j->next, 64, __ATOMIC_RELAXED);
        if (i >= j->count) return NULL;
        for (uint32_t k = i; k < i + 64 && k < j->count; ++k)
            nxo_dataset_item_2048(j->c, j->first + k, j->out + 256 * (size_t)k);
    }
}

static void run_threads(void* (*fn)(void*), void* arg, int threads)
{
    pthread_t th[256];
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    for (int t = 1; t < threads; ++t) pthread_create(&th[t], NULL, fn, arg);
    fn(arg);
    for (int t = 1; t < threads; ++t) pthread_join(th[t], NULL);
}

void nxo_dataset_items_2048(const nxo_ctx* c, uint32_t first, uint32_t count, uint8_t* out, int threads)
{
    volatile uint32_t next = 0;
    items_job j = {c, first, count, out, &next};
    run_threads(items_worker, &j, threads);
}

/* lib/ethash/ethash.cpp:131-172: sizes, seed, light cache, then L1 = first 64 256-byte DAG items */
nxo_ctx* nxo_ctx_create(int epoch)
{
    nxo_ctx* c = (nxo_ctx*)calloc(1, sizeof *c);
    uint8_t seed[32];
    if (!c) return NULL;
    c->epoch = epoch;
    c->light_items = nxo_light_cache_num_items(epoch);
    c->dataset_items = nxo_full_dataset_num_items(epoch);
    c->light = (uint8_t*)malloc((size_t)c->light_items * 64);
    if (!c->light) {
        free(c);
        return NULL;
    }
    nxo_epoch_seed(epoch, seed);
    nxo_build_light_cache(c->light, c->light_items, seed);
    for (uint32_t i = 0; i < 64; ++i) nxo_dataset_item_2048(c, i, (uint8_t*)c->l1 + 256 * i);
    return c;
}

void nxo_ctx_destroy(nxo_ctx* c)
{
    if (!c) return;
    free(c->light);
    free(c);
}

void nxo_ctx_set_dag(nxo_ctx* c, const uint8_t* dag256) { c->dag = dag256; }

/* ------------------------------------------------------------------ KawPow */

enum { REGS = 32, LANES = 16, CACHE_OPS = 11, MATH_OPS = 18, DAG_WORDS = 4 }; /* include/ethash/progpow.hpp:21-27 */

/* "rAVENCOINKAWPOW" absorb padding, progpow.cpp:157-173 */
static const uint32_t KAWPOW_PAD[15] = {0x72, 0x41, 0x56, 0x45, 0x4E, 0x43, 0x4F, 0x49,
                                        0x4E, 0x4B, 0x41, 0x57, 0x50, 0x4F, 0x57};

typedef struct
{
    kiss99_t rng;
    uint32_t dst_seq[REGS], src_seq[REGS];
    uint32_t dst_n, src_n;
} prog_rng;

/* progpow.cpp:63-88 */
static void prog_rng_init(prog_rng* p, uint64_t prog_seed)
{
    const uint32_t lo = (uint32_t)prog_seed, hi = (uint32_t)(prog_seed >> 32);
    p->rng.z = fnv1a(FNV_BASIS, lo);
    p->rng.w = fnv1a(p->rng.z, hi);
    p->rng.jsr = fnv1a(p->rng.w, lo);
    p->rng.jcong = fnv1a(p->rng.jsr, hi);
    for (uint32_t i = 0; i < REGS; ++i) p->dst_seq[i] = p->src_seq[i] = i;
    for (uint32_t i = REGS; i > 1; --i) {
        uint32_t j = kiss99(&p->rng) % i, t = p->dst_seq[i - 1];
        p->dst_seq[i - 1] = p->dst_seq[j];
        p->dst_seq[j] = t;
        j = kiss99(&p->rng) % i;
        t = p->src_seq[i - 1];
        p->src_seq[i - 1] = p->src_seq[j];
        p->src_seq[j] = t;
    }
    p->dst_n = p->src_n = 0;
}
static inline uint32_t next_dst(prog_rng* p) { return p->dst_seq[(p->dst_n++) % REGS]; } /* progpow.cpp:51 */
static inline uint32_t next_src(prog_rng* p) { return p->src_seq[(p->src_n++) % REGS]; } /* progpow.cpp:52 */

static inline uint32_t clz32(uint32_t x) { return x ? (uint32_t)__builtin_clz(x) : 32; } /* bit_manipulation.h:35-38 */

/* progpow.cpp:92-120 */
static uint32_t random_math(uint32_t a, uint32_t b, uint32_t sel)
{
    switch (sel % 11) {
    default:
    case 0: return a + b;
    case 1: return a * b;
    case 2: return (uint32_t)(((uint64_t)a * b) >> 32);
    case 3: return a < b ? a : b;
    case 4: return rol32(a, b);
    case 5: return ror32(a, b);
    case 6: return a & b;
    case 7: return a | b;
    case 8: return a ^ b;
    case 9: return clz32(a) + clz32(b);
    case 10: return (uint32_t)__builtin_popcount(a) + (uint32_t)__builtin_popcount(b);
    }
}

/* progpow.cpp:126-144 */
static uint32_t random_merge(uint32_t a, uint32_t b, uint32_t sel)
{
    const uint32_t x = ((sel >> 16) % 31) + 1;
    switch (sel % 4) {
    case 0: return a * 33 + b;
    case 1: return (a ^ b) * 33;
    case 2: return rol32(a, x) ^ b;
    default: return ror32(a, x) ^ b;
    }
}

static void dag_lookup(const nxo_ctx* c, uint32_t index, uint32_t item[64])
{
    if (c->dag)
        memcpy(item, c->dag + 256 * (size_t)index, 256);
    else
        nxo_dataset_item_2048(c, index, (uint8_t*)item);
}

/* progpow.cpp:179-245; `st` arrives by value so every round replays the same program */
static void kawpow_round(const nxo_ctx* c, uint32_t r, uint32_t mix[LANES][REGS], prog_rng st)
{
    const uint32_t num_items = (uint32_t)(c->dataset_items / 2);
    uint32_t item[64];
    dag_lookup(c, mix[r % LANES][0] % num_items, item);

    for (int i = 0; i < MATH_OPS; ++i) {
        if (i < CACHE_OPS) {
            const uint32_t src = next_src(&st), dst = next_dst(&st), sel = kiss99(&st.rng);
            for (int l = 0; l < LANES; ++l)
                mix[l][dst] = random_merge(mix[l][dst], c->l1[mix[l][src] % 4096], sel);
        }
        {
            const uint32_t rnd = kiss99(&st.rng) % (REGS * (REGS - 1));
            const uint32_t src1 = rnd % REGS;
            uint32_t src2 = rnd / REGS;
            if (src2 >= src1) ++src2;
            const uint32_t sel1 = kiss99(&st.rng), dst = next_dst(&st), sel2 = kiss99(&st.rng);
            for (int l = 0; l < LANES; ++l)
                mix[l][dst] = random_merge(mix[l][dst], random_math(mix[l][src1], mix[l][src2], sel1), sel2);
        }
    }
    uint32_t dsts[DAG_WORDS], sels[DAG_WORDS];
    for (int i = 0; i < DAG_WORDS; ++i) {
        dsts[i] = i == 0 ? 0 : next_dst(&st);
        sels[i] = kiss99(&st.rng);
    }
    for (uint32_t l = 0; l < LANES; ++l) {
        const uint32_t off = ((l ^ r) % LANES) * DAG_WORDS;
        for (int i = 0; i < DAG_WORDS; ++i) mix[l][dsts[i]] = random_merge(mix[l][dsts[i]], item[off + i], sels[i]);
    }
}

/* Flat op list in the reference's draw order (progpow.cpp:191-233), for the product's code generator tests. */
int nxo_kawpow_program(uint64_t prog_seed, uint32_t ops[33][6])
{
    prog_rng st;
    int n = 0;
    prog_rng_init(&st, prog_seed);
    for (int i = 0; i < MATH_OPS; ++i) {
        if (i < CACHE_OPS) {
            ops[n][0] = 0;
            ops[n][1] = next_src(&st);
            ops[n][2] = next_dst(&st);
            ops[n][3] = kiss99(&st.rng);
            ops[n][4] = ops[n][5] = 0;
            ++n;
        }
        const uint32_t rnd = kiss99(&st.rng) % (REGS * (REGS - 1));
        const uint32_t src1 = rnd % REGS;
        uint32_t src2 = rnd / REGS;
        if (src2 >= src1) ++src2;
        ops[n][0] = 1;
        ops[n][1] = src1;
        ops[n][2] = src2;
        ops[n][3] = kiss99(&st.rng);
        ops[n][4] = next_dst(&st);
        ops[n][5] = kiss99(&st.rng);
        ++n;
    }
    for (int i = 0; i < DAG_WORDS; ++i) {
        ops[n][0] = 2;
        ops[n][1] = i == 0 ? 0 : next_dst(&st);
        ops[n][2] = kiss99(&st.rng);
        ops[n][3] = ops[n][4] = ops[n][5] = 0;
        ++n;
    }
    return n;
}

/* progpow.cpp:247-295 */
static void kawpow_mix(const nxo_ctx* c, int height, const uint32_t seed[2], uint32_t mix_hash[8])
{
    uint32_t mix[LANES][REGS];
    const uint32_t z = fnv1a(FNV_BASIS, seed[0]), w = fnv1a(z, seed[1]);
    for (uint32_t l = 0; l < LANES; ++l) {
        kiss99_t g;
        g.z = z;
        g.w = w;
        g.jsr = fnv1a(w, l);
        g.jcong = fnv1a(g.jsr, l);
        for (int i = 0; i < REGS; ++i) mix[l][i] = kiss99(&g);
    }
    prog_rng st;
    prog_rng_init(&st, (uint64_t)(height / 3));
    for (uint32_t r = 0; r < 64; ++r) kawpow_round(c, r, mix, st);

    uint32_t lane_hash[LANES];
    for (int l = 0; l < LANES; ++l) {
        lane_hash[l] = FNV_BASIS;
        for (int i = 0; i < REGS; ++i) lane_hash[l] = fnv1a(lane_hash[l], mix[l][i]);
    }
    for (int i = 0; i < 8; ++i) mix_hash[i] = FNV_BASIS;
    for (int l = 0; l < LANES; ++l) mix_hash[l % 8] = fnv1a(mix_hash[l % 8], lane_hash[l]);
}

/* progpow.cpp:305-325: header(8) | nonce lo,hi | 15 pad words -> f800 -> first 8 words carried */
static void kawpow_absorb_first(const uint8_t header[32], uint64_t nonce, uint32_t carry[8])
{
    uint32_t st[25];
    memcpy(st, header, 32);
    st[8] = (uint32_t)nonce;
    st[9] = (uint32_t)(nonce >> 32);
    for (int i = 10; i < 25; ++i) st[i] = KAWPOW_PAD[i - 10];
    nxo_keccakf800(st);
    memcpy(carry, st, 32);
}

/* progpow.cpp:333-352: carry(8) | mix(8) | 9 pad words -> f800 -> first 8 words = final */
static void kawpow_absorb_last(const uint32_t carry[8], const uint32_t mix[8], uint8_t final_out[32])
{
    uint32_t st[25];
    memcpy(st, carry, 32);
    memcpy(st + 8, mix, 32);
    for (int i = 16; i < 25; ++i) st[i] = KAWPOW_PAD[i - 16];
    nxo_keccakf800(st);
    memcpy(final_out, st, 32);
}

/* progpow.cpp:298-355 (light) and :357-429 (full; here when c->dag is set) */
void nxo_kawpow_hash(const nxo_ctx* c, int height, const uint8_t header[32], uint64_t nonce, uint8_t final_out[32],
                     uint8_t mix_out[32])
{
    uint32_t carry[8], mix[8];
    kawpow_absorb_first(header, nonce, carry);
    kawpow_mix(c, height, carry, mix);
    kawpow_absorb_last(carry, mix, final_out);
    memcpy(mix_out, mix, 32);
}

/* lib/ethash/ethash-internal.hpp:32-43: compare as big-endian byte strings */
int nxo_less_or_equal(const uint8_t a[32], const uint8_t b[32]) { return memcmp(a, b, 32) <= 0; }

/* progpow.cpp:431-495 */
int nxo_kawpow_verify(const nxo_ctx* c, int height, const uint8_t header[32], const uint8_t mix[32], uint64_t nonce,
                      const uint8_t boundary[32])
{
    uint32_t carry[8], claimed[8], expect[8];
    uint8_t fin[32];
    kawpow_absorb_first(header, nonce, carry);
    memcpy(claimed, mix, 32);
    kawpow_absorb_last(carry, claimed, fin);
    if (!nxo_less_or_equal(fin, boundary)) return 0;
    kawpow_mix(c, height, carry, expect);
    return memcmp(expect, mix, 32) == 0;
}

/* progpow.cpp:498-550 (height is unused there too) */
void nxo_kawpow_hash_no_verify(int height, const uint8_t header[32], const uint8_t mix[32], uint64_t nonce,
                               uint8_t out[32])
{
    uint32_t carry[8], m[8];
    (void)height;
    kawpow_absorb_first(header, nonce, carry);
    memcpy(m, mix, 32);
    kawpow_absorb_last(carry, m, out);
}

/* progpow.cpp:553-579: ascending scan, first hit wins, all-zero result when none */
int nxo_kawpow_search(const nxo_ctx* c, int height, const uint8_t header[32], const uint8_t boundary[32],
                      uint64_t start, uint64_t iterations, uint64_t* nonce, uint8_t final_out[32], uint8_t mix_out[32])
{
    for (uint64_t n = start; n < start + iterations; ++n) {
        nxo_kawpow_hash(c, height, header, n, final_out, mix_out);
        if (nxo_less_or_equal(final_out, boundary)) {
            *nonce = n;
            return 1;
        }
    }
    *nonce = 0;
    memset(final_out, 0, 32);
    memset(mix_out, 0, 32);
    return 0;
}

typedef struct
{
    const nxo_ctx* c;
    int height;
    const uint8_t *header, *boundary;
    uint64_t start, count;
    uint8_t* out;
    volatile uint64_t* next;
    volatile uint64_t* best; /* search: lowest hit offset so far */
} range_job;

static void* range_worker(void* p)
{
    range_job* j = (range_job*)p;
    for (;;) {
        const uint64_t i = __atomic_fetch_add(j->next, 1, __ATOMIC_RELAXED);
        if (i >= j->count) return NULL;
        if (j->best) {
            uint8_t f[32], m[32];
            if (i > __atomic_load_n(j->best, __ATOMIC_RELAXED)) return NULL;
            nxo_kawpow_hash(j->c, j->height, j->header, j->start + i, f, m);
            if (nxo_less_or_equal(f, j->boundary)) {
                uint64_t cur = __atomic_load_n(j->best, __ATOMIC_RELAXED);
                while (i < cur && !__atomic_compare_exchange_n(j->best, &cur, i, 0, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {
                }
            }
        } else {
            nxo_kawpow_hash(j->c, j->height, j->header, j->start + i, j->out + 64 * i, j->out + 64 * i + 32);
        }
    }
}

void nxo_kawpow_hash_range(const nxo_ctx* c, int height, const uint8_t header[32], uint64_t start, uint64_t count,
                           uint8_t* out, int threads)
{
    volatile uint64_t next = 0;
    range_job j = {c, height, header, NULL, start, count, out, &next, NULL};
    run_threads(range_worker, &j, threads);
}

int nxo_kawpow_search_mt(const nxo_ctx* c, int height, const uint8_t header[32], const uint8_t boundary[32],
                         uint64_t start, uint64_t iterations, uint64_t* nonce, uint8_t final_out[32],
                         uint8_t mix_out[32], int threads)
{
    volatile uint64_t next = 0, best = UINT64_MAX;
    range_job j = {c, height, header, boundary, start, iterations, NULL, &next, &best};
    run_threads(range_worker, &j, threads);
    if (best == UINT64_MAX) {
        *nonce = 0;
        memset(final_out, 0, 32);
        memset(mix_out, 0, 32);
        return 0;
    }
    *nonce = start + best;
    nxo_kawpow_hash(c, height, header, *nonce, final_out, mix_out);
    return 1;
}

/* ------------------------------------------------------------------ classic Ethash */

/* lib/ethash/ethash.cpp:221-233 (1024-bit item = two 512-bit items) + :257-303 + :416-422 */
void nxo_ethash_hash(const nxo_ctx* c, const uint8_t header[32], uint64_t nonce, uint8_t final_out[32],
                     uint8_t mix_out[32])
{
    uint8_t init[40], seed[64], fin[96];
    uint32_t s[16], mix[32], mh[8];
    memcpy(init, header, 32);
    memcpy(init + 32, &nonce, 8);
    nxo_keccak512(init, 40, seed);
    memcpy(s, seed, 64);
    memcpy(mix, s, 64);
    memcpy(mix + 16, s, 64);
    const uint32_t limit = (uint32_t)c->dataset_items;
    for (uint32_t i = 0; i < 64; ++i) {
        const uint32_t p = fnv1(i ^ s[0], mix[i % 32]) % limit;
        uint32_t item[32];
        nxo_dataset_item_512(c, (int64_t)p * 2, (uint8_t*)item);
        nxo_dataset_item_512(c, (int64_t)p * 2 + 1, (uint8_t*)(item + 16));
        for (int k = 0; k < 32; ++k) mix[k] = fnv1(mix[k], item[k]);
    }
    for (int i = 0; i < 32; i += 4) mh[i / 4] = fnv1(fnv1(fnv1(mix[i], mix[i + 1]), mix[i + 2]), mix[i + 3]);
    memcpy(mix_out, mh, 32);
    memcpy(fin, seed, 64);
    memcpy(fin + 64, mh, 32);
    nxo_keccak256(fin, 96, final_out);
}

/* lib/ethash/ethash.cpp:424-429 */
int nxo_ethash_verify_final(const uint8_t header[32], const uint8_t mix[32], uint64_t nonce, const uint8_t boundary[32])
{
    uint8_t init[40], fin[96], out[32];
    memcpy(init, header, 32);
    memcpy(init + 32, &nonce, 8);
    nxo_keccak512(init, 40, fin);
    memcpy(fin + 64, mix, 32);
    nxo_keccak256(fin, 96, out);
    return nxo_less_or_equal(out, boundary);
}

/* ------------------------------------------------------------------ node glue */

static const uint32_t SHA_K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01,
    0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc,
    0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147,
    0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08,
    0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
    0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

/* FIPS 180-4 SHA-256 (the node's CSHA256, src/crypto/sha256.cpp, is the same function) */
void nxo_sha256(const uint8_t* data, size_t n, uint8_t out[32])
{
    uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    const size_t total = ((n + 9 + 63) / 64) * 64;
    uint8_t* buf = (uint8_t*)calloc(1, total);
    memcpy(buf, data, n);
    buf[n] = 0x80;
    for (int i = 0; i < 8; ++i) buf[total - 1 - i] = (uint8_t)(((uint64_t)n * 8) >> (8 * i));
    for (size_t off = 0; off < total; off += 64) {
        uint32_t w[64], a[8];
        for (int i = 0; i < 16; ++i)
            w[i] = (uint32_t)buf[off + 4 * i] << 24 | (uint32_t)buf[off + 4 * i + 1] << 16 |
                   (uint32_t)buf[off + 4 * i + 2] << 8 | buf[off + 4 * i + 3];
        for (int i = 16; i < 64; ++i) {
            const uint32_t s0 = ror32(w[i - 15], 7) ^ ror32(w[i - 15], 18) ^ (w[i - 15] >> 3);
            const uint32_t s1 = ror32(w[i - 2], 17) ^ ror32(w[i - 2], 19) ^ (w[i - 2] >> 10);
            w[i] = w[i - 16] + s0 + w[i - 7] + s1;
        }
        memcpy(a, h, sizeof a);
        for (int i = 0; i < 64; ++i) {
            const uint32_t S1 = ror32(a[4], 6) ^ ror32(a[4], 11) ^ ror32(a[4], 25);
            const uint32_t ch = (a[4] & a[5]) ^ (~a[4] & a[6]);
            const uint32_t t1 = a[7] + S1 + ch + SHA_K[i] + w[i];
            const uint32_t S0 = ror32(a[0], 2) ^ ror32(a[0], 13) ^ ror32(a[0], 22);
            const uint32_t mj = (a[0] & a[1]) ^ (a[0] & a[2]) ^ (a[1] & a[2]);
            const uint32_t t2 = S0 + mj;
            a[7] = a[6];
            a[6] = a[5];
            a[5] = a[4];
            a[4] = a[3] + t1;
            a[3] = a[2];
            a[2] = a[1];
            a[1] = a[0];
            a[0] = t1 + t2;
        }
        for (int i = 0; i < 8; ++i) h[i] += a[i];
    }
    free(buf);
    for (int i = 0; i < 8; ++i) {
        out[4 * i] = (uint8_t)(h[i] >> 24);
        out[4 * i + 1] = (uint8_t)(h[i] >> 16);
        out[4 * i + 2] = (uint8_t)(h[i] >> 8);
        out[4 * i + 3] = (uint8_t)h[i];
    }
}

/* src/primitives/block.h:213-233 (CKAWPOWInput: version,prev,merkle,time,bits,height, all LE) ->
 * src/primitives/block.cpp:94-99 (SerializeHash = SHA256d) -> src/hash.cpp:269-270
 * (to_hash256(GetHex()): GetHex prints the uint256 bytes reversed, src/uint256.cpp:24-27). */
void nxo_kawpow_header_hash(int32_t version, const uint8_t prev[32], const uint8_t merkle[32], uint32_t time,
                            uint32_t bits, uint32_t height, uint8_t out[32])
{
    uint8_t in[80], d1[32], d2[32];
    memcpy(in, &version, 4);
    memcpy(in + 4, prev, 32);
    memcpy(in + 36, merkle, 32);
    memcpy(in + 68, &time, 4);
    memcpy(in + 72, &bits, 4);
    memcpy(in + 76, &height, 4);
    nxo_sha256(in, 80, d1);
    nxo_sha256(d1, 32, d2);
    for (int i = 0; i < 32; ++i) out[i] = d2[31 - i];
}

/* src/arith_uint256.cpp:209-228 SetCompact + the range checks of src/pow.cpp:182-192 that do not need powLimit.
 * Output is the target as 32 big-endian bytes (what is_less_or_equal compares against after the node's hex round trip). */
int nxo_compact_to_boundary(uint32_t bits, uint8_t out[32])
{
    const int size = (int)(bits >> 24);
    uint32_t word = bits & 0x007fffff;
    uint8_t le[64];
    memset(le, 0, sizeof le);
    memset(out, 0, 32);
    if (size <= 3) {
        word >>= 8 * (3 - size);
        le[0] = (uint8_t)word;
        le[1] = (uint8_t)(word >> 8);
        le[2] = (uint8_t)(word >> 16);
    } else if (size - 3 < 32) {
        le[size - 3] = (uint8_t)word;
        le[size - 2] = (uint8_t)(word >> 8);
        le[size - 1] = (uint8_t)(word >> 16);
    }
    const int negative = word != 0 && (bits & 0x00800000) != 0;
    const int overflow = word != 0 && ((size > 34) || (word > 0xff && size > 33) || (word > 0xffff && size > 32));
    for (int i = 0; i < 32; ++i) out[i] = le[31 - i];
    if (negative || overflow) return 0;
    for (int i = 0; i < 32; ++i)
        if (out[i]) return 1;
    return 0;
}
