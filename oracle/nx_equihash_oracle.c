/* TEST INFRASTRUCTURE ONLY — SELF-CONSISTENCY oracle for the gated Equihash engine.
 *
 * *** NOT derived from DeonDavisV/Nodexa-Chain-Core: the reference contains no Equihash code, no (n,k), no
 * *** personalization string, no validator, no test vector (SURVEY.md §0 Fact 1, §8 a28).  Parity is UNPINNED.
 *
 * What this restates instead is the PUBLISHED algorithm (Biryukov & Khovratovich, "Equihash", NDSS 2016; the
 * parameterisation popularised by Zcash): BLAKE2b with a 16-byte personalization (8-byte tag || LE32(n) ||
 * LE32(k)), 512/n strings per digest, Wagner's algorithm with k collision rounds of n/(k+1) bits, the ordering
 * rule "left subtree's first index < right subtree's first index", distinct indices, minimal big-endian
 * (n/(k+1)+1)-bit index encoding.  It exists so that the GPU solver can be checked against an independent,
 * exhaustive CPU implementation of the same rules; (n,k) and the tag are inputs, nothing is assumed about the chain.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ BLAKE2b (RFC 7693) */
static const uint64_t B2IV[8] = {0x6a09e667f3bcc908ULL, 0xbb67ae8584caa73bULL, 0x3c6ef372fe94f82bULL, 0xa54ff53a5f1d36f1ULL,
                                 0x510e527fade682d1ULL, 0x9b05688c2b3e6c1fULL, 0x1f83d9abfb41bd6bULL, 0x5be0cd19137e2179ULL};
static const uint8_t B2S[12][16] = {{0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3},
                                    {11, 8, 12, 0, 5, 2, 15, 13, 10, 14, 3, 6, 7, 1, 9, 4}, {7, 9, 3, 1, 13, 12, 11, 14, 2, 6, 5, 10, 4, 0, 15, 8},
                                    {9, 0, 5, 7, 2, 4, 10, 15, 14, 1, 11, 12, 6, 8, 3, 13}, {2, 12, 6, 10, 0, 11, 8, 3, 4, 13, 7, 5, 15, 14, 1, 9},
                                    {12, 5, 1, 15, 14, 13, 4, 10, 0, 7, 6, 3, 9, 2, 8, 11}, {13, 11, 7, 14, 12, 1, 3, 9, 5, 0, 15, 4, 8, 6, 2, 10},
                                    {6, 15, 14, 9, 11, 3, 0, 8, 12, 2, 13, 7, 1, 4, 10, 5}, {10, 2, 8, 4, 7, 6, 1, 5, 15, 11, 9, 14, 3, 12, 13, 0},
                                    {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15}, {14, 10, 4, 8, 9, 15, 13, 6, 1, 12, 0, 2, 11, 7, 5, 3}};
static inline uint64_t rotr64(uint64_t x, int n) { return (x >> n) | (x << (64 - n)); }

static void b2_compress(uint64_t h[8], const uint8_t block[128], uint64_t t, int last)
{
    uint64_t m[16], v[16];
    memcpy(m, block, 128);
    for (int i = 0; i < 8; ++i) {
        v[i] = h[i];
        v[i + 8] = B2IV[i];
    }
    v[12] ^= t;
    if (last) v[14] = ~v[14];
#define G(a, b, c, d, x, y)                 \
    v[a] += v[b] + (x);                     \
    v[d] = rotr64(v[d] ^ v[a], 32);         \
    v[c] += v[d];                           \
    v[b] = rotr64(v[b] ^ v[c], 24);         \
    v[a] += v[b] + (y);                     \
    v[d] = rotr64(v[d] ^ v[a], 16);         \
    v[c] += v[d];                           \
    v[b] = rotr64(v[b] ^ v[c], 63)
    for (int r = 0; r < 12; ++r) {
        const uint8_t* s = B2S[r];
        G(0, 4, 8, 12, m[s[0]], m[s[1]]);
        G(1, 5, 9, 13, m[s[2]], m[s[3]]);
        G(2, 6, 10, 14, m[s[4]], m[s[5]]);
        G(3, 7, 11, 15, m[s[6]], m[s[7]]);
        G(0, 5, 10, 15, m[s[8]], m[s[9]]);
        G(1, 6, 11, 12, m[s[10]], m[s[11]]);
        G(2, 7, 8, 13, m[s[12]], m[s[13]]);
        G(3, 4, 9, 14, m[s[14]], m[s[15]]);
    }
#undef G
    for (int i = 0; i < 8; ++i) h[i] ^= v[i] ^ v[i + 8];
}

/* one-shot: out = BLAKE2b-outlen(personal; data) */
void nxeq_blake2b(const uint8_t* data, size_t len, const uint8_t personal[16], uint8_t* out, size_t outlen)
{
    uint64_t h[8], p[2];
    memcpy(h, B2IV, sizeof h);
    h[0] ^= 0x01010000ULL ^ (uint64_t)outlen;
    memcpy(p, personal, 16);
    h[6] ^= p[0];
    h[7] ^= p[1];
    uint64_t t = 0;
    uint8_t block[128];
    while (len > 128) {
        t += 128;
        b2_compress(h, data, t, 0);
        data += 128;
        len -= 128;
    }
    memset(block, 0, 128);
    memcpy(block, data, len);
    t += len;
    b2_compress(h, block, t, 1);
    memcpy(out, h, outlen);
}

/* ------------------------------------------------------------------ parameters */
typedef struct
{
    int n, k, cbl, chunks, indices_per_hash, hash_bytes, digest_bytes;
    uint32_t N; /* number of initial strings = 2^(cbl+1) */
    uint8_t personal[16];
} eq_params;

static int eq_init(eq_params* p, int n, int k, const uint8_t tag[8])
{
    if (n <= 0 || k <= 0 || n % 8 || n % (k + 1) || n / (k + 1) > 24 || n / (k + 1) < 4 || n > 512) return 0;
    p->n = n;
    p->k = k;
    p->cbl = n / (k + 1);
    p->chunks = k + 1;
    p->indices_per_hash = 512 / n;
    p->hash_bytes = n / 8;
    p->digest_bytes = p->indices_per_hash * p->hash_bytes;
    p->N = 1u << (p->cbl + 1);
    memcpy(p->personal, tag, 8);
    const uint32_t le_n = (uint32_t)n, le_k = (uint32_t)k;
    memcpy(p->personal + 8, &le_n, 4);
    memcpy(p->personal + 12, &le_k, 4);
    return 1;
}

/* chunk c (0..k) of the n-bit string, big-endian bit order: bits [c*cbl, (c+1)*cbl) */
static uint32_t chunk_of(const uint8_t* str, int cbl, int c)
{
    uint32_t v = 0;
    for (int b = 0; b < cbl; ++b) {
        const int bit = c * cbl + b;
        v = (v << 1) | ((str[bit >> 3] >> (7 - (bit & 7))) & 1u);
    }
    return v;
}

/* all k+1 chunks of initial string i */
static void initial_chunks(const eq_params* p, const uint8_t* header, size_t header_len, uint32_t i, uint32_t* out)
{
    uint8_t* buf = (uint8_t*)malloc(header_len + 4);
    uint8_t digest[64];
    const uint32_t g = i / (uint32_t)p->indices_per_hash;
    memcpy(buf, header, header_len);
    memcpy(buf + header_len, &g, 4);
    nxeq_blake2b(buf, header_len + 4, p->personal, digest, (size_t)p->digest_bytes);
    free(buf);
    const uint8_t* str = digest + (i % (uint32_t)p->indices_per_hash) * p->hash_bytes;
    for (int c = 0; c <= p->k; ++c) out[c] = chunk_of(str, p->cbl, c);
}

/* ------------------------------------------------------------------ validator */
/* indices: 2^k values.  Returns 1 when they form a valid solution under the rules in the header comment. */
int nxeq_validate_indices(int n, int k, const uint8_t tag[8], const uint8_t* header, size_t header_len,
                          const uint32_t* indices)
{
    eq_params p;
    if (!eq_init(&p, n, k, tag)) return 0;
    const uint32_t count = 1u << k;
    const int C = p.chunks;
    uint32_t* rows = (uint32_t*)malloc((size_t)count * C * 4);
    int ok = 1;
    for (uint32_t i = 0; i < count && ok; ++i) {
        if (indices[i] >= p.N) ok = 0;
        else initial_chunks(&p, header, header_len, indices[i], rows + (size_t)i * C);
    }
    /* distinct indices */
    for (uint32_t i = 0; i < count && ok; ++i)
        for (uint32_t j = i + 1; j < count; ++j)
            if (indices[i] == indices[j]) {
                ok = 0;
                break;
            }
    /* level by level: combine neighbours; row i of level L covers indices [i*2^L, (i+1)*2^L) */
    for (int level = 0; level < k && ok; ++level) {
        const uint32_t width = 1u << level, pairs = count >> (level + 1);
        for (uint32_t q = 0; q < pairs; ++q) {
            uint32_t* a = rows + (size_t)(2 * q) * width * C;       /* row of the left subtree lives at its first leaf */
            uint32_t* b = rows + (size_t)(2 * q + 1) * width * C;
            if (a[level] != b[level]) ok = 0;                        /* collision on this round's chunk */
            if (indices[(2 * q) * width] >= indices[(2 * q + 1) * width]) ok = 0; /* ordering rule */
            for (int c = 0; c < C; ++c) a[c] ^= b[c];
        }
    }
    if (ok) {
        for (int c = 0; c < C; ++c)
            if (rows[c]) ok = 0; /* the last round cancels the final two chunks at once */
    }
    free(rows);
    return ok;
}

/* minimal encoding: 2^k indices, (cbl+1) bits each, most significant bit first */
size_t nxeq_solution_bytes(int n, int k) { return ((size_t)1 << k) * (size_t)(n / (k + 1) + 1) / 8; }

void nxeq_encode(int n, int k, const uint32_t* indices, uint8_t* out)
{
    const int bits = n / (k + 1) + 1;
    const size_t total = nxeq_solution_bytes(n, k);
    memset(out, 0, total);
    size_t pos = 0;
    for (uint32_t i = 0; i < (1u << k); ++i)
        for (int b = bits - 1; b >= 0; --b, ++pos)
            if ((indices[i] >> b) & 1u) out[pos >> 3] |= (uint8_t)(0x80u >> (pos & 7));
}

void nxeq_decode(int n, int k, const uint8_t* in, uint32_t* indices)
{
    const int bits = n / (k + 1) + 1;
    size_t pos = 0;
    for (uint32_t i = 0; i < (1u << k); ++i) {
        uint32_t v = 0;
        for (int b = 0; b < bits; ++b, ++pos) v = (v << 1) | ((in[pos >> 3] >> (7 - (pos & 7))) & 1u);
        indices[i] = v;
    }
}

int nxeq_validate(int n, int k, const uint8_t tag[8], const uint8_t* header, size_t header_len, const uint8_t* solution)
{
    uint32_t* idx = (uint32_t*)malloc(((size_t)1 << k) * 4);
    nxeq_decode(n, k, solution, idx);
    const int ok = nxeq_validate_indices(n, k, tag, header, header_len, idx);
    free(idx);
    return ok;
}

/* ------------------------------------------------------------------ exhaustive Wagner solver */
typedef struct
{
    uint32_t a, b; /* slots of the previous round (round 0: a = initial index) */
} eq_pair;

static int C_sort;
static int chunk_sort_c;
static const uint32_t* chunk_sort_rows;
static int cmp_rows(const void* x, const void* y)
{
    const uint32_t i = *(const uint32_t*)x, j = *(const uint32_t*)y;
    const uint32_t a = chunk_sort_rows[(size_t)i * C_sort + chunk_sort_c], b = chunk_sort_rows[(size_t)j * C_sort + chunk_sort_c];
    if (a != b) return a < b ? -1 : 1;
    return i < j ? -1 : (i > j);
}

static void expand(eq_pair* const* tree, int round, uint32_t slot, uint32_t* out) /* writes 2^round leaves */
{
    if (round == 0) {
        out[0] = tree[0][slot].a;
        return;
    }
    const uint32_t half = 1u << (round - 1);
    expand(tree, round - 1, tree[round][slot].a, out);
    expand(tree, round - 1, tree[round][slot].b, out + half);
    if (out[0] > out[half]) { /* ordering rule: the subtree with the smaller first index goes left */
        for (uint32_t i = 0; i < half; ++i) {
            const uint32_t t = out[i];
            out[i] = out[half + i];
            out[half + i] = t;
        }
    }
}

static int cmp_u32(const void* a, const void* b)
{
    const uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
    return x < y ? -1 : (x > y);
}

/* Finds every solution for this header; writes up to max_solutions encoded solutions, sorted bytewise, without
 * duplicates; returns the number found (which may exceed max_solutions), or -1 on bad parameters / memory. */
int nxeq_solve(int n, int k, const uint8_t tag[8], const uint8_t* header, size_t header_len, uint8_t* solutions,
               int max_solutions)
{
    eq_params p;
    if (!eq_init(&p, n, k, tag)) return -1;
    const int C = p.chunks;
    C_sort = C;
    const size_t cap = (size_t)p.N * 4; /* rows per round hover around N (count' ~ count^2/N); 4x is generous */
    uint32_t* rows = (uint32_t*)malloc(cap * C * 4);
    uint32_t* next = (uint32_t*)malloc(cap * C * 4);
    uint32_t* order = (uint32_t*)malloc(cap * 4);
    eq_pair** tree = (eq_pair**)calloc((size_t)k + 1, sizeof(eq_pair*));
    if (!rows || !next || !order || !tree) return -1;
    size_t count = p.N;
    tree[0] = (eq_pair*)malloc(count * sizeof(eq_pair));
    for (uint32_t i = 0; i < p.N; ++i) {
        initial_chunks(&p, header, header_len, i, rows + (size_t)i * C);
        tree[0][i].a = i;
        tree[0][i].b = 0;
    }
    const size_t sol_bytes = nxeq_solution_bytes(n, k);
    const uint32_t leaves = 1u << k;
    size_t found_cap = 64, found = 0;
    uint8_t* all = (uint8_t*)malloc(found_cap * sol_bytes);
    uint32_t* idx = (uint32_t*)malloc((size_t)leaves * 4 * 2);

    for (int r = 0; r < k; ++r) {
        const int last = r == k - 1;
        for (size_t i = 0; i < count; ++i) order[i] = (uint32_t)i;
        chunk_sort_c = r;
        chunk_sort_rows = rows;
        qsort(order, count, 4, cmp_rows);
        tree[r + 1] = (eq_pair*)malloc(cap * sizeof(eq_pair));
        size_t out = 0;
        for (size_t g = 0; g < count;) {
            size_t e = g + 1;
            while (e < count && rows[(size_t)order[e] * C + r] == rows[(size_t)order[g] * C + r]) ++e;
            for (size_t i = g; i < e; ++i)
                for (size_t j = i + 1; j < e; ++j) {
                    const uint32_t* a = rows + (size_t)order[i] * C;
                    const uint32_t* b = rows + (size_t)order[j] * C;
                    if (last) {
                        if (a[k] != b[k]) continue; /* final round: both remaining chunks must cancel */
                        if (out < cap) {
                            tree[r + 1][out].a = order[i];
                            tree[r + 1][out].b = order[j];
                            ++out;
                        }
                    } else {
                        uint32_t any = 0;
                        for (int c = r + 1; c < C; ++c) any |= a[c] ^ b[c];
                        if (!any) continue; /* identical remainders: the two subtrees share their index set */
                        if (out < cap) {
                            for (int c = 0; c < C; ++c) next[out * C + c] = c <= r ? 0 : a[c] ^ b[c];
                            tree[r + 1][out].a = order[i];
                            tree[r + 1][out].b = order[j];
                            ++out;
                        }
                    }
                }
            g = e;
        }
        if (last) {
            for (size_t s = 0; s < out; ++s) {
                expand(tree, k, (uint32_t)s, idx);
                memcpy(idx + leaves, idx, (size_t)leaves * 4);
                qsort(idx + leaves, leaves, 4, cmp_u32);
                int distinct = 1;
                for (uint32_t i = 1; i < leaves; ++i)
                    if (idx[leaves + i] == idx[leaves + i - 1]) distinct = 0;
                if (!distinct) continue;
                if (found == found_cap) {
                    found_cap *= 2;
                    all = (uint8_t*)realloc(all, found_cap * sol_bytes);
                }
                nxeq_encode(n, k, idx, all + found * sol_bytes);
                ++found;
            }
        } else {
            uint32_t* t = rows;
            rows = next;
            next = t;
            count = out;
        }
    }
    /* sort + dedupe (the same index set can be reached through more than one pairing order) */
    size_t uniq = 0;
    if (found) {
        /* insertion sort: the number of solutions per header is tiny */
        for (size_t i = 1; i < found; ++i)
            for (size_t j = i; j > 0 && memcmp(all + (j - 1) * sol_bytes, all + j * sol_bytes, sol_bytes) > 0; --j) {
                uint8_t* tmp = (uint8_t*)malloc(sol_bytes);
                memcpy(tmp, all + j * sol_bytes, sol_bytes);
                memcpy(all + j * sol_bytes, all + (j - 1) * sol_bytes, sol_bytes);
                memcpy(all + (j - 1) * sol_bytes, tmp, sol_bytes);
                free(tmp);
            }
        for (size_t i = 0; i < found; ++i) {
            if (i && memcmp(all + i * sol_bytes, all + (i - 1) * sol_bytes, sol_bytes) == 0) continue;
            if ((int)uniq < max_solutions) memcpy(solutions + uniq * sol_bytes, all + i * sol_bytes, sol_bytes);
            ++uniq;
        }
    }
    for (int r = 0; r <= k; ++r) free(tree[r]);
    free(tree);
    free(rows);
    free(next);
    free(order);
    free(all);
    free(idx);
    return (int)uniq;
}
