// GATED "next" row: a parameterised Equihash engine for sm_100a.
//
// *** The reference node contains NO Equihash: no solver, no validator, no (n,k), no personalization, no header
// *** field, no test vector (SURVEY.md §0 Fact 1).  Nothing here can be checked against it; (n,k) and the 8-byte
// *** personalization tag are therefore INPUTS, and correctness is established against an independent CPU
// *** implementation of the published algorithm (oracle/nx_equihash_oracle.c) - self-consistency, parity unpinned.
//
// Algorithm (Biryukov-Khovratovich Equihash, Wagner's generalised birthday): N = 2^(n/(k+1)+1) strings of n bits
// from BLAKE2b(personal = tag||LE32(n)||LE32(k); header||LE32(i / (512/n))), k collision rounds of cbl = n/(k+1)
// bits, the last one cancelling two chunks at once; a solution is 2^k distinct indices whose tree satisfies
// "left subtree's first index < right subtree's first index".
//
// GPU mapping: a round is (1) a scatter of the live rows into buckets by the top cbl-8 bits of the round's chunk
// (one atomic slot counter per bucket, ~512 rows per bucket whatever (n,k) is, because N / 2^(cbl-8) = 512), then
// (2) one CTA per bucket that counting-sorts its rows by the remaining 8 bits in shared memory and emits the XOR
// of every colliding pair together with a two-slot back-reference.  Candidates of the last round are expanded,
// canonically ordered and checked for distinct indices by one CTA each; an independent validator kernel recomputes
// the BLAKE2b strings of a solution and re-checks every rule.
#include "../../include/nx_equihash.h"
#include "nx_blake2b.cuh"
#include "nx_internal.hpp"

#include <algorithm>
#include <cstring>

namespace nx
{
namespace
{
constexpr int EQ_MAX_CHUNKS = 12;      // k+1
constexpr int EQ_BUCKET_CAP = 2048;    // slots per bucket (mean load 512; tiny parameter sets have ONE bucket
                                       // whose load swings widely, hence the generous headroom)
constexpr int EQ_COLLIDE_THREADS = 256;
constexpr int EQ_MAX_CANDIDATES = 4096;

struct EqParams
{
    int n, k, cbl, chunks, ipp /* strings per digest */, hash_bytes, digest_bytes;
    uint32_t N;
    int bucket_bits;      // cbl - 8 (0 when cbl == 8)
    uint32_t num_buckets;
};

__host__ __device__ inline uint32_t eq_chunk(const uint8_t* str, int cbl, int c)
{
    uint32_t v = 0;
    for (int b = 0; b < cbl; ++b) {
        const int bit = c * cbl + b;
        v = (v << 1) | ((str[bit >> 3] >> (7 - (bit & 7))) & 1u);
    }
    return v;
}

// chunks of initial string i (shared by the generator and the validator)
__device__ void eq_string_chunks(const EqParams& p, const Blake2bState& mid, uint32_t i, uint32_t* out)
{
    Blake2bState s = mid;
    const uint32_t g = i / (uint32_t)p.ipp;
    uint8_t le[4] = {(uint8_t)g, (uint8_t)(g >> 8), (uint8_t)(g >> 16), (uint8_t)(g >> 24)};
    b2_update(s, le, 4);
    uint8_t digest[64];
    b2_final(s, digest);
    const uint8_t* str = digest + (i % (uint32_t)p.ipp) * p.hash_bytes;
    for (int c = 0; c < p.chunks; ++c) out[c] = eq_chunk(str, p.cbl, c);
}

// ---- round 0: N strings -> rows[i][chunks] ----
__global__ void eq_generate_kernel(EqParams p, Blake2bState mid, uint32_t* __restrict__ rows)
{
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;  // one digest per thread
    if (g >= (p.N + (uint32_t)p.ipp - 1) / (uint32_t)p.ipp) return;  // N need not be a multiple of 512/n
    Blake2bState s = mid;
    uint8_t le[4] = {(uint8_t)g, (uint8_t)(g >> 8), (uint8_t)(g >> 16), (uint8_t)(g >> 24)};
    b2_update(s, le, 4);
    uint8_t digest[64];
    b2_final(s, digest);
    for (int j = 0; j < p.ipp; ++j) {
        const uint32_t i = g * (uint32_t)p.ipp + (uint32_t)j;
        if (i >= p.N) break;
        const uint8_t* str = digest + j * p.hash_bytes;
        for (int c = 0; c < p.chunks; ++c) rows[(size_t)i * p.chunks + c] = eq_chunk(str, p.cbl, c);
    }
}

// ---- scatter live rows into buckets by the top bits of chunk r ----
__global__ void eq_scatter_kernel(EqParams p, int r, const uint32_t* __restrict__ rows, const uint32_t* __restrict__ count_ptr,
                                  uint32_t count_cap, uint32_t* __restrict__ bucket_cnt, uint32_t* __restrict__ bucket_rows)
{
    const uint32_t count = min(*count_ptr, count_cap);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const uint32_t c = rows[(size_t)i * p.chunks + r];
        const uint32_t b = p.bucket_bits ? (c >> 8) : 0u;
        const uint32_t slot = atomicAdd(&bucket_cnt[b], 1u);
        if (slot < EQ_BUCKET_CAP) bucket_rows[(size_t)b * EQ_BUCKET_CAP + slot] = i;
    }
}

// ---- one CTA per bucket: counting sort on the low 8 bits in shared memory, emit every colliding pair ----
__global__ void __launch_bounds__(EQ_COLLIDE_THREADS)
eq_collide_kernel(EqParams p, int r, const uint32_t* __restrict__ rows, const uint32_t* __restrict__ bucket_cnt,
                  const uint32_t* __restrict__ bucket_rows, uint32_t* __restrict__ next_rows, uint2* __restrict__ tree,
                  uint32_t* __restrict__ out_count, uint32_t out_cap)
{
    __shared__ uint32_t ids[EQ_BUCKET_CAP];
    __shared__ uint32_t sorted[EQ_BUCKET_CAP];
    __shared__ uint16_t keys[EQ_BUCKET_CAP];
    __shared__ uint32_t hist[257];
    __shared__ uint32_t start[257];
    const uint32_t b = blockIdx.x;
    const uint32_t cnt = min(bucket_cnt[b], (uint32_t)EQ_BUCKET_CAP);
    const bool last = r == p.k - 1;
    for (uint32_t i = threadIdx.x; i < 257; i += EQ_COLLIDE_THREADS) hist[i] = 0;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < cnt; i += EQ_COLLIDE_THREADS) {
        const uint32_t id = bucket_rows[(size_t)b * EQ_BUCKET_CAP + i];
        const uint32_t key = rows[(size_t)id * p.chunks + r] & 0xffu;
        ids[i] = id;
        keys[i] = (uint16_t)key;
        atomicAdd(&hist[key], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t acc = 0;
        for (int i = 0; i < 256; ++i) {
            start[i] = acc;
            acc += hist[i];
        }
        start[256] = acc;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 256; i += EQ_COLLIDE_THREADS) hist[i] = start[i];  // running insert positions
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < cnt; i += EQ_COLLIDE_THREADS) sorted[atomicAdd(&hist[keys[i]], 1u)] = ids[i];
    __syncthreads();
    // position i pairs with every later position of its key group
    for (uint32_t i = threadIdx.x; i < cnt; i += EQ_COLLIDE_THREADS) {
        const uint32_t ia = sorted[i];
        const uint32_t key = rows[(size_t)ia * p.chunks + r] & 0xffu;
        const uint32_t end = start[key + 1];
        const uint32_t* ra = rows + (size_t)ia * p.chunks;
        for (uint32_t j = i + 1; j < end; ++j) {
            const uint32_t ib = sorted[j];
            const uint32_t* rb = rows + (size_t)ib * p.chunks;
            uint32_t x[EQ_MAX_CHUNKS];
            uint32_t any = 0;
            for (int c = r + 1; c < p.chunks; ++c) {
                x[c] = ra[c] ^ rb[c];
                any |= x[c];
            }
            if (last ? any != 0 : any == 0) continue;  // last round: both remaining chunks must cancel; earlier: drop
                                                       // identical remainders (the subtrees share their index set)
            const uint32_t slot = atomicAdd(out_count, 1u);
            if (slot >= out_cap) continue;
            tree[slot] = make_uint2(ia, ib);
            if (!last)
                for (int c = 0; c < p.chunks; ++c) next_rows[(size_t)slot * p.chunks + c] = c <= r ? 0u : x[c];
        }
    }
}

// ---- candidates -> index lists: expand the tree, order canonically, require distinct indices ----
// grid = candidates, block = 2^k threads (<= 512); trees[r] is the pair table written by round r-1 (r = 1..k).
struct EqTrees
{
    const uint2* t[EQ_MAX_CHUNKS];
};

__global__ void eq_recover_kernel(EqParams p, EqTrees trees, const uint32_t* __restrict__ cand_count, uint32_t cand_cap,
                                  uint32_t* __restrict__ out_indices, uint32_t* __restrict__ out_valid)
{
    __shared__ uint32_t idx[512];
    __shared__ uint32_t tmp[512];
    __shared__ int ok;
    const uint32_t cand = blockIdx.x;
    const uint32_t leaves = 1u << p.k;
    const uint32_t t = threadIdx.x;
    if (cand >= min(*cand_count, cand_cap)) return;
    if (t == 0) ok = 1;
    // walk from the root to leaf t: bit (k-1-d) of t picks the child at depth d
    uint32_t slot = cand;
    for (int round = p.k; round >= 1; --round) {
        const uint2 pr = trees.t[round][slot];
        slot = ((t >> (round - 1)) & 1u) ? pr.y : pr.x;
    }
    idx[t] = slot;  // a round-0 slot IS the initial index
    __syncthreads();
    // ordering rule, bottom-up: at every node the subtree with the smaller first index goes left
    for (int level = 0; level < p.k; ++level) {
        const uint32_t width = 1u << level;
        const uint32_t node = t / (2 * width);
        const uint32_t left = node * 2 * width;
        const bool swap = idx[left] > idx[left + width];
        const uint32_t v = idx[t];
        __syncthreads();
        if (swap) idx[t ^ width] = v;  // exchange the two halves of the node
        __syncthreads();
    }
    // distinct indices: every thread compares its index with all later ones
    const uint32_t mine = idx[t];
    tmp[t] = mine;
    __syncthreads();
    for (uint32_t j = t + 1; j < leaves; ++j)
        if (tmp[j] == mine) ok = 0;
    __syncthreads();
    out_indices[(size_t)cand * leaves + t] = idx[t];
    if (t == 0) out_valid[cand] = (uint32_t)ok;
}

// ---- GPU-side validator: recompute the strings of a claimed solution and re-check every rule ----
// grid = solutions, block = 2^k threads.  rows live in dynamic shared memory: leaves * chunks words.
__global__ void eq_validate_kernel(EqParams p, Blake2bState mid, const uint32_t* __restrict__ indices, uint32_t* __restrict__ verdict)
{
    extern __shared__ uint32_t vrows[];
    __shared__ uint32_t vidx[512];
    __shared__ int ok;
    const uint32_t leaves = 1u << p.k;
    const uint32_t t = threadIdx.x;
    const uint32_t* sol = indices + (size_t)blockIdx.x * leaves;
    if (t == 0) ok = 1;
    __syncthreads();
    const uint32_t mine = sol[t];
    vidx[t] = mine;
    if (mine >= p.N) {
        ok = 0;
    } else {
        uint32_t ch[EQ_MAX_CHUNKS];
        eq_string_chunks(p, mid, mine, ch);
        for (int c = 0; c < p.chunks; ++c) vrows[t * p.chunks + c] = ch[c];
    }
    __syncthreads();
    for (uint32_t j = t + 1; j < leaves; ++j)
        if (vidx[j] == mine) ok = 0;
    __syncthreads();
    if (ok) {
        for (int level = 0; level < p.k; ++level) {
            const uint32_t width = 1u << level;
            if (t % (2 * width) == 0) {  // t is the first leaf of a left subtree at this level
                uint32_t* a = vrows + (size_t)t * p.chunks;
                const uint32_t* b = vrows + (size_t)(t + width) * p.chunks;
                if (a[level] != b[level]) ok = 0;            // collision on this round's chunk
                if (vidx[t] >= vidx[t + width]) ok = 0;      // ordering rule
                for (int c = 0; c < p.chunks; ++c) a[c] ^= b[c];
            }
            __syncthreads();
        }
        if (t == 0)
            for (int c = 0; c < p.chunks; ++c)
                if (vrows[c]) ok = 0;
    }
    __syncthreads();
    if (t == 0) verdict[blockIdx.x] = (uint32_t)ok;
}

bool eq_make_params(EqParams& p, int n, int k)
{
    if (n <= 0 || k <= 0 || n % 8 || n % (k + 1) || n > 512 || k + 1 > EQ_MAX_CHUNKS || k > 9) return false;
    const int cbl = n / (k + 1);
    if (cbl < 8 || cbl > 24) return false;
    p.n = n;
    p.k = k;
    p.cbl = cbl;
    p.chunks = k + 1;
    p.ipp = 512 / n;
    p.hash_bytes = n / 8;
    p.digest_bytes = p.ipp * p.hash_bytes;
    p.N = 1u << (cbl + 1);
    p.bucket_bits = cbl - 8;
    p.num_buckets = 1u << p.bucket_bits;
    return true;
}

void eq_midstate(const EqParams& p, const uint8_t tag[8], const uint8_t* header, size_t len, Blake2bState& s)
{
    uint8_t personal[16];
    std::memcpy(personal, tag, 8);
    const uint32_t le_n = (uint32_t)p.n, le_k = (uint32_t)p.k;
    std::memcpy(personal + 8, &le_n, 4);
    std::memcpy(personal + 12, &le_k, 4);
    b2_init(s, (uint32_t)p.digest_bytes, personal);
    b2_update(s, header, len);
}
}  // namespace
}  // namespace nx

using namespace nx;

struct nx_equihash
{
    EqParams p;
    int device;
    uint8_t tag[8];
    uint32_t cap;  // rows per round
    uint32_t* rows[2] = {nullptr, nullptr};
    uint32_t* bucket_cnt = nullptr;
    uint32_t* bucket_rows = nullptr;
    uint2* tree[EQ_MAX_CHUNKS] = {nullptr};
    uint32_t* counts = nullptr;       // [k+1] row counts per round (device)
    uint32_t* cand_indices = nullptr; // EQ_MAX_CANDIDATES x 2^k
    uint32_t* cand_valid = nullptr;
    uint32_t* verdict = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0;
    int last_launches = 0;
    std::mutex mu;
};

extern "C" {

size_t nx_equihash_solution_bytes(int n, int k)
{
    EqParams p;
    if (!eq_make_params(p, n, k)) return 0;
    return ((size_t)1 << k) * (size_t)(p.cbl + 1) / 8;
}

void nx_equihash_destroy(nx_equihash* s)
{
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    for (int i = 0; i < 2; ++i)
        if (s->rows[i]) cudaFree(s->rows[i]);
    if (s->bucket_cnt) cudaFree(s->bucket_cnt);
    if (s->bucket_rows) cudaFree(s->bucket_rows);
    for (int i = 0; i < EQ_MAX_CHUNKS; ++i)
        if (s->tree[i]) cudaFree(s->tree[i]);
    if (s->counts) cudaFree(s->counts);
    if (s->cand_indices) cudaFree(s->cand_indices);
    if (s->cand_valid) cudaFree(s->cand_valid);
    if (s->verdict) cudaFree(s->verdict);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

int nx_equihash_create(int device, int n, int k, const uint8_t tag[8], nx_equihash** out)
{
    if (!out || !tag) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    *out = nullptr;
    EqParams p;
    if (!eq_make_params(p, n, k)) {
        set_error("unsupported Equihash parameters (n=%d, k=%d): need n %% 8 == 0, n %% (k+1) == 0, 8 <= n/(k+1) <= 24, k <= 9", n, k);
        return NX_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: this library has no CPU fallback");
        return NX_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
        set_error("device %d out of range (have %d)", device, ndev);
        return NX_ERR_INVALID;
    }
    NX_CUDA(cudaSetDevice(device));
    nx_equihash* s = new (std::nothrow) nx_equihash;
    if (!s) return NX_ERR_OOM;
    s->p = p;
    s->device = device;
    std::memcpy(s->tag, tag, 8);
    // rows per round hover around N (count' ~ count^2 / N); relative swings are large only for tiny N
    s->cap = p.N <= (1u << 12) ? 4 * p.N : p.N + p.N / 2;
    const size_t row_bytes = (size_t)s->cap * p.chunks * 4;
    const uint32_t leaves = 1u << k;
    bool ok = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) == cudaSuccess && cudaEventCreate(&s->ev0) == cudaSuccess &&
              cudaEventCreate(&s->ev1) == cudaSuccess && cudaMalloc((void**)&s->rows[0], row_bytes) == cudaSuccess &&
              cudaMalloc((void**)&s->rows[1], row_bytes) == cudaSuccess &&
              cudaMalloc((void**)&s->bucket_cnt, (size_t)p.num_buckets * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->bucket_rows, (size_t)p.num_buckets * EQ_BUCKET_CAP * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->counts, (size_t)(k + 2) * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->cand_indices, (size_t)EQ_MAX_CANDIDATES * leaves * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->cand_valid, (size_t)EQ_MAX_CANDIDATES * 4) == cudaSuccess &&
              cudaMalloc((void**)&s->verdict, (size_t)EQ_MAX_CANDIDATES * 4) == cudaSuccess;
    for (int r = 1; ok && r <= k; ++r)
        ok = cudaMalloc((void**)&s->tree[r], (size_t)(r == k ? EQ_MAX_CANDIDATES : s->cap) * sizeof(uint2)) == cudaSuccess;
    if (!ok) {
        set_error("Equihash(%d,%d): device allocation failed: %s", n, k, cudaGetErrorString(cudaGetLastError()));
        nx_equihash_destroy(s);
        return NX_ERR_OOM;
    }
    *out = s;
    return NX_OK;
}

// All solutions of one header (header = everything that is hashed before the LE32 string counter, nonce included).
// solutions_out receives up to max_solutions minimal encodings, sorted bytewise; *n_solutions the number found.
int nx_equihash_solve(nx_equihash* s, const uint8_t* header, size_t header_len, uint8_t* solutions_out, int max_solutions,
                      int* n_solutions)
{
    if (!s || !header || !n_solutions || (!solutions_out && max_solutions > 0)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    *n_solutions = 0;
    std::lock_guard<std::mutex> lk(s->mu);
    NX_CUDA(cudaSetDevice(s->device));
    const EqParams& p = s->p;
    Blake2bState mid;
    eq_midstate(p, s->tag, header, header_len, mid);
    const int k = p.k;
    int launches = 0;

    NX_CUDA(cudaEventRecord(s->ev0, s->stream));
    NX_CUDA(cudaMemsetAsync(s->counts, 0, (size_t)(k + 2) * 4, s->stream));
    const uint32_t digests = (p.N + (uint32_t)p.ipp - 1) / (uint32_t)p.ipp;
    eq_generate_kernel<<<(digests + 127) / 128, 128, 0, s->stream>>>(p, mid, s->rows[0]);
    ++launches;
    NX_CUDA(cudaMemcpyAsync(s->counts, &p.N, 4, cudaMemcpyHostToDevice, s->stream));  // p lives in *s: stays valid
    for (int r = 0; r < k; ++r) {
        const bool last = r == k - 1;
        const uint32_t* in = s->rows[r & 1];
        uint32_t* outr = s->rows[(r + 1) & 1];
        NX_CUDA(cudaMemsetAsync(s->bucket_cnt, 0, (size_t)p.num_buckets * 4, s->stream));
        eq_scatter_kernel<<<148 * 8, 256, 0, s->stream>>>(p, r, in, s->counts + r, s->cap, s->bucket_cnt, s->bucket_rows);
        eq_collide_kernel<<<p.num_buckets, EQ_COLLIDE_THREADS, 0, s->stream>>>(
            p, r, in, s->bucket_cnt, s->bucket_rows, outr, s->tree[r + 1], s->counts + r + 1,
            last ? (uint32_t)EQ_MAX_CANDIDATES : s->cap);
        launches += 2;
    }
    EqTrees trees;
    for (int i = 0; i < EQ_MAX_CHUNKS; ++i) trees.t[i] = s->tree[i];
    const uint32_t leaves = 1u << k;
    eq_recover_kernel<<<EQ_MAX_CANDIDATES, leaves, 0, s->stream>>>(p, trees, s->counts + k, EQ_MAX_CANDIDATES, s->cand_indices,
                                                                   s->cand_valid);
    ++launches;
    NX_CUDA(cudaGetLastError());
    NX_CUDA(cudaEventRecord(s->ev1, s->stream));
    std::vector<uint32_t> counts((size_t)k + 2);
    NX_CUDA(cudaMemcpyAsync(counts.data(), s->counts, (size_t)(k + 2) * 4, cudaMemcpyDeviceToHost, s->stream));
    NX_CUDA(cudaStreamSynchronize(s->stream));
    NX_CUDA(cudaEventElapsedTime(&s->last_ms, s->ev0, s->ev1));
    s->last_launches = launches;
    const uint32_t ncand = std::min<uint32_t>(counts[(size_t)k], EQ_MAX_CANDIDATES);
    if (ncand == 0) return NX_OK;
    std::vector<uint32_t> valid(ncand), idx((size_t)ncand * leaves);
    NX_CUDA(cudaMemcpy(valid.data(), s->cand_valid, (size_t)ncand * 4, cudaMemcpyDeviceToHost));
    NX_CUDA(cudaMemcpy(idx.data(), s->cand_indices, (size_t)ncand * leaves * 4, cudaMemcpyDeviceToHost));
    const size_t sb = ((size_t)1 << k) * (size_t)(p.cbl + 1) / 8;
    std::vector<std::vector<uint8_t>> sols;
    for (uint32_t c = 0; c < ncand; ++c) {
        if (!valid[c]) continue;
        std::vector<uint8_t> enc(sb, 0);
        size_t pos = 0;
        for (uint32_t i = 0; i < leaves; ++i)
            for (int b = p.cbl; b >= 0; --b, ++pos)
                if ((idx[(size_t)c * leaves + i] >> b) & 1u) enc[pos >> 3] |= (uint8_t)(0x80u >> (pos & 7));
        sols.push_back(std::move(enc));
    }
    std::sort(sols.begin(), sols.end());
    sols.erase(std::unique(sols.begin(), sols.end()), sols.end());
    *n_solutions = (int)sols.size();
    for (size_t i = 0; i < sols.size() && (int)i < max_solutions; ++i) std::memcpy(solutions_out + i * sb, sols[i].data(), sb);
    return NX_OK;
}

// GPU-side validator for `count` encoded solutions of one header: verdict[i] = 1 when solution i is valid.
int nx_equihash_validate(nx_equihash* s, const uint8_t* header, size_t header_len, const uint8_t* solutions, int count,
                         uint8_t* verdict)
{
    if (!s || !header || ((!solutions || !verdict) && count > 0)) {
        set_error("null argument");
        return NX_ERR_INVALID;
    }
    if (count <= 0) return NX_OK;
    if (count > EQ_MAX_CANDIDATES) {
        set_error("at most %d solutions per call", EQ_MAX_CANDIDATES);
        return NX_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(s->mu);
    NX_CUDA(cudaSetDevice(s->device));
    const EqParams& p = s->p;
    const uint32_t leaves = 1u << p.k;
    const size_t sb = (size_t)leaves * (size_t)(p.cbl + 1) / 8;
    std::vector<uint32_t> idx((size_t)count * leaves);
    for (int c = 0; c < count; ++c) {
        size_t pos = 0;
        const uint8_t* in = solutions + (size_t)c * sb;
        for (uint32_t i = 0; i < leaves; ++i) {
            uint32_t v = 0;
            for (int b = 0; b <= p.cbl; ++b, ++pos) v = (v << 1) | ((in[pos >> 3] >> (7 - (pos & 7))) & 1u);
            idx[(size_t)c * leaves + i] = v;
        }
    }
    Blake2bState mid;
    eq_midstate(p, s->tag, header, header_len, mid);
    NX_CUDA(cudaMemcpyAsync(s->cand_indices, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, s->stream));
    eq_validate_kernel<<<count, leaves, (size_t)leaves * p.chunks * 4, s->stream>>>(p, mid, s->cand_indices, s->verdict);
    NX_CUDA(cudaGetLastError());
    std::vector<uint32_t> v((size_t)count);
    NX_CUDA(cudaMemcpyAsync(v.data(), s->verdict, (size_t)count * 4, cudaMemcpyDeviceToHost, s->stream));
    NX_CUDA(cudaStreamSynchronize(s->stream));
    for (int c = 0; c < count; ++c) verdict[c] = (uint8_t)(v[(size_t)c] ? 1 : 0);
    return NX_OK;
}

int nx_equihash_last_timing(const nx_equihash* s, float* kernel_ms, int* launches)
{
    if (!s) return NX_ERR_INVALID;
    if (kernel_ms) *kernel_ms = s->last_ms;
    if (launches) *launches = s->last_launches;
    return NX_OK;
}

}  // extern "C"
