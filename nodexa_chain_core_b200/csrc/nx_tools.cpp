// Introspection entry points used by the tests and by tools/ (no device needed):
// the generated program text and an NVRTC compile that stops at the cubin.
#include "nx_internal.hpp"

#include <cstring>

using namespace nx;

extern "C" {

// Writes the generated CUDA text for (prog_seed, dag_items) into buf (NUL-terminated, truncated to
// cap); returns the full length.  with_skeleton=0 gives only the two NX_ROUND_* macros.
size_t nx_kawpow_program_source(uint64_t prog_seed, uint32_t dag_items_256, int with_skeleton, char* buf, size_t cap)
{
    const std::string s = with_skeleton ? kawpow_source(prog_seed, dag_items_256) : kawpow_program_text(prog_seed);
    if (buf && cap) {
        const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
        std::memcpy(buf, s.data(), n);
        buf[n] = 0;
    }
    return s.size();
}

// Flat op list of the period program: ops[i] = {kind, src1, src2, dst, sel_math, sel_merge}; returns count (33).
int nx_kawpow_program_ops(uint64_t prog_seed, uint32_t (*ops)[6], int cap)
{
    const std::vector<ProgramOp> v = kawpow_program(prog_seed);
    int n = 0;
    for (const ProgramOp& op : v) {
        if (n >= cap) break;
        ops[n][0] = (uint32_t)op.kind;
        ops[n][1] = op.src1;
        ops[n][2] = op.src2;
        ops[n][3] = op.dst;
        ops[n][4] = op.sel_math;
        ops[n][5] = op.sel_merge;
        ++n;
    }
    return n;
}

// NVRTC-compile the period program for sm_100a without touching a device; the cubin is copied to
// buf when it fits.  Returns the cubin size, or a negative nx_status.
long nx_kawpow_compile_cubin(uint64_t prog_seed, uint32_t dag_items_256, char* buf, size_t cap)
{
    std::vector<char> cubin;
    std::string log;
    const int rc = kawpow_compile_only(prog_seed, dag_items_256, cubin, log);
    if (rc != NX_OK) return rc;
    if (buf && cap >= cubin.size()) std::memcpy(buf, cubin.data(), cubin.size());
    return (long)cubin.size();
}

}  // extern "C"

extern "C" {
// The product's host light-cache builder on its own (CPU tests compare it with the reference-made digest).
void nx_host_light_cache(int epoch, uint8_t* out)
{
    uint8_t seed[32];
    nx_epoch_seed(epoch, seed);
    host_build_light_cache(out, nx_light_cache_num_items(epoch), seed);
}
}

extern "C" {
// which = 0: dispatching entry, 1: scalar, 2: AVX-512 (returns -1 when the CPU lacks it).  CPU tests time and compare them.
int nx_host_light_cache_variant(int epoch, int which, uint8_t* out)
{
    uint8_t seed[32];
    nx_epoch_seed(epoch, seed);
    const int n = nx_light_cache_num_items(epoch);
    if (which == 2) {
        if (!cpu_has_avx512f()) return -1;
        host_build_light_cache_avx512(out, n, seed);
    } else if (which == 1) {
        host_build_light_cache_scalar(out, n, seed);
    } else {
        host_build_light_cache(out, n, seed);
    }
    return 0;
}

int nx_keccakf1600_avx512(uint64_t state[25])
{
    if (!cpu_has_avx512f()) return -1;
    keccakf1600_avx512(state);
    return 0;
}
}

extern "C" {
// Light cache through the NX_CACHE_DIR file cache (tests): returns 1 when it was served from disk, 0 when built.
int nx_host_light_cache_cached(int epoch, uint8_t* out)
{
    const int n = nx_light_cache_num_items(epoch);
    if (host_load_light_cache(out, n, epoch)) return 1;
    uint8_t seed[32];
    nx_epoch_seed(epoch, seed);
    host_build_light_cache(out, n, seed);
    host_store_light_cache(out, n, epoch);
    return 0;
}
}

extern "C" {
// The launch shape the KawPow kernel gets for `count` nonces on a device with `sms` multiprocessors (CPU tests).
void nx_kawpow_launch_shape(uint64_t count, int threads, int sms, int minblocks, int persist, uint32_t* per_hw,
                            uint32_t* n_units, uint32_t* grid)
{
    kawpow_launch_shape(count, threads, sms, minblocks, persist != 0, per_hw, n_units, grid);
}
}
