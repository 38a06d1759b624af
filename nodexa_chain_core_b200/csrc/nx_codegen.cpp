// Per-period ProgPoW program generator: turns prog_seed = block_number / 3 into straight-line
// CUDA text in which every register index and selector is a literal.
//
// Restates, as a code generator, what the reference INTERPRETS for every hash:
//   mix_rng_state            src/crypto/ethash/lib/ethash/progpow.cpp:46-88
//   the draw order of round  progpow.cpp:191-233
//   random_math / merge      progpow.cpp:92-144
// The reference re-derives this program for each of the 64 rounds of each hash (the rng state is
// passed to round() by value, progpow.cpp:179-180); here it is derived once per period.
#include "nx_internal.hpp"

#include <sstream>

namespace nx
{
namespace
{
constexpr uint32_t FNV_PRIME = 0x01000193u, FNV_BASIS = 0x811c9dc5u;
inline uint32_t fnv1a(uint32_t h, uint32_t d) { return (h ^ d) * FNV_PRIME; }

struct Kiss99
{
    uint32_t z, w, jsr, jcong;
    uint32_t next()
    {
        z = 36969u * (z & 0xffffu) + (z >> 16);
        w = 18000u * (w & 0xffffu) + (w >> 16);
        jcong = 69069u * jcong + 1234567u;
        jsr ^= jsr << 17;
        jsr ^= jsr >> 13;
        jsr ^= jsr << 5;
        return (((z << 16) + w) ^ jcong) + jsr;
    }
};

struct ProgramRng
{
    Kiss99 rng;
    uint32_t dst_seq[32], src_seq[32];
    uint32_t dst_n = 0, src_n = 0;

    explicit ProgramRng(uint64_t prog_seed)
    {
        const uint32_t lo = (uint32_t)prog_seed, hi = (uint32_t)(prog_seed >> 32);
        rng.z = fnv1a(FNV_BASIS, lo);
        rng.w = fnv1a(rng.z, hi);
        rng.jsr = fnv1a(rng.w, lo);
        rng.jcong = fnv1a(rng.jsr, hi);
        for (uint32_t i = 0; i < 32; ++i) dst_seq[i] = src_seq[i] = i;
        for (uint32_t i = 32; i > 1; --i) {
            std::swap(dst_seq[i - 1], dst_seq[rng.next() % i]);
            std::swap(src_seq[i - 1], src_seq[rng.next() % i]);
        }
    }
    uint32_t next_dst() { return dst_seq[dst_n++ % 32]; }
    uint32_t next_src() { return src_seq[src_n++ % 32]; }
};

std::string reg(const char* arr, uint32_t i)
{
    std::ostringstream s;
    s << arr << "[" << i << "]";
    return s.str();
}

// a = merge(a, b, sel)   (progpow.cpp:126-144)
std::string merge_expr(const std::string& a, const std::string& b, uint32_t sel, int& rot_ordinal)
{
    const uint32_t x = ((sel >> 16) % 31) + 1;
    std::ostringstream s;
    switch (sel % 4) {
    case 0: s << a << " = " << a << " * 33u + " << b << ";"; break;
    case 1: s << a << " = (" << a << " ^ " << b << ") * 33u;"; break;
    case 2: s << a << " = nx_rotl_c(" << a << ", " << x << "u, " << rot_ordinal++ << "u) ^ " << b << ";"; break;
    default: s << a << " = nx_rotr_c(" << a << ", " << x << "u, " << rot_ordinal++ << "u) ^ " << b << ";"; break;
    }
    return s.str();
}

// math(a, b, sel)   (progpow.cpp:92-120)
std::string math_expr(const std::string& a, const std::string& b, uint32_t sel)
{
    switch (sel % 11) {
    default:
    case 0: return a + " + " + b;
    case 1: return a + " * " + b;
    case 2: return "nx_mulhi(" + a + ", " + b + ")";
    case 3: return "nx_min(" + a + ", " + b + ")";
    case 4: return "nx_rotl(" + a + ", " + b + ")";
    case 5: return "nx_rotr(" + a + ", " + b + ")";
    case 6: return a + " & " + b;
    case 7: return a + " | " + b;
    case 8: return a + " ^ " + b;
    case 9: return "nx_clz(" + a + ") + nx_clz(" + b + ")";
    case 10: return "nx_popc(" + a + ") + nx_popc(" + b + ")";
    }
}
}  // namespace

std::vector<ProgramOp> kawpow_program(uint64_t prog_seed)
{
    ProgramRng st(prog_seed);
    std::vector<ProgramOp> ops;
    for (int i = 0; i < 18; ++i) {
        if (i < 11) {
            ProgramOp op{};
            op.kind = ProgramOp::CACHE;
            op.src1 = st.next_src();
            op.dst = st.next_dst();
            op.sel_merge = st.rng.next();
            ops.push_back(op);
        }
        ProgramOp op{};
        op.kind = ProgramOp::MATH;
        const uint32_t rnd = st.rng.next() % (32 * 31);
        op.src1 = rnd % 32;
        op.src2 = rnd / 32;
        if (op.src2 >= op.src1) ++op.src2;
        op.sel_math = st.rng.next();
        op.dst = st.next_dst();
        op.sel_merge = st.rng.next();
        ops.push_back(op);
    }
    for (int i = 0; i < 4; ++i) {
        ProgramOp op{};
        op.kind = ProgramOp::DAG;
        op.dst = i == 0 ? 0 : st.next_dst();
        op.sel_merge = st.rng.next();
        op.src1 = (uint32_t)i;
        ops.push_back(op);
    }
    return ops;
}

std::string kawpow_program_text(uint64_t prog_seed)
{
    const std::vector<ProgramOp> ops = kawpow_program(prog_seed);
    std::ostringstream cm, dg;
    cm << "#define NX_ROUND_CACHE_MATH(m, L1) do { \\\n";
    dg << "#define NX_ROUND_DAG(m, d0, d1, d2, d3) do { \\\n";
    static const char* dag_words[4] = {"(d0)", "(d1)", "(d2)", "(d3)"};
    int n = 0, rot = 0;
    for (const ProgramOp& op : ops) {
        switch (op.kind) {
        case ProgramOp::CACHE:
            cm << "  { const u32 t" << n << " = L1(" << reg("m", op.src1) << "); "
               << merge_expr(reg("m", op.dst), "t" + std::to_string(n), op.sel_merge, rot) << " } \\\n";
            break;
        case ProgramOp::MATH:
            cm << "  { const u32 t" << n << " = " << math_expr(reg("m", op.src1), reg("m", op.src2), op.sel_math) << "; "
               << merge_expr(reg("m", op.dst), "t" + std::to_string(n), op.sel_merge, rot) << " } \\\n";
            break;
        case ProgramOp::DAG:
            dg << "  " << merge_expr(reg("m", op.dst), dag_words[op.src1], op.sel_merge, rot) << " \\\n";
            break;
        }
        ++n;
    }
    cm << "} while (0)\n";
    dg << "} while (0)\n";
    return cm.str() + dg.str();
}

std::string kawpow_source(uint64_t prog_seed, uint32_t dag_items_256)
{
    std::ostringstream s;
    s << "// generated: prog_seed=" << prog_seed << " dag_items=" << dag_items_256 << "\n";
    s << "#define NX_PROG_SEED " << prog_seed << "ull\n";
    s << "#define NX_DAG_ITEMS " << dag_items_256 << "u\n";
    s << kawpow_program_text(prog_seed);
    s << kawpow_kernel_skeleton();
    return s.str();
}

}  // namespace nx
