// TEST HARNESS: runs the PRODUCT's generated per-period program (and the scalar helpers of
// kawpow_kernel.cuh: keccak-f800, fill_mix, lane reduce) on the CPU, lanes emulated by a loop, so that
// the code generator can be checked against the oracle without a GPU.  Compiled by
// tests/test_codegen.py with  g++ -include <generated source> kawpow_emul.cpp.
// DAG items come from the oracle (this is test code; the product never does this).
#include <stdint.h>
#include <string.h>

extern "C" void nxo_dataset_item_2048(const void* ctx, uint32_t index, uint8_t out[256]);

extern "C" void emul_kawpow_hash(const void* oracle_ctx, const uint32_t* l1, const uint8_t header[32], uint64_t nonce,
                                 uint8_t final_out[32], uint8_t mix_out[32])
{
    u32 st[25];
    memcpy(st, header, 32);
    st[8] = (u32)nonce;
    st[9] = (u32)(nonce >> 32);
    NX_PAD_FIRST(st);
    nxk_f800(st);
    u32 carry[8];
    memcpy(carry, st, 32);

    static u32 m[16][32];
    for (u32 l = 0; l < 16; ++l) nx_fill_mix(carry[0], carry[1], l, m[l]);
    for (u32 r = 0; r < 64; ++r) {
        const u32 item = m[r & 15][0] % NX_DAG_ITEMS;
        u32 words[64];
        nxo_dataset_item_2048(oracle_ctx, item, (uint8_t*)words);
        for (u32 l = 0; l < 16; ++l) {
            const u32* d = words + ((l ^ r) & 15) * 4;
#define EMUL_L1(x) l1[(x) & 4095u]
            NX_ROUND_CACHE_MATH(m[l], EMUL_L1);
            NX_ROUND_DAG(m[l], d[0], d[1], d[2], d[3]);
        }
    }
    u32 lh[16], dg[8];
    for (u32 l = 0; l < 16; ++l) lh[l] = nx_lane_reduce(m[l]);
    for (int i = 0; i < 8; ++i) dg[i] = nx_fnv1a(nx_fnv1a(NX_FNV_BASIS, lh[i]), lh[i + 8]);

    memcpy(st, carry, 32);
    memcpy(st + 8, dg, 32);
    NX_PAD_LAST(st);
    nxk_f800(st);
    memcpy(final_out, st, 32);
    memcpy(mix_out, dg, 32);
}
