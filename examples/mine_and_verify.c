/* Pure C99 against the C ABI (include/nx_kawpow.h): mine one synthetic KawPow block on every visible GPU, serialise it
 * as the node's 120-byte wire header, and check it the way CheckBlockHeader would.
 *
 *   gcc -std=c99 -Iinclude examples/mine_and_verify.c -Lnodexa_chain_core_b200 -lnxkawpow \
 *       -Wl,-rpath,$PWD/nodexa_chain_core_b200 -o /tmp/mine_and_verify && /tmp/mine_and_verify
 */
#include <nx_kawpow.h>

#include <stdio.h>
#include <string.h>

int main(void)
{
    nx_miner* miner = NULL;
    if (nx_miner_create(NULL, 0, &miner) != NX_OK) { /* all visible GPUs */
        fprintf(stderr, "miner: %s\n", nx_last_error());
        return 2;
    }
    nx_block_header h;
    memset(&h, 0, sizeof h);
    h.version = 0x30000000;
    memset(h.prev_block, 0x11, 32);
    memset(h.merkle_root, 0x22, 32);
    h.time = 1651444217u + 600u; /* after nKAWPOWActivationTime (chainparams.cpp:264) */
    h.bits = 0x1f00ffffu;        /* target 0000ffff00..: about one nonce in 65536 qualifies */
    h.height = 123456u;          /* epoch 16 */

    nx_search_result sr;
    if (nx_miner_mine_header(miner, h.version, h.prev_block, h.merkle_root, h.time, h.bits, h.height,
                             /*start*/ 0, /*max tries*/ 1ull << 24, /*cancel*/ NULL, &sr) != NX_OK) {
        fprintf(stderr, "mine: %s\n", nx_last_error());
        return 2;
    }
    if (!sr.found) {
        fprintf(stderr, "no nonce below the target in 2^24 tries (probability e^-256)\n");
        return 1;
    }
    h.nonce64 = sr.nonce;
    for (int i = 0; i < 32; ++i) h.mix_hash[i] = sr.mix_hash[31 - i]; /* uint256 storage order, as the node keeps it */

    uint8_t raw[120], id[32], status = 0xff;
    nx_block_header_serialize(&h, raw);
    nx_block_hash(raw, id);
    const uint8_t pow_limit[32] = {0x00, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff,
                                   0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff, 0xff};
    if (nx_check_block_headers(nx_miner_context(miner, 0), raw, 1, pow_limit, /*mix_only*/ 0, &status) != NX_OK) {
        fprintf(stderr, "check: %s\n", nx_last_error());
        return 2;
    }
    uint64_t hashes = 0;
    double secs = 0;
    nx_miner_last_stats(miner, &hashes, &secs);
    printf("devices=%d nonce=%llu status=%u hashes=%llu seconds=%.3f block_id=", nx_miner_device_count(miner),
           (unsigned long long)sr.nonce, (unsigned)status, (unsigned long long)hashes, secs);
    for (int i = 31; i >= 0; --i) printf("%02x", id[i]); /* printed like uint256::GetHex() */
    printf("\n");
    /* the block id of a valid block is its PoW hash; tampering with the nonce must be caught */
    int ok = status == NX_HEADER_OK;
    for (int i = 0; i < 32; ++i) ok = ok && id[31 - i] == sr.final_hash[i];
    raw[80] ^= 1;
    nx_check_block_headers(nx_miner_context(miner, 0), raw, 1, pow_limit, 0, &status);
    ok = ok && status != NX_HEADER_OK;
    nx_miner_destroy(miner);
    puts(ok ? "OK" : "MISMATCH");
    return ok ? 0 : 1;
}
