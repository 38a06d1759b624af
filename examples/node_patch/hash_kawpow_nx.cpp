// The node-side binding of INTEGRATION.md section 1 as a compilable file: drop-in definitions of KAWPOWHash and
// KAWPOWHash_OnlyMix (declared in the node's src/hash.h:608-609, defined in src/hash.cpp:258-291) over the nx_* C ABI.
// A maintainer deletes those two functions from src/hash.cpp and adds this file to libclore_common; nothing else in the
// node changes.  It includes the NODE's headers, so it builds only inside the node tree (or, for the tests of this
// repository, against /root/reference/src - see oracle/Makefile, target `ref_node_explicit`).
#include <nx_kawpow.h>

#include "hash.h"
#include "primitives/block.h"
#include "uint256.h"
#include "utilstrencodings.h"

#include <algorithm>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <string>

// Replaces the unsynchronised function-static epoch_context_ptr (hash.cpp:260).  A caller copies the shared_ptr under
// the lock and keeps it for the duration of its hash, so a thread that validates a header of another epoch can swap the
// context without pulling it from under a hash in flight; the old DAG is freed when its last user lets go.
static std::shared_ptr<nx_epoch> g_nx_ctx;
static std::mutex g_nx_ctx_mu;

static std::shared_ptr<nx_epoch> NxContextFor(int nHeight)
{
    const int epoch = nHeight / NX_EPOCH_LENGTH;  // ethash::get_epoch_number (ethash.hpp:75-78)
    std::lock_guard<std::mutex> lk(g_nx_ctx_mu);
    if (!g_nx_ctx || nx_epoch_number(g_nx_ctx.get()) != epoch) {
        g_nx_ctx.reset();
        const char* dev = std::getenv("NX_DEVICE");
        nx_epoch* fresh = nullptr;
        if (nx_epoch_create(epoch, dev ? std::atoi(dev) : 0, &fresh) != NX_OK)
            LogPrintf("nx: %s\n", nx_last_error());
        else
            g_nx_ctx.reset(fresh, nx_epoch_destroy);
    }
    return g_nx_ctx;
}

uint256 KAWPOWHash(const CBlockHeader& blockHeader, uint256& mix_hash)
{
    const std::shared_ptr<nx_epoch> ctx = NxContextFor((int)blockHeader.nHeight);
    uint8_t header_hash[32];
    // = to_hash256(GetKAWPOWHeaderHash().GetHex()), hash.cpp:269-270
    nx_kawpow_header_hash(blockHeader.nVersion, blockHeader.hashPrevBlock.begin(), blockHeader.hashMerkleRoot.begin(),
                          blockHeader.nTime, blockHeader.nBits, blockHeader.nHeight, header_hash);
    nx_result r;
    if (nx_kawpow_hash(ctx.get(), (int)blockHeader.nHeight, header_hash, blockHeader.nNonce64, &r) != NX_OK) {  // progpow::hash, :273
        // no context or a failed device call: fail closed - a final hash no target accepts, a mix no header carries
        LogPrintf("nx: %s\n", nx_last_error());
        mix_hash = uint256S(std::string(64, 'f'));
        return uint256S(std::string(64, 'f'));
    }
    mix_hash = uint256S(HexStr(r.mix_hash, r.mix_hash + 32));  // uint256S(to_hex(...)), hash.cpp:275-276
    return uint256S(HexStr(r.final_hash, r.final_hash + 32));
}

uint256 KAWPOWHash_OnlyMix(const CBlockHeader& blockHeader)
{
    uint8_t header_hash[32], mix[32], out[32];
    nx_kawpow_header_hash(blockHeader.nVersion, blockHeader.hashPrevBlock.begin(), blockHeader.hashMerkleRoot.begin(),
                          blockHeader.nTime, blockHeader.nBits, blockHeader.nHeight, header_hash);
    std::reverse_copy(blockHeader.mix_hash.begin(), blockHeader.mix_hash.end(), mix);  // to_hash256(mix_hash.GetHex())
    nx_kawpow_hash_no_verify((int)blockHeader.nHeight, header_hash, mix, blockHeader.nNonce64, out);  // hash.cpp:287
    return uint256S(HexStr(out, out + 32));
}
