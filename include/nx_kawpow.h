/* nx_kawpow.h — C ABI of the B200-native KawPow / Ethash proof-of-work engine.
 *
 * Drop-in boundary for the PoW path of DeonDavisV/Nodexa-Chain-Core.  Every entry point names the
 * reference interface it replaces (paths relative to the reference's src/).  Plain pointers and
 * sizes only; no C++ or torch types cross this boundary.  All functions are thread-safe unless
 * noted; none throws.  Functions returning int return NX_OK (0) or a negative nx_status, and leave a
 * human-readable message for the calling thread in nx_last_error().
 *
 * Byte conventions (identical to the reference, see DESIGN.md "byte order"):
 *   - header_hash, mix_hash, final_hash, boundary are 32-byte strings exactly as the reference's
 *     ethash::hash256::bytes (crypto/ethash/include/ethash/hash_types.h:14-20);
 *   - boundary comparison is big-endian over those bytes (crypto/ethash/lib/ethash/ethash-internal.hpp:32-43).
 *
 * The library needs a CUDA device (sm_100a) for everything that hashes; there is no CPU fallback:
 * without a usable device those calls fail with NX_ERR_CUDA.
 */
#ifndef NX_KAWPOW_H
#define NX_KAWPOW_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum nx_status
{
    NX_OK = 0,
    NX_ERR_INVALID = -1, /* bad argument */
    NX_ERR_CUDA = -2,    /* CUDA runtime/driver failure, or no device */
    NX_ERR_NVRTC = -3,   /* period program failed to compile */
    NX_ERR_OOM = -4,     /* host or device allocation failed */
    NX_ERR_STATE = -5    /* context not ready (e.g. sharded DAG not finalized) */
} nx_status;

/* Message for the last failed call on this thread ("" if none). */
const char* nx_last_error(void);
const char* nx_version(void);
int nx_device_count(void);

/* ------------------------------------------------------------------------------------------------
 * Epoch geometry — host only.
 * Replaces ethash_calculate_light_cache_num_items / _full_dataset_num_items / _epoch_seed
 * (crypto/ethash/lib/ethash/ethash.cpp:357-393), ethash_find_largest_prime
 * (crypto/ethash/lib/ethash/primes.c:24-43) and ethash::find_epoch_number (ethash.cpp:57-97). */
#define NX_EPOCH_LENGTH 7500 /* crypto/ethash/include/ethash/ethash.h:29 */
#define NX_PERIOD_LENGTH 3   /* crypto/ethash/include/ethash/progpow.hpp:21 */
int nx_find_largest_prime(int upper_bound);
int nx_light_cache_num_items(int epoch);   /* 64-byte items  */
int nx_full_dataset_num_items(int epoch);  /* 128-byte items; KawPow addresses floor(n/2) items of 256 bytes */
void nx_epoch_seed(int epoch, uint8_t out[32]);
int nx_find_epoch_number(const uint8_t seed[32]); /* -1 when not found within 30000 epochs */

/* Host Keccak helpers, exported because the reference exports them
 * (crypto/ethash/include/ethash/keccak.h:30-45). */
void nx_keccakf1600(uint64_t state[25]);
void nx_keccakf800(uint32_t state[25]);
void nx_keccak256(const uint8_t* data, size_t size, uint8_t out[32]);
void nx_keccak512(const uint8_t* data, size_t size, uint8_t out[64]);

/* ------------------------------------------------------------------------------------------------
 * Epoch context: light cache (host + device), 16 KiB L1 cache (host mirror), full DAG in HBM.
 * Replaces ethash_create_epoch_context / _full / ethash_destroy_epoch_context(_full)
 * (crypto/ethash/lib/ethash/ethash.cpp:131-172,395-414).  On B200 the DAG (1-5 GiB) is always
 * materialised: generating it (K3) is faster than one CPU light-cache build. */
typedef struct nx_epoch nx_epoch;

/* Light cache (CUDA kernel K4 or host) + whole DAG generated on `device`. */
int nx_epoch_create(int epoch, int device, nx_epoch** out);

/* Same, but with the light cache (nx_light_cache_num_items(epoch) x 64 bytes) supplied by the caller, who ran
 * the serial chain once for several devices; only the DAG kernel runs. */
int nx_epoch_create_from_light(int epoch, int device, const uint8_t* light_cache, nx_epoch** out);

/* Multi-GPU build: this rank generates only DAG items [rank*share, min(M,(rank+1)*share)) with
 * share = ceil(M/world), written at their final offset inside a buffer padded to world*share items.
 * The caller then all-gathers nx_epoch_dag_ptr() in place (sendcount = nx_epoch_dag_share_bytes(),
 * e.g. ncclAllGather / torch.distributed.all_gather_into_tensor) and calls nx_epoch_finalize().
 * This is the only collective on the path. */
int nx_epoch_create_sharded(int epoch, int device, int rank, int world, nx_epoch** out);
int nx_epoch_finalize(nx_epoch* ctx);
void nx_epoch_destroy(nx_epoch* ctx);

/* Fused generate + all-gather over NVLink peer memory (no NCCL): every rank's DAG kernel stores its slice
 * straight into the DAG buffer of EVERY rank, so the exchange overlaps the generation instead of following it.
 *   1. nx_epoch_create_deferred(): allocations + light cache (pass light_cache = NULL to build it), no DAG yet;
 *   2. exchange buffers: ranks in separate processes trade nx_epoch_ipc_handle() blobs (64 bytes each, any
 *      transport) and call nx_epoch_generate_fused_ipc(); ranks inside one process enable peer access and call
 *      nx_epoch_generate_fused_peers() with the nx_epoch_dag_ptr() of every rank;
 *   3. barrier across ranks (every kernel has finished), then nx_epoch_finalize() on each. */
int nx_epoch_create_deferred(int epoch, int device, int rank, int world, const uint8_t* light_cache, nx_epoch** out);
int nx_epoch_ipc_handle(const nx_epoch* ctx, uint8_t handle[64]);
int nx_epoch_generate_fused_ipc(nx_epoch* ctx, const uint8_t* handles_by_rank /* world x 64 bytes; NULL once opened */);
/* One serial chain for a whole multi-process job (build_light_cache, ethash.cpp:101-129, cannot be split: every rank
 * running it redundantly makes an N-GPU epoch build SLOWER than a 1-GPU one).  Every rank creates a shell (allocations
 * only) and, while the root rank runs nx_epoch_build_light_cache(), trades IPC handles and maps its peers with
 * nx_epoch_open_ipc_peers(); the root's light cache is then broadcast in place into nx_epoch_light_ptr() of the others
 * (ncclBroadcast / torch.distributed.broadcast, nx_epoch_light_items() x 64 bytes), which call
 * nx_epoch_adopt_light_cache(); nx_epoch_generate_fused_ipc(ctx, NULL), a barrier and nx_epoch_finalize() follow. */
int nx_epoch_create_shell(int epoch, int device, int rank, int world, nx_epoch** out);
int nx_epoch_build_light_cache(nx_epoch* ctx);
void* nx_epoch_light_ptr(const nx_epoch* ctx);        /* DEVICE pointer to the light cache */
int nx_epoch_adopt_light_cache(nx_epoch* ctx);
int nx_epoch_open_ipc_peers(nx_epoch* ctx, const uint8_t* handles_by_rank /* world x 64 bytes */);
/* A shell (or deferred) context on the NCCL route: generate this rank's slice at its final offset in the own buffer;
 * the caller's in-place all-gather of nx_epoch_dag_ptr() and nx_epoch_finalize() follow, as after
 * nx_epoch_create_sharded(). */
int nx_epoch_generate_slice(nx_epoch* ctx);
int nx_epoch_generate_fused_peers(nx_epoch* ctx, void* const* peer_dag_ptrs /* world entries */);

/* Public fields of the reference's struct ethash_epoch_context (ethash.h:35-42). */
int nx_epoch_number(const nx_epoch* ctx);
int nx_epoch_light_items(const nx_epoch* ctx);
int nx_epoch_dataset_items(const nx_epoch* ctx);      /* 128-byte items, as in the reference */
const uint8_t* nx_epoch_light_cache(const nx_epoch* ctx); /* host copy, light_items*64 bytes */
const uint32_t* nx_epoch_l1_cache(const nx_epoch* ctx);   /* host mirror, 4096 words */
int nx_epoch_device(const nx_epoch* ctx);
void* nx_epoch_dag_ptr(const nx_epoch* ctx);          /* DEVICE pointer to the (padded) DAG */
size_t nx_epoch_dag_bytes(const nx_epoch* ctx);       /* addressable DAG bytes = floor(n/2)*256 */
size_t nx_epoch_dag_padded_bytes(const nx_epoch* ctx);
size_t nx_epoch_dag_share_bytes(const nx_epoch* ctx); /* bytes each rank contributes to the all-gather */
/* Timings of the build phases in milliseconds (CUDA events / host clock): light cache, DAG kernel. */
void nx_epoch_build_times(const nx_epoch* ctx, double* light_ms, double* dag_ms);

/* Regenerate `count` 256-byte DAG items starting at `first` into a caller-provided DEVICE buffer
 * (K3 alone; used by benchmarks and parity tests).  Replaces calculate_dataset_item_2048
 * (ethash.cpp:235-251) applied to a range. */
int nx_dag_generate_range(const nx_epoch* ctx, uint32_t first, uint32_t count, void* out_device);
/* Copy `count` items starting at `first` from the resident DAG to host memory. */
int nx_dag_read_items(const nx_epoch* ctx, uint32_t first, uint32_t count, uint8_t* out_host);
/* Position-weighted 64-bit sum over the addressable DAG: sum_i u64word_i*(2i+1) mod 2^64.
 * No reference counterpart (test/ops aid: equal on every GPU after the all-gather). */
int nx_dag_checksum(const nx_epoch* ctx, uint64_t* out);

/* The same weighted sum for callers that must not allocate.  No counterpart in the reference.
 *
 * Who needs it: code that runs next to a live search - cudaMalloc/cudaFree stall every stream of the
 * device while they run, so a DAG integrity check from a watchdog or validation thread would punch a
 * hole into the hash rate - and plug-ins living inside a host application that owns the whole HBM
 * budget and hands out pointers but forbids allocations.  nx_dag_checksum above uses a buffer the
 * epoch context cudaMalloc'ed; this variant keeps its per-block partial sums in a fixed 16.25 KiB array
 * that is part of the library image (each GPU that loads the library holds it until the process exits),
 * and calls no cudaMalloc, cudaFree or any allocating library routine.
 *
 *   data_device  device memory of the CURRENT device, 8-byte aligned (e.g. nx_epoch_dag_ptr())
 *   n_bytes      multiple of 8, at most NX_NOALLOC_MAX_BYTES (larger inputs are refused with
 *                NX_ERR_INVALID and a message; use nx_dag_checksum for those); a range that runs past the
 *                end of the allocation data_device points into is refused too, where the driver can tell
 *   stream       cudaStream_t to run on (NULL = default stream); the call returns after it completed
 *   out_host     receives the sum (same value as nx_dag_checksum over the same bytes)
 *
 * The first call on each device looks up the device address of the array (one-time set-up), so make
 * that first call before starting any stream capture on that device (after a cudaDeviceReset the next
 * call repeats the look-up, so the same holds for it).  All calls share the one
 * array and it keeps whatever the previous call left in it; nothing in it needs to be zero at entry
 * (every slot is rewritten before it is read), but two calls that overlap on the same GPU - from two
 * threads or two streams - race on it: serialise them. */
#define NX_NOALLOC_MAX_BYTES (8ull << 30)
int nx_dag_checksum_noalloc(const void* data_device, size_t n_bytes, void* stream, uint64_t* out_host);

/* ------------------------------------------------------------------------------------------------
 * KawPow. */
typedef struct nx_result
{
    uint8_t final_hash[32];
    uint8_t mix_hash[32];
} nx_result; /* layout of ethash_result, ethash.h:48-52 */

typedef struct nx_search_result
{
    int32_t found;
    uint32_t reserved;
    uint64_t nonce;
    uint8_t final_hash[32];
    uint8_t mix_hash[32];
} nx_search_result; /* fields of ethash::search_result, ethash.hpp:49-61 */

/* Compile (NVRTC, sm_100a) and cache the random program of the period block_number/3 for this
 * context's device, so that the first hash/search of a period does not pay for it.  Optional. */
int nx_kawpow_prepare(const nx_epoch* ctx, int block_number);

/* One nonce.  Replaces progpow::hash(ctx, block_number, header_hash, nonce)
 * (crypto/ethash/lib/ethash/progpow.cpp:298-355 and :357-429).  Runs on the interpreted kernel (K6): no NVRTC
 * compilation is ever waited for, 0.3 ms per call; many nonces of one header belong in nx_kawpow_hash_batch / _search. */
int nx_kawpow_hash(const nx_epoch* ctx, int block_number, const uint8_t header_hash[32], uint64_t nonce,
                   nx_result* out);

/* `count` consecutive nonces; out = count x {final_hash[32], mix_hash[32]} in HOST memory
 * (the period's compiled kernel in batch mode; the copy back is part of the call). */
int nx_kawpow_hash_batch(const nx_epoch* ctx, int block_number, const uint8_t header_hash[32],
                         uint64_t start_nonce, uint64_t count, uint8_t* out_host);
/* Same, results left in DEVICE memory (count*64 bytes); returns after the kernel completed. */
int nx_kawpow_hash_batch_device(const nx_epoch* ctx, int block_number, const uint8_t header_hash[32],
                                uint64_t start_nonce, uint64_t count, void* out_device);

/* Lowest nonce in [start_nonce, start_nonce+iterations) whose final hash <= boundary, or found=0 and
 * all fields zero.  Replaces progpow::search / search_light (progpow.cpp:553-579) and the inner
 * nonce loop of miner.cpp:663-690 / rpc/mining.cpp:141-149.  Like the reference's loop it stops at the hit: the
 * kernel's warps look at the lowest nonce reported so far before each 32-nonce unit and skip everything above it, so
 * a call whose range holds an early hit returns in about a millisecond however long the range is.  Any size is
 * efficient: small calls are spread over the whole GPU (65 536 nonces: 0.37 ms). */
int nx_kawpow_search(const nx_epoch* ctx, int block_number, const uint8_t header_hash[32],
                     const uint8_t boundary[32], uint64_t start_nonce, uint64_t iterations,
                     nx_search_result* out);

/* Throughput form of the search for benchmarking/pipelining: enqueue the scan of one batch on the
 * context's stream without synchronising.  nx_kawpow_search_sync() waits and reports the lowest hit
 * of everything enqueued since the last sync.  Batches enqueued between two syncs must share
 * header_hash and block_number and must not start below the first batch (NX_ERR_STATE otherwise).
 * One batch is one kernel launch and holds at most 2^32 nonces (NX_ERR_INVALID beyond); batches queued
 * behind a hit return at once. */
int nx_kawpow_search_enqueue(const nx_epoch* ctx, int block_number, const uint8_t header_hash[32],
                             const uint8_t boundary[32], uint64_t start_nonce, uint64_t count);
int nx_kawpow_search_sync(const nx_epoch* ctx, nx_search_result* out);
/* Device time in ms of the kernels enqueued between the last two syncs (CUDA events on the
 * launching stream) and how many kernels that was. */
int nx_kawpow_last_timing(const nx_epoch* ctx, float* kernel_ms, int* launches);

/* 1 = valid, 0 = invalid, <0 = error.  Replaces progpow::verify (progpow.cpp:431-495): final hash from
 * the claimed mix must be <= boundary, then the recomputed mix must equal the claimed one. */
int nx_kawpow_verify(const nx_epoch* ctx, int block_number, const uint8_t header_hash[32],
                     const uint8_t mix_hash[32], uint64_t nonce, const uint8_t boundary[32]);

/* Final hash from a CLAIMED mix (two keccak-f800, no DAG) — host only.  Replaces
 * progpow::hash_no_verify (progpow.cpp:498-550); block_number is unused there as well. */
void nx_kawpow_hash_no_verify(int block_number, const uint8_t header_hash[32], const uint8_t mix_hash[32],
                              uint64_t nonce, uint8_t out[32]);

/* Many independent records, each with its own header, nonce and height (any mix of ProgPoW periods inside
 * the context's epoch): out_host = n x {final_hash[32], mix_hash[32]}.  Runs the interpreted kernel K6, which
 * needs no per-period compilation - the shape of header validation, where every block is its own period.
 * Replaces progpow::hash called once per header (validation.cpp:11652-11662, rpc/mining.cpp:811,880). */
typedef struct nx_hash_record
{
    int32_t block_number;
    uint32_t pad;
    uint64_t nonce;
    uint8_t header_hash[32];
} nx_hash_record;
int nx_kawpow_hash_many(const nx_epoch* ctx, const nx_hash_record* records, size_t n, uint8_t* out_host);

/* Batched header verification (kernel K6): n independent {block_number(all in ctx's epoch), header,
 * mix, nonce, boundary} records; ok[i] = 1/0.  Replaces the per-header progpow::verify calls of
 * validation.cpp:11652-11662 during header sync.  Records are 112 bytes:
 * {int32 block_number, uint32 pad, uint64 nonce, header[32], mix[32], boundary[32]}. */
typedef struct nx_verify_record
{
    int32_t block_number;
    uint32_t pad;
    uint64_t nonce;
    uint8_t header_hash[32];
    uint8_t mix_hash[32];
    uint8_t boundary[32];
} nx_verify_record;
int nx_kawpow_verify_batch(const nx_epoch* ctx, const nx_verify_record* records, size_t n, uint8_t* ok);

/* ------------------------------------------------------------------------------------------------
 * Classic Ethash (exported by the reference, no node caller; kernel K7).
 * Replaces ethash_hash (crypto/ethash/lib/ethash/ethash.cpp:416-422, kernel :257-303) for one nonce or
 * `count` consecutive nonces (out = count x {final_hash[32], mix_hash[32]}, HOST memory), and
 * ethash_verify_final_hash (:424-429; host only, 1 = final hash <= boundary). */
int nx_ethash_hash(const nx_epoch* ctx, const uint8_t header_hash[32], uint64_t nonce, nx_result* out);
int nx_ethash_hash_batch(const nx_epoch* ctx, const uint8_t header_hash[32], uint64_t start_nonce, uint64_t count,
                         uint8_t* out_host);
int nx_ethash_verify_final_hash(const uint8_t header_hash[32], const uint8_t mix_hash[32], uint64_t nonce,
                                const uint8_t boundary[32]);

/* ------------------------------------------------------------------------------------------------
 * Node glue — host only. */

/* header_hash that the node feeds to progpow: SHA256d over the 80-byte CKAWPOWInput
 * {nVersion,hashPrevBlock,hashMerkleRoot,nTime,nBits,nHeight} (primitives/block.h:213-233,
 * primitives/block.cpp:94-99), byte-reversed by the GetHex()/to_hash256 round trip of hash.cpp:269-270. */
void nx_kawpow_header_hash(int32_t version, const uint8_t prev_block[32], const uint8_t merkle_root[32],
                           uint32_t time, uint32_t bits, uint32_t height, uint8_t out[32]);
void nx_sha256d(const uint8_t* data, size_t size, uint8_t out[32]);

/* Compact nBits -> 32-byte big-endian boundary (arith_uint256::SetCompact, arith_uint256.cpp:209-228).
 * Returns 1 when the target is usable, 0 when negative / zero / overflowing (pow.cpp:190-191). */
int nx_compact_to_boundary(uint32_t bits, uint8_t out[32]);
/* CheckProofOfWork (pow.cpp:182-199) on a final hash in progpow byte order: 1 = passes. */
int nx_check_proof_of_work(const uint8_t final_hash[32], uint32_t bits, const uint8_t pow_limit_be[32]);

/* ------------------------------------------------------------------------------------------------
 * Wire format: the 120-byte block header of KawPow blocks (primitives/block.h:36-74 with
 * nTime >= nKAWPOWActivationTime): nVersion i32 | hashPrevBlock[32] | hashMerkleRoot[32] | nTime u32 | nBits u32 |
 * nHeight u32 | nNonce64 u64 | mix_hash[32], integers little-endian, uint256 fields in storage order (i.e. the
 * byte-reverse of the hash256 strings the progpow functions use, see hash.cpp:269-276). */
typedef struct nx_block_header
{
    int32_t version;
    uint8_t prev_block[32];
    uint8_t merkle_root[32];
    uint32_t time;
    uint32_t bits;
    uint32_t height;
    uint64_t nonce64;
    uint8_t mix_hash[32]; /* uint256 storage order, as on the wire */
} nx_block_header;
void nx_block_header_parse(const uint8_t raw[120], nx_block_header* out);
void nx_block_header_serialize(const nx_block_header* in, uint8_t raw[120]);
/* Block id of a KawPow block = CBlockHeader::GetHash() = KAWPOWHash_OnlyMix (primitives/block.cpp:38-55,
 * hash.cpp:280-291): final hash from the header's own mix_hash, returned in uint256 storage order. Host only. */
void nx_block_hash(const uint8_t raw[120], uint8_t out_uint256[32]);

/* PoW part of CheckBlockHeader (validation.cpp:11638-11665) for n wire headers of ONE epoch, in one K6 launch.
 *   mix_only = 1: the below-checkpoint branch (:11641-11649): CheckProofOfWork(GetHash()) only, no GPU work;
 *   mix_only = 0: CheckProofOfWork(GetHashFull(mix)) with the RECOMPUTED final hash, then mix == header.mix_hash.
 * status[i]: */
enum
{
    NX_HEADER_OK = 0,
    NX_HEADER_HIGH_HASH = 1,        /* "high-hash": proof of work failed */
    NX_HEADER_INVALID_MIX = 2,      /* "invalid-mix-hash" */
    NX_HEADER_BAD_BITS = 3,         /* nBits negative / zero / overflow / above pow_limit (pow.cpp:190-191) */
    NX_HEADER_WRONG_EPOCH = 4       /* nHeight is not in the context's epoch: not checked */
};
int nx_check_block_headers(const nx_epoch* ctx, const uint8_t* raw_headers /* n x 120 */, size_t n,
                           const uint8_t pow_limit_be[32], int mix_only, uint8_t* status);

/* Block-index records as the node keeps them in LevelDB: CDiskBlockIndex::SerializationOp (chain.h:398-431) over the
 * VARINT coding of serialize.h:290-366.  Record = VARINT client version | VARINT nHeight | VARINT nStatus | VARINT nTx |
 * [VARINT nFile if status has DATA or UNDO] | [VARINT nDataPos if DATA] | [VARINT nUndoPos if UNDO] | nVersion i32 |
 * hashPrev[32] | hashMerkleRoot[32] | nTime u32 | nBits u32 | then nNonce u32 for blocks before the KawPow activation
 * time, else nNonce64 u64 | mix_hash[32].  (nHeight is NOT repeated in the header part.)  Host only. */
#define NX_BLOCK_HAVE_DATA 8u  /* chain.h:156 */
#define NX_BLOCK_HAVE_UNDO 16u /* chain.h:157 */
typedef struct nx_disk_block_index
{
    int32_t client_version;
    int32_t height;
    uint32_t status;
    uint32_t tx_count;
    int32_t file;
    uint32_t data_pos;
    uint32_t undo_pos;
    int32_t version;
    uint8_t prev_block[32]; /* uint256 storage order */
    uint8_t merkle_root[32];
    uint32_t time;
    uint32_t bits;
    uint32_t nonce;       /* blocks before the activation time */
    uint64_t nonce64;     /* KawPow blocks */
    uint8_t mix_hash[32]; /* KawPow blocks, uint256 storage order */
} nx_disk_block_index;
/* WriteVarInt / ReadVarInt<uint64_t>: bytes written (1..10) / bytes consumed, 0 for a truncated or oversized number. */
size_t nx_varint_write(uint64_t n, uint8_t out[10]);
size_t nx_varint_read(const uint8_t* in, size_t len, uint64_t* n);
/* Bytes written, 0 when `cap` is too small or a signed field is negative. */
size_t nx_disk_block_index_serialize(const nx_disk_block_index* in, uint32_t kawpow_activation_time, uint8_t* out, size_t cap);
/* Bytes consumed, 0 for a malformed record (truncated, or a VARINT beyond its field's type - where ReadVarInt throws). */
size_t nx_disk_block_index_parse(const uint8_t* in, size_t len, uint32_t kawpow_activation_time, nx_disk_block_index* out);
/* The check LoadBlockIndexGuts makes on every record at start-up (txdb.cpp:470-492): rebuild the header, take
 * GetBlockHash() - for a KawPow block the mix-only hash - and CheckProofOfWork it against nBits.  Returns 1 / 0, or -1 for
 * a block before the activation time (X16R: not this library's).  block_id (may be NULL) receives GetBlockHash(). */
int nx_disk_block_index_check(const nx_disk_block_index* rec, uint32_t kawpow_activation_time, const uint8_t pow_limit_be[32],
                              uint8_t block_id_uint256[32]);

/* ------------------------------------------------------------------------------------------------
 * Miner: one process, all GPUs.  Replaces the nonce loops of miner.cpp:656-713 (CloreMiner) and
 * rpc/mining.cpp:117-173 (generateBlocks), and the epoch-context handling of hash.cpp:260-266 /
 * crypto/ethash/lib/ethash/managed.cpp.  Devices claim nonce super-batches in ascending order as they become
 * free; the result is the LOWEST qualifying nonce of the range (host-side min reduction, progpow.cpp:572-577).
 * Threading: one miner serves one mining loop - do not overlap nx_miner_search / _mine_header / _set_epoch calls on the
 * same miner; nx_miner_prepare_async may run beside them (that is its purpose), and `cancel` may be set from any thread. */
typedef struct nx_miner nx_miner;
int nx_miner_create(const int* devices, int n_devices, nx_miner** out); /* devices NULL / n 0 = all visible GPUs */
void nx_miner_destroy(nx_miner* m);
int nx_miner_device_count(const nx_miner* m);
int nx_miner_set_batch(nx_miner* m, uint64_t nonces_per_batch);          /* default 2^22 */
int nx_miner_set_epoch(nx_miner* m, int epoch);   /* light cache once on the host, one DAG per device, in parallel */
/* Background: compile the next period's program; prebuild the next epoch when it starts within
 * lookahead_blocks.  Returns immediately; nx_miner_wait_prepared() joins and reports the outcome. */
int nx_miner_prepare_async(nx_miner* m, int block_number, int lookahead_blocks);
int nx_miner_wait_prepared(nx_miner* m);
int nx_miner_epoch(const nx_miner* m);
int nx_miner_prepared_epoch(const nx_miner* m);
const nx_epoch* nx_miner_context(const nx_miner* m, int index);          /* borrowed: the index-th device's context */
int nx_miner_search(nx_miner* m, int block_number, const uint8_t header_hash[32], const uint8_t boundary[32],
                    uint64_t start_nonce, uint64_t iterations, const volatile int* cancel, nx_search_result* out);
int nx_miner_last_stats(const nx_miner* m, uint64_t* hashes, double* seconds);
/* Devices that failed during the last search; their windows were re-issued to the surviving devices (the call
 * only fails when no device is left). */
int nx_miner_last_failed_devices(const nx_miner* m);
/* Header fields in, nBits decoded, nNonce64 scanned: the body of miner.cpp:656-690 in one call. */
int nx_miner_mine_header(nx_miner* m, int32_t version, const uint8_t prev_block[32], const uint8_t merkle_root[32],
                         uint32_t time, uint32_t bits, uint32_t height, uint64_t start_nonce, uint64_t max_tries,
                         const volatile int* cancel, nx_search_result* out);

#ifdef __cplusplus
}
#endif
#endif /* NX_KAWPOW_H */
