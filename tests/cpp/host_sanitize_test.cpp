// Host-side product code under AddressSanitizer + UndefinedBehaviorSanitizer (SURVEY.md §5 "race detection /
// sanitizers").  Built by tests/test_host_sanitizers.py from the product's own host sources with
// -fsanitize=address,undefined -fno-sanitize-recover; prints "key=value" lines that the test compares with the
// regular (unsanitized) library and with the oracle.  Touches no device.
#include "nx_internal.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace nx;

static std::string hex(const uint8_t* p, size_t n)
{
    static const char* d = "0123456789abcdef";
    std::string s;
    for (size_t i = 0; i < n; ++i) {
        s += d[p[i] >> 4];
        s += d[p[i] & 15];
    }
    return s;
}
static std::string digest(const void* p, size_t n)
{
    uint8_t out[32];
    nx_sha256d(static_cast<const uint8_t*>(p), n, out);
    return hex(out, 32);
}

int main()
{
    // Keccak sponges around every block-boundary length of both rates (136 and 72 bytes)
    std::vector<uint8_t> msg(300);
    for (size_t i = 0; i < msg.size(); ++i) msg[i] = (uint8_t)(i * 7 + 1);
    for (size_t len : {0, 1, 31, 32, 64, 71, 72, 73, 135, 136, 137, 144, 200, 272, 300}) {
        uint8_t h256[32], h512[64];
        // exact-size heap copies: an over-read of the message by even one byte trips ASan
        std::vector<uint8_t> m(msg.begin(), msg.begin() + (long)len);
        nx_keccak256(m.data(), len, h256);
        nx_keccak512(m.data(), len, h512);
        printf("keccak256_%zu=%s\nkeccak512_%zu=%s\n", len, hex(h256, 32).c_str(), len, hex(h512, 64).c_str());
    }
    for (int e : {0, 1, 2, 22, 448}) {
        uint8_t s[32];
        nx_epoch_seed(e, s);
        printf("seed_%d=%s\nlight_items_%d=%d\ndataset_items_%d=%d\n", e, hex(s, 32).c_str(), e, nx_light_cache_num_items(e), e,
               nx_full_dataset_num_items(e));
        if (e <= 22) printf("find_epoch_%d=%d\n", e, nx_find_epoch_number(s));
    }
    for (int ub : {-5, 0, 1, 2, 3, 4, 100, 262144}) printf("prime_%d=%d\n", ub, nx_find_largest_prime(ub));

    // the serial light-cache chain on a small prime item count, exact-size heap buffer; scalar and (if present) AVX-512
    {
        const int n = 1021;
        uint8_t seed[32];
        nx_epoch_seed(3, seed);
        std::vector<uint8_t> a((size_t)n * 64), b((size_t)n * 64);
        host_build_light_cache_scalar(a.data(), n, seed);
        printf("light_scalar_1021=%s\n", digest(a.data(), a.size()).c_str());
        if (cpu_has_avx512f()) {
            host_build_light_cache_avx512(b.data(), n, seed);
            printf("light_avx512_1021=%s\n", digest(b.data(), b.size()).c_str());
        }
        // on-disk cache: store, load, reject a corrupted file (NX_CACHE_DIR is set by the test)
        if (std::getenv("NX_CACHE_DIR")) {
            host_store_light_cache(a.data(), n, 77777);
            std::vector<uint8_t> c((size_t)n * 64);
            printf("cache_load=%d\n", (int)host_load_light_cache(c.data(), n, 77777));
            printf("cache_same=%d\n", (int)(c == a));
            printf("cache_wrong_size=%d\n", (int)host_load_light_cache(c.data(), n - 2, 77777));
            const std::string path = std::string(std::getenv("NX_CACHE_DIR")) + "/light_cache_epoch77777.nxlc";
            FILE* f = std::fopen(path.c_str(), "r+b");
            if (f) {
                std::fseek(f, 1000, SEEK_SET);
                std::fputc(0x5a ^ a[1000 - 24], f);
                std::fclose(f);
            }
            printf("cache_corrupt=%d\n", (int)host_load_light_cache(c.data(), n, 77777));
        }
    }

    // period program generator: op lists and generated text for a spread of seeds (0, a mainnet-sized one, > 2^32)
    for (uint64_t seed : {0ull, 1ull, 2499ull, 1120000ull, (1ull << 40) + 3}) {
        const std::vector<ProgramOp> ops = kawpow_program(seed);
        std::vector<uint32_t> flat;
        for (const ProgramOp& o : ops) {
            flat.push_back((uint32_t)o.kind);
            flat.push_back(o.src1);
            flat.push_back(o.src2);
            flat.push_back(o.dst);
            flat.push_back(o.sel_math);
            flat.push_back(o.sel_merge);
        }
        const std::string text = kawpow_program_text(seed);
        const std::string full = kawpow_source(seed, 18874358u);
        printf("program_ops_%llu=%zu:%s\nprogram_text_%llu=%zu:%s\nprogram_full_%llu=%zu\n", (unsigned long long)seed, ops.size(),
               digest(flat.data(), flat.size() * 4).c_str(), (unsigned long long)seed, text.size(),
               digest(text.data(), text.size()).c_str(), (unsigned long long)seed, full.size());
        uint32_t packed[36 * 4];
        (void)packed;
    }

    // node glue
    for (uint32_t bits : {0u, 0x00123456u, 0x01003456u, 0x01123456u, 0x02008000u, 0x03123456u, 0x04123456u, 0x04923456u, 0x1b0404cbu,
                          0x1f00ffffu, 0x2000ffffu, 0x207fffffu, 0x20123456u, 0x21123456u, 0x22123456u, 0x23000001u, 0xff123456u}) {
        uint8_t t[32];
        const int ok = nx_compact_to_boundary(bits, t);
        printf("compact_%08x=%d:%s\n", bits, ok, hex(t, 32).c_str());
    }
    {
        nx_block_header h;
        std::memset(&h, 0, sizeof h);
        h.version = 0x30000000;
        for (int i = 0; i < 32; ++i) {
            h.prev_block[i] = (uint8_t)i;
            h.merkle_root[i] = (uint8_t)(255 - i);
            h.mix_hash[i] = (uint8_t)(i * 3);
        }
        h.time = 1651444217u;
        h.bits = 0x1b0404cbu;
        h.height = 3360000u;
        h.nonce64 = 0x123456789abcdef0ull;
        std::vector<uint8_t> raw(120);  // exact size on the heap
        nx_block_header_serialize(&h, raw.data());
        nx_block_header back;
        nx_block_header_parse(raw.data(), &back);
        uint8_t hh[32], id[32], fin[32];
        nx_kawpow_header_hash(h.version, h.prev_block, h.merkle_root, h.time, h.bits, h.height, hh);
        nx_block_hash(raw.data(), id);
        nx_kawpow_hash_no_verify((int)h.height, hh, h.mix_hash, h.nonce64, fin);
        printf("wire=%s\nroundtrip=%d\nheader_hash=%s\nblock_id=%s\nno_verify=%s\n", hex(raw.data(), 120).c_str(),
               (int)(back.nonce64 == h.nonce64 && back.height == h.height && std::memcmp(back.mix_hash, h.mix_hash, 32) == 0),
               hex(hh, 32).c_str(), hex(id, 32).c_str(), hex(fin, 32).c_str());
        printf("pow_check=%d\n", nx_check_proof_of_work(fin, 0x207fffffu, nullptr));
        printf("ethash_final=%d\n", nx_ethash_verify_final_hash(hh, h.mix_hash, 5, fin));
    }
    // block-index records: parse exact-size heap copies of random bytes, of a valid record cut at every length, and of
    // that record with every single bit of its first 16 bytes flipped - any read past the buffer trips ASan
    {
        nx_disk_block_index rec;
        std::memset(&rec, 0, sizeof rec);
        rec.client_version = 4040402;
        rec.height = 3360000;
        rec.status = NX_BLOCK_HAVE_DATA | NX_BLOCK_HAVE_UNDO | 5;
        rec.tx_count = 77;
        rec.file = 12;
        rec.data_pos = 0xfedcba98u;
        rec.undo_pos = 123456;
        rec.version = 0x30000000;
        rec.time = 1700000000u;
        rec.bits = 0x1b0404cbu;
        rec.nonce64 = 0x0123456789abcdefull;
        for (int i = 0; i < 32; ++i) rec.prev_block[i] = rec.merkle_root[i] = rec.mix_hash[i] = (uint8_t)(i + 1);
        uint8_t full[256];
        const size_t n = nx_disk_block_index_serialize(&rec, 1651444217u, full, sizeof full);
        printf("disk_record=%zu:%s\n", n, hex(full, n).c_str());
        unsigned long long accepted = 0, state = 88172645463325252ull;
        auto rnd = [&state]() {
            state ^= state << 13;
            state ^= state >> 7;
            state ^= state << 17;
            return state;
        };
        nx_disk_block_index out;
        for (size_t cut = 0; cut <= n; ++cut) {
            std::vector<uint8_t> b(full, full + cut);
            accepted += nx_disk_block_index_parse(b.data(), b.size(), 1651444217u, &out) != 0;
            uint8_t small[10];
            if (nx_disk_block_index_serialize(&rec, 1651444217u, small, cut < sizeof small ? cut : sizeof small) != 0) return 3;
        }
        for (int bit = 0; bit < 128; ++bit) {
            std::vector<uint8_t> b(full, full + n);
            b[(size_t)bit / 8] ^= (uint8_t)(1u << (bit % 8));
            accepted += nx_disk_block_index_parse(b.data(), b.size(), 1651444217u, &out) != 0;
        }
        for (int i = 0; i < 20000; ++i) {
            std::vector<uint8_t> b(rnd() % 141);
            for (auto& x : b) x = (uint8_t)rnd();
            if (!b.empty() && (i & 1)) b[0] = (uint8_t)(b[0] & 0x7f);  // half of them start with a one-byte version
            accepted += nx_disk_block_index_parse(b.data(), b.size(), (uint32_t)rnd(), &out) != 0;
            uint64_t v;
            accepted += nx_varint_read(b.data(), b.size(), &v) != 0;
        }
        printf("disk_fuzz_accepted=%llu\n", accepted);
        uint8_t id[32];
        printf("disk_check=%d\n", nx_disk_block_index_check(&rec, 1651444217u, nullptr, id));
    }
    puts("done=1");
    return 0;
}
