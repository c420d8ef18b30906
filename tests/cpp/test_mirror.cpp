// The reference's own unit tests for the path, replayed in C++ against gpu_ec.hpp (which sits on the
// C ABI of libgpu_ec.so): dfs/common/src/erasure.rs:61-109, dfs/chunkserver/src/chunkserver.rs:1099-1198,
// dfs/client/src/mod.rs:1541-1557, 1637-1664. Exit code 0 and "ok" = all passed.
#include <cstdio>
#include <cstdlib>
#include <string>

#include "../../rust-hadoop-generated-by-llm_b200/cpp/gpu_ec.hpp"

using namespace gpu_ec;
static int g_checks = 0;
#define REQUIRE(c)                                                                  \
    do {                                                                            \
        g_checks++;                                                                 \
        if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } \
    } while (0)

static Bytes bytes_of(const char *s) { return Bytes(s, s + std::strlen(s)); }
static OptShards some(const std::vector<Bytes> &v) { OptShards o; for (auto &b : v) o.emplace_back(b); return o; }

int main(int argc, char **argv) {
    const std::string tmp = argc > 1 ? argv[1] : "/tmp";
    // erasure.rs:66-73 test_encode_decode_roundtrip
    {
        const Bytes data = bytes_of("Hello, Erasure Coding World!");
        auto shards = erasure::encode(data, 4, 2);
        REQUIRE(shards.size() == 6);
        REQUIRE(shards[4] == (Bytes{0xa9, 0x35, 0xea, 0x10, 0x3a, 0x13, 0x81}));  // golden parity (tests/golden)
        auto opt = some(shards);
        REQUIRE(erasure::decode(opt, 4, 2, data.size()) == data);
        // erasure.rs:75-84 test_decode_with_missing_shards
        opt = some(shards);
        opt[1].reset();
        opt[4].reset();
        REQUIRE(erasure::decode(opt, 4, 2, data.size()) == data);
        REQUIRE(*opt[1] == shards[1] && *opt[4] == shards[4]);
    }
    // erasure.rs:86-94 test_encode_large_data
    {
        Bytes data(10000);
        for (size_t i = 0; i < data.size(); i++) data[i] = static_cast<uint8_t>(i % 256);
        auto shards = erasure::encode(data, 4, 2);
        REQUIRE(shards.size() == 6);
        auto opt = some(shards);
        REQUIRE(erasure::decode(opt, 4, 2, data.size()) == data);
    }
    // erasure.rs:96-101 test_shard_len ; client/src/mod.rs:1553-1556
    REQUIRE(erasure::shard_len(28, 4) == 7 && erasure::shard_len(10000, 4) == 2500 && erasure::shard_len(1, 4) == 1);
    REQUIRE(erasure::shard_len(1000, 4) == 250 && erasure::shard_len(1001, 4) == 251 && erasure::shard_len(4, 4) == 1);
    // erasure.rs:103-108 test_encode_empty_data_returns_error (+ the two bail!() texts)
    {
        bool threw = false;
        try { erasure::encode(Bytes{}, 4, 2); } catch (const Error &e) { threw = std::string(e.what()) == "data must not be empty"; }
        REQUIRE(threw);
        threw = false;
        try { erasure::encode(bytes_of("x"), 0, 2); } catch (const Error &e) { threw = std::string(e.what()) == "data_shards and parity_shards must both be > 0"; }
        REQUIRE(threw);
    }
    // chunkserver.rs:1099-1115 test_ec_reconstruct_shard_logic
    {
        const Bytes data(4096, 7);
        auto shards = erasure::encode(data, 4, 2);
        auto opt = some(shards);
        opt[4].reset();
        erasure::reconstruct(opt, 4, 2);
        REQUIRE(*opt[4] == shards[4]);
    }
    // repair path: only the wanted shard is rebuilt, the other missing slot stays empty
    {
        Bytes data(100003);
        for (size_t i = 0; i < data.size(); i++) data[i] = static_cast<uint8_t>(i * 7 + 3);
        auto shards = erasure::encode(data, 6, 3);
        auto opt = some(shards);
        opt[1].reset();
        opt[7].reset();
        erasure::reconstruct_only(opt, 6, 3, {7});
        REQUIRE(opt[7] && *opt[7] == shards[7] && !opt[1]);
    }
    // client/src/mod.rs:1541-1551: 1024 B at RS(4,2) -> 6 shards of 256 B
    {
        auto shards = erasure::encode(Bytes(1024, 0), 4, 2);
        REQUIRE(shards.size() == 6);
        for (auto &s : shards) REQUIRE(s.size() == 256);
    }
    // client/src/mod.rs:1637-1664: fast-path concat and degraded decode
    {
        const Bytes d(5000, 42);
        auto sh = erasure::encode(d, 4, 2);
        Bytes cat;
        for (int i = 0; i < 4; i++) cat.insert(cat.end(), sh[i].begin(), sh[i].end());
        cat.resize(5000);
        REQUIRE(cat == d);
        const Bytes e(3333, 99);
        auto sh2 = erasure::encode(e, 4, 2);
        auto opt = some(sh2);
        opt[1].reset();
        opt[3].reset();
        REQUIRE(erasure::decode(opt, 4, 2, 3333) == e);
    }
    // client write path in one call + CRC check values
    {
        REQUIRE(crc32(bytes_of("123456789")) == 0xCBF43926u);
        const Bytes data = bytes_of("Hello, Erasure Coding World!");
        auto blk = erasure::encode_with_checksums(data, 4, 2);
        REQUIRE(blk.full_checksum == crc32(data));
        for (size_t i = 0; i < blk.shards.size(); i++) REQUIRE(blk.shard_checksums[i] == crc32(blk.shards[i]));
    }
    // chunkserver.rs:1117-1140 test_checksum_verification
    {
        chunkserver::BlockStore server(tmp);
        const Bytes data = bytes_of("Hello, world! This is a test block for checksum verification.");
        server.write_block("test_block", data);
        REQUIRE(!server.verify_block("test_block", data));
        REQUIRE(chunkserver::BlockStore::read_file(std::filesystem::path(tmp) / "test_block.meta") == (Bytes{0x3c, 0x7e, 0x92, 0xbb}));
        Bytes corrupted = data;
        std::memcpy(corrupted.data(), "Corrupted", 9);
        auto err = server.verify_block("test_block", corrupted);
        REQUIRE(err && *err == "Checksum mismatch at chunk 0");
        REQUIRE(*server.verify_block("no_such_block", data) == "Checksum file missing");
        REQUIRE(!chunkserver::check_expected(data, 0));  // 0 = skip (chunkserver.rs:746)
        REQUIRE(chunkserver::check_expected(data, 1)->rfind("Checksum mismatch: expected 1, actual ", 0) == 0);
    }
    // chunkserver.rs:1146-1198: 64 KiB (i % 256) block, .meta non-empty, partial-range verify; then the scrubber
    {
        chunkserver::BlockStore server(tmp);
        Bytes data(64 * 1024);
        for (size_t i = 0; i < data.size(); i++) data[i] = static_cast<uint8_t>(i % 256);
        server.write_block("block_64k", data);
        REQUIRE(std::filesystem::file_size(std::filesystem::path(tmp) / "block_64k.meta") == 4 * 128);
        REQUIRE(!server.verify_partial_read("block_64k", 32768, 4096));
        REQUIRE(server.scrub().empty());
        Bytes raw = data;
        raw[33000] ^= 1;
        chunkserver::BlockStore::write_file(std::filesystem::path(tmp) / "block_64k", raw);
        auto err = server.verify_partial_read("block_64k", 32768, 4096);
        REQUIRE(err && *err == "Checksum mismatch at chunk 64");
        REQUIRE(!server.verify_partial_read("block_64k", 0, 32768));
        auto bad = server.scrub();
        REQUIRE(bad.size() == 1 && bad[0] == "block_64k");
    }
    std::printf("ok %d checks, %llu kernel launches\n", g_checks, static_cast<unsigned long long>(gpu_ec_kernel_launch_count()));
    return 0;
}
