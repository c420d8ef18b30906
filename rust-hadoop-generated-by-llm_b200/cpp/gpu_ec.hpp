// gpu_ec.hpp — header-only C++17 host side above the C ABI of libgpu_ec.so, mirroring the reference's
// Rust interface for the byte path name for name (same argument meaning, same error text):
//
//   gpu_ec::erasure::{shard_len, encode, decode}          dfs/common/src/erasure.rs:56, 7, 30
//   gpu_ec::erasure::reconstruct                          ReedSolomon::reconstruct, chunkserver.rs:620-623
//   gpu_ec::erasure::encode_with_checksums                client/src/mod.rs:600-611 (encode + k+m CRCs + CRC(original))
//   gpu_ec::chunkserver::{calculate_checksums, BlockStore::{write_block, verify_block,
//       verify_partial_read, scrub}}                      dfs/chunkserver/src/chunkserver.rs:182-351, 642-718
//   gpu_ec::crc32                                         crc32fast::Hasher::new/update/finalize
//
// anyhow::Result<T> becomes "returns T or throws gpu_ec::Error(what() = the reference's message)";
// Result<(), String> becomes std::optional<std::string> (nullopt = Ok). All compute runs in the CUDA
// kernels behind the C ABI; there is no CPU fallback here either.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/gpu_ec.h"

namespace gpu_ec {

struct Error : std::runtime_error {
    int status;
    Error(int st, const std::string &msg) : std::runtime_error(msg), status(st) {}
};

inline void check(int st) {
    if (st == GPU_EC_OK) return;
    std::string msg = gpu_ec_strerror(st);
    if (st == GPU_EC_ERR_CUDA || st == GPU_EC_ERR_NO_DEVICE) {
        const char *d = gpu_ec_last_error();
        if (d && *d) msg += std::string(": ") + d;
    }
    throw Error(st, msg);
}

using Bytes = std::vector<uint8_t>;
using OptShards = std::vector<std::optional<Bytes>>;

/// crc32fast::Hasher::new(); update(data); finalize()
inline uint32_t crc32(const uint8_t *data, size_t len) {
    uint32_t out = 0;
    check(gpu_ec_crc32(data, len, GPU_EC_POLY_IEEE, &out));
    return out;
}
inline uint32_t crc32(const Bytes &b) { return crc32(b.data(), b.size()); }

namespace erasure {

/// dfs/common/src/erasure.rs:56 — panics (here: std::logic_error) when data_shards == 0
inline size_t shard_len(size_t data_len, size_t data_shards) {
    if (data_shards == 0) throw std::logic_error("data_shards must be > 0");
    return gpu_ec_shard_len(data_len, data_shards);
}

/// dfs/common/src/erasure.rs:7
inline std::vector<Bytes> encode(const uint8_t *data, size_t len, size_t data_shards, size_t parity_shards) {
    if (data_shards == 0 || parity_shards == 0) throw Error(GPU_EC_ERR_ZERO_SHARDS, gpu_ec_strerror(GPU_EC_ERR_ZERO_SHARDS));
    if (len == 0) throw Error(GPU_EC_ERR_EMPTY_DATA, gpu_ec_strerror(GPU_EC_ERR_EMPTY_DATA));
    if (data_shards + parity_shards > 256) throw Error(GPU_EC_ERR_TOO_MANY_SHARDS, gpu_ec_strerror(GPU_EC_ERR_TOO_MANY_SHARDS));
    const size_t s = shard_len(len, data_shards);
    std::vector<Bytes> shards(data_shards + parity_shards, Bytes(s));
    std::vector<uint8_t *> ptrs;
    for (auto &v : shards) ptrs.push_back(v.data());
    check(gpu_ec_encode(data, len, static_cast<int>(data_shards), static_cast<int>(parity_shards), ptrs.data(), nullptr, nullptr, nullptr));
    return shards;
}
inline std::vector<Bytes> encode(const Bytes &data, size_t k, size_t m) { return encode(data.data(), data.size(), k, m); }

struct EncodedBlock {
    std::vector<Bytes> shards;
    std::vector<uint32_t> shard_checksums;  // WriteBlockRequest.expected_checksum_crc32c per shard
    uint32_t full_checksum = 0;              // BlockChecksumInfo.checksum_crc32c
};
/// client/src/mod.rs:600-611 in one call
inline EncodedBlock encode_with_checksums(const Bytes &data, size_t k, size_t m) {
    if (k == 0 || m == 0) throw Error(GPU_EC_ERR_ZERO_SHARDS, gpu_ec_strerror(GPU_EC_ERR_ZERO_SHARDS));
    if (data.empty()) throw Error(GPU_EC_ERR_EMPTY_DATA, gpu_ec_strerror(GPU_EC_ERR_EMPTY_DATA));
    EncodedBlock b;
    const size_t s = shard_len(data.size(), k);
    b.shards.assign(k + m, Bytes(s));
    b.shard_checksums.assign(k + m, 0);
    std::vector<uint8_t *> ptrs;
    for (auto &v : b.shards) ptrs.push_back(v.data());
    check(gpu_ec_encode(data.data(), data.size(), static_cast<int>(k), static_cast<int>(m), ptrs.data(), b.shard_checksums.data(),
                        &b.full_checksum, nullptr));
    return b;
}

/// ReedSolomon::new(k,m)?.reconstruct(&mut shards): every nullopt slot, data or parity, is rebuilt in place
inline void reconstruct(OptShards &shards, size_t k, size_t m) {
    if (k == 0 || m == 0) throw Error(GPU_EC_ERR_ZERO_SHARDS, gpu_ec_strerror(GPU_EC_ERR_ZERO_SHARDS));
    if (shards.size() != k + m)
        throw Error(GPU_EC_ERR_INVALID_ARG, shards.size() < k + m ? "RS reconstruct error: TooFewShards" : "RS reconstruct error: TooManyShards");
    size_t s = 0;
    bool have = false;
    for (auto &o : shards)
        if (o) {
            if (have && o->size() != s) throw Error(GPU_EC_ERR_SHARD_SIZE, gpu_ec_strerror(GPU_EC_ERR_SHARD_SIZE));
            s = o->size();
            have = true;
        }
    std::vector<uint8_t> present(shards.size());
    std::vector<Bytes> fresh(shards.size());
    std::vector<uint8_t *> ptrs(shards.size());
    for (size_t i = 0; i < shards.size(); i++) {
        present[i] = shards[i] ? 1 : 0;
        if (shards[i]) ptrs[i] = shards[i]->data();
        else { fresh[i].assign(s, 0); ptrs[i] = fresh[i].data(); }
    }
    check(gpu_ec_reconstruct(ptrs.data(), present.data(), s, static_cast<int>(k), static_cast<int>(m)));
    for (size_t i = 0; i < shards.size(); i++)
        if (!shards[i]) shards[i] = std::move(fresh[i]);
}

/// the repair path's variant: rebuild only the listed missing slots (chunkserver.rs:626-631 keeps one)
inline void reconstruct_only(OptShards &shards, size_t k, size_t m, const std::vector<int> &want) {
    if (k == 0 || m == 0) throw Error(GPU_EC_ERR_ZERO_SHARDS, gpu_ec_strerror(GPU_EC_ERR_ZERO_SHARDS));
    if (shards.size() != k + m) throw Error(GPU_EC_ERR_INVALID_ARG, "RS reconstruct error: TooFewShards");
    size_t s = 0;
    for (auto &o : shards) if (o) s = o->size();
    std::vector<uint8_t> present(shards.size());
    std::vector<Bytes> fresh(shards.size());
    std::vector<uint8_t *> ptrs(shards.size());
    for (size_t i = 0; i < shards.size(); i++) {
        present[i] = shards[i] ? 1 : 0;
        if (shards[i]) ptrs[i] = shards[i]->data();
        else { fresh[i].assign(s, 0); ptrs[i] = fresh[i].data(); }
    }
    check(gpu_ec_reconstruct_only(ptrs.data(), present.data(), s, static_cast<int>(k), static_cast<int>(m), want.data(),
                                  static_cast<int>(want.size())));
    for (int i : want) shards[static_cast<size_t>(i)] = std::move(fresh[static_cast<size_t>(i)]);
}

/// dfs/common/src/erasure.rs:30
inline Bytes decode(OptShards &shards, size_t k, size_t m, size_t original_len) {
    reconstruct(shards, k, m);
    Bytes result;
    for (size_t i = 0; i < k; i++) {
        if (!shards[i]) throw Error(GPU_EC_ERR_INVALID_ARG, "data shard " + std::to_string(i) + " missing after reconstruction");
        result.insert(result.end(), shards[i]->begin(), shards[i]->end());
    }
    if (result.size() > original_len) result.resize(original_len);  // Vec::truncate never grows
    return result;
}

}  // namespace erasure

namespace chunkserver {

constexpr size_t CHECKSUM_CHUNK_SIZE = 512;  // chunkserver.rs:16

/// MyChunkServer::calculate_checksums (chunkserver.rs:182-190)
inline std::vector<uint32_t> calculate_checksums(const uint8_t *data, size_t len) {
    std::vector<uint32_t> out((len + CHECKSUM_CHUNK_SIZE - 1) / CHECKSUM_CHUNK_SIZE);
    check(gpu_ec_crc32_chunks(data, len, CHECKSUM_CHUNK_SIZE, GPU_EC_POLY_IEEE, out.data()));
    return out;
}
inline std::vector<uint32_t> calculate_checksums(const Bytes &b) { return calculate_checksums(b.data(), b.size()); }

/// chunkserver.rs:196
inline Bytes meta_bytes(const std::vector<uint32_t> &crcs) {
    Bytes out(4 * crcs.size());
    gpu_ec_meta_encode_be(crcs.data(), crcs.size(), out.data());
    return out;
}

/// the compare of verify_block (chunkserver.rs:265-289): nullopt = Ok(()), else the Err string
inline std::optional<std::string> verify_chunks(const uint8_t *data, size_t len, const Bytes &meta) {
    int64_t bad = -1;
    const int st = gpu_ec_verify_chunks(data, len, CHECKSUM_CHUNK_SIZE, GPU_EC_POLY_IEEE, meta.data(), meta.size(), &bad);
    if (st == GPU_EC_ERR_COUNT_MISMATCH) return std::string("Checksum count mismatch");
    if (st == GPU_EC_ERR_CRC_MISMATCH) return "Checksum mismatch at chunk " + std::to_string(bad);
    check(st);
    return std::nullopt;
}

/// WriteBlock's in-flight check (chunkserver.rs:746-766); expected == 0 means "not provided"
inline std::optional<std::string> check_expected(const Bytes &data, uint32_t expected) {
    if (expected == 0) return std::nullopt;
    const uint32_t actual = crc32(data);
    if (actual != expected) return "Checksum mismatch: expected " + std::to_string(expected) + ", actual " + std::to_string(actual);
    return std::nullopt;
}

/// the part of MyChunkServer that owns <dir>/<block_id> and <dir>/<block_id>.meta
class BlockStore {
public:
    explicit BlockStore(std::filesystem::path storage_dir, std::optional<std::filesystem::path> cold = std::nullopt)
        : dir_(std::move(storage_dir)), cold_(std::move(cold)) {}

    /// write_block_async (chunkserver.rs:192-209)
    void write_block(const std::string &block_id, const Bytes &data) const {
        const Bytes meta = meta_bytes(calculate_checksums(data));
        write_file(dir_ / block_id, data);
        write_file(dir_ / (block_id + ".meta"), meta);
    }
    /// verify_block (chunkserver.rs:238-292)
    std::optional<std::string> verify_block(const std::string &block_id, const Bytes &data) const {
        std::filesystem::path mp = dir_ / (block_id + ".meta");
        if (!std::filesystem::exists(mp) && cold_ && std::filesystem::exists(*cold_ / (block_id + ".meta"))) mp = *cold_ / (block_id + ".meta");
        if (!std::filesystem::exists(mp)) return std::string("Checksum file missing");
        return verify_chunks(data.data(), data.size(), read_file(mp));
    }
    /// verify_partial_read (chunkserver.rs:296-351): only the chunks the range touches, hot dir only
    std::optional<std::string> verify_partial_read(const std::string &block_id, uint64_t offset, uint64_t length) const {
        const std::filesystem::path mp = dir_ / (block_id + ".meta");
        if (!std::filesystem::exists(mp)) return std::string("Checksum file missing");
        const Bytes meta = read_file(mp);
        const size_t n_all = meta.size() / 4;
        if (n_all == 0) return std::nullopt;
        const size_t start = offset / CHECKSUM_CHUNK_SIZE;
        size_t end = (offset + length == 0 ? 0 : offset + length - 1) / CHECKSUM_CHUNK_SIZE;
        end = std::min(end, n_all - 1);
        if (end < start) return std::nullopt;
        const Bytes file = read_file(dir_ / block_id);
        const size_t lo = start * CHECKSUM_CHUNK_SIZE, hi = std::min((end + 1) * CHECKSUM_CHUNK_SIZE, file.size());
        const std::vector<uint32_t> got = calculate_checksums(file.data() + lo, hi - lo);
        for (size_t i = 0; i < got.size(); i++) {
            const uint8_t *e = meta.data() + 4 * (start + i);
            const uint32_t want = (uint32_t(e[0]) << 24) | (uint32_t(e[1]) << 16) | (uint32_t(e[2]) << 8) | e[3];
            if (want != got[i]) return "Checksum mismatch at chunk " + std::to_string(start + i);
        }
        return std::nullopt;
    }
    /// one pass of run_background_scrubber (chunkserver.rs:652-701) as ONE batched verify: ids of bad blocks
    std::vector<std::string> scrub() const {
        std::vector<std::string> ids, bad;
        for (auto &e : std::filesystem::directory_iterator(dir_))
            if (e.is_regular_file() && e.path().extension() != ".meta") ids.push_back(e.path().filename().string());
        std::sort(ids.begin(), ids.end());
        std::vector<Bytes> blocks, metas;
        std::vector<std::string> have;
        for (auto &id : ids) {
            const auto mp = dir_ / (id + ".meta");
            if (!std::filesystem::exists(mp)) { bad.push_back(id); continue; }
            blocks.push_back(read_file(dir_ / id));
            metas.push_back(read_file(mp));
            have.push_back(id);
        }
        const size_t n = have.size();
        std::vector<const uint8_t *> dp(n), mp(n);
        std::vector<size_t> dl(n), ml(n);
        std::vector<int> st(n);
        for (size_t i = 0; i < n; i++) { dp[i] = blocks[i].data(); dl[i] = blocks[i].size(); mp[i] = metas[i].data(); ml[i] = metas[i].size(); }
        check(gpu_ec_verify_chunks_batch(n, dp.data(), dl.data(), CHECKSUM_CHUNK_SIZE, GPU_EC_POLY_IEEE, mp.data(), ml.data(), st.data(), nullptr));
        for (size_t i = 0; i < n; i++) if (st[i] != GPU_EC_OK) bad.push_back(have[i]);
        std::sort(bad.begin(), bad.end());
        return bad;
    }

    static Bytes read_file(const std::filesystem::path &p) {
        std::ifstream f(p, std::ios::binary);
        return Bytes((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    }
    static void write_file(const std::filesystem::path &p, const Bytes &b) {
        std::ofstream f(p, std::ios::binary | std::ios::trunc);
        f.write(reinterpret_cast<const char *>(b.data()), static_cast<std::streamsize>(b.size()));
    }

private:
    std::filesystem::path dir_;
    std::optional<std::filesystem::path> cold_;
};

}  // namespace chunkserver
}  // namespace gpu_ec
