// host_math.hpp — small host-side algebra for libgpu_ec: GF(2^8) coding matrices and CRC-32
// table/shift constants. Everything here works on <= 256x256-byte matrices or 4-byte values;
// the data path itself only ever runs in the CUDA kernels.
//
// Matches reed-solomon-erasure 6.0.0 galois_8 (field 0x11D, generator 2; systematic matrix =
// Vandermonde x inverse(top square); reconstruct uses the first k present shards) as called at
// dfs/common/src/erasure.rs:14,21,36,38 and dfs/chunkserver/src/chunkserver.rs:620-623.
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace gpu_ec {

struct GF {
    uint8_t exp[512];
    uint8_t log[256];
    GF() {
        unsigned x = 1;
        for (int i = 0; i < 255; i++) {
            exp[i] = static_cast<uint8_t>(x);
            log[x] = static_cast<uint8_t>(i);
            x <<= 1;
            if (x & 0x100) x ^= 0x11D;
        }
        for (int i = 255; i < 512; i++) exp[i] = exp[i - 255];
        log[0] = 0;
    }
    uint8_t mul(uint8_t a, uint8_t b) const { return (a && b) ? exp[log[a] + log[b]] : 0; }
    uint8_t inv(uint8_t a) const { return exp[255 - log[a]]; }
    uint8_t pow(uint8_t a, unsigned n) const {
        if (n == 0) return 1;
        if (a == 0) return 0;
        return exp[(static_cast<unsigned>(log[a]) * n) % 255];
    }
};
inline const GF &gf() { static const GF g; return g; }

using Mat = std::vector<uint8_t>;  // row-major

// in-place Gauss-Jordan on [A | I]; false when singular
inline bool invert(const Mat &a, int n, Mat &out) {
    const GF &g = gf();
    const int W = 2 * n;
    Mat w(static_cast<size_t>(n) * W, 0);
    for (int r = 0; r < n; r++) {
        std::memcpy(&w[static_cast<size_t>(r) * W], &a[static_cast<size_t>(r) * n], n);
        w[static_cast<size_t>(r) * W + n + r] = 1;
    }
    for (int c = 0; c < n; c++) {
        int piv = -1;
        for (int r = c; r < n; r++) if (w[static_cast<size_t>(r) * W + c]) { piv = r; break; }
        if (piv < 0) return false;
        if (piv != c) for (int j = 0; j < W; j++) std::swap(w[static_cast<size_t>(c) * W + j], w[static_cast<size_t>(piv) * W + j]);
        uint8_t s = g.inv(w[static_cast<size_t>(c) * W + c]);
        for (int j = 0; j < W; j++) w[static_cast<size_t>(c) * W + j] = g.mul(s, w[static_cast<size_t>(c) * W + j]);
        for (int r = 0; r < n; r++) {
            uint8_t f = w[static_cast<size_t>(r) * W + c];
            if (r == c || !f) continue;
            for (int j = 0; j < W; j++) w[static_cast<size_t>(r) * W + j] ^= g.mul(f, w[static_cast<size_t>(c) * W + j]);
        }
    }
    out.assign(static_cast<size_t>(n) * n, 0);
    for (int r = 0; r < n; r++) std::memcpy(&out[static_cast<size_t>(r) * n], &w[static_cast<size_t>(r) * W + n], n);
    return true;
}

// (k+m) x k systematic coding matrix; cached per (k,m) (the reference rebuilds it on every call)
inline const Mat &coding_matrix(int k, int m) {
    static std::mutex mu;
    static std::map<std::pair<int, int>, Mat> cache;
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find({k, m});
    if (it != cache.end()) return it->second;
    const GF &g = gf();
    const int n = k + m;
    Mat v(static_cast<size_t>(n) * k);
    for (int r = 0; r < n; r++)
        for (int c = 0; c < k; c++) v[static_cast<size_t>(r) * k + c] = g.pow(static_cast<uint8_t>(r), static_cast<unsigned>(c));
    Mat top(v.begin(), v.begin() + static_cast<size_t>(k) * k), top_inv;
    invert(top, k, top_inv);  // a Vandermonde top square with distinct points is never singular
    Mat M(static_cast<size_t>(n) * k, 0);
    for (int r = 0; r < n; r++)
        for (int c = 0; c < k; c++) {
            uint8_t acc = 0;
            for (int t = 0; t < k; t++) acc ^= g.mul(v[static_cast<size_t>(r) * k + t], top_inv[static_cast<size_t>(t) * k + c]);
            M[static_cast<size_t>(r) * k + c] = acc;
        }
    return cache.emplace(std::make_pair(k, m), std::move(M)).first->second;
}

struct DecodePlan {
    std::vector<int> survivors;  // k shard indices read (first k present, index order)
    std::vector<int> missing;    // absent shard indices written
    Mat rows;                    // missing.size() x k, over the survivors
};

// One-pass decode rows, cached per (k, m, present-bitmap) like the crate's LRU of inverses.
// returns 0, or -1 when fewer than k shards are present, or -2 when singular.
inline int decode_plan(int k, int m, const uint8_t *present, DecodePlan &plan) {
    static std::mutex mu;
    static std::map<std::vector<uint8_t>, DecodePlan> cache;
    const int n = k + m;
    std::vector<uint8_t> key(static_cast<size_t>(n) + 2);
    key[0] = static_cast<uint8_t>(k - 1);
    key[1] = static_cast<uint8_t>(m - 1);
    int np = 0;
    for (int i = 0; i < n; i++) { key[i + 2] = present[i] ? 1 : 0; np += key[i + 2]; }
    if (np < k) return -1;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) { plan = it->second; return 0; }
    }
    const Mat &M = coding_matrix(k, m);
    const GF &g = gf();
    DecodePlan p;
    for (int i = 0; i < n; i++) {
        if (present[i]) { if (static_cast<int>(p.survivors.size()) < k) p.survivors.push_back(i); }
        else p.missing.push_back(i);
    }
    Mat sub(static_cast<size_t>(k) * k), inv;
    for (int r = 0; r < k; r++) std::memcpy(&sub[static_cast<size_t>(r) * k], &M[static_cast<size_t>(p.survivors[r]) * k], k);
    if (!invert(sub, k, inv)) return -2;
    p.rows.assign(p.missing.size() * static_cast<size_t>(k), 0);
    for (size_t i = 0; i < p.missing.size(); i++) {
        int idx = p.missing[i];
        if (idx < k) std::memcpy(&p.rows[i * k], &inv[static_cast<size_t>(idx) * k], k);
        else
            for (int c = 0; c < k; c++) {
                uint8_t acc = 0;
                for (int t = 0; t < k; t++) acc ^= g.mul(M[static_cast<size_t>(idx) * k + t], inv[static_cast<size_t>(t) * k + c]);
                p.rows[i * k + c] = acc;
            }
    }
    {
        std::lock_guard<std::mutex> lk(mu);
        if (cache.size() > 4096) cache.clear();
        cache[key] = p;
    }
    plan = std::move(p);
    return 0;
}

// ---- CRC-32 constants (reflected representation: bit 31 = x^0) -----------------------------
constexpr uint32_t kPolyReflected[2] = {0xEDB88320u, 0x82F63B78u};

inline uint32_t multmodp(uint32_t a, uint32_t b, uint32_t poly) {
    uint32_t p = 0;
    for (int i = 0; i < 32; i++) {
        if (a & 0x80000000u) p ^= b;
        a <<= 1;
        b = (b >> 1) ^ ((b & 1) ? poly : 0);
    }
    return p;
}
// x^(8 n) mod P
inline uint32_t x8n_modp(uint64_t n, uint32_t poly) {
    uint32_t p = 0x80000000u, sq = 0x00800000u;
    while (n) {
        if (n & 1) p = multmodp(sq, p, poly);
        sq = multmodp(sq, sq, poly);
        n >>= 1;
    }
    return p;
}
// x^(-8 n) mod P: x^-1 = (P + 1)/x exists because P has a constant term
inline uint32_t xinv8n_modp(uint64_t n, uint32_t poly) {
    uint32_t r = 0x80000000u, sq = ((0x80000000u ^ poly) << 1) | 1u;
    for (uint64_t e = 8ull * n; e; e >>= 1) {
        if (e & 1) r = multmodp(sq, r, poly);
        sq = multmodp(sq, sq, poly);
    }
    return r;
}
// CRC-32 (init and xorout 0xFFFFFFFF) of n zero bytes in O(log n): the register is just multiplied by x^8 per byte
inline uint32_t crc_of_zeros(uint64_t n, uint32_t poly) {
    return n ? (multmodp(0xFFFFFFFFu, x8n_modp(n, poly), poly) ^ 0xFFFFFFFFu) : 0u;
}
// slice-by-4 tables T[t][b] uploaded once per device
inline void crc_tables(uint32_t poly, uint32_t (*T)[256]) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int j = 0; j < 8; j++) c = (c >> 1) ^ ((c & 1) ? poly : 0);
        T[0][i] = c;
    }
    for (int t = 1; t < 4; t++)
        for (uint32_t i = 0; i < 256; i++) T[t][i] = (T[t - 1][i] >> 8) ^ T[0][T[t - 1][i] & 0xff];
}

}  // namespace gpu_ec
