// crc_kernels.cuh — CRC-32 kernels for sm_100a (IEEE = what the reference's crc32fast computes;
// Castagnoli as the labelled extra variant).
//
// Replaces: MyChunkServer::calculate_checksums (dfs/chunkserver/src/chunkserver.rs:182-190), the
// verify_block compare loop (:275-289) and the inline whole-buffer crc32fast::Hasher sites
// (:746-750, :1006-1010; dfs/client/src/mod.rs:329-341,426-428,601-611).
//
// B200 has no carry-less multiply, so CRC is table work and the binding resource is the
// shared-memory crossbar (32 lookups/clk/SM = at most ~1 byte per lookup). Design:
//   * slice-by-4 with the four 256-entry tables replicated per lane (word = (t*256+b)*32+lane,
//     128 KiB) so every lookup is bank-conflict-free; exactly 4 LDS per 4 input bytes;
//   * lane = segment: each lane runs the plain serial CRC of its own G-byte segment, so no
//     per-lane shift tables and no cross-lane fold in the hot loop. A warp owns 32 consecutive
//     segments and walks them in 128-byte column slices: the slice (32 rows x 128 B) is fetched
//     with fully coalesced 16-byte loads (8 lanes per 128-byte line), parked in a swizzled
//     per-warp shared-memory box, and read back row-per-lane with conflict-free LDS.128;
//     the next slice's loads are in flight while the current one is hashed;
//   * chunk sizes above G (64 KiB blocks, whole shards) are hashed as G-byte sub-segments and
//     folded by crc_combine_kernel using carry-less shift constants (x^(8n) mod P) and warp
//     shuffles: crc(A||B) = crc(A)*x^(8|B|) xor crc(B) on finalised CRCs;
//   * the ragged tail of a buffer (len % G bytes) is hashed by one lane straight from global.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "lut_common.cuh"

namespace gpu_ec {

constexpr int CRC_WARPS = 16;
constexpr int CRC_THREADS = CRC_WARPS * 32;
constexpr int CRC_TAB_BYTES = 4 * 256 * 32 * 4;            // 128 KiB
constexpr int CRC_BOX_BYTES = 32 * 128;                    // per-warp staging box
constexpr int CRC_SMEM_BYTES = CRC_TAB_BYTES + CRC_WARPS * CRC_BOX_BYTES;  // 192 KiB

// slice-by-4 tables for both polynomials, filled once per device by the host (host_math.hpp)
__device__ uint32_t g_crc_tab[2][4][256];

struct CrcParams {
    const uint8_t *base;
    unsigned long long buf_stride;
    unsigned long long n_buf;
    unsigned long long len;            // bytes per buffer
    unsigned long long nfull;          // full G-byte segments per buffer
    unsigned long long segs_per_buf;   // nfull + (rag ? 1 : 0)
    unsigned long long tiles_per_buf;  // ceil(nfull / 32)
    unsigned int G;                    // segment bytes, multiple of 128
    unsigned int rag;                  // len % G
    uint32_t *out;                     // out[b*out_stride + seg]
    unsigned long long out_stride;
    int poly;
    int big_endian;                    // store byte-swapped (.meta layout, chunkserver.rs:196)
    const uint8_t *expect_be;          // verify mode: big-endian expected words (4-byte aligned) or null
    unsigned long long *first_bad;     // verify mode: atomicMin target
};

__device__ __forceinline__ uint32_t bswap32(uint32_t v) { return __byte_perm(v, 0, 0x0123); }

__device__ __forceinline__ void crc_emit(const CrcParams &p, unsigned long long buf, unsigned long long seg, uint32_t crc) {
    const uint32_t stored = p.big_endian ? bswap32(crc) : crc;
    if (p.out) p.out[buf * p.out_stride + seg] = stored;
    if (p.expect_be) {
        const uint32_t want = *reinterpret_cast<const uint32_t *>(p.expect_be + 4ull * seg);  // BE bytes read as LE word
        if (want != bswap32(crc)) atomicMin(p.first_bad, seg);
    }
}

__global__ void __launch_bounds__(CRC_THREADS, 1) crc_segments_kernel(const __grid_constant__ CrcParams p) {
    extern __shared__ __align__(128) unsigned char crc_smem[];
    uint32_t *tab = reinterpret_cast<uint32_t *>(crc_smem);
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    // replicate the 4 KiB table 32 times: thread i of each group of 32 writes lane copy i
    for (int e = warp; e < 1024; e += CRC_WARPS) tab[e * 32 + lane] = g_crc_tab[p.poly][e >> 8][e & 255];
    __syncthreads();

    const uint32_t tl = static_cast<uint32_t>(__cvta_generic_to_shared(tab)) + lane * 4;
    uint4 *box = reinterpret_cast<uint4 *>(crc_smem + CRC_TAB_BYTES + warp * CRC_BOX_BYTES);
    auto step_word = [&](uint32_t c, uint32_t w) -> uint32_t {
        c ^= w;
        const uint32_t a = lut_lds<3 * 32768>(tl, byte_of<0>(c));
        const uint32_t b = lut_lds<2 * 32768>(tl, byte_of<1>(c));
        const uint32_t d = lut_lds<1 * 32768>(tl, byte_of<2>(c));
        const uint32_t e = lut_lds<0>(tl, c >> 24);
        return a ^ b ^ d ^ e;
    };

    const unsigned int ncol = p.G >> 7;  // 128-byte column slices per segment
    const unsigned long long n_tiles = p.n_buf * p.tiles_per_buf;
    const unsigned long long wstep = static_cast<unsigned long long>(gridDim.x) * CRC_WARPS;
    const unsigned long long w0 = static_cast<unsigned long long>(blockIdx.x) * CRC_WARPS + warp;
    // staging roles: 8 lanes cover one 128-byte row slice, 4 rows per load instruction
    const int srow = lane >> 3, scol = lane & 7;

    unsigned long long tile = w0;
    unsigned int col = 0;
    uint4 pre[8];
    auto issue_loads = [&](unsigned long long tl_, unsigned int c_) {
        const unsigned long long buf = tl_ / p.tiles_per_buf;
        const unsigned long long seg0 = (tl_ - buf * p.tiles_per_buf) * 32ull;
        const uint8_t *src = p.base + buf * p.buf_stride + seg0 * p.G + static_cast<unsigned long long>(c_) * 128ull + scol * 16;
        const unsigned long long left = p.nfull - seg0;  // rows available in this tile (may exceed 32)
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const unsigned int row = i * 4 + srow;
            pre[i] = (row < left) ? __ldg(reinterpret_cast<const uint4 *>(src + static_cast<unsigned long long>(row) * p.G))
                                  : make_uint4(0, 0, 0, 0);
        }
    };
    if (tile < n_tiles) issue_loads(tile, 0);
    uint32_t c = 0xFFFFFFFFu;
    while (tile < n_tiles) {
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int row = i * 4 + srow;
            box[row * 8 + (scol ^ (row & 7))] = pre[i];
        }
        __syncwarp();
        // advance and prefetch
        const unsigned long long cur_tile = tile;
        const bool last_col = (col + 1 == ncol);
        if (last_col) { tile += wstep; col = 0; } else { col++; }
        if (tile < n_tiles) issue_loads(tile, col);
        // hash this lane's 128-byte row slice
#pragma unroll
        for (int q = 0; q < 8; q++) {
            const uint4 v = box[lane * 8 + (q ^ (lane & 7))];
            c = step_word(c, v.x);
            c = step_word(c, v.y);
            c = step_word(c, v.z);
            c = step_word(c, v.w);
        }
        if (last_col) {
            const unsigned long long buf = cur_tile / p.tiles_per_buf;
            const unsigned long long seg = (cur_tile - buf * p.tiles_per_buf) * 32ull + lane;
            if (seg < p.nfull) crc_emit(p, buf, seg, ~c);
            c = 0xFFFFFFFFu;
        }
    }

    // ragged tails: one per buffer, hashed serially by lane 0 of a warp straight from global
    if (p.rag) {
        for (unsigned long long b = w0; b < p.n_buf; b += wstep) {
            if (lane == 0) {
                const uint8_t *src = p.base + b * p.buf_stride + p.nfull * p.G;
                uint32_t cc = 0xFFFFFFFFu;
                unsigned int i = 0;
                for (; i + 4 <= p.rag; i += 4) cc = step_word(cc, *reinterpret_cast<const uint32_t *>(src + i));
                for (; i < p.rag; i++) {
                    cc = (cc >> 8) ^ lut_lds<0>(tl, (cc ^ src[i]) & 0xffu);
                }
                crc_emit(p, b, p.nfull, ~cc);
            }
        }
    }
}

// Generic fallback for chunk sizes that are not a multiple of 128 or unaligned buffers: one
// thread per chunk, byte-at-a-time from global with the (non-replicated) global table.
struct CrcSlowParams {
    const uint8_t *base;
    unsigned long long buf_stride, n_buf, len, chunk, chunks_per_buf;
    uint32_t *out;
    unsigned long long out_stride;
    int poly, big_endian;
    const uint8_t *expect_be;
    unsigned long long *first_bad;
};
__global__ void crc_chunks_slow_kernel(const CrcSlowParams p) {
    const unsigned long long total = p.n_buf * p.chunks_per_buf;
    for (unsigned long long i = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
        const unsigned long long b = i / p.chunks_per_buf, ch = i - b * p.chunks_per_buf;
        const unsigned long long off = ch * p.chunk;
        const unsigned long long n = (p.len - off < p.chunk) ? p.len - off : p.chunk;
        const uint8_t *src = p.base + b * p.buf_stride + off;
        uint32_t c = 0xFFFFFFFFu;
        for (unsigned long long j = 0; j < n; j++) c = (c >> 8) ^ g_crc_tab[p.poly][0][(c ^ src[j]) & 0xffu];
        c = ~c;
        if (p.out) p.out[b * p.out_stride + ch] = p.big_endian ? bswap32(c) : c;
        if (p.expect_be) {
            const uint8_t *e = p.expect_be + 4ull * ch;
            const uint32_t want = (uint32_t(e[0]) << 24) | (uint32_t(e[1]) << 16) | (uint32_t(e[2]) << 8) | e[3];
            if (want != c) atomicMin(p.first_bad, ch);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// carry-less arithmetic mod P (reflected: bit 31 = x^0)
__device__ __forceinline__ uint32_t multmodp_dev(uint32_t a, uint32_t b, uint32_t poly) {
    uint32_t r = 0;
#pragma unroll 8
    for (int i = 0; i < 32; i++) {
        r ^= (a & 0x80000000u) ? b : 0u;
        a <<= 1;
        b = (b >> 1) ^ ((b & 1u) ? poly : 0u);
    }
    return r;
}

constexpr int COMBINE_MAX_LOG = 10;  // <= 1024 threads per output

struct CombineParams {
    const uint32_t *partial;           // [n_buf][segs_per_buf] finalised CRCs of seg_bytes-byte segments
    unsigned long long n_buf, segs_per_buf, outs_per_buf;
    unsigned int group;                // partials per regular output
    unsigned int last_cnt;             // partials in the last output of a buffer
    unsigned int last_has_rag;         // the buffer's final partial covers fewer than seg_bytes bytes
    uint32_t x_rag;                    // x^(8*rag) mod P
    uint32_t x1;                       // x^(8*seg_bytes) mod P
    // regular output: run length R, runs, x^(8*seg_bytes*R)^(2^j), x^(8*seg_bytes*len_of_last_run)
    unsigned int R_reg, runs_reg;  uint32_t xr_reg[COMBINE_MAX_LOG];  uint32_t xl_reg;
    unsigned int R_last, runs_last; uint32_t xr_last[COMBINE_MAX_LOG]; uint32_t xl_last;
    uint32_t poly_bits;
    uint32_t *out;                     // out[b*out_stride + o]
    unsigned long long out_stride;
    int big_endian;
    const uint8_t *expect_be;
    unsigned long long *first_bad;
};

// One CTA per output CRC. Thread t folds a contiguous run of partials by Horner with x1, is
// shifted to the end of the output with x^(8*seg_bytes*distance), and the CTA XOR-reduces with
// warp shuffles.
__global__ void crc_combine_kernel(const __grid_constant__ CombineParams p) {
    __shared__ uint32_t red[32];
    const unsigned int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t P = p.poly_bits;
    auto run_mult = [&](unsigned int runs, const uint32_t *xr, uint32_t xl) -> uint32_t {
        if (t + 1 >= runs) return 0x80000000u;  // last run (or idle thread): distance 0
        unsigned int e = runs - 2 - t;          // full runs between this run and the last one
        uint32_t m = xl;
        for (int j = 0; j < COMBINE_MAX_LOG && e; j++, e >>= 1)
            if (e & 1u) m = multmodp_dev(m, xr[j], P);
        return m;
    };
    const uint32_t mult_reg = run_mult(p.runs_reg, p.xr_reg, p.xl_reg);
    const unsigned long long total = p.n_buf * p.outs_per_buf;
    for (unsigned long long o = blockIdx.x; o < total; o += gridDim.x) {
        const unsigned long long b = o / p.outs_per_buf, oi = o - b * p.outs_per_buf;
        const bool is_last = (oi + 1 == p.outs_per_buf);
        const unsigned int cnt = is_last ? p.last_cnt : p.group;
        const bool rag = is_last && p.last_has_rag;
        const unsigned int nreg = cnt - (rag ? 1u : 0u);
        const unsigned int R = is_last ? p.R_last : p.R_reg;
        const uint32_t mult = is_last ? run_mult(p.runs_last, p.xr_last, p.xl_last) : mult_reg;
        const uint32_t *src = p.partial + b * p.segs_per_buf + oi * p.group;
        const unsigned long long i0 = static_cast<unsigned long long>(t) * R;
        uint32_t acc = 0;
        if (i0 < nreg) {
            const unsigned int i1 = (i0 + R < nreg) ? static_cast<unsigned int>(i0 + R) : nreg;
            acc = src[i0];
            for (unsigned int i = static_cast<unsigned int>(i0) + 1; i < i1; i++) acc = multmodp_dev(acc, p.x1, P) ^ src[i];
            acc = multmodp_dev(acc, mult, P);
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, s);
        __syncthreads();
        if (lane == 0) red[warp] = acc;
        __syncthreads();
        if (t == 0) {
            uint32_t r = 0;
            for (unsigned int w = 0; w < (blockDim.x + 31) / 32; w++) r ^= red[w];
            if (rag) r = multmodp_dev(r, p.x_rag, P) ^ src[cnt - 1];
            if (p.out) p.out[b * p.out_stride + oi] = p.big_endian ? bswap32(r) : r;
            if (p.expect_be) {
                const uint32_t want = *reinterpret_cast<const uint32_t *>(p.expect_be + 4ull * oi);
                if (want != bswap32(r)) atomicMin(p.first_bad, oi);
            }
        }
    }
}

// CRC of the original (unpadded) buffer from the data-shard CRCs of one stripe
// (client/src/mod.rs:339-341, 609-611) with 32-bit algebra only. The original covers n_full whole
// shards and, when has_part, a prefix of shard n_full whose tail is `pad` zero bytes:
//   crc(prefix) = (crc(shard) ^ crc(zeros(pad))) * x^(-8*pad)
//   full = (Horner over whole shards with x^(8S)) * x^(8*rem) ^ crc(prefix)
struct FullCrcParams {
    const uint32_t *shard_crc;  // [n_stripes][n_shards]
    uint32_t *full_crc;         // [n_stripes]
    unsigned long long n_stripes;
    int n_shards, n_full, has_part;
    uint32_t xS, xLp, x_inv_pad, crc_zeros_pad, poly_bits;
};
__global__ void stripe_full_crc_kernel(const FullCrcParams p) {
    for (unsigned long long s = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; s < p.n_stripes;
         s += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
        const uint32_t *c = p.shard_crc + s * p.n_shards;
        uint32_t acc = 0;
        for (int j = 0; j < p.n_full; j++) acc = multmodp_dev(acc, p.xS, p.poly_bits) ^ c[j];
        if (p.has_part) {
            const uint32_t part = multmodp_dev(c[p.n_full] ^ p.crc_zeros_pad, p.x_inv_pad, p.poly_bits);
            acc = multmodp_dev(acc, p.xLp, p.poly_bits) ^ part;
        }
        p.full_crc[s] = acc;
    }
}

__global__ void first_bad_finish_kernel(unsigned long long *first_bad) {
    if (*first_bad == 0xFFFFFFFFFFFFFFFFull) *reinterpret_cast<long long *>(first_bad) = -1;
}

}  // namespace gpu_ec
