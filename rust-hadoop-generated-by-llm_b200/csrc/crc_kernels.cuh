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
//     with fully coalesced 16-byte cp.async.cg copies (8 lanes per 128-byte line, straight into a
//     swizzled per-warp shared-memory box: no register staging, no L1 allocation — with ~200 KiB
//     of tables + boxes the L1 is tiny and an LDG path through it throttles), and read back
//     row-per-lane with conflict-free LDS.128; the next slice streams in while this one is hashed;
//   * chunk sizes above G (64 KiB blocks, whole shards) are hashed as G-byte sub-segments and
//     folded by crc_fold_kernel using carry-less shift constants (x^(8n) mod P) and warp
//     shuffles: crc(A||B) = crc(A)*x^(8|B|) xor crc(B) on finalised CRCs;
//   * the ragged tail of a buffer (len % G bytes) rides along as one more row of the last tile
//     (fewer column slices), only its final < 128 bytes are hashed serially by that lane.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "lut_common.cuh"

namespace gpu_ec {

#ifndef GPU_EC_CRC_WARPS
#define GPU_EC_CRC_WARPS 16
#endif
constexpr int CRC_WARPS = GPU_EC_CRC_WARPS;
constexpr int CRC_THREADS = CRC_WARPS * 32;
constexpr int CRC_ROW_BYTES = 128;                         // XOR-swizzled 16-byte chunks, no padding
constexpr int CRC_BOX_BYTES = 32 * CRC_ROW_BYTES;          // per-warp staging box
// Shared-memory map (shared-window addresses; the dynamic window starts at s0, 1 KiB on sm_100):
//   [s0, 0x10000)        staging boxes of the first warps
//   [0x10000, 0x20000)   table block A: T3 and T2 interleaved, entry b at A + b*256 + {0,128} + lane*4
//   [0x20000, 0x30000)   table block B: T1 and T0 interleaved the same way
//   [0x30000, ...)       staging boxes of the remaining warps
// Blocks are 64 KiB aligned and rows 256 B, so an address is {block base | data byte | lane*4}
// byte for byte: ONE PRMT assembles it (no shift, no multiply-add) and the table half is the LDS
// immediate: 2 instructions per lookup instead of 3.
constexpr uint32_t CRC_BLK_A = 0x10000u, CRC_BLK_B = 0x20000u, CRC_REGION2 = 0x30000u;
constexpr int CRC_TAB_BYTES = 4 * 256 * 32 * 4;            // 128 KiB
constexpr int CRC_SMEM_BYTES = 232448;                     // the whole 227 KiB opt-in window

// slice-by-4 tables for both polynomials, filled once per device by the host (host_math.hpp)
__device__ uint32_t g_crc_tab[2][4][256];

struct CrcParams {
    const uint8_t *base;
    unsigned long long buf_stride;
    unsigned long long n_buf;
    unsigned long long len;            // bytes per buffer
    unsigned long long nfull;          // full G-byte segments per buffer
    unsigned long long segs_per_buf;   // nfull + (rag ? 1 : 0)
    unsigned long long tiles_per_buf;  // ceil(segs_per_buf / 32): the ragged segment is a row too
    unsigned int G;                    // segment bytes, multiple of 128
    unsigned int rag;                  // len % G
    uint32_t *out;                     // out[b*out_stride + seg]
    unsigned long long out_stride;
    int poly;
    int big_endian;                    // store byte-swapped (.meta layout, chunkserver.rs:196)
    const uint8_t *expect_be;          // verify mode: big-endian expected words (4-byte aligned) or null
    unsigned long long *first_bad;     // verify mode: atomicMin target
    // fold32 mode (chunk == 32 segments): a warp's 32 lanes hold exactly the segments of ONE chunk, so the
    // warp folds them itself (one carry-less multiply per lane + shuffles) and emits the chunk CRC:
    // no per-segment scratch, no second kernel. f_c1[e] = x^(8*G*e), f_xrag = x^(8*rag).
    int fold32;
    uint32_t f_c1[32];
    uint32_t f_xrag, f_poly;
};

__device__ __forceinline__ uint32_t multmodp_dev(uint32_t a, uint32_t b, uint32_t poly);
__device__ __forceinline__ uint32_t bswap32(uint32_t v) { return __byte_perm(v, 0, 0x0123); }

// `want` = the stored big-endian word of this segment, read as a little-endian u32; in verify mode
// the kernel loads it when it ENTERS the segment so the compare at the end never waits on memory
__device__ __forceinline__ void crc_emit(const CrcParams &p, unsigned long long buf, unsigned long long seg, uint32_t crc, uint32_t want) {
    const uint32_t stored = p.big_endian ? bswap32(crc) : crc;
    if (p.out) p.out[buf * p.out_stride + seg] = stored;
    if (p.expect_be && want != bswap32(crc)) atomicMin(p.first_bad, seg);
}

// FOLD32: the in-warp fold of CrcParams::fold32, a compile-time variant so the plain path keeps its schedule
template <bool FOLD32>
__global__ void __launch_bounds__(CRC_THREADS, 1) crc_segments_kernel(const __grid_constant__ CrcParams p) {
    extern __shared__ __align__(128) unsigned char crc_smem[];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(crc_smem));
    if (s0 > 0x1000u) __trap();  // the map above assumes the dynamic window starts in the first 4 KiB
    // replicate the 4 KiB table 32 times: one coalesced global load per lane, then each entry is
    // broadcast into its 32 lane copies. Table t (T3,T2 -> block A; T1,T0 -> block B), half = !(t&1).
    for (int e0 = warp * 32; e0 < 1024; e0 += CRC_WARPS * 32) {
        const int e = e0 + lane;
        const uint32_t v = g_crc_tab[p.poly][e >> 8][e & 255];
#pragma unroll 8
        for (int q = 0; q < 32; q++) {
            const int eq = e0 + q, t = eq >> 8, bq = eq & 255;
            const uint32_t addr = ((t >= 2) ? CRC_BLK_A : CRC_BLK_B) + bq * 256 + ((t & 1) ? 0 : 128) + lane * 4;
            const uint32_t vq = __shfl_sync(0xffffffffu, v, q);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(vq) : "memory");
        }
    }
    // staging box of this warp: below block A while there is room, the rest above block B
    const uint32_t s0a = (s0 + 127u) & ~127u;
    const int n1 = static_cast<int>((CRC_BLK_A - s0a) / CRC_BOX_BYTES);
    // fold32 constants behind the last box (a lane-indexed read of the parameter bank would replay 32x)
    const uint32_t cc_s = CRC_REGION2 + (CRC_WARPS > n1 ? (CRC_WARPS - n1) : 0) * CRC_BOX_BYTES;
    if (FOLD32 && threadIdx.x < 32) asm volatile("st.shared.u32 [%0], %1;" ::"r"(cc_s + threadIdx.x * 4), "r"(p.f_c1[threadIdx.x]) : "memory");
    __syncthreads();

    const uint32_t box_s = (warp < n1) ? (s0a + warp * CRC_BOX_BYTES) : (CRC_REGION2 + (warp - n1) * CRC_BOX_BYTES);
    const uint4 *box = reinterpret_cast<const uint4 *>(crc_smem + (box_s - s0));
    const uint32_t lb_a = CRC_BLK_A | static_cast<uint32_t>(lane * 4), lb_b = CRC_BLK_B | static_cast<uint32_t>(lane * 4);
    // table[byte I of c]: address = {block | byte | lane*4} by one PRMT, table half in the immediate
#define CRC_LUT(LB, HALF, C, I)                                                                         \
    ([&]() -> uint32_t {                                                                                \
        uint32_t v_;                                                                                    \
        asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v_) : "r"(__byte_perm((C), (LB), 0x7604 | ((I) << 4))), "n"((HALF) * 128)); \
        return v_;                                                                                      \
    }())
    auto step_word = [&](uint32_t c, uint32_t w) -> uint32_t {
        c ^= w;
        const uint32_t a = CRC_LUT(lb_a, 0, c, 0);  // T3[byte 0]
        const uint32_t b = CRC_LUT(lb_a, 1, c, 1);  // T2[byte 1]
        const uint32_t d = CRC_LUT(lb_b, 0, c, 2);  // T1[byte 2]
        const uint32_t e = CRC_LUT(lb_b, 1, c, 3);  // T0[byte 3]
        return a ^ b ^ d ^ e;
    };

    const unsigned int ncol = p.G >> 7;        // 128-byte column slices per full segment
    const unsigned int ncol_rag = p.rag >> 7;  // whole slices of the ragged segment
    const unsigned long long n_tiles = p.n_buf * p.tiles_per_buf;
    const unsigned long long wstep = static_cast<unsigned long long>(gridDim.x) * CRC_WARPS;
    // staging roles: 8 lanes cover one 128-byte row slice, 4 rows per copy instruction
    const int srow = lane >> 3, scol = lane & 7;

    unsigned long long tile = static_cast<unsigned long long>(blockIdx.x) * CRC_WARPS + warp;
    unsigned int col = 0;
    // (buffer, first segment) of `tile`, advanced incrementally: no division in the loop
    const unsigned long long tpb = p.tiles_per_buf;
    const unsigned long long step_buf = wstep / tpb;
    const unsigned long long step_til = wstep - step_buf * tpb;
    unsigned long long t_buf = tpb ? tile / tpb : 0;
    unsigned long long t_til = tile - t_buf * tpb;
    // number of column slices row `seg` of a buffer has
    auto cols_of = [&](unsigned long long seg) -> unsigned int {
        return seg < p.nfull ? ncol : (seg == p.nfull ? ncol_rag : 0u);
    };
    // global -> shared without touching registers or allocating in L1 (cp.async.cg, 16 bytes per
    // lane): the box is filled for slice (tile, c_) while the previous slice is hashed from registers.
    // Per tile the lane's 8 staging rows are classified once (bit i of full_mask: row i*4+srow is a
    // full segment; rag_i: which of them, if any, is the buffer's ragged segment), so a slice costs
    // 8 predicated copies and a handful of 32-bit adds.
    uint32_t dst_of[8];
    unsigned int row_off[8];  // row * G (G <= 64 KiB, 32 rows: fits 32 bits)
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const unsigned int row = i * 4 + srow;
        dst_of[i] = box_s + (row * 8 + (scol ^ (row & 7))) * 16;
        row_off[i] = row * p.G;
    }
    const uint8_t *tile_src = nullptr;
    unsigned int full_mask = 0;
    int rag_i = -1;
    auto enter_tile = [&](unsigned long long buf, unsigned long long til) {
        const unsigned long long seg0 = til * 32ull;
        tile_src = p.base + buf * p.buf_stride + seg0 * p.G + scol * 16;
        const unsigned long long left = p.nfull > seg0 ? p.nfull - seg0 : 0;  // full rows from seg0 on
        full_mask = 0;
        rag_i = -1;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const unsigned int row = i * 4 + srow;
            if (row < left) full_mask |= 1u << i;
            else if (p.rag && row == left) rag_i = i;
        }
    };
    auto issue_loads = [&](unsigned int c_) {
        const uint8_t *src = tile_src + (static_cast<unsigned long long>(c_) << 7);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (((full_mask >> i) & 1u) || (i == rag_i && c_ < ncol_rag))
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_of[i]), "l"(src + row_off[i]) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto wait_loads = [&]() { asm volatile("cp.async.wait_group 0;" ::: "memory"); };
    auto box_read = [&](int q) -> uint4 { return box[lane * 8 + (q ^ (lane & 7))]; };
    // a tile whose only row is a short ragged segment may have zero full slices: still one pass
    const unsigned int ncol_loop = ncol ? ncol : 1u;
    if (tile < n_tiles) { enter_tile(t_buf, t_til); issue_loads(0); }
    uint32_t c = 0xFFFFFFFFu, want = 0;
    while (tile < n_tiles) {
        wait_loads();
        __syncwarp();
        // this lane's 128-byte row slice -> registers (conflict-free LDS.128 through the swizzle)
        uint4 v[8];
#pragma unroll
        for (int q = 0; q < 8; q++) v[q] = box_read(q);
        __syncwarp();
        // advance; the next slice streams into the box while this one is hashed
        const unsigned long long buf = t_buf;
        const unsigned long long seg = t_til * 32ull + lane;
        const unsigned int cur_col = col;
        const bool last_col = (col + 1 == ncol_loop);
        if (last_col) {
            tile += wstep; col = 0;
            t_buf += step_buf; t_til += step_til;
            if (t_til >= tpb) { t_til -= tpb; t_buf++; }
            if (tile < n_tiles) enter_tile(t_buf, t_til);
        } else {
            col++;
        }
        if (tile < n_tiles) issue_loads(col);
        if constexpr (FOLD32) {  // expected word of the CHUNK, kept by lane 0
            if (cur_col == 0 && p.expect_be && lane == 0) want = __ldg(reinterpret_cast<const uint32_t *>(p.expect_be + 4ull * (seg / 32ull)));
        } else {
            if (cur_col == 0 && p.expect_be && seg < p.segs_per_buf)
                want = __ldg(reinterpret_cast<const uint32_t *>(p.expect_be + 4ull * seg));
        }
        if (cur_col < cols_of(seg)) {
#pragma unroll
            for (int q = 0; q < 8; q++) {
                c = step_word(c, v[q].x);
                c = step_word(c, v[q].y);
                c = step_word(c, v[q].z);
                c = step_word(c, v[q].w);
            }
        }
        if (last_col) {
            auto ragged_tail = [&]() {  // the ragged segment's last (rag % 128) bytes straight from global
                const uint8_t *src = p.base + buf * p.buf_stride + p.nfull * p.G + (static_cast<unsigned long long>(ncol_rag) << 7);
                const unsigned int left = p.rag & 127u;
                unsigned int i = 0;
                for (; i + 4 <= left; i += 4) c = step_word(c, *reinterpret_cast<const uint32_t *>(src + i));
                for (; i < left; i++) c = (c >> 8) ^ CRC_LUT(lb_b, 1, c ^ src[i], 0);
            };
            if constexpr (!FOLD32) {
                if (seg < p.nfull) {
                    crc_emit(p, buf, seg, ~c, want);
                } else if (seg == p.nfull && p.rag) {
                    ragged_tail();
                    crc_emit(p, buf, seg, ~c, want);
                }
            } else {
                // this warp's lanes are the segments of one chunk: weight lane l by x^(8*G*(nft-1-l)), XOR-reduce,
                // then append the ragged segment (if the buffer ends in this chunk): fold * x^(8*rag) ^ p_rag
                if (seg == p.nfull && p.rag) ragged_tail();
                const unsigned long long seg0 = seg - lane;
                const unsigned int nft = (p.nfull > seg0) ? static_cast<unsigned int>(p.nfull - seg0 < 32 ? p.nfull - seg0 : 32) : 0u;
                const uint32_t mine = ~c;
                uint32_t f = 0;
                if (static_cast<unsigned int>(lane) < nft) {
                    uint32_t w_;
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w_) : "r"(cc_s + (nft - 1u - lane) * 4u));
                    f = multmodp_dev(mine, w_, p.f_poly);
                }
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) f ^= __shfl_xor_sync(0xffffffffu, f, sft);
                const bool rag_here = p.rag && nft < 32 && (seg0 + nft == p.nfull);
                const uint32_t p_rag = __shfl_sync(0xffffffffu, mine, static_cast<int>(nft & 31u));
                if (rag_here) f = multmodp_dev(f, p.f_xrag, p.f_poly) ^ p_rag;
                if (lane == 0 && (nft || rag_here)) crc_emit(p, buf, seg0 / 32ull, f, want);
            }
            c = 0xFFFFFFFFu;
        }
    }
#undef CRC_LUT
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

// Fold per-segment CRCs into one CRC per output (chunk or whole buffer), one level per launch.
// Elements are weighted by their distance from the END of the output, counted in elements:
// element e (0 = last full element) contributes p_e * x^(8*U*e), U = bytes per element. A team of
// T threads folds T consecutive elements: lane l multiplies by the per-level constant
// c1[l] = x^(8*U*l), the warp XOR-reduces with shuffles, warp w's value is multiplied by
// c2[w] = x^(8*U*32*w) and reduced again. Outputs with more than T elements produce one value per
// block of T, which the next launch folds with U' = U*T. All constants are 32-bit values computed
// on the host; nothing is irregular: a ragged final element only enters at the last level as
//   result = fold * x^(8*rag) xor p_rag.
struct FoldParams {
    const uint32_t *cur;                // cur[b*cur_stride_buf + oi*cur_group + i], position order
    unsigned long long n_buf, outs_per_buf, cur_stride_buf;
    unsigned int cur_group;             // element stride between the outputs of one buffer
    unsigned int cnt_reg, cnt_last;     // FULL elements in a regular / in the last output of a buffer
    unsigned int T, nb;                 // team threads; blocks of T elements per output at this level
    uint32_t c1[32], c2[32];
    uint32_t poly_bits;
    int final_level;
    uint32_t *nxt;                      // !final: nxt[b*nxt_stride_buf + oi*nb + j]
    unsigned long long nxt_stride_buf;
    uint32_t x_rag;                     // final: x^(8*rag)
    int last_has_rag;
    const uint32_t *rag_src;            // final: level-1 array, ragged element = last of the buffer's last output
    unsigned long long rag_stride_buf;
    unsigned int rag_group, rag_last_cnt;
    uint32_t *out;                      // out[b*out_stride + oi]
    unsigned long long out_stride;
    int big_endian;
    const uint8_t *expect_be;
    unsigned long long *first_bad;
    int cur_be, rag_be;                 // `cur` / `rag_src` hold byte-swapped words (a .meta array folded in place)
};

__global__ void crc_fold_kernel(const __grid_constant__ FoldParams p) {
    __shared__ uint32_t red[32];
    const unsigned int T = p.T, W = T >> 5;
    const unsigned int tpc = blockDim.x / T;  // teams per CTA
    const unsigned int team = threadIdx.x / T, tau = threadIdx.x - team * T;
    const unsigned int lane = tau & 31, wteam = tau >> 5, gwarp = threadIdx.x >> 5;
    const uint32_t P = p.poly_bits;
    // per-lane constants through shared memory: a lane-indexed read of the parameter bank would be
    // replayed 32 times per warp, and most threads fold exactly one element
    __shared__ uint32_t cc[64];
    if (threadIdx.x < 64) cc[threadIdx.x] = (threadIdx.x < 32) ? p.c1[threadIdx.x] : p.c2[threadIdx.x - 32];
    __syncthreads();
    const uint32_t my_c1 = cc[lane];
    const uint32_t my_c2 = cc[32 + lane];
    const unsigned long long total = p.n_buf * p.outs_per_buf * p.nb;
    for (unsigned long long item0 = static_cast<unsigned long long>(blockIdx.x) * tpc; item0 < total;
         item0 += static_cast<unsigned long long>(gridDim.x) * tpc) {
        const unsigned long long item = item0 + team;
        const bool valid = item < total;
        const unsigned long long o = item / p.nb;
        const unsigned int blk = static_cast<unsigned int>(item - o * p.nb);
        const unsigned long long b = o / p.outs_per_buf, oi = o - b * p.outs_per_buf;
        const bool is_last = (oi + 1 == p.outs_per_buf);
        const unsigned int cnt = is_last ? p.cnt_last : p.cnt_reg;
        const unsigned long long e = static_cast<unsigned long long>(blk) * T + tau;
        const uint32_t *src = p.cur + b * p.cur_stride_buf + oi * p.cur_group;
        uint32_t v = 0;
        if (valid && e < cnt) {
            const uint32_t w = src[cnt - 1 - e];
            v = multmodp_dev(p.cur_be ? bswap32(w) : w, my_c1, P);
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) v ^= __shfl_xor_sync(0xffffffffu, v, s);
        if (W > 1) {
            __syncthreads();
            if (lane == 0) red[gwarp] = v;
            __syncthreads();
            if (wteam == 0) {
                v = (lane < W) ? multmodp_dev(red[team * W + lane], my_c2, P) : 0u;
#pragma unroll
                for (int s = 16; s > 0; s >>= 1) v ^= __shfl_xor_sync(0xffffffffu, v, s);
            }
        }
        if (valid && tau == 0) {
            const unsigned int nb_o = (cnt + T - 1) / T;
            if (!p.final_level) {
                if (blk < nb_o) p.nxt[b * p.nxt_stride_buf + oi * p.nb + (nb_o - 1 - blk)] = v;
            } else {
                uint32_t r = v;
                if (is_last && p.last_has_rag) {
                    const uint32_t w = p.rag_src[b * p.rag_stride_buf + oi * p.rag_group + p.rag_last_cnt - 1];
                    r = multmodp_dev(r, p.x_rag, P) ^ (p.rag_be ? bswap32(w) : w);
                }
                if (p.out) p.out[b * p.out_stride + oi] = p.big_endian ? bswap32(r) : r;
                if (p.expect_be) {
                    const uint32_t want = *reinterpret_cast<const uint32_t *>(p.expect_be + 4ull * oi);
                    if (want != bswap32(r)) atomicMin(p.first_bad, oi);
                }
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

// ---------------------------------------------------------------------------------------------
// Batched small-block checksums (many WriteBlocks in one go: chunkserver.rs:746-769 + :195 per block).
// The blocks sit back to back in one staging buffer, each starting on a chunk boundary and zero-padded to
// one, so ONE crc_segments launch hashes every chunk of every block; this kernel then, per block (one CTA),
//   * repairs the CRC of a short last chunk, which was hashed with its zero padding:
//       crc(A) = (crc(A || 0^p) ^ crc(0^chunk)) * x^(-8p) ^ crc(0^r),   r = |A|, p = chunk - r
//   * folds the chunk CRCs into the whole-block CRC: crc(A||B) = crc(A) * x^(8|B|) ^ crc(B).
struct BlockDesc {
    unsigned long long first_chunk;  // index of the block's first chunk CRC in `crcs`
    unsigned int n_full;             // full chunks
    unsigned int rag;                // bytes of the short last chunk (0 = none)
    uint32_t x_inv_pad;              // x^(-8*(chunk - rag))
    uint32_t cz_rag;                 // crc(zeros(rag))
    uint32_t x_rag;                  // x^(8*rag)
    uint32_t pad_;
};
struct BlocksFinishParams {
    const BlockDesc *desc;
    unsigned long long n_blocks;
    uint32_t *crcs;       // chunk CRCs of all blocks, repaired in place
    uint32_t *whole;      // [n_blocks] or null
    uint32_t cz_chunk;    // crc(zeros(chunk))
    uint32_t x_chunk;     // x^(8*chunk)
    uint32_t poly_bits;
};
__global__ void __launch_bounds__(128) crc_blocks_finish_kernel(const __grid_constant__ BlocksFinishParams p) {
    __shared__ uint32_t z[8], wsum[4];
    const unsigned int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t P = p.poly_bits;
    for (unsigned long long b = blockIdx.x; b < p.n_blocks; b += gridDim.x) {
        const BlockDesc d = p.desc[b];
        const unsigned int pt = (d.n_full + 127u) / 128u;  // full chunks per thread, right-aligned over 128*pt slots
        __syncthreads();
        if (t == 0) {  // tree constants z[L] = X^(pt * 2^L), X = x^(8*chunk)
            uint32_t r = 0x80000000u, sq = p.x_chunk;
            for (unsigned int e = pt; e; e >>= 1) { if (e & 1u) r = multmodp_dev(sq, r, P); sq = multmodp_dev(sq, sq, P); }
            for (int L = 0; L < 7; L++) { z[L] = r; r = multmodp_dev(r, r, P); }
        }
        __syncthreads();
        const unsigned int off = pt * 128u - d.n_full;
        uint32_t val = 0;
        for (unsigned int i = 0; i < pt; i++) {
            const unsigned int slot = t * pt + i;
            val = multmodp_dev(val, p.x_chunk, P);
            if (slot >= off) val ^= p.crcs[d.first_chunk + (slot - off)];
        }
#pragma unroll
        for (int L = 0; L < 5; L++) {
            const uint32_t left = __shfl_up_sync(0xffffffffu, val, 1u << L);
            if ((lane & ((2u << L) - 1u)) == (2u << L) - 1u) val ^= multmodp_dev(left, z[L], P);
        }
        if (lane == 31) wsum[warp] = val;
        __syncthreads();
        if (t == 0) {
            uint32_t w = 0;
            for (int i = 0; i < 4; i++) w = multmodp_dev(w, z[5], P) ^ wsum[i];  // z[5] = X^(32*pt): one warp further
            if (d.rag) {
                const uint32_t c = multmodp_dev(p.crcs[d.first_chunk + d.n_full] ^ p.cz_chunk, d.x_inv_pad, P) ^ d.cz_rag;
                p.crcs[d.first_chunk + d.n_full] = c;
                w = multmodp_dev(w, d.x_rag, P) ^ c;
            }
            if (p.whole) p.whole[b] = w;
        }
    }
}

__global__ void first_bad_finish_kernel(unsigned long long *first_bad) {
    if (*first_bad == 0xFFFFFFFFFFFFFFFFull) *reinterpret_cast<long long *>(first_bad) = -1;
}

}  // namespace gpu_ec
