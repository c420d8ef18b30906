// sfold_kernels.cuh — "sparse fold" CRC-32: whole-buffer CRCs with no tables and no carry-less
// multiply in the streaming loop, only funnel shifts and XORs on registers.
//
// Replaces the same reference sites as crc_kernels.cuh's whole-buffer mode (inline crc32fast::Hasher
// over an entire block/shard: dfs/chunkserver/src/chunkserver.rs:746-750,1006-1010;
// dfs/client/src/mod.rs:329-341,601-611). Why a second algorithm: B200 has no CLMUL, so a table CRC
// costs one shared-memory lookup per byte and owns 128 KiB of the SM; it cannot share an SM with the
// RS byte-LUT tables, which is what kept "encode + shard CRCs" a two-pass job. This one needs no
// shared memory at all, so it rides inside the RS kernel on the bytes already in registers.
//
// The algebra. A buffer is walked as rows of Dv 16-byte vectors (rows are aligned to the END of the
// buffer, virtual zeros in front). Column c keeps a 128-bit accumulator A_c; going one row down the
// buffer is A_c <- A_c * x^(128*Dv) + vector (Horner over the rows). Dv is picked offline
// (tools/sfold_search.c) so that
//        x^(128*Dv) == x^k0 + x^k1 + x^k2 + x^k3   (mod P),   all k < 64:
// the multiply is then 4 funnel shifts + XORs per 32-bit word (CRC-32 is reflected: x^k is a right
// shift of the little-endian bit stream). Bits shifted across the 128-bit boundary come from the
// next column, i.e. the next lane (2 SHFL per step); a warp's 32 columns form one 4096-bit
// polynomial. What leaves the warp at its low end (<= 63 bits per step) is not reduced in the loop:
// lane 0 logs its two low words before each step (8 bytes per 512 bytes streamed), and the finisher
// replays the lost bits as virtual data words sitting just in front of the warp's columns in that
// row. After the last row the accumulators are a Dv*16-byte message whose plain CRC (table kernel,
// 1/R of the data) ends the job. tools/sfold_model.py is this file's arithmetic on the CPU, checked
// against zlib.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "crc_kernels.cuh"

namespace gpu_ec {

// one offline-found stride: x^(128*dv) == sum x^k[t] (mod P_IEEE). Shape (k>>5) must be (0,0,1,1):
// that is what sfold_mul is unrolled for (own words + the next lane's two low words).
struct SfoldStride { unsigned int dv; unsigned int k[4]; };
constexpr SfoldStride kSfoldIeee[] = {
    {1651, {3, 22, 38, 49}},  {11102, {3, 14, 33, 45}}, {25291, {2, 18, 33, 47}},
    {33645, {2, 22, 41, 50}}, {47326, {7, 16, 48, 49}}, {57663, {15, 22, 48, 56}},
};
constexpr int kSfoldIeeeCount = sizeof(kSfoldIeee) / sizeof(kSfoldIeee[0]);

// geometry + fold constants shared by the streaming kernels and the finisher
struct SfoldGeom {
    unsigned int dv;       // row length in 16-byte vectors (the stride)
    unsigned int ng;       // 32-column groups per row = ceil(dv / 32)
    unsigned int rseg;     // rows per segment
    unsigned int n_seg;    // segments per buffer (rows = n_seg * rseg)
    unsigned int skip_rows;  // leading rows of segment 0 that are wholly virtual (never walked, never logged)
    unsigned int lead;     // virtual zero vectors in front of the first walked row, < dv
    unsigned int sh[4];    // k & 31 of the four terms (terms 0,1: k < 32; terms 2,3: 32 <= k < 64)
};

// A <- A * (x^k0 + x^k1 + x^k2 + x^k3) on the warp's 4096-bit row; (a.x .. a.w) are little-endian words
// at increasing addresses, the next lane holds the following 16 bytes. Lane 31 shifts in zeros.
__device__ __forceinline__ void sfold_mul(uint4 &a, const unsigned int (&sh)[4], unsigned int lane) {
    uint32_t n0 = __shfl_down_sync(0xffffffffu, a.x, 1);
    uint32_t n1 = __shfl_down_sync(0xffffffffu, a.y, 1);
    if (lane == 31) { n0 = 0; n1 = 0; }
    const uint32_t w0 = a.x, w1 = a.y, w2 = a.z, w3 = a.w;
    a.x = __funnelshift_r(w0, w1, sh[0]) ^ __funnelshift_r(w0, w1, sh[1]) ^ __funnelshift_r(w1, w2, sh[2]) ^ __funnelshift_r(w1, w2, sh[3]);
    a.y = __funnelshift_r(w1, w2, sh[0]) ^ __funnelshift_r(w1, w2, sh[1]) ^ __funnelshift_r(w2, w3, sh[2]) ^ __funnelshift_r(w2, w3, sh[3]);
    a.z = __funnelshift_r(w2, w3, sh[0]) ^ __funnelshift_r(w2, w3, sh[1]) ^ __funnelshift_r(w3, n0, sh[2]) ^ __funnelshift_r(w3, n0, sh[3]);
    a.w = __funnelshift_r(w3, n0, sh[0]) ^ __funnelshift_r(w3, n0, sh[1]) ^ __funnelshift_r(n0, n1, sh[2]) ^ __funnelshift_r(n0, n1, sh[3]);
}
__device__ __forceinline__ void xor4(uint4 &a, const uint4 &d) { a.x ^= d.x; a.y ^= d.y; a.z ^= d.z; a.w ^= d.w; }

// The same multiply with its work split over both integer pipes. SHF/LOP3/PRMT share the alu pipe, which
// issues one warp instruction per two cycles per scheduler, and IMAD has the fma pipe (same rate) to itself;
// the fused RS+CRC kernel is alu-bound (ncu: alu pipe 85 %, fma 16 %), so terms 2 and 3 are computed as
//     funnel_r(lo, hi, k) = (lo >> k) | (hi << (32-k)) = umulhi(lo, 2^(32-k)) + hi * 2^(32-k)
// (IMAD.HI + IMAD, no carry: the two parts have no bit in common). m2/m3 = 2^(32-k) of terms 2/3 (k & 31 must
// not be 0: true for every stride in the table); m2n/m3n are the same for the words that come from the next
// lane and are 0 on lane 31, which is how lane 31 shifts in zeros here.
struct SfoldMulC { uint32_t m2, m3, m2n, m3n; };
__device__ __forceinline__ SfoldMulC sfold_mul_consts(const unsigned int (&sh)[4], unsigned int lane) {
    SfoldMulC c;
    c.m2 = 1u << (32u - sh[2]);
    c.m3 = 1u << (32u - sh[3]);
    c.m2n = lane == 31 ? 0u : c.m2;
    c.m3n = lane == 31 ? 0u : c.m3;
    return c;
}
__device__ __forceinline__ uint32_t fsr_mul(uint32_t lo, uint32_t hi, uint32_t mlo, uint32_t mhi) { return __umulhi(lo, mlo) + hi * mhi; }
__device__ __forceinline__ void sfold_mul2(uint4 &a, unsigned int sh0, unsigned int sh1, const SfoldMulC &c, unsigned int lane) {
    uint32_t n0 = __shfl_down_sync(0xffffffffu, a.x, 1);
    const uint32_t n1 = __shfl_down_sync(0xffffffffu, a.y, 1);
    const uint32_t n0z = lane == 31 ? 0u : n0;  // the SHF terms need the zero itself
    const uint32_t w0 = a.x, w1 = a.y, w2 = a.z, w3 = a.w;
    a.x = __funnelshift_r(w0, w1, sh0) ^ __funnelshift_r(w0, w1, sh1) ^ fsr_mul(w1, w2, c.m2, c.m2) ^ fsr_mul(w1, w2, c.m3, c.m3);
    a.y = __funnelshift_r(w1, w2, sh0) ^ __funnelshift_r(w1, w2, sh1) ^ fsr_mul(w2, w3, c.m2, c.m2) ^ fsr_mul(w2, w3, c.m3, c.m3);
    a.z = __funnelshift_r(w2, w3, sh0) ^ __funnelshift_r(w2, w3, sh1) ^ fsr_mul(w3, n0, c.m2, c.m2n) ^ fsr_mul(w3, n0, c.m3, c.m3n);
    a.w = __funnelshift_r(w3, n0z, sh0) ^ __funnelshift_r(w3, n0z, sh1) ^ fsr_mul(n0, n1, c.m2n, c.m2n) ^ fsr_mul(n0, n1, c.m3n, c.m3n);
}

// ---------------------------------------------------------------------------------------------
// Stand-alone streaming kernel: whole-buffer CRC partials of n_buf equally long buffers.
// A job is (buffer, segment, class): `class` = blockDim.x consecutive columns of the row, `segment` = a
// range of rseg rows. Per job every thread walks its column down the segment's rows.
// Algorithmic bytes: 16 per thread per row; extra traffic 16/rseg (dump) + 8/32 per row (log).
struct SfoldParams {
    const uint8_t *base;
    unsigned long long buf_stride, n_buf;
    unsigned long long nvec;      // 16-byte vectors per buffer = ceil(len / 16)
    SfoldGeom g;
    unsigned int n_class;         // ceil(dv / blockDim.x)
    uint32_t tail_mask[4];        // keeps the bytes of the buffer's last vector that lie inside len
    uint4 *dump;                  // [n_buf * n_seg][dv]   accumulators after each segment's last row
    uint2 *qlog;                  // [n_buf * n_seg][rseg][ng]  lane-0 words before each step
};

// inactive columns (past the row's end, or in the virtual zeros in front of the buffer) read this instead of
// branching around their loads: the streaming loop stays straight-line and convergent
__device__ const uint4 g_sfold_zero = {0u, 0u, 0u, 0u};

__global__ void __launch_bounds__(512, 2) crc_sfold_kernel(const __grid_constant__ SfoldParams p) {
    const unsigned int tid = threadIdx.x, lane = tid & 31;
    const unsigned int tw = blockDim.x;
    const unsigned int dv = p.g.dv, rseg = p.g.rseg, ng = p.g.ng;
    const unsigned long long jobs_per_buf = static_cast<unsigned long long>(p.g.n_seg) * p.n_class;
    const unsigned long long n_jobs = p.n_buf * jobs_per_buf;
    const unsigned int sh[4] = {p.g.sh[0], p.g.sh[1], p.g.sh[2], p.g.sh[3]};
    const uint8_t *zero = reinterpret_cast<const uint8_t *>(&g_sfold_zero);
#pragma unroll 1
    for (unsigned long long job = blockIdx.x; job < n_jobs; job += gridDim.x) {
        const unsigned long long b = job / jobs_per_buf;
        const unsigned int r = static_cast<unsigned int>(job - b * jobs_per_buf);
        const unsigned int seg = r / p.n_class, cls = r - seg * p.n_class;
        const unsigned int o = cls * tw + tid;
        const bool active = o < dv;
        const bool logger = (lane == 0) && active;  // a group exists iff its first column does
        const uint8_t *src = p.base + b * p.buf_stride;
        const unsigned long long stream = b * p.g.n_seg + seg;
        const unsigned int s_begin = seg ? 0u : p.g.skip_rows;  // first walked row of this segment
        uint2 *ql = p.qlog + stream * rseg * ng + (o >> 5) + static_cast<unsigned long long>(s_begin + 1) * ng;  // log of row s_begin+1
        // vector index of (row s_begin of this segment, column o); only that row can reach into the virtual zeros
        const long long v0 = static_cast<long long>((static_cast<unsigned long long>(seg) * rseg + s_begin) * dv + o) -
                             static_cast<long long>(static_cast<unsigned long long>(p.g.skip_rows) * dv + p.g.lead);
        const long long last_v = static_cast<long long>(p.nvec) - 1;
        auto ld_edge = [&](long long vv) -> uint4 {  // first and last row of a segment: may be virtual, may hold the tail
            uint4 d = make_uint4(0, 0, 0, 0);
            if (active && vv >= 0) {
                d = __ldg(reinterpret_cast<const uint4 *>(src + static_cast<unsigned long long>(vv) * 16ull));
                if (vv == last_v) { d.x &= p.tail_mask[0]; d.y &= p.tail_mask[1]; d.z &= p.tail_mask[2]; d.w &= p.tail_mask[3]; }
            }
            return d;
        };
        uint4 a = ld_edge(v0);
        if (rseg - s_begin > 1) {
            // rows s_begin+1 .. rseg-2: every active column is inside the buffer (lead < dv)
            const uint8_t *ptr = active ? src + static_cast<unsigned long long>(v0 + dv) * 16ull : zero;
            const unsigned long long step = active ? static_cast<unsigned long long>(dv) * 16ull : 0ull;
            unsigned int s = s_begin + 1;
            // four rows in flight per thread; the folds are serial, the loads are not
#pragma unroll 1
            for (; s + 4 < rseg; s += 4) {
                uint4 d[4];
#pragma unroll
                for (int i = 0; i < 4; i++) d[i] = __ldg(reinterpret_cast<const uint4 *>(ptr + i * step));
                ptr += 4 * step;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    if (logger) ql[static_cast<unsigned long long>(i) * ng] = make_uint2(a.x, a.y);
                    sfold_mul(a, sh, lane);
                    xor4(a, d[i]);
                }
                ql += 4ull * ng;
            }
#pragma unroll 1
            for (; s + 1 < rseg; s++) {
                const uint4 d = __ldg(reinterpret_cast<const uint4 *>(ptr));
                ptr += step;
                if (logger) *ql = make_uint2(a.x, a.y);
                ql += ng;
                sfold_mul(a, sh, lane);
                xor4(a, d);
            }
            const uint4 d = ld_edge(v0 + static_cast<long long>(rseg - 1 - s_begin) * dv);
            if (logger) *ql = make_uint2(a.x, a.y);
            sfold_mul(a, sh, lane);
            xor4(a, d);
        }
        if (active) p.dump[stream * dv + o] = a;
    }
}

// ---------------------------------------------------------------------------------------------
// Finisher: one CTA (256 threads) per buffer at a time.
//   crc0(segment) = crc0(dump row) ^ x^(128*tail+32) * SUM_g y^(ng-1-g) * SUM_{s>=1} (x^D)^(rseg-1-s) * lost(s, g)
//   crc0(buffer)  = Horner over the segments with x^(D*rseg);   crc = crc0 * x^(-8*pad) ^ crc(zeros(len))
// crc0 = CRC with init 0 / xorout 0 (linear); y = x^4096 (one group); D = 128*dv; tail = columns of the
// last group; lost(s,g) = the (<= 63) bits the group lost at step s, as two words hm1 + hm2*x^32.
// Multiplications by the three constants that occur per log entry use byte tables in shared memory
// (built per CTA, 12 KiB); the 8-level tree over the threads uses the bitwise multiply.
struct SfoldFinishParams {
    const uint2 *qlog;
    const uint32_t *dump_crc;   // standard CRC-32 of each dump row, [n_buf * n_seg]
    uint32_t *out;              // out[b * out_stride]
    unsigned long long out_stride, n_buf;
    SfoldGeom g;
    unsigned int q_group;       // streams interleaved per log entry: 1 = [stream][s][g]; n = [stream / n][s][g][stream % n]
    unsigned int per_thread;    // groups per thread = ceil(ng / 256)
    uint32_t poly_bits;
    uint32_t xD, x32, y;        // x^(128*dv), x^32, x^4096
    uint32_t z[8];              // tree level constants: y^(per_thread * 2^L)
    uint32_t x_tail;            // x^(128*tail_cols + 32)
    uint32_t cz_dump;           // crc(zeros(dv*16)): standard crc -> crc0 of a dump row
    uint32_t x_seg;             // x^(128*dv*rseg): one segment further down the buffer
    uint32_t x_inv_pad;         // x^(-8*pad), pad = nvec*16 - len
    uint32_t cz_len;            // crc(zeros(len))
};

__device__ __forceinline__ uint32_t mul_tab(const uint32_t (*t)[256], uint32_t a) {
    return t[0][a & 255u] ^ t[1][(a >> 8) & 255u] ^ t[2][(a >> 16) & 255u] ^ t[3][a >> 24];
}

__global__ void __launch_bounds__(256) crc_sfold_finish_kernel(const __grid_constant__ SfoldFinishParams p) {
    __shared__ uint32_t tD[4][256], t32[4][256], tY[4][256];
    __shared__ uint32_t wsum[8];
    const unsigned int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t P = p.poly_bits;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const uint32_t v = t << (8 * b);
        tD[b][t] = multmodp_dev(v, p.xD, P);
        t32[b][t] = multmodp_dev(v, p.x32, P);
        tY[b][t] = multmodp_dev(v, p.y, P);
    }
    __syncthreads();
    const unsigned int ng = p.g.ng, rseg = p.g.rseg, pt = p.per_thread;
    const unsigned int off = pt * 256u - ng;  // groups are right-aligned over the 256*pt slots
    const unsigned int s0 = p.g.sh[0], s1 = p.g.sh[1], s2 = p.g.sh[2], s3 = p.g.sh[3];
    for (unsigned long long b = blockIdx.x; b < p.n_buf; b += gridDim.x) {
        uint32_t total = 0;  // crc0 of the buffer so far (thread 0 of the CTA keeps it)
        for (unsigned int seg = 0; seg < p.g.n_seg; seg++) {
            const unsigned long long stream = b * p.g.n_seg + seg;
            const uint2 *ql = p.qlog + (stream / p.q_group) * rseg * ng * p.q_group + (stream % p.q_group);
            const unsigned long long qs_g = p.q_group, qs_s = static_cast<unsigned long long>(ng) * p.q_group;
            uint32_t val = 0;
            for (unsigned int i = 0; i < pt; i++) {
                const unsigned int slot = t * pt + i;
                uint32_t T = 0;
                if (slot >= off) {
                    const unsigned int g = slot - off;
#pragma unroll 8
                    for (unsigned int s = (seg ? 0u : p.g.skip_rows) + 1; s < rseg; s++) {
                        const uint2 w = __ldg(ql + s * qs_s + g * qs_g);
                        // bits that left the group at step s: virtual words -1 (hm1) and -2 (hm2) in front of it
                        const uint32_t hm1 = (s0 ? (w.x << (32 - s0)) : 0u) ^ (s1 ? (w.x << (32 - s1)) : 0u) ^
                                             __funnelshift_r(w.x, w.y, s2) ^ __funnelshift_r(w.x, w.y, s3);
                        const uint32_t hm2 = (s2 ? (w.x << (32 - s2)) : 0u) ^ (s3 ? (w.x << (32 - s3)) : 0u);
                        T = mul_tab(tD, T) ^ hm1 ^ mul_tab(t32, hm2);
                    }
                }
                val = mul_tab(tY, val) ^ T;
            }
            // tree over the 256 threads: the right end of every block of 2^(L+1) absorbs its left half * z[L]
#pragma unroll
            for (int L = 0; L < 5; L++) {
                const uint32_t left = __shfl_up_sync(0xffffffffu, val, 1u << L);
                if ((lane & ((2u << L) - 1u)) == (2u << L) - 1u) val ^= multmodp_dev(left, p.z[L], P);
            }
            __syncthreads();
            if (lane == 31) wsum[warp] = val;
            __syncthreads();
            if (warp == 0) {
                uint32_t v = (lane < 8) ? wsum[lane] : 0u;
#pragma unroll
                for (int L = 0; L < 3; L++) {
                    const uint32_t left = __shfl_up_sync(0xffffffffu, v, 1u << L);
                    if ((lane & ((2u << L) - 1u)) == (2u << L) - 1u) v ^= multmodp_dev(left, p.z[5 + L], P);
                }
                if (lane == 7) {
                    const uint32_t seg0 = (p.dump_crc[stream] ^ p.cz_dump) ^ multmodp_dev(v, p.x_tail, P);
                    total = multmodp_dev(total, p.x_seg, P) ^ seg0;
                }
            }
        }
        if (t == 7) p.out[b * p.out_stride] = multmodp_dev(total, p.x_inv_pad, P) ^ p.cz_len;
    }
}

}  // namespace gpu_ec
