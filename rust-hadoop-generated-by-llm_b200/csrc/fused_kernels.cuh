// fused_kernels.cuh — RS encode + whole-shard CRC-32 of all k+m shards in ONE pass over HBM.
//
// Replaces the client write sequence dfs/client/src/mod.rs:600-611 (erasure::encode, then a
// crc32fast::Hasher over each of the k+m shards, then one over the original buffer), which reads
// the same bytes three times. Here every data byte is loaded once, every parity byte stored once,
// and both are checksummed while they sit in registers.
//
// RS side: the 2-CTA-cluster byte-LUT kernel of rs_kernels.cuh (one shared-memory lookup per input
// byte, tables split over an SM pair, partial parity traded over distributed shared memory).
// CRC side: the sparse-fold accumulators of sfold_kernels.cuh — no tables, no shared memory, only
// SHF/LOP3/SHFL on registers — which is what lets both run on one SM (a table CRC needs 128 KiB of the
// same shared memory the RS tables fill). Each CTA folds the data shards it loads and the parity rows
// it finishes: 7 of the 14 streams of RS(10,4) per SM.
//
// Tile walk: a job is (stripe, class): NT consecutive columns of the stride-Dv row structure; the
// cluster walks the job's rows top to bottom (row s = vectors s*Dv - lead + column), so a thread
// always meets the same column and its accumulators follow one stream down the shard. After the last
// row they are dumped (16 bytes per column per shard) and cleared. Lost bits are logged by lane 0.
// crc_sfold_finish_kernel turns dump + log into the k+m CRCs; CRC(original) follows by 32-bit algebra
// (stripe_full_crc_kernel).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "rs_kernels.cuh"
#include "sfold_kernels.cuh"

namespace gpu_ec {

struct FusedParams {
    RsParams rs;                 // base, strides, nvec, coefficient rows, slot offsets (n_tiles/tiles_per_stripe unused)
    SfoldGeom g;                 // n_seg = 1, skip_rows = 0: a shard is one segment of g.rseg rows
    unsigned int n_class;        // ceil(dv / NT)
    unsigned int n_shards;       // k + m: stream index = stripe * n_shards + slot
    unsigned long long n_jobs;   // n_stripes * n_class
    uint32_t tail_mask[4];       // keeps the bytes of a shard's last vector that lie inside shard_len
    int prefetch;                // plain apply only: L2-prefetch the tile after next (helps short bursts, hurts sustained runs)
    uint4 *dump;                 // [n_stripes * n_shards][dv]
    uint2 *qlog;                 // [n_stripes][rseg][ng][n_shards]: lane 0 of a warp logs its streams side by side
};

#ifndef GPU_EC_FUSED_THREADS
#define GPU_EC_FUSED_THREADS 384
#endif
#ifndef GPU_EC_SFOLD_MODE
#define GPU_EC_SFOLD_MODE 0   // 0: funnel shifts on the alu pipe; 1: terms 2,3 as IMAD.HI + IMAD on the fma pipe
#endif
constexpr int FUSED_THREADS = GPU_EC_FUSED_THREADS;  // 12 warps: 168 registers per thread for 7 accumulators + two tiles of inputs
__host__ __device__ constexpr int fused_xbuf_bytes(int nt) { return 2 * 2 * nt * 16; }
__host__ __device__ constexpr int fused_mbar_bytes(int nt) { return 2 * 2 * (nt / 32) * 8; }
constexpr int fused_smem_bytes(int kh, int nt) { return kh <= 4 ? LUT_FAST_RS_BYTES : kh * 256 * 32 * 4 + fused_xbuf_bytes(nt) + fused_mbar_bytes(nt); }

__device__ __forceinline__ void and4(uint4 &a, uint32_t m) { a.x &= m; a.y &= m; a.z &= m; a.w &= m; }

// tile position of one thread. flags: 1 = this thread's vector exists, 2 = edge tile (first row or last
// class: some threads have no vector, or the shard's ragged tail is here), 4 = last row of the job
struct FusedPos {
    const uint8_t *stripe;
    unsigned int v, o, s, stripe_i, flags;
};

template <int KH, int NT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT, 1) rs_lut8x2_crc_kernel(const __grid_constant__ FusedParams fp) {
    extern __shared__ __align__(16) uint32_t rs_tab[];  // tables | exchange buffer | mbarriers
    const RsParams &p = fp.rs;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int NW = NT / 32;
    constexpr int XBUF_BYTES = fused_xbuf_bytes(NT);
    const uint32_t rank = cluster_ctarank();
    const int j_base = static_cast<int>(rank) * KH;
    const int ni = (p.n_in - j_base < KH) ? (p.n_in - j_base) : KH;
    const int h0 = (p.n_out + 1) >> 1;
    const int own0 = rank ? h0 : 0, n_own = rank ? (p.n_out - h0) : h0;
    const int oth0 = rank ? 0 : h0, n_oth = rank ? h0 : (p.n_out - h0);

    // ---- tables, exchange buffer, mbarriers: as rs_lut8x2_kernel
    constexpr bool FAST = (KH <= 4);
    const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(rs_tab));
    const uint32_t wb = s0 & 0xFF000000u;
    if (FAST && (s0 - wb) + 128u + XBUF_BYTES + fused_mbar_bytes(NT) > LUT_BLK_A) __trap();
    for (int e0 = warp * 32; e0 < KH * 256; e0 += NW * 32) {
        const int e = e0 + lane;
        const int j = e >> 8;
        const uint32_t v = (j < ni) ? packed_products(p, j_base + j, static_cast<uint32_t>(e & 255)) : 0u;
#pragma unroll 8
        for (int q = 0; q < 32; q++) {
            const int eq = e0 + q;
            const uint32_t addr = FAST ? wb + lut_fast_addr(eq >> 8, static_cast<uint32_t>(eq & 255), lane) : s0 + (eq * 32 + lane) * 4;
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(__shfl_sync(0xffffffffu, v, q)) : "memory");
        }
    }
    const uint32_t xbuf_s = FAST ? ((s0 + 127u) & ~127u) : s0 + KH * 256 * 32 * 4;
    const uint4 *xbuf = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(rs_tab) + (xbuf_s - s0));
    const uint32_t mbar_s = xbuf_s + XBUF_BYTES;
    if (threadIdx.x < 4 * NW) mbar_init(mbar_s + threadIdx.x * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    cluster_sync_all();
    const uint32_t xbuf_peer = map_to_peer(xbuf_s, rank ^ 1u);
    const uint32_t mbar_peer = map_to_peer(mbar_s, rank ^ 1u);
    const uint32_t my_full = mbar_s + warp * 8, my_empty = mbar_s + (2 * NW + warp) * 8;
    const uint32_t peer_full = mbar_peer + warp * 8, peer_empty = mbar_peer + (2 * NW + warp) * 8;
    const uint32_t tl = s0 + lane * 4;
    const uint32_t lane4 = wb | static_cast<uint32_t>(lane * 4);

    // ---- the walk: jobs cluster, cluster + n_clusters, ...; rows 0 .. R-1 inside a job
    const unsigned long long n_clusters = gridDim.x >> 1;
    const unsigned long long cluster = blockIdx.x >> 1;
    const unsigned int dv = fp.g.dv, R = fp.g.rseg, ng = fp.g.ng, lead = fp.g.lead;
    const unsigned int last_vec = p.nvec - 1;
    const unsigned long long my_jobs = cluster < fp.n_jobs ? (fp.n_jobs - cluster + n_clusters - 1) / n_clusters : 0ull;
    const unsigned long long my_tiles = my_jobs * R;
    const unsigned int sh[4] = {fp.g.sh[0], fp.g.sh[1], fp.g.sh[2], fp.g.sh[3]};

    unsigned long long it_job = cluster;
    unsigned int it_s = 0;
    unsigned long long it_stripe = fp.n_class ? it_job / fp.n_class : 0;
    unsigned int it_cls = static_cast<unsigned int>(it_job - it_stripe * fp.n_class);
    auto locate_next = [&]() -> FusedPos {
        FusedPos q;
        q.o = it_cls * NT + threadIdx.x;
        q.s = it_s;
        const long long vv = static_cast<long long>(it_s) * dv + q.o - static_cast<long long>(lead);
        const bool ok = (q.o < dv) && (vv >= 0);
        q.v = ok ? static_cast<unsigned int>(vv) : (vv > static_cast<long long>(last_vec) ? last_vec : 0u);
        q.stripe = p.base + it_stripe * p.stripe_stride;
        q.stripe_i = static_cast<unsigned int>(it_stripe);
        q.flags = (ok ? 1u : 0u) | ((it_s == 0 || it_cls + 1 == fp.n_class) ? 2u : 0u) | ((it_s + 1 == R) ? 4u : 0u);
        if (++it_s == R) {
            it_s = 0;
            it_job += n_clusters;
            it_stripe = it_job / fp.n_class;
            it_cls = static_cast<unsigned int>(it_job - it_stripe * fp.n_class);
        }
        return q;
    };
    unsigned int my_in_off[KH];
#pragma unroll
    for (int j = 0; j < KH; j++) my_in_off[j] = p.in_off16[(j < ni) ? (j_base + j) : j_base];
    const unsigned int own_off0 = p.out_off16[n_own > 0 ? own0 : 0], own_off1 = p.out_off16[n_own > 1 ? own0 + 1 : 0];
    auto load = [&](uint4 (&d)[KH], const FusedPos &q) {
#pragma unroll
        for (int j = 0; j < KH; j++) d[j] = ld_vec(q.stripe + static_cast<unsigned long long>(q.v + my_in_off[j]) * 16ull);
    };

    // ---- sparse-fold CRC state: the data shards this CTA loads + the parity rows it finishes
    uint4 A[KH], AP0 = make_uint4(0, 0, 0, 0), AP1 = AP0;
#pragma unroll
    for (int j = 0; j < KH; j++) A[j] = make_uint4(0, 0, 0, 0);
    // lost-bit log of a tile: entry ((stripe*R + s)*ng + group) holds one uint2 per shard slot, so a warp's
    // lane 0 writes its streams' words side by side (one or two sectors per tile, immediate offsets)
    auto log_row = [&](const FusedPos &q) -> uint2 * {
        return fp.qlog + ((static_cast<unsigned long long>(q.stripe_i) * R + q.s) * ng + (q.o >> 5)) * fp.n_shards;
    };
    // one stream, one tile, common case (every thread has a vector, not the job's last row):
    // log (lane 0), multiply by x^(128*dv), add the tile's vector. No branches, no address math.
    auto crc_fast = [&](uint4 &acc, const uint4 &d, uint2 *lp) {
        if (lane == 0) *lp = make_uint2(acc.x, acc.y);
        sfold_mul(acc, sh, lane);
        xor4(acc, d);
    };
    // edge tile (first row / last class: threads without a vector add nothing, the ragged tail is cut at
    // shard_len) or last row of the job (dump the accumulator, start the next job from zero)
    auto crc_slow = [&](uint4 &acc, uint4 d, const FusedPos &q, uint2 *lp, unsigned int slot) {
        and4(d, (q.flags & 1u) ? 0xFFFFFFFFu : 0u);
        if (q.v == last_vec) { d.x &= fp.tail_mask[0]; d.y &= fp.tail_mask[1]; d.z &= fp.tail_mask[2]; d.w &= fp.tail_mask[3]; }
        if (lane == 0 && q.o < dv) *lp = make_uint2(acc.x, acc.y);
        sfold_mul(acc, sh, lane);
        xor4(acc, d);
        if (q.flags & 4u) {
            if (q.o < dv) fp.dump[(static_cast<unsigned long long>(q.stripe_i) * fp.n_shards + slot) * dv + q.o] = acc;
            acc = make_uint4(0, 0, 0, 0);
        }
    };

    uint4 cur[KH];
    FusedPos cur_pos = {p.base, 0u, 0u, 0u, 0u, 0u};
    if (my_tiles) { cur_pos = locate_next(); load(cur, cur_pos); }
    uint4 keep0 = make_uint4(0, 0, 0, 0), keep1 = keep0;  // the rows this CTA owns, from the previous tile
    FusedPos keep_pos = cur_pos;
    bool have_prev = false;
    unsigned int it = 0;
    const unsigned int slot_own0 = static_cast<unsigned int>(p.out_idx[n_own > 0 ? own0 : 0]);
    // finish tile it-1: wait for the partner's partial rows, fold, store the parity, checksum it, free the slot
    auto finish_prev = [&]() {
        const unsigned int ps = (it - 1) & 1u, pu = (it - 1) >> 1;
        if (n_own > 0) {
            mbar_wait(my_full + ps * (NW * 8), pu & 1u);
            const uint4 *x = xbuf + ps * (2 * NT) + threadIdx.x;
            uint4 o0 = x[0], o1 = make_uint4(0, 0, 0, 0);
            if (n_own > 1) o1 = x[NT];
            xor4(o0, keep0);
            xor4(o1, keep1);
            __syncwarp();
            if (lane == 0) mbar_arrive_peer(peer_empty + ps * (NW * 8));
            if (keep_pos.flags & 1u) {
                uint8_t *ws = const_cast<uint8_t *>(keep_pos.stripe);
                *reinterpret_cast<uint4 *>(ws + static_cast<unsigned long long>(keep_pos.v + own_off0) * 16ull) = o0;
                if (n_own > 1) *reinterpret_cast<uint4 *>(ws + static_cast<unsigned long long>(keep_pos.v + own_off1) * 16ull) = o1;
            }
#ifndef GPU_EC_FUSED_NOCRC
            uint2 *lp = log_row(keep_pos) + slot_own0;
            if ((keep_pos.flags & 6u) == 0u) {
                crc_fast(AP0, o0, lp);
                if (n_own > 1) crc_fast(AP1, o1, lp + 1);
            } else {
                crc_slow(AP0, o0, keep_pos, lp, slot_own0);
                if (n_own > 1) crc_slow(AP1, o1, keep_pos, lp + 1, slot_own0 + 1);
            }
#endif
        }
    };

    for (unsigned long long ti = 0; ti < my_tiles; ti++, it++) {
        uint4 nxt[KH];
        FusedPos nxt_pos = cur_pos;
        if (ti + 1 < my_tiles) { nxt_pos = locate_next(); load(nxt, nxt_pos); }

        uint32_t acc[16];
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] = 0;
        rs_accumulate<KH, FAST>(cur, acc, tl, lane4);
        uint32_t o[4][4];  // o[q][row]
#pragma unroll
        for (int q = 0; q < 4; q++) transpose4x4(&acc[q * 4], o[q]);

        const unsigned int st_ = it & 1u, use_ = it >> 1;
        if (n_own > 0 && lane == 0) mbar_expect_tx(my_full + st_ * (NW * 8), static_cast<uint32_t>(n_own) * 32u * 16u);
        if (n_oth > 0) mbar_wait(my_empty + st_ * (NW * 8), (use_ & 1u) ^ 1u);
        const uint32_t dst = xbuf_peer + (st_ * (2 * NT) + threadIdx.x) * 16u;
        const uint32_t sig = peer_full + st_ * (NW * 8);
        uint4 own_a, own_b;
        if (p.n_out == RS_MAX_OUT) {
            const uint4 r0 = make_uint4(o[0][0], o[1][0], o[2][0], o[3][0]), r1 = make_uint4(o[0][1], o[1][1], o[2][1], o[3][1]);
            const uint4 r2 = make_uint4(o[0][2], o[1][2], o[2][2], o[3][2]), r3 = make_uint4(o[0][3], o[1][3], o[2][3], o[3][3]);
            const bool hi = rank != 0;
            st_async_peer(dst, hi ? r0 : r2, sig);
            st_async_peer(dst + NT * 16, hi ? r1 : r3, sig);
            own_a = hi ? r2 : r0;
            own_b = hi ? r3 : r1;
        } else {
            own_a = own_b = make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int row = 0; row < RS_MAX_OUT; row++) {
                const uint4 v = make_uint4(o[0][row], o[1][row], o[2][row], o[3][row]);
                if (n_oth > 0 && row == oth0) st_async_peer(dst, v, sig);
                if (n_oth > 1 && row == oth0 + 1) st_async_peer(dst + NT * 16, v, sig);
                if (n_own > 0 && row == own0) own_a = v;
                if (n_own > 1 && row == own0 + 1) own_b = v;
            }
        }
        // checksum the data shards of this tile while the partner's rows are in flight
#ifndef GPU_EC_FUSED_NOCRC
        {
            uint2 *lp = log_row(cur_pos) + j_base;  // data slot of input j is j (encode: in_idx[j] == j)
            if ((cur_pos.flags & 6u) == 0u) {
#pragma unroll
                for (int j = 0; j < KH; j++)
                    if (j < ni) crc_fast(A[j], cur[j], lp + j);
            } else {
#pragma unroll
                for (int j = 0; j < KH; j++)
                    if (j < ni) crc_slow(A[j], cur[j], cur_pos, lp + j, static_cast<unsigned int>(j_base + j));
            }
        }
#endif

        if (have_prev) finish_prev();
        keep0 = own_a;
        keep1 = own_b;
        keep_pos = cur_pos;
        have_prev = true;
#pragma unroll
        for (int j = 0; j < KH; j++) cur[j] = nxt[j];
        cur_pos = nxt_pos;
    }
    if (have_prev) finish_prev();
    cluster_sync_all();  // neither CTA may retire while its partner can still write into it or signal it
}

// ---------------------------------------------------------------------------------------------
// The same single pass, specialised and trimmed for the shapes that matter (n_in == 2*KH, n_out == 4:
// RS(10,4), RS(8,4)). ncu on the generic kernel above showed it bound by the half-rate integer ALU
// pipe and by issue slots (853 warp instructions per tile, a third of them bookkeeping), so this one
// removes bookkeeping rather than memory traffic:
//   * no second register tile: a shard's next vector is loaded into the registers of the current
//     one as soon as its 16 lookups are done (the CRC of the tile runs first), so there is no
//     cur = nxt copy and no copy of the tile position;
//   * job-level state (stripe pointer, column, log pointer) lives in registers and changes once per
//     R tiles; the parity pipeline never crosses a job boundary (the last tile of a job is finished
//     at once), so the previous tile's position is just "one row up";
//   * the packed products are built per rank with the rows to SEND in bytes 0,1 and the rows to KEEP
//     in bytes 2,3, so nothing selects by rank in the loop;
//   * tables 0..3 use the single-PRMT layout (the 24 KiB exchange buffer of a 384-thread CTA fits
//     below them), table 4 the classic one: 64 fewer address IMADs per tile;
//   * n_in, n_out, the rows owned are compile-time: no uniform-predicate tests per stream.
__device__ __forceinline__ uint32_t packed_products_rot(const RsParams &p, int j, uint32_t x, int rot) {
    uint32_t c = 0;
#pragma unroll
    for (int r = 0; r < RS_MAX_OUT; r++) c |= static_cast<uint32_t>(p.coef[(r + rot) & 3][j]) << (8 * r);
    uint32_t v = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        v ^= ((x >> i) & 1u) ? c : 0u;
        uint32_t hi;
        asm("prmt.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(c), "r"(0u), "r"(0xBA98u));
        c = ((c & 0x7f7f7f7fu) << 1) ^ (hi & 0x1d1d1d1du);
    }
    return v;
}

// the same for either table layout (rs_lut8's classic [K][256][32] when K > 5)
template <int K, bool FAST, int J, typename F>
__device__ __forceinline__ void rs_accumulate_reload_any(uint4 (&d)[K], uint32_t (&acc)[16], uint32_t tl, uint32_t lane4, F &&reload) {
    if constexpr (J < K) {
        const uint32_t w[4] = {d[J].x, d[J].y, d[J].z, d[J].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if constexpr (FAST) {
                acc[q * 4 + 0] ^= lut_fast<J, 0>(lane4, w[q]);
                acc[q * 4 + 1] ^= lut_fast<J, 1>(lane4, w[q]);
                acc[q * 4 + 2] ^= lut_fast<J, 2>(lane4, w[q]);
                acc[q * 4 + 3] ^= lut_fast<J, 3>(lane4, w[q]);
            } else {
                const uint32_t tj = tl + J * (256 * 32 * 4);
                acc[q * 4 + 0] ^= lut_lds<0>(tj, byte_of<0>(w[q]));
                acc[q * 4 + 1] ^= lut_lds<0>(tj, byte_of<1>(w[q]));
                acc[q * 4 + 2] ^= lut_lds<0>(tj, byte_of<2>(w[q]));
                acc[q * 4 + 3] ^= lut_lds<0>(tj, w[q] >> 24);
            }
        }
        reload(J);
        rs_accumulate_reload_any<K, FAST, J + 1>(d, acc, tl, lane4, reload);
    }
}

// one table word of shard J (compile time), byte I of w, either layout
template <int J, int I, bool FAST>
__device__ __forceinline__ uint32_t rs_lut_word(uint32_t tl, uint32_t lane4, uint32_t w) {
    if constexpr (FAST) return lut_fast<J, I>(lane4, w);
    else return lut_lds<0>(tl + J * (256 * 32 * 4), I == 3 ? (w >> 24) : byte_of<I>(w));
}
// all lookups of a tile, XORs merged three table words at a time (one LOP3 each): shards J, J+1, J+2 per step;
// FIRST starts the accumulators instead of adding to them. `reload(j)` is called once shard j's registers are dead.
template <int K, bool FAST, int J, typename F>
__device__ __forceinline__ void rs_lookup_merged3(uint4 (&d)[K], uint32_t (&acc)[16], uint32_t tl, uint32_t lane4, F &&reload) {
    if constexpr (J < K) {
        constexpr int N = (K - J >= 3) ? 3 : (K - J);
        const uint32_t a[4] = {d[J].x, d[J].y, d[J].z, d[J].w};
        const uint32_t b[4] = {d[J + (N > 1 ? 1 : 0)].x, d[J + (N > 1 ? 1 : 0)].y, d[J + (N > 1 ? 1 : 0)].z, d[J + (N > 1 ? 1 : 0)].w};
        const uint32_t c[4] = {d[J + (N > 2 ? 2 : 0)].x, d[J + (N > 2 ? 2 : 0)].y, d[J + (N > 2 ? 2 : 0)].z, d[J + (N > 2 ? 2 : 0)].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint32_t t0 = rs_lut_word<J, 0, FAST>(tl, lane4, a[q]), t1 = rs_lut_word<J, 1, FAST>(tl, lane4, a[q]);
            uint32_t t2 = rs_lut_word<J, 2, FAST>(tl, lane4, a[q]), t3 = rs_lut_word<J, 3, FAST>(tl, lane4, a[q]);
            if constexpr (N > 1) {
                t0 ^= rs_lut_word<J + 1, 0, FAST>(tl, lane4, b[q]); t1 ^= rs_lut_word<J + 1, 1, FAST>(tl, lane4, b[q]);
                t2 ^= rs_lut_word<J + 1, 2, FAST>(tl, lane4, b[q]); t3 ^= rs_lut_word<J + 1, 3, FAST>(tl, lane4, b[q]);
            }
            if constexpr (N > 2) {
                t0 ^= rs_lut_word<J + 2, 0, FAST>(tl, lane4, c[q]); t1 ^= rs_lut_word<J + 2, 1, FAST>(tl, lane4, c[q]);
                t2 ^= rs_lut_word<J + 2, 2, FAST>(tl, lane4, c[q]); t3 ^= rs_lut_word<J + 2, 3, FAST>(tl, lane4, c[q]);
            }
            if constexpr (J == 0) { acc[q * 4 + 0] = t0; acc[q * 4 + 1] = t1; acc[q * 4 + 2] = t2; acc[q * 4 + 3] = t3; }
            else { acc[q * 4 + 0] ^= t0; acc[q * 4 + 1] ^= t1; acc[q * 4 + 2] ^= t2; acc[q * 4 + 3] ^= t3; }
        }
        reload(J);
        if constexpr (N > 1) reload(J + 1);
        if constexpr (N > 2) reload(J + 2);
        rs_lookup_merged3<K, FAST, J + 3>(d, acc, tl, lane4, reload);
    }
}

// lookups of input shard J, then `reload(J)`: d[J] is dead from here on, the next tile's vector may land in it
template <int KH, int J, typename F>
__device__ __forceinline__ void rs_accumulate_reload(uint4 (&d)[KH], uint32_t (&acc)[16], uint32_t lane4, F &&reload) {
    if constexpr (J < KH) {
        const uint32_t w[4] = {d[J].x, d[J].y, d[J].z, d[J].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            acc[q * 4 + 0] ^= lut_fast<J, 0>(lane4, w[q]);
            acc[q * 4 + 1] ^= lut_fast<J, 1>(lane4, w[q]);
            acc[q * 4 + 2] ^= lut_fast<J, 2>(lane4, w[q]);
            acc[q * 4 + 3] ^= lut_fast<J, 3>(lane4, w[q]);
        }
        reload(J);
        rs_accumulate_reload<KH, J + 1>(d, acc, lane4, reload);
    }
}

// all lookups of a tile with the XORs merged three at a time (one LOP3 per three table words): shards 0,1,2
// start the accumulators, the rest are folded in; `reload(j)` is called once shard j's registers are dead
template <int KH, typename F>
__device__ __forceinline__ void rs_lookup_merged(uint4 (&d)[KH], uint32_t (&acc)[16], uint32_t lane4, F &&reload) {
    static_assert(KH == 4 || KH == 5, "");
    {
        const uint32_t a[4] = {d[0].x, d[0].y, d[0].z, d[0].w}, b[4] = {d[1].x, d[1].y, d[1].z, d[1].w}, c[4] = {d[2].x, d[2].y, d[2].z, d[2].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            acc[q * 4 + 0] = lut_fast<0, 0>(lane4, a[q]) ^ lut_fast<1, 0>(lane4, b[q]) ^ lut_fast<2, 0>(lane4, c[q]);
            acc[q * 4 + 1] = lut_fast<0, 1>(lane4, a[q]) ^ lut_fast<1, 1>(lane4, b[q]) ^ lut_fast<2, 1>(lane4, c[q]);
            acc[q * 4 + 2] = lut_fast<0, 2>(lane4, a[q]) ^ lut_fast<1, 2>(lane4, b[q]) ^ lut_fast<2, 2>(lane4, c[q]);
            acc[q * 4 + 3] = lut_fast<0, 3>(lane4, a[q]) ^ lut_fast<1, 3>(lane4, b[q]) ^ lut_fast<2, 3>(lane4, c[q]);
        }
        reload(0); reload(1); reload(2);
    }
    if constexpr (KH == 5) {
        const uint32_t a[4] = {d[3].x, d[3].y, d[3].z, d[3].w}, b[4] = {d[4].x, d[4].y, d[4].z, d[4].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            acc[q * 4 + 0] ^= lut_fast<3, 0>(lane4, a[q]) ^ lut_fast<4, 0>(lane4, b[q]);
            acc[q * 4 + 1] ^= lut_fast<3, 1>(lane4, a[q]) ^ lut_fast<4, 1>(lane4, b[q]);
            acc[q * 4 + 2] ^= lut_fast<3, 2>(lane4, a[q]) ^ lut_fast<4, 2>(lane4, b[q]);
            acc[q * 4 + 3] ^= lut_fast<3, 3>(lane4, a[q]) ^ lut_fast<4, 3>(lane4, b[q]);
        }
        reload(3); reload(4);
    } else {
        const uint32_t a[4] = {d[3].x, d[3].y, d[3].z, d[3].w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            acc[q * 4 + 0] ^= lut_fast<3, 0>(lane4, a[q]);
            acc[q * 4 + 1] ^= lut_fast<3, 1>(lane4, a[q]);
            acc[q * 4 + 2] ^= lut_fast<3, 2>(lane4, a[q]);
            acc[q * 4 + 3] ^= lut_fast<3, 3>(lane4, a[q]);
        }
        reload(3);
    }
}

// CRC = false: the same walk and exchange as a plain apply-matrix kernel (encode, or reconstruct of 4 shards)
template <int KH, int NT, bool CRC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT, 1) rs_x2_crc_fast_kernel(const __grid_constant__ FusedParams fp) {
    static_assert(KH == 4 || KH == 5, "single-PRMT layout: tables 0..3 in the two 64 KiB blocks, table 4 below them");
    extern __shared__ __align__(16) uint32_t rs_tab[];
    const RsParams &p = fp.rs;
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = NT / 32;
    constexpr int XBUF_BYTES = fused_xbuf_bytes(NT);
    const uint32_t rank = cluster_ctarank();
    const int j_base = static_cast<int>(rank) * KH;
    const int own0 = rank ? 2 : 0;  // rank 0 finishes parity rows 0,1; rank 1 rows 2,3

    // ---- tables (rows to send in bytes 0,1; rows to keep in bytes 2,3), exchange buffer, mbarriers
    const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(rs_tab));
    const uint32_t wb = s0 & 0xFF000000u;
    if ((s0 - wb) + 128u + XBUF_BYTES + fused_mbar_bytes(NT) > (KH == 5 ? LUT_FRAG1 : LUT_BLK_A)) __trap();
    const int rot = rank ? 0 : 2;
    for (int e0 = warp * 32; e0 < KH * 256; e0 += NW * 32) {
        const int e = e0 + lane;
        const uint32_t v = packed_products_rot(p, j_base + (e >> 8), static_cast<uint32_t>(e & 255), rot);
#pragma unroll 8
        for (int q = 0; q < 32; q++) {
            const int eq = e0 + q;
            const uint32_t addr = wb + lut_fast_addr(eq >> 8, static_cast<uint32_t>(eq & 255), lane);
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(__shfl_sync(0xffffffffu, v, q)) : "memory");
        }
    }
    const uint32_t xbuf_s = (s0 + 127u) & ~127u;
    const uint4 *xbuf = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned char *>(rs_tab) + (xbuf_s - s0));
    const uint32_t mbar_s = xbuf_s + XBUF_BYTES;
    if (threadIdx.x < 4 * NW) mbar_init(mbar_s + threadIdx.x * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    cluster_sync_all();
    const uint32_t xbuf_peer = map_to_peer(xbuf_s, rank ^ 1u);
    const uint32_t mbar_peer = map_to_peer(mbar_s, rank ^ 1u);
    const uint32_t my_full = mbar_s + warp * 8, my_empty = mbar_s + (2 * NW + warp) * 8;
    const uint32_t peer_full = mbar_peer + warp * 8, peer_empty = mbar_peer + (2 * NW + warp) * 8;
    const uint32_t lane4 = wb | (lane * 4u);

    const unsigned long long n_clusters = gridDim.x >> 1;
    const unsigned long long cluster = blockIdx.x >> 1;
    const unsigned int dv = fp.g.dv, R = fp.g.rseg, ng = fp.g.ng, lead = fp.g.lead;
    const unsigned int last_vec = p.nvec - 1;
    const unsigned long long my_jobs = cluster < fp.n_jobs ? (fp.n_jobs - cluster + n_clusters - 1) / n_clusters : 0ull;
    const unsigned long long my_tiles = my_jobs * R;
    const unsigned int sh[4] = {fp.g.sh[0], fp.g.sh[1], fp.g.sh[2], fp.g.sh[3]};
    const SfoldMulC mc = sfold_mul_consts(sh, lane);
    const unsigned long long lp_stride = static_cast<unsigned long long>(ng) * fp.n_shards;  // log entries per row
    unsigned int my_in_off[KH];
#pragma unroll
    for (int j = 0; j < KH; j++) my_in_off[j] = p.in_off16[j_base + j];
    const unsigned int own_off0 = p.out_off16[own0], own_off1 = p.out_off16[own0 + 1];
    const unsigned int slot_own0 = static_cast<unsigned int>(p.out_idx[own0]);

    // ---- job state (changes once per R tiles) and tile state
    struct Job {
        const uint8_t *stripe;   // stripe base
        uint2 *lp;               // log entry of (stripe, row 0, this warp's group), slot 0
        unsigned int o, stripe_i;
        bool col, lastcls;       // column exists (o < dv); last class of the row (some columns do not)
    };
    auto setup_job = [&](unsigned long long job) -> Job {
        Job J;
        const unsigned long long st = job / fp.n_class;
        const unsigned int cls = static_cast<unsigned int>(job - st * fp.n_class);
        J.stripe_i = static_cast<unsigned int>(st);
        J.o = cls * NT + threadIdx.x;
        J.col = J.o < dv;
        J.lastcls = (cls + 1 == fp.n_class);
        J.stripe = p.base + st * p.stripe_stride;
        J.lp = fp.qlog + ((st * R) * ng + (J.o >> 5)) * fp.n_shards;
        return J;
    };

    uint4 A[KH], AP0 = make_uint4(0, 0, 0, 0), AP1 = AP0;
#pragma unroll
    for (int j = 0; j < KH; j++) A[j] = make_uint4(0, 0, 0, 0);
    auto crc_fast = [&](uint4 &acc, const uint4 &d, uint2 *lp) {
        if (lane == 0) *lp = make_uint2(acc.x, acc.y);
        if constexpr (GPU_EC_SFOLD_MODE == 1) sfold_mul2(acc, fp.g.sh[0], fp.g.sh[1], mc, lane);
        else sfold_mul(acc, fp.g.sh, lane);
        xor4(acc, d);
    };
    // first row / last class / last row: threads without a vector add nothing, the ragged tail is cut at
    // shard_len, and after the job's last row the accumulator is dumped and cleared
    auto crc_slow = [&](uint4 &acc, uint4 d, const Job &J, bool ok, unsigned int v, bool last, uint2 *lp, unsigned int slot) {
        and4(d, ok ? 0xFFFFFFFFu : 0u);
        if (v == last_vec) { d.x &= fp.tail_mask[0]; d.y &= fp.tail_mask[1]; d.z &= fp.tail_mask[2]; d.w &= fp.tail_mask[3]; }
        if (lane == 0 && J.col) *lp = make_uint2(acc.x, acc.y);
        if constexpr (GPU_EC_SFOLD_MODE == 1) sfold_mul2(acc, fp.g.sh[0], fp.g.sh[1], mc, lane);
        else sfold_mul(acc, fp.g.sh, lane);
        xor4(acc, d);
        if (last) {
            if (J.col) fp.dump[(static_cast<unsigned long long>(J.stripe_i) * fp.n_shards + slot) * dv + J.o] = acc;
            acc = make_uint4(0, 0, 0, 0);
        }
    };

    uint4 cur[KH];
    uint4 keep0 = make_uint4(0, 0, 0, 0), keep1 = keep0;  // rows 2,3 of the packed products: the parity rows this CTA finishes
    Job J = setup_job(cluster < fp.n_jobs ? cluster : 0ull);
    unsigned long long job = cluster;
    unsigned int s = 0;                       // row inside the job
    bool ok = J.col && J.o >= lead;           // this thread has a vector in the current tile
    unsigned int v = ok ? J.o - lead : 0u;    // its index in the shard (0 when it has none: a harmless address)
    uint2 *lp = J.lp;                         // log entry of the current row
    bool have_prev = false, keep_ok = false;
    unsigned int keep_v = 0;
    unsigned int it = 0;
    if (my_tiles) {
#pragma unroll
        for (int j = 0; j < KH; j++) cur[j] = ld_vec(J.stripe + static_cast<unsigned long long>(v + my_in_off[j]) * 16ull);
    }
    // finish exchange tile `tit` (row `row` of job J): wait for the partner's halves, fold, store the two
    // parity vectors, checksum them, hand the slot back
    auto finish = [&](unsigned int tit, const Job &JJ, unsigned int row, bool last, bool kok, unsigned int kv, uint2 *klp) {
        const unsigned int ps = tit & 1u, pu = tit >> 1;
        mbar_wait(my_full + ps * (NW * 8), pu & 1u);
        const uint4 *x = xbuf + ps * (2 * NT) + threadIdx.x;
        uint4 p0 = x[0], p1 = x[NT];
        xor4(p0, keep0);
        xor4(p1, keep1);
        __syncwarp();
        if (lane == 0) mbar_arrive_peer(peer_empty + ps * (NW * 8));
        // pairs (lo = positions 4q, 4q+1; hi = 4q+2, 4q+3) -> the two finished parity rows
        const uint4 o0 = make_uint4(__byte_perm(p0.x, p0.y, 0x6420), __byte_perm(p0.z, p0.w, 0x6420), __byte_perm(p1.x, p1.y, 0x6420), __byte_perm(p1.z, p1.w, 0x6420));
        const uint4 o1 = make_uint4(__byte_perm(p0.x, p0.y, 0x7531), __byte_perm(p0.z, p0.w, 0x7531), __byte_perm(p1.x, p1.y, 0x7531), __byte_perm(p1.z, p1.w, 0x7531));
        if (kok) {
            uint8_t *ws = const_cast<uint8_t *>(JJ.stripe);
            *reinterpret_cast<uint4 *>(ws + static_cast<unsigned long long>(kv + own_off0) * 16ull) = o0;
            *reinterpret_cast<uint4 *>(ws + static_cast<unsigned long long>(kv + own_off1) * 16ull) = o1;
        }
        if constexpr (CRC) {
            uint2 *l2 = klp + slot_own0;
            if (row != 0 && !JJ.lastcls && !last) {
                crc_fast(AP0, o0, l2);
                crc_fast(AP1, o1, l2 + 1);
            } else {
                crc_slow(AP0, o0, JJ, kok, kv, last, l2, slot_own0);
                crc_slow(AP1, o1, JJ, kok, kv, last, l2 + 1, slot_own0 + 1);
            }
        }
    };

    for (unsigned long long ti = 0; ti < my_tiles; ti++, it++) {
        const bool last = (s + 1 == R);
        // ---- checksum the data vectors of this tile (cur is still whole)
        if constexpr (CRC) {
            uint2 *l2 = lp + j_base;  // data slot of input j is j (encode)
            if (s != 0 && !J.lastcls && !last) {
#pragma unroll
                for (int j = 0; j < KH; j++) crc_fast(A[j], cur[j], l2 + j);
            } else {
#pragma unroll
                for (int j = 0; j < KH; j++) crc_slow(A[j], cur[j], J, ok, v, last, l2 + j, static_cast<unsigned int>(j_base + j));
            }
        }
        // ---- where the next tile is: one row down, or row 0 of the next job
        Job NJ = J;
        bool nok = J.col;
        unsigned int nv = (s + 1) * dv + J.o - lead;  // one row down (lead < dv: never negative there)
        if (last) {
            NJ = setup_job(job + n_clusters < fp.n_jobs ? job + n_clusters : job);
            nok = NJ.col && NJ.o >= lead;
            nv = NJ.o - lead;
        }
        if (!nok) nv = 0u;
        const bool more = ti + 1 < my_tiles;
        const uint8_t *nsrc = NJ.stripe;

        if constexpr (!CRC) {
            // without the checksum work the kernel waits on memory in short runs at full clock (ncu: long_scoreboard
            // 2.1 per issue): pulling the tile after next into L2 gave +6 % there (0.80 -> 0.85) but -12 % in long
            // runs under the power cap (0.77 -> 0.67, 1 024 stripes), so it is off unless the caller's plan asks
            if (fp.prefetch && s + 2 < R && J.col) {
                const unsigned int pv = nv + dv;
#pragma unroll
                for (int j = 0; j < KH; j++)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(J.stripe + static_cast<unsigned long long>(pv + my_in_off[j]) * 16ull));
            }
        }
        // ---- lookups; each shard's registers are refilled with the next tile as soon as they are dead
        uint32_t acc[16];
        rs_lookup_merged<KH>(cur, acc, lane4, [&](int j) {
            if (more) cur[j] = ld_vec(nsrc + static_cast<unsigned long long>(nv + my_in_off[j]) * 16ull);
        });
        // ---- planes 0,1 go to the partner, planes 2,3 stay. No full 4x4 byte transposes: both halves travel as
        // byte pairs (word = [pos a: plane A, plane B | pos a+1: plane A, plane B]), the receiver XORs like with like
        // and only then separates the two rows: 16 + 8 PRMT per tile instead of 32.
        uint32_t snd[8], kp[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            snd[i] = __byte_perm(acc[2 * i], acc[2 * i + 1], 0x5410);
            kp[i] = __byte_perm(acc[2 * i], acc[2 * i + 1], 0x7632);
        }
        const unsigned int st_ = it & 1u, use_ = it >> 1;
        if (lane == 0) mbar_expect_tx(my_full + st_ * (NW * 8), 2u * 32u * 16u);
        mbar_wait(my_empty + st_ * (NW * 8), (use_ & 1u) ^ 1u);
        const uint32_t dst = xbuf_peer + (st_ * (2 * NT) + threadIdx.x) * 16u;
        const uint32_t sig = peer_full + st_ * (NW * 8);
        st_async_peer(dst, make_uint4(snd[0], snd[1], snd[2], snd[3]), sig);
        st_async_peer(dst + NT * 16, make_uint4(snd[4], snd[5], snd[6], snd[7]), sig);
        const uint4 own_a = make_uint4(kp[0], kp[1], kp[2], kp[3]), own_b = make_uint4(kp[4], kp[5], kp[6], kp[7]);

        // ---- the previous tile of this job (sent a whole tile of work ago), then bookkeeping
        if (have_prev) finish(it - 1, J, s - 1, false, keep_ok, keep_v, lp - lp_stride);
        keep0 = own_a;
        keep1 = own_b;
        if (last) {
            finish(it, J, s, true, ok, v, lp);  // the parity pipeline does not cross a job boundary
            have_prev = false;
            J = NJ;
            job += n_clusters;
            s = 0;
            lp = J.lp;
        } else {
            keep_ok = ok;
            keep_v = v;
            have_prev = true;
            s++;
            lp += lp_stride;
        }
        ok = nok;
        v = nv;
    }
    cluster_sync_all();  // neither CTA may retire while its partner can still write into it or signal it
}

// ---------------------------------------------------------------------------------------------
// The single-SM form for n_in = K <= 7, n_out <= 4 (RS(6,3), RS(4,2), ...): rs_lut8_kernel's tables (one
// lookup per input byte, all output rows in one packed word) with the same walk, register reload and
// sparse-fold accumulators as above; no cluster, no exchange: a CTA loads all K data vectors of a column,
// stores its n_out parity vectors and checksums all K + n_out of them in the same tile.
template <int K, int NT>
__global__ void __launch_bounds__(NT, 1) rs_lut8_crc_kernel(const __grid_constant__ FusedParams fp) {
    extern __shared__ __align__(16) uint32_t rs_tab[];  // [K][256][32]
    const RsParams &p = fp.rs;
    const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = NT / 32;
    constexpr bool FAST = (K <= 5);
    const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(rs_tab));
    if (FAST && s0 > LUT_FRAG1) __trap();
    for (int e0 = warp * 32; e0 < K * 256; e0 += NW * 32) {
        const int e = e0 + lane;
        const uint32_t v = packed_products(p, e >> 8, static_cast<uint32_t>(e & 255));
#pragma unroll 8
        for (int q = 0; q < 32; q++) {
            const int eq = e0 + q;
            const uint32_t addr = FAST ? lut_fast_addr(eq >> 8, static_cast<uint32_t>(eq & 255), lane) : s0 + (eq * 32 + lane) * 4;
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(__shfl_sync(0xffffffffu, v, q)) : "memory");
        }
    }
    __syncthreads();
    const uint32_t tl = s0 + lane * 4;
    const uint32_t lane4 = lane * 4u;

    const unsigned long long n_ctas = gridDim.x, cta = blockIdx.x;
    const unsigned int dv = fp.g.dv, R = fp.g.rseg, ng = fp.g.ng, lead = fp.g.lead;
    const unsigned int last_vec = p.nvec - 1;
    const int n_out = p.n_out;
    const unsigned long long my_jobs = cta < fp.n_jobs ? (fp.n_jobs - cta + n_ctas - 1) / n_ctas : 0ull;
    const unsigned long long my_tiles = my_jobs * R;
    const unsigned long long lp_stride = static_cast<unsigned long long>(ng) * fp.n_shards;

    struct Job {
        const uint8_t *stripe;
        uint2 *lp;
        unsigned int o, stripe_i;
        bool col, lastcls;
    };
    auto setup_job = [&](unsigned long long job) -> Job {
        Job J;
        const unsigned long long st = job / fp.n_class;
        const unsigned int cls = static_cast<unsigned int>(job - st * fp.n_class);
        J.stripe_i = static_cast<unsigned int>(st);
        J.o = cls * NT + threadIdx.x;
        J.col = J.o < dv;
        J.lastcls = (cls + 1 == fp.n_class);
        J.stripe = p.base + st * p.stripe_stride;
        J.lp = fp.qlog + ((st * R) * ng + (J.o >> 5)) * fp.n_shards;
        return J;
    };
    uint4 A[K], AP[RS_MAX_OUT];
#pragma unroll
    for (int j = 0; j < K; j++) A[j] = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int r = 0; r < RS_MAX_OUT; r++) AP[r] = make_uint4(0, 0, 0, 0);
    auto crc_fast = [&](uint4 &acc, const uint4 &d, uint2 *lp) {
        if (lane == 0) *lp = make_uint2(acc.x, acc.y);
        sfold_mul(acc, fp.g.sh, lane);
        xor4(acc, d);
    };
    auto crc_slow = [&](uint4 &acc, uint4 d, const Job &J, bool ok, unsigned int v, bool last, uint2 *lp, unsigned int slot) {
        and4(d, ok ? 0xFFFFFFFFu : 0u);
        if (v == last_vec) { d.x &= fp.tail_mask[0]; d.y &= fp.tail_mask[1]; d.z &= fp.tail_mask[2]; d.w &= fp.tail_mask[3]; }
        if (lane == 0 && J.col) *lp = make_uint2(acc.x, acc.y);
        sfold_mul(acc, fp.g.sh, lane);
        xor4(acc, d);
        if (last) {
            if (J.col) fp.dump[(static_cast<unsigned long long>(J.stripe_i) * fp.n_shards + slot) * dv + J.o] = acc;
            acc = make_uint4(0, 0, 0, 0);
        }
    };

    uint4 cur[K];
    Job J = setup_job(cta < fp.n_jobs ? cta : 0ull);
    unsigned long long job = cta;
    unsigned int s = 0;
    bool ok = J.col && J.o >= lead;
    unsigned int v = ok ? J.o - lead : 0u;
    uint2 *lp = J.lp;
    if (my_tiles) {
#pragma unroll
        for (int j = 0; j < K; j++) cur[j] = ld_vec(J.stripe + static_cast<unsigned long long>(v + p.in_off16[j]) * 16ull);
    }
    for (unsigned long long ti = 0; ti < my_tiles; ti++) {
        const bool last = (s + 1 == R);
        const bool slow = (s == 0) || J.lastcls || last;
        if (!slow) {
#pragma unroll
            for (int j = 0; j < K; j++) crc_fast(A[j], cur[j], lp + j);  // data slot of input j is j (encode)
        } else {
#pragma unroll
            for (int j = 0; j < K; j++) crc_slow(A[j], cur[j], J, ok, v, last, lp + j, static_cast<unsigned int>(j));
        }
        Job NJ = J;
        bool nok = J.col;
        unsigned int nv = (s + 1) * dv + J.o - lead;
        if (last) {
            NJ = setup_job(job + n_ctas < fp.n_jobs ? job + n_ctas : job);
            nok = NJ.col && NJ.o >= lead;
            nv = NJ.o - lead;
        }
        if (!nok) nv = 0u;
        const bool more = ti + 1 < my_tiles;
        const uint8_t *nsrc = NJ.stripe;

        uint32_t acc[16];
        rs_lookup_merged3<K, FAST, 0>(cur, acc, tl, lane4, [&](int j) {
            if (more) cur[j] = ld_vec(nsrc + static_cast<unsigned long long>(nv + p.in_off16[j]) * 16ull);
        });
        uint32_t o[4][4];
#pragma unroll
        for (int q = 0; q < 4; q++) transpose4x4(&acc[q * 4], o[q]);
        uint8_t *ws = const_cast<uint8_t *>(J.stripe);
#pragma unroll
        for (int r = 0; r < RS_MAX_OUT; r++) {
            if (r < n_out) {
                const uint4 pv = make_uint4(o[0][r], o[1][r], o[2][r], o[3][r]);
                if (ok) *reinterpret_cast<uint4 *>(ws + static_cast<unsigned long long>(v + p.out_off16[r]) * 16ull) = pv;
                uint2 *l2 = lp + p.out_idx[r];
                if (!slow) crc_fast(AP[r], pv, l2);
                else crc_slow(AP[r], pv, J, ok, v, last, l2, static_cast<unsigned int>(p.out_idx[r]));
            }
        }
        if (last) {
            J = NJ;
            job += n_ctas;
            s = 0;
            lp = J.lp;
        } else {
            s++;
            lp += lp_stride;
        }
        ok = nok;
        v = nv;
    }
}

}  // namespace gpu_ec
