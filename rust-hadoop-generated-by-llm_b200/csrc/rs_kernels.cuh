// rs_kernels.cuh — Reed-Solomon GF(2^8) "apply matrix" kernels for sm_100a.
//
// One kernel family serves both encode (reference: reed-solomon-erasure `encode`, call site
// dfs/common/src/erasure.rs:21) and reconstruct (`reconstruct`, erasure.rs:38 and
// dfs/chunkserver/src/chunkserver.rs:620-623): out_r = XOR_j coef[r][j] * in_j over GF(2^8),
// with the host-composed decode rows for reconstruct, so both are ONE pass: read n_in shards,
// write n_out shards.
//
// The path is HBM-bound byte work (no tensor cores). GF multiply-by-constant is a table lookup;
// B200 has no GF/CLMUL instruction, so the binding resource is the shared-memory crossbar
// (32 lookups/clk/SM). Design choices that follow from that:
//   * rs_lut8: ONE lookup per data byte regardless of n_out<=4: a 256-entry table per input
//     shard whose u32 entry packs the products for all 4 output rows; the table is replicated
//     per lane (word index = entry*32 + lane) so every LDS is bank-conflict-free. 32 KiB per
//     input shard -> n_in <= 7 in the 227 KiB of one SM.
//   * rs_lut8x2: the same for 8 <= n_in <= 12 (RS(10,4)) on a 2-CTA thread-block cluster: the SM
//     pair splits the tables and trades partial parity over distributed shared memory.
//   * rs_lut4: split-nibble variant (two 16-entry packed tables per input shard, 4 KiB per
//     shard, lane-replicated, 2 lookups per byte) for very wide codes, n_in up to 56 per pass,
//     with an accumulate flag so k+m <= 256 is covered in several passes.
//   * where the tables fit below 192 KiB in 64 KiB-aligned blocks (n_in <= 5, cluster KH <= 4) a
//     lookup is PRMT + LDS (lut_common.cuh); otherwise PRMT/SHF + IMAD + LDS.
//   * 16-byte coalesced loads straight into registers (no reuse -> no smem staging needed),
//     next tile prefetched before the current one is computed, XOR-accumulate in registers,
//     4x4 byte transpose with PRMT, 16-byte coalesced stores: each byte crosses HBM once.
//   * persistent grid: one CTA per SM (the table owns the SM's shared memory), tiles strided.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "lut_common.cuh"

namespace gpu_ec {

constexpr int RS_THREADS = 512;
constexpr int RS_MAX_OUT = 4;        // output rows per pass (packed into one u32 entry)
constexpr int RS_LUT8_MAX_IN = 7;    // 7 * 32 KiB = 224 KiB
constexpr int RS_LUT4_MAX_IN = 56;   // 56 * 4 KiB = 224 KiB

struct RsParams {
    uint8_t *base;
    unsigned long long shard_stride;
    unsigned long long stripe_stride;
    unsigned long long n_tiles;       // n_stripes * tiles_per_stripe
    unsigned int nvec;                // 16-byte vectors per shard = ceil(shard_len / 16)
    unsigned int tiles_per_stripe;    // ceil(nvec / RS_THREADS)
    int n_in;
    int n_out;
    int accumulate;                   // XOR into the existing output instead of overwriting
    int in_idx[RS_LUT4_MAX_IN];
    int out_idx[RS_MAX_OUT];
    uint8_t coef[RS_MAX_OUT][RS_LUT4_MAX_IN];
    // slot offsets within a stripe in 16-byte units (host: idx * shard_stride / 16; a stripe spans < 64 GiB),
    // so a vector's address is stripe_ptr + (v + off16) * 16: one 32-bit add with a constant-bank operand
    // and one IMAD.WIDE per slot, nothing to keep in registers and nothing worth rematerialising
    unsigned int in_off16[RS_LUT4_MAX_IN];
    unsigned int out_off16[RS_MAX_OUT];
};

// packed products of byte value `x` with the coefficients of input column j for all output rows:
// the four coefficients sit in the four bytes of one word and are doubled together (packed
// multiply-by-x in GF(2^8), polynomial 0x11D), so one Russian-peasant loop yields all four products
__device__ __forceinline__ uint32_t packed_products(const RsParams &p, int j, uint32_t x) {
    uint32_t c = 0;
#pragma unroll
    for (int r = 0; r < RS_MAX_OUT; r++)
        if (r < p.n_out) c |= static_cast<uint32_t>(p.coef[r][j]) << (8 * r);
    uint32_t v = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        v ^= ((x >> i) & 1u) ? c : 0u;
        uint32_t hi;  // 0xFF in every byte whose top bit is set (prmt with sign-replicate selectors)
        asm("prmt.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(c), "r"(0u), "r"(0xBA98u));
        c = ((c & 0x7f7f7f7fu) << 1) ^ (hi & 0x1d1d1d1du);
    }
    return v;
}

// acc[4] hold, for 4 consecutive data byte positions, {row0,row1,row2,row3} product bytes.
// Returns word `r` of the transposed 4x4: the 4 consecutive bytes of output row r.
__device__ __forceinline__ void transpose4x4(const uint32_t a[4], uint32_t o[4]) {
    uint32_t t0 = __byte_perm(a[0], a[1], 0x5140);  // a0.b0 a1.b0 a0.b1 a1.b1
    uint32_t t1 = __byte_perm(a[0], a[1], 0x7362);  // a0.b2 a1.b2 a0.b3 a1.b3
    uint32_t t2 = __byte_perm(a[2], a[3], 0x5140);
    uint32_t t3 = __byte_perm(a[2], a[3], 0x7362);
    o[0] = __byte_perm(t0, t2, 0x5410);
    o[1] = __byte_perm(t0, t2, 0x7632);
    o[2] = __byte_perm(t1, t3, 0x5410);
    o[3] = __byte_perm(t1, t3, 0x7632);
}

// 16-byte read-only load. The RS kernels keep their shared-memory request at <= 192.5 KiB: their
// LDG path needs the remaining L1 (with the whole 227 KiB window requested RS(6,3) fell from 0.98 to
// 0.77 of HBM peak, with or without L1::no_allocate; measured).
__device__ __forceinline__ uint4 ld_vec(const uint8_t *p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }

// acc[q*4+b] ^= table_j[byte b of word q of d[j]] for all j < K: the lookup phase shared by the
// byte-LUT kernels. FAST selects the single-PRMT layout of lut_common.cuh, else tables sit at
// tl + j*32 KiB with 128-byte rows.
template <int K, bool FAST, int J = 0>
__device__ __forceinline__ void rs_accumulate(const uint4 (&d)[K], uint32_t (&acc)[16], uint32_t tl, uint32_t lane4) {
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
        rs_accumulate<K, FAST, J + 1>(d, acc, tl, lane4);
    }
}

// ---------------------------------------------------------------------------------------------
// rs_lut8: full-byte packed LUT, lane-replicated, n_in = K <= 7.
// Algorithmic bytes per tile: (K + n_out) * 16 * RS_THREADS; one LDS per input byte.
template <int K>
__global__ void __launch_bounds__(RS_THREADS, 1) rs_lut8_kernel(const __grid_constant__ RsParams p) {
    extern __shared__ __align__(16) uint32_t rs_tab[];  // [K][256][32]
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int NW = RS_THREADS / 32;

    // K <= 5: fast layout (single-PRMT addressing for tables 0..3, table 4 in [0x8000,0x10000); the
    // window ends at 0x30000 = 191 KiB). K = 6, 7 keep the classic layout [K][256][32] from the window
    // start: a sixth table would push the request past 192 KiB and starve the L1 the loads need.
    constexpr bool FAST = (K <= 5);
    const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(rs_tab));
    if (FAST && s0 > LUT_FRAG1) __trap();  // not a cluster kernel: the window starts at 0
    // build the table: each lane computes one entry, the warp broadcasts it into the 32 lane copies
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
    const uint32_t lane4 = lane * 4;
    const unsigned long long step = gridDim.x;
    unsigned long long t = blockIdx.x;
    if (t >= p.n_tiles) return;

    // Loop-invariant addressing. The tile position is kept as (sbase = stripe * stripe_stride, ntt =
    // tile within the stripe) and advanced incrementally: no division, no 64-bit index math per tile.
    // Loads are unconditional — a thread past the end of the shard re-reads the last vector, its
    // result is never used — so no registers are zeroed and no load is predicated.
    const unsigned int tps = p.tiles_per_stripe;
    const unsigned long long step_stripe = step / tps;
    const unsigned int step_tt = static_cast<unsigned int>(step - step_stripe * tps);
    const unsigned long long d_sbase = step_stripe * p.stripe_stride;
    unsigned long long sbase = (t / tps) * p.stripe_stride;
    unsigned int ntt = static_cast<unsigned int>(t % tps);
    const unsigned int last_vec = p.nvec - 1;

    // a tile position = (stripe pointer, this thread's vector index in the shard); then advance
    struct Pos { const uint8_t *stripe; unsigned int v; bool ok; };
    auto locate_next = [&]() -> Pos {
        Pos q;
        const unsigned int v = ntt * RS_THREADS + threadIdx.x;
        q.ok = v <= last_vec;
        q.v = q.ok ? v : last_vec;
        q.stripe = p.base + sbase;
        sbase += d_sbase;
        ntt += step_tt;
        if (ntt >= tps) { ntt -= tps; sbase += p.stripe_stride; }
        return q;
    };
    auto load = [&](uint4 (&d)[K], const Pos &q) {
#pragma unroll
        for (int j = 0; j < K; j++) d[j] = ld_vec(q.stripe + static_cast<unsigned long long>(q.v + p.in_off16[j]) * 16ull);
    };
    auto process = [&](const uint4 (&d)[K], const Pos &q) {
        if (!q.ok) return;
        uint32_t acc[16];
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] = 0;
        rs_accumulate<K, FAST>(d, acc, tl, lane4);
        uint32_t o[4][4];  // o[q][r]
#pragma unroll
        for (int qq = 0; qq < 4; qq++) transpose4x4(&acc[qq * 4], o[qq]);
        uint8_t *wstripe = const_cast<uint8_t *>(q.stripe);
#pragma unroll
        for (int r = 0; r < RS_MAX_OUT; r++) {
            if (r < p.n_out) {
                uint4 *dst = reinterpret_cast<uint4 *>(wstripe + static_cast<unsigned long long>(q.v + p.out_off16[r]) * 16ull);
                uint4 v = make_uint4(o[0][r], o[1][r], o[2][r], o[3][r]);
                if (p.accumulate) { const uint4 old = *dst; v.x ^= old.x; v.y ^= old.y; v.z ^= old.z; v.w ^= old.w; }
                *dst = v;
            }
        }
    };

    // (a ping-pong variant without the cur = nxt copies was measured: 0.88 vs 1.02 of HBM peak — the
    // doubled loop body costs more in instruction fetch than the copies do)
    uint4 cur[K];
    Pos pos = locate_next();
    load(cur, pos);
    for (; t < p.n_tiles; t += step) {
        uint4 nxt[K];
        Pos npos = pos;
        if (t + step < p.n_tiles) { npos = locate_next(); load(nxt, npos); }
        process(cur, pos);
#pragma unroll
        for (int j = 0; j < K; j++) cur[j] = nxt[j];
        pos = npos;
    }
}

// ---------------------------------------------------------------------------------------------
// rs_lut8x2: the byte-LUT kernel for 8 <= n_in <= 12 on a thread-block CLUSTER of two CTAs.
// One SM cannot hold more than 7 lane-replicated 32 KiB tables, so the two SMs of a cluster split
// the input shards: CTA r owns the tables of inputs [r*KH, r*KH + ni_r), both CTAs walk the same
// tile sequence, each computes the partial products of ALL output rows from its own inputs (still
// one lookup per input byte), and they trade halves through distributed shared memory: CTA 0
// finishes rows [0, h0), CTA 1 rows [h0, n_out); the rows a CTA does not own are sent into its
// partner's exchange buffer, the rows it owns wait in registers.
// The hand-over is warp-to-warp and fence-free: warp w of the producer writes only the slots warp
// w of the consumer reads, with st.async (remote store that completes a transaction count on the
// consumer's mbarrier full[stage][w]); the consumer's warp waits on that mbarrier one tile later
// (so the DSMEM latency hides behind a tile of lookups) and hands the slot back with a remote
// arrive on the producer's empty[stage][w]. No cluster-wide barrier and no memory fence in the loop
// (a barrier.cluster.arrive.release per tile was measured: its MEMBAR drains the prefetched loads
// and halves throughput). HBM traffic stays algorithmic: every input byte read once, every output
// byte written once; DSMEM carries ~0.4 B per input byte.
constexpr int RS_X2_MAX_KH = 6;  // 6 * 32 KiB tables + 32 KiB exchange + mbarriers <= 227 KiB
#ifndef GPU_EC_X2_THREADS
#define GPU_EC_X2_THREADS 512
#endif
constexpr int RS_X2_THREADS = GPU_EC_X2_THREADS;
constexpr int RS_X2_NW = RS_X2_THREADS / 32;
constexpr int RS_X2_XBUF_BYTES = 2 * 2 * RS_X2_THREADS * 16;   // [2 stages][2 rows][RS_X2_THREADS] uint4
constexpr int RS_X2_MBAR_BYTES = 2 * 2 * RS_X2_NW * 8;         // full[2][NW], empty[2][NW]

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_peer(uint32_t local_smem_addr, uint32_t peer) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(peer));
    return r;
}
// remote 16-byte store that signals 16 bytes of transaction completion on the remote mbarrier
__device__ __forceinline__ void st_async_peer(uint32_t addr, const uint4 &v, uint32_t mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(addr),
                 "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(mbar)
                 : "memory");
}
__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t a, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_peer(uint32_t remote) { asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(a),
        "r"(parity)
        : "memory");
}

template <int KH>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(RS_X2_THREADS, 1) rs_lut8x2_kernel(const __grid_constant__ RsParams p) {
    extern __shared__ __align__(16) uint32_t rs_tab[];  // tables | exchange buffer | mbarriers
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int NW = RS_X2_NW;
    const uint32_t rank = cluster_ctarank();
    const int j_base = static_cast<int>(rank) * KH;
    const int ni = (p.n_in - j_base < KH) ? (p.n_in - j_base) : KH;  // rank 1 may own one input fewer
    const int h0 = (p.n_out + 1) >> 1;                               // rows [0,h0) finish on rank 0
    const int own0 = rank ? h0 : 0, n_own = rank ? (p.n_out - h0) : h0;
    const int oth0 = rank ? 0 : h0, n_oth = rank ? h0 : (p.n_out - h0);

    // KH <= 4: fast layout (tables 0..3 single-PRMT addressed in [0x10000,0x30000), exchange buffer +
    // mbarriers below them); KH = 5, 6 keep the classic contiguous layout (same L1 reason as rs_lut8)
    constexpr bool FAST = (KH <= 4);
    const uint32_t s0 = static_cast<uint32_t>(__cvta_generic_to_shared(rs_tab));
    const uint32_t wb = s0 & 0xFF000000u;  // this CTA's window in the cluster shared address space (rank << 24)
    if (FAST && (s0 - wb) + 128u + RS_X2_XBUF_BYTES + RS_X2_MBAR_BYTES > LUT_BLK_A) __trap();
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
    const uint32_t mbar_s = xbuf_s + RS_X2_XBUF_BYTES;  // full[stage][w] at +(stage*NW+w)*8, empty at +(2*NW + stage*NW+w)*8
    if (threadIdx.x < 4 * NW) mbar_init(mbar_s + threadIdx.x * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    cluster_sync_all();  // both CTAs resident, tables built, mbarriers initialised before any remote access
    const uint32_t xbuf_peer = map_to_peer(xbuf_s, rank ^ 1u);
    const uint32_t mbar_peer = map_to_peer(mbar_s, rank ^ 1u);
    const uint32_t my_full = mbar_s + warp * 8, my_empty = mbar_s + (2 * NW + warp) * 8;          // + stage*NW*8
    const uint32_t peer_full = mbar_peer + warp * 8, peer_empty = mbar_peer + (2 * NW + warp) * 8;

    const uint32_t tl = s0 + lane * 4;
    const uint32_t lane4 = wb | static_cast<uint32_t>(lane * 4);
    const unsigned long long n_clusters = gridDim.x >> 1;
    unsigned long long t = blockIdx.x >> 1;

    // Loop-invariant addressing as in rs_lut8: slot offsets are host-precomputed 32-bit vector indices,
    // the tile position is (stripe pointer, vector index) advanced incrementally, loads are
    // unconditional (a thread past the end re-reads the last vector; its result is never stored).
    const unsigned int tps = p.tiles_per_stripe;
    const unsigned long long step_stripe = n_clusters / tps;
    const unsigned int step_tt = static_cast<unsigned int>(n_clusters - step_stripe * tps);
    const unsigned long long d_sbase = step_stripe * p.stripe_stride;
    unsigned long long sbase = (t / tps) * p.stripe_stride;
    unsigned int ntt = static_cast<unsigned int>(t % tps);
    const unsigned int last_vec = p.nvec - 1;
    const unsigned int own_off0 = p.out_off16[n_own > 0 ? own0 : 0], own_off1 = p.out_off16[n_own > 1 ? own0 + 1 : 0];
    unsigned int my_in_off[KH];
#pragma unroll
    for (int j = 0; j < KH; j++) my_in_off[j] = p.in_off16[(j < ni) ? (j_base + j) : j_base];

    struct Pos { const uint8_t *stripe; unsigned int v; bool ok; };
    uint4 cur[KH];
    auto locate_next = [&]() -> Pos {
        Pos q;
        const unsigned int v = ntt * RS_X2_THREADS + threadIdx.x;
        q.ok = v <= last_vec;
        q.v = q.ok ? v : last_vec;
        q.stripe = p.base + sbase;
        sbase += d_sbase;
        ntt += step_tt;
        if (ntt >= tps) { ntt -= tps; sbase += p.stripe_stride; }
        return q;
    };
    auto load = [&](uint4 (&d)[KH], const Pos &q) {
#pragma unroll
        for (int j = 0; j < KH; j++) d[j] = ld_vec(q.stripe + static_cast<unsigned long long>(q.v + my_in_off[j]) * 16ull);
    };
    Pos cur_pos = {p.base, 0, false};
    if (t < p.n_tiles) { cur_pos = locate_next(); load(cur, cur_pos); }

    uint4 keep0 = make_uint4(0, 0, 0, 0), keep1 = keep0;  // the rows this CTA owns, from the previous tile
    Pos keep_pos = {p.base, 0, false};
    bool have_prev = false;
    unsigned int it = 0;
    // finish tile `it-1` (stage (it-1)&1, use (it-1)>>1): wait for the partner's rows, fold, store, free the slot
#define RS_X2_FINISH_PREV()                                                                                          \
    do {                                                                                                             \
        const unsigned int ps_ = (it - 1) & 1u, pu_ = (it - 1) >> 1;                                                 \
        if (n_own > 0) {                                                                                             \
            mbar_wait(my_full + ps_ * (NW * 8), pu_ & 1u);                                                           \
            const uint4 *x = xbuf + ps_ * (2 * RS_X2_THREADS) + threadIdx.x;                                            \
            uint4 o0_ = x[0], o1_ = make_uint4(0, 0, 0, 0);                                                          \
            if (n_own > 1) o1_ = x[RS_X2_THREADS];                                                                      \
            o0_.x ^= keep0.x; o0_.y ^= keep0.y; o0_.z ^= keep0.z; o0_.w ^= keep0.w;                                  \
            o1_.x ^= keep1.x; o1_.y ^= keep1.y; o1_.z ^= keep1.z; o1_.w ^= keep1.w;                                  \
            __syncwarp();                                                                                            \
            if (lane == 0) mbar_arrive_peer(peer_empty + ps_ * (NW * 8)); /* slot may be overwritten */              \
            if (keep_pos.ok) {                                                                                       \
                uint8_t *ws_ = const_cast<uint8_t *>(keep_pos.stripe);                                               \
                *reinterpret_cast<uint4 *>(ws_ + static_cast<unsigned long long>(keep_pos.v + own_off0) * 16ull) = o0_; \
                if (n_own > 1) *reinterpret_cast<uint4 *>(ws_ + static_cast<unsigned long long>(keep_pos.v + own_off1) * 16ull) = o1_; \
            }                                                                                                        \
        }                                                                                                            \
    } while (0)

    for (; t < p.n_tiles; t += n_clusters, it++) {
        uint4 nxt[KH];
        Pos nxt_pos = cur_pos;
        if (t + n_clusters < p.n_tiles) { nxt_pos = locate_next(); load(nxt, nxt_pos); }

        uint32_t acc[16];
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] = 0;
        rs_accumulate<KH, FAST>(cur, acc, tl, lane4);
        uint32_t o[4][4];  // o[q][row]
#pragma unroll
        for (int q = 0; q < 4; q++) transpose4x4(&acc[q * 4], o[q]);

        // send first (the partner gets its rows as early as possible), then finish the previous tile
        const unsigned int st_ = it & 1u, use_ = it >> 1;
        if (n_own > 0 && lane == 0) mbar_expect_tx(my_full + st_ * (NW * 8), static_cast<uint32_t>(n_own) * 32u * 16u);
        // the partner's warp must have drained this stage's previous use (first use passes at once)
        if (n_oth > 0) mbar_wait(my_empty + st_ * (NW * 8), (use_ & 1u) ^ 1u);
        // rows the partner finishes go into its exchange buffer, rows we finish wait in registers
        const uint32_t dst = xbuf_peer + (st_ * (2 * RS_X2_THREADS) + threadIdx.x) * 16u;
        const uint32_t sig = peer_full + st_ * (NW * 8);
        uint4 own_a, own_b;
        if (p.n_out == RS_MAX_OUT) {  // the common 4-row case: rows are compile-time per rank
            const uint4 r0 = make_uint4(o[0][0], o[1][0], o[2][0], o[3][0]), r1 = make_uint4(o[0][1], o[1][1], o[2][1], o[3][1]);
            const uint4 r2 = make_uint4(o[0][2], o[1][2], o[2][2], o[3][2]), r3 = make_uint4(o[0][3], o[1][3], o[2][3], o[3][3]);
            const bool hi = rank != 0;
            st_async_peer(dst, hi ? r0 : r2, sig);
            st_async_peer(dst + RS_X2_THREADS * 16, hi ? r1 : r3, sig);
            own_a = hi ? r2 : r0;
            own_b = hi ? r3 : r1;
        } else {
            own_a = own_b = make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int row = 0; row < RS_MAX_OUT; row++) {
                const uint4 v = make_uint4(o[0][row], o[1][row], o[2][row], o[3][row]);
                if (n_oth > 0 && row == oth0) st_async_peer(dst, v, sig);
                if (n_oth > 1 && row == oth0 + 1) st_async_peer(dst + RS_X2_THREADS * 16, v, sig);
                if (n_own > 0 && row == own0) own_a = v;
                if (n_own > 1 && row == own0 + 1) own_b = v;
            }
        }
        if (have_prev) RS_X2_FINISH_PREV();  // the partner sent these a whole tile of work ago
        keep0 = own_a;
        keep1 = own_b;
        keep_pos = cur_pos;
        have_prev = true;

#pragma unroll
        for (int j = 0; j < KH; j++) cur[j] = nxt[j];
        cur_pos = nxt_pos;
    }
    if (have_prev) RS_X2_FINISH_PREV();
#undef RS_X2_FINISH_PREV
    cluster_sync_all();  // neither CTA may retire while its partner can still write into it or signal it
}

// ---------------------------------------------------------------------------------------------
// rs_lut4: split-nibble packed LUTs, lane-replicated, any n_in <= 56 per pass.
// smem word index: ((j*2 + half)*16 + nibble)*32 + lane. Two LDS per input byte.
__global__ void __launch_bounds__(RS_THREADS, 1) rs_lut4_kernel(const __grid_constant__ RsParams p) {
    extern __shared__ __align__(16) uint32_t rs_tab[];  // [n_in][2][16][32]
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int NW = RS_THREADS / 32;
    const int n_entries = p.n_in * 32;
    for (int e0 = warp * 32; e0 < n_entries; e0 += NW * 32) {
        const int e = e0 + lane;  // e = (j*2 + half)*16 + nibble
        const int j = e >> 5;
        const uint32_t x = (e & 16) ? static_cast<uint32_t>((e & 15) << 4) : static_cast<uint32_t>(e & 15);
        const uint32_t v = packed_products(p, j, x);
#pragma unroll 8
        for (int q = 0; q < 32; q++) rs_tab[(e0 + q) * 32 + lane] = __shfl_sync(0xffffffffu, v, q);
    }
    __syncthreads();

    const uint32_t tl = static_cast<uint32_t>(__cvta_generic_to_shared(rs_tab)) + lane * 4;
    for (unsigned long long t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
        const unsigned long long stripe = t / p.tiles_per_stripe;
        const unsigned int tt = static_cast<unsigned int>(t - stripe * p.tiles_per_stripe);
        const unsigned int v = tt * RS_THREADS + threadIdx.x;
        if (v >= p.nvec) continue;
        const unsigned long long off = stripe * p.stripe_stride + static_cast<unsigned long long>(v) * 16ull;
        uint32_t acc[16];
#pragma unroll
        for (int i = 0; i < 16; i++) acc[i] = 0;
        for (int j0 = 0; j0 < p.n_in; j0 += 4) {
            uint4 d[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                d[u] = (j0 + u < p.n_in)
                           ? ld_vec(p.base + off + static_cast<unsigned long long>(p.in_idx[j0 + u]) * p.shard_stride)
                           : make_uint4(0, 0, 0, 0);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (j0 + u < p.n_in) {
                    const uint32_t tj = tl + (j0 + u) * (2 * 16 * 32 * 4);
                    const uint32_t w[4] = {d[u].x, d[u].y, d[u].z, d[u].w};
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const uint32_t lo4 = w[q] & 0x0f0f0f0fu, hi4 = (w[q] >> 4) & 0x0f0f0f0fu;
                        acc[q * 4 + 0] ^= lut_lds<0>(tj, byte_of<0>(lo4)) ^ lut_lds<2048>(tj, byte_of<0>(hi4));
                        acc[q * 4 + 1] ^= lut_lds<0>(tj, byte_of<1>(lo4)) ^ lut_lds<2048>(tj, byte_of<1>(hi4));
                        acc[q * 4 + 2] ^= lut_lds<0>(tj, byte_of<2>(lo4)) ^ lut_lds<2048>(tj, byte_of<2>(hi4));
                        acc[q * 4 + 3] ^= lut_lds<0>(tj, lo4 >> 24) ^ lut_lds<2048>(tj, hi4 >> 24);
                    }
                }
            }
        }
        uint32_t o[4][4];
#pragma unroll
        for (int q = 0; q < 4; q++) transpose4x4(&acc[q * 4], o[q]);
#pragma unroll
        for (int r = 0; r < RS_MAX_OUT; r++) {
            if (r < p.n_out) {
                uint4 *dst = reinterpret_cast<uint4 *>(p.base + off + static_cast<unsigned long long>(p.out_idx[r]) * p.shard_stride);
                uint4 val = make_uint4(o[0][r], o[1][r], o[2][r], o[3][r]);
                if (p.accumulate) { const uint4 old = *dst; val.x ^= old.x; val.y ^= old.y; val.z ^= old.z; val.w ^= old.w; }
                *dst = val;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// synthetic data: byte i = byte (i%8) of splitmix64(seed ^ (i/8))
__device__ __forceinline__ unsigned long long splitmix64_dev(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// out must be 8-byte aligned and byte_offset a multiple of 8 (checked by the host wrapper)
__global__ void fill_synthetic_kernel(uint8_t *out, unsigned long long len, unsigned long long seed,
                                      unsigned long long byte_offset) {
    const unsigned long long nwords = (len + 7) / 8;
    for (unsigned long long w = blockIdx.x * static_cast<unsigned long long>(blockDim.x) + threadIdx.x; w < nwords;
         w += static_cast<unsigned long long>(gridDim.x) * blockDim.x) {
        const unsigned long long v = splitmix64_dev(seed ^ ((byte_offset >> 3) + w));
        if ((w + 1) * 8 <= len) {
            reinterpret_cast<unsigned long long *>(out)[w] = v;
        } else {
            for (unsigned long long i = w * 8; i < len; i++) out[i] = static_cast<uint8_t>(v >> (8 * (i & 7)));
        }
    }
}

}  // namespace gpu_ec
