// gpu_ec.cu — libgpu_ec.so: context, launch logic and the extern "C" entry points declared in
// include/gpu_ec.h. The data path runs only in the CUDA kernels of rs_kernels.cuh and
// crc_kernels.cuh; there is no CPU fallback (no device -> GPU_EC_ERR_NO_DEVICE).
#include "../../include/gpu_ec.h"

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include "crc_kernels.cuh"
#include "host_math.hpp"
#include "rs_kernels.cuh"
#include "sfold_kernels.cuh"
#include "fused_kernels.cuh"

namespace gpu_ec {

static thread_local std::string t_last_error;
static std::atomic<uint64_t> g_launches{0};

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            char buf_[512];                                                                        \
            snprintf(buf_, sizeof buf_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            t_last_error = buf_;                                                                   \
            return (e_ == cudaErrorMemoryAllocation) ? GPU_EC_ERR_OOM : GPU_EC_ERR_CUDA;           \
        }                                                                                          \
    } while (0)
#define RET(call)                   \
    do {                            \
        int rc_ = (call);           \
        if (rc_ != GPU_EC_OK) return rc_; \
    } while (0)

// NVTX range around every compute entry point (the reference wraps its RPCs in tracing spans,
// dfs/chunkserver/src/chunkserver.rs:727-728; here the spans show up in nsys/ncu timelines). Header-only
// NVTX3: without a profiler attached a push/pop is one predictable branch.
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};
#define GPU_EC_TRACE() NvtxRange nvtx_range_(__func__)
// the _dev entry points work on whichever device owns the caller's pointer: they switch to it for their launches
// and put the caller's current device back on return
struct DeviceGuard {
    int prev = -1;
    DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); } }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

static inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline size_t div_up(size_t a, size_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------
constexpr int NSLOT = 4;
struct Slot {
    cudaStream_t stream = nullptr;
    uint8_t *dbuf = nullptr;
    size_t cap = 0;
    uint8_t *hbuf = nullptr;   // pinned host staging (batched small blocks: one H2D / one D2H per batch)
    size_t hcap = 0;
    cudaStream_t stream2 = nullptr;   // second stream + events: D2H of a large single call overlaps its own H2D
    cudaEvent_t ev[8] = {};
    bool busy = false;
};
struct Dev {
    int ordinal = -1;
    bool ready = false;
    int sm_count = 0;
    std::mutex mu;
    std::condition_variable cv;
    Slot slots[NSLOT];
};
struct Ctx {
    std::mutex mu;
    bool inited = false;
    int status = GPU_EC_OK;
    std::vector<std::unique_ptr<Dev>> devs;
    std::vector<int> active;
    std::atomic<uint64_t> rr{0};
};
static Ctx &ctx() { static Ctx c; return c; }

// kernels in the single-PRMT layout (lut_common.cuh) address the shared window absolutely, up to 0x30000
static constexpr int x2_smem_bytes(int kh) { return kh <= 4 ? LUT_FAST_RS_BYTES : kh * 256 * 32 * 4 + RS_X2_XBUF_BYTES + RS_X2_MBAR_BYTES; }
static constexpr int lut8_smem_bytes(int k) { return k <= 5 ? LUT_FAST_RS_BYTES : k * 256 * 32 * 4; }

template <int K>
static int set_lut8_attr() {
    CK(cudaFuncSetAttribute(rs_lut8_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(K)));
    return GPU_EC_OK;
}

static int dev_prepare(Dev &d) {
    std::lock_guard<std::mutex> lk(d.mu);
    if (d.ready) return GPU_EC_OK;
    CK(cudaSetDevice(d.ordinal));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, d.ordinal));
    if (prop.major < 10) {
        t_last_error = "gpu_ec needs an sm_100a (B200) device";
        return GPU_EC_ERR_NO_DEVICE;
    }
    d.sm_count = prop.multiProcessorCount;
    static uint32_t tabs[2][4][256];
    static std::once_flag once;
    std::call_once(once, [] { for (int p = 0; p < 2; p++) crc_tables(kPolyReflected[p], tabs[p]); });
    CK(cudaMemcpyToSymbol(g_crc_tab, tabs, sizeof tabs));
    RET(set_lut8_attr<1>()); RET(set_lut8_attr<2>()); RET(set_lut8_attr<3>()); RET(set_lut8_attr<4>());
    RET(set_lut8_attr<5>()); RET(set_lut8_attr<6>()); RET(set_lut8_attr<7>());
    CK(cudaFuncSetAttribute(rs_lut4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RS_LUT4_MAX_IN * 4096));
    CK(cudaFuncSetAttribute(rs_lut8x2_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, x2_smem_bytes(4)));
    CK(cudaFuncSetAttribute(rs_lut8x2_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, x2_smem_bytes(5)));
    CK(cudaFuncSetAttribute(rs_lut8x2_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, x2_smem_bytes(6)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<1, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(1)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<2, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(2)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<3, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(3)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<4, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(4)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<5, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(5)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<6, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(6)));
    CK(cudaFuncSetAttribute(rs_lut8_crc_kernel<7, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, lut8_smem_bytes(7)));
    CK(cudaFuncSetAttribute(rs_x2_crc_fast_kernel<4, FUSED_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, LUT_FAST_RS_BYTES));
    CK(cudaFuncSetAttribute(rs_x2_crc_fast_kernel<4, FUSED_THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, LUT_FAST_RS_BYTES));
    CK(cudaFuncSetAttribute(rs_x2_crc_fast_kernel<5, FUSED_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, LUT_FAST_RS_BYTES));
    CK(cudaFuncSetAttribute(rs_x2_crc_fast_kernel<5, FUSED_THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, LUT_FAST_RS_BYTES));
    CK(cudaFuncSetAttribute(rs_lut8x2_crc_kernel<4, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, fused_smem_bytes(4, FUSED_THREADS)));
    CK(cudaFuncSetAttribute(rs_lut8x2_crc_kernel<5, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, fused_smem_bytes(5, FUSED_THREADS)));
    CK(cudaFuncSetAttribute(rs_lut8x2_crc_kernel<6, FUSED_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, fused_smem_bytes(6, FUSED_THREADS)));
    CK(cudaFuncSetAttribute(crc_segments_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CRC_SMEM_BYTES));
    CK(cudaFuncSetAttribute(crc_segments_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CRC_SMEM_BYTES));
    for (auto &s : d.slots) CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    // keep stream-ordered scratch cached instead of returning it to the OS after every call
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, d.ordinal) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    d.ready = true;
    return GPU_EC_OK;
}

static int ensure_init() {
    Ctx &c = ctx();
    std::lock_guard<std::mutex> lk(c.mu);
    if (c.inited) return c.status;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        t_last_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
        cudaGetLastError();
        return GPU_EC_ERR_NO_DEVICE;  // not latched: a later call may find a device
    }
    for (int i = 0; i < n; i++) {
        c.devs.emplace_back(new Dev());
        c.devs.back()->ordinal = i;
    }
    if (c.active.empty()) for (int i = 0; i < n; i++) c.active.push_back(i);
    c.inited = true;
    c.status = GPU_EC_OK;
    return GPU_EC_OK;
}

static int get_dev(int ordinal, Dev **out) {
    RET(ensure_init());
    Ctx &c = ctx();
    if (ordinal < 0 || ordinal >= static_cast<int>(c.devs.size())) return GPU_EC_ERR_INVALID_ARG;
    Dev &d = *c.devs[ordinal];
    RET(dev_prepare(d));
    *out = &d;
    return GPU_EC_OK;
}

// device that owns a device pointer (for the _dev entry points)
static int dev_of_ptr(const void *p, Dev **out) {
    RET(ensure_init());
    cudaPointerAttributes a;
    CK(cudaPointerGetAttributes(&a, p));
    if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged) {
        t_last_error = "pointer is not device memory";
        return GPU_EC_ERR_INVALID_ARG;
    }
    RET(get_dev(a.device, out));
    CK(cudaSetDevice(a.device));
    return GPU_EC_OK;
}

struct Lease {
    Dev *d = nullptr;
    Slot *s = nullptr;
    bool alone = false;  // no other caller held a slot of this device when this one was leased
    ~Lease() { release(); }
    void release() {
        if (d && s) {
            // every success path has already synchronised; on an early error return this drains copies that
            // may still target caller memory before the slot (and its staging buffer) is handed on
            if (cudaSetDevice(d->ordinal) == cudaSuccess) {
                cudaStreamSynchronize(s->stream);
                if (s->stream2) cudaStreamSynchronize(s->stream2);
            }
            std::lock_guard<std::mutex> lk(d->mu);
            s->busy = false;
            d->cv.notify_all();
        }
        d = nullptr; s = nullptr;
    }
};

static int lease_on(Dev &d, Lease &l) {
    {
        std::unique_lock<std::mutex> lk(d.mu);
        Slot *s = nullptr;
        d.cv.wait(lk, [&] {
            for (auto &x : d.slots) if (!x.busy) { s = &x; return true; }
            return false;
        });
        s->busy = true;
        l.d = &d; l.s = s;
        int busy = 0;
        for (auto &x : d.slots) busy += x.busy ? 1 : 0;
        l.alone = (busy == 1);
    }
    CK(cudaSetDevice(d.ordinal));
    return GPU_EC_OK;
}

// second stream and events of a slot, created on first use (device current)
static int slot_overlap_init(Slot &s) {
    if (s.stream2) return GPU_EC_OK;
    CK(cudaStreamCreateWithFlags(&s.stream2, cudaStreamNonBlocking));
    for (auto &e : s.ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    return GPU_EC_OK;
}

// n slots of one device, all or nothing: two batch callers that each grabbed some and waited for the rest
// would deadlock, so the wait is for n free slots at once
static int lease_n_on(Dev &d, int n, Lease *out) {
    {
        std::unique_lock<std::mutex> lk(d.mu);
        d.cv.wait(lk, [&] {
            int free_slots = 0;
            for (auto &x : d.slots) free_slots += x.busy ? 0 : 1;
            return free_slots >= n;
        });
        int got = 0;
        for (auto &x : d.slots) {
            if (got == n) break;
            if (!x.busy) { x.busy = true; out[got].d = &d; out[got].s = &x; got++; }
        }
    }
    CK(cudaSetDevice(d.ordinal));
    return GPU_EC_OK;
}

static int lease_any(Lease &l) {
    RET(ensure_init());
    Ctx &c = ctx();
    int ord;
    {
        std::lock_guard<std::mutex> lk(c.mu);
        ord = c.active[c.rr.fetch_add(1) % c.active.size()];
    }
    Dev *d;
    RET(get_dev(ord, &d));
    return lease_on(*d, l);
}

static int slot_reserve(Slot &s, size_t bytes) {
    if (s.cap >= bytes) return GPU_EC_OK;
    if (s.dbuf) { CK(cudaStreamSynchronize(s.stream)); CK(cudaFree(s.dbuf)); s.dbuf = nullptr; s.cap = 0; }
    size_t want = round_up(bytes + bytes / 8, 1 << 20);
    CK(cudaMalloc(&s.dbuf, want));
    CK(cudaMemsetAsync(s.dbuf, 0, want, s.stream));
    s.cap = want;
    return GPU_EC_OK;
}

static int slot_reserve_host(Slot &s, size_t bytes) {
    if (s.hcap >= bytes) return GPU_EC_OK;
    if (s.hbuf) { CK(cudaStreamSynchronize(s.stream)); CK(cudaFreeHost(s.hbuf)); s.hbuf = nullptr; s.hcap = 0; }
    const size_t want = round_up(bytes + bytes / 4, 1 << 20);
    CK(cudaHostAlloc(reinterpret_cast<void **>(&s.hbuf), want, cudaHostAllocPortable));
    s.hcap = want;
    return GPU_EC_OK;
}

// ---------------------------------------------------------------------------------------------
// launch logic

static int check_stripes(const gpu_ec_stripes *s) {
    if (!s || !s->base) return GPU_EC_ERR_INVALID_ARG;
    if ((reinterpret_cast<uintptr_t>(s->base) & 15) || (s->shard_stride & 15) || (s->stripe_stride & 15)) return GPU_EC_ERR_UNALIGNED;
    if (s->shard_stride < round_up(s->shard_len, 16)) return GPU_EC_ERR_INVALID_ARG;
    return GPU_EC_OK;
}

template <int K>
static void launch_lut8(const RsParams &p, int grid, cudaStream_t st) {
    rs_lut8_kernel<K><<<grid, RS_THREADS, lut8_smem_bytes(K), st>>>(p);
}

// one kernel launch: out slot out_idx[r] (r < no) (^)= XOR_{j<ni} rows[r][j] * in slot in_idx[j]
static int launch_apply_pass(Dev &d, const gpu_ec_stripes &s, const uint8_t *rows, size_t row_stride, const int *in_idx,
                             int ni, const int *out_idx, int no, bool accumulate, bool use_lut8, cudaStream_t st) {
    RsParams p;
    memset(&p, 0, sizeof p);
    p.base = s.base;
    p.shard_stride = s.shard_stride;
    p.stripe_stride = s.stripe_stride;
    p.nvec = static_cast<unsigned int>(div_up(s.shard_len, 16));
    p.tiles_per_stripe = static_cast<unsigned int>(div_up(p.nvec, RS_THREADS));
    p.n_tiles = static_cast<unsigned long long>(s.n_stripes) * p.tiles_per_stripe;
    p.n_in = ni;
    p.n_out = no;
    p.accumulate = accumulate ? 1 : 0;
    for (int r = 0; r < no; r++) { p.out_idx[r] = out_idx[r]; p.out_off16[r] = static_cast<unsigned int>(out_idx[r] * (s.shard_stride / 16)); }
    for (int j = 0; j < ni; j++) { p.in_idx[j] = in_idx[j]; p.in_off16[j] = static_cast<unsigned int>(in_idx[j] * (s.shard_stride / 16)); }
    for (int r = 0; r < no; r++)
        for (int j = 0; j < ni; j++) p.coef[r][j] = rows[static_cast<size_t>(r) * row_stride + j];
    const int grid = static_cast<int>(std::min<unsigned long long>(p.n_tiles, static_cast<unsigned long long>(d.sm_count)));
    if (use_lut8) {
        switch (ni) {
            case 1: launch_lut8<1>(p, grid, st); break;
            case 2: launch_lut8<2>(p, grid, st); break;
            case 3: launch_lut8<3>(p, grid, st); break;
            case 4: launch_lut8<4>(p, grid, st); break;
            case 5: launch_lut8<5>(p, grid, st); break;
            case 6: launch_lut8<6>(p, grid, st); break;
            default: launch_lut8<7>(p, grid, st); break;
        }
    } else {
        rs_lut4_kernel<<<grid, RS_THREADS, ni * 4096, st>>>(p);
    }
    g_launches++;
    CK(cudaGetLastError());
    return GPU_EC_OK;
}

template <int KH>
static cudaError_t launch_x2(const RsParams &p, int n_clusters, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * n_clusters);
    cfg.blockDim = dim3(RS_X2_THREADS);
    cfg.dynamicSmemBytes = x2_smem_bytes(KH);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, rs_lut8x2_kernel<KH>, p);
}

static int apply_x2_fast(Dev &d, const RsParams &rp, size_t n_stripes, cudaStream_t st);

static int launch_apply_x2(Dev &d, const gpu_ec_stripes &s, const uint8_t *rows, size_t row_stride, const int *in_idx,
                           int ni, const int *out_idx, int no, cudaStream_t st) {
    RsParams p;
    memset(&p, 0, sizeof p);
    p.base = s.base;
    p.shard_stride = s.shard_stride;
    p.stripe_stride = s.stripe_stride;
    p.nvec = static_cast<unsigned int>(div_up(s.shard_len, 16));
    p.tiles_per_stripe = static_cast<unsigned int>(div_up(p.nvec, RS_X2_THREADS));
    p.n_tiles = static_cast<unsigned long long>(s.n_stripes) * p.tiles_per_stripe;
    p.n_in = ni;
    p.n_out = no;
    for (int r = 0; r < no; r++) { p.out_idx[r] = out_idx[r]; p.out_off16[r] = static_cast<unsigned int>(out_idx[r] * (s.shard_stride / 16)); }
    for (int j = 0; j < ni; j++) { p.in_idx[j] = in_idx[j]; p.in_off16[j] = static_cast<unsigned int>(in_idx[j] * (s.shard_stride / 16)); }
    for (int r = 0; r < no; r++)
        for (int j = 0; j < ni; j++) p.coef[r][j] = rows[static_cast<size_t>(r) * row_stride + j];
    const int n_clusters = static_cast<int>(std::min<unsigned long long>(p.n_tiles, static_cast<unsigned long long>(d.sm_count / 2)));
    const int kh = (ni + 1) / 2;
    static const bool old_x2 = getenv("GPU_EC_X2_OLD") != nullptr;  // kernel experiments only
    if (ni == 2 * kh && no == RS_MAX_OUT && (kh == 4 || kh == 5) && !old_x2) return apply_x2_fast(d, p, s.n_stripes, st);
    cudaError_t e = kh <= 4 ? launch_x2<4>(p, n_clusters, st) : kh == 5 ? launch_x2<5>(p, n_clusters, st) : launch_x2<6>(p, n_clusters, st);
    g_launches++;
    CK(e);
    return GPU_EC_OK;
}

// slot offsets travel to the kernels as 32-bit indices of 16-byte vectors: every used slot must start and end
// below 64 GiB from the stripe's base (a 7 GiB shard at RS(6,3); the reference caps shards at 100 MiB)
static bool slots_addressable(const gpu_ec_stripes &s, const int *in_idx, int n_in, const int *out_idx, int n_out) {
    unsigned long long hi = 0;
    for (int j = 0; j < n_in; j++) hi = std::max<unsigned long long>(hi, static_cast<unsigned long long>(in_idx[j]));
    for (int r = 0; r < n_out; r++) hi = std::max<unsigned long long>(hi, static_cast<unsigned long long>(out_idx[r]));
    return (hi + 1) * (s.shard_stride / 16) < (1ull << 32);
}

// out slot out_idx[r] = XOR_j rows[r*n_in+j] * in slot in_idx[j].
// n_in <= 7 and n_out <= 4: ONE pass of the byte-LUT kernel over everything.
// Otherwise several passes (output rows in groups of 4; inputs in balanced groups of <= 7 for the
// byte-LUT kernel, or of <= 56 for the split-nibble kernel when the code is very wide), run block by
// block: a block is a range of stripes (or of columns of one stripe) whose 4 output rows fit well
// inside the 126 MB L2, so the partial parity a pass accumulates into is still L2-resident and
// only the inputs stream from HBM — one lookup per input byte and ~algorithmic HBM traffic.
static int launch_apply(Dev &d, const gpu_ec_stripes &s, const uint8_t *rows, const int *in_idx, int n_in,
                        const int *out_idx, int n_out, cudaStream_t st) {
    if (n_in <= 0 || n_out <= 0) return GPU_EC_ERR_INVALID_ARG;
    if (s.n_stripes == 0 || s.shard_len == 0) return GPU_EC_OK;
    for (int j = 0; j < n_in; j++) if (in_idx[j] < 0) return GPU_EC_ERR_INVALID_ARG;
    for (int r = 0; r < n_out; r++) if (out_idx[r] < 0) return GPU_EC_ERR_INVALID_ARG;
    if (!slots_addressable(s, in_idx, n_in, out_idx, n_out)) {
        t_last_error = "stripe too large: a used shard slot lies 64 GiB or more past the stripe base";
        return GPU_EC_ERR_INVALID_ARG;
    }
    if (n_in <= RS_LUT8_MAX_IN && n_out <= RS_MAX_OUT)
        return launch_apply_pass(d, s, rows, n_in, in_idx, n_in, out_idx, n_out, false, true, st);
    static const bool no_cluster = getenv("GPU_EC_NO_CLUSTER") != nullptr;  // kernel experiments only
    if (n_in <= 2 * RS_X2_MAX_KH && !no_cluster) {
        // 8..12 inputs: one pass on 2-CTA clusters (tables split over the SM pair, DSMEM exchange)
        for (int o0 = 0; o0 < n_out; o0 += RS_MAX_OUT)
            RET(launch_apply_x2(d, s, rows + static_cast<size_t>(o0) * n_in, n_in, in_idx, n_in, out_idx + o0,
                                std::min(RS_MAX_OUT, n_out - o0), st));
        return GPU_EC_OK;
    }
    static const bool force_lut4 = getenv("GPU_EC_FORCE_LUT4") != nullptr;  // kernel experiments only
    const bool lut8 = n_in <= 2 * RS_LUT8_MAX_IN && !force_lut4;
    const int max_in = lut8 ? RS_LUT8_MAX_IN : RS_LUT4_MAX_IN;
    const int n_groups = static_cast<int>(div_up(n_in, max_in));
    const int per_group = static_cast<int>(div_up(n_in, n_groups));  // balanced: 10 -> 5 + 5
    // measured on B200: per-block launches cost more than the L2 hits save (tools/exp_kernels.py), so by
    // default a pass covers every stripe; GPU_EC_L2_BLOCK_MB re-enables blocking for experiments
    static const size_t L2_BLOCK = [] { const char *e = getenv("GPU_EC_L2_BLOCK_MB"); return e ? (static_cast<size_t>(atoi(e)) << 20) : ~size_t(0) / 4; }();
    const size_t tile_bytes = static_cast<size_t>(RS_THREADS) * 16;
    for (int o0 = 0; o0 < n_out; o0 += RS_MAX_OUT) {
        const int no = std::min(RS_MAX_OUT, n_out - o0);
        const size_t out_per_stripe = static_cast<size_t>(no) * s.shard_len;
        const bool split_cols = (n_groups > 1) && out_per_stripe > L2_BLOCK;
        const size_t stripes_per_block = (n_groups == 1) ? s.n_stripes
                                         : split_cols     ? 1
                                                          : std::max<size_t>(1, L2_BLOCK / out_per_stripe);
        const size_t cols_per_block = split_cols ? std::max(tile_bytes, L2_BLOCK / no / tile_bytes * tile_bytes) : s.shard_len;
        for (size_t i0 = 0; i0 < s.n_stripes; i0 += stripes_per_block) {
            for (size_t c0 = 0; c0 < s.shard_len; c0 += cols_per_block) {
                gpu_ec_stripes blk = s;
                blk.base = s.base + i0 * s.stripe_stride + c0;
                blk.n_stripes = std::min(stripes_per_block, s.n_stripes - i0);
                blk.shard_len = std::min(cols_per_block, s.shard_len - c0);
                for (int g = 0; g < n_groups; g++) {
                    const int j0 = g * per_group, ni = std::min(per_group, n_in - j0);
                    if (ni <= 0) break;
                    RET(launch_apply_pass(d, blk, rows + static_cast<size_t>(o0) * n_in + j0, n_in, in_idx + j0, ni,
                                          out_idx + o0, no, g > 0, lut8, st));
                }
            }
        }
    }
    return GPU_EC_OK;
}

struct CrcJob {
    const uint8_t *base = nullptr;
    size_t n_buf = 0, buf_stride = 0, len = 0;
    size_t chunk = 0;          // 0 = whole-buffer mode (one CRC per buffer)
    int poly = 0;
    uint32_t *out = nullptr;   // may be null in verify mode
    size_t out_stride = 0;
    int big_endian = 0;
    const uint8_t *expect_be = nullptr;
    unsigned long long *first_bad = nullptr;
    bool no_sfold = false;     // keep whole-buffer mode on the table kernel (the sparse-fold finisher's own dump rows)
};

// words of second-level scratch launch_combine needs for outputs of up to `maxcnt` full elements
static size_t fold_scratch_words(size_t n_outputs, size_t maxcnt) {
    size_t words = 0;
    while (maxcnt > 1024) {
        maxcnt = div_up(maxcnt, 1024);
        words += n_outputs * maxcnt;
    }
    return words;
}

// Fold `segs_per_buf` per-segment CRCs per buffer (seg_bytes each, the buffer's last one maybe
// shorter) into one CRC per `group` segments. scratch2 (fold_scratch_words) is only touched when
// an output has more than 1024 segments; pass nullptr to take it from the stream-ordered allocator.
static int launch_combine(Dev &d, const uint32_t *partial, size_t n_buf, size_t segs_per_buf, size_t seg_bytes,
                          size_t len, size_t group, const CrcJob &j, cudaStream_t st, uint32_t *scratch2 = nullptr,
                          bool partial_be = false) {
    const uint32_t P = kPolyReflected[j.poly];
    FoldParams f;
    memset(&f, 0, sizeof f);
    f.n_buf = n_buf;
    f.outs_per_buf = div_up(segs_per_buf, group);
    const size_t last_cnt = segs_per_buf - (f.outs_per_buf - 1) * group;
    const size_t rag = len % seg_bytes;
    f.last_has_rag = rag ? 1 : 0;
    f.x_rag = x8n_modp(rag, P);
    f.rag_src = partial; f.rag_stride_buf = segs_per_buf;
    f.rag_group = static_cast<unsigned int>(group); f.rag_last_cnt = static_cast<unsigned int>(last_cnt);
    f.poly_bits = P;
    f.out = j.out; f.out_stride = j.out_stride; f.big_endian = j.big_endian;
    f.expect_be = j.expect_be; f.first_bad = j.first_bad;
    size_t cnt_last = last_cnt - (rag ? 1 : 0);
    size_t cnt_reg = f.outs_per_buf > 1 ? group : cnt_last;
    const size_t n_outputs = n_buf * f.outs_per_buf;
    const size_t need2 = fold_scratch_words(n_outputs, std::max(cnt_reg, cnt_last));
    uint32_t *owned = nullptr;
    if (need2 && !scratch2) { CK(cudaMallocAsync(&owned, 4 * need2, st)); scratch2 = owned; }
    f.cur = partial; f.cur_stride_buf = segs_per_buf; f.cur_group = static_cast<unsigned int>(group);
    f.cur_be = partial_be ? 1 : 0; f.rag_be = f.cur_be;  // only the first level reads the caller's array
    uint64_t U = seg_bytes;
    for (;;) {
        const size_t maxcnt = std::max(cnt_reg, cnt_last);
        unsigned int T = 32;
        while (T < 1024 && T < maxcnt) T <<= 1;
        f.T = T;
        f.nb = static_cast<unsigned int>(std::max<size_t>(1, div_up(maxcnt, T)));
        f.final_level = (f.nb == 1);
        f.cnt_reg = static_cast<unsigned int>(cnt_reg);
        f.cnt_last = static_cast<unsigned int>(cnt_last);
        {   // c1[l] = x^(8*U*l), c2[w] = x^(8*U*32*w): one power, then 62 products
            const uint32_t xu = x8n_modp(U, P);
            f.c1[0] = 0x80000000u;
            for (int l = 1; l < 32; l++) f.c1[l] = multmodp(f.c1[l - 1], xu, P);
            const uint32_t xu32 = multmodp(f.c1[31], xu, P);
            f.c2[0] = 0x80000000u;
            for (int l = 1; l < 32; l++) f.c2[l] = multmodp(f.c2[l - 1], xu32, P);
        }
        if (!f.final_level) { f.nxt = scratch2; f.nxt_stride_buf = f.outs_per_buf * f.nb; }
        const unsigned int block = std::max(T, 256u);
        const unsigned long long items = static_cast<unsigned long long>(n_outputs) * f.nb;
        const unsigned long long ctas = div_up(items, block / T);
        const int grid = static_cast<int>(std::min<unsigned long long>(ctas, static_cast<unsigned long long>(d.sm_count) * (2048 / block) * 4));
        crc_fold_kernel<<<grid, block, 0, st>>>(f);
        g_launches++;
        if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) {
            if (owned) cudaFreeAsync(owned, st);
            CK(e);
        }
        if (f.final_level) break;
        f.cur = scratch2; f.cur_stride_buf = f.nxt_stride_buf; f.cur_group = f.nb; f.cur_be = 0;
        scratch2 += n_outputs * f.nb;
        cnt_reg = div_up(cnt_reg, T); cnt_last = div_up(cnt_last, T);
        U *= T;
    }
    if (owned) CK(cudaFreeAsync(owned, st));
    return GPU_EC_OK;
}

// whole-buffer mode: segment length that fills an integer number of waves of (SMs x warps) tiles
static size_t pick_whole_segment(const Dev &d, size_t n_buf, size_t len) {
    const size_t wave_tiles = static_cast<size_t>(d.sm_count) * CRC_WARPS;
    const size_t G_CAP = 16384;
    for (size_t w = 1; w <= 64; w++) {
        const size_t tpb = w * wave_tiles / n_buf;  // 32-segment tiles each buffer may use
        if (tpb == 0) continue;
        const size_t G = round_up(div_up(len, 32 * tpb), 128);
        if (G <= G_CAP) return std::max<size_t>(G, 512);
    }
    return G_CAP;
}

// ---------------------------------------------------------------------------------------------
// sparse-fold whole-buffer CRC (sfold_kernels.cuh)

struct SfoldPlan {
    SfoldGeom g;
    unsigned int tw = 0, n_class = 0;     // columns per CTA, CTAs per row
    unsigned long long nvec = 0;
    size_t n_streams = 0;                 // n_buf * n_seg
    size_t dump_bytes = 0, qlog_bytes = 0;  // scratch, whole job
};

// GPU_EC_SFOLD=0 switches the sparse-fold kernels off everywhere (tests pin each path with it). The fused
// encode + CRC pass uses them by default. Plain whole-buffer CRCs stay on the table kernel unless
// GPU_EC_SFOLD=2: stand-alone, the sparse-fold kernel measured 0.77 of HBM peak against the table kernel's
// 0.84 (four launches instead of two, and nothing else in the SM for its shifts to hide behind).
static int sfold_mode() {
    const char *e = getenv("GPU_EC_SFOLD");
    return e ? atoi(e) : 1;
}
static bool sfold_enabled() { return sfold_mode() != 0; }

// geometry for n_buf buffers of `len` bytes walked by CTAs of `tw_fixed` threads (0 = pick 256..512).
// want_ctas = CTAs the launch can keep busy at once. false = not worth it (too short for any stride).
static bool sfold_plan(size_t n_buf, size_t len, unsigned int tw_fixed, size_t want_ctas, bool allow_segments, SfoldPlan &pl) {
    const unsigned long long nvec = div_up(len, 16);
    double best = -1.0;
    for (int e = 0; e < kSfoldIeeeCount; e++) {
        const SfoldStride &sd = kSfoldIeee[e];
        const unsigned long long rows = div_up(nvec, sd.dv);
        if (rows < 6) break;
        unsigned int tw = tw_fixed;
        if (!tw) {
            double bf = 0;
            for (unsigned int c = 256; c <= 512; c += 32) {
                const double f = static_cast<double>(sd.dv) / (static_cast<double>(c) * div_up(sd.dv, c));
                if (f > bf + 1e-9) { bf = f; tw = c; }
            }
        }
        const unsigned int n_class = static_cast<unsigned int>(div_up(sd.dv, tw));
        const double fill = static_cast<double>(sd.dv) / (static_cast<double>(tw) * n_class);
        unsigned long long n_seg = 1;
        if (allow_segments) {
            const unsigned long long max_seg = std::max<unsigned long long>(1, rows / 24);
            n_seg = std::min<unsigned long long>(max_seg, std::max<unsigned long long>(1, div_up(4 * want_ctas, n_buf * n_class)));
        }
        const unsigned long long rseg = div_up(rows, n_seg);
        n_seg = div_up(rows, rseg);
        const double jobs = static_cast<double>(n_buf) * n_seg * n_class;
        const double busy = std::min(1.0, jobs / static_cast<double>(want_ctas));
        const double score = fill * busy / (1.0 + 2.0 / static_cast<double>(rseg) + 0.03);
        if (score > best) {
            best = score;
            pl.g.dv = sd.dv;
            pl.g.ng = static_cast<unsigned int>(div_up(sd.dv, 32));
            pl.g.rseg = static_cast<unsigned int>(rseg);
            pl.g.n_seg = static_cast<unsigned int>(n_seg);
            const unsigned long long lead = n_seg * rseg * sd.dv - nvec;  // virtual zero vectors in front of the buffer
            pl.g.skip_rows = static_cast<unsigned int>(lead / sd.dv);
            pl.g.lead = static_cast<unsigned int>(lead % sd.dv);
            for (int t = 0; t < 4; t++) pl.g.sh[t] = sd.k[t] & 31u;
            pl.tw = tw;
            pl.n_class = n_class;
        }
    }
    if (best < 0) return false;
    pl.nvec = nvec;
    pl.n_streams = n_buf * pl.g.n_seg;
    pl.dump_bytes = pl.n_streams * static_cast<size_t>(pl.g.dv) * 16;
    pl.qlog_bytes = pl.n_streams * static_cast<size_t>(pl.g.rseg) * pl.g.ng * 8;
    return true;
}

static int launch_crc(Dev &d, const CrcJob &j, cudaStream_t st);

// finisher for n_buf buffers whose dump rows and lost-bit logs are in place: CRC the dump rows with the table
// kernel, then fold everything into out[b * out_stride]
static int launch_sfold_finish(Dev &d, const SfoldPlan &pl, size_t n_buf, size_t len, const uint4 *dump, const uint2 *qlog,
                               uint32_t *dump_crc, uint32_t *out, size_t out_stride, cudaStream_t st, unsigned int q_group = 1) {
    const uint32_t P = kPolyReflected[0];
    CrcJob j;
    j.base = reinterpret_cast<const uint8_t *>(dump); j.n_buf = pl.n_streams; j.buf_stride = static_cast<size_t>(pl.g.dv) * 16;
    j.len = j.buf_stride; j.chunk = 0; j.poly = 0; j.out = dump_crc; j.out_stride = 1; j.no_sfold = true;
    RET(launch_crc(d, j, st));
    SfoldFinishParams f;
    memset(&f, 0, sizeof f);
    f.qlog = qlog; f.dump_crc = dump_crc; f.out = out; f.out_stride = out_stride; f.n_buf = n_buf;
    f.g = pl.g;
    f.q_group = q_group;
    f.per_thread = static_cast<unsigned int>(div_up(pl.g.ng, 256));
    f.poly_bits = P;
    f.xD = x8n_modp(16ull * pl.g.dv, P);
    f.x32 = x8n_modp(4, P);
    f.y = x8n_modp(512, P);
    uint32_t zp = 0x80000000u;
    for (unsigned int i = 0; i < f.per_thread; i++) zp = multmodp(zp, f.y, P);
    for (int L = 0; L < 8; L++) { f.z[L] = zp; zp = multmodp(zp, zp, P); }
    const unsigned int tail_cols = pl.g.dv - 32u * (pl.g.ng - 1);
    f.x_tail = x8n_modp(16ull * tail_cols + 4, P);
    f.cz_dump = crc_of_zeros(16ull * pl.g.dv, P);
    f.x_seg = x8n_modp(16ull * pl.g.dv * pl.g.rseg, P);
    f.x_inv_pad = xinv8n_modp(pl.nvec * 16 - len, P);
    f.cz_len = crc_of_zeros(len, P);
    const int grid = static_cast<int>(std::min<size_t>(n_buf, static_cast<size_t>(d.sm_count) * 8));  // latency-bound: fill the SMs with warps
    crc_sfold_finish_kernel<<<grid, 256, 0, st>>>(f);
    g_launches++;
    CK(cudaGetLastError());
    return GPU_EC_OK;
}

static void sfold_tail_mask(size_t len, uint32_t (&mask)[4]) {
    const size_t keep = len % 16 ? len % 16 : 16;
    for (int w = 0; w < 4; w++) {
        uint32_t m = 0;
        for (int b = 0; b < 4; b++) if (static_cast<size_t>(4 * w + b) < keep) m |= 0xFFu << (8 * b);
        mask[w] = m;
    }
}

// whole-buffer CRC-32 (IEEE) of n_buf 16-byte-aligned buffers, out[b * out_stride]
static int launch_sfold_crc(Dev &d, const SfoldPlan &pl, const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len,
                            uint32_t *out, size_t out_stride, cudaStream_t st) {
    uint8_t *scratch = nullptr;
    const size_t dump_off = 0, qlog_off = round_up(pl.dump_bytes, 256), crc_off = qlog_off + round_up(pl.qlog_bytes, 256);
    CK(cudaMallocAsync(&scratch, crc_off + 4 * pl.n_streams, st));
    SfoldParams p;
    memset(&p, 0, sizeof p);
    p.base = base; p.buf_stride = buf_stride; p.n_buf = n_buf; p.nvec = pl.nvec; p.g = pl.g; p.n_class = pl.n_class;
    sfold_tail_mask(len, p.tail_mask);
    p.dump = reinterpret_cast<uint4 *>(scratch + dump_off);
    p.qlog = reinterpret_cast<uint2 *>(scratch + qlog_off);
    const unsigned long long jobs = static_cast<unsigned long long>(n_buf) * pl.g.n_seg * pl.n_class;
    const int grid = static_cast<int>(std::min<unsigned long long>(jobs, static_cast<unsigned long long>(d.sm_count) * 2));
    crc_sfold_kernel<<<grid, pl.tw, 0, st>>>(p);
    g_launches++;
    int rc = (cudaGetLastError() == cudaSuccess) ? GPU_EC_OK : GPU_EC_ERR_CUDA;
    if (rc == GPU_EC_OK)
        rc = launch_sfold_finish(d, pl, n_buf, len, p.dump, p.qlog, reinterpret_cast<uint32_t *>(scratch + crc_off), out, out_stride, st);
    cudaFreeAsync(scratch, st);
    return rc;
}

static int launch_crc(Dev &d, const CrcJob &j, cudaStream_t st) {
    if (j.n_buf == 0) return GPU_EC_OK;
    const bool whole = (j.chunk == 0);
    if (j.len == 0) {
        if (whole && j.out) {  // crc32fast of nothing is 0
            if (j.out_stride == 1) CK(cudaMemsetAsync(j.out, 0, 4 * j.n_buf, st));
            else CK(cudaMemset2DAsync(j.out, 4 * j.out_stride, 0, 4, j.n_buf, st));
        }
        return GPU_EC_OK;
    }
    const bool aligned = !(reinterpret_cast<uintptr_t>(j.base) & 15) && !(j.buf_stride & 15) &&
                         !(j.expect_be && (reinterpret_cast<uintptr_t>(j.expect_be) & 3));
    if (whole && aligned && j.poly == 0 && !j.expect_be && j.out && !j.no_sfold && j.len >= (256u << 10) &&
        j.n_buf * j.len >= (8u << 20) && sfold_mode() == 2) {
        SfoldPlan pl;
        if (sfold_plan(j.n_buf, j.len, 0, static_cast<size_t>(d.sm_count) * 2, true, pl))
            return launch_sfold_crc(d, pl, j.base, j.n_buf, j.buf_stride, j.len, j.out, j.out_stride, st);
    }
    const unsigned long long lanes_wave = static_cast<unsigned long long>(d.sm_count) * CRC_THREADS;
    size_t G = 0;
    bool fold32 = false;
    if (aligned) {
        const unsigned long long total = static_cast<unsigned long long>(j.n_buf) * j.len;
        if (whole) {
            G = pick_whole_segment(d, j.n_buf, j.len);
        } else if (j.chunk % 128 == 0) {
            if (j.chunk <= 1024) {
                G = j.chunk;  // one lane per chunk, CRC written directly
            } else if (j.chunk % 4096 == 0 && j.chunk / 32 <= 4096) {
                G = j.chunk / 32;  // 16..128 KiB chunks: a warp's 32 lanes are one chunk, folded in the kernel
                fold32 = true;
            } else {
                for (size_t g : {size_t(4096), size_t(2048), size_t(1024), size_t(512)})
                    if (j.chunk % g == 0 && total / g >= 4 * lanes_wave) { G = g; break; }
                if (!G)
                    for (size_t g : {size_t(512), size_t(256), size_t(128)})
                        if (j.chunk % g == 0) { G = g; break; }
            }
        }
    }
    if (G == 0) {  // generic fallback kernel (odd chunk size / unaligned): one thread per chunk
        const size_t chunk = whole ? j.len : j.chunk;
        CrcSlowParams sp;
        memset(&sp, 0, sizeof sp);
        sp.base = j.base; sp.buf_stride = j.buf_stride; sp.n_buf = j.n_buf; sp.len = j.len; sp.chunk = chunk;
        sp.chunks_per_buf = div_up(j.len, chunk);
        sp.out = j.out; sp.out_stride = j.out_stride; sp.poly = j.poly; sp.big_endian = j.big_endian;
        sp.expect_be = j.expect_be; sp.first_bad = j.first_bad;
        const unsigned long long total = sp.n_buf * sp.chunks_per_buf;
        const int grid = static_cast<int>(std::min<unsigned long long>(div_up(total, 128), static_cast<unsigned long long>(d.sm_count) * 16));
        crc_chunks_slow_kernel<<<grid, 128, 0, st>>>(sp);
        g_launches++;
        CK(cudaGetLastError());
        return GPU_EC_OK;
    }
    CrcParams p;
    memset(&p, 0, sizeof p);
    p.base = j.base; p.buf_stride = j.buf_stride; p.n_buf = j.n_buf; p.len = j.len;
    p.G = static_cast<unsigned int>(G);
    p.nfull = j.len / G;
    p.rag = static_cast<unsigned int>(j.len % G);
    p.segs_per_buf = p.nfull + (p.rag ? 1 : 0);
    p.tiles_per_buf = div_up(p.segs_per_buf, 32);
    p.poly = j.poly;
    const bool direct = !whole && (G == j.chunk || fold32);
    if (fold32) {
        const uint32_t P = kPolyReflected[j.poly];
        p.fold32 = 1;
        p.f_poly = P;
        p.f_xrag = x8n_modp(p.rag, P);
        const uint32_t xg = x8n_modp(G, P);
        p.f_c1[0] = 0x80000000u;
        for (int e = 1; e < 32; e++) p.f_c1[e] = multmodp(p.f_c1[e - 1], xg, P);
    }
    uint32_t *scratch = nullptr;
    const size_t group = whole ? static_cast<size_t>(p.segs_per_buf) : j.chunk / G;
    const size_t n_part = static_cast<size_t>(j.n_buf) * p.segs_per_buf;
    if (direct) {
        p.out = j.out; p.out_stride = j.out_stride; p.big_endian = j.big_endian;
        p.expect_be = j.expect_be; p.first_bad = j.first_bad;
    } else {
        const size_t need2 = fold_scratch_words(j.n_buf * div_up(p.segs_per_buf, group), group);
        CK(cudaMallocAsync(&scratch, 4ull * (n_part + need2), st));
        p.out = scratch; p.out_stride = p.segs_per_buf;
    }
    const unsigned long long warps = j.n_buf * p.tiles_per_buf;
    const int grid = static_cast<int>(std::max<unsigned long long>(1, std::min<unsigned long long>(div_up(warps, CRC_WARPS), static_cast<unsigned long long>(d.sm_count))));
    if (fold32) crc_segments_kernel<true><<<grid, CRC_THREADS, CRC_SMEM_BYTES, st>>>(p);
    else crc_segments_kernel<false><<<grid, CRC_THREADS, CRC_SMEM_BYTES, st>>>(p);
    g_launches++;
    if (cudaError_t e = cudaGetLastError(); e != cudaSuccess) {
        if (scratch) cudaFreeAsync(scratch, st);
        CK(e);
    }
    if (!direct) {
        int rc = launch_combine(d, scratch, j.n_buf, p.segs_per_buf, G, j.len, group, j, st, scratch + n_part);
        cudaError_t fe = cudaFreeAsync(scratch, st);
        RET(rc);
        CK(fe);
    }
    return GPU_EC_OK;
}

// ---------------------------------------------------------------------------------------------
// fused encode + shard CRCs (fused_kernels.cuh): one pass over HBM for 8 <= k <= 12, m <= 4, IEEE.
template <int KH>
static cudaError_t launch_fused_x2(const FusedParams &fp, int n_clusters, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * n_clusters);
    cfg.blockDim = dim3(FUSED_THREADS);
    cfg.dynamicSmemBytes = fused_smem_bytes(KH, FUSED_THREADS);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, rs_lut8x2_crc_kernel<KH, FUSED_THREADS>, fp);
}
// n_in == 2*KH, n_out == 4 (RS(8,4), RS(10,4)): the trimmed kernel
template <int KH, bool CRC>
static cudaError_t launch_fused_fast(const FusedParams &fp, int n_clusters, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * n_clusters);
    cfg.blockDim = dim3(FUSED_THREADS);
    cfg.dynamicSmemBytes = LUT_FAST_RS_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, rs_x2_crc_fast_kernel<KH, FUSED_THREADS, CRC>, fp);
}

// plain apply-matrix (encode, or reconstruct of 4 shards) on the trimmed kernel: same walk, no checksums.
// Without a CRC the row length is free: rows of n_class * FUSED_THREADS vectors, about 32 rows per job,
// more classes (shorter jobs) when there are few stripes so that every cluster has work.
static int apply_x2_fast(Dev &d, const RsParams &rp, size_t n_stripes, cudaStream_t st) {
    FusedParams fp;
    memset(&fp, 0, sizeof fp);
    fp.rs = rp;
    const unsigned long long nvec = rp.nvec;
    const unsigned long long n_clusters_max = static_cast<unsigned long long>(d.sm_count / 2);
    unsigned long long n_class = div_up(nvec, 32ull * FUSED_THREADS);
    n_class = std::max<unsigned long long>(n_class, div_up(2 * n_clusters_max, n_stripes));
    n_class = std::min<unsigned long long>(n_class, div_up(nvec, FUSED_THREADS));
    const unsigned long long dv = n_class * FUSED_THREADS, R = div_up(nvec, dv);
    fp.g.dv = static_cast<unsigned int>(dv);
    fp.g.ng = static_cast<unsigned int>(dv / 32);
    fp.g.rseg = static_cast<unsigned int>(R);
    fp.g.n_seg = 1;
    fp.g.lead = static_cast<unsigned int>(R * dv - nvec);
    fp.n_class = static_cast<unsigned int>(n_class);
    fp.n_shards = 1;
    fp.n_jobs = n_stripes * n_class;
    static const bool prefetch = getenv("GPU_EC_X2_PREFETCH") != nullptr;  // kernel experiments only (see fused_kernels.cuh)
    fp.prefetch = prefetch ? 1 : 0;
    const int n_clusters = static_cast<int>(std::min<unsigned long long>(fp.n_jobs, n_clusters_max));
    const int kh = rp.n_in / 2;
    cudaError_t e = kh == 4 ? launch_fused_fast<4, false>(fp, n_clusters, st) : launch_fused_fast<5, false>(fp, n_clusters, st);
    g_launches++;
    CK(e);
    return GPU_EC_OK;
}

// -> *done = false when the shape is not one the fused kernel serves (caller runs the two-pass path)
static int encode_crc_fused(Dev &d, const gpu_ec_stripes &s, int k, int m, uint32_t *shard_crc, cudaStream_t st, bool *done) {
    *done = false;
    if (k > 2 * RS_X2_MAX_KH || m > RS_MAX_OUT || !sfold_enabled()) return GPU_EC_OK;
    const bool single = k <= RS_LUT8_MAX_IN;  // one CTA per SM holds all k tables: no cluster
    if (s.n_stripes == 0 || s.shard_len == 0 || s.n_stripes * static_cast<size_t>(k + m) >= (1ull << 32)) return GPU_EC_OK;
    const int n = k + m;
    const int n_clusters_max = single ? d.sm_count : d.sm_count / 2;  // CTAs (single) or 2-CTA clusters that run at once
    SfoldPlan pl;
    if (!sfold_plan(s.n_stripes, s.shard_len, FUSED_THREADS, static_cast<size_t>(n_clusters_max), false, pl)) return GPU_EC_OK;
    const unsigned long long n_jobs = static_cast<unsigned long long>(s.n_stripes) * pl.n_class;
    if (n_jobs < static_cast<unsigned long long>(n_clusters_max)) return GPU_EC_OK;  // too few jobs to fill the GPU
    std::vector<int> in(k), out(m);
    for (int i = 0; i < k; i++) in[i] = i;
    for (int i = 0; i < m; i++) out[i] = k + i;
    if (!slots_addressable(s, in.data(), k, out.data(), m)) return GPU_EC_OK;
    const Mat &M = coding_matrix(k, m);
    const uint8_t *rows = &M[static_cast<size_t>(k) * k];
    FusedParams fp;
    memset(&fp, 0, sizeof fp);
    RsParams &p = fp.rs;
    p.base = s.base; p.shard_stride = s.shard_stride; p.stripe_stride = s.stripe_stride;
    p.nvec = static_cast<unsigned int>(div_up(s.shard_len, 16));
    p.n_in = k; p.n_out = m;
    for (int r = 0; r < m; r++) { p.out_idx[r] = out[r]; p.out_off16[r] = static_cast<unsigned int>(out[r] * (s.shard_stride / 16)); }
    for (int j = 0; j < k; j++) { p.in_idx[j] = in[j]; p.in_off16[j] = static_cast<unsigned int>(in[j] * (s.shard_stride / 16)); }
    for (int r = 0; r < m; r++) for (int j = 0; j < k; j++) p.coef[r][j] = rows[static_cast<size_t>(r) * k + j];
    fp.g = pl.g;
    fp.n_class = pl.n_class;
    fp.n_shards = static_cast<unsigned int>(n);
    fp.n_jobs = n_jobs;
    sfold_tail_mask(s.shard_len, fp.tail_mask);
    const size_t n_streams = s.n_stripes * static_cast<size_t>(n);
    pl.n_streams = n_streams;
    const size_t dump_bytes = n_streams * static_cast<size_t>(pl.g.dv) * 16, qlog_bytes = n_streams * static_cast<size_t>(pl.g.rseg) * pl.g.ng * 8;
    const size_t qlog_off = round_up(dump_bytes, 256), crc_off = qlog_off + round_up(qlog_bytes, 256);
    uint8_t *scratch = nullptr;
    CK(cudaMallocAsync(&scratch, crc_off + 4 * n_streams, st));
    fp.dump = reinterpret_cast<uint4 *>(scratch);
    fp.qlog = reinterpret_cast<uint2 *>(scratch + qlog_off);
    const int n_clusters = static_cast<int>(std::min<unsigned long long>(n_jobs, static_cast<unsigned long long>(n_clusters_max)));
    const int kh = (k + 1) / 2;
    if (single) {
        const size_t smem = lut8_smem_bytes(k);
        switch (k) {
            case 1: rs_lut8_crc_kernel<1, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
            case 2: rs_lut8_crc_kernel<2, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
            case 3: rs_lut8_crc_kernel<3, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
            case 4: rs_lut8_crc_kernel<4, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
            case 5: rs_lut8_crc_kernel<5, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
            case 6: rs_lut8_crc_kernel<6, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
            default: rs_lut8_crc_kernel<7, FUSED_THREADS><<<n_clusters, FUSED_THREADS, smem, st>>>(fp); break;
        }
    }
    static const bool generic_only = getenv("GPU_EC_FUSED_GENERIC") != nullptr;  // kernel experiments only
    const bool fast = (k == 2 * kh) && m == RS_MAX_OUT && (kh == 4 || kh == 5) && !generic_only;
    cudaError_t e = single ? cudaGetLastError()
                  : fast ? (kh == 4 ? launch_fused_fast<4, true>(fp, n_clusters, st) : launch_fused_fast<5, true>(fp, n_clusters, st))
                  : kh <= 4 ? launch_fused_x2<4>(fp, n_clusters, st) : kh == 5 ? launch_fused_x2<5>(fp, n_clusters, st) : launch_fused_x2<6>(fp, n_clusters, st);
    g_launches++;
    int rc = GPU_EC_OK;
    if (e != cudaSuccess) { t_last_error = std::string("fused encode+crc launch failed: ") + cudaGetErrorString(e); rc = GPU_EC_ERR_CUDA; }
    if (rc == GPU_EC_OK)
        rc = launch_sfold_finish(d, pl, n_streams, s.shard_len, fp.dump, fp.qlog, reinterpret_cast<uint32_t *>(scratch + crc_off), shard_crc, 1, st,
                                 static_cast<unsigned int>(n));
    cudaFreeAsync(scratch, st);
    if (rc == GPU_EC_OK) *done = true;
    return rc;
}

// CRC of the original buffer from the data-shard CRCs: constants for stripe_full_crc_kernel.
// The original's `len` bytes cover n_full whole data shards plus (maybe) a `rem`-byte prefix of
// the next one, whose remaining S-rem bytes are the zero padding of erasure.rs:18.
static void full_crc_constants(FullCrcParams &f, size_t S, size_t original_len, int poly) {
    const uint32_t P = kPolyReflected[poly];
    const size_t rem = original_len % S;
    const size_t pad = rem ? S - rem : 0;
    f.n_full = static_cast<int>(original_len / S);
    f.has_part = rem ? 1 : 0;
    f.xS = x8n_modp(S, P);
    f.xLp = x8n_modp(rem, P);
    f.x_inv_pad = xinv8n_modp(pad, P);
    f.crc_zeros_pad = crc_of_zeros(pad, P);  // O(log pad): original_len may be far below k*S on the _dev entry
    f.poly_bits = P;
}
static uint32_t full_crc_host(const uint32_t *shard_crc, size_t S, size_t original_len, int poly) {
    FullCrcParams f;
    full_crc_constants(f, S, original_len, poly);
    uint32_t acc = 0;
    for (int j = 0; j < f.n_full; j++) acc = multmodp(acc, f.xS, f.poly_bits) ^ shard_crc[j];
    if (f.has_part) {
        const uint32_t part = multmodp(shard_crc[f.n_full] ^ f.crc_zeros_pad, f.x_inv_pad, f.poly_bits);
        acc = multmodp(acc, f.xLp, f.poly_bits) ^ part;
    }
    return acc;
}

static int check_km(int k, int m) {
    if (k <= 0 || m <= 0) return GPU_EC_ERR_ZERO_SHARDS;
    if (k + m > 256) return GPU_EC_ERR_TOO_MANY_SHARDS;
    return GPU_EC_OK;
}

static int encode_dev_impl(Dev &d, const gpu_ec_stripes &s, int k, int m, cudaStream_t st) {
    const Mat &M = coding_matrix(k, m);
    std::vector<int> in(k), out(m);
    for (int i = 0; i < k; i++) in[i] = i;
    for (int i = 0; i < m; i++) out[i] = k + i;
    return launch_apply(d, s, &M[static_cast<size_t>(k) * k], in.data(), k, out.data(), m, st);
}

static int reconstruct_plan(int k, int m, const uint8_t *present, DecodePlan &plan, bool &noop) {
    RET(check_km(k, m));
    int np = 0;
    for (int i = 0; i < k + m; i++) np += present[i] ? 1 : 0;
    noop = (np == k + m);
    if (noop) return GPU_EC_OK;
    if (np < k) return GPU_EC_ERR_TOO_FEW_PRESENT;
    int rc = decode_plan(k, m, present, plan);
    if (rc == -1) return GPU_EC_ERR_TOO_FEW_PRESENT;
    if (rc) return GPU_EC_ERR_SINGULAR;
    return GPU_EC_OK;
}

}  // namespace gpu_ec

using namespace gpu_ec;

// =============================================================================================
extern "C" {

int gpu_ec_init(void) {
    RET(ensure_init());
    Ctx &c = ctx();
    std::vector<int> act;
    { std::lock_guard<std::mutex> lk(c.mu); act = c.active; }
    for (int o : act) { Dev *d; RET(get_dev(o, &d)); }
    return GPU_EC_OK;
}

void gpu_ec_shutdown(void) {
    Ctx &c = ctx();
    std::lock_guard<std::mutex> lk(c.mu);
    for (auto &dp : c.devs) {
        Dev &d = *dp;
        std::lock_guard<std::mutex> lk2(d.mu);
        if (!d.ready) continue;
        cudaSetDevice(d.ordinal);
        for (auto &s : d.slots) {
            if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); s.stream = nullptr; }
            if (s.stream2) {
                cudaStreamSynchronize(s.stream2); cudaStreamDestroy(s.stream2); s.stream2 = nullptr;
                for (auto &e : s.ev) if (e) { cudaEventDestroy(e); e = nullptr; }
            }
            if (s.dbuf) { cudaFree(s.dbuf); s.dbuf = nullptr; s.cap = 0; }
            if (s.hbuf) { cudaFreeHost(s.hbuf); s.hbuf = nullptr; s.hcap = 0; }
        }
        d.ready = false;
    }
}

int gpu_ec_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int gpu_ec_set_devices(const int *devices, int n) {
    if (!devices || n <= 0) return GPU_EC_ERR_INVALID_ARG;
    const int have = gpu_ec_device_count();
    if (have <= 0) return GPU_EC_ERR_NO_DEVICE;
    for (int i = 0; i < n; i++) if (devices[i] < 0 || devices[i] >= have) return GPU_EC_ERR_INVALID_ARG;
    Ctx &c = ctx();
    std::lock_guard<std::mutex> lk(c.mu);
    c.active.clear();
    for (int i = 0; i < n; i++)  // a repeated ordinal would make a batch lease the same device's slots twice and wait on itself
        if (std::find(c.active.begin(), c.active.end(), devices[i]) == c.active.end()) c.active.push_back(devices[i]);
    return GPU_EC_OK;
}

const char *gpu_ec_strerror(int status) {
    switch (status) {
        case GPU_EC_OK: return "ok";
        case GPU_EC_ERR_ZERO_SHARDS: return "data_shards and parity_shards must both be > 0";
        case GPU_EC_ERR_EMPTY_DATA: return "data must not be empty";
        case GPU_EC_ERR_TOO_MANY_SHARDS: return "RS init error: TooManyShards";
        case GPU_EC_ERR_TOO_FEW_PRESENT: return "RS reconstruct error: TooFewShardsPresent";
        case GPU_EC_ERR_EMPTY_SHARD: return "RS reconstruct error: EmptyShard";
        case GPU_EC_ERR_SHARD_SIZE: return "RS reconstruct error: IncorrectShardSize";
        case GPU_EC_ERR_COUNT_MISMATCH: return "Checksum count mismatch";
        case GPU_EC_ERR_CRC_MISMATCH: return "Checksum mismatch at chunk";
        case GPU_EC_ERR_CUDA: return "CUDA error";
        case GPU_EC_ERR_OOM: return "out of device memory";
        case GPU_EC_ERR_INVALID_ARG: return "invalid argument";
        case GPU_EC_ERR_NO_DEVICE: return "no usable CUDA device (gpu_ec has no CPU fallback)";
        case GPU_EC_ERR_UNALIGNED: return "device pointer or stride not 16-byte aligned";
        case GPU_EC_ERR_SINGULAR: return "RS reconstruct error: singular decode matrix";
        default: return "unknown gpu_ec status";
    }
}

const char *gpu_ec_last_error(void) { return t_last_error.c_str(); }

int gpu_ec_alloc_pinned(void **ptr, size_t bytes) {
    if (!ptr) return GPU_EC_ERR_INVALID_ARG;
    RET(ensure_init());
    CK(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocPortable));
    return GPU_EC_OK;
}
int gpu_ec_free_pinned(void *ptr) {
    if (!ptr) return GPU_EC_OK;
    CK(cudaFreeHost(ptr));
    return GPU_EC_OK;
}

size_t gpu_ec_shard_len(size_t len, size_t k) { return k ? (len + k - 1) / k : 0; }

int gpu_ec_rs_matrix(int k, int m, uint8_t *out) {
    RET(check_km(k, m));
    if (!out) return GPU_EC_ERR_INVALID_ARG;
    const Mat &M = coding_matrix(k, m);
    memcpy(out, M.data(), M.size());
    return GPU_EC_OK;
}

int gpu_ec_rs_decode_matrix(int k, int m, const uint8_t *present, uint8_t *rows_out, int *missing_out,
                            int *survivors_out, int *n_missing_out) {
    if (!present || !rows_out || !missing_out || !survivors_out || !n_missing_out) return GPU_EC_ERR_INVALID_ARG;
    DecodePlan plan;
    bool noop = false;
    RET(reconstruct_plan(k, m, present, plan, noop));
    if (noop) {
        *n_missing_out = 0;
        for (int i = 0; i < k; i++) survivors_out[i] = i;
        return GPU_EC_OK;
    }
    *n_missing_out = static_cast<int>(plan.missing.size());
    memcpy(rows_out, plan.rows.data(), plan.rows.size());
    for (size_t i = 0; i < plan.missing.size(); i++) missing_out[i] = plan.missing[i];
    for (int i = 0; i < k; i++) survivors_out[i] = plan.survivors[i];
    return GPU_EC_OK;
}

void gpu_ec_meta_encode_be(const uint32_t *crcs, size_t n, uint8_t *out) {
    for (size_t i = 0; i < n; i++) {
        out[4 * i + 0] = static_cast<uint8_t>(crcs[i] >> 24);
        out[4 * i + 1] = static_cast<uint8_t>(crcs[i] >> 16);
        out[4 * i + 2] = static_cast<uint8_t>(crcs[i] >> 8);
        out[4 * i + 3] = static_cast<uint8_t>(crcs[i]);
    }
}

uint32_t gpu_ec_crc32_combine(uint32_t crc1, uint32_t crc2, size_t len2, int poly) {
    const uint32_t P = kPolyReflected[poly ? 1 : 0];
    return multmodp(x8n_modp(len2, P), crc1, P) ^ crc2;
}

uint64_t gpu_ec_kernel_launch_count(void) { return g_launches.load(); }

// ---------------------------------------------------------------------------------------------
// device-resident entry points

int gpu_ec_encode_dev(const gpu_ec_stripes *s, int k, int m, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    RET(check_km(k, m));
    RET(check_stripes(s));
    Dev *d;
    RET(dev_of_ptr(s->base, &d));
    return encode_dev_impl(*d, *s, k, m, static_cast<cudaStream_t>(stream));
}

int gpu_ec_reconstruct_dev(const gpu_ec_stripes *s, const uint8_t *present, int k, int m, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!present) return GPU_EC_ERR_INVALID_ARG;
    DecodePlan plan;
    bool noop = false;
    RET(reconstruct_plan(k, m, present, plan, noop));
    if (noop) return GPU_EC_OK;
    RET(check_stripes(s));
    if (s->shard_len == 0) return GPU_EC_ERR_EMPTY_SHARD;
    Dev *d;
    RET(dev_of_ptr(s->base, &d));
    return launch_apply(*d, *s, plan.rows.data(), plan.survivors.data(), k, plan.missing.data(),
                        static_cast<int>(plan.missing.size()), static_cast<cudaStream_t>(stream));
}

int gpu_ec_apply_matrix_dev(const gpu_ec_stripes *s, const uint8_t *rows, const int *in_idx, int n_in,
                            const int *out_idx, int n_out, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!rows || !in_idx || !out_idx || n_in <= 0 || n_out <= 0 || n_in > 256 || n_out > 256) return GPU_EC_ERR_INVALID_ARG;
    RET(check_stripes(s));
    Dev *d;
    RET(dev_of_ptr(s->base, &d));
    return launch_apply(*d, *s, rows, in_idx, n_in, out_idx, n_out, static_cast<cudaStream_t>(stream));
}

int gpu_ec_crc32_chunks_dev(const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len, size_t chunk,
                            int poly, uint32_t *out, size_t out_stride, int big_endian, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!base || !out || chunk == 0 || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    Dev *d;
    RET(dev_of_ptr(base, &d));
    CrcJob j;
    j.base = base; j.n_buf = n_buf; j.buf_stride = buf_stride; j.len = len; j.chunk = chunk; j.poly = poly;
    j.out = out; j.out_stride = out_stride; j.big_endian = big_endian;
    return launch_crc(*d, j, static_cast<cudaStream_t>(stream));
}

int gpu_ec_crc32_dev(const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len, int poly, uint32_t *out,
                     void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!base || !out || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    Dev *d;
    RET(dev_of_ptr(base, &d));
    CrcJob j;
    j.base = base; j.n_buf = n_buf; j.buf_stride = buf_stride; j.len = len; j.chunk = 0; j.poly = poly;
    j.out = out; j.out_stride = 1;
    return launch_crc(*d, j, static_cast<cudaStream_t>(stream));
}

int gpu_ec_verify_chunks_dev(const uint8_t *data, size_t len, size_t chunk, int poly, const uint8_t *meta_be,
                             long long *first_bad, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!data || !meta_be || !first_bad || chunk == 0 || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    Dev *d;
    RET(dev_of_ptr(data, &d));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CK(cudaMemsetAsync(first_bad, 0xFF, 8, st));
    CrcJob j;
    j.base = data; j.n_buf = 1; j.buf_stride = round_up(len, 16); j.len = len; j.chunk = chunk; j.poly = poly;
    j.expect_be = meta_be; j.first_bad = reinterpret_cast<unsigned long long *>(first_bad);
    RET(launch_crc(*d, j, st));
    first_bad_finish_kernel<<<1, 1, 0, st>>>(reinterpret_cast<unsigned long long *>(first_bad));
    g_launches++;
    CK(cudaGetLastError());
    return GPU_EC_OK;
}

int gpu_ec_encode_crc_dev(const gpu_ec_stripes *s, int k, int m, size_t original_len, int poly,
                          uint32_t *shard_crc, uint32_t *full_crc, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    RET(check_km(k, m));
    RET(check_stripes(s));
    if (!shard_crc || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    const size_t S = s->shard_len;
    if (original_len == 0) original_len = S * static_cast<size_t>(k);
    if (S == 0 || original_len > S * static_cast<size_t>(k)) return GPU_EC_ERR_INVALID_ARG;
    Dev *d;
    RET(dev_of_ptr(s->base, &d));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int n = k + m;
    // encode every stripe, then checksum every shard: two streaming passes. (Per-stripe-group launch
    // chains that keep the parity L2-warm were measured slower: small launches underfill the GPU.)
    bool fused = false;
    if (poly == GPU_EC_POLY_IEEE) RET(encode_crc_fused(*d, *s, k, m, shard_crc, st, &fused));
    if (!fused) RET(encode_dev_impl(*d, *s, k, m, st));
    if (!fused) {
        CrcJob j;
        j.buf_stride = s->shard_stride; j.len = S; j.chunk = 0; j.poly = poly; j.out_stride = 1;
        if (s->stripe_stride == static_cast<size_t>(n) * s->shard_stride) {  // all shards at one stride
            j.base = s->base; j.n_buf = s->n_stripes * n; j.out = shard_crc;
            RET(launch_crc(*d, j, st));
        } else {
            for (size_t i = 0; i < s->n_stripes; i++) {
                j.base = s->base + i * s->stripe_stride; j.n_buf = n; j.out = shard_crc + i * n;
                RET(launch_crc(*d, j, st));
            }
        }
    }
    if (full_crc) {
        FullCrcParams f;
        memset(&f, 0, sizeof f);
        full_crc_constants(f, S, original_len, poly);
        f.shard_crc = shard_crc; f.full_crc = full_crc; f.n_stripes = s->n_stripes; f.n_shards = n;
        const int grid = static_cast<int>(std::min<size_t>(div_up(s->n_stripes, 128), 1024));
        stripe_full_crc_kernel<<<grid, 128, 0, st>>>(f);
        g_launches++;
        CK(cudaGetLastError());
    }
    return GPU_EC_OK;
}

int gpu_ec_fill_synthetic_dev(uint8_t *out, size_t len, uint64_t seed, uint64_t byte_offset, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!out) return GPU_EC_ERR_INVALID_ARG;
    if ((reinterpret_cast<uintptr_t>(out) & 7) || (byte_offset & 7)) return GPU_EC_ERR_UNALIGNED;
    if (len == 0) return GPU_EC_OK;
    Dev *d;
    RET(dev_of_ptr(out, &d));
    const size_t words = div_up(len, 8);
    const int grid = static_cast<int>(std::min<size_t>(div_up(words, 256), static_cast<size_t>(d->sm_count) * 16));
    fill_synthetic_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(out, len, seed, byte_offset);
    g_launches++;
    CK(cudaGetLastError());
    return GPU_EC_OK;
}

// ---------------------------------------------------------------------------------------------
// host-pointer entry points

int gpu_ec_encode(const uint8_t *data, size_t len, int k, int m, uint8_t *const *shards_out,
                  uint32_t *shard_crc_out, uint32_t *full_crc_out, uint8_t *const *meta_out) {
    GPU_EC_TRACE();
    if (k <= 0 || m <= 0) return GPU_EC_ERR_ZERO_SHARDS;   // erasure.rs:8-10 (checked first there too)
    if (len == 0) return GPU_EC_ERR_EMPTY_DATA;            // erasure.rs:11-13
    if (k + m > 256) return GPU_EC_ERR_TOO_MANY_SHARDS;    // ReedSolomon::new
    if (!data || !shards_out) return GPU_EC_ERR_INVALID_ARG;
    for (int r = 0; r < m; r++) if (!shards_out[k + r]) return GPU_EC_ERR_INVALID_ARG;  // before any async copy is queued
    const int n = k + m;
    const size_t S = gpu_ec_shard_len(len, k);
    const size_t pitch = round_up(S, 256);
    const size_t n_meta = div_up(S, 512);
    const size_t crc_off = static_cast<size_t>(n) * pitch;
    const size_t meta_off = crc_off + round_up(4 * static_cast<size_t>(n), 256);
    const size_t fold_off = meta_off + round_up(meta_out ? 4 * n_meta * n : 0, 256);
    const size_t need = fold_off + (meta_out ? 4 * fold_scratch_words(n, n_meta) : 0);
    Lease L;
    RET(lease_any(L));
    RET(slot_reserve(*L.s, need));
    cudaStream_t st = L.s->stream;
    uint8_t *db = L.s->dbuf;
    // H2D: the contiguous input becomes k pitched rows; the tail of the last rows is zero padding
    const size_t full_rows = len / S, rem = len - full_rows * S;
    gpu_ec_stripes sv{db, pitch, static_cast<size_t>(n) * pitch, S, 1};
    const bool want_crc = shard_crc_out || full_crc_out;
    uint32_t *d_crc = reinterpret_cast<uint32_t *>(db + crc_off);
    bool fused = false;  // one pass for parity + shard CRCs where the fused kernel serves the shape
    // A large call that has the device to itself is latency-bound by its own copies (256 MiB: 4.8 ms in, 2.4 ms out,
    // 0.1 ms of kernels): walk the stripe in column chunks so the parity of chunk c goes home on a second stream
    // while chunk c+1 comes in. With other callers on the device the slots already overlap each other.
    static const bool no_chunks = getenv("GPU_EC_NO_CHUNKED_COPY") != nullptr;  // experiments only
    const bool chunked = L.alone && len >= (64u << 20) && !no_chunks;
    bool parity_home = false;
    if (chunked) {
        RET(slot_overlap_init(*L.s));
        cudaStream_t st2 = L.s->stream2;
        for (size_t r = full_rows; r < static_cast<size_t>(k); r++) {
            const size_t o = (r == full_rows) ? rem : 0;
            CK(cudaMemsetAsync(db + r * pitch + o, 0, S - o, st));
        }
        const size_t width = round_up(div_up(S, 8), 4096);
        int c = 0;
        for (size_t c0 = 0; c0 < S; c0 += width, c++) {
            const size_t w = std::min(width, S - c0);
            if (full_rows) CK(cudaMemcpy2DAsync(db + c0, pitch, data + c0, S, w, full_rows, cudaMemcpyHostToDevice, st));
            if (rem > c0) CK(cudaMemcpyAsync(db + full_rows * pitch + c0, data + full_rows * S + c0, std::min(w, rem - c0), cudaMemcpyHostToDevice, st));
            gpu_ec_stripes blk{db + c0, pitch, static_cast<size_t>(n) * pitch, w, 1};
            RET(encode_dev_impl(*L.d, blk, k, m, st));
            CK(cudaEventRecord(L.s->ev[c], st));
            CK(cudaStreamWaitEvent(st2, L.s->ev[c], 0));
            for (int r = 0; r < m; r++)
                CK(cudaMemcpyAsync(shards_out[k + r] + c0, db + static_cast<size_t>(k + r) * pitch + c0, w, cudaMemcpyDeviceToHost, st2));
        }
        parity_home = true;
    } else {
        if (full_rows) CK(cudaMemcpy2DAsync(db, pitch, data, S, S, full_rows, cudaMemcpyHostToDevice, st));
        if (rem) CK(cudaMemcpyAsync(db + full_rows * pitch, data + full_rows * S, rem, cudaMemcpyHostToDevice, st));
        for (size_t r = full_rows; r < static_cast<size_t>(k); r++) {
            const size_t o = (r == full_rows) ? rem : 0;
            CK(cudaMemsetAsync(db + r * pitch + o, 0, S - o, st));
        }
        if (want_crc && !meta_out) RET(encode_crc_fused(*L.d, sv, k, m, d_crc, st, &fused));
        if (!fused) RET(encode_dev_impl(*L.d, sv, k, m, st));
    }
    std::vector<uint32_t> crc_tmp;
    if (meta_out) {
        CrcJob j;
        j.base = db; j.n_buf = n; j.buf_stride = pitch; j.len = S; j.chunk = 512;
        j.out = reinterpret_cast<uint32_t *>(db + meta_off); j.out_stride = n_meta; j.big_endian = 1;
        RET(launch_crc(*L.d, j, st));
        for (int i = 0; i < n; i++)
            if (meta_out[i]) CK(cudaMemcpyAsync(meta_out[i], db + meta_off + 4 * n_meta * i, 4 * n_meta, cudaMemcpyDeviceToHost, st));
        if (want_crc) {  // whole-shard CRCs folded from the .meta words: no second pass over the shards
            CrcJob f;
            f.poly = GPU_EC_POLY_IEEE; f.out = d_crc; f.out_stride = 1;
            RET(launch_combine(*L.d, j.out, n, n_meta, 512, S, n_meta, f, st, reinterpret_cast<uint32_t *>(db + fold_off), true));
        }
    } else if (want_crc && !fused) {
        CrcJob j;
        j.base = db; j.n_buf = n; j.buf_stride = pitch; j.len = S; j.chunk = 0;
        j.out = d_crc; j.out_stride = 1;
        RET(launch_crc(*L.d, j, st));
    }
    if (want_crc) {
        crc_tmp.resize(n);
        CK(cudaMemcpyAsync(crc_tmp.data(), d_crc, 4 * static_cast<size_t>(n), cudaMemcpyDeviceToHost, st));
    }
    for (int r = 0; r < m && !parity_home; r++) {
        CK(cudaMemcpyAsync(shards_out[k + r], db + static_cast<size_t>(k + r) * pitch, S, cudaMemcpyDeviceToHost, st));
    }
    // systematic data shards are the padded input verbatim (erasure.rs:17-19): host copy while the GPU works
    for (int j = 0; j < k; j++) {
        if (!shards_out[j]) continue;  // caller keeps its own view of the systematic shards
        const size_t off = static_cast<size_t>(j) * S;
        const size_t have = off < len ? std::min(S, len - off) : 0;
        if (have) memcpy(shards_out[j], data + off, have);
        if (have < S) memset(shards_out[j] + have, 0, S - have);
    }
    CK(cudaStreamSynchronize(st));
    if (parity_home) CK(cudaStreamSynchronize(L.s->stream2));
    if (shard_crc_out) memcpy(shard_crc_out, crc_tmp.data(), 4 * static_cast<size_t>(n));
    if (full_crc_out) *full_crc_out = full_crc_host(crc_tmp.data(), S, len, GPU_EC_POLY_IEEE);
    return GPU_EC_OK;
}

int gpu_ec_reconstruct(uint8_t *const *shards, const uint8_t *present, size_t shard_len, int k, int m) {
    GPU_EC_TRACE();
    if (!shards || !present) return GPU_EC_ERR_INVALID_ARG;
    DecodePlan plan;
    bool noop = false;
    RET(reconstruct_plan(k, m, present, plan, noop));
    if (noop) return GPU_EC_OK;
    if (shard_len == 0) return GPU_EC_ERR_EMPTY_SHARD;
    const int n = k + m;
    const size_t S = shard_len, pitch = round_up(S, 256);
    for (int idx : plan.survivors) if (!shards[idx]) return GPU_EC_ERR_INVALID_ARG;
    for (int idx : plan.missing) if (!shards[idx]) return GPU_EC_ERR_INVALID_ARG;
    Lease L;
    RET(lease_any(L));
    RET(slot_reserve(*L.s, static_cast<size_t>(n) * pitch));
    cudaStream_t st = L.s->stream;
    uint8_t *db = L.s->dbuf;
    static const bool no_chunks = getenv("GPU_EC_NO_CHUNKED_COPY") != nullptr;  // experiments only
    if (L.alone && S * static_cast<size_t>(k) >= (64u << 20) && !no_chunks) {
        // alone on the device: column chunks, rebuilt shards go home on a second stream while the next chunk comes in
        // (same scheme as gpu_ec_encode)
        RET(slot_overlap_init(*L.s));
        cudaStream_t st2 = L.s->stream2;
        const size_t width = round_up(div_up(S, 8), 4096);
        int c = 0;
        for (size_t c0 = 0; c0 < S; c0 += width, c++) {
            const size_t w = std::min(width, S - c0);
            for (int j = 0; j < k; j++) {
                const int idx = plan.survivors[j];
                CK(cudaMemcpyAsync(db + static_cast<size_t>(idx) * pitch + c0, shards[idx] + c0, w, cudaMemcpyHostToDevice, st));
            }
            gpu_ec_stripes blk{db + c0, pitch, static_cast<size_t>(n) * pitch, w, 1};
            RET(launch_apply(*L.d, blk, plan.rows.data(), plan.survivors.data(), k, plan.missing.data(),
                             static_cast<int>(plan.missing.size()), st));
            CK(cudaEventRecord(L.s->ev[c], st));
            CK(cudaStreamWaitEvent(st2, L.s->ev[c], 0));
            for (int idx : plan.missing)
                CK(cudaMemcpyAsync(shards[idx] + c0, db + static_cast<size_t>(idx) * pitch + c0, w, cudaMemcpyDeviceToHost, st2));
        }
        CK(cudaStreamSynchronize(st));
        CK(cudaStreamSynchronize(st2));
        return GPU_EC_OK;
    }
    for (int j = 0; j < k; j++) {
        const int idx = plan.survivors[j];
        CK(cudaMemcpyAsync(db + static_cast<size_t>(idx) * pitch, shards[idx], S, cudaMemcpyHostToDevice, st));
    }
    gpu_ec_stripes sv{db, pitch, static_cast<size_t>(n) * pitch, S, 1};
    RET(launch_apply(*L.d, sv, plan.rows.data(), plan.survivors.data(), k, plan.missing.data(),
                     static_cast<int>(plan.missing.size()), st));
    for (int idx : plan.missing) {
        CK(cudaMemcpyAsync(shards[idx], db + static_cast<size_t>(idx) * pitch, S, cudaMemcpyDeviceToHost, st));
    }
    CK(cudaStreamSynchronize(st));
    return GPU_EC_OK;
}

int gpu_ec_reconstruct_only(uint8_t *const *shards, const uint8_t *present, size_t shard_len, int k, int m,
                            const int *want, int n_want) {
    GPU_EC_TRACE();
    if (!shards || !present || !want || n_want <= 0) return GPU_EC_ERR_INVALID_ARG;
    DecodePlan plan;
    bool noop = false;
    RET(reconstruct_plan(k, m, present, plan, noop));
    const int n = k + m;
    for (int i = 0; i < n_want; i++)
        if (want[i] < 0 || want[i] >= n || present[want[i]] || !shards[want[i]]) return GPU_EC_ERR_INVALID_ARG;
    if (shard_len == 0) return GPU_EC_ERR_EMPTY_SHARD;
    // the decode rows of just the wanted slots, in the caller's order
    std::vector<uint8_t> rows(static_cast<size_t>(n_want) * k);
    std::vector<int> outs(n_want);
    for (int i = 0; i < n_want; i++) {
        const auto it = std::find(plan.missing.begin(), plan.missing.end(), want[i]);
        if (it == plan.missing.end()) return GPU_EC_ERR_INVALID_ARG;
        memcpy(&rows[static_cast<size_t>(i) * k], &plan.rows[static_cast<size_t>(it - plan.missing.begin()) * k], k);
        outs[i] = want[i];
    }
    const size_t S = shard_len, pitch = round_up(S, 256);
    Lease L;
    RET(lease_any(L));
    RET(slot_reserve(*L.s, static_cast<size_t>(n) * pitch));
    cudaStream_t st = L.s->stream;
    uint8_t *db = L.s->dbuf;
    for (int j = 0; j < k; j++) {
        const int idx = plan.survivors[j];
        if (!shards[idx]) return GPU_EC_ERR_INVALID_ARG;
        CK(cudaMemcpyAsync(db + static_cast<size_t>(idx) * pitch, shards[idx], S, cudaMemcpyHostToDevice, st));
    }
    gpu_ec_stripes sv{db, pitch, static_cast<size_t>(n) * pitch, S, 1};
    RET(launch_apply(*L.d, sv, rows.data(), plan.survivors.data(), k, outs.data(), n_want, st));
    for (int i = 0; i < n_want; i++)
        CK(cudaMemcpyAsync(shards[want[i]], db + static_cast<size_t>(want[i]) * pitch, S, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return GPU_EC_OK;
}

int gpu_ec_decode(uint8_t *const *shards, uint8_t *present, size_t shard_len, int k, int m, size_t original_len,
                  uint8_t *out) {
    GPU_EC_TRACE();
    if (!shards || !present || (!out && original_len)) return GPU_EC_ERR_INVALID_ARG;
    RET(gpu_ec_reconstruct(shards, present, shard_len, k, m));
    for (int i = 0; i < k + m; i++) present[i] = 1;
    size_t left = std::min(original_len, shard_len * static_cast<size_t>(k));  // Vec::truncate never grows
    for (int j = 0; j < k && left; j++) {
        const size_t c = std::min(left, shard_len);
        memcpy(out + static_cast<size_t>(j) * shard_len, shards[j], c);
        left -= c;
    }
    return GPU_EC_OK;
}

static int crc_host_common(const uint8_t *data, size_t len, size_t chunk, int poly, uint32_t *whole_out,
                           uint32_t *chunk_out, const uint8_t *meta_be, int64_t *first_bad) {
    const size_t nchunks = chunk ? div_up(len, chunk) : 0;
    const size_t data_cap = round_up(len, 256);
    const size_t crc_off = data_cap;
    const size_t whole_off = crc_off + round_up(4 * nchunks, 256);
    const size_t meta_off = whole_off + 256;
    const size_t fold_off = meta_off + round_up(4 * nchunks, 256) + 256;
    const size_t need = fold_off + 4 * fold_scratch_words(1, nchunks) + 256;
    Lease L;
    RET(lease_any(L));
    RET(slot_reserve(*L.s, need));
    cudaStream_t st = L.s->stream;
    uint8_t *db = L.s->dbuf;
    if (len) CK(cudaMemcpyAsync(db, data, len, cudaMemcpyHostToDevice, st));
    uint32_t *d_chunks = reinterpret_cast<uint32_t *>(db + crc_off);
    uint32_t *d_whole = reinterpret_cast<uint32_t *>(db + whole_off);
    unsigned long long *d_bad = reinterpret_cast<unsigned long long *>(db + whole_off + 8);
    if (meta_be) {
        if (nchunks) CK(cudaMemcpyAsync(db + meta_off, meta_be, 4 * nchunks, cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(d_bad, 0xFF, 8, st));
    }
    if (chunk && nchunks) {
        CrcJob j;
        j.base = db; j.n_buf = 1; j.buf_stride = data_cap; j.len = len; j.chunk = chunk; j.poly = poly;
        j.out = d_chunks; j.out_stride = nchunks;
        if (meta_be) { j.expect_be = db + meta_off; j.first_bad = d_bad; }
        RET(launch_crc(*L.d, j, st));
        if (chunk_out) CK(cudaMemcpyAsync(chunk_out, d_chunks, 4 * nchunks, cudaMemcpyDeviceToHost, st));
    }
    if (whole_out) {
        if (chunk && nchunks) {  // fold the chunk CRCs: the bytes were read once (chunkserver.rs:746-769 + 195)
            CrcJob j;
            j.poly = poly; j.out = d_whole; j.out_stride = 1;
            RET(launch_combine(*L.d, d_chunks, 1, nchunks, chunk, len, nchunks, j, st, reinterpret_cast<uint32_t *>(db + fold_off)));
        } else {
            CrcJob j;
            j.base = db; j.n_buf = 1; j.buf_stride = data_cap; j.len = len; j.chunk = 0; j.poly = poly;
            j.out = d_whole; j.out_stride = 1;
            RET(launch_crc(*L.d, j, st));
        }
        CK(cudaMemcpyAsync(whole_out, d_whole, 4, cudaMemcpyDeviceToHost, st));
    }
    unsigned long long bad = ~0ull;
    if (meta_be) CK(cudaMemcpyAsync(&bad, d_bad, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (meta_be) {
        if (first_bad) *first_bad = (bad == ~0ull) ? -1 : static_cast<int64_t>(bad);
        if (bad != ~0ull) return GPU_EC_ERR_CRC_MISMATCH;
    }
    return GPU_EC_OK;
}

int gpu_ec_crc32(const uint8_t *data, size_t len, int poly, uint32_t *out) {
    GPU_EC_TRACE();
    if (!out || (!data && len) || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    return crc_host_common(data, len, 0, poly, out, nullptr, nullptr, nullptr);
}

int gpu_ec_crc32_chunks(const uint8_t *data, size_t len, size_t chunk, int poly, uint32_t *out) {
    GPU_EC_TRACE();
    if ((!out && len) || (!data && len) || chunk == 0 || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    if (len == 0) { RET(ensure_init()); return GPU_EC_OK; }  // zero chunks, empty .meta (chunkserver.rs:184)
    return crc_host_common(data, len, chunk, poly, nullptr, out, nullptr, nullptr);
}

int gpu_ec_verify_chunks(const uint8_t *data, size_t len, size_t chunk, int poly, const uint8_t *meta_be,
                         size_t meta_len, int64_t *first_bad) {
    GPU_EC_TRACE();
    if ((!data && len) || (!meta_be && meta_len) || chunk == 0 || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    if (first_bad) *first_bad = -1;
    RET(ensure_init());
    const size_t n_exp = meta_len / 4, n_act = div_up(len, chunk);  // chunks_exact(4) drops a ragged tail
    if (n_exp != n_act) return GPU_EC_ERR_COUNT_MISMATCH;           // chunkserver.rs:277-279
    if (n_act == 0) return GPU_EC_OK;
    return crc_host_common(data, len, chunk, poly, nullptr, nullptr, meta_be, first_bad);
}

int gpu_ec_block_checksums(const uint8_t *data, size_t len, size_t chunk, int poly, uint32_t *whole_crc_out,
                           uint32_t *chunk_crc_out) {
    GPU_EC_TRACE();
    if ((!data && len) || chunk == 0 || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    if (len == 0) { RET(ensure_init()); if (whole_crc_out) *whole_crc_out = 0; return GPU_EC_OK; }
    return crc_host_common(data, len, chunk, poly, whole_crc_out, chunk_crc_out, nullptr, nullptr);
}

int gpu_ec_block_checksums_batch(size_t n, const uint8_t *const *data, const size_t *lens, size_t chunk, int poly,
                                 uint32_t *whole_crc_out, uint32_t *const *chunk_crc_out) {
    GPU_EC_TRACE();
    if (n == 0) { RET(ensure_init()); return GPU_EC_OK; }
    if (!data || !lens || chunk == 0 || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    for (size_t b = 0; b < n; b++) if (!data[b] && lens[b]) return GPU_EC_ERR_INVALID_ARG;
    if (chunk % 128 || chunk > 1024) {  // chunk sizes the one-launch layout does not serve: block by block
        for (size_t b = 0; b < n; b++)
            RET(gpu_ec_block_checksums(data[b], lens[b], chunk, poly, whole_crc_out ? whole_crc_out + b : nullptr,
                                       chunk_crc_out ? chunk_crc_out[b] : nullptr));
        return GPU_EC_OK;
    }
    const uint32_t P = kPolyReflected[poly];
    // layout: blocks back to back, each starting on a chunk boundary (so chunk i of the staging buffer is a chunk
    // of exactly one block); a short last chunk is hashed with its zero padding and repaired afterwards
    std::vector<BlockDesc> desc(n);
    std::vector<size_t> off(n);
    size_t total = 0, n_chunks = 0;
    for (size_t b = 0; b < n; b++) {
        off[b] = total;
        BlockDesc &d = desc[b];
        memset(&d, 0, sizeof d);
        d.first_chunk = n_chunks;
        d.n_full = static_cast<unsigned int>(lens[b] / chunk);
        d.rag = static_cast<unsigned int>(lens[b] % chunk);
        if (d.rag) {
            d.x_inv_pad = xinv8n_modp(chunk - d.rag, P);
            d.cz_rag = crc_of_zeros(d.rag, P);
            d.x_rag = x8n_modp(d.rag, P);
        }
        const size_t nc = div_up(lens[b], chunk);
        if (lens[b] / chunk >= (1ull << 32)) return GPU_EC_ERR_INVALID_ARG;
        n_chunks += nc;
        total += nc * chunk;
    }
    const size_t crc_off = round_up(std::max<size_t>(total, 256), 256);
    const size_t desc_off = crc_off + round_up(4 * n_chunks + 4, 256);
    const size_t whole_off = desc_off + round_up(sizeof(BlockDesc) * n, 256);
    const size_t need = whole_off + round_up(4 * n, 256);
    Lease L;
    RET(lease_any(L));
    RET(slot_reserve(*L.s, need));
    cudaStream_t st = L.s->stream;
    uint8_t *db = L.s->dbuf;
    // small batches are packed on the host into pinned staging: one H2D for all blocks (and one D2H for all CRCs)
    // instead of a copy per block; large ones copy block by block straight from the caller's memory
    const bool packed = total <= (64u << 20);
    if (packed) {
        RET(slot_reserve_host(*L.s, std::max(total, 4 * n_chunks + 4 * n)));
        for (size_t b = 0; b < n; b++) {
            if (lens[b]) memcpy(L.s->hbuf + off[b], data[b], lens[b]);
            const size_t pad = div_up(lens[b], chunk) * chunk - lens[b];
            if (pad) memset(L.s->hbuf + off[b] + lens[b], 0, pad);
        }
        if (total) CK(cudaMemcpyAsync(db, L.s->hbuf, total, cudaMemcpyHostToDevice, st));
    } else {
        for (size_t b = 0; b < n; b++) {
            if (!lens[b]) continue;
            CK(cudaMemcpyAsync(db + off[b], data[b], lens[b], cudaMemcpyHostToDevice, st));
            const size_t pad = div_up(lens[b], chunk) * chunk - lens[b];
            if (pad) CK(cudaMemsetAsync(db + off[b] + lens[b], 0, pad, st));
        }
    }
    CK(cudaMemcpyAsync(db + desc_off, desc.data(), sizeof(BlockDesc) * n, cudaMemcpyHostToDevice, st));
    uint32_t *d_crc = reinterpret_cast<uint32_t *>(db + crc_off);
    uint32_t *d_whole = reinterpret_cast<uint32_t *>(db + whole_off);
    if (n_chunks) {
        CrcJob j;  // ONE launch over every chunk of every block
        j.base = db; j.n_buf = 1; j.buf_stride = crc_off; j.len = total; j.chunk = chunk; j.poly = poly;
        j.out = d_crc; j.out_stride = n_chunks;
        RET(launch_crc(*L.d, j, st));
    }
    BlocksFinishParams f;
    memset(&f, 0, sizeof f);
    f.desc = reinterpret_cast<const BlockDesc *>(db + desc_off);
    f.n_blocks = n; f.crcs = d_crc; f.whole = d_whole;
    f.cz_chunk = crc_of_zeros(chunk, P); f.x_chunk = x8n_modp(chunk, P); f.poly_bits = P;
    const int grid = static_cast<int>(std::min<size_t>(n, static_cast<size_t>(L.d->sm_count) * 8));
    crc_blocks_finish_kernel<<<grid, 128, 0, st>>>(f);
    g_launches++;
    CK(cudaGetLastError());
    if (packed) {  // hbuf was copied to the device before the kernels ran on this stream: it can take the results now
        uint32_t *h_crc = reinterpret_cast<uint32_t *>(L.s->hbuf), *h_whole = h_crc + n_chunks;
        if (chunk_crc_out && n_chunks) CK(cudaMemcpyAsync(h_crc, d_crc, 4 * n_chunks, cudaMemcpyDeviceToHost, st));
        if (whole_crc_out) CK(cudaMemcpyAsync(h_whole, d_whole, 4 * n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (whole_crc_out) memcpy(whole_crc_out, h_whole, 4 * n);
        if (chunk_crc_out)
            for (size_t b = 0; b < n; b++) {
                const size_t nc = div_up(lens[b], chunk);
                if (nc && chunk_crc_out[b]) memcpy(chunk_crc_out[b], h_crc + desc[b].first_chunk, 4 * nc);
            }
        return GPU_EC_OK;
    }
    if (whole_crc_out) CK(cudaMemcpyAsync(whole_crc_out, d_whole, 4 * n, cudaMemcpyDeviceToHost, st));
    if (chunk_crc_out)
        for (size_t b = 0; b < n; b++) {
            const size_t nc = div_up(lens[b], chunk);
            if (nc && chunk_crc_out[b]) CK(cudaMemcpyAsync(chunk_crc_out[b], d_crc + desc[b].first_chunk, 4 * nc, cudaMemcpyDeviceToHost, st));
        }
    CK(cudaStreamSynchronize(st));
    return GPU_EC_OK;
}

// ---- batches: stripe i -> device i mod G, several slots (streams) per device in flight --------
namespace {
struct BatchLanes {
    std::vector<Lease> leases;  // [dev][slot]
    int per_dev = 0;
    std::vector<Dev *> devs;
    int acquire(int slots_per_dev) {
        RET(ensure_init());
        std::vector<int> act;
        { Ctx &c = ctx(); std::lock_guard<std::mutex> lk(c.mu); act = c.active; }
        per_dev = slots_per_dev;
        leases.resize(act.size() * slots_per_dev);
        for (size_t a = 0; a < act.size(); a++) {
            Dev *d;
            RET(get_dev(act[a], &d));
            devs.push_back(d);
            RET(lease_n_on(*d, slots_per_dev, &leases[a * slots_per_dev]));
        }
        return GPU_EC_OK;
    }
    int sync_all() {
        for (auto &l : leases) if (l.s) { CK(cudaSetDevice(l.d->ordinal)); CK(cudaStreamSynchronize(l.s->stream)); }
        return GPU_EC_OK;
    }
};
}  // namespace

int gpu_ec_encode_batch(const uint8_t *data, size_t stripe_len, size_t n_stripes, int k, int m,
                        uint8_t *parity_out, uint32_t *shard_crc_out, uint32_t *full_crc_out) {
    GPU_EC_TRACE();
    if (k <= 0 || m <= 0) return GPU_EC_ERR_ZERO_SHARDS;
    if (stripe_len == 0) return GPU_EC_ERR_EMPTY_DATA;
    if (k + m > 256) return GPU_EC_ERR_TOO_MANY_SHARDS;
    if (!data || !parity_out) return GPU_EC_ERR_INVALID_ARG;
    if (n_stripes == 0) return GPU_EC_OK;
    const int n = k + m;
    const size_t S = gpu_ec_shard_len(stripe_len, k), pitch = round_up(S, 256);
    const size_t crc_off = static_cast<size_t>(n) * pitch;
    const size_t need = crc_off + round_up(4 * static_cast<size_t>(n) + 4, 256);
    const bool want_crc = shard_crc_out || full_crc_out;
    BatchLanes B;
    RET(B.acquire(3));
    for (auto &l : B.leases) { CK(cudaSetDevice(l.d->ordinal)); RET(slot_reserve(*l.s, need)); }
    FullCrcParams f;
    memset(&f, 0, sizeof f);
    if (full_crc_out) full_crc_constants(f, S, stripe_len, GPU_EC_POLY_IEEE);
    const size_t G = B.devs.size();
    const size_t full_rows = stripe_len / S, rem = stripe_len - full_rows * S;
    for (size_t i = 0; i < n_stripes; i++) {
        const size_t di = i % G;
        Lease &l = B.leases[di * B.per_dev + (i / G) % B.per_dev];
        CK(cudaSetDevice(l.d->ordinal));
        cudaStream_t st = l.s->stream;
        uint8_t *db = l.s->dbuf;
        const uint8_t *src = data + i * stripe_len;
        if (full_rows) CK(cudaMemcpy2DAsync(db, pitch, src, S, S, full_rows, cudaMemcpyHostToDevice, st));
        if (rem) CK(cudaMemcpyAsync(db + full_rows * pitch, src + full_rows * S, rem, cudaMemcpyHostToDevice, st));
        for (size_t r = full_rows; r < static_cast<size_t>(k); r++) {
            const size_t o = (r == full_rows) ? rem : 0;
            CK(cudaMemsetAsync(db + r * pitch + o, 0, S - o, st));
        }
        gpu_ec_stripes sv{db, pitch, static_cast<size_t>(n) * pitch, S, 1};
        RET(encode_dev_impl(*l.d, sv, k, m, st));
        CK(cudaMemcpy2DAsync(parity_out + i * static_cast<size_t>(m) * S, S, db + static_cast<size_t>(k) * pitch, pitch, S, m,
                             cudaMemcpyDeviceToHost, st));
        if (want_crc) {
            uint32_t *d_crc = reinterpret_cast<uint32_t *>(db + crc_off);
            CrcJob j;
            j.base = db; j.n_buf = n; j.buf_stride = pitch; j.len = S; j.chunk = 0; j.out = d_crc; j.out_stride = 1;
            RET(launch_crc(*l.d, j, st));
            if (shard_crc_out) CK(cudaMemcpyAsync(shard_crc_out + i * n, d_crc, 4 * static_cast<size_t>(n), cudaMemcpyDeviceToHost, st));
            if (full_crc_out) {
                f.shard_crc = d_crc; f.full_crc = d_crc + n; f.n_stripes = 1; f.n_shards = n;
                stripe_full_crc_kernel<<<1, 32, 0, st>>>(f);
                g_launches++;
                CK(cudaGetLastError());
                CK(cudaMemcpyAsync(full_crc_out + i, d_crc + n, 4, cudaMemcpyDeviceToHost, st));
            }
        }
    }
    return B.sync_all();
}

int gpu_ec_reconstruct_batch(uint8_t *shards, const uint8_t *present, size_t shard_len, size_t n_stripes, int k, int m) {
    GPU_EC_TRACE();
    if (!shards || !present) return GPU_EC_ERR_INVALID_ARG;
    DecodePlan plan;
    bool noop = false;
    RET(reconstruct_plan(k, m, present, plan, noop));
    if (noop || n_stripes == 0) return GPU_EC_OK;
    if (shard_len == 0) return GPU_EC_ERR_EMPTY_SHARD;
    const int n = k + m;
    const size_t S = shard_len, pitch = round_up(S, 256);
    const size_t host_stripe = static_cast<size_t>(n) * S, dev_stripe = static_cast<size_t>(n) * pitch;
    // groups of stripes per transfer: one 2-D copy per surviving / rebuilt slot per group
    const size_t grp = std::max<size_t>(1, std::min<size_t>(n_stripes, (48ull << 20) / host_stripe));
    BatchLanes B;
    RET(B.acquire(3));
    for (auto &l : B.leases) { CK(cudaSetDevice(l.d->ordinal)); RET(slot_reserve(*l.s, grp * dev_stripe)); }
    const size_t G = B.devs.size();
    size_t gi = 0;
    for (size_t i0 = 0; i0 < n_stripes; i0 += grp, gi++) {
        const size_t cnt = std::min(grp, n_stripes - i0);
        Lease &l = B.leases[(gi % G) * B.per_dev + (gi / G) % B.per_dev];
        CK(cudaSetDevice(l.d->ordinal));
        cudaStream_t st = l.s->stream;
        uint8_t *db = l.s->dbuf;
        uint8_t *hb = shards + i0 * host_stripe;
        for (int j = 0; j < k; j++) {
            const size_t idx = plan.survivors[j];
            CK(cudaMemcpy2DAsync(db + idx * pitch, dev_stripe, hb + idx * S, host_stripe, S, cnt, cudaMemcpyHostToDevice, st));
        }
        gpu_ec_stripes sv{db, pitch, dev_stripe, S, cnt};
        RET(launch_apply(*l.d, sv, plan.rows.data(), plan.survivors.data(), k, plan.missing.data(),
                         static_cast<int>(plan.missing.size()), st));
        for (int idx : plan.missing)
            CK(cudaMemcpy2DAsync(hb + static_cast<size_t>(idx) * S, host_stripe, db + static_cast<size_t>(idx) * pitch, dev_stripe, S, cnt,
                                 cudaMemcpyDeviceToHost, st));
    }
    return B.sync_all();
}


int gpu_ec_verify_chunks_batch(size_t n_blocks, const uint8_t *const *data, const size_t *lens, size_t chunk, int poly,
                               const uint8_t *const *meta_be, const size_t *meta_lens, int *status_out,
                               int64_t *first_bad_out) {
    GPU_EC_TRACE();
    if ((n_blocks && (!data || !lens || !meta_be || !meta_lens || !status_out)) || chunk == 0 || (poly != 0 && poly != 1))
        return GPU_EC_ERR_INVALID_ARG;
    if (n_blocks == 0) return GPU_EC_OK;
    BatchLanes B;
    RET(B.acquire(3));
    const size_t lanes = B.leases.size();
    // per lane: the verdict of the block it verified last (device int64 at the start of its buffer)
    std::vector<unsigned long long> verdict(n_blocks, ~0ull);
    std::vector<size_t> lane_last(lanes, SIZE_MAX);
    for (size_t i = 0; i < n_blocks; i++) {
        if (first_bad_out) first_bad_out[i] = -1;
        const size_t n_exp = meta_lens[i] / 4, n_act = div_up(lens[i], chunk);
        if (n_exp != n_act) { status_out[i] = GPU_EC_ERR_COUNT_MISMATCH; continue; }   // chunkserver.rs:277-279
        status_out[i] = GPU_EC_OK;
        if (n_act == 0) continue;
        if (!data[i] || !meta_be[i]) return GPU_EC_ERR_INVALID_ARG;
        Lease &l = B.leases[i % lanes];
        CK(cudaSetDevice(l.d->ordinal));
        cudaStream_t st = l.s->stream;
        const size_t data_off = 256, meta_off = data_off + round_up(lens[i], 256);
        RET(slot_reserve(*l.s, meta_off + round_up(4 * n_act, 256) + 256));
        uint8_t *db = l.s->dbuf;
        unsigned long long *d_bad = reinterpret_cast<unsigned long long *>(db);
        // the slot's previous verdict was copied out in stream order before this memset runs
        CK(cudaMemsetAsync(d_bad, 0xFF, 8, st));
        CK(cudaMemcpyAsync(db + data_off, data[i], lens[i], cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(db + meta_off, meta_be[i], 4 * n_act, cudaMemcpyHostToDevice, st));
        CrcJob j;
        j.base = db + data_off; j.n_buf = 1; j.buf_stride = round_up(lens[i], 256); j.len = lens[i]; j.chunk = chunk; j.poly = poly;
        j.expect_be = db + meta_off; j.first_bad = d_bad;
        RET(launch_crc(*l.d, j, st));
        CK(cudaMemcpyAsync(&verdict[i], d_bad, 8, cudaMemcpyDeviceToHost, st));
        lane_last[i % lanes] = i;
    }
    RET(B.sync_all());
    for (size_t i = 0; i < n_blocks; i++) {
        if (status_out[i] == GPU_EC_OK && verdict[i] != ~0ull) {
            status_out[i] = GPU_EC_ERR_CRC_MISMATCH;
            if (first_bad_out) first_bad_out[i] = static_cast<int64_t>(verdict[i]);
        }
    }
    return GPU_EC_OK;
}

// ---------------------------------------------------------------------------------------------
// gpu_ec_crc32_dev_noalloc — whole-buffer CRCs for callers that are not allowed to allocate.
//
// gpu_ec_crc32_dev hashes each buffer as many short segments and parks the per-segment CRCs in
// scratch that it takes from the stream-ordered allocator. Some callers cannot do that: code
// that runs where the process's device-memory budget is fixed up front (a storage daemon that
// pre-sizes its HBM pools and treats any later allocation as a fault), code on a latency path
// where an allocator call may block behind another thread or trim/grow a pool, and code that
// records its work for replay, where an allocation would be baked into the recording. For them
// the per-segment CRCs go into the fixed array below, which is part of the library image: each
// GPU that loads libgpu_ec.so holds these 4 MiB from load until the process exits.
//
// Consequences of sharing one array between all calls:
//   * the capacity is fixed: n_buf * ceil(len / segment) must not exceed NOALLOC_MAX_PARTIALS.
//     The segment length is raised (up to 64 KiB) to make large inputs fit; beyond that the call
//     is refused with GPU_EC_ERR_INVALID_ARG and gpu_ec_last_error() says why;
//   * the array still holds the previous call's values. Nothing in it has to start at zero:
//     every slot the fold kernel reads is written by the segment kernel earlier in the same call,
//     on the same stream, so stale contents are never observed and no clearing pass is issued;
//   * two calls whose kernels can be in flight at the same time on one GPU (different streams,
//     or different host threads) would overwrite each other's partials. Serialise them.
//
// The first call on each device does a one-time set-up with that device current: it resolves the
// device address of the array (the host-side name is only a handle, not something a kernel can
// dereference), uploads the CRC tables into their static symbol and raises the kernel's
// shared-memory limit. Make that first call before starting any stream capture.
// It never calls cudaMalloc / cudaFree / cudaMallocAsync / cudaFreeAsync or any other allocating
// routine, and it creates no streams or events.
namespace {
constexpr size_t NOALLOC_MAX_PARTIALS = size_t(1) << 20;  // 4 MiB of u32 per-segment CRCs
constexpr size_t NOALLOC_MAX_SEGMENT = 65536;
}  // namespace
static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_MAX_PARTIALS];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
struct NoAllocDev {
    std::mutex mu;
    bool ready = false;
    uint32_t *scratch = nullptr;  // device address of g_noalloc_scratch on this device
    Dev shim;                     // only ordinal + sm_count are used (launch_combine sizes its grid from it)
};
NoAllocDev g_noalloc_dev[NOALLOC_MAX_DEVICES];

int noalloc_setup(int ordinal, NoAllocDev **out) {
    if (ordinal < 0 || ordinal >= NOALLOC_MAX_DEVICES) return GPU_EC_ERR_INVALID_ARG;
    NoAllocDev &nd = g_noalloc_dev[ordinal];
    std::lock_guard<std::mutex> lk(nd.mu);
    CK(cudaSetDevice(ordinal));
    if (!nd.ready) {
        void *addr = nullptr;
        CK(cudaGetSymbolAddress(&addr, g_noalloc_scratch));  // per device: each GPU has its own copy
        if (!addr) { t_last_error = "cudaGetSymbolAddress returned a null scratch address"; return GPU_EC_ERR_CUDA; }
        int sms = 0;
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ordinal));
        static uint32_t tabs[2][4][256];
        static std::once_flag once;
        std::call_once(once, [] { for (int p = 0; p < 2; p++) crc_tables(kPolyReflected[p], tabs[p]); });
        CK(cudaMemcpyToSymbol(g_crc_tab, tabs, sizeof tabs));  // same bytes dev_prepare uploads; no allocation
        CK(cudaFuncSetAttribute(crc_segments_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CRC_SMEM_BYTES));
        // load all three kernels this entry point can launch now (module loading is lazy, and the first launch of a
        // kernel is where its code gets device memory): no later call, on whichever path, triggers a load
        cudaFuncAttributes fa;
        CK(cudaFuncGetAttributes(&fa, crc_segments_kernel<false>));
        CK(cudaFuncGetAttributes(&fa, crc_fold_kernel));
        CK(cudaFuncGetAttributes(&fa, crc_chunks_slow_kernel));
        nd.scratch = static_cast<uint32_t *>(addr);
        nd.shim.ordinal = ordinal;
        nd.shim.sm_count = sms;
        nd.ready = true;
    }
    *out = &nd;
    return GPU_EC_OK;
}
}  // namespace

int gpu_ec_crc32_dev_noalloc(const uint8_t *base, size_t n_buf, size_t buf_stride, size_t len, int poly,
                             uint32_t *out, void *stream) {
    GPU_EC_TRACE();
    DeviceGuard device_guard_;
    if (!base || !out || (poly != 0 && poly != 1)) return GPU_EC_ERR_INVALID_ARG;
    cudaPointerAttributes attr;
    CK(cudaPointerGetAttributes(&attr, base));
    if (attr.type != cudaMemoryTypeDevice && attr.type != cudaMemoryTypeManaged) {
        t_last_error = "gpu_ec_crc32_dev_noalloc: base is not device memory";
        return GPU_EC_ERR_INVALID_ARG;
    }
    NoAllocDev *nd = nullptr;
    RET(noalloc_setup(attr.device, &nd));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_buf == 0) return GPU_EC_OK;
    if (len == 0) {  // crc32fast of nothing is 0
        CK(cudaMemsetAsync(out, 0, 4 * n_buf, st));
        return GPU_EC_OK;
    }
    CrcJob j;
    j.base = base; j.n_buf = n_buf; j.buf_stride = buf_stride; j.len = len; j.chunk = 0; j.poly = poly;
    j.out = out; j.out_stride = 1;
    const int sms = nd->shim.sm_count;
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (buf_stride & 15)) {
        // unaligned buffers take the one-thread-per-buffer kernel, which needs no scratch at all
        CrcSlowParams sp;
        memset(&sp, 0, sizeof sp);
        sp.base = base; sp.buf_stride = buf_stride; sp.n_buf = n_buf; sp.len = len; sp.chunk = len; sp.chunks_per_buf = 1;
        sp.out = out; sp.out_stride = 1; sp.poly = poly;
        const int grid = static_cast<int>(std::min<size_t>(div_up(n_buf, 128), static_cast<size_t>(sms) * 16));
        crc_chunks_slow_kernel<<<grid, 128, 0, st>>>(sp);
        g_launches++;
        CK(cudaGetLastError());
        return GPU_EC_OK;
    }
    // shortest segment (most parallelism) whose per-segment CRCs, plus the few words the fold needs
    // beyond 1024 segments per buffer, fit the fixed array
    auto words_for = [&](size_t g) { const size_t segs = div_up(len, g); return n_buf * segs + fold_scratch_words(n_buf, segs); };
    // start from the segment length gpu_ec_crc32_dev would pick (tiles fill whole waves of the GPU), then grow it
    // until the partials fit
    size_t G = pick_whole_segment(nd->shim, n_buf, len);
    while (G <= NOALLOC_MAX_SEGMENT && words_for(G) > NOALLOC_MAX_PARTIALS) G = round_up(G * 2, 128);
    if (G > NOALLOC_MAX_SEGMENT) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "gpu_ec_crc32_dev_noalloc: %zu buffers x %zu bytes need more than the %zu per-segment slots of the "
                 "built-in scratch even at %zu-byte segments; split the batch or use gpu_ec_crc32_dev",
                 n_buf, len, NOALLOC_MAX_PARTIALS, NOALLOC_MAX_SEGMENT);
        t_last_error = msg;
        return GPU_EC_ERR_INVALID_ARG;
    }
    CrcParams p;
    memset(&p, 0, sizeof p);
    p.base = base; p.buf_stride = buf_stride; p.n_buf = n_buf; p.len = len;
    p.G = static_cast<unsigned int>(G);
    p.nfull = len / G;
    p.rag = static_cast<unsigned int>(len % G);
    p.segs_per_buf = p.nfull + (p.rag ? 1 : 0);
    p.tiles_per_buf = div_up(p.segs_per_buf, 32);
    p.poly = poly;
    p.out = nd->scratch;  // the device address, never the host-side symbol
    p.out_stride = p.segs_per_buf;
    const unsigned long long warps = n_buf * p.tiles_per_buf;
    const int grid = static_cast<int>(std::max<unsigned long long>(1, std::min<unsigned long long>(div_up(warps, CRC_WARPS), static_cast<unsigned long long>(sms))));
    crc_segments_kernel<false><<<grid, CRC_THREADS, CRC_SMEM_BYTES, st>>>(p);
    g_launches++;
    CK(cudaGetLastError());
    return launch_combine(nd->shim, nd->scratch, n_buf, p.segs_per_buf, G, len, p.segs_per_buf, j, st,
                          nd->scratch + n_buf * p.segs_per_buf);
}

}  // extern "C"
