// lut_common.cuh — the lane-replicated shared-memory lookup both kernel families are built on.
#pragma once
#include <cstdint>

namespace gpu_ec {

// Table lookup tab[entry*32 + lane]: byte extraction with PRMT (alu pipe), address with an integer
// multiply-add (fma pipe) so the two issue pipes share the per-byte work, then one LDS.
template <int OFF>
__device__ __forceinline__ uint32_t lut_lds(uint32_t lane_base, uint32_t entry) {
    uint32_t addr, v;
    asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(addr) : "r"(entry), "r"(lane_base));
    asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
    return v;
}
template <int B>
__device__ __forceinline__ uint32_t byte_of(uint32_t w) { return __byte_perm(w, 0, 0x4440 | B); }



// ---------------------------------------------------------------------------------------------
// Single-PRMT addressing (see crc_kernels.cuh for the full story): tables 0..3 live in two 64 KiB
// aligned blocks at shared-window addresses 0x10000 / 0x20000 with 256-byte rows (two tables
// interleaved per block), so a lookup address is {block base | data byte | lane*4} byte for byte:
// one PRMT, no multiply-add. Tables 4 and 5 (when present) keep the classic 128-byte-row form in
// the two free fragments of the window: [0x8000, 0x10000) and [0x30000, 0x38000).
// In a cluster kernel a CTA's shared window sits at (cluster rank << 24) in the shared address
// space, so every absolute address carries that window base `wb` (0 outside clusters): callers
// pass lane4w = wb | lane*4 and the PRMT copies the rank byte along.
constexpr uint32_t LUT_BLK_A = 0x10000u, LUT_BLK_B = 0x20000u, LUT_FRAG1 = 0x8000u, LUT_FRAG2 = 0x30000u;
constexpr int LUT_WINDOW_BYTES = 232448;  // the whole 227 KiB opt-in window (CRC kernel: its staging bypasses L1)
constexpr int LUT_FAST_RS_BYTES = 0x30000 - 0x400;  // RS kernels in the fast layout: window up to block B's end

// shared-window address of table j's entry b for `lane` in the fast layout
__device__ __forceinline__ uint32_t lut_fast_addr(int j, uint32_t b, uint32_t lane) {
    if (j < 4) return ((j < 2) ? LUT_BLK_A : LUT_BLK_B) + b * 256u + ((j & 1) ? 128u : 0u) + lane * 4u;
    return ((j == 4) ? LUT_FRAG1 : LUT_FRAG2) + b * 128u + lane * 4u;
}
// table J (compile time), byte I (compile time) of word w; lane4 = window base | lane*4
template <int J, int I>
__device__ __forceinline__ uint32_t lut_fast(uint32_t lane4, uint32_t w) {
    uint32_t v;
    if constexpr (J < 4) {
        const uint32_t lb = ((J < 2) ? LUT_BLK_A : LUT_BLK_B) | lane4;
#ifdef GPU_EC_LUT_BYTE3_FMA
        if constexpr (I == 3) {
            // top byte on the fma pipe (idle next to a saturated alu pipe): w >> 24 = umulhi(w, 2^8), then * 256 + base
            asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(__umulhi(w, 256u) * 256u + lb), "n"((J & 1) ? 128 : 0));
            return v;
        }
#endif
        asm("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(__byte_perm(w, lb, 0x7604 | (I << 4))), "n"((J & 1) ? 128 : 0));
    } else {
        const uint32_t base = ((J == 4) ? LUT_FRAG1 : LUT_FRAG2) + lane4;
        const uint32_t idx = (I == 3) ? (w >> 24) : __byte_perm(w, 0, 0x4440 | I);
        uint32_t addr;
        asm("mad.lo.u32 %0, %1, 128, %2;" : "=r"(addr) : "r"(idx), "r"(base));
        asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    }
    return v;
}

}  // namespace gpu_ec
