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


}  // namespace gpu_ec
