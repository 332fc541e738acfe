// pressure_coarsen_fast.cuh -- helpers shared by the FAST fused pressure-coarsening kernels
// (pressure_coarsen_sweep.cuh, pressure_coarsen_tma.cuh): the cp.async primitive, the transposing
// butterfly over the 8 lanes of a window row, numpy-order row sums, and the eligibility test of
// the specialised kernels (fp32, 8-wide windows, kord <= 3).  The first FAST kernel
// (output-layer-major, 493 warp instructions per level; profiles/r01_fast_v*_ncu_summary.txt)
// lived here until round 2; the sweep kernel superseded it.
#pragma once
#include "common.cuh"
#include "ppm_fast.cuh"
#include "pressure_coarsen.cuh"

namespace b200cs {

constexpr int kFastKA = 8;  // Phase A levels per barrier


__device__ __forceinline__ void cp_async4(unsigned smem_addr, const void* gptr) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr), "l"(gptr) : "memory");
}

// sums v[0..NF) over the 8 lanes of a window row; afterwards lane l of the group holds the total
// of field fast_owner<NF>(l) in the return value.  Pairwise order (numpy's for 8 elements).
template <int NF>
__device__ __forceinline__ float reduce_fields8(const float (&v)[NF], int lane) {
    const unsigned m = 0xffffffffu;
    if (NF == 4) {
        const bool b0 = lane & 1, b1 = lane & 2;
        // xor 1: even lanes collect fields 0,1; odd lanes fields 2,3
        const float k0 = b0 ? v[2] : v[0], k1 = b0 ? v[3] : v[1];
        const float s0 = b0 ? v[0] : v[2], s1 = b0 ? v[1] : v[3];
        const float r0 = k0 + __shfl_xor_sync(m, s0, 1);
        const float r1 = k1 + __shfl_xor_sync(m, s1, 1);
        // xor 2: lanes with bit1 clear keep the first of their pair
        const float k = b1 ? r1 : r0, s = b1 ? r0 : r1;
        const float t = k + __shfl_xor_sync(m, s, 2);
        return t + __shfl_xor_sync(m, t, 4);
    } else if (NF == 2) {
        const bool b0 = lane & 1;
        const float k = b0 ? v[1] : v[0], s = b0 ? v[0] : v[1];
        float t = k + __shfl_xor_sync(m, s, 1);
        t = t + __shfl_xor_sync(m, t, 2);
        return t + __shfl_xor_sync(m, t, 4);
    } else {
        float t = v[0];
        t = t + __shfl_xor_sync(m, t, 1);
        t = t + __shfl_xor_sync(m, t, 2);
        return t + __shfl_xor_sync(m, t, 4);
    }
}
// seq: numpy's order for a window whose elements were strided in the reference's array
// (sequential onto 0) -- keeps the coarse delp, hence the mask, bit-identical for D-grid v lines
__device__ __forceinline__ float reduce_row8(float t, bool seq = false, int lane = 0) {
    const unsigned m = 0xffffffffu;
    if (seq) {
        const int base = lane & ~7;
        float r = 0.f;
#pragma unroll
        for (int c = 0; c < 8; c++) r = r + __shfl_sync(m, t, base + c);
        return r;
    }
    t = t + __shfl_xor_sync(m, t, 1);
    t = t + __shfl_xor_sync(m, t, 2);
    return t + __shfl_xor_sync(m, t, 4);
}
template <int NF>
__device__ __forceinline__ int fast_owner(int lane) {
    if (NF == 4) return ((lane & 1) << 1) | ((lane >> 1) & 1);
    if (NF == 2) return lane & 1;
    return 0;
}

// true when the specialised kernel covers the call (everything else takes the generic kernel)
inline bool pc_fast_eligible(const PcArgs& a, int field_dtype, int p_dtype, int w_dtype) {
    const int64_t span_q = (int64_t)a.km * a.fstride[1] * 4, span_d = (int64_t)a.km * a.dstride[1] * 4;
    return field_dtype == B200CS_F32 && p_dtype == B200CS_F32 && w_dtype == B200CS_F32 && a.fx == 8 &&
           (a.fy == 8 || a.fy == 1) && a.kord <= 3 && a.fstride[1] >= 0 && a.dstride[1] >= 0 &&
           span_q < (int64_t)0xffffffff && span_d < (int64_t)0xffffffff;
}

}  // namespace b200cs
