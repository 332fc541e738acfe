// pressure_coarsen_fast.cuh -- K6-fast: the fused pressure-level coarse-graining kernel rebuilt
// around the SM's issue rate (B200CS_FAST arithmetic, fp32 data, 8-wide windows, kord <= 3).
//
// Same dataflow and results (to the stated 1e-6 tolerance; masks, level search and coarse delp
// bit-identical) as pressure_coarsen.cuh, which stays the bit-exact statement.  Replaces
// regrid_to_area_weighted_pressure -> regrid_vertical/mappm -> _mask_weights ->
// weighted_block_average (external/vcm/vcm/cubedsphere/regridz.py:31-220,
// external/mappm/mappm/mappm.f90, external/vcm/vcm/cubedsphere/coarsen_restarts.py:485-534).
//
// What differs from the exact kernel, and why (profiles/r01_base_exact_k6_ncu_summary.txt:
// issue-bound at 6 % of the HBM roofline, 2.1e9 warp instructions, 92 KB smem -> 24 % occupancy):
//   * the column machine is PpmFast (ppm_fast.cuh): hoisted geometry, no divisions, one
//     advance() site, register prefetch of the next source layer;
//   * fine interface pressures are NOT staged in shared memory (80 KB/CTA): Phase B re-forms the
//     fp32 cumsum from delp, which it re-reads through L2 (the CTA's 80 KB of delp were streamed
//     by Phase A moments earlier); shared memory drops to ~12 KB and occupancy is set by
//     registers alone;
//   * the NF per-level partial sums are reduced with a transposing butterfly (4 shuffles for 4
//     fields instead of 12) and the weight sum is reduced only on levels where some lane of the
//     warp is masked (warp vote), otherwise it is the precomputed row sum;
//   * per-cell 1/(coarse layer thickness) comes from Phase A instead of a per-column running sum.
#pragma once
#include "common.cuh"
#include "ppm_fast.cuh"
#include "pressure_coarsen.cuh"

namespace b200cs {

constexpr int kFastKA = 8;  // Phase A levels per barrier
constexpr int kFastKC = 4;  // Phase B levels per barrier

constexpr int kRing = 4;  // cp.async ring depth (levels): kRing-1 levels in flight per thread

__device__ __forceinline__ void cp_async4(unsigned smem_addr, const void* gptr) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr), "l"(gptr) : "memory");
}

// Level source of one thread's column: NF field values + delp per level, streamed global ->
// shared with cp.async into a private ring (slot = level mod kRing) kRing-1 levels ahead of use.
// No registers are tied up by data in flight and nothing waits on a load in the iteration that
// issued it.
template <int NF>
struct FastSrc {
    const char* f[NF];
    const char* delp;
    unsigned off_q, off_d;    // byte offsets of the next level to ISSUE (this thread's column)
    unsigned step_q, step_d;  // level strides in bytes
    int ji, jc, km;           // next level to issue / level made current by the last fetch()
    unsigned ring;            // shared-space address of this thread's element of slot 0, value 0
    float toa;
    float qv[NF], dv;
    static constexpr unsigned kValStride = kThreads * 4, kSlotStride = (NF + 1) * kThreads * 4;
    __device__ __forceinline__ void issue() {
        const unsigned dst = ring + (unsigned)(ji & (kRing - 1)) * kSlotStride;
#pragma unroll
        for (int i = 0; i < NF; i++) cp_async4(dst + i * kValStride, f[i] + off_q);
        cp_async4(dst + NF * kValStride, delp + off_d);
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (ji < km - 1) { off_q += step_q; off_d += step_d; }  // beyond the surface: re-issue the last level
        ji++;
    }
    __device__ __forceinline__ void prologue() {
#pragma unroll
        for (int i = 0; i < kRing - 1; i++) issue();
    }
    // levels are requested in increasing order, one at a time (the arguments document which)
    __device__ __forceinline__ void fetch(int, int) {
        jc++;
        if (ji < jc + kRing) issue();  // level jc+kRing-1 replaces the slot of level jc-1; false only right after rewind()
        asm volatile("cp.async.wait_group %0;" ::"n"(kRing - 1) : "memory");
        const unsigned src = ring + (unsigned)(jc & (kRing - 1)) * kSlotStride;
#pragma unroll
        for (int i = 0; i < NF; i++) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(qv[i]) : "r"(src + i * kValStride));
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(dv) : "r"(src + NF * kValStride));
    }
    __device__ __forceinline__ void rewind() { jc--; }
    __device__ __forceinline__ float q(int i) const { return qv[i]; }
    __device__ __forceinline__ float pe(float prev) const { return prev + nan_to_zero(dv); }
    __device__ __forceinline__ float pe0() const { return toa; }
};

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

template <int NF, int FY>
__global__ void __launch_bounds__(kThreads, 2)
pressure_coarsen_fast_kernel(FieldPtrs<NF> fp, int64_t fs0, int64_t fsl, int64_t fsy,
                             const float* __restrict__ delp, int64_t ds0, int64_t dsl, int64_t dsy,
                             const float* __restrict__ w, int64_t ws0, int64_t wsy,
                             float* __restrict__ delp_coarse, int km, int64_t ny, int64_t nx, float toa,
                             int mask_mode, int iv, int seq_i) {
    constexpr int FX = 8;
    const bool seq = seq_i != 0;
    constexpr int CW = kTileX / FX;   // cells per tile row
    constexpr int CH = kTileY / FY;   // cell rows per tile
    constexpr int NCELL = CW * CH;
    const int tid = threadIdx.x, lane = tid & 31, ty = tid >> 5;
    const int64_t xg = (int64_t)blockIdx.x * kTileX + lane, yg = (int64_t)blockIdx.y * kTileY + ty;
    const int64_t b0 = blockIdx.z;
    const bool active = (xg < nx) && (yg < ny);
    // out-of-range threads shadow an in-range column with zero weight (keeps the warp converged)
    const int64_t x = xg < nx ? xg : nx - 1, y = yg < ny ? yg : ny - 1;
    const int64_t cxN = nx / FX, cyN = ny / FY;
    const int my_cell = (ty / FY) * CW + lane / FX;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* s_lvl = reinterpret_cast<float4*>(smem_raw);                       // [NCELL][km]: pb, 1/(pb-pt), mask threshold, coarse delp
    float* s_rows = reinterpret_cast<float*>(s_lvl + (size_t)NCELL * km);      // [2][KA][8][CW]
    float* s_part = s_rows + 2 * kFastKA * kTileY * CW;                        // [2][KC][8][CW][NF+1]
    float* s_den = s_part + 2 * kFastKC * kTileY * CW * (NF + 1);              // [NCELL]
    float* s_ring = s_den + ((NCELL + 3) & ~3);                                // [kRing][NF+1][256]

    // ---------------- Phase A: coarse pressure levels of the tile's cells -------------------
    const float wv = active ? __ldg(w + b0 * ws0 + y * wsy + x) : 0.f;
    const float wrow = reduce_row8(nan_to_zero(wv), seq, lane);  // weight sum of this window row
    if ((lane & (FX - 1)) == 0) s_part[ty * CW + lane / FX] = wrow;
    __syncthreads();
    if (tid < NCELL) {
        const int cyi = tid / CW, cxi = tid % CW;
        float d = 0.f;
        for (int r = 0; r < FY; r++) d = d + s_part[(cyi * FY + r) * CW + cxi];
        s_den[tid] = d;
    }
    __syncthreads();
    const float* dcol = delp + b0 * ds0 + y * dsy + x;
    float pacc = toa;
    for (int k0 = 0; k0 < km; k0 += kFastKA) {
        const int buf = (k0 / kFastKA) & 1;
        float d[kFastKA];
#pragma unroll
        for (int kk = 0; kk < kFastKA; kk++)
            d[kk] = (k0 + kk < km) ? __ldg(dcol + (int64_t)(k0 + kk) * dsl) : 0.f;
#pragma unroll
        for (int kk = 0; kk < kFastKA; kk++) {
            if (k0 + kk < km) {
                pacc = pacc + nan_to_zero(d[kk]);  // xarray cumsum == nancumsum
                const float rs = reduce_row8(nan_to_zero(d[kk] * wv), seq, lane);
                if ((lane & (FX - 1)) == 0) s_rows[((buf * kFastKA + kk) * kTileY + ty) * CW + lane / FX] = rs;
            }
        }
        __syncthreads();
        for (int t = tid; t < kFastKA * NCELL; t += kThreads) {
            const int kk = t / NCELL, c = t % NCELL, k = k0 + kk;
            if (k < km) {
                const int cyi = c / CW, cxi = c % CW;
                float num = 0.f;
                for (int r = 0; r < FY; r++) num = num + s_rows[((buf * kFastKA + kk) * kTileY + cyi * FY + r) * CW + cxi];
                const float dc = num / s_den[c];
                s_lvl[c * km + k].w = dc;
                const int64_t cyg = (int64_t)blockIdx.y * CH + cyi, cxg = (int64_t)blockIdx.x * CW + cxi;
                if (delp_coarse && cyg < cyN && cxg < cxN) delp_coarse[((b0 * km + k) * cyN + cyg) * cxN + cxg] = dc;
            }
        }
    }
    __syncthreads();
    if (tid < NCELL) {  // coarse interfaces: sequential cumsum (exact), thresholds, 1/thickness
        float pc = toa;
        for (int k = 0; k < km; k++) {
            const float dc = s_lvl[tid * km + k].w;
            const float pn = pc + nan_to_zero(dc);
            float thr = pn;  // PHALF: bottom interface of coarse layer k
            if (mask_mode == B200CS_MASK_PFULL) thr = dc / (log(pn) - log(pc));
            s_lvl[tid * km + k] = make_float4(pn, 1.0f / (pn - pc), thr, dc);
            pc = pn;
        }
    }
    __syncthreads();
    const float psurf = pacc;  // fine surface pressure, phalf_fine[km]

    // ---------------- Phase B: remap + masked block average ---------------------------------
    FastSrc<NF> src;
    {
        const int64_t fbase = b0 * fs0 + y * fsy + x;
#pragma unroll
        for (int i = 0; i < NF; i++) src.f[i] = (const char*)((const float*)fp.in[i] + fbase);
        src.delp = (const char*)dcol;
        src.off_q = 0; src.off_d = 0;
        src.step_q = (unsigned)(fsl * 4); src.step_d = (unsigned)(dsl * 4);
        src.ji = 0; src.jc = -1; src.km = km;
        src.ring = (unsigned)__cvta_generic_to_shared(s_ring + tid);
        src.toa = toa;
        src.prologue();
    }
    PpmFast<NF> st;
    st.init(src, km, iv, psurf);
    const float4* lvl = s_lvl + my_cell * km;
    float pt = toa;
    const bool writer = (lane & (FX - 1)) < NF;          // first NF lanes of each window row
    const int own = fast_owner<NF>(lane);
    const bool den_writer = (lane & (FX - 1)) == 0;
    const int item_lane = lane;                          // finalisation: 8 items per warp
    for (int k0 = 0; k0 < km; k0 += kFastKC) {
        const int buf = (k0 / kFastKC) & 1;
#pragma unroll 1
        for (int kk = 0; kk < kFastKC; kk++) {
            const int k = k0 + kk;
            if (k < km) {
                const float4 lv = lvl[k];
                float o[NF];
                st.remap(src, pt, lv.x, lv.y, o);
                pt = lv.x;
                const bool keep = (mask_mode == B200CS_MASK_NONE) || (lv.z < psurf);
                const float wk = keep ? wv : 0.f;
                float prod[NF];
#pragma unroll
                for (int i = 0; i < NF; i++) prod[i] = nan_to_zero(o[i] * wk);
                const float tot = reduce_fields8<NF>(prod, lane);
                float* dst = s_part + (size_t)((buf * kFastKC + kk) * kTileY + ty) * CW * (NF + 1) + (lane / FX) * (NF + 1);
                if (writer) dst[own] = tot;
                float rd = wrow;
                if (!__all_sync(0xffffffffu, keep)) rd = reduce_row8(nan_to_zero(wk), seq, lane);
                if (den_writer) dst[NF] = rd;
            }
        }
        __syncthreads();
        // finalise: kFastKC * NCELL * NF coarse values; 8 per warp so that no warp lags the barrier
        constexpr int ITEMS = kFastKC * NCELL * NF;
        for (int it = ty * 8 + (item_lane & 7) + (item_lane >> 3) * 64; it < ITEMS; it += 256) {
            const int i = it % NF, c = (it / NF) % NCELL, kk = it / (NF * NCELL), k = k0 + kk;
            const int cyi = c / CW, cxi = c % CW;
            const int64_t cyg = (int64_t)blockIdx.y * CH + cyi, cxg = (int64_t)blockIdx.x * CW + cxi;
            if (k < km && cyg < cyN && cxg < cxN) {
                float num = 0.f, den = 0.f;
#pragma unroll
                for (int r = 0; r < FY; r++) {
                    const float* srcp = s_part + (size_t)((buf * kFastKC + kk) * kTileY + cyi * FY + r) * CW * (NF + 1) + cxi * (NF + 1);
                    num = num + srcp[i];
                    den = den + srcp[NF];
                }
                ((float*)fp.out[i])[((b0 * km + k) * cyN + cyg) * cxN + cxg] = num / den;
            }
        }
    }
}

template <int NF, int FY>
size_t pressure_coarsen_fast_smem(int km) {
    constexpr int CW = kTileX / 8, NCELL = CW * (kTileY / FY);
    return (size_t)NCELL * km * sizeof(float4) +
           sizeof(float) * (2 * kFastKA * kTileY * CW + 2 * kFastKC * kTileY * CW * (NF + 1) + ((NCELL + 3) & ~3) +
                            kRing * (NF + 1) * kThreads);
}

template <int NF, int FY>
int launch_pc_fast(const PcArgs& a, int first) {
    auto kern = pressure_coarsen_fast_kernel<NF, FY>;
    const size_t smem = pressure_coarsen_fast_smem<NF, FY>(a.km);
    if (smem > 227 * 1024)
        return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen(fast): km=%d needs %zu B of shared memory (> 227 KB)", a.km, smem);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen(fast): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    FieldPtrs<NF> fp;
    for (int i = 0; i < NF; i++) { fp.in[i] = a.fields[first + i]; fp.out[i] = a.out[first + i]; }
    dim3 grid((unsigned)((a.nx + kTileX - 1) / kTileX), (unsigned)((a.ny + kTileY - 1) / kTileY), (unsigned)a.nb0);
    kern<<<grid, kThreads, smem, a.st>>>(fp, a.fstride[0], a.fstride[1], a.fstride[2], (const float*)a.delp,
                                         a.dstride[0], a.dstride[1], a.dstride[2], (const float*)a.w, a.wstride[0],
                                         a.wstride[1], first == 0 ? (float*)a.delp_coarse : nullptr, a.km, a.ny, a.nx,
                                         (float)a.toa, a.mask_mode, a.iv, a.seq);
    return check_launch("pressure_coarsen(fast)");
}

// true when the specialised kernel covers the call (everything else takes the generic kernel)
inline bool pc_fast_eligible(const PcArgs& a, int field_dtype, int p_dtype, int w_dtype) {
    const int64_t span_q = (int64_t)a.km * a.fstride[1] * 4, span_d = (int64_t)a.km * a.dstride[1] * 4;
    return field_dtype == B200CS_F32 && p_dtype == B200CS_F32 && w_dtype == B200CS_F32 && a.fx == 8 &&
           (a.fy == 8 || a.fy == 1) && a.kord <= 3 && a.fstride[1] >= 0 && a.dstride[1] >= 0 &&
           span_q < (int64_t)0xffffffff && span_d < (int64_t)0xffffffff;
}

template <int FY>
int launch_pc_fast_fields(const PcArgs& a) {
    int first = 0;
    while (first < a.nfields) {
        const int left = a.nfields - first;
        int rc;
        if (left >= 4) { rc = launch_pc_fast<4, FY>(a, first); first += 4; }
        else if (left >= 2) { rc = launch_pc_fast<2, FY>(a, first); first += 2; }
        else { rc = launch_pc_fast<1, FY>(a, first); first += 1; }
        if (rc != B200CS_OK) return rc;
    }
    return B200CS_OK;
}

}  // namespace b200cs
