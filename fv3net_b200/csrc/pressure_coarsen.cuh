// pressure_coarsen.cuh -- K6: fused pressure-level coarse-graining (one pass over the fine data).
//
// Replaces regrid_to_area_weighted_pressure / _regrid_given_delp / regrid_vertical / mappm /
// _mask_weights followed by weighted_block_average with the masked weights
// (external/vcm/vcm/cubedsphere/regridz.py:31-220, external/mappm/mappm/mappm.f90,
// external/vcm/vcm/cubedsphere/coarsen_restarts.py:485-534,934-959) -- a chain that in the
// reference materialises >= 2N+6 fine-size temporaries and re-derives the pressure geometry for
// each of the N variables.
//
// One CTA owns a tile of 8 rows x 32 columns of fine columns (whole coarse cells), one thread per
// column, all km levels:
//   Phase A  streams delp once: per-column interface pressures (sequential cumsum in delp's dtype)
//            -> shared memory as fp32 (what f2py hands to mappm); block average of delp in numpy
//            order -> coarse delp -> coarse interface pressures per cell (never upsampled: the
//            cell's threads read them from shared memory); the mask threshold per (cell, level).
//   Phase B  loops over the km coarse layers; every thread remaps its column with the forward-only
//            PPM machine (ppm_stream.cuh: 3-level register window per field, geometry shared by the
//            NF fields), applies the below-surface mask to its weight and the CTA reduces
//            sum(q*w), sum(w) per cell in numpy order (pairwise along x by warp shuffles, rows
//            sequentially through shared memory) and writes the coarse value.
// HBM traffic = delp + NF fields + weights read once, coarse outputs written once: the
// algorithmic bytes of SURVEY.md 8d (1608.7 B per fine column at km=79, NF=4, fp32).
#pragma once
#include "common.cuh"
#include "ppm_stream.cuh"

namespace b200cs {

constexpr int kTileX = 32, kTileY = 8, kThreads = kTileX * kTileY;
constexpr int kKC = 4;  // levels handled between two CTA barriers

// Peer-mapped copies of the coarse outputs (fused all-gather over NVLink): slot layout
// [nfields + 1][nb0, km, cy, cx], the coarse delp last.
struct PeerOut {
    int n;
    int nfields;
    int64_t field_elems;
    int* counters;  // [nb0 * gridDim.y], zeroed before the launch: CTAs finished per coarse row of tiles
    float* base[8];
};

// Wide windows (f = 16, 32, 64; b200cs_pressure_coarsen_wide): the coarse delp comes from a
// block-average pre-pass, every CTA leaves partial sums that a second kernel adds in a fixed order.
struct WideArgs {
    int f;
    const float* delp_c;  // [nb0, km, ny/f, nx/f]
    float* partial;       // [CTA][km][cells per CTA][NF + 1]
};

template <int NF>
struct FieldPtrs {
    const void* in[NF];
    void* out[NF];
};

// numpy order along a window row held by FX adjacent lanes: FX==8 -> pairwise tree
// ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)); FX<8 -> sequential from 0 (numpy pairwise_sum, n<8)
// seq: the window axis was NOT the contiguous axis of the reference's array (D-grid v lines are
// gathered columns): numpy then adds the FX elements one after the other onto 0.
template <int FX, typename T>
__device__ __forceinline__ T row_reduce(T v, int lane, bool seq = false) {
    if (FX == 8 && !seq) {
        v = v + __shfl_xor_sync(0xffffffffu, v, 1);
        v = v + __shfl_xor_sync(0xffffffffu, v, 2);
        v = v + __shfl_xor_sync(0xffffffffu, v, 4);
        return v;
    } else {
        const int base = lane & ~(FX - 1);
        T r = T(0);
#pragma unroll
        for (int c = 0; c < FX; c++) r = r + __shfl_sync(0xffffffffu, v, base + c);
        return r;
    }
}

template <int NF>
struct TileAcc {
    const float* s_pe1;  // [km+1][kThreads] + tid
    const void* f[NF];   // field base of this column (level 1)
    int64_t ls;          // level stride (elements)
    bool f64;
    __device__ __forceinline__ float pe(int j) const { return s_pe1[(j - 1) * kThreads]; }
    __device__ __forceinline__ float q(int i, int j) const {
        const int64_t o = (int64_t)(j - 1) * ls;
        return f64 ? (float)__ldg((const double*)f[i] + o) : __ldg((const float*)f[i] + o);
    }
};

template <typename A, typename B>
__device__ __forceinline__ bool less_promoted(A a, B b) {
    using T = typename Promote<A, B>::type;
    return (T)a < (T)b;
}

// TP: dtype of delp (fine pressures); TW: dtype of the 2-D weights (also accumulate / output
// dtype: promote(float32 remap output, TW) == TW).  FX: window width.  fy: FX or 1 (runtime).
template <int MODE, int NF, typename TP, typename TW, int FX>
__global__ void __launch_bounds__(kThreads, 2)
pressure_coarsen_kernel(FieldPtrs<NF> fp, int field_f64, int64_t fs0, int64_t fsl, int64_t fsy,
                        const TP* __restrict__ delp, int64_t ds0, int64_t dsl, int64_t dsy,
                        const TW* __restrict__ w, int64_t ws0, int64_t wsy,
                        typename Promote<TP, TW>::type* __restrict__ delp_coarse, int km, int64_t ny,
                        int64_t nx, int fy, double toa, int mask_mode, int iv, int kord, int seq_i) {
    const bool seq = seq_i != 0;
    using TC = typename Promote<TP, TW>::type;  // dtype of coarse delp / coarse pressures
    constexpr int CW = kTileX / FX;             // cells per tile row
    const int CH = kTileY / fy;                 // cell rows per tile
    const int ncell = CW * CH;
    const int tid = threadIdx.x, lane = tid & 31, ty = tid >> 5;
    const int64_t x = (int64_t)blockIdx.x * kTileX + lane, y = (int64_t)blockIdx.y * kTileY + ty;
    const int64_t b0 = blockIdx.z;
    const bool active = (x < nx) && (y < ny);
    const int64_t cxN = nx / FX, cyN = ny / fy;
    const int my_cell = (ty / fy) * CW + lane / FX;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    // layout: s_pe1 float[(km+1)*256] | s_pe2 float[ncell*(km+1)] | s_pm TC[ncell*km] |
    //         s_den TW[ncell] | s_rowsA TC[2*kKC*8*CW] | s_part TW[2*kKC*8*CW*(NF+1)]
    float* s_pe1 = reinterpret_cast<float*>(smem_raw);
    float* s_pe2 = s_pe1 + (size_t)(km + 1) * kThreads;
    size_t off = ((size_t)(km + 1) * kThreads + (size_t)ncell * (km + 1)) * sizeof(float);
    off = (off + 7) & ~size_t(7);
    TC* s_pm = reinterpret_cast<TC*>(smem_raw + off);
    off += (size_t)ncell * km * sizeof(TC);
    TC* s_rowsA = reinterpret_cast<TC*>(smem_raw + off);
    off += (size_t)2 * kKC * kTileY * CW * sizeof(TC);
    TW* s_part = reinterpret_cast<TW*>(smem_raw + off);
    off += (size_t)2 * kKC * kTileY * CW * (NF + 1) * sizeof(TW);
    TW* s_den = reinterpret_cast<TW*>(smem_raw + off);

    // ---------------- Phase A: pressures -------------------------------------------------
    const TW wv = active ? w[b0 * ws0 + y * wsy + x] : TW(0);
    {   // weight sum per cell (2-D weights: one value per cell), numpy order
        TW rs = row_reduce<FX>(nan_to_zero(wv), lane, seq);
        if ((lane & (FX - 1)) == 0) s_part[ty * CW + lane / FX] = rs;
        __syncthreads();
        if (tid < ncell) {
            const int cyi = tid / CW, cxi = tid % CW;
            TW d = TW(0);
            for (int r = 0; r < fy; r++) d = d + s_part[(cyi * fy + r) * CW + cxi];
            s_den[tid] = d;
        }
        __syncthreads();
    }
    const TP* dcol = delp + b0 * ds0 + y * dsy + x;
    TP pacc = (TP)toa;
    s_pe1[tid] = (float)pacc;
    for (int k0 = 0; k0 < km; k0 += kKC) {
        const int buf = (k0 / kKC) & 1;
        TP d[kKC];
#pragma unroll
        for (int kk = 0; kk < kKC; kk++)
            d[kk] = (active && k0 + kk < km) ? __ldg(dcol + (int64_t)(k0 + kk) * dsl) : TP(0);
#pragma unroll
        for (int kk = 0; kk < kKC; kk++) {
            if (k0 + kk < km) {
                pacc = pacc + nan_to_zero(d[kk]);  // xarray cumsum == nancumsum
                s_pe1[(k0 + kk + 1) * kThreads + tid] = (float)pacc;
                TC rs = row_reduce<FX>(nan_to_zero((TC)d[kk] * (TC)wv), lane, seq);
                if ((lane & (FX - 1)) == 0) s_rowsA[((buf * kKC + kk) * kTileY + ty) * CW + lane / FX] = rs;
            }
        }
        __syncthreads();
        for (int t = tid; t < kKC * ncell; t += kThreads) {
            const int kk = t / ncell, c = t % ncell, k = k0 + kk;
            if (k < km) {
                const int cyi = c / CW, cxi = c % CW;
                TC num = TC(0);
                for (int r = 0; r < fy; r++) num = num + s_rowsA[((buf * kKC + kk) * kTileY + cyi * fy + r) * CW + cxi];
                const TC dc = num / (TC)s_den[c];
                s_pm[c * km + k] = dc;
                const int64_t cyg = (int64_t)blockIdx.y * CH + cyi, cxg = (int64_t)blockIdx.x * CW + cxi;
                if (delp_coarse && cyg < cyN && cxg < cxN)
                    delp_coarse[((b0 * km + k) * cyN + cyg) * cxN + cxg] = dc;
            }
        }
        // s_rowsA is double-buffered: the next chunk writes the other half, and the barrier of the
        // next chunk orders these reads before the chunk after that overwrites this half.
    }
    __syncthreads();
    if (tid < ncell) {  // coarse interfaces: sequential cumsum in the coarse delp's dtype
        TC pc = (TC)toa;
        s_pe2[tid * (km + 1)] = (float)pc;
        for (int k = 0; k < km; k++) {
            const TC dc = s_pm[tid * km + k];
            const TC pn = pc + nan_to_zero(dc);
            s_pe2[tid * (km + 1) + k + 1] = (float)pn;
            TC thr = pn;  // PHALF: bottom interface of coarse layer k
            if (mask_mode == B200CS_MASK_PFULL) thr = dc / (log(pn) - log(pc));
            s_pm[tid * km + k] = thr;
            pc = pn;
        }
    }
    __syncthreads();
    const TP psurf = pacc;  // fine surface pressure, phalf_fine[km]

    // ---------------- Phase B: remap + masked block average -------------------------------
    TileAcc<NF> acc;
    acc.s_pe1 = s_pe1 + tid;
    acc.ls = fsl;
    acc.f64 = field_f64 != 0;
    const int64_t fbase = b0 * fs0 + y * fsy + x;
#pragma unroll
    for (int i = 0; i < NF; i++)
        acc.f[i] = acc.f64 ? (const void*)((const double*)fp.in[i] + fbase) : (const void*)((const float*)fp.in[i] + fbase);
    PpmStream<NF> st;
    if (active) st.init(acc, km, iv, kord);
    const float* pe2 = s_pe2 + my_cell * (km + 1);
    const TC* pm = s_pm + my_cell * km;
    constexpr int kPrefetch = 8;

    for (int k0 = 0; k0 < km; k0 += kKC) {
        const int buf = (k0 / kKC) & 1;
#pragma unroll 1
        for (int kk = 0; kk < kKC; kk++) {
            const int k = k0 + kk;
            if (k < km) {
                float o[NF];
                TW wk = TW(0);
#pragma unroll
                for (int i = 0; i < NF; i++) o[i] = 0.f;
                if (active) {
                    if (k + kPrefetch < km) {  // pull the level the window will need soon into L2
#pragma unroll
                        for (int i = 0; i < NF; i++) {
                            const char* p = (const char*)acc.f[i] + (int64_t)(k + kPrefetch) * fsl * (acc.f64 ? 8 : 4);
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
                        }
                    }
                    st.remap(acc, pe2[k], pe2[k + 1], o);
                    const bool keep = (mask_mode == B200CS_MASK_NONE) || less_promoted(pm[k], psurf);
                    wk = keep ? wv : TW(0);
                }
                TW* dst = s_part + (size_t)((buf * kKC + kk) * kTileY + ty) * CW * (NF + 1) + (lane / FX) * (NF + 1);
#pragma unroll
                for (int i = 0; i < NF; i++) {
                    TW rs = row_reduce<FX>(nan_to_zero((TW)o[i] * wk), lane, seq);
                    if ((lane & (FX - 1)) == 0) dst[i] = rs;
                }
                TW rd = row_reduce<FX>(nan_to_zero(wk), lane, seq);
                if ((lane & (FX - 1)) == 0) dst[NF] = rd;
            }
        }
        __syncthreads();
        for (int t = tid; t < kKC * ncell * NF; t += kThreads) {
            const int i = t % NF, c = (t / NF) % ncell, kk = t / (NF * ncell), k = k0 + kk;
            const int cyi = c / CW, cxi = c % CW;
            const int64_t cyg = (int64_t)blockIdx.y * CH + cyi, cxg = (int64_t)blockIdx.x * CW + cxi;
            if (k < km && cyg < cyN && cxg < cxN) {
                TW num = TW(0), den = TW(0);
                for (int r = 0; r < fy; r++) {
                    const TW* src = s_part + (size_t)((buf * kKC + kk) * kTileY + cyi * fy + r) * CW * (NF + 1) + cxi * (NF + 1);
                    num = num + src[i];
                    den = den + src[NF];
                }
                ((TW*)fp.out[i])[((b0 * km + k) * cyN + cyg) * cxN + cxg] = num / den;
            }
        }
    }
}

template <int NF, typename TP, typename TW, int FX>
size_t pressure_coarsen_smem(int km, int fy) {
    using TC = typename Promote<TP, TW>::type;
    constexpr int CW = kTileX / FX;
    const int ncell = CW * (kTileY / fy);
    size_t off = ((size_t)(km + 1) * kThreads + (size_t)ncell * (km + 1)) * sizeof(float);
    off = (off + 7) & ~size_t(7);
    off += (size_t)ncell * km * sizeof(TC);
    off += (size_t)2 * kKC * kTileY * CW * sizeof(TC);
    off += (size_t)2 * kKC * kTileY * CW * (NF + 1) * sizeof(TW);
    off += (size_t)ncell * sizeof(TW);
    return off;
}

struct PcArgs {
    int nfields;
    const void* const* fields;
    int field_dtype;
    const int64_t* fstride;
    const void* delp;
    const int64_t* dstride;
    const void* w;
    const int64_t* wstride;
    void* const* out;
    void* delp_coarse;
    int64_t nb0;
    int km;
    int64_t ny, nx;
    int fy, fx;
    double toa;
    int mask_mode, iv, kord;
    cudaStream_t st;
    int seq;  // B200CS_SEQ_WINDOW
    int npeer = 0;                 // b200cs_pressure_coarsen_gather: peer-mapped destinations
    void* const* peer = nullptr;   // [npeer] this rank's slot in each peer's gathered buffer
    int* counters = nullptr;       // workspace of the gather (see PeerOut)
};

template <int MODE, int NF, typename TP, typename TW, int FX>
int launch_pc(const PcArgs& a, int first) {
    using TC = typename Promote<TP, TW>::type;
    auto kern = pressure_coarsen_kernel<MODE, NF, TP, TW, FX>;
    const size_t smem = pressure_coarsen_smem<NF, TP, TW, FX>(a.km, a.fy);
    if (smem > 227 * 1024)
        return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen: km=%d, window %dx%d needs %zu B of shared memory (> 227 KB)",
                       a.km, a.fy, FX, smem);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    FieldPtrs<NF> fp;
    for (int i = 0; i < NF; i++) { fp.in[i] = a.fields[first + i]; fp.out[i] = a.out[first + i]; }
    dim3 grid((unsigned)((a.nx + kTileX - 1) / kTileX), (unsigned)((a.ny + kTileY - 1) / kTileY), (unsigned)a.nb0);
    // delp_coarse is written by the first launch of a multi-launch call only
    kern<<<grid, kThreads, smem, a.st>>>(fp, a.field_dtype == B200CS_F64, a.fstride[0], a.fstride[1], a.fstride[2],
                                         (const TP*)a.delp, a.dstride[0], a.dstride[1], a.dstride[2], (const TW*)a.w,
                                         a.wstride[0], a.wstride[1], first == 0 ? (TC*)a.delp_coarse : nullptr, a.km,
                                         a.ny, a.nx, a.fy, a.toa, a.mask_mode, a.iv, a.kord, a.seq);
    return check_launch("pressure_coarsen");
}

template <int MODE, typename TP, typename TW, int FX>
int launch_pc_fields(const PcArgs& a) {
    int first = 0;
    while (first < a.nfields) {
        const int left = a.nfields - first;
        int rc;
        if (left >= 4) { rc = launch_pc<MODE, 4, TP, TW, FX>(a, first); first += 4; }
        else if (left >= 2) { rc = launch_pc<MODE, 2, TP, TW, FX>(a, first); first += 2; }
        else { rc = launch_pc<MODE, 1, TP, TW, FX>(a, first); first += 1; }
        if (rc != B200CS_OK) return rc;
    }
    return B200CS_OK;
}

template <int MODE, int FX>
int launch_pc_types(const PcArgs& a, int p_dtype, int w_dtype) {
    if (p_dtype == B200CS_F32 && w_dtype == B200CS_F32) return launch_pc_fields<MODE, float, float, FX>(a);
    if (p_dtype == B200CS_F64 && w_dtype == B200CS_F32) return launch_pc_fields<MODE, double, float, FX>(a);
    if (p_dtype == B200CS_F64 && w_dtype == B200CS_F64) return launch_pc_fields<MODE, double, double, FX>(a);
    return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen: (delp f32, weights f64) is not instantiated");
}

}  // namespace b200cs
