// block_ops.cu -- horizontal block kernels of libb200cs (sm_100a):
//   K1/K2  weighted block average / edge-weighted block average
//   K3     generic block reduce (sum/mean/min/max/median/mode, NaN-skipping, padding)
//   K4     block upsample
//   plus the small layout helpers used by the D-grid pressure path.
// All are HBM-bound streaming kernels: 128-bit coalesced loads, every fine byte read once,
// numerator / denominator / divide fused in one pass.  Compiled with -fmad=false so that
// x*w is rounded before it is summed, exactly like numpy's materialised `obj * weights`.
#include <stdlib.h>

#include "common.cuh"

namespace b200cs {

// ---------------------------------------------------------------------------------------
// K1 generic: one thread per (output cell, level).  Any window, any dtype mix.
// ---------------------------------------------------------------------------------------
template <typename TO, typename TW, typename TW2, bool HAS_W2>
__global__ void __launch_bounds__(256)
wba_generic_kernel(const TO* __restrict__ obj, int64_t so0, int64_t so1, int64_t soy,
                   const TW* __restrict__ w, int64_t sw0, int64_t sw1, int64_t swy,
                   const TW2* __restrict__ w2, int64_t sv0, int64_t sv1, int64_t svy,
                   typename Promote<TO, typename Promote<TW, TW2>::type>::type* __restrict__ out,
                   int64_t nb0, int64_t nb1, int64_t cy, int64_t cx, int fy, int fx, int ystep,
                   int xstep) {
    using TD = typename Promote<TW, TW2>::type;  // weight dtype (mass = delp*area promotes)
    using TN = typename Promote<TO, TD>::type;   // product / output dtype
    const int64_t total = nb0 * nb1 * cy * cx;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = idx % cx;
        const int64_t j = (idx / cx) % cy;
        const int64_t b1 = (idx / (cx * cy)) % nb1;
        const int64_t b0 = idx / (cx * cy * nb1);
        TN acc = TN(0);
        TD dacc = TD(0);
        for (int r = 0; r < fy; r++) {
            const int64_t y = j * ystep + r, x = i * xstep;
            const TO* xo = obj + b0 * so0 + b1 * so1 + y * soy + x;
            const TW* xw = w + b0 * sw0 + b1 * sw1 + y * swy + x;
            const TW2* xv = HAS_W2 ? (w2 + b0 * sv0 + b1 * sv1 + y * svy + x) : nullptr;
            auto wt = [&](int c) -> TD {
                if (HAS_W2) return (TD)xv[c] * (TD)xw[c];
                return (TD)xw[c];
            };
            acc += rowsum_numpy<TN>(fx, [&](int c) { return nan_to_zero((TN)xo[c] * (TN)wt(c)); });
            dacc += rowsum_numpy<TD>(fx, [&](int c) { return nan_to_zero(wt(c)); });
        }
        out[idx] = acc / (TN)dacc;
    }
}

// ---------------------------------------------------------------------------------------
// K1 fast path: 8x8 windows, fp32 everywhere, 16-byte aligned rows.  One thread owns one
// coarse cell for LV consecutive levels, so a 2-D weight row (two float4) is loaded once and
// reused LV times, and the weight sum is formed once.  Row sums use the 8-wide pairwise tree
// ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)), rows are added in order: numpy's order exactly.
// W3D: weights carry a level stride (masked 3-D weights) -> per-level weight rows and sums.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float tree8(const float4& a, const float4& b) {
    return ((a.x + a.y) + (a.z + a.w)) + ((b.x + b.y) + (b.z + b.w));
}
__device__ __forceinline__ float4 mul_nz(const float4& x, const float4& w) {
    return make_float4(nan_to_zero(x.x * w.x), nan_to_zero(x.y * w.y), nan_to_zero(x.z * w.z),
                       nan_to_zero(x.w * w.w));
}
__device__ __forceinline__ float4 nz4(const float4& w) {
    return make_float4(nan_to_zero(w.x), nan_to_zero(w.y), nan_to_zero(w.z), nan_to_zero(w.w));
}
__device__ __forceinline__ float4 ldg4(const float* p) {
    return __ldg(reinterpret_cast<const float4*>(p));
}

// K1, wide windows (F = 16, 32, 64; fp32): nansum(obj * w) / nansum(w) over F x F blocks in
// numpy's order (rowsum_numpy along x, rows added one after the other), streaming.  numpy sums a
// row of F <= 128 elements with eight interleaved accumulators r[j] += a[8m + j] and the tree
// ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)).  TWO lanes share one (window row, cell): lane q keeps
// accumulators 4q..4q+3, i.e. every second float4 of the row, so the chains need no shuffles and
// the pair reads 32 contiguous bytes per load (a full sector; one lane per row touched 32
// different lines per load instruction and was bound by L1 wavefronts at 30 % of HBM peak).
// CTA tile = F rows x 128 columns.  C3072 -> C48 is factor 64 (workflows' coarsen_surface_c3072,
// BASELINE configs[4]).
template <int F>
__global__ void __launch_bounds__(256)
wba_wide_f32_kernel(const float* __restrict__ obj, int64_t so0, int64_t so1, int64_t soy,
                    const float* __restrict__ w, int64_t sw0, int64_t sw1, int64_t swy, float* __restrict__ out,
                    int64_t nb1, int64_t cyN, int64_t cxN) {
    constexpr int NCX = 128 / F;  // cells per CTA
    __shared__ float s_num[NCX][F + 1], s_den[NCX][F + 1];
    const int tid = threadIdx.x, q = tid & 1, pr = tid >> 1, r = pr / NCX, c = pr % NCX;
    const int64_t b0 = blockIdx.z / nb1, b1 = blockIdx.z % nb1;
    const int64_t cyg = blockIdx.y, cxg = (int64_t)blockIdx.x * NCX + c;
    float num = 0.f, den = 0.f;
    if (cxg < cxN) {
        const int64_t y = cyg * F + r, x0 = cxg * F;
        const float4* dp = reinterpret_cast<const float4*>(obj + b0 * so0 + b1 * so1 + y * soy + x0) + q;
        const float4* wp = reinterpret_cast<const float4*>(w + b0 * sw0 + b1 * sw1 + y * swy + x0) + q;
        float rn[4], rd[4];
#pragma unroll
        for (int m = 0; m < F / 8; m++) {
            const float4 d0 = __ldg(dp + 2 * m), w0 = __ldg(wp + 2 * m);
            const float dv[4] = {d0.x, d0.y, d0.z, d0.w};
            const float wv[4] = {w0.x, w0.y, w0.z, w0.w};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float p = nan_to_zero(__fmul_rn(dv[j], wv[j]));
                const float z = nan_to_zero(wv[j]);
                rn[j] = m == 0 ? p : __fadd_rn(rn[j], p);
                rd[j] = m == 0 ? z : __fadd_rn(rd[j], z);
            }
        }
        num = __fadd_rn(__fadd_rn(rn[0], rn[1]), __fadd_rn(rn[2], rn[3]));
        den = __fadd_rn(__fadd_rn(rd[0], rd[1]), __fadd_rn(rd[2], rd[3]));
    }
    // (r0..r3) + (r4..r7): the odd lane's half is added onto the even lane's
    const float num1 = __shfl_down_sync(0xffffffffu, num, 1), den1 = __shfl_down_sync(0xffffffffu, den, 1);
    if (q == 0) {
        s_num[c][r] = __fadd_rn(num, num1);
        s_den[c][r] = __fadd_rn(den, den1);
    }
    __syncthreads();
    if (tid < NCX) {
        const int64_t cx = (int64_t)blockIdx.x * NCX + tid;
        if (cx < cxN) {
            float n = 0.f, d = 0.f;
            for (int rr = 0; rr < F; rr++) { n = __fadd_rn(n, s_num[tid][rr]); d = __fadd_rn(d, s_den[tid][rr]); }
            out[((b0 * nb1 + b1) * cyN + cyg) * cxN + cx] = __fdiv_rn(n, d);
        }
    }
}

template <int F>
static int launch_wba_wide(const void* obj, const int64_t* so, const void* w, const int64_t* sw, void* out, int64_t nb0,
                           int64_t nb1, int64_t cy, int64_t cx, cudaStream_t st) {
    dim3 grid((unsigned)((cx + 128 / F - 1) / (128 / F)), (unsigned)cy, (unsigned)(nb0 * nb1));
    wba_wide_f32_kernel<F><<<grid, 256, 0, st>>>((const float*)obj, so[0], so[1], so[2], (const float*)w, sw[0], sw[1],
                                                sw[2], (float*)out, nb1, cy, cx);
    return check_launch("weighted_block_average(wide f32)");
}

template <int LV, bool W3D, bool HAS_W2>
__global__ void __launch_bounds__(256)
wba8_f32_kernel(const float* __restrict__ obj, int64_t so0, int64_t so1, int64_t soy,
                const float* __restrict__ w, int64_t sw0, int64_t sw1, int64_t swy,
                const float* __restrict__ w2, int64_t sv0, int64_t sv1, int64_t svy,
                float* __restrict__ out, int64_t nb0, int64_t nb1, int64_t cy, int64_t cx) {
    const int64_t ncell = nb0 * cy * cx;
    const int64_t cell = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (cell >= ncell) return;
    const int64_t i = cell % cx, j = (cell / cx) % cy, b0 = cell / (cx * cy);
    const int64_t l0 = (int64_t)blockIdx.y * LV;
    const float* xo = obj + b0 * so0 + l0 * so1 + (j * 8) * soy + i * 8;
    const float* xw = w + b0 * sw0 + (W3D ? l0 * sw1 : 0) + (j * 8) * swy + i * 8;
    const float* xv = HAS_W2 ? (w2 + b0 * sv0 + l0 * sv1 + (j * 8) * svy + i * 8) : nullptr;
    float acc[LV], den[LV];
#pragma unroll
    for (int l = 0; l < LV; l++) { acc[l] = 0.f; den[l] = 0.f; }
#pragma unroll 2
    for (int r = 0; r < 8; r++) {
        float4 wa, wb;
        if (!W3D && !HAS_W2) {
            wa = ldg4(xw + r * swy);
            wb = ldg4(xw + r * swy + 4);
            den[0] += tree8(nz4(wa), nz4(wb));
        }
#pragma unroll
        for (int l = 0; l < LV; l++) {
            if (l0 + l < nb1) {
                if (W3D || HAS_W2) {
                    wa = ldg4(xw + (W3D ? l * sw1 : 0) + r * swy);
                    wb = ldg4(xw + (W3D ? l * sw1 : 0) + r * swy + 4);
                    if (HAS_W2) {  // mass = delp * area, rounded before use
                        float4 va = ldg4(xv + l * sv1 + r * svy), vb = ldg4(xv + l * sv1 + r * svy + 4);
                        wa = make_float4(va.x * wa.x, va.y * wa.y, va.z * wa.z, va.w * wa.w);
                        wb = make_float4(vb.x * wb.x, vb.y * wb.y, vb.z * wb.z, vb.w * wb.w);
                    }
                    den[l] += tree8(nz4(wa), nz4(wb));
                }
                float4 xa = ldg4(xo + l * so1 + r * soy), xb = ldg4(xo + l * so1 + r * soy + 4);
                acc[l] += tree8(mul_nz(xa, wa), mul_nz(xb, wb));
            }
        }
    }
#pragma unroll
    for (int l = 0; l < LV; l++)
        if (l0 + l < nb1)
            out[((b0 * nb1 + l0 + l) * cy + j) * cx + i] = acc[l] / ((W3D || HAS_W2) ? den[l] : den[0]);
}

// 8 x 8 windows, fp32, 2-D weights shared by all levels (the area): a CTA keeps the 64 weights of its
// 128 cells in shared memory (32 KB, conflict-free 16-byte accesses) while it walks consecutive levels,
// so the weights are read once per run of levels, not once per level (with one level per thread the
// re-read weights doubled the L2 -> SM traffic: 77 % of the HBM peak at C384 against 94 % for the
// unweighted sum), and a thread's registers hold the 16 loads of a level in flight at once.  The
// (block of 128 cells, level) items, level fastest, are dealt out in equal contiguous runs to a grid of
// one wave.  Same operations in the same order as wba8_f32_kernel<1, false, false>: bit-identical.
__global__ void __launch_bounds__(128, 5)
wba8_f32_w2d_kernel(const float* __restrict__ obj, int64_t so0, int64_t so1, int64_t soy,
                    const float* __restrict__ w, int64_t sw0, int64_t swy,
                    float* __restrict__ out, int64_t nb0, int nb1, int64_t cy, int64_t cx, int64_t nitem) {
    __shared__ float4 s_w[16 * 128];
    const int64_t ncell = nb0 * cy * cx;
    const int64_t it0 = nitem * blockIdx.x / gridDim.x, it1 = nitem * (blockIdx.x + 1) / gridDim.x;
    int64_t cb = it0 / nb1;
    int l = (int)(it0 - cb * nb1);
    const int tid = threadIdx.x;
    float den = 0.f;
    const float* xo = nullptr;
    float* po = nullptr;
    bool live = false, fresh = true;
    for (int64_t it = it0; it < it1; it++) {
        if (fresh) {  // (uniform per CTA) a new block of cells: its weights into shared memory
            fresh = false;
            const int64_t cell = cb * 128 + tid;
            live = cell < ncell;
            if (live) {
                const int64_t i = cell % cx, j = (cell / cx) % cy, b0 = cell / (cx * cy);
                const float* xw = w + b0 * sw0 + (j * 8) * swy + i * 8;
                float4 wa[8], wb[8];
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    wa[r] = ldg4(xw + r * swy);
                    wb[r] = ldg4(xw + r * swy + 4);
                }
                den = 0.f;
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    den += tree8(nz4(wa[r]), nz4(wb[r]));
                    s_w[(2 * r) * 128 + tid] = wa[r];   // read back by the same thread only: no barrier
                    s_w[(2 * r + 1) * 128 + tid] = wb[r];
                }
                xo = obj + b0 * so0 + (int64_t)l * so1 + (j * 8) * soy + i * 8;
                po = out + ((b0 * nb1 + l) * cy + j) * cx + i;
            }
        }
        if (live) {
            float4 xa[8], xb[8];
#pragma unroll
            for (int r = 0; r < 8; r++) {
                xa[r] = ldg4(xo + r * soy);
                xb[r] = ldg4(xo + r * soy + 4);
            }
            float acc = 0.f;
#pragma unroll
            for (int r = 0; r < 8; r++) acc += tree8(mul_nz(xa[r], s_w[(2 * r) * 128 + tid]), mul_nz(xb[r], s_w[(2 * r + 1) * 128 + tid]));
            *po = acc / den;
            xo += so1;
            po += cy * cx;
        }
        if (++l == nb1) { l = 0; cb++; fresh = true; }
    }
}

template <typename TO, typename TW, typename TW2, bool HAS_W2>
static int launch_wba_generic(const void* obj, const int64_t* so, const void* w, const int64_t* sw,
                              const void* w2, const int64_t* sv, void* out, int64_t nb0, int64_t nb1,
                              int64_t cy, int64_t cx, int fy, int fx, int ystep, int xstep,
                              cudaStream_t st) {
    using TD = typename Promote<TW, TW2>::type;
    using TN = typename Promote<TO, TD>::type;
    const int64_t total = nb0 * nb1 * cy * cx;
    if (total == 0) return B200CS_OK;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148LL * 64) blocks = 148LL * 64;
    wba_generic_kernel<TO, TW, TW2, HAS_W2><<<(unsigned)blocks, 256, 0, st>>>(
        (const TO*)obj, so[0], so[1], so[2], (const TW*)w, sw[0], sw[1], sw[2], (const TW2*)w2,
        HAS_W2 ? sv[0] : 0, HAS_W2 ? sv[1] : 0, HAS_W2 ? sv[2] : 0, (TN*)out, nb0, nb1, cy, cx, fy, fx,
        ystep, xstep);
    return check_launch("weighted_block_average");
}

template <typename TO, typename TW>
static int dispatch_w2(const void* obj, const int64_t* so, const void* w, const int64_t* sw,
                       const void* w2, int w2_dtype, const int64_t* sv, void* out, int64_t nb0,
                       int64_t nb1, int64_t cy, int64_t cx, int fy, int fx, int ystep, int xstep,
                       cudaStream_t st) {
    if (!w2)
        return launch_wba_generic<TO, TW, TW, false>(obj, so, w, sw, nullptr, sv, out, nb0, nb1, cy, cx,
                                                     fy, fx, ystep, xstep, st);
    if (w2_dtype == B200CS_F32)
        return launch_wba_generic<TO, TW, float, true>(obj, so, w, sw, w2, sv, out, nb0, nb1, cy, cx, fy,
                                                       fx, ystep, xstep, st);
    return launch_wba_generic<TO, TW, double, true>(obj, so, w, sw, w2, sv, out, nb0, nb1, cy, cx, fy,
                                                    fx, ystep, xstep, st);
}

// VEC consecutive elements with the alignment of one 16-byte (32-byte for 4 doubles) access
template <typename T, int N>
struct alignas(sizeof(T) * N) Pack { T v[N]; };

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---------------------------------------------------------------------------------------
// K3 generic block reduce: thread per output cell; window gathered into local memory for
// median / mode (window <= 256 elements), streamed otherwise.
// ---------------------------------------------------------------------------------------
constexpr int kMaxSortWindow = 256;

template <typename T>
__global__ void __launch_bounds__(128)
block_reduce_kernel(const T* __restrict__ obj, int64_t s0, int64_t sy, T* __restrict__ out, int64_t nb,
                    int64_t ny, int64_t nx, int64_t cy, int64_t cx, int fy, int fx, int ystep,
                    int xstep, int op, int skipna, T cval) {
    const int64_t total = nb * cy * cx;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = idx % cx, j = (idx / cx) % cy, b = idx / (cx * cy);
        const T* base = obj + b * s0;
        auto at = [&](int r, int c) -> T {
            const int64_t y = j * ystep + r, x = i * (int64_t)xstep + c;
            return (y < ny && x < nx) ? base[y * sy + x] : cval;  // trailing-edge padding
        };
        T res;
        if (op == B200CS_OP_SUM || op == B200CS_OP_MEAN) {
            // numpy order: pairwise along the window row, rows added sequentially
            T acc = T(0);
            int64_t cnt = 0;
            for (int r = 0; r < fy; r++) {
                acc += rowsum_numpy<T>(fx, [&](int c) {
                    T v = at(r, c);
                    if (skipna) { if (v == v) cnt++; return nan_to_zero(v); }
                    return v;
                });
            }
            if (op == B200CS_OP_SUM) res = acc;
            else res = skipna ? acc / (T)cnt : acc / (T)((int64_t)fy * fx);
        } else if (op == B200CS_OP_MIN || op == B200CS_OP_MAX) {
            bool any = false, sawnan = false;
            T m = T(0);
            for (int r = 0; r < fy; r++)
                for (int c = 0; c < fx; c++) {
                    T v = at(r, c);
                    if (v != v) { sawnan = true; continue; }
                    if (!any) { m = v; any = true; }
                    else if (op == B200CS_OP_MIN ? (v < m) : (v > m)) m = v;
                }
            res = ((sawnan && !skipna) || !any) ? (T)NAN : m;
        } else {
            // median / mode: insertion sort of the (non-NaN) window in local memory
            T buf[kMaxSortWindow];
            int n = 0;
            bool sawnan = false;
            for (int r = 0; r < fy; r++)
                for (int c = 0; c < fx; c++) {
                    T v = at(r, c);
                    if (v != v) { sawnan = true; continue; }
                    int p = n++;
                    while (p > 0 && buf[p - 1] > v) { buf[p] = buf[p - 1]; p--; }
                    buf[p] = v;
                }
            if ((sawnan && !skipna) || n == 0) {
                res = (T)NAN;
            } else if (op == B200CS_OP_MEDIAN) {
                // np.median: mean of the two middle elements for even n
                res = (n & 1) ? buf[n / 2] : (buf[n / 2 - 1] + buf[n / 2]) / T(2);
            } else {  // mode: most frequent, ties -> smallest (scipy.stats.mode)
                T best = buf[0];
                int bestc = 0, run = 0;
                for (int p = 0; p < n; p++) {
                    run = (p > 0 && buf[p] == buf[p - 1]) ? run + 1 : 1;
                    if (run > bestc) { bestc = run; best = buf[p]; }
                }
                res = best;
            }
        }
        out[idx] = res;
    }
}

// ---------------------------------------------------------------------------------------
// K3 median / mode of square power-of-two windows (2x2, 4x4, 8x8; non-overlapping, no padding):
// one thread per window, the window lives in REGISTERS and is sorted by Batcher's odd-even merge
// network, fully unrolled (543 compare-exchanges for 64 values, each one min + one max).  The
// insertion sort above keeps its window in local memory and moves ~n^2/4 elements through L1;
// this is ~50x fewer instructions and none of them touches memory.  NaNs sort last (they enter
// the network as +inf and are counted), so the first n sorted values are the non-NaN ones.
// ---------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void cmp_exchange(T& a, T& b) {
    const T lo = a < b ? a : b, hi = a < b ? b : a;  // no NaNs inside the network
    a = lo; b = hi;
}
template <>
__device__ __forceinline__ void cmp_exchange<float>(float& a, float& b) {
    const float lo = fminf(a, b), hi = fmaxf(a, b);
    a = lo; b = hi;
}

// Batcher's odd-even merge sort as recursive templates: every index is a template parameter, so
// the network unrolls into straight-line code on registers (a loop nest with data-independent
// but non-affine bounds was left partly rolled by the compiler and put the window in local memory).
template <typename T, int N, int LO, int LEN, int R>
struct OddEvenMerge {  // merges the two sorted halves of a[LO], a[LO+R], ..., LEN elements spanned
    __device__ __forceinline__ static void run(T (&a)[N]) {
        constexpr int M = R * 2;
        if constexpr (M < LEN) {
            OddEvenMerge<T, N, LO, LEN, M>::run(a);
            OddEvenMerge<T, N, LO + R, LEN, M>::run(a);
#pragma unroll
            for (int i = LO + R; i + R < LO + LEN; i += M) cmp_exchange<T>(a[i], a[i + R]);
        } else {
            cmp_exchange<T>(a[LO], a[LO + R]);
        }
    }
};
template <typename T, int N, int LO, int LEN>
struct OddEvenSort {
    __device__ __forceinline__ static void run(T (&a)[N]) {
        if constexpr (LEN > 1) {
            OddEvenSort<T, N, LO, LEN / 2>::run(a);
            OddEvenSort<T, N, LO + LEN / 2, LEN / 2>::run(a);
            OddEvenMerge<T, N, LO, LEN, 1>::run(a);
        }
    }
};
template <typename T, int N>
__device__ __forceinline__ void sort_network(T (&a)[N]) { OddEvenSort<T, N, 0, N>::run(a); }

// VEC: window rows start on 16-byte boundaries and are read with 16-byte loads.
// SORT = false: sum / mean / min / max on the same register window (numpy's summation order:
// pairwise tree for an 8-wide row, sequential below 8, rows added in order).
template <typename T, int F, bool VEC, bool SORT>
__global__ void __launch_bounds__(128)
block_sort_kernel(const T* __restrict__ obj, int64_t s0, int64_t sy, T* __restrict__ out, int64_t nb, int64_t cy,
                  int64_t cx, int op, int skipna) {
    constexpr int N = F * F;
    const int64_t total = nb * cy * cx;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = idx % cx, j = (idx / cx) % cy, b = idx / (cx * cy);
        const T* base = obj + b * s0 + (j * F) * sy + i * F;
        T a[N];
        constexpr int PV = 16 / sizeof(T);  // elements per 16-byte load
        if constexpr (VEC) {
#pragma unroll
            for (int r = 0; r < F; r++)
#pragma unroll
                for (int c = 0; c < F; c += PV) {
                    const Pack<T, PV> pk = *reinterpret_cast<const Pack<T, PV>*>(base + r * sy + c);
#pragma unroll
                    for (int v = 0; v < PV; v++) a[r * F + c + v] = pk.v[v];
                }
        } else {
#pragma unroll
            for (int r = 0; r < F; r++)
#pragma unroll
                for (int c = 0; c < F; c++) a[r * F + c] = __ldg(base + r * sy + c);
        }
        int n = 0;
#pragma unroll
        for (int q = 0; q < N; q++) n += (a[q] == a[q]) ? 1 : 0;
        T res;
        if constexpr (!SORT) {
            if (op == B200CS_OP_SUM || op == B200CS_OP_MEAN) {
                T acc = T(0);
#pragma unroll
                for (int r = 0; r < F; r++) {
                    T v[F];
#pragma unroll
                    for (int c = 0; c < F; c++) v[c] = skipna ? nan_to_zero(a[r * F + c]) : a[r * F + c];
                    T rs;
                    if constexpr (F == 8) {
                        rs = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
                    } else {
                        rs = T(0);
#pragma unroll
                        for (int c = 0; c < F; c++) rs += v[c];
                    }
                    acc += rs;
                }
                res = (op == B200CS_OP_SUM) ? acc : acc / (T)(skipna ? n : N);
            } else {
                T m = T(0);
                bool any = false;
#pragma unroll
                for (int q = 0; q < N; q++) {
                    const T v = a[q];
                    const bool ok = (v == v);
                    const bool take = ok && (!any || (op == B200CS_OP_MIN ? (v < m) : (v > m)));
                    m = take ? v : m;
                    any = any || ok;
                }
                res = ((n < N && !skipna) || !any) ? (T)NAN : m;
            }
            out[idx] = res;
            continue;
        }
#pragma unroll
        for (int q = 0; q < N; q++) a[q] = (a[q] == a[q]) ? a[q] : (T)INFINITY;
        sort_network<T, N>(a);
        if ((n < N && !skipna) || n == 0) {
            res = (T)NAN;
        } else if (op == B200CS_OP_MEDIAN) {
            // np.median: mean of the two middle elements for even n
            T lo = a[N / 2 - 1], hi = a[N / 2];
            if (n < N) {
                const int k1 = (n - 1) / 2, k2 = n / 2;
#pragma unroll
                for (int q = 0; q <= N / 2; q++) {
                    lo = (q == k1) ? a[q] : lo;
                    hi = (q == k2) ? a[q] : hi;
                }
            }
            res = (n & 1) ? hi : (lo + hi) / T(2);
        } else {  // mode: most frequent, ties -> smallest (scipy.stats.mode)
            T best = a[0];
            int bestc = 1, run = 1;
#pragma unroll
            for (int q = 1; q < N; q++) {
                run = (a[q] == a[q - 1]) ? run + 1 : 1;
                if (q < n && run > bestc) { bestc = run; best = a[q]; }
            }
            res = best;
        }
        out[idx] = res;
    }
}

template <typename T>
static bool launch_block_sort(const T* obj, const int64_t s[2], T* out, int64_t nb, int64_t cy, int64_t cx, int f,
                              int op, int skipna, cudaStream_t st) {
    const int64_t total = nb * cy * cx;
    int64_t blocks = (total + 127) / 128;
    if (blocks > 148LL * 64) blocks = 148LL * 64;
    const unsigned g = (unsigned)blocks;
    constexpr int PV = 16 / (int)sizeof(T);
    const bool vec = f % PV == 0 && s[0] % PV == 0 && s[1] % PV == 0 && ((uintptr_t)obj & 15) == 0;
    const bool sort = (op == B200CS_OP_MEDIAN || op == B200CS_OP_MODE);
#define B200CS_SORT_LAUNCH1(F, V, S) block_sort_kernel<T, F, V, S><<<g, 128, 0, st>>>(obj, s[0], s[1], out, nb, cy, cx, op, skipna)
#define B200CS_SORT_LAUNCH(F)                                                                      \
    do {                                                                                           \
        constexpr bool kV = ((F) % PV == 0);                                                       \
        if (vec && kV) { if (sort) B200CS_SORT_LAUNCH1(F, kV, true); else B200CS_SORT_LAUNCH1(F, kV, false); } \
        else { if (sort) B200CS_SORT_LAUNCH1(F, false, true); else B200CS_SORT_LAUNCH1(F, false, false); }     \
    } while (0)
    if (f == 8) B200CS_SORT_LAUNCH(8);
    else if (f == 4) B200CS_SORT_LAUNCH(4);
    else if (f == 2) B200CS_SORT_LAUNCH(2);
    else return false;
#undef B200CS_SORT_LAUNCH
#undef B200CS_SORT_LAUNCH1
    return true;
}

// ---------------------------------------------------------------------------------------
// K4 block upsample (repeat); 4- or 8-byte elements.
// ---------------------------------------------------------------------------------------
// One CTA per output row: the row decomposition (64-bit division) happens once per CTA, the
// coarse row is read once (cx values), the fine row is written with full-width coalesced stores.
template <typename T>
__global__ void __launch_bounds__(128)
upsample_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t nb, int64_t cy, int64_t cx,
                int64_t ny, int64_t nx, int fy, int fx) {
    const int64_t rows = nb * ny;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const int64_t b = row / ny, y = row - b * ny;
        // a staggered axis has (c-1)*f+1 points: index/f maps the last point onto c-1
        const T* src = in + (b * cy + y / fy) * cx;
        T* dst = out + row * nx;
        for (int x = threadIdx.x; x < (int)nx; x += blockDim.x) dst[x] = __ldg(src + x / fx);
    }
}

// Unstaggered windows whose width is a multiple of the 16-byte vector: one CTA per COARSE row; a
// thread reads one coarse value and writes it as 16-byte stores into the fy fine rows below it
// (fy * nx elements per CTA instead of nx: fewer, fatter CTAs and a quarter of the store
// instructions).
template <typename T, int VEC>
__global__ void __launch_bounds__(128)
upsample_rows_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t rows, int64_t cx, int64_t nx, int fy,
                     int fx) {
    const int nv = (int)(nx / VEC);
    for (int64_t crow = blockIdx.x; crow < rows; crow += gridDim.x) {
        const T* src = in + crow * cx;
        T* dst = out + crow * fy * nx;
        for (int xv = threadIdx.x; xv < nv; xv += blockDim.x) {
            const T sv = __ldg(src + (xv * VEC) / fx);
            Pack<T, VEC> pk;
#pragma unroll
            for (int v = 0; v < VEC; v++) pk.v[v] = sv;
            for (int r = 0; r < fy; r++) *reinterpret_cast<Pack<T, VEC>*>(dst + (int64_t)r * nx + (int64_t)xv * VEC) = pk;
        }
    }
}

// gathers columns 0, step, 2*step.. of [nb, ny, nx] into line-major [nb, nlines, ny]
template <typename T>
__global__ void __launch_bounds__(256)
gather_columns_kernel(const T* __restrict__ in, int64_t s0, int64_t sy, T* __restrict__ out, int64_t nb,
                      int64_t ny, int64_t nlines, int step) {
    // tile: 32 lines x 32 rows through shared memory so both sides stay coalesced-ish
    __shared__ T tile[32][33];
    const int64_t b = blockIdx.z;
    const int64_t y0 = (int64_t)blockIdx.y * 32, l0 = (int64_t)blockIdx.x * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t y = y0 + r, l = l0 + threadIdx.x;
        if (y < ny && l < nlines) tile[r][threadIdx.x] = in[b * s0 + y * sy + l * step];
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int64_t l = l0 + r, y = y0 + threadIdx.x;
        if (y < ny && l < nlines) out[(b * nlines + l) * ny + y] = tile[threadIdx.x][r];
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
pressure_interface_kernel(const T* __restrict__ delp, int64_t s0, int64_t sl, T* __restrict__ out,
                          int64_t nb0, int64_t km, int64_t ncol, T toa) {
    const int64_t total = nb0 * ncol;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t c = idx % ncol, b = idx / ncol;
        T acc = toa;
        T* o = out + b * (km + 1) * ncol + c;
        o[0] = acc;
        for (int64_t k = 0; k < km; k++) {
            acc += nan_to_zero(delp[b * s0 + k * sl + c]);  // xarray cumsum == nancumsum
            o[(k + 1) * ncol] = acc;
        }
    }
}

// grid = (column chunks, level chunks, batch).  A thread owns VEC consecutive columns: the
// surface pressure and the weight are loaded once and kept in registers while it walks kMwLevels
// levels, each one vector load of the coarse pressure and one vector store.
constexpr int kMwLevels = 8;

template <typename TW, typename TP, int VEC>
__global__ void __launch_bounds__(256)
mask_weights_kernel(const TW* __restrict__ w, const TP* __restrict__ pc, const TP* __restrict__ pf,
                    TW* __restrict__ out, int64_t nb0, int64_t km, int64_t ncol, int mode) {
    using PW = Pack<TW, VEC>;
    using PP = Pack<TP, VEC>;
    const int64_t k0 = (int64_t)blockIdx.y * kMwLevels;
    const int64_t k1 = (k0 + kMwLevels < km) ? k0 + kMwLevels : km;
    for (int64_t b = blockIdx.z; b < nb0; b += gridDim.z) {
        for (int64_t c = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * VEC; c < ncol;
             c += (int64_t)gridDim.x * blockDim.x * VEC) {
            const PP surf = *reinterpret_cast<const PP*>(pf + (b * (km + 1) + km) * ncol + c);
            const PW wv = *reinterpret_cast<const PW*>(w + b * ncol + c);
            const TP* p = (mode == B200CS_MASK_PHALF) ? pc + (b * (km + 1) + k0 + 1) * ncol + c : pc + (b * km + k0) * ncol + c;
            TW* o = out + (b * km + k0) * ncol + c;
#pragma unroll 4
            for (int64_t k = k0; k < k1; k++, p += ncol, o += ncol) {
                PW r = wv;
                if (mode != B200CS_MASK_NONE) {
                    const PP pv = *reinterpret_cast<const PP*>(p);
#pragma unroll
                    for (int v = 0; v < VEC; v++) r.v[v] = (pv.v[v] < surf.v[v]) ? wv.v[v] : TW(0);
                }
                *reinterpret_cast<PW*>(o) = r;
            }
        }
    }
}

template <typename TW, typename TP>
static void launch_mask_weights(const void* w, const void* pc, const void* pf, void* out, int64_t nb0, int64_t km,
                                int64_t ncol, int mode, cudaStream_t st) {
    const unsigned gy = (unsigned)((km + kMwLevels - 1) / kMwLevels), gz = (unsigned)(nb0 < 65535 ? nb0 : 65535);
    const uintptr_t bits = (uintptr_t)w | (uintptr_t)pc | (uintptr_t)pf | (uintptr_t)out;
    if (ncol % 4 == 0 && (bits & (4 * sizeof(TP) - 1)) == 0 && (bits & (4 * sizeof(TW) - 1)) == 0) {
        const int64_t gx = (ncol / 4 + 255) / 256;
        mask_weights_kernel<TW, TP, 4><<<dim3((unsigned)(gx < 65535 ? gx : 65535), gy, gz), 256, 0, st>>>(
            (const TW*)w, (const TP*)pc, (const TP*)pf, (TW*)out, nb0, km, ncol, mode);
    } else {
        const int64_t gx = (ncol + 255) / 256;
        mask_weights_kernel<TW, TP, 1><<<dim3((unsigned)(gx < 65535 ? gx : 65535), gy, gz), 256, 0, st>>>(
            (const TW*)w, (const TP*)pc, (const TP*)pf, (TW*)out, nb0, km, ncol, mode);
    }
}

static unsigned grid_for(int64_t total, int threads) {
    int64_t blocks = (total + threads - 1) / threads;
    const int64_t cap = 148LL * 32;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}

}  // namespace b200cs

using namespace b200cs;

extern "C" int b200cs_weighted_block_average(const void* obj, int obj_dtype, const int64_t so[3],
                                             const void* w, int w_dtype, const int64_t sw[3],
                                             const void* w2, int w2_dtype, const int64_t sv[3],
                                             void* out, int out_dtype, int64_t nb0, int64_t nb1,
                                             int64_t ny, int64_t nx, int fy, int fx, int ystep,
                                             int xstep, void* stream) {
    B200CS_REQUIRE(obj && w && out && so && sw, "weighted_block_average: NULL argument");
    B200CS_REQUIRE(!w2 || sv, "weighted_block_average: w2 given without strides");
    B200CS_REQUIRE(fy >= 1 && fx >= 1 && ystep >= 1 && xstep >= 1, "weighted_block_average: bad window");
    B200CS_REQUIRE(fx <= 128, "weighted_block_average: fx > 128 not supported (numpy pairwise recursion)");
    B200CS_REQUIRE(nb0 >= 0 && nb1 >= 0 && ny >= fy && nx >= fx, "weighted_block_average: bad extents");
    B200CS_REQUIRE((obj_dtype | 1) == 1 && (w_dtype | 1) == 1 && (!w2 || (w2_dtype | 1) == 1),
                   "weighted_block_average: dtype must be F32 or F64");
    const int want = (obj_dtype == B200CS_F64 || w_dtype == B200CS_F64 || (w2 && w2_dtype == B200CS_F64))
                         ? B200CS_F64 : B200CS_F32;
    B200CS_REQUIRE(out_dtype == want, "weighted_block_average: out_dtype must be the numpy promotion (%d)", want);
    const int64_t cy = (ny - fy) / ystep + 1, cx = (nx - fx) / xstep + 1;
    cudaStream_t st = as_stream(stream);

    // fast path: 8x8 fp32 blocks with 16-byte aligned rows
    const bool all32 = obj_dtype == B200CS_F32 && w_dtype == B200CS_F32 && (!w2 || w2_dtype == B200CS_F32);
    const bool geo8 = fy == 8 && fx == 8 && ystep == 8 && xstep == 8 && ny % 8 == 0 && nx % 8 == 0;
    auto s4 = [](const int64_t* s) { return s[0] % 4 == 0 && s[1] % 4 == 0 && s[2] % 4 == 0; };
    if (all32 && geo8 && aligned16(obj) && aligned16(w) && s4(so) && s4(sw) &&
        (!w2 || (aligned16(w2) && s4(sv))) && nb0 * nb1 * cy * cx > 0) {
        const int64_t ncell = nb0 * cy * cx;
        const bool w3d = sw[1] != 0;
        // levels per thread: 1 keeps the grid at several waves at C384 (13,824 cells x 79 levels); the
        // 2-D weights are re-read from L2
        dim3 grid((unsigned)((ncell + 127) / 128), 1, 1);
#define B200CS_WBA8(LVV, W3, H2)                                                                     \
    grid.y = (unsigned)((nb1 + LVV - 1) / LVV);                                                      \
    wba8_f32_kernel<LVV, W3, H2><<<grid, 128, 0, st>>>((const float*)obj, so[0], so[1], so[2],       \
                                                       (const float*)w, sw[0], sw[1], sw[2],         \
                                                       (const float*)w2, w2 ? sv[0] : 0,             \
                                                       w2 ? sv[1] : 0, w2 ? sv[2] : 0, (float*)out,  \
                                                       nb0, nb1, cy, cx)
        if (w2) { if (w3d) { B200CS_WBA8(2, true, true); } else { B200CS_WBA8(2, false, true); } }
        else if (w3d) { B200CS_WBA8(2, true, false); }
        else {
            const int64_t nitem = (int64_t)grid.x * nb1;  // (block of 128 cells, level)
            int64_t nct = 148 * 5;                        // one wave: five CTAs per SM
            if (nct * 4 > nitem) nct = (nitem + 3) / 4;   // at least ~4 levels per run
            wba8_f32_w2d_kernel<<<(unsigned)nct, 128, 0, st>>>((const float*)obj, so[0], so[1], so[2], (const float*)w, sw[0],
                                                               sw[2], (float*)out, nb0, (int)nb1, cy, cx, nitem);
        }
#undef B200CS_WBA8
        return check_launch("weighted_block_average(8x8 f32)");
    }

    // wide square windows (16, 32, 64), fp32, no second weight: streaming kernel
    const bool geow = fy == fx && ystep == fx && xstep == fx && (fx == 16 || fx == 32 || fx == 64) && ny % fx == 0 &&
                      nx % fx == 0;
    if (all32 && geow && !w2 && aligned16(obj) && aligned16(w) && s4(so) && s4(sw) && cy <= 65535 &&
        nb0 * nb1 <= 65535 && nb0 * nb1 * cy * cx > 0) {
        if (fx == 64) return launch_wba_wide<64>(obj, so, w, sw, out, nb0, nb1, cy, cx, st);
        if (fx == 32) return launch_wba_wide<32>(obj, so, w, sw, out, nb0, nb1, cy, cx, st);
        return launch_wba_wide<16>(obj, so, w, sw, out, nb0, nb1, cy, cx, st);
    }

    if (obj_dtype == B200CS_F32 && w_dtype == B200CS_F32)
        return dispatch_w2<float, float>(obj, so, w, sw, w2, w2_dtype, sv, out, nb0, nb1, cy, cx, fy, fx, ystep, xstep, st);
    if (obj_dtype == B200CS_F32 && w_dtype == B200CS_F64)
        return dispatch_w2<float, double>(obj, so, w, sw, w2, w2_dtype, sv, out, nb0, nb1, cy, cx, fy, fx, ystep, xstep, st);
    if (obj_dtype == B200CS_F64 && w_dtype == B200CS_F32)
        return dispatch_w2<double, float>(obj, so, w, sw, w2, w2_dtype, sv, out, nb0, nb1, cy, cx, fy, fx, ystep, xstep, st);
    return dispatch_w2<double, double>(obj, so, w, sw, w2, w2_dtype, sv, out, nb0, nb1, cy, cx, fy, fx, ystep, xstep, st);
}

extern "C" int b200cs_block_reduce(const void* obj, int dtype, const int64_t s[2], void* out, int64_t nb,
                                   int64_t ny, int64_t nx, int fy, int fx, int ystep, int xstep, int op,
                                   int skipna, int pad, double cval, void* stream) {
    B200CS_REQUIRE(obj && out && s, "block_reduce: NULL argument");
    B200CS_REQUIRE(fy >= 1 && fx >= 1 && ystep >= 1 && xstep >= 1, "block_reduce: bad window");
    B200CS_REQUIRE(fx <= 128, "block_reduce: fx > 128 not supported");
    B200CS_REQUIRE(op >= B200CS_OP_SUM && op <= B200CS_OP_MODE, "block_reduce: unknown op %d", op);
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "block_reduce: dtype must be F32 or F64");
    if ((op == B200CS_OP_MEDIAN || op == B200CS_OP_MODE) && (int64_t)fy * fx > kMaxSortWindow)
        return set_err(B200CS_EUNSUPPORTED, "block_reduce: median/mode window %d x %d exceeds %d elements",
                       fy, fx, kMaxSortWindow);
    int64_t cy, cx;
    if (pad) {
        B200CS_REQUIRE(ystep == fy && xstep == fx, "block_reduce: padding needs non-overlapping windows");
        cy = (ny + fy - 1) / fy; cx = (nx + fx - 1) / fx;
    } else {
        B200CS_REQUIRE(ny >= fy && nx >= fx, "block_reduce: window larger than array");
        cy = (ny - fy) / ystep + 1; cx = (nx - fx) / xstep + 1;
    }
    const int64_t total = nb * cy * cx;
    if (total == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (!pad && fy == fx && ystep == fy && xstep == fx) {  // register-window kernels (2x2, 4x4, 8x8)
        const bool done = (dtype == B200CS_F32)
                              ? launch_block_sort<float>((const float*)obj, s, (float*)out, nb, cy, cx, fy, op, skipna, st)
                              : launch_block_sort<double>((const double*)obj, s, (double*)out, nb, cy, cx, fy, op, skipna, st);
        if (done) return check_launch("block_reduce(register window)");
    }
    if (dtype == B200CS_F32)
        block_reduce_kernel<float><<<grid_for(total, 128), 128, 0, st>>>((const float*)obj, s[0], s[1], (float*)out,
            nb, ny, nx, cy, cx, fy, fx, ystep, xstep, op, skipna, (float)cval);
    else
        block_reduce_kernel<double><<<grid_for(total, 128), 128, 0, st>>>((const double*)obj, s[0], s[1], (double*)out,
            nb, ny, nx, cy, cx, fy, fx, ystep, xstep, op, skipna, cval);
    return check_launch("block_reduce");
}

extern "C" int b200cs_block_upsample(const void* in, void* out, int elem_size, int64_t nb, int64_t cy,
                                     int64_t cx, int fy, int fx, int y_staggered, int x_staggered,
                                     void* stream) {
    B200CS_REQUIRE(in && out, "block_upsample: NULL argument");
    B200CS_REQUIRE(fy >= 1 && fx >= 1 && cy >= 1 && cx >= 1, "block_upsample: bad extents");
    B200CS_REQUIRE(elem_size == 1 || elem_size == 2 || elem_size == 4 || elem_size == 8,
                   "block_upsample: elem_size must be 1, 2, 4 or 8");
    const int64_t ny = y_staggered ? (cy - 1) * fy + 1 : cy * fy;
    const int64_t nx = x_staggered ? (cx - 1) * fx + 1 : cx * fx;
    const int64_t total = nb * ny * nx;
    if (total == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    const int vec = 16 / elem_size;
    if (elem_size >= 4 && !y_staggered && !x_staggered && fx % vec == 0 && ((uintptr_t)out & 15) == 0) {
        const int64_t rows = nb * cy;
        const unsigned g = (unsigned)(rows < (1 << 30) ? rows : (1 << 30));
        if (elem_size == 4)
            upsample_rows_kernel<uint32_t, 4><<<g, 128, 0, st>>>((const uint32_t*)in, (uint32_t*)out, rows, cx, nx, fy, fx);
        else
            upsample_rows_kernel<uint64_t, 2><<<g, 128, 0, st>>>((const uint64_t*)in, (uint64_t*)out, rows, cx, nx, fy, fx);
        return check_launch("block_upsample");
    }
    if (elem_size == 1)  // masks (bool / int8) and 2-byte dtypes keep their width: the op only moves bits
        upsample_kernel<uint8_t><<<(unsigned)(nb * ny < (1 << 30) ? nb * ny : (1 << 30)), 128, 0, st>>>((const uint8_t*)in, (uint8_t*)out, nb, cy, cx, ny, nx, fy, fx);
    else if (elem_size == 2)
        upsample_kernel<uint16_t><<<(unsigned)(nb * ny < (1 << 30) ? nb * ny : (1 << 30)), 128, 0, st>>>((const uint16_t*)in, (uint16_t*)out, nb, cy, cx, ny, nx, fy, fx);
    else if (elem_size == 4)
        upsample_kernel<uint32_t><<<(unsigned)(nb * ny < (1 << 30) ? nb * ny : (1 << 30)), 128, 0, st>>>((const uint32_t*)in, (uint32_t*)out, nb, cy, cx, ny, nx, fy, fx);
    else
        upsample_kernel<uint64_t><<<(unsigned)(nb * ny < (1 << 30) ? nb * ny : (1 << 30)), 128, 0, st>>>((const uint64_t*)in, (uint64_t*)out, nb, cy, cx, ny, nx, fy, fx);
    return check_launch("block_upsample");
}

extern "C" int b200cs_gather_columns(const void* in, const int64_t s[2], void* out, int elem_size,
                                     int64_t nb, int64_t ny, int64_t nx, int step, void* stream) {
    B200CS_REQUIRE(in && out && s && step >= 1 && nx >= 1, "gather_columns: bad argument");
    B200CS_REQUIRE(elem_size == 4 || elem_size == 8, "gather_columns: elem_size must be 4 or 8");
    B200CS_REQUIRE(nb <= 65535, "gather_columns: nb > 65535");
    const int64_t nlines = (nx - 1) / step + 1;
    if (nb * ny * nlines == 0) return B200CS_OK;
    dim3 grid((unsigned)((nlines + 31) / 32), (unsigned)((ny + 31) / 32), (unsigned)nb), block(32, 8);
    cudaStream_t st = as_stream(stream);
    if (elem_size == 4)
        gather_columns_kernel<uint32_t><<<grid, block, 0, st>>>((const uint32_t*)in, s[0], s[1], (uint32_t*)out, nb, ny, nlines, step);
    else
        gather_columns_kernel<uint64_t><<<grid, block, 0, st>>>((const uint64_t*)in, s[0], s[1], (uint64_t*)out, nb, ny, nlines, step);
    return check_launch("gather_columns");
}

extern "C" int b200cs_transpose_last2(const void* in, void* out, int elem_size, int64_t nb, int64_t ny,
                                      int64_t nx, void* stream) {
    const int64_t s[2] = {ny * nx, nx};
    return b200cs_gather_columns(in, s, out, elem_size, nb, ny, nx, 1, stream);
}

extern "C" int b200cs_pressure_at_interface(const void* delp, int dtype, const int64_t s[2], void* out,
                                            int64_t nb0, int64_t km, int64_t ncol, double toa, void* stream) {
    B200CS_REQUIRE(delp && out && s && km >= 0, "pressure_at_interface: bad argument");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "pressure_at_interface: dtype must be F32 or F64");
    const int64_t total = nb0 * ncol;
    if (total == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        pressure_interface_kernel<float><<<grid_for(total, 256), 256, 0, st>>>((const float*)delp, s[0], s[1], (float*)out, nb0, km, ncol, (float)toa);
    else
        pressure_interface_kernel<double><<<grid_for(total, 256), 256, 0, st>>>((const double*)delp, s[0], s[1], (double*)out, nb0, km, ncol, toa);
    return check_launch("pressure_at_interface");
}

extern "C" int b200cs_mask_weights(const void* w, int w_dtype, const void* pc, const void* pf, int p_dtype,
                                   void* out, int64_t nb0, int64_t km, int64_t ncol, int mode, void* stream) {
    B200CS_REQUIRE(w && pc && pf && out, "mask_weights: NULL argument");
    B200CS_REQUIRE(mode >= B200CS_MASK_NONE && mode <= B200CS_MASK_PFULL, "mask_weights: bad mask_mode");
    const int64_t total = nb0 * km * ncol;
    if (total == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (w_dtype == B200CS_F32 && p_dtype == B200CS_F32) launch_mask_weights<float, float>(w, pc, pf, out, nb0, km, ncol, mode, st);
    else if (w_dtype == B200CS_F32 && p_dtype == B200CS_F64) launch_mask_weights<float, double>(w, pc, pf, out, nb0, km, ncol, mode, st);
    else if (w_dtype == B200CS_F64 && p_dtype == B200CS_F32) launch_mask_weights<double, float>(w, pc, pf, out, nb0, km, ncol, mode, st);
    else if (w_dtype == B200CS_F64 && p_dtype == B200CS_F64) launch_mask_weights<double, double>(w, pc, pf, out, nb0, km, ncol, mode, st);
    else
        return set_err(B200CS_EINVAL, "mask_weights: dtype must be F32 or F64");
    return check_launch("mask_weights");
}
