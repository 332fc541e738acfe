// stagger.cu -- K5: centre -> edge interpolation of a 6-tile cubed-sphere field with ghost
// lines taken from the neighbouring faces.  Replaces xgcm 0.6.1 Grid.interp(delp, axis) built on
// FV3_FACE_CONNECTIONS (external/vcm/vcm/cubedsphere/xgcm.py:7-34; call sites
// external/vcm/vcm/cubedsphere/regridz.py:118-132 and coarsen_restarts.py:825-853).
//
//   out[j] = 0.5 * (ext[j-1] + ext[j]),  j = 0..n,   ext = tile padded by one ghost line a side
//   left ghost  = LAST  line of the left  neighbour along ITS connecting axis
//   right ghost = FIRST line of the right neighbour along ITS connecting axis
//   a ghost line is index-reversed when the neighbour's axis differs from ours
// (rule pinned bit-exactly against the reference's golden u, v: tests/test_oracle.py).
//
// Only lines 0, step, 2*step.. are produced when line_step > 1 (the D-grid pressure path keeps
// only those, external/vcm/vcm/cubedsphere/coarsen.py:263-264), written compactly and, for the
// x axis, line-major, so each kept line is contiguous.  Pure streaming: every needed input line
// is read once, coalesced along the line.
#include "common.cuh"

namespace b200cs {

// connection table: [tile][axis: 0=x,1=y][side: 0=left,1=right] -> {neighbour tile, neighbour axis}
__constant__ int kConn[6][2][2][2] = {
    {{{4, 1}, {1, 0}}, {{5, 1}, {2, 0}}},
    {{{0, 0}, {3, 1}}, {{5, 0}, {2, 1}}},
    {{{0, 1}, {3, 0}}, {{1, 1}, {4, 0}}},
    {{{2, 0}, {5, 1}}, {{1, 0}, {4, 1}}},
    {{{2, 1}, {5, 0}}, {{3, 1}, {0, 0}}},
    {{{4, 0}, {1, 1}}, {{3, 0}, {0, 1}}},
};

// element `i` (position along the line) of interior/ghost line `j` in [-1, n] of tile t, level b1
template <typename T>
__device__ __forceinline__ T ext_at(const T* in, int64_t nb1, int64_t n, int t, int64_t b1, int axis_xy,
                                    int64_t j, int64_t i) {
    // axis_xy: 0 = lines are columns (stagger along x), 1 = lines are rows (stagger along y)
    int tt = t;
    int ax = axis_xy;
    int64_t jj = j, ii = i;
    if (j < 0 || j >= n) {
        const int side = (j < 0) ? 0 : 1;
        tt = kConn[t][axis_xy][side][0];
        ax = kConn[t][axis_xy][side][1];
        jj = (side == 0) ? n - 1 : 0;
        if (ax != axis_xy) ii = n - 1 - i;
    }
    const T* base = in + ((int64_t)tt * nb1 + b1) * n * n;
    // ax == 0: line index is the x coordinate, position runs along y
    return (ax == 0) ? base[ii * n + jj] : base[jj * n + ii];
}

// One CTA per output line (line-major outputs) or per output row (the full x-stagger): the line
// decomposition happens once per CTA; interior lines are two coalesced row reads (or two column
// reads for x lines), only lines 0 and n go through the face-connection table.
template <typename T>
__global__ void __launch_bounds__(128)
stagger_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t nb1, int64_t n, int axis, int step) {
    // axis: 0 = y (lines are rows), 1 = x (lines are columns)
    const int axis_xy = (axis == 0) ? 1 : 0;
    const int64_t nlines = n / step + 1;
    if (axis == 0 || step > 1) {  // line-major output [6, nb1, nlines, n]
        const int64_t total = 6 * nb1 * nlines;
        for (int64_t line = blockIdx.x; line < total; line += gridDim.x) {
            const int64_t l = line % nlines, b1 = (line / nlines) % nb1;
            const int t = (int)(line / (nlines * nb1));
            const int64_t j = l * step;
            T* dst = out + line * n;
            if (j > 0 && j < n) {
                const T* base = in + ((int64_t)t * nb1 + b1) * n * n;
                if (axis == 0) {
                    const T* r0 = base + (j - 1) * n;
                    const T* r1 = base + j * n;
                    for (int i = threadIdx.x; i < (int)n; i += blockDim.x) dst[i] = T(0.5) * (__ldg(r0 + i) + __ldg(r1 + i));
                } else {
                    for (int i = threadIdx.x; i < (int)n; i += blockDim.x)
                        dst[i] = T(0.5) * (__ldg(base + (int64_t)i * n + j - 1) + __ldg(base + (int64_t)i * n + j));
                }
            } else {
                for (int i = threadIdx.x; i < (int)n; i += blockDim.x)
                    dst[i] = T(0.5) * (ext_at(in, nb1, n, t, b1, axis_xy, j - 1, (int64_t)i) +
                                       ext_at(in, nb1, n, t, b1, axis_xy, j, (int64_t)i));
            }
        }
    } else {  // full x-stagger keeps the natural [6, nb1, n, n+1] layout: one CTA per row
        const int64_t total = 6 * nb1 * n;
        for (int64_t row = blockIdx.x; row < total; row += gridDim.x) {
            const int64_t y = row % n, b1 = (row / n) % nb1;
            const int t = (int)(row / (n * nb1));
            const T* src = in + (((int64_t)t * nb1 + b1) * n + y) * n;
            T* dst = out + row * (n + 1);
            for (int jx = threadIdx.x; jx <= (int)n; jx += blockDim.x) {
                T a, b;
                if (jx > 0 && jx < (int)n) {
                    a = __ldg(src + jx - 1);
                    b = __ldg(src + jx);
                } else {
                    a = ext_at(in, nb1, n, t, b1, axis_xy, (int64_t)jx - 1, y);
                    b = ext_at(in, nb1, n, t, b1, axis_xy, (int64_t)jx, y);
                }
                dst[jx] = T(0.5) * (a + b);
            }
        }
    }
}

// y-stagger of EVERY line (line_step 1): a CTA owns a band of kBand consecutive output lines of
// one (tile, level); a thread owns one 16-byte column group, walks down the band and keeps the
// previous row in registers, so each input row is read once per band (9 row reads for 8 lines
// instead of 16) with 16-byte loads and stores.  Lines 0 and n take the face-connection path.
constexpr int kBand = 8;

template <typename T, int VEC>
struct alignas(sizeof(T) * VEC) StagPack { T v[VEC]; };

template <typename T, int VEC>
__global__ void __launch_bounds__(128)
stagger_y_band_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t nb1, int64_t n) {
    using P = StagPack<T, VEC>;
    const int64_t tb = blockIdx.y, b1 = tb % nb1;
    const int t = (int)(tb / nb1);
    const int64_t j0 = (int64_t)blockIdx.x * kBand;
    const T* base = in + tb * n * n;
    T* dst = out + tb * (n + 1) * n;
    for (int64_t x = (int64_t)threadIdx.x * VEC; x < n; x += (int64_t)blockDim.x * VEC) {
        P prev;
        if (j0 > 0) prev = *reinterpret_cast<const P*>(base + (j0 - 1) * n + x);
        for (int dj = 0; dj < kBand; dj++) {
            const int64_t j = j0 + dj;
            if (j > n) break;
            P o;
            if (j == 0 || j == n) {
#pragma unroll
                for (int v = 0; v < VEC; v++)
                    o.v[v] = T(0.5) * (ext_at(in, nb1, n, t, b1, 1, j - 1, x + v) + ext_at(in, nb1, n, t, b1, 1, j, x + v));
                if (j == 0) prev = *reinterpret_cast<const P*>(base + x);
            } else {
                const P cur = *reinterpret_cast<const P*>(base + j * n + x);
#pragma unroll
                for (int v = 0; v < VEC; v++) o.v[v] = T(0.5) * (prev.v[v] + cur.v[v]);
                prev = cur;
            }
            *reinterpret_cast<P*>(dst + j * n + x) = o;
        }
    }
}

}  // namespace b200cs

using namespace b200cs;

extern "C" int b200cs_stagger(const void* in, void* out, int dtype, int64_t nb1, int64_t n, int axis,
                              int line_step, void* stream) {
    B200CS_REQUIRE(in && out, "stagger: NULL argument");
    B200CS_REQUIRE(axis == 0 || axis == 1, "stagger: axis must be 0 (y) or 1 (x)");
    B200CS_REQUIRE(line_step >= 1 && n >= 1 && n % line_step == 0, "stagger: line_step must divide n");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "stagger: dtype must be F32 or F64");
    const int64_t total = 6 * nb1 * (n / line_step + 1) * n;
    if (total == 0) return B200CS_OK;
    int64_t blocks = (axis == 0 || line_step > 1) ? 6 * nb1 * (n / line_step + 1) : 6 * nb1 * n;
    if (blocks > (1LL << 30)) blocks = 1LL << 30;
    cudaStream_t st = as_stream(stream);
    const int64_t vec = dtype == B200CS_F32 ? 4 : 2;
    if (axis == 0 && line_step == 1 && n % vec == 0 && (((uintptr_t)in | (uintptr_t)out) & 15) == 0 && 6 * nb1 <= 65535) {
        const dim3 grid((unsigned)((n + 1 + kBand - 1) / kBand), (unsigned)(6 * nb1));
        if (dtype == B200CS_F32) stagger_y_band_kernel<float, 4><<<grid, 128, 0, st>>>((const float*)in, (float*)out, nb1, n);
        else stagger_y_band_kernel<double, 2><<<grid, 128, 0, st>>>((const double*)in, (double*)out, nb1, n);
        return check_launch("stagger(y band)");
    }
    if (dtype == B200CS_F32)
        stagger_kernel<float><<<(unsigned)blocks, 128, 0, st>>>((const float*)in, (float*)out, nb1, n, axis, line_step);
    else
        stagger_kernel<double><<<(unsigned)blocks, 128, 0, st>>>((const double*)in, (double*)out, nb1, n, axis, line_step);
    return check_launch("stagger");
}
