// elementwise.cu -- masks and selections of the 'complex' sfc_data coarse-graining.
// Replaces the xarray expressions `.where(mask)`, `xr.where(cond, a, b)`, `.fillna(v)`,
// `xarray_utils.isclose(a, b)` and `block_upsample_like` + `isclose` of
// external/vcm/vcm/cubedsphere/coarsen_restarts.py:1162-1470 (external/vcm/vcm/xarray_utils.py:7-35).
// Masks are one byte per element (0 / 1).
#include "common.cuh"

namespace b200cs {

template <typename T>
__device__ __forceinline__ bool is_close(T a, T b, T rtol, T atol) {
    // numpy.isclose, equal_nan=False: finite values within atol + rtol*|b|; infinities only when equal
    if (a != a || b != b) return false;
    if (isinf(a) || isinf(b)) return a == b;
    return fabs(a - b) <= atol + rtol * fabs(b);
}

template <typename T>
__global__ void elementwise_kernel(int op, const T* __restrict__ a, const T* __restrict__ b,
                                   const uint8_t* __restrict__ mask, T* __restrict__ out, int64_t n, T s) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    T r;
    switch (op) {
        case B200CS_EW_MUL: r = a[i] * b[i]; break;
        case B200CS_EW_WHERE: r = mask[i] ? a[i] : s; break;
        case B200CS_EW_WHERE_NOT: r = mask[i] ? s : a[i]; break;
        case B200CS_EW_SELECT: r = mask[i] ? a[i] : b[i]; break;
        case B200CS_EW_FILLNA: { const T v = a[i]; r = (v != v) ? s : v; break; }
        case B200CS_EW_DIFF_OVER: r = (a[i] - b[i]) / s; break;
        case B200CS_EW_MIN_SCALAR: { const T v = a[i]; r = (v < s) ? v : s; break; }
        default: r = a[i]; break;
    }
    out[i] = r;
}

template <typename T>
__global__ void compare_scalar_kernel(int cmp, const T* __restrict__ a, T s, T rtol, T atol,
                                      const uint8_t* __restrict__ and_mask, uint8_t* __restrict__ out, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const T v = a[i];
    bool r;
    if (cmp == B200CS_CMP_ISCLOSE) r = is_close(v, s, rtol, atol);
    else if (cmp == B200CS_CMP_LT) r = v < s;
    else r = v > s;
    if (and_mask) r = r && and_mask[i];
    out[i] = r ? 1 : 0;
}

// out[b, y, x] = isclose(fine[b, y, x], coarse[b, y / f, x / f]) (& and_mask): block_upsample_like +
// isclose without the fine-size upsampled temporary (coarsen_restarts.py:1173-1190)
template <typename T>
__global__ void dominant_mask_kernel(const T* __restrict__ fine, const T* __restrict__ coarse,
                                     const uint8_t* __restrict__ and_mask, uint8_t* __restrict__ out, int64_t nb,
                                     int64_t ny, int64_t nx, int f, T rtol, T atol) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb * ny * nx) return;
    const int64_t x = i % nx, y = (i / nx) % ny, b = i / (nx * ny);
    const int64_t cy = ny / f, cx = nx / f;
    bool r = is_close(fine[i], coarse[(b * cy + y / f) * cx + x / f], rtol, atol);
    if (and_mask) r = r && and_mask[i];
    out[i] = r ? 1 : 0;
}

inline unsigned ew_blocks(int64_t n) { return (unsigned)((n + 255) / 256); }

}  // namespace b200cs

using namespace b200cs;

extern "C" {

int b200cs_elementwise(int op, const void* a, const void* b, const uint8_t* mask, void* out, int dtype, int64_t n,
                       double scalar, void* stream) {
    B200CS_REQUIRE(op >= B200CS_EW_MUL && op <= B200CS_EW_DIFF_OVER, "elementwise: unknown op %d", op);
    B200CS_REQUIRE(a && out, "elementwise: NULL pointer");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "elementwise: bad dtype %d", dtype);
    B200CS_REQUIRE(n >= 0, "elementwise: bad extent");
    if (op == B200CS_EW_MUL || op == B200CS_EW_SELECT || op == B200CS_EW_DIFF_OVER) B200CS_REQUIRE(b, "elementwise: op %d needs b", op);
    if (op == B200CS_EW_WHERE || op == B200CS_EW_WHERE_NOT || op == B200CS_EW_SELECT)
        B200CS_REQUIRE(mask, "elementwise: op %d needs a mask", op);
    if (n == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        elementwise_kernel<float><<<ew_blocks(n), 256, 0, st>>>(op, (const float*)a, (const float*)b, mask, (float*)out, n,
                                                              (float)scalar);
    else
        elementwise_kernel<double><<<ew_blocks(n), 256, 0, st>>>(op, (const double*)a, (const double*)b, mask,
                                                               (double*)out, n, scalar);
    return check_launch("elementwise");
}

int b200cs_compare_scalar(int cmp, const void* a, int dtype, double scalar, double rtol, double atol,
                          const uint8_t* and_mask, uint8_t* out, int64_t n, void* stream) {
    B200CS_REQUIRE(cmp >= B200CS_CMP_ISCLOSE && cmp <= B200CS_CMP_GT, "compare_scalar: unknown comparison %d", cmp);
    B200CS_REQUIRE(a && out, "compare_scalar: NULL pointer");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "compare_scalar: bad dtype %d", dtype);
    B200CS_REQUIRE(n >= 0, "compare_scalar: bad extent");
    if (n == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        compare_scalar_kernel<float><<<ew_blocks(n), 256, 0, st>>>(cmp, (const float*)a, (float)scalar, (float)rtol,
                                                                 (float)atol, and_mask, out, n);
    else
        compare_scalar_kernel<double><<<ew_blocks(n), 256, 0, st>>>(cmp, (const double*)a, scalar, rtol, atol, and_mask,
                                                                  out, n);
    return check_launch("compare_scalar");
}

int b200cs_dominant_mask(const void* fine, const void* coarse, int dtype, int64_t nb, int64_t ny, int64_t nx,
                         int factor, double rtol, double atol, const uint8_t* and_mask, uint8_t* out, void* stream) {
    B200CS_REQUIRE(fine && coarse && out, "dominant_mask: NULL pointer");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "dominant_mask: bad dtype %d", dtype);
    B200CS_REQUIRE(factor >= 1 && nb >= 0 && ny >= 0 && nx >= 0 && ny % factor == 0 && nx % factor == 0,
                   "dominant_mask: factor %d must divide the extents", factor);
    const int64_t n = nb * ny * nx;
    if (n == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        dominant_mask_kernel<float><<<ew_blocks(n), 256, 0, st>>>((const float*)fine, (const float*)coarse, and_mask, out,
                                                                nb, ny, nx, factor, (float)rtol, (float)atol);
    else
        dominant_mask_kernel<double><<<ew_blocks(n), 256, 0, st>>>((const double*)fine, (const double*)coarse, and_mask,
                                                                 out, nb, ny, nx, factor, rtol, atol);
    return check_launch("dominant_mask");
}

}  // extern "C"
