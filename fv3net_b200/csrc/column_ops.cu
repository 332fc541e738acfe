// column_ops.cu -- column-wise post-processing of the coarse-grained state:
//   surface pressure, blending weights, blend, hydrostatic balance.
// Replaces the xarray expressions of
//   external/vcm/vcm/calc/thermo/vertically_dependent.py:69-98,153-235 and
//   external/vcm/vcm/cubedsphere/coarsen_restarts.py:559-667,990-1017
// One thread per column, levels walked sequentially in the reference's order (xarray's cumsum /
// sum over a non-contiguous axis add level slices one after the other; NaNs count as 0 in both).
// Arrays are [nb, km, ncol] with the column index contiguous.
#include "common.cuh"

namespace b200cs {

template <typename T>
__device__ __forceinline__ T ld_lev(const T* p, int64_t b, int64_t k, int64_t c, int64_t sb, int64_t sl) {
    return p[b * sb + k * sl + c];
}

// ps[b, c] = nansum_k delp[b, k, c] + toa   (surface_pressure_from_delp, vertically_dependent.py:189-208)
template <typename T>
__global__ void surface_pressure_kernel(const T* __restrict__ delp, int64_t sb, int64_t sl, T* __restrict__ ps,
                                        int64_t nb, int km, int64_t ncol, T toa) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb * ncol) return;
    const int64_t b = i / ncol, c = i % ncol;
    T s = T(0);
    for (int k = 0; k < km; k++) s = s + nan_to_zero(ld_lev(delp, b, k, c, sb, sl));
    ps[i] = s + toa;
}

// compute_blending_weights (coarsen_restarts.py:559-577) with its inputs formed in place:
//   pfull = delp_c / diff(log(phalf_c))      (pressure_at_midpoint_log, vertically_dependent.py:153-179)
//   ps_c  = nansum(delp_c) + toa
//   pb    = sigma_blend * ps_min             (ps_min = block minimum of the fine surface pressure)
//   w     = pfull > pb ? (ps_c - pfull) / (ps_c - pb) : 1
template <typename T>
__global__ void blending_weights_kernel(const T* __restrict__ delp_c, const T* __restrict__ ps_min,
                                        T* __restrict__ out, int64_t nb, int km, int64_t ncol, T toa, T sigma) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb * ncol) return;
    const int64_t b = i / ncol, c = i % ncol;
    const int64_t sb = (int64_t)km * ncol;
    T s = T(0);
    for (int k = 0; k < km; k++) s = s + nan_to_zero(delp_c[b * sb + k * ncol + c]);
    const T ps = s + toa;
    const T pb = sigma * ps_min[i];
    T p = toa, lp = log(toa);
    for (int k = 0; k < km; k++) {
        const T d = delp_c[b * sb + k * ncol + c];
        p = p + nan_to_zero(d);
        const T ln = log(p);
        const T pfull = d / (ln - lp);
        lp = ln;
        const T w = (ps - pfull) / (ps - pb);
        out[b * sb + k * ncol + c] = (pfull > pb) ? w : T(1);
    }
}

// blend (coarsen_restarts.py:651-667): w * pressure_level + (1 - w) * model_level
template <typename T>
__global__ void blend_kernel(const T* __restrict__ w, const T* __restrict__ a, const T* __restrict__ m,
                             T* __restrict__ out, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const T wi = w[i];
    out[i] = wi * a[i] + (T(1) - wi) * m[i];
}

// _impose_hydrostatic_balance (coarsen_restarts.py:990-1017):
//   height_top = phis/g + sum over levels, bottom up, of (-DZ)   (height_at_interface, :69-98)
//   dz[k]      = -(log p[k+1] - log p[k]) * Rd * T (1 + (Rv/Rd - 1) q) / g   (hydrostatic_dz, :210-235)
//   phis_out   = g * (height_top + nansum_k dz[k])                (dz_and_top_to_phis, :182-186)
template <typename T>
__global__ void hydrostatic_kernel(const T* __restrict__ dz_in, const T* __restrict__ phis, const T* __restrict__ temp,
                                   const T* __restrict__ q, const T* __restrict__ delp, T* __restrict__ dz_out,
                                   T* __restrict__ phis_out, int64_t nb, int km, int64_t ncol, T toa, T grav,
                                   T rdgas, T zvir) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb * ncol) return;
    const int64_t b = i / ncol, c = i % ncol;
    const int64_t sb = (int64_t)km * ncol;
    T h = nan_to_zero(phis[i] / grav);
    for (int k = km - 1; k >= 0; k--) h = h + nan_to_zero(-dz_in[b * sb + k * ncol + c]);
    T p = toa, lp = log(toa), s = T(0);
    for (int k = 0; k < km; k++) {
        const int64_t o = b * sb + k * ncol + c;
        p = p + nan_to_zero(delp[o]);
        const T ln = log(p);
        const T dlogp = ln - lp;
        lp = ln;
        const T tv = temp[o] * (T(1) + zvir * q[o]);
        const T dz = -dlogp * rdgas * tv / grav;
        dz_out[o] = dz;
        s = s + nan_to_zero(dz);
    }
    phis_out[i] = grav * (h + s);
}


// pressure_at_midpoint_log (vertically_dependent.py:153-179): delp / diff(log(p_interface))
template <typename T>
__global__ void midpoint_log_kernel(const T* __restrict__ delp, T* __restrict__ out, int64_t nb, int km, int64_t ncol,
                                    T toa) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb * ncol) return;
    const int64_t b = i / ncol, c = i % ncol;
    const int64_t sb = (int64_t)km * ncol;
    T p = toa, lp = log(toa);
    for (int k = 0; k < km; k++) {
        const T d = delp[b * sb + k * ncol + c];
        p = p + nan_to_zero(d);
        const T ln = log(p);
        out[b * sb + k * ncol + c] = d / (ln - lp);
        lp = ln;
    }
}

// Grid.area_above_fine_surface (workflows/fine_res_budget/budget/budgets.py:62-77), the fine-size
// part: out[b, k, y, x] = pfull_coarse[b, k, y/f, x/f] <= ps[b, y, x] ? area[b, y, x] : 0
template <typename TP, typename TA>
__global__ void exposed_area_kernel(const TP* __restrict__ pfull_c, const TP* __restrict__ ps,
                                    const TA* __restrict__ area, TA* __restrict__ out, int64_t nb, int km, int64_t ny,
                                    int64_t nx, int f) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb * km * ny * nx) return;
    const int64_t x = i % nx, y = (i / nx) % ny, k = (i / (nx * ny)) % km, b = i / (nx * ny * km);
    const int64_t cy = ny / f, cx = nx / f;
    const TP pc = pfull_c[((b * km + k) * cy + y / f) * cx + x / f];
    const int64_t col = (b * ny + y) * nx + x;
    out[i] = (pc <= ps[col]) ? area[col] : TA(0);
}

inline unsigned blocks_for(int64_t n) { return (unsigned)((n + 255) / 256); }

}  // namespace b200cs

using namespace b200cs;

extern "C" {

int b200cs_surface_pressure(const void* delp, int dtype, const int64_t stride[2], void* ps, int64_t nb, int km,
                            int64_t ncol, double toa_pressure, void* stream) {
    B200CS_REQUIRE(delp && ps && stride, "surface_pressure: NULL pointer");
    B200CS_REQUIRE(nb >= 0 && km >= 1 && ncol >= 0, "surface_pressure: bad extents");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "surface_pressure: bad dtype %d", dtype);
    if (nb * ncol == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        surface_pressure_kernel<float><<<blocks_for(nb * ncol), 256, 0, st>>>((const float*)delp, stride[0], stride[1],
                                                                             (float*)ps, nb, km, ncol, (float)toa_pressure);
    else
        surface_pressure_kernel<double><<<blocks_for(nb * ncol), 256, 0, st>>>((const double*)delp, stride[0], stride[1],
                                                                              (double*)ps, nb, km, ncol, toa_pressure);
    return check_launch("surface_pressure");
}

int b200cs_blending_weights(const void* delp_coarse, const void* ps_min, void* out, int dtype, int64_t nb, int km,
                            int64_t ncol, double toa_pressure, double sigma_blend, void* stream) {
    B200CS_REQUIRE(delp_coarse && ps_min && out, "blending_weights: NULL pointer");
    B200CS_REQUIRE(nb >= 0 && km >= 1 && ncol >= 0, "blending_weights: bad extents");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "blending_weights: bad dtype %d", dtype);
    if (nb * ncol == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        blending_weights_kernel<float><<<blocks_for(nb * ncol), 256, 0, st>>>(
            (const float*)delp_coarse, (const float*)ps_min, (float*)out, nb, km, ncol, (float)toa_pressure, (float)sigma_blend);
    else
        blending_weights_kernel<double><<<blocks_for(nb * ncol), 256, 0, st>>>(
            (const double*)delp_coarse, (const double*)ps_min, (double*)out, nb, km, ncol, toa_pressure, sigma_blend);
    return check_launch("blending_weights");
}

int b200cs_blend(const void* weights, const void* pressure_level, const void* model_level, void* out, int dtype,
                 int64_t n, void* stream) {
    B200CS_REQUIRE(weights && pressure_level && model_level && out, "blend: NULL pointer");
    B200CS_REQUIRE(n >= 0, "blend: bad extent");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "blend: bad dtype %d", dtype);
    if (n == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        blend_kernel<float><<<blocks_for(n), 256, 0, st>>>((const float*)weights, (const float*)pressure_level,
                                                          (const float*)model_level, (float*)out, n);
    else
        blend_kernel<double><<<blocks_for(n), 256, 0, st>>>((const double*)weights, (const double*)pressure_level,
                                                           (const double*)model_level, (double*)out, n);
    return check_launch("blend");
}

int b200cs_hydrostatic_balance(const void* dz, const void* phis, const void* temperature, const void* sphum,
                               const void* delp, void* dz_out, void* phis_out, int dtype, int64_t nb, int km,
                               int64_t ncol, double toa_pressure, double gravity, double rdgas, double rvgas,
                               void* stream) {
    B200CS_REQUIRE(dz && phis && temperature && sphum && delp && dz_out && phis_out, "hydrostatic_balance: NULL pointer");
    B200CS_REQUIRE(nb >= 0 && km >= 1 && ncol >= 0, "hydrostatic_balance: bad extents");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "hydrostatic_balance: bad dtype %d", dtype);
    if (nb * ncol == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    const double zvir = rvgas / rdgas - 1.0;
    if (dtype == B200CS_F32)
        hydrostatic_kernel<float><<<blocks_for(nb * ncol), 256, 0, st>>>(
            (const float*)dz, (const float*)phis, (const float*)temperature, (const float*)sphum, (const float*)delp,
            (float*)dz_out, (float*)phis_out, nb, km, ncol, (float)toa_pressure, (float)gravity, (float)rdgas, (float)zvir);
    else
        hydrostatic_kernel<double><<<blocks_for(nb * ncol), 256, 0, st>>>(
            (const double*)dz, (const double*)phis, (const double*)temperature, (const double*)sphum, (const double*)delp,
            (double*)dz_out, (double*)phis_out, nb, km, ncol, toa_pressure, gravity, rdgas, zvir);
    return check_launch("hydrostatic_balance");
}

int b200cs_pressure_at_midpoint_log(const void* delp, void* out, int dtype, int64_t nb, int km, int64_t ncol,
                                    double toa_pressure, void* stream) {
    B200CS_REQUIRE(delp && out, "pressure_at_midpoint_log: NULL pointer");
    B200CS_REQUIRE(nb >= 0 && km >= 1 && ncol >= 0, "pressure_at_midpoint_log: bad extents");
    B200CS_REQUIRE(dtype == B200CS_F32 || dtype == B200CS_F64, "pressure_at_midpoint_log: bad dtype %d", dtype);
    if (nb * ncol == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (dtype == B200CS_F32)
        midpoint_log_kernel<float><<<blocks_for(nb * ncol), 256, 0, st>>>((const float*)delp, (float*)out, nb, km, ncol,
                                                                         (float)toa_pressure);
    else
        midpoint_log_kernel<double><<<blocks_for(nb * ncol), 256, 0, st>>>((const double*)delp, (double*)out, nb, km, ncol,
                                                                          toa_pressure);
    return check_launch("pressure_at_midpoint_log");
}

int b200cs_exposed_area_weights(const void* pfull_coarse, const void* ps_fine, int p_dtype, const void* area,
                                int a_dtype, void* out, int64_t nb, int km, int64_t ny, int64_t nx, int factor,
                                void* stream) {
    B200CS_REQUIRE(pfull_coarse && ps_fine && area && out, "exposed_area_weights: NULL pointer");
    B200CS_REQUIRE(factor >= 1 && ny % factor == 0 && nx % factor == 0, "exposed_area_weights: factor %d must divide the extents", factor);
    B200CS_REQUIRE((p_dtype == B200CS_F32 || p_dtype == B200CS_F64) && (a_dtype == B200CS_F32 || a_dtype == B200CS_F64),
                   "exposed_area_weights: bad dtype");
    const int64_t n = nb * km * ny * nx;
    if (n <= 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    const unsigned g = blocks_for(n);
    if (p_dtype == B200CS_F32 && a_dtype == B200CS_F32)
        exposed_area_kernel<float, float><<<g, 256, 0, st>>>((const float*)pfull_coarse, (const float*)ps_fine, (const float*)area, (float*)out, nb, km, ny, nx, factor);
    else if (p_dtype == B200CS_F64 && a_dtype == B200CS_F32)
        exposed_area_kernel<double, float><<<g, 256, 0, st>>>((const double*)pfull_coarse, (const double*)ps_fine, (const float*)area, (float*)out, nb, km, ny, nx, factor);
    else if (p_dtype == B200CS_F32 && a_dtype == B200CS_F64)
        exposed_area_kernel<float, double><<<g, 256, 0, st>>>((const float*)pfull_coarse, (const float*)ps_fine, (const double*)area, (double*)out, nb, km, ny, nx, factor);
    else
        exposed_area_kernel<double, double><<<g, 256, 0, st>>>((const double*)pfull_coarse, (const double*)ps_fine, (const double*)area, (double*)out, nb, km, ny, nx, factor);
    return check_launch("exposed_area_weights");
}

}  // extern "C"
