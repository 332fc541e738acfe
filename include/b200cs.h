/*
 * b200cs.h -- C ABI of libb200cs.so: cubed-sphere coarse-graining for NVIDIA B200 (sm_100a).
 *
 * Drop-in boundary for ONE path of VulcanClimateModeling/fv3net: vcm.cubedsphere's block
 * averages and the mappm piecewise-parabolic vertical remap.  Paths below are relative to
 * the reference tree; VCM = external/vcm/vcm, MAPPM = external/mappm/mappm.
 *
 * Conventions:
 *   - every pointer is a DEVICE pointer on the current CUDA device, caller-allocated;
 *     the library never allocates, frees or retains memory;
 *   - arrays are addressed [b0, b1, y, x] with explicit ELEMENT strides for b0, b1, y;
 *     the x stride is always 1 (x is the contiguous axis of FV3 restart files,
 *     [tile, z, y, x]); a stride of 0 broadcasts (2-D weights over levels);
 *   - every export is asynchronous on `stream` (a cudaStream_t passed as void*; NULL =
 *     legacy default stream), re-entrant, and returns 0 or a negative status;
 *     b200cs_last_error() returns a thread-local message for the last failure;
 *   - dtypes are B200CS_F32 / B200CS_F64.  Arithmetic dtypes follow numpy promotion of the
 *     reference expressions: products in promote(obj, weights), the weight sum in the
 *     weights' own dtype, mappm always fp32 (the f2py build casts to `real`);
 *   - sums follow numpy's order for `a.reshape(.., cy, fy, cx, fx).sum((-3,-1))` (pairwise
 *     along x inside a window row, rows added sequentially) so block sums are
 *     bit-identical to the reference's on the same inputs; NaNs are skipped exactly as
 *     xarray's coarsen().sum() does (numerator: NaN of obj*w; denominator: NaN of w).
 */
#ifndef B200CS_H
#define B200CS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200CS_VERSION 100 /* 0.1.0 */
#define B200CS_MAX_FIELDS 16

enum { B200CS_F32 = 0, B200CS_F64 = 1 };

enum {
    B200CS_OK = 0,
    B200CS_EINVAL = -1,       /* bad argument (shape, stride, factor, NULL pointer) */
    B200CS_EUNSUPPORTED = -2, /* valid in the reference but not implemented on device */
    B200CS_ECUDA = -3,        /* CUDA runtime error; text in b200cs_last_error() */
    B200CS_EWORKSPACE = -4    /* workspace missing or too small */
};

/* block_reduce operators: the callables the reference itself passes to
 * horizontal_block_reduce (VCM/cubedsphere/coarsen.py:581-588,743-747) and the method
 * names of block_coarsen (:795-840). */
enum {
    B200CS_OP_SUM = 0,
    B200CS_OP_MEAN = 1,
    B200CS_OP_MIN = 2,
    B200CS_OP_MAX = 3,
    B200CS_OP_MEDIAN = 4,
    B200CS_OP_MODE = 5
};

/* pressure_coarsen weight masking (VCM/cubedsphere/regridz.py:200-220 and
 * workflows/fine_res_budget/budget/budgets.py:49-57) */
enum {
    B200CS_MASK_NONE = 0,  /* fine_res_budget: plain area weights */
    B200CS_MASK_PHALF = 1, /* extrapolate=False: coarse bottom interface < fine surface */
    B200CS_MASK_PFULL = 2  /* extrapolate=True : coarse log-midpoint      < fine surface */
};

/* arithmetic mode of the fused remap */
enum {
    B200CS_EXACT = 0, /* no FMA contraction, IEEE division: bit-identical to the fp32 no-FMA
                         restatement of mappm.f90 (oracle/mappm_ref.c) */
    B200CS_FAST = 1,  /* hoisted-geometry arithmetic with FMA and rcp.approx; within 1e-6 of the
                         field scale (tests/test_gpu_parity.py) */
    /* flag, OR-ed into mode: the fx window elements were NOT contiguous in the reference's array
       (D-grid v lines arrive as gathered columns), so numpy adds them one after the other instead
       of pairwise; only matters for bit-exactness in B200CS_EXACT */
    B200CS_SEQ_WINDOW = 2
};

int b200cs_version(void);
const char* b200cs_last_error(void);

/*
 * Weighted block average.  Replaces weighted_block_average (VCM/cubedsphere/coarsen.py:183-218)
 * and edge_weighted_block_average (:221-273).
 *   out[b0,b1,j,i] = nansum_{window}(obj*W) / nansum_{window}(W),   W = w  or  w2*w
 * window (j,i) = rows [j*ystep, j*ystep+fy) x cols [i*xstep, i*xstep+fx):
 *   block average      fy=fx=ystep=xstep=f
 *   edge='x' (u, dx)   fy=1, ystep=f, fx=xstep=f  (only every f-th y line is read at all)
 *   edge='y' (v, dy)   fy=ystep=f, fx=1, xstep=f
 * obj [nb0,nb1,ny,nx]; w same shape, w_stride[1]==0 for 2-D weights; w2 optional (NULL) second
 * factor, e.g. mass = delp*area formed on the fly (VCM/cubedsphere/coarsen_restarts.py:400).
 * out: contiguous [nb0,nb1,cy,cx], cy=(ny-fy)/ystep+1, cx=(nx-fx)/xstep+1, dtype
 * promote(obj,w,w2) which the caller states in out_dtype (checked).
 * den_out (optional, may be NULL): the weight sums, contiguous [nb0,(nb1 or 1),cy,cx] in the
 * weights' dtype.
 */
int b200cs_weighted_block_average(const void* obj, int obj_dtype, const int64_t obj_stride[3],
                                  const void* w, int w_dtype, const int64_t w_stride[3],
                                  const void* w2, int w2_dtype, const int64_t w2_stride[3],
                                  void* out, int out_dtype, int64_t nb0, int64_t nb1, int64_t ny,
                                  int64_t nx, int fy, int fx, int ystep, int xstep, void* stream);

/*
 * Generic block reduction with trailing-edge padding.  Replaces horizontal_block_reduce /
 * xarray_block_reduce (VCM/cubedsphere/coarsen.py:393-554) -> block_reduce
 * (VCM/cubedsphere/_skimage.py:125-202) for the callables the reference passes, and
 * block_coarsen / block_edge_coarsen (coarsen.py:629-683,795-840).
 * obj [nb, ny, nx] (stride for nb and y) -> out contiguous [nb, cy, cx] of the same dtype,
 * cy = ceil(ny/fy) when padding (pad != 0) else windows as above with ystep/xstep.
 * skipna: NaN-skipping form (np.nansum / xarray coarsen defaults, scipy mode nan_policy='omit').
 */
int b200cs_block_reduce(const void* obj, int dtype, const int64_t obj_stride[2], void* out,
                        int64_t nb, int64_t ny, int64_t nx, int fy, int fx, int ystep, int xstep,
                        int op, int skipna, int pad, double cval, void* stream);

/*
 * Block upsample.  Replaces block_upsample / block_upsample_like
 * (VCM/cubedsphere/coarsen.py:843-938, VCM/xarray_utils.py:37-82): each element repeated f
 * times along y and x; a staggered (interface) axis repeats all but its last element.
 * in contiguous [nb, cy, cx] -> out contiguous [nb, ny, nx]; elem_size 1, 2, 4 or 8 bytes (bits are
 * moved, never converted: a bool mask stays a bool mask).
 */
int b200cs_block_upsample(const void* in, void* out, int elem_size, int64_t nb, int64_t cy,
                          int64_t cx, int fy, int fx, int y_staggered, int x_staggered,
                          void* stream);

/*
 * Centre -> edge stagger of a 6-tile field across cube faces.  Replaces xgcm
 * Grid.interp(delp, axis) with FV3_FACE_CONNECTIONS (VCM/cubedsphere/xgcm.py:7-34; call sites
 * VCM/cubedsphere/regridz.py:118-132, VCM/cubedsphere/coarsen_restarts.py:825-853).
 * in contiguous [6, nb1, n, n]; axis 0 = y (out [6,nb1,n+1,n]), 1 = x (out [6,nb1,n,n+1]).
 * line_step > 1 writes only lines 0, step, 2*step, .. of the staggered axis, compactly:
 * axis 0 -> out [6,nb1,n/step+1,n]; axis 1 -> out TRANSPOSED [6,nb1,n/step+1,n] (line-major)
 * so that the lines the D-grid pressure path needs are contiguous.
 */
int b200cs_stagger(const void* in, void* out, int dtype, int64_t nb1, int64_t n, int axis,
                   int line_step, void* stream);

/*
 * Gathers every step-th x column of in[nb, ny, nx] into out[nb, (nx-1)/step+1, ny]
 * (column-major lines); the D-grid v / dy inputs of the pressure path.  The inverse
 * (scatter of coarse lines back to [.., cy, cx]) is done with transpose=1 on the small output.
 */
int b200cs_gather_columns(const void* in, const int64_t in_stride[2], void* out, int elem_size,
                          int64_t nb, int64_t ny, int64_t nx, int step, void* stream);
int b200cs_transpose_last2(const void* in, void* out, int elem_size, int64_t nb, int64_t ny,
                           int64_t nx, void* stream);

/*
 * Interface pressure.  Replaces pressure_at_interface
 * (VCM/calc/thermo/vertically_dependent.py:41-66): out[b,0]=toa, out[b,k+1]=out[b,k]+delp[b,k]
 * (NaN counted as 0), sequential in the input dtype.  delp [nb0, km, ncol] with strides for
 * nb0 and level (column stride 1) -> out contiguous [nb0, km+1, ncol].
 */
int b200cs_pressure_at_interface(const void* delp, int dtype, const int64_t stride[2], void* out,
                                 int64_t nb0, int64_t km, int64_t ncol, double toa_pressure,
                                 void* stream);

/*
 * mappm column remap.  Replaces the f2py extension `mappm.mappm(pe1, q1, pe2, i1, i2, iv, kord,
 * ptop)` (MAPPM/mappm.f90:10-126; call site VCM/cubedsphere/regridz.py:329-334).  fp32 like the
 * Fortran build.  Layout-agnostic: element (column c, level k) of an array lives at
 * ptr[c*col_stride + k*lev_stride], so the reference's [ncol, nz] (col_stride=nz, lev_stride=1)
 * and the restart layout [z, ncol] (col_stride=1, lev_stride=ncol) both work without a
 * transpose.  All kord of the Fortran are supported (ppm_profile kord<=7, cs_profile 8..16+).
 * workspace: b200cs_mappm_workspace_bytes(ncol, km) bytes of device memory.
 * Defined where the Fortran is undefined: a failed level search (non-monotone pressures) yields
 * NaN for that and all later layers of the column; km < 4 (or < 5 for kord > 7) -> EINVAL.
 */
int64_t b200cs_mappm_workspace_bytes(int64_t ncol, int64_t km);
int b200cs_mappm(const float* pe1, int64_t pe1_cs, int64_t pe1_ls, const float* q1, int64_t q1_cs,
                 int64_t q1_ls, const float* pe2, int64_t pe2_cs, int64_t pe2_ls, float* q2,
                 int64_t q2_cs, int64_t q2_ls, int64_t ncol, int km, int kn, int iv, int kord,
                 void* workspace, int64_t workspace_bytes, void* stream);

/*
 * b200cs_mappm without a caller-provided workspace and WITHOUT ANY ALLOCATION.  No counterpart in
 * the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, kernels and results as b200cs_mappm; the workspace of the kord > 6 machine is a
 * file-scope device array of B200CS_MAPPM_STATIC_MAX_BYTES that every GPU that loads the library
 * holds until exit.  Limit: b200cs_mappm_workspace_bytes(ncol, km) <= B200CS_MAPPM_STATIC_MAX_BYTES
 * (6553 columns of km = 79 per call), else B200CS_EWORKSPACE -- split the columns over calls or
 * use b200cs_mappm.  kord <= 6 needs no workspace and has no limit.  All calls on one device
 * share the array: calls that overlap on the GPU (different streams) race -- serialise them.
 * The array keeps the previous call's contents; the kernel writes every element before it reads
 * it, so nothing needs clearing.
 */
#define B200CS_MAPPM_STATIC_MAX_BYTES (16 * 1024 * 1024)
int b200cs_mappm_static(const float* pe1, int64_t pe1_cs, int64_t pe1_ls, const float* q1, int64_t q1_cs,
                        int64_t q1_ls, const float* pe2, int64_t pe2_cs, int64_t pe2_ls, float* q2,
                        int64_t q2_cs, int64_t q2_ls, int64_t ncol, int km, int kn, int iv, int kord,
                        void* stream);

/*
 * mappm in B200CS_FAST arithmetic: the level search and the parabolas of MAPPM/mappm.f90:10-126,
 * 614-931 with the pressure geometry hoisted, rcp.approx for its reciprocals and FMA contraction
 * -- within 1e-6 relative of b200cs_mappm.  Only the lmt=0 profiles (kord <= 3; every reference
 * caller passes kord=1, VCM/cubedsphere/regridz.py:227-228); pe2 must be strictly increasing.
 * Other kord -> EINVAL.  No workspace.
 * Batched: element (slab b, column c, level k) of an array lives at
 * ptr[b*s[0] + c*s[1] + k*s[2]] (element strides), so the restart layout [tile, z, y, x] is ONE
 * launch with s = {nz*ny*nx, 1, ny*nx} and the reference's [ncol, nz] is s = {0, nz, 1}, nbatch 1.
 */
int b200cs_mappm_fast(const float* pe1, const int64_t pe1_stride[3], const float* q1,
                      const int64_t q1_stride[3], const float* pe2, const int64_t pe2_stride[3], float* q2,
                      const int64_t q2_stride[3], int64_t nbatch, int64_t ncol, int km, int kn, int iv,
                      int kord, void* stream);

/*
 * Fused pressure-level coarse-graining.  Replaces, in ONE pass over the fine data,
 *   regrid_to_area_weighted_pressure (VCM/cubedsphere/regridz.py:31-78) =
 *     weighted_block_average(delp, w) -> block_upsample_like -> pressure_at_interface x2 ->
 *     regrid_vertical per variable (Python loop :180-183 -> mappm) -> _mask_weights (:200-220)
 *   followed by weighted_block_average(regridded, masked weights)
 *     (VCM/cubedsphere/coarsen_restarts.py:528-534, 945-959),
 * and Grid.pressure_level_average of workflows/fine_res_budget/budget/budgets.py:49-57
 * (mask_mode NONE).  With fy=1 and inputs holding only the kept lines it is also the body of
 * regrid_to_edge_weighted_pressure + edge_weighted_block_average for D-grid winds
 * (regridz.py:81-146).
 *
 *   fields[i]  [nb0, km, ny, nx]  strides (b0, level, y); dtype field_dtype (cast to fp32
 *              exactly as f2py does)
 *   delp       [nb0, km, ny, nx]  dtype p_dtype; interface pressures are its sequential cumsum
 *   w          [nb0, ny, nx]      2-D weights (area or edge length), dtype w_dtype
 *   out[i]     contiguous [nb0, km, cy, cx] dtype promote(f32, w_dtype), cy=ny/fy, cx=nx/fx
 *   delp_coarse (optional) contiguous [nb0, km, cy, cx] dtype promote(p_dtype, w_dtype):
 *              the block average of delp that defines the coarse pressure levels
 * fy x fx is the window (fx in {2,4,8}; fy == fx, or fy == 1 for edge lines); ny % fy == 0 and
 * nx % fx == 0.  kord <= 6 (every reference caller uses kord=1, iv=1); km >= 4.
 * mode: B200CS_EXACT or B200CS_FAST.
 */
int b200cs_pressure_coarsen(int nfields, const void* const* fields, int field_dtype,
                            const int64_t field_stride[3], const void* delp, int p_dtype,
                            const int64_t delp_stride[3], const void* w, int w_dtype,
                            const int64_t w_stride[2], void* const* out, void* delp_coarse,
                            int64_t nb0, int km, int64_t ny, int64_t nx, int fy, int fx,
                            double toa_pressure, int mask_mode, int iv, int kord, int mode,
                            void* stream);

/*
 * Fused compute + all-gather.  The same kernel as b200cs_pressure_coarsen (FAST mode, fp32, 8 x 8
 * windows) additionally stores every coarse value into `npeers` (1..8) peer-mapped buffers as it
 * is produced -- plain stores over NVLink peer memory, no separate collective and no SMs taken
 * from the coarsening.  peer_out[p] points at THIS rank's slot inside rank p's gathered buffer;
 * a slot is contiguous [nfields + 1][nb0, km, cy, cx] fp32 with the coarse delp last.  The caller
 * maps the buffers (CUDA IPC / cuMem fabric handles; torch symmetric memory in the Python layer)
 * and orders readers after the writers (a barrier after the kernel).  Replaces the NCCL
 * all-gather of the coarse outputs of SURVEY.md 8e.  The copies are made by the last CTA of each
 * row of tiles (complete coarse rows, 16-byte stores); `workspace` holds the per-row arrival
 * counters (b200cs_pressure_coarsen_gather_workspace_bytes(nb0, ny) bytes, zeroed by the call).
 */
int64_t b200cs_pressure_coarsen_gather_workspace_bytes(int64_t nb0, int64_t ny);
int b200cs_pressure_coarsen_gather(int nfields, const void* const* fields, int field_dtype,
                                   const int64_t field_stride[3], const void* delp, int p_dtype,
                                   const int64_t delp_stride[3], const void* w, int w_dtype,
                                   const int64_t w_stride[2], void* const* out, void* delp_coarse,
                                   int64_t nb0, int km, int64_t ny, int64_t nx, int fy, int fx,
                                   double toa_pressure, int mask_mode, int iv, int kord, int mode,
                                   int npeers, void* const* peer_out, void* workspace,
                                   int64_t workspace_bytes, void* stream);

/*
 * Wide windows: the fused pressure-level coarse-graining for factor 16, 32 or 64 (C3072 -> C48 is
 * factor 64).  fp32 data, FAST arithmetic, kord <= 3, square windows.  A coarse cell now spans
 * several CTAs: (1) b200cs_weighted_block_average forms the coarse delp (numpy order, bit-exact),
 * (2) the sweep kernel remaps / masks / weights every fine column against it and leaves per-CTA
 * partial sums, (3) a finalize kernel adds the partial sums of a cell in a fixed order and divides.
 * Deterministic (no atomics); values within 1e-6 of the unfused EXACT operators.  Arguments as
 * b200cs_pressure_coarsen; delp_coarse may be NULL.
 */
int64_t b200cs_pressure_coarsen_wide_workspace_bytes(int64_t nb0, int km, int64_t ny, int64_t nx,
                                                     int factor);
int b200cs_pressure_coarsen_wide(int nfields, const void* const* fields,
                                 const int64_t field_stride[3], const void* delp,
                                 const int64_t delp_stride[3], const void* w,
                                 const int64_t w_stride[2], void* const* out, void* delp_coarse,
                                 int64_t nb0, int km, int64_t ny, int64_t nx, int factor,
                                 double toa_pressure, int mask_mode, int iv, int kord,
                                 void* workspace, int64_t workspace_bytes, void* stream);

/*
 * Below-surface weight mask for the unfused path and API parity with _mask_weights
 * (VCM/cubedsphere/regridz.py:200-220): out[b,k,col] = cond ? w[b,col] : 0 with
 *   PHALF: phalf_coarse_on_fine[b,k+1,col] < phalf_fine[b,km,col]
 *   PFULL: pfull_coarse_on_fine[b,k,col]   < phalf_fine[b,km,col]
 * p_coarse is the array compared (phalf [nb0,km+1,ncol] or pfull [nb0,km,ncol]), contiguous;
 * phalf_fine contiguous [nb0,km+1,ncol]; w [nb0,ncol]; out [nb0,km,ncol] in w's dtype.
 */
int b200cs_mask_weights(const void* w, int w_dtype, const void* p_coarse, const void* phalf_fine,
                        int p_dtype, void* out, int64_t nb0, int64_t km, int64_t ncol,
                        int mask_mode, void* stream);

/*
 * Column-wise post-processing of the coarse state.  Arrays are contiguous
 * [nb, km, ncol] (column index fastest) unless strides are given; one dtype per call.
 *
 * b200cs_surface_pressure: ps[nb, ncol] = nansum_k delp + toa (surface_pressure_from_delp,
 *   VCM/calc/thermo/vertically_dependent.py:189-208); delp strided (batch, level).
 * b200cs_blending_weights: compute_blending_weights with pfull = pressure_at_midpoint_log(delp_c),
 *   ps = surface pressure of delp_c and blending pressure sigma_blend * ps_min, ps_min[nb, ncol]
 *   being the block minimum of the fine surface pressure
 *   (VCM/cubedsphere/coarsen_restarts.py:559-648).
 * b200cs_blend: w * pressure_level + (1 - w) * model_level (coarsen_restarts.py:651-667).
 * b200cs_hydrostatic_balance: _impose_hydrostatic_balance (coarsen_restarts.py:990-1017):
 *   DZ from T, sphum, delp in hydrostatic balance (hydrostatic_dz, vertically_dependent.py:210-235)
 *   and phis adjusted to keep the model-top height of (dz, phis)
 *   (height_at_interface :69-98, dz_and_top_to_phis :182-186); phis [nb, ncol].
 */
/* pressure_at_midpoint_log (VCM/calc/thermo/vertically_dependent.py:153-179) of contiguous
 * [nb, km, ncol]; and the fine-size half of Grid.area_above_fine_surface
 * (workflows/fine_res_budget/budget/budgets.py:62-77): out[nb, km, ny, nx] =
 * pfull_coarse[nb, km, ny/f, nx/f] <= ps_fine[nb, ny, nx] ? area[nb, ny, nx] : 0 (block-summed by
 * b200cs_block_reduce afterwards). */
int b200cs_pressure_at_midpoint_log(const void* delp, void* out, int dtype, int64_t nb, int km,
                                    int64_t ncol, double toa_pressure, void* stream);
int b200cs_exposed_area_weights(const void* pfull_coarse, const void* ps_fine, int p_dtype,
                                const void* area, int a_dtype, void* out, int64_t nb, int km,
                                int64_t ny, int64_t nx, int factor, void* stream);
int b200cs_surface_pressure(const void* delp, int dtype, const int64_t stride[2], void* ps,
                            int64_t nb, int km, int64_t ncol, double toa_pressure, void* stream);
int b200cs_blending_weights(const void* delp_coarse, const void* ps_min, void* out, int dtype,
                            int64_t nb, int km, int64_t ncol, double toa_pressure,
                            double sigma_blend, void* stream);
int b200cs_blend(const void* weights, const void* pressure_level, const void* model_level,
                 void* out, int dtype, int64_t n, void* stream);
int b200cs_hydrostatic_balance(const void* dz, const void* phis, const void* temperature,
                               const void* sphum, const void* delp, void* dz_out, void* phis_out,
                               int dtype, int64_t nb, int km, int64_t ncol, double toa_pressure,
                               double gravity, double rdgas, double rvgas, void* stream);

/*
 * Masks and selections of the 'complex' sfc_data method (SURVEY.md 8f N1;
 * VCM/cubedsphere/coarsen_restarts.py:1162-1470).  Masks are one byte per element (0/1).
 *   b200cs_elementwise   MUL a*b | WHERE mask?a:scalar (`.where(mask)` with scalar = NaN) |
 *                        WHERE_NOT mask?scalar:a (`xr.where(mask, scalar, a)`) | SELECT mask?a:b
 *                        (`xr.where`) | FILLNA | MIN_SCALAR a<scalar?a:scalar
 *                        (`a.where(a < scalar, other=scalar)`)
 *   b200cs_compare_scalar  numpy.isclose(a, scalar) (VCM/xarray_utils.py:7-35) / a < scalar /
 *                        a > scalar, optionally AND-ed with a mask
 *   b200cs_dominant_mask isclose(fine, block_upsample_like(coarse)) without the upsampled
 *                        temporary (coarsen_restarts.py:1173-1190), optionally AND-ed with a mask;
 *                        fine [nb, ny, nx], coarse [nb, ny/f, nx/f]
 */
enum {
    B200CS_EW_MUL = 0,
    B200CS_EW_WHERE = 1,
    B200CS_EW_WHERE_NOT = 2,
    B200CS_EW_SELECT = 3,
    B200CS_EW_FILLNA = 4,
    B200CS_EW_MIN_SCALAR = 5,
    B200CS_EW_DIFF_OVER = 6 /* (a - b) / scalar: budget storage terms, budgets.py:80-85 */
};
enum { B200CS_CMP_ISCLOSE = 0, B200CS_CMP_LT = 1, B200CS_CMP_GT = 2 };
int b200cs_elementwise(int op, const void* a, const void* b, const uint8_t* mask, void* out,
                       int dtype, int64_t n, double scalar, void* stream);
int b200cs_compare_scalar(int cmp, const void* a, int dtype, double scalar, double rtol,
                          double atol, const uint8_t* and_mask, uint8_t* out, int64_t n,
                          void* stream);
int b200cs_dominant_mask(const void* fine, const void* coarse, int dtype, int64_t nb, int64_t ny,
                         int64_t nx, int factor, double rtol, double atol,
                         const uint8_t* and_mask, uint8_t* out, void* stream);

/*
 * Host-side staging for the drop-in entry points, which receive plain (pageable) numpy arrays
 * (VCM/cubedsphere/regridz.py:31-78, coarsen.py:183-218): copies `bytes` from src to dst with up to
 * `nthreads` host threads.  dst is a pinned staging buffer the caller then hands to
 * cudaMemcpyAsync; the copy of slab i+1 overlaps the DMA of slab i (fv3net_b200/core.py).
 * No CUDA call is made.
 */
int b200cs_host_copy(void* dst, const void* src, int64_t bytes, int nthreads);

#ifdef __cplusplus
}
#endif
#endif /* B200CS_H */
