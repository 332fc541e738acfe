// pressure_coarsen_fast.cu -- instantiations of the specialised FAST fused kernel
// (pressure_coarsen_fast.cuh).  Built with FMA contraction on (-fmad=true).
#include "pressure_coarsen_fast.cuh"
#include "pressure_coarsen_sweep.cuh"
#include "pressure_coarsen_tma.cuh"
#ifdef B200CS_WITH_WS  // experimental warp-specialised kernel: developer builds only
#include <stdlib.h>

#include "experimental/pressure_coarsen_ws.cuh"
#endif

namespace b200cs {
bool pc_fast_covers(const PcArgs& a, int field_dtype, int p_dtype, int w_dtype) {
    return pc_fast_eligible(a, field_dtype, p_dtype, w_dtype);
}
int pc_fast_entry(const PcArgs& a) {
#ifdef B200CS_WITH_WS
    static const int ws = [] { const char* e = getenv("B200CS_DEV_WS"); return e ? atoi(e) : 0; }();
    if (ws && pc_tma_eligible(a)) {
        const int rc = launch_pc_ws_fields(a);
        if (rc != -1) return rc;
    }
#endif
    if (pc_tma_eligible(a)) {
        // TMA-fed kernel; -1: the arrays cannot be described to TMA (alignment, strides) -> sweep kernel
        const int rc = launch_pc_tma_fields(a);
        if (rc != -1) return rc;
    }
    if (a.fy == 8) return launch_pc_sweep_fields<8>(a);
    return launch_pc_sweep_fields<1>(a);
}
// wide windows (f = 16, 32, 64): block-average pre-pass for the coarse delp, sweep kernel with
// per-CTA partial sums, fixed-order finalize (pressure_coarsen_sweep.cuh)
int64_t pc_wide_workspace_floats(int64_t nb0, int km, int64_t ny, int64_t nx, int f) {
    return nb0 * km * (ny / f) * (nx / f) + pc_wide_partial_floats(nb0, km, ny, nx, f);
}
int pc_wide_entry(const PcArgs& a, float* ws) {
    const int f = a.fx;
    float* delp_c = a.delp_coarse ? (float*)a.delp_coarse : ws;
    float* partial = ws + a.nb0 * a.km * (a.ny / f) * (a.nx / f);
    const int64_t ds[3] = {a.dstride[0], a.dstride[1], a.dstride[2]};
    const int64_t wst[3] = {a.wstride[0], 0, a.wstride[1]};
    // coarse delp: the streaming wide-window block average of block_ops.cu (numpy order, bit-exact)
    int rc = b200cs_weighted_block_average(a.delp, B200CS_F32, ds, a.w, B200CS_F32, wst, nullptr, 0, nullptr, delp_c,
                                           B200CS_F32, a.nb0, a.km, a.ny, a.nx, f, f, f, f, (void*)a.st);
    if (rc != B200CS_OK) return rc;
    return launch_pc_sweep_wide_fields(a, delp_c, partial);
}
}  // namespace b200cs
