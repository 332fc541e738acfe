// pressure_coarsen_fast.cu -- instantiations of the specialised FAST fused kernel
// (pressure_coarsen_fast.cuh).  Built with FMA contraction on (-fmad=true).
#include <stdlib.h>

#include "pressure_coarsen_fast.cuh"
#include "pressure_coarsen_sweep.cuh"

namespace b200cs {
bool pc_fast_covers(const PcArgs& a, int field_dtype, int p_dtype, int w_dtype) {
    return pc_fast_eligible(a, field_dtype, p_dtype, w_dtype);
}
// B200CS_FAST_KERNEL=v1 selects the first (output-layer-major) fast kernel, kept for A/B profiling
static bool use_v1() {
    static const bool v = [] { const char* e = getenv("B200CS_FAST_KERNEL"); return e && e[0] == 'v' && e[1] == '1'; }();
    return v;
}
int pc_fast_entry(const PcArgs& a) {
    if (use_v1() && a.npeer == 0) return a.fy == 8 ? launch_pc_fast_fields<8>(a) : launch_pc_fast_fields<1>(a);
    if (a.fy == 8) return launch_pc_sweep_fields<8>(a);
    return launch_pc_sweep_fields<1>(a);
}
}  // namespace b200cs
