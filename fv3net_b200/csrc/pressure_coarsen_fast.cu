// pressure_coarsen_fast.cu -- instantiations of the specialised FAST fused kernel
// (pressure_coarsen_fast.cuh).  Built with FMA contraction on (-fmad=true).
#include "pressure_coarsen_fast.cuh"

namespace b200cs {
bool pc_fast_covers(const PcArgs& a, int field_dtype, int p_dtype, int w_dtype) {
    return pc_fast_eligible(a, field_dtype, p_dtype, w_dtype);
}
int pc_fast_entry(const PcArgs& a) {
    if (a.fy == 8) return launch_pc_fast_fields<8>(a);
    return launch_pc_fast_fields<1>(a);
}
}  // namespace b200cs
