// pressure_coarsen_inst.cu -- one translation unit per (window width, arithmetic mode) of the
// fused pressure-coarsening kernel, so the instantiations build in parallel and the EXACT
// variants are compiled with -fmad=false while the FAST variants allow FMA contraction.
// Build: nvcc -DB200CS_FX={2,4,8} -DB200CS_MODE={0,1} [-fmad=false|true] ...
#include "pressure_coarsen.cuh"

#ifndef B200CS_FX
#error "define B200CS_FX (2, 4 or 8)"
#endif
#ifndef B200CS_MODE
#error "define B200CS_MODE (0 exact, 1 fast)"
#endif

#define B200CS_CAT2(a, b, c, d) a##b##c##d
#define B200CS_CAT(a, b, c, d) B200CS_CAT2(a, b, c, d)

namespace b200cs {
int B200CS_CAT(pc_entry_fx, B200CS_FX, _m, B200CS_MODE)(const PcArgs& a, int p_dtype, int w_dtype) {
    return launch_pc_types<B200CS_MODE, B200CS_FX>(a, p_dtype, w_dtype);
}
}  // namespace b200cs
