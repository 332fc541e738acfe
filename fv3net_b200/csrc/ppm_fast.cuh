// ppm_fast.cuh -- the B200CS_FAST arithmetic of the PPM column remap: what is re-derived and why.
// (The column machines that implement it are ppm_sweep.cuh and ppm_pack.cuh; the first one,
// PpmFast, output-layer-major, was retired in round 2 together with its kernel.)
//
// Same algorithm as the reference's mappm + ppm_profile + ppm_limiters
// (external/mappm/mappm/mappm.f90:10-126, 614-931) for the profile every reference caller uses:
// kord <= 3 (limiter lmt = 0 on every layer; regridz.py:227-228 passes kord=1), any iv.
// ppm_stream.cuh is the bit-exact statement of that algorithm; this file re-derives it for the
// SM's issue rate (hard part 1):
//   * every quantity that depends only on the pressure geometry is formed once per source layer
//     and shared by the NF fields: the slope factors A,B (mappm.f90:658-662 with the divisions
//     hoisted), the edge factors G1,G2,G3 (:674-683 expanded in d01, dc(k-1), dc(k)), 1/dp for the
//     sub-layer coordinates;
//   * the integrals of :76-112 are taken in "mean + deviation" form.  With s the position of an
//     output edge inside source layer L (0..1) and D(s) = dp*s*(1-s)*[(A6/3)(1+s) - B],
//     B = (A6+AR-AL)/2, the mass of layer L above the edge is q*dp*s + D(s) (D(0) = D(1) = 0).
//     Every output edge is evaluated ONCE (it closes one output layer and opens the next: the
//     running sum R keeps -D for the layer it opens), whole and partial layers enter through
//     q*dP, and no difference of large numbers is formed.  A field costs 5 flops per output
//     layer plus one FMA per source layer, and only (q, A6/3, B) are kept per layer;
//   * divisions by geometry become MUFU.RCP (rcp.approx, <= 1 ulp); no per-field division is left;
//   * max3(q)-q and q-min3(q) of :664-666 are evaluated on the differences, max3(0,-d01,d12) and
//     max3(0,d01,-d12): identical values (rounding is monotone), two subtractions fewer;
//   * ONE call site of advance() inside the per-output-layer loop: lanes of a warp that sit on
//     different source layers stay converged (the exact machine has three);
//   * level data arrives through the accessor's fetch(), which the fused kernel backs with a
//     per-thread cp.async ring in shared memory several source layers deep (a register prefetch
//     ends up as a phi copy that waits for the load inside the same iteration).
// Every level-search and mask decision is made on the same fp32 pressures as the reference
// (sequential cumsum), so which source layers feed which output layer is identical; only the
// rounding of the parabola coefficients and integrals differs (a few ulp; tests state 1e-6).
//
// Accessor concept (stateful; levels are requested in increasing order, one at a time):
//   void  fetch(int jq, int jpe)   make q[.][min(jq,km-1)] and the datum that defines edge
//                                  pe[min(jpe,km)] current
//   float q(int f)                 the current layer mean of field f
//   float pe(float prev)           the current edge pressure given the previous edge
//                                  (prev + delp for a cumsum source, or the stored edge itself)
//   float pe0()                    edge 0
//   void  rewind()                 the next fetch() returns the current level once more
#pragma once
#include "ppm_stream.cuh"

namespace b200cs {

B200CS_HD float fast_rcp(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

}  // namespace b200cs
