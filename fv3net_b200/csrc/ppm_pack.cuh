// ppm_pack.cuh -- layer-synchronous PPM column remap on PAIRS of fields (B200CS_FAST arithmetic,
// third design): the per-thread machine of the TMA-fed fused kernel (pressure_coarsen_tma.cuh).
//
// Same algorithm as ppm_sweep.cuh (the reference's mappm + ppm_profile + ppm_limiters,
// external/mappm/mappm/mappm.f90:10-126, 614-931, kord <= 3: limiter lmt = 0 on every layer) and
// the same organisation (source layer = outer loop, output edges emitted from an inner loop).
// What changes is the instruction stream (profiles/r01_sweep_v8_ncu_summary.txt: 405 warp
// instructions per level for 4 fields, ALU pipe -- FSEL/FSETP/FMNMX, half rate -- the busiest):
//
//   * two fields share every fp32 instruction: sm_100a's packed FFMA2 / FADD2 / FMUL2
//     (fma.rn.f32x2 ...) take a 64-bit register pair per operand and a scalar register for the
//     per-column geometry factors, so the per-field arithmetic of step() and edge() issues once
//     per PAIR;
//   * the lmt=0 limiter (mappm.f90:875-894) is evaluated as two clamps.  With dl = AL-q, dr = AR-q
//     the Fortran's two tests are "dr overshoots -2 dl" / "dl overshoots -2 dr", and the limited
//     deviations are dl' = -s min(|dl|, 2|dr|), dr' = s min(|dr|, 2|dl|), s = sign(dr), whenever
//     dl and dr have opposite signs -- which the limited slopes guarantee for interior layers
//     (Colella & Woodward 1984, eq. 1.6-1.8: the interpolated edge value lies between the two
//     neighbouring means; rounding can violate it by an ulp of q, which moves the result by the
//     same ulp).  2 FMNMX3 + 2 LOP3 per field instead of 3 FSETP + 5 FSEL.  The flat case
//     (dm == 0 -> A6 = 0, AL = AR = q) enters the same min3 as a third operand |dm| * 2^100.
//     The two boundary layers at either end, whose edge values are extrapolated and CAN lie
//     outside the neighbouring means, keep the literal tests of ppm_sweep.cuh -- and so does any
//     layer within two levels of a NaN (`general` argument of step(); the caller watches the
//     incoming level values): the reference's NaN pattern next to a NaN level follows from
//     those tests being false on NaN and from Fortran MIN/MAX dropping NaN operands (a layer
//     whose neighbour is NaN can come out FLAT and finite), which no clamp reproduces;
//   * the parabola is stored as its two edge values and their limited deviations from the mean
//     (AL, AR, dl = AL-q, dr = AR-q), and every piece of an output layer is integrated ANCHORED at
//     the end of the source layer it touches: from the layer top, I(s) = h (AL - s(1+t) dl - s t dr),
//     h = p - pe_lo, s = h/dp, t = 1-s; up to the layer bottom, J(t) = g (AR - t(1+s) dr - s t dl),
//     g = pe_hi - p.  A sliver at either end of a layer whose profile runs into zero there (cloud
//     water next to a cloud-free layer) keeps the relative accuracy of its own small value; the
//     "layer mean minus deviation integral" form of the first two designs cancelled there
//     (tests/test_gpu_parity.py::test_pressure_coarsen_fast_tracer_decades: 4e-3 of the coarse
//     cell's value).  edge() is 2 fma + fma + 2 mul per pair, close() 2 fma + fma.
//
// The functions compile for the host (f2 = two floats) so that tests/harness/ppm_pack_host.cpp
// runs the same machine on the CPU against the oracle.
#pragma once
#include "ppm_sweep.cuh"

namespace b200cs {

#if defined(__CUDA_ARCH__)
struct f2 { unsigned long long v; };
B200CS_HD f2 f2_make(float a, float b) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(a), "f"(b)); return r; }
B200CS_HD float f2_lo(f2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a.v)); return x; }
B200CS_HD float f2_hi(f2 a) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(a.v)); return y; }
B200CS_HD f2 f2_add(f2 a, f2 b) { f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
B200CS_HD f2 f2_sub(f2 a, f2 b) { f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
B200CS_HD f2 f2_mul(f2 a, f2 b) { f2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v)); return r; }
B200CS_HD f2 f2_fma(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v)); return r; }
B200CS_HD float min3abs(float a, float b, float c) { return fminf(fminf(fabsf(a), fabsf(b)), fabsf(c)); }
#else
struct f2 { float x, y; };
B200CS_HD f2 f2_make(float a, float b) { f2 r; r.x = a; r.y = b; return r; }
B200CS_HD float f2_lo(f2 a) { return a.x; }
B200CS_HD float f2_hi(f2 a) { return a.y; }
B200CS_HD f2 f2_add(f2 a, f2 b) { return f2_make(a.x + b.x, a.y + b.y); }
B200CS_HD f2 f2_sub(f2 a, f2 b) { return f2_make(a.x - b.x, a.y - b.y); }
B200CS_HD f2 f2_mul(f2 a, f2 b) { return f2_make(a.x * b.x, a.y * b.y); }
B200CS_HD f2 f2_fma(f2 a, f2 b, f2 c) { return f2_make(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
B200CS_HD float min3abs(float a, float b, float c) { return fminf(fminf(fabsf(a), fabsf(b)), fabsf(c)); }
#endif
B200CS_HD f2 f2_bc(float a) { return f2_make(a, a); }
// a with the sign bit of s xor-ed in (a * sign(s) for s != 0)
B200CS_HD float xor_sign(float a, float s) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__float_as_uint(a) ^ (__float_as_uint(s) & 0x80000000u));
#else
    return signbit(s) ? -a : a;
#endif
}

// a with the inverted sign bit of s xor-ed in (-a * sign(s) for s != 0)
B200CS_HD float xor_nsign(float a, float s) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__float_as_uint(a) ^ (~__float_as_uint(s) & 0x80000000u));
#else
    return signbit(s) ? a : -a;
#endif
}

// ---- per-column geometry, shared by every field of the column (and, in the warp-specialised kernel,
// computed by other threads than the field math: experimental/pressure_coarsen_ws.cuh) ---------------------------
struct PpmStepGeo { float A, B, G1, G2, nG3; };
// slope factors of layer L+1 and edge-value factors of interface L+1 from dp[L-1..L+2]
// (mappm.f90:658-683); r01 = 1/(dp0+dp1) formed by the previous step, rn = 1/(dp1+dp2) for the next
B200CS_HD PpmStepGeo ppm_step_geo(float dpm, float dp0, float dp1, float dp2, float r01, float& rn) {
    const float Pm0 = dpm + dp0, P01 = dp0 + dp1, P12 = dp1 + dp2;
    rn = fast_rcp(P12);
    const float t = dp1 * fast_rcp(P01 + dp2);
    PpmStepGeo o;
    o.A = t * (dp0 + 0.5f * dp1) * rn;
    o.B = t * (dp2 + 0.5f * dp1) * r01;
    const float a1 = Pm0 * fast_rcp(P01 + dp0);
    const float a2 = P12 * fast_rcp(P01 + dp1);
    const float g = 2.f * fast_rcp(Pm0 + P12);
    const float gd0 = g * dp1;
    o.G1 = (dp0 * r01) * (1.f + gd0 * (a1 - a2));
    o.G2 = gd0 * a2;
    o.nG3 = -(g * dp0 * a1);
    return o;
}
// an output edge pb inside (or, last layer, below) the source layer [pe_lo, pe_hi]: what the
// anchored integrals of the fields need (header): h, c1 = s(1+t), c2 = s t; last layer: the
// stretch below the surface and the two whole-layer cases
struct PpmEdgeGeo { float h, c1, c2, ext; bool whole, below; };
template <bool LAST>
B200CS_HD PpmEdgeGeo ppm_edge_geo(float pt, float pb, float pe_lo, float pe_hi, float rdp0) {
    PpmEdgeGeo o;
    o.h = (LAST ? fminf(pb, pe_hi) : pb) - pe_lo;
    const float g = LAST ? fmaxf(pe_hi - pb, 0.f) : pe_hi - pb;
    const float s = o.h * rdp0, t = g * rdp0;
    o.c2 = s * t;
    o.c1 = s + o.c2;
    // below the surface the last layer's mean carries on (:118-123); a last layer that lies
    // entirely inside the output layer enters through its mean (:99-104), whatever its edge values
    o.ext = LAST ? fmaxf(pb - fmaxf(pt, pe_hi), 0.f) : 0.f;
    o.whole = LAST && !(pt > pe_lo) && pb > pe_hi;
    o.below = LAST && !(pt < pe_hi);  // output layer entirely below the surface: q[km-1] whatever lies above (:52-54)
    return o;
}
// leaving the layer: what lies below its last output edge pt
struct PpmCloseGeo { float g, e1, e2; bool whole; };
B200CS_HD PpmCloseGeo ppm_close_geo(float pt, float pe_lo, float pe_hi, float rdp0) {
    PpmCloseGeo o;
    o.g = pe_hi - fmaxf(pt, pe_lo);
    const float t = o.g * rdp0, sv = 1.f - t;
    o.e2 = sv * t;
    o.e1 = t + o.e2;
    o.whole = !(pt > pe_lo);  // an edge AT pe_lo belongs to the layer above (edges are emitted while pb <= pe_hi)
    return o;
}

template <int NP>
struct PpmPack {
    static constexpr int NF = 2 * NP;
    f2 Q[4][NP];
    float PE[4], DP[4];
    f2 E[2][NP], DC[2][NP];
    f2 DQ[2][NP];           // slot j%2: q[j+1] - q[j]
    f2 AL[NP], AR[NP];      // limited edge values of the current layer
    f2 DL[NP], DR[NP];      // ... minus the layer mean
    f2 R[NP];               // mass of the output layer being filled: what lies above the current source layer
                            // (no edge in this layer yet), else minus I(s) of the last edge
    float r12;              // 1/(dp[L+1]+dp[L+2]) formed by the previous step
    float rdp0;             // 1/dp[L]
    float pe_lo, pe_hi;     // pe[L], pe[L+1]

    // boundary layers: ppm_limiters lmt=0 with the Fortran's tests kept literally
    // (mappm.f90:875-894), one pair.  In deviations: "A6 = 3(AL-q), AR = AL-A6" is dr = -2 dl,
    // "A6 = 3(AR-q), AL = AR-A6" is dl = -2 dr.
    B200CS_HD void parabola_general(int p, f2 l, f2 r, f2 a1, f2 dm) {
        float a[2], b[2];
        const float lv[2] = {f2_lo(l), f2_hi(l)}, rv[2] = {f2_lo(r), f2_hi(r)};
        const float av[2] = {f2_lo(a1), f2_hi(a1)}, dv[2] = {f2_lo(dm), f2_hi(dm)};
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const float dl = lv[j] - av[j], dr = rv[j] - av[j];
            const float sm = dl + dr, df = rv[j] - lv[j];
            const float ta = (3.f * sm) * df;
            const bool lim = fabsf(ta) > df * df;
            const bool neg = ta > 0.f;
            const float dl2 = lim ? (neg ? dl : -2.f * dr) : dl;
            const float dr2 = lim ? (neg ? -2.f * dl : dr) : dr;
            const bool flat = (dv[j] == 0.f);
            a[j] = flat ? 0.f : dl2;
            b[j] = flat ? 0.f : dr2;
        }
        DL[p] = f2_make(a[0], a[1]);
        DR[p] = f2_make(b[0], b[1]);
        AL[p] = f2_add(a1, DL[p]);
        AR[p] = f2_add(a1, DR[p]);
    }

    // interior layers: the limiter as two clamps (header)
    B200CS_HD void parabola_interior(int p, f2 l, f2 r, f2 a1, f2 dm) {
        const f2 dl = f2_sub(l, a1), dr = f2_sub(r, a1);
        const f2 tl2 = f2_add(dl, dl), tr2 = f2_add(dr, dr);
        const f2 X = f2_mul(dm, f2_bc(1.2676506e30f));  // 2^100: 0 when dm == 0, else beyond any |dl|
        const float ml0 = min3abs(f2_lo(dl), f2_lo(tr2), f2_lo(X));
        const float ml1 = min3abs(f2_hi(dl), f2_hi(tr2), f2_hi(X));
        const float mr0 = min3abs(f2_lo(dr), f2_lo(tl2), f2_lo(X));
        const float mr1 = min3abs(f2_hi(dr), f2_hi(tl2), f2_hi(X));
        // dl and dr have opposite signs (or the layer is flat): -s = sign(dl), one LOP3 each
        DL[p] = f2_make(xor_sign(ml0, f2_lo(dl)), xor_sign(ml1, f2_hi(dl)));  // -s min(|dl|, 2|dr|)
        DR[p] = f2_make(xor_sign(mr0, f2_lo(dr)), xor_sign(mr1, f2_hi(dr)));  //  s min(|dr|, 2|dl|)
        AL[p] = f2_add(a1, DL[p]);
        AR[p] = f2_add(a1, DR[p]);
    }

    // Layer 0.  p[0..4] = pe[0..4]; qA..qD = q[0..3] as NF scalars (km >= 4).  mappm.f90:689-725.
    B200CS_HD void init(const float (&p)[5], const float (&qA)[NF], const float (&qB)[NF],
                        const float (&qC)[NF], const float (&qD)[NF], int iv) {
        const float d1 = p[1] - p[0], d2 = p[2] - p[1], d3 = p[3] - p[2], d4 = p[4] - p[3];
        init_fields(d1, d2, d3, d4, qA, qB, qC, qD, iv);
        PE[0] = p[0]; PE[1] = p[1]; PE[2] = p[2]; PE[3] = p[3];
        DP[0] = d1; DP[1] = d2; DP[2] = d3; DP[3] = d4;
        r12 = fast_rcp(d2 + d3);
        rdp0 = fast_rcp(d1);
        pe_lo = p[0]; pe_hi = p[1];
    }
    // ... the field part alone (d1..d4 = dp[0..3])
    B200CS_HD void init_fields(float d1, float d2, float d3, float d4, const float (&qA)[NF], const float (&qB)[NF],
                               const float (&qC)[NF], const float (&qD)[NF], int iv) {
        float s2[NF], s3[NF], e3[NF];
        PpmStream<NF>::slope(qA, qB, qC, d1, d2, d3, s2);          // dc[1]
        PpmStream<NF>::slope(qB, qC, qD, d2, d3, d4, s3);          // dc[2]
        PpmStream<NF>::edge(qB, qC, d1, d2, d3, d4, s2, s3, e3);   // edge value at interface 2
        const float s = d1 + d2;
        const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float t3 = d2 + 3.f * d1;
        const float d1sq = d1 * d1;
        float x1v[NF], x2v[NF], dc0v[NF];
#pragma unroll
        for (int f = 0; f < NF; f++) {
            const float qa = qA[f], qb = qB[f];
            float qm = (d2 * qa + d1 * qb) / s;
            float dq = 2.f * (qb - qa) / s;
            float c1 = 4.f * (e3[f] - qm - d2 * dq) / den;
            float c3 = dq - 0.5f * c1 * t5;
            float x2 = qm - 0.25f * c1 * d1 * d2 * t3;
            float x1 = d1 * (2.f * c1 * d1sq - c3) + x2;
            x2 = fmaxf(x2, fminf(qa, qb));
            x2 = fminf(x2, fmaxf(qa, qb));
            dc0v[f] = 0.5f * (x2 - qa);
            if (iv == 0) {
                x1 = fmaxf(0.f, x1);
                x2 = fmaxf(0.f, x2);
            } else if (iv == -1) {
                if (x1 * qa <= 0.f) x1 = 0.f;
            } else if (iv == 2 || iv == -2) {
                x1 = qa;
            }
            x1v[f] = x1; x2v[f] = x2;
        }
#pragma unroll
        for (int j = 0; j < NP; j++) {
            const int a = 2 * j, b = 2 * j + 1;
            Q[0][j] = f2_make(qA[a], qA[b]); Q[1][j] = f2_make(qB[a], qB[b]);
            Q[2][j] = f2_make(qC[a], qC[b]); Q[3][j] = f2_make(qD[a], qD[b]);
            E[1][j] = f2_make(x2v[a], x2v[b]);       // interface 1
            DC[1][j] = f2_make(s2[a], s2[b]);        // layer 1
            DQ[1][j] = f2_sub(Q[2][j], Q[1][j]);
            DQ[0][j] = f2_sub(Q[1][j], Q[0][j]);
            R[j] = f2_bc(0.f);
            parabola_general(j, f2_make(x1v[a], x1v[b]), E[1][j], Q[0][j], f2_make(dc0v[a], dc0v[b]));
        }
    }

    // Enter interior layer L (1 <= L <= km-3), PH = L%4.  pn = pe[L+3], qn = q[L+2].
    // mappm.f90:658-668 (slope of layer L+1), :674-683 (edge value at interface L+1), then the
    // limited parabola of layer L.  Geometry factors as in PpmFast::advance.
    // general: layer 1 (its upper edge value comes from the top boundary formula) and layers near
    // a NaN level (header); uniform across the warp.
    template <int PH>
    B200CS_HD void step(float pn, const f2 (&qn)[NP], bool general) {
#pragma unroll
        for (int p = 0; p < NP; p++) Q[(PH + 2) & 3][p] = qn[p];
        step_inplace<PH>(pn, general);
    }
    // ... with q[L+2] already in Q[(PH+2)%4] (the kernel loads it there one step ahead)
    template <int PH>
    B200CS_HD void step_inplace(float pn, bool general) {
        constexpr int i0 = PH, i1 = (PH + 1) & 3, i2 = (PH + 2) & 3, i3 = (PH + 3) & 3;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
        const float dpm = DP[i3], dp0 = DP[i0], dp1 = DP[i1];
        const float dp2 = pn - PE[i2];
        DP[i2] = dp2;
        PE[i3] = pn;
        pe_lo = PE[i0]; pe_hi = PE[i1];
        const float r01 = r12;
        rdp0 = fast_rcp(dp0);
        const PpmStepGeo geo = ppm_step_geo(dpm, dp0, dp1, dp2, r01, r12);
        step_fields<PH>(geo, general);
    }
    // ... the field part alone
    template <int PH>
    B200CS_HD void step_fields(const PpmStepGeo& geo, bool general) {
        constexpr int i0 = PH, i1 = (PH + 1) & 3, i2 = (PH + 2) & 3;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
        const float A = geo.A, B = geo.B, G1 = geo.G1, G2 = geo.G2, nG3 = geo.nG3;
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const f2 q0 = Q[i0][p], q1 = Q[i1][p];
            const f2 d01 = DQ[e0][p], d12 = f2_sub(Q[i2][p], q1);
            DQ[e1][p] = d12;
            const f2 df2 = f2_fma(f2_bc(A), d12, f2_mul(f2_bc(B), d01));
            float dcv[2];
            const float d01v[2] = {f2_lo(d01), f2_hi(d01)}, d12v[2] = {f2_lo(d12), f2_hi(d12)};
            const float dfv[2] = {f2_lo(df2), f2_hi(df2)};
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const float hi = max3f(0.f, -d01v[j], d12v[j]);
                const float lo = max3f(0.f, d01v[j], -d12v[j]);
                dcv[j] = copysignf(min3f(fabsf(dfv[j]), hi, lo), dfv[j]);
            }
            const f2 dcn = f2_make(dcv[0], dcv[1]);
            const f2 dm = DC[e0][p];
            const f2 en = f2_fma(f2_bc(nG3), dcn, f2_fma(f2_bc(G2), dm, f2_fma(f2_bc(G1), d01, q0)));
            if (general) parabola_general(p, E[e0][p], en, q0, dm);
            else parabola_interior(p, E[e0][p], en, q0, dm);
            E[e1][p] = en;
            DC[e1][p] = dcn;
        }
    }

    // Renames the window slots so that the layer about to be entered has phase 0 (PH = its phase):
    // the two bottom layers, whose code is large (divisions), then exist once instead of four
    // times; run once per column.  DQ is not used below the interior layers.
    template <int PH>
    B200CS_HD void rotate_to0() {
        if (PH == 0) return;
        f2 q[4][NP], e[2][NP], dc[2][NP];
        float pe[4], dp[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            pe[j] = PE[(PH + j) & 3]; dp[j] = DP[(PH + j) & 3];
#pragma unroll
            for (int p = 0; p < NP; p++) q[j][p] = Q[(PH + j) & 3][p];
        }
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int p = 0; p < NP; p++) { e[j][p] = E[(PH + j) & 1][p]; dc[j][p] = DC[(PH + j) & 1][p]; }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            PE[j] = pe[j]; DP[j] = dp[j];
#pragma unroll
            for (int p = 0; p < NP; p++) Q[j][p] = q[j][p];
        }
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int p = 0; p < NP; p++) { E[j][p] = e[j][p]; DC[j][p] = dc[j][p]; }
    }

    // Enter layer km-2 (PH = (km-2)%4): bottom cubic, mappm.f90:729-761.
    template <int PH>
    B200CS_HD void step_bottom(int iv) {
        constexpr int i0 = PH, i1 = (PH + 1) & 3;
        const float d2 = DP[i0], d1 = DP[i1];  // dp[km-2], dp[km-1]
        pe_lo = PE[i0]; pe_hi = PE[i1];
        rdp0 = fast_rcp(d2);
        step_bottom_fields<PH>(d2, d1, iv);
    }
    // ... the field part alone (d2 = dp[km-2], d1 = dp[km-1])
    template <int PH>
    B200CS_HD void step_bottom_fields(float d2, float d1, int iv) {
        constexpr int i0 = PH, i1 = (PH + 1) & 3;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
        const float s = d1 + d2;
        const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float t3 = d2 + 3.f * d1;
        const float d1sq = d1 * d1;
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const float q1v[2] = {f2_lo(Q[i0][p]), f2_hi(Q[i0][p])}, q2v[2] = {f2_lo(Q[i1][p]), f2_hi(Q[i1][p])};
            const float ev[2] = {f2_lo(E[e0][p]), f2_hi(E[e0][p])};
            float lv[2], rv[2], dcv[2];
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const float qkm1 = q1v[j], qkm = q2v[j], ekm2 = ev[j];
                float qm = (d2 * qkm + d1 * qkm1) / s;
                float dq = 2.f * (qkm1 - qkm) / s;
                float c1 = (ekm2 - qm - d2 * dq) / den;
                float c3 = dq - 2.0f * c1 * t5;
                float l = qm - c1 * d1 * d2 * t3;
                float r = d1 * (8.f * c1 * d1sq - c3) + l;
                l = fmaxf(l, fminf(qkm, qkm1));
                l = fminf(l, fmaxf(qkm, qkm1));
                dcv[j] = 0.5f * (qkm - l);
                if (iv == 0) {
                    l = fmaxf(0.f, l);
                    r = fmaxf(0.f, r);
                } else if (iv < 0) {
                    if (qkm * r <= 0.f) r = 0.f;
                }
                lv[j] = l; rv[j] = r;
            }
            const f2 l2 = f2_make(lv[0], lv[1]);
            parabola_general(p, E[e0][p], l2, Q[i0][p], DC[e0][p]);
            E[e1][p] = l2;
            DC[e1][p] = f2_make(dcv[0], dcv[1]);
            E[e0][p] = f2_make(rv[0], rv[1]);  // interface km has the parity of interface km-2
        }
    }

    // Enter layer km-1 (PH = (km-1)%4).
    template <int PH>
    B200CS_HD void step_last() {
        constexpr int i0 = PH, i1 = (PH + 1) & 3;
        pe_lo = PE[i0]; pe_hi = PE[i1];
        rdp0 = fast_rcp(DP[i0]);
        step_last_fields<PH>();
    }
    template <int PH>
    B200CS_HD void step_last_fields() {
        constexpr int i0 = PH;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
#pragma unroll
        for (int p = 0; p < NP; p++) parabola_general(p, E[e0][p], E[e1][p], Q[i0][p], DC[e0][p]);
    }

    // Output edge pb inside the current layer (pb <= pe_hi), or below the surface when the current
    // layer is the last one: closes the output layer [pt, pb] and opens the next.  scale =
    // 1/(pb-pt) (times whatever weight the caller folds in).  mappm.f90:76-120 with every piece
    // anchored at the layer end it touches (header).  `on` = false: evaluate, change nothing (the
    // lanes of a warp that have no edge in this layer run along).
    template <bool LAST = true>
    B200CS_HD void edge(float pt, float pb, float scale, f2 (&out)[NP], bool on = true, const f2* qm = nullptr) {
        edge_fields<LAST>(ppm_edge_geo<LAST>(pt, pb, pe_lo, pe_hi, rdp0), scale, out, on, qm);
    }
    // ... the field part alone
    template <bool LAST = true>
    B200CS_HD void edge_fields(const PpmEdgeGeo& eg, float scale, f2 (&out)[NP], bool on, const f2* qm) {
#pragma unroll
        for (int p = 0; p < NP; p++) {
            f2 P = f2_fma(f2_bc(-eg.c2), DR[p], f2_fma(f2_bc(-eg.c1), DL[p], AL[p]));
            if (LAST) P = eg.whole ? qm[p] : P;
            f2 m = f2_fma(f2_bc(eg.h), P, R[p]);
            if (LAST) m = eg.below ? f2_mul(f2_bc(eg.ext), qm[p]) : f2_fma(f2_bc(eg.ext), qm[p], m);
            out[p] = f2_mul(m, f2_bc(scale));
            const f2 nI = f2_mul(f2_bc(-eg.h), P);
            if (on) R[p] = nI;
        }
    }

    // Leaving the current layer: what lies below its last output edge pt (the whole layer when pt
    // is not inside it, :99-104) opens the mass of the output layer being filled.
    B200CS_HD void close(float pt) { close_fields(ppm_close_geo(pt, pe_lo, pe_hi, rdp0)); }
    B200CS_HD void close_fields(const PpmCloseGeo& cg) {
        const float keep = cg.whole ? 1.f : 0.f;
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const f2 P = f2_fma(f2_bc(-cg.e2), DL[p], f2_fma(f2_bc(-cg.e1), DR[p], AR[p]));
            R[p] = f2_fma(f2_bc(cg.g), P, f2_mul(f2_bc(keep), R[p]));
        }
    }
    // ... for layers whose edge values may be NaN (a NaN level inside the stencil): a layer without
    // an output edge enters through its mean q, as in the reference, and after an edge the mass
    // above it is dropped, not multiplied by zero
    B200CS_HD void close_general(float pt, const f2 (&q)[NP]) { close_general_fields(ppm_close_geo(pt, pe_lo, pe_hi, rdp0), q); }
    B200CS_HD void close_general_fields(const PpmCloseGeo& cg, const f2 (&q)[NP]) {
#pragma unroll
        for (int p = 0; p < NP; p++) {
            const f2 P = f2_fma(f2_bc(-cg.e2), DL[p], f2_fma(f2_bc(-cg.e1), DR[p], AR[p]));
            R[p] = cg.whole ? f2_fma(f2_bc(cg.g), q[p], R[p]) : f2_mul(f2_bc(cg.g), P);
        }
    }

    // First output edge p when it does not coincide with pe[0] (general grids).  Call once p <= pe_hi
    // (or in the last layer).  An edge above the column leaves the state as it is.
    B200CS_HD void open_edge(float p) {
        if (!(p > pe_lo)) return;
        f2 dummy[NP], qz[NP];
#pragma unroll
        for (int j = 0; j < NP; j++) qz[j] = f2_sub(AL[j], DL[j]);
        edge<true>(p, p, 0.f, dummy, true, qz);  // below the surface (last layer): the mass of the whole layer, taken off again by the next edge
    }
};

}  // namespace b200cs
