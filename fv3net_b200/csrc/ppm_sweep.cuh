// ppm_sweep.cuh -- layer-synchronous PPM column remap (B200CS_FAST arithmetic, second design).
//
// Same algorithm and arithmetic as ppm_fast.cuh (the reference's mappm + ppm_profile +
// ppm_limiters, external/mappm/mappm/mappm.f90:10-126, 614-931, for kord <= 3: limiter lmt = 0 on
// every layer), but organised the other way round: the OUTER loop runs over SOURCE layers and is
// the same for every column, the output edges that fall into the current source layer are
// emitted from an inner loop.
//
// Why (profiles/r01_fast_v3_*): with the output layer as the outer loop, the lanes of a warp sit
// on different source layers, so advance() runs 1.24x per level with 25/32 lanes active, and
// because the trip count is data dependent the compiler rotates the (q, dp, pe, edge, slope)
// windows with ~25 register moves per layer.  With the source layer as the outer loop every lane
// advances exactly once per iteration: no divergence in the expensive part, the loop is unrolled
// by 4 so the windows rotate by renaming (compile-time slot indices, zero moves), the top and
// bottom boundary formulas become uniform branches, and the shared-memory level ring is
// addressed with compile-time slots.  The data-dependent part shrinks to the edge evaluation
// (~10 flops per field).
//
// State layout: slot j%4 of Q / PE / DP holds q[j] / pe[j] / dp[j]; slot j%2 of E / DC holds the
// pre-limiter edge value at interface j / the limited slope of layer j.  All indices are
// compile-time (template parameter PH = L%4).
#pragma once
#include "ppm_fast.cuh"

namespace b200cs {

template <int NF>
struct PpmSweep {
    float Q[4][NF];
    float PE[4], DP[4];
    float E[2][NF], DC[2][NF];
    float DQ[2][NF];        // slot j%2: q[j+1] - q[j] (formed once, used by the steps of layers j-1 and j)
    float a63[NF], bb[NF];  // limited parabola of the current layer: A6/3, B = (A6+AR-AL)/2
    float R[NF];            // running mass of the output layer being filled (ppm_fast.cuh)
    float r12;              // 1/(dp[L+1]+dp[L+2]) formed by the previous step
    float rdp0;             // 1/dp[L]
    float pe_lo, pe_hi;     // pe[L], pe[L+1] of the current layer (copies: used by every edge)
    float qc[NF];           // q[L] of the current layer (copy: the edge code is phase-independent,
                            // one instance instead of four keeps the hot loop inside the i-cache)

    // ppm_limiters lmt=0 in deviation form; see PpmFast::make_parabola
    B200CS_HD void make_parabola(int f, float l, float r, float a1, float dm) {
        const float dl = l - a1, dr = r - a1;
        const float sm = dl + dr, df = r - l;
        const float ta = (3.f * sm) * df;
        const bool lim = fabsf(ta) > df * df;
        const bool neg = ta > 0.f;
        const float x = neg ? dl : dr;
        const float a = lim ? x : -sm;
        const float b = lim ? (neg ? 0.f : 3.f * x) : (-2.f * dl - dr);
        const bool flat = (dm == 0.f);
        a63[f] = flat ? 0.f : a;
        bb[f] = flat ? 0.f : b;
    }

    // Layer 0.  p[0..4] = pe[0..4]; q0..q3 = q[0..3] (km >= 4).  mappm.f90:689-725 (top cubic).
    B200CS_HD void init(const float (&p)[5], const float (&qA)[NF], const float (&qB)[NF],
                        const float (&qC)[NF], const float (&qD)[NF], int iv) {
        const float d1 = p[1] - p[0], d2 = p[2] - p[1], d3 = p[3] - p[2], d4 = p[4] - p[3];
        float s2[NF], s3[NF], e3[NF];
        PpmStream<NF>::slope(qA, qB, qC, d1, d2, d3, s2);          // dc[1]
        PpmStream<NF>::slope(qB, qC, qD, d2, d3, d4, s3);          // dc[2]
        PpmStream<NF>::edge(qB, qC, d1, d2, d3, d4, s2, s3, e3);   // edge value at interface 2
        const float s = d1 + d2;
        const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float t3 = d2 + 3.f * d1;
        const float d1sq = d1 * d1;
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
            const float dc0 = 0.5f * (x2 - qa);
            if (iv == 0) {
                x1 = fmaxf(0.f, x1);
                x2 = fmaxf(0.f, x2);
            } else if (iv == -1) {
                if (x1 * qa <= 0.f) x1 = 0.f;
            } else if (iv == 2 || iv == -2) {
                x1 = qa;
            }
            Q[0][f] = qa; Q[1][f] = qb; Q[2][f] = qC[f]; Q[3][f] = qD[f];
            E[1][f] = x2;          // interface 1
            DC[1][f] = s2[f];      // layer 1
            DQ[1][f] = qC[f] - qb;
            R[f] = 0.f;
            qc[f] = qa;
            make_parabola(f, x1, x2, qa, dc0);
        }
        PE[0] = p[0]; PE[1] = p[1]; PE[2] = p[2]; PE[3] = p[3];
        DP[0] = d1; DP[1] = d2; DP[2] = d3; DP[3] = d4;
        r12 = fast_rcp(d2 + d3);
        rdp0 = fast_rcp(d1);
        pe_lo = p[0]; pe_hi = p[1];
    }

    // Enter interior layer L (1 <= L <= km-3), PH = L%4.  pn = pe[L+3], qn = q[L+2].
    // mappm.f90:658-668 (slope of layer L+1), :674-683 (edge value at interface L+1), then the
    // limited parabola of layer L.  Geometry factors as in PpmFast::advance.
    template <int PH>
    B200CS_HD void step(float pn, const float (&qn)[NF]) {
        constexpr int i0 = PH, i1 = (PH + 1) & 3, i2 = (PH + 2) & 3, i3 = (PH + 3) & 3;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
        const float dpm = DP[i3], dp0 = DP[i0], dp1 = DP[i1];
        const float dp2 = pn - PE[i2];
        DP[i2] = dp2;
        PE[i3] = pn;
        pe_lo = PE[i0]; pe_hi = PE[i1];
        const float r01 = r12;
        rdp0 = fast_rcp(dp0);
        const float Pm0 = dpm + dp0, P01 = dp0 + dp1, P12 = dp1 + dp2;
        const float rn = fast_rcp(P12);
        r12 = rn;
        const float t = dp1 * fast_rcp(P01 + dp2);
        const float A = t * (dp0 + 0.5f * dp1) * rn;
        const float B = t * (dp2 + 0.5f * dp1) * r01;
        const float a1 = Pm0 * fast_rcp(P01 + dp0);
        const float a2 = P12 * fast_rcp(P01 + dp1);
        const float g = 2.f * fast_rcp(Pm0 + P12);
        const float gd0 = g * dp1;
        const float G1 = (dp0 * r01) * (1.f + gd0 * (a1 - a2));
        const float G2 = gd0 * a2;
        const float G3 = g * dp0 * a1;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            Q[i2][f] = qn[f];
            const float q0 = Q[i0][f], q1 = Q[i1][f];
            const float d01 = DQ[e0][f], d12 = qn[f] - q1;
            DQ[e1][f] = d12;
            const float df2 = A * d12 + B * d01;
            const float hi = max3f(0.f, -d01, d12);
            const float lo = max3f(0.f, d01, -d12);
            const float dcn = copysignf(min3f(fabsf(df2), hi, lo), df2);
            const float dm = DC[e0][f];
            const float en = q0 + G1 * d01 + G2 * dm - G3 * dcn;
            make_parabola(f, E[e0][f], en, q0, dm);
            qc[f] = q0;
            E[e1][f] = en;
            DC[e1][f] = dcn;
        }
    }

    // Enter layer km-2 (PH = (km-2)%4): bottom cubic, mappm.f90:729-761.  Leaves the edge value at
    // interface km-1 in E[(PH+1)&1], dc[km-1] in DC[(PH+1)&1] and the surface edge value in E[PH&1].
    template <int PH>
    B200CS_HD void step_bottom(int iv) {
        constexpr int i0 = PH, i1 = (PH + 1) & 3;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
        const float d2 = DP[i0], d1 = DP[i1];  // dp[km-2], dp[km-1]
        pe_lo = PE[i0]; pe_hi = PE[i1];
        rdp0 = fast_rcp(d2);
        const float s = d1 + d2;
        const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float t3 = d2 + 3.f * d1;
        const float d1sq = d1 * d1;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            const float qkm1 = Q[i0][f], qkm = Q[i1][f];
            const float ekm2 = E[e0][f];
            float qm = (d2 * qkm + d1 * qkm1) / s;
            float dq = 2.f * (qkm1 - qkm) / s;
            float c1 = (ekm2 - qm - d2 * dq) / den;
            float c3 = dq - 2.0f * c1 * t5;
            float l = qm - c1 * d1 * d2 * t3;
            float r = d1 * (8.f * c1 * d1sq - c3) + l;
            l = fmaxf(l, fminf(qkm, qkm1));
            l = fminf(l, fmaxf(qkm, qkm1));
            const float dckm = 0.5f * (qkm - l);
            if (iv == 0) {
                l = fmaxf(0.f, l);
                r = fmaxf(0.f, r);
            } else if (iv < 0) {
                if (qkm * r <= 0.f) r = 0.f;
            }
            make_parabola(f, ekm2, l, qkm1, DC[e0][f]);
            qc[f] = qkm1;
            E[e1][f] = l;
            DC[e1][f] = dckm;
            E[e0][f] = r;  // interface km has the parity of interface km-2
        }
    }

    // Enter layer km-1 (PH = (km-1)%4).
    template <int PH>
    B200CS_HD void step_last() {
        constexpr int i0 = PH, i1 = (PH + 1) & 3;
        constexpr int e0 = PH & 1, e1 = (PH + 1) & 1;
        pe_lo = PE[i0]; pe_hi = PE[i1];
        rdp0 = fast_rcp(DP[i0]);
#pragma unroll
        for (int f = 0; f < NF; f++) {
            make_parabola(f, E[e0][f], E[e1][f], Q[i0][f], DC[e0][f]);
            qc[f] = Q[i0][f];
        }
    }

    // Output edge pb inside the current layer (pb <= pe_hi), or below the surface when the current
    // layer is the last one: closes the output layer [pt, pb] (rdpc = 1/(pb-pt)) and opens the
    // next.  mappm.f90:76-120 in the mean + deviation form of ppm_fast.cuh.
    // LAST = false: pb <= pe_hi is known (the caller tested it), the edge is inside the layer.
    template <bool LAST = true>
    B200CS_HD void edge(float pt, float pb, float rdpc, float (&out)[NF]) {
        const float d = pb - fmaxf(pt, pe_lo);
        const float h = (LAST ? fminf(pb, pe_hi) : pb) - pe_lo;
        const float s = h * rdp0;
        const float u = 1.f + s;
        const bool inside = !LAST || pb <= pe_hi;
        const float nc = h * (s - 1.f);  // -dp*s*(1-s)
#pragma unroll
        for (int f = 0; f < NF; f++) {
            const float t = a63[f] * u - bb[f];
            const float nD = inside ? nc * t : 0.f;
            out[f] = ((R[f] + d * qc[f]) - nD) * rdpc;
            R[f] = nD;
        }
    }

    // Leaving the current layer: what lies below the last output edge pt enters through the mean
    // (:99-104).  Nothing is added while the level search is still above pt (:70-74).
    B200CS_HD void close(float pt) {
        if (pt < pe_hi) {
            const float d = pe_hi - fmaxf(pt, pe_lo);
#pragma unroll
            for (int f = 0; f < NF; f++) R[f] = R[f] + d * qc[f];
        }
    }

    // First output edge p when it does not coincide with pe[0] (general grids; the fused kernel's
    // output grid starts at pe[0] where R = 0 is already right).  Call once p <= pe_hi.
    B200CS_HD void open_edge(float p) {
        const float h = fminf(fmaxf(p, pe_lo), pe_hi) - pe_lo;
        const float s = h * rdp0;
        const float nc = h * (s - 1.f);
        const float u = 1.f + s;
        const bool inside = (p > pe_lo) && (p <= pe_hi);
#pragma unroll
        for (int f = 0; f < NF; f++) R[f] = inside ? nc * (a63[f] * u - bb[f]) : 0.f;
    }
};

}  // namespace b200cs
