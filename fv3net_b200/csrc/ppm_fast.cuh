// ppm_fast.cuh -- throughput-oriented forward-only PPM column remap (B200CS_FAST arithmetic).
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

template <int NF>
struct PpmFast {
    int L, km, iv;               // current source layer (0-based)
    float pe_lo, pe_hi;          // pe[L], pe[L+1]
    float pe_b, pe_a;            // pe[L+2], pe[L+3]
    float dp0, dp1, dp2;         // dp[L], dp[L+1], dp[L+2]
    float rP1;                   // 1/(dp[L+1]+dp[L+2])
    float rdp0;                  // 1/dp[L]
    float pe_top, pe_surf;       // pe[0], pe[km]
    float q0[NF], q1[NF], q2[NF];  // q[L], q[L+1], q[L+2]
    float e1[NF];                // pre-limiter edge value at interface L+1
    float dc1[NF];               // limited slope of layer L+1
    float a63[NF], bb[NF];       // limited parabola of layer L: A6/3 and B = (A6+AR-AL)/2
    float R[NF];                 // running mass of the output layer being filled (see remap)
    float ebot[NF];              // edge value at the surface, valid once L >= km-2
    float qtop[NF];

    // parabola of layer L from its edge values (l, r), mean a1 and limited slope dm;
    // mappm.f90:773-778,844-849 with ppm_limiters lmt=0 (:875-894), in deviations from the mean:
    // dl = AL-a1, dr = AR-a1.  Unlimited: A6 = -3(dl+dr), B = -2dl-dr.  The limiter's tests
    // A6*dA < -dA^2 / > dA^2 (dA = AR-AL; kept literally: they are discontinuous at dA = 0) pick
    // the edge x to keep: A6 = 3x and B = 0 (x = dl) or 3x (x = dr).
    // dm == 0 (local extremum): flat, A6 = B = 0.
    B200CS_HD void make_parabola(int f, float l, float r, float a1, float dm) {
        const float dl = l - a1, dr = r - a1;
        const float sm = dl + dr, df = r - l;
        const float ta = (3.f * sm) * df;  // -A6*dA
        const bool lim = fabsf(ta) > df * df;
        const bool neg = ta > 0.f;         // A6*dA < 0: keep AL
        const float x = neg ? dl : dr;
        const float a = lim ? x : -sm;
        const float b = lim ? (neg ? 0.f : 3.f * x) : (-2.f * dl - dr);
        const bool flat = (dm == 0.f);
        a63[f] = flat ? 0.f : a;
        bb[f] = flat ? 0.f : b;
    }

    template <class Acc>
    B200CS_HD void init(Acc& a, int km_, int iv_, float pe_surf_) {
        km = km_; iv = iv_;
        pe_surf = pe_surf_;
        const float p0 = a.pe0();
        float qA[NF], qB[NF], qC[NF], qD[NF];
        a.fetch(0, 1);
        const float p1 = a.pe(p0);
#pragma unroll
        for (int f = 0; f < NF; f++) qA[f] = a.q(f);
        a.fetch(1, 2);
        const float p2 = a.pe(p1);
#pragma unroll
        for (int f = 0; f < NF; f++) qB[f] = a.q(f);
        a.fetch(2, 3);
        const float p3 = a.pe(p2);
#pragma unroll
        for (int f = 0; f < NF; f++) qC[f] = a.q(f);
        a.fetch(3, 4);
        const float p4 = a.pe(p3);
#pragma unroll
        for (int f = 0; f < NF; f++) qD[f] = a.q(f);
        a.rewind();  // the first advance needs q[3] and pe[4] again
        pe_top = p0;
        const float d1 = p1 - p0, d2 = p2 - p1, d3 = p3 - p2, d4 = p4 - p3;
        float s2[NF], s3[NF], e3[NF], dc0[NF], x1v[NF];
        PpmStream<NF>::slope(qA, qB, qC, d1, d2, d3, s2);  // dc[1]
        PpmStream<NF>::slope(qB, qC, qD, d2, d3, d4, s3);  // dc[2]
        PpmStream<NF>::edge(qB, qC, d1, d2, d3, d4, s2, s3, e3);  // edge at interface 2
        {   // top boundary, mappm.f90:689-725
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
                dc0[f] = 0.5f * (x2 - qa);
                if (iv == 0) {
                    x1 = fmaxf(0.f, x1);
                    x2 = fmaxf(0.f, x2);
                } else if (iv == -1) {
                    if (x1 * qa <= 0.f) x1 = 0.f;
                } else if (iv == 2 || iv == -2) {
                    x1 = qa;
                }
                x1v[f] = x1;
                e1[f] = x2;
                dc1[f] = s2[f];
                q0[f] = qa; q1[f] = qb; q2[f] = qC[f];
                qtop[f] = qa;
                ebot[f] = 0.f;
                R[f] = 0.f;
            }
        }
        L = 0;
        pe_lo = p0; pe_hi = p1; pe_b = p2; pe_a = p3;
        dp0 = d1; dp1 = d2; dp2 = d3;
        rP1 = fast_rcp(d2 + d3);
        rdp0 = fast_rcp(d1);
#pragma unroll
        for (int f = 0; f < NF; f++) make_parabola(f, x1v[f], e1[f], q0[f], dc0[f]);
    }

    // bottom boundary, mappm.f90:729-761: edges at interfaces km-1 and km and dc[km-1] from the
    // edge at km-2.  qa=q[km-2], qb=q[km-1], da=dp[km-2], db=dp[km-1].
    B200CS_HD void bottom(float da, float db, float (&btm_l)[NF], float (&btm_dc)[NF]) {
        const float d1 = db, d2 = da;
        const float s = d1 + d2;
        const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float t3 = d2 + 3.f * d1;
        const float d1sq = d1 * d1;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            const float qkm1 = q0[f], qkm = q1[f];
            float qm = (d2 * qkm + d1 * qkm1) / s;
            float dq = 2.f * (qkm1 - qkm) / s;
            float c1 = (e1[f] - qm - d2 * dq) / den;  // e1 = edge at interface km-2 (not yet replaced)
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
            ebot[f] = r;
            btm_l[f] = l;
            btm_dc[f] = dckm;
        }
    }

    // L -> L+1.  Precondition L < km-1.
    template <class Acc>
    B200CS_HD void advance(Acc& a) {
        const int Ln = L + 1;
        L = Ln;
        // geometry window: dpm, dp0, dp1, dp2 = dp[Ln-1], dp[Ln], dp[Ln+1], dp[Ln+2]
        const float dpm = dp0;
        dp0 = dp1; dp1 = dp2;
        pe_lo = pe_hi; pe_hi = pe_b; pe_b = pe_a;
        a.fetch(Ln + 2, Ln + 3);
        const float pe_n = a.pe(pe_a);   // pe[Ln+3] (garbage beyond the surface, unused)
        dp2 = pe_n - pe_a;
        pe_a = pe_n;
        const float r01 = rP1;           // 1/(dp[Ln]+dp[Ln+1]), formed by the previous advance
        rdp0 = fast_rcp(dp0);
        float dl[NF], dm[NF];            // left edge value and limited slope of layer Ln
#pragma unroll
        for (int f = 0; f < NF; f++) {
            q0[f] = q1[f]; q1[f] = q2[f]; q2[f] = a.q(f);
            dl[f] = e1[f]; dm[f] = dc1[f];
        }
        if (Ln + 1 <= km - 2) {
            // layer Ln+1 has a centred slope and interface Ln+1 an interior edge value
            const float Pm0 = dpm + dp0, P01 = dp0 + dp1, P12 = dp1 + dp2;
            const float r12 = fast_rcp(P12);
            rP1 = r12;
            // slope of layer Ln+1, window (Ln, Ln+1, Ln+2), mappm.f90:658-662 with the divisions
            // hoisted: df2 = A*(q2-q1) + B*(q1-q0)
            const float t = dp1 * fast_rcp(P01 + dp2);
            const float A = t * (dp0 + 0.5f * dp1) * r12;
            const float B = t * (dp2 + 0.5f * dp1) * r01;
            // edge value at interface Ln+1, :674-683 with d2m,d1m,d0,d1p = dpm,dp0,dp1,dp2 and
            // d4(k-2),d4(k-1),d4(k) = Pm0,P01,P12, expanded in d01, dc[Ln], dc[Ln+1]:
            //   E = q0 + G1*d01 + G2*dc[Ln] - G3*dc[Ln+1]
            const float a1 = Pm0 * fast_rcp(P01 + dp0);
            const float a2 = P12 * fast_rcp(P01 + dp1);
            const float g = 2.f * fast_rcp(Pm0 + P12);
            const float gd0 = g * dp1;
            const float G1 = (dp0 * r01) * (1.f + gd0 * (a1 - a2));
            const float G2 = gd0 * a2;
            const float G3 = g * dp0 * a1;
#pragma unroll
            for (int f = 0; f < NF; f++) {
                const float d01 = q1[f] - q0[f], d12 = q2[f] - q1[f];
                const float df2 = A * d12 + B * d01;
                const float hi = max3f(0.f, -d01, d12);
                const float lo = max3f(0.f, d01, -d12);
                const float dcn = copysignf(min3f(fabsf(df2), hi, lo), df2);
                const float en = q0[f] + G1 * d01 + G2 * dm[f] - G3 * dcn;
                e1[f] = en;
                dc1[f] = dcn;
            }
        } else if (Ln + 1 == km - 1) {
            float btm_l[NF], btm_dc[NF];
            bottom(dp0, dp1, btm_l, btm_dc);
#pragma unroll
            for (int f = 0; f < NF; f++) { e1[f] = btm_l[f]; dc1[f] = btm_dc[f]; }
        } else {  // Ln == km-1
#pragma unroll
            for (int f = 0; f < NF; f++) e1[f] = ebot[f];
        }
#pragma unroll
        for (int f = 0; f < NF; f++) make_parabola(f, dl[f], e1[f], q0[f], dm[f]);
    }

    // Positions the machine on the first output edge p (only needed when p > pe[0]; the fused
    // kernel's output grid starts at pe[0], where R = 0 from init() is already right).
    template <class Acc>
    B200CS_HD void open(Acc& a, float p) {
        while (p > pe_hi && L < km - 1) advance(a);
        const float h = fminf(fmaxf(p, pe_lo), pe_hi) - pe_lo;
        const float s = h * rdp0;
        const float nc = h * (s - 1.f);
        const float u = 1.f + s;
        const bool inside = (p > pe_lo) && (p <= pe_hi);
#pragma unroll
        for (int f = 0; f < NF; f++) R[f] = inside ? nc * (a63[f] * u - bb[f]) : 0.f;
    }

    // One output layer [pt, pb] (rdpc = 1/(pb-pt)); mappm.f90:58-124.  Must be called for
    // consecutive output layers (pt = the previous call's pb), starting at pt = pe[0] or after
    // open(pt).
    template <class Acc>
    B200CS_HD void remap(Acc& a, float pt, float pb, float rdpc, float (&out)[NF]) {
        // source layers that end above pb: the part of them below pt enters through the mean
        // (:99-104 for whole layers; for the layer holding pt, R already carries -D(pt))
        while (pb > pe_hi && L < km - 1) {
            if (pt < pe_hi) {  // false while the level search skips layers above pt (:70-74)
                const float d = pe_hi - fmaxf(pt, pe_lo);
#pragma unroll
                for (int f = 0; f < NF; f++) R[f] = R[f] + d * q0[f];
            }
            advance(a);
        }
        // pb lies in layer L -- or below the surface (L == km-1): the mean then extends downwards
        // (:65-67,115-120) and the edge sits at s = 1 where D vanishes
        const float d = pb - fmaxf(pt, pe_lo);
        const float h = fminf(pb, pe_hi) - pe_lo;
        const float s = h * rdp0;
        const float nc = h * (s - 1.f);  // -dp*s*(1-s)
        const float u = 1.f + s;
        const bool inside = pb <= pe_hi;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            const float t = a63[f] * u - bb[f];
            const float nD = inside ? nc * t : 0.f;  // -D(s)
            out[f] = ((R[f] + d * q0[f]) - nD) * rdpc;
            R[f] = nD;
        }
        if (pt <= pe_top) {  // :62-64: the top output layer copies the top source layer
#pragma unroll
            for (int f = 0; f < NF; f++) out[f] = qtop[f];
        }
    }
};

}  // namespace b200cs
