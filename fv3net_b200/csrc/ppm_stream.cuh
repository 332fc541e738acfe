// ppm_stream.cuh -- forward-only ("streaming") piecewise-parabolic column remap.
//
// Replaces, for one column, the reference's mappm + ppm_profile + ppm_limiters
// (external/mappm/mappm/mappm.f90:10-126, 614-931) for kord <= 6 (the non-Huynh
// branch; every reference caller uses kord=1, iv=1 -- regridz.py:227-228).
//
// The reference materialises a4(4,km) per column and rescans it for every output
// layer.  Here a column is swept ONCE from the model top: the thread keeps a
// 3-level register window of the field(s) and a 4-level window of layer
// thicknesses, builds the edge values E(L), E(L+1) and the limited parabola of the
// CURRENT source layer L only, and moves L forward as the (monotone) target edges
// pass by.  Everything that depends only on pressure geometry is computed once per
// level and shared by the NF fields the thread carries.
//
// All arithmetic is fp32 in the operation order of the Fortran text, so with FMA
// contraction off (-fmad=false) results are bit-identical to the CPU oracle
// (oracle/mappm_ref.c).  The functions are __host__ __device__ so that the host-side
// harness in tests/ can run the very same state machine on the CPU.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define B200CS_HD __host__ __device__ __forceinline__
#else
#define B200CS_HD inline
#endif

namespace b200cs {

B200CS_HD float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }
B200CS_HD float max3f(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
// Fortran SIGN(a, b) for a >= 0; a NaN b counts as positive (its sign bit is platform noise)
B200CS_HD float sign_of(float a, float b) { return (b != b) ? a : copysignf(a, b); }

// ppm_limiters (mappm.f90:854-931) for one layer.
B200CS_HD void ppm_limit(float dm, float a1, float& al, float& ar, float& a6, int lmt) {
    if (lmt == 0) {
        if (dm == 0.f) {
            al = a1; ar = a1; a6 = 0.f;
        } else {
            float da1 = ar - al;
            float da2 = da1 * da1;
            float a6da = a6 * da1;
            if (a6da < -da2) {
                a6 = 3.f * (al - a1);
                ar = al - a6;
            } else if (a6da > da2) {
                a6 = 3.f * (ar - a1);
                al = ar - a6;
            }
        }
    } else if (lmt == 1) {
        float qmp = 2.f * dm;
        al = a1 - sign_of(fminf(fabsf(qmp), fabsf(al - a1)), qmp);
        ar = a1 + sign_of(fminf(fabsf(qmp), fabsf(ar - a1)), qmp);
        a6 = 3.f * (2.f * a1 - (al + ar));
    } else if (lmt == 2) {
        if (fabsf(ar - al) < -a6) {
            float d = ar - al;
            float fmin_ = a1 + 0.25f * (d * d) / a6 + a6 * (1.f / 12.f);
            if (fmin_ < 0.f) {
                if (a1 < ar && a1 < al) {
                    ar = a1; al = a1; a6 = 0.f;
                } else if (ar > al) {
                    a6 = 3.f * (al - a1);
                    ar = al - a6;
                } else {
                    a6 = 3.f * (ar - a1);
                    al = ar - a6;
                }
            }
        }
    }
}

// Accessor concept:  float pe(int j)  edge pressure j = 1..km+1 (fp32, as f2py casts it)
//                    float q(int f, int j)  layer mean of field f, j = 1..km
template <int NF>
struct PpmStream {
    int L, km, iv, kord, lmt_mid;
    bool poisoned;
    float pe_lo, pe_hi;        // pe(L), pe(L+1)
    float pe_top, pe_surf;     // pe(1), pe(km+1)
    float dpm, dp0, dp1, dp2;  // dp(L-1), dp(L), dp(L+1), dp(L+2)
    float q0[NF], q1[NF], q2[NF];  // q(L), q(L+1), q(L+2)
    float e0[NF], e1[NF];          // pre-limiter edge values E(L), E(L+1)
    float dc0[NF], dc1[NF];        // limited slopes dc(L), dc(L+1)
    float ebot[NF];                // E(km+1) (valid once L >= km-1)
    float al[NF], ar[NF], a6[NF];  // limited parabola of layer L
    float qtop[NF];                // q(1)

    // limited slope dc(k) from the (k-1,k,k+1) window; mappm.f90:658-668
    B200CS_HD static void slope(const float (&qa)[NF], const float (&qb)[NF], const float (&qc)[NF],
                                float da, float db, float dcn, float (&out)[NF]) {
        const float c1 = (da + 0.5f * db) / (db + dcn);
        const float c2 = (dcn + 0.5f * db) / (da + db);
        const float den = (da + db) + dcn;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            float df2 = db * (c1 * (qc[f] - qb[f]) + c2 * (qb[f] - qa[f])) / den;
            float hi = max3f(qa[f], qb[f], qc[f]) - qb[f];
            float lo = qb[f] - min3f(qa[f], qb[f], qc[f]);
            out[f] = sign_of(min3f(fabsf(df2), hi, lo), df2);
        }
    }

    // interior edge E(k), 3 <= k <= km-1, mappm.f90:674-683.
    // qa=q(k-1), qb=q(k); d2m=dp(k-2), d1m=dp(k-1), d0=dp(k), d1p=dp(k+1); sa=dc(k-1), sb=dc(k)
    B200CS_HD static void edge(const float (&qa)[NF], const float (&qb)[NF], float d2m, float d1m,
                               float d0, float d1p, const float (&sa)[NF], const float (&sb)[NF],
                               float (&out)[NF]) {
        const float d4m = d2m + d1m, d4c = d1m + d0, d4p = d0 + d1p;
        const float a1 = d4m / (d4c + d1m);
        const float a2 = d4p / (d4c + d0);
        const float g = 2.f / (d4m + d4p);
        const float a12 = a1 - a2;
        const float da1 = d1m * a1;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            float c1 = (qb[f] - qa[f]) * d1m / d4c;
            out[f] = qa[f] + c1 + g * (d0 * (c1 * a12 + a2 * sa[f]) - da1 * sb[f]);
        }
    }

    B200CS_HD void make_parabola() {
        const int lmt = (L <= 2 || L >= km - 1) ? 0 : lmt_mid;
        const bool interior = !(L <= 2 || L >= km - 1);
#pragma unroll
        for (int f = 0; f < NF; f++) {
            float l = e0[f], r = e1[f], a = 0.f;
            // kord == 4 leaves A6 to the lmt=1 limiter (mappm.f90:835-839)
            if (!(interior && kord == 4)) a = 3.f * (2.f * q0[f] - (l + r));
            if (!(interior && kord == 6)) ppm_limit(dc0[f], q0[f], l, r, a, lmt);
            al[f] = l; ar[f] = r; a6[f] = a;
        }
    }

    // bottom boundary, mappm.f90:729-761: E(km), E(km+1), dc(km) from E(km-1)
    B200CS_HD void bottom(const float (&qkm1)[NF], const float (&qkm)[NF], float dkm1, float dkm,
                          const float (&ekm1)[NF], float (&ekm)[NF], float (&dckm)[NF]) {
        const float d1 = dkm, d2 = dkm1;
        const float s = d1 + d2;
        const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
        const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
        const float t3 = d2 + 3.f * d1;
        const float d1sq = d1 * d1;
#pragma unroll
        for (int f = 0; f < NF; f++) {
            float qm = (d2 * qkm[f] + d1 * qkm1[f]) / s;
            float dq = 2.f * (qkm1[f] - qkm[f]) / s;
            float c1 = (ekm1[f] - qm - d2 * dq) / den;
            float c3 = dq - 2.0f * c1 * t5;
            float l = qm - c1 * d1 * d2 * t3;
            float r = d1 * (8.f * c1 * d1sq - c3) + l;
            l = fmaxf(l, fminf(qkm[f], qkm1[f]));
            l = fminf(l, fmaxf(qkm[f], qkm1[f]));
            dckm[f] = 0.5f * (qkm[f] - l);
            if (iv == 0) {
                l = fmaxf(0.f, l);
                r = fmaxf(0.f, r);
            } else if (iv < 0) {
                if (qkm[f] * r <= 0.f) r = 0.f;
            }
            ekm[f] = l;
            ebot[f] = r;
        }
    }

    // prologue: positions the machine on layer L = 1.  Requires km >= 4.
    template <class Acc>
    B200CS_HD void init(const Acc& a, int km_, int iv_, int kord_) {
        km = km_; iv = iv_; kord = kord_;
        int lm = kord - 3;
        if (lm < 0) lm = 0;
        if (iv == 0 && lm > 2) lm = 2;
        lmt_mid = lm;
        poisoned = false;
        const float p1 = a.pe(1), p2 = a.pe(2), p3 = a.pe(3), p4 = a.pe(4), p5 = a.pe(5);
        pe_top = p1;
        pe_surf = a.pe(km + 1);
        const float d1 = p2 - p1, d2 = p3 - p2, d3 = p4 - p3, d4 = p5 - p4;
        float qa[NF], qb[NF], qc[NF], qd[NF], s2[NF], s3[NF], e3[NF];
#pragma unroll
        for (int f = 0; f < NF; f++) {
            qa[f] = a.q(f, 1); qb[f] = a.q(f, 2); qc[f] = a.q(f, 3); qd[f] = a.q(f, 4);
        }
        slope(qa, qb, qc, d1, d2, d3, s2);  // dc(2)
        slope(qb, qc, qd, d2, d3, d4, s3);  // dc(3) (k=3 <= km-1 since km >= 4)
        edge(qb, qc, d1, d2, d3, d4, s2, s3, e3);  // E(3)
        {  // top boundary, mappm.f90:689-725
            const float s = d1 + d2;
            const float den = d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1));
            const float t5 = d2 * (5.f * d1 + d2) - 3.f * d1 * d1;
            const float t3 = d2 + 3.f * d1;
            const float d1sq = d1 * d1;
#pragma unroll
            for (int f = 0; f < NF; f++) {
                float qm = (d2 * qa[f] + d1 * qb[f]) / s;
                float dq = 2.f * (qb[f] - qa[f]) / s;
                float c1 = 4.f * (e3[f] - qm - d2 * dq) / den;
                float c3 = dq - 0.5f * c1 * t5;
                float x2 = qm - 0.25f * c1 * d1 * d2 * t3;
                float x1 = d1 * (2.f * c1 * d1sq - c3) + x2;
                x2 = fmaxf(x2, fminf(qa[f], qb[f]));
                x2 = fminf(x2, fmaxf(qa[f], qb[f]));
                dc0[f] = 0.5f * (x2 - qa[f]);
                if (iv == 0) {
                    x1 = fmaxf(0.f, x1);
                    x2 = fmaxf(0.f, x2);
                } else if (iv == -1) {
                    if (x1 * qa[f] <= 0.f) x1 = 0.f;
                } else if (iv == 2 || iv == -2) {
                    x1 = qa[f];
                }
                e0[f] = x1; e1[f] = x2;
                dc1[f] = s2[f];
                q0[f] = qa[f]; q1[f] = qb[f]; q2[f] = qc[f];
                qtop[f] = qa[f];
                ebot[f] = 0.f;
            }
        }
        L = 1;
        pe_lo = p1; pe_hi = p2;
        dpm = 0.f; dp0 = d1; dp1 = d2; dp2 = d3;
        make_parabola();
    }

    // L -> L+1.  Precondition L < km.
    template <class Acc>
    B200CS_HD void advance(const Acc& a) {
        const int Ln = L + 1;
        // shift windows
        dpm = dp0; dp0 = dp1; dp1 = dp2;
        pe_lo = pe_hi;
        pe_hi = a.pe(Ln + 1);
        float qn[NF];
        if (Ln + 2 <= km) {
            dp2 = a.pe(Ln + 3) - a.pe(Ln + 2);
#pragma unroll
            for (int f = 0; f < NF; f++) qn[f] = a.q(f, Ln + 2);
        } else {
            dp2 = 0.f;
#pragma unroll
            for (int f = 0; f < NF; f++) qn[f] = 0.f;
        }
#pragma unroll
        for (int f = 0; f < NF; f++) {
            q0[f] = q1[f]; q1[f] = q2[f]; q2[f] = qn[f];
            e0[f] = e1[f];
            dc0[f] = dc1[f];
        }
        L = Ln;
        if (Ln + 1 <= km - 1) {  // E(L+1) is an interior edge
            slope(q0, q1, q2, dp0, dp1, dp2, dc1);
            edge(q0, q1, dpm, dp0, dp1, dp2, dc0, dc1, e1);
        } else if (Ln + 1 == km) {  // E(km), E(km+1), dc(km) from the bottom cubic
            float ekm1[NF];
#pragma unroll
            for (int f = 0; f < NF; f++) ekm1[f] = e0[f];
            bottom(q0, q1, dp0, dp1, ekm1, e1, dc1);
        } else {  // Ln == km
#pragma unroll
            for (int f = 0; f < NF; f++) e1[f] = ebot[f];
        }
        make_parabola();
    }

    // One output layer with top edge pt and bottom edge pb; mappm.f90:58-124.
    template <class Acc>
    B200CS_HD void remap(const Acc& a, float pt, float pb, float (&out)[NF]) {
        const float r3 = 1.f / 3.f, r23 = 2.f / 3.f;
        if (pt <= pe_top) {
#pragma unroll
            for (int f = 0; f < NF; f++) out[f] = qtop[f];
            return;
        }
        if (pt >= pe_surf) {
            while (L < km) advance(a);
#pragma unroll
            for (int f = 0; f < NF; f++) out[f] = q0[f];
            return;
        }
        if (!poisoned) {
            while (!(pt >= pe_lo && pt <= pe_hi)) {
                if (L == km) { poisoned = true; break; }
                advance(a);
            }
        }
        if (poisoned) {
#pragma unroll
            for (int f = 0; f < NF; f++) out[f] = NAN;
            return;
        }
        const float PL = (pt - pe_lo) / dp0;
        if (pb <= pe_hi) {
            const float PR = (pb - pe_lo) / dp0;
            const float TT = r3 * (PR * (PR + PL) + PL * PL);
            const float PS = PR + PL;
#pragma unroll
            for (int f = 0; f < NF; f++)
                out[f] = al[f] + 0.5f * (a6[f] + ar[f] - al[f]) * PS - a6[f] * TT;
            return;
        }
        float qsum[NF];
        float delp = pe_hi - pt;
        {
            const float TT = r3 * (1.f + PL * (1.f + PL));
            const float P1 = 1.f + PL;
#pragma unroll
            for (int f = 0; f < NF; f++)
                qsum[f] = delp * (al[f] + 0.5f * (a6[f] + ar[f] - al[f]) * P1 - a6[f] * TT);
        }
        float dpsum = delp;
        bool closed = false;
        while (L < km) {
            advance(a);
            if (pb > pe_hi) {
#pragma unroll
                for (int f = 0; f < NF; f++) qsum[f] = qsum[f] + dp0 * q0[f];
                dpsum = dpsum + dp0;
            } else {
                delp = pb - pe_lo;
                const float esl = delp / dp0;
                const float w = 1.f - r23 * esl;
                const float hs = 0.5f * esl;
#pragma unroll
                for (int f = 0; f < NF; f++)
                    qsum[f] = qsum[f] + delp * (al[f] + hs * (ar[f] - al[f] + a6[f] * w));
                dpsum = dpsum + delp;
                closed = true;
                break;
            }
        }
        if (!closed) {
            delp = pb - pe_surf;
            if (delp > 0.f) {
#pragma unroll
                for (int f = 0; f < NF; f++) qsum[f] = qsum[f] + delp * q0[f];
                dpsum = dpsum + delp;
            }
        }
#pragma unroll
        for (int f = 0; f < NF; f++) out[f] = qsum[f] / dpsum;
    }
};

}  // namespace b200cs
