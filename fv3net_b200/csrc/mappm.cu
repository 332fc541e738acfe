// mappm.cu -- K7: standalone column remap, the device replacement of the f2py extension
// `mappm.mappm` (external/mappm/mappm/mappm.f90:10-126; call site
// external/vcm/vcm/cubedsphere/regridz.py:329-334).
//
// Two kernels, one thread per column, coalesced across columns when the level stride is the
// large one (restart layout [z, ncol]); the reference's [ncol, nz] layout works through the
// same strides (the three transposing copies of regridz.py:270-272,376-384 are never made):
//   * kord <= 6: the forward-only streaming machine of ppm_stream.cuh -- no workspace, each
//     input element is read once;
//   * kord >= 7 (Huynh's constraint, mappm.f90:780-826) and kord >= 8 (cs_profile,
//     mappm.f90:132-611) need two sweeps over the column, so the profile is built in a
//     per-thread slot of a caller-provided global workspace laid out [level][slot]
//     (coalesced across the threads of a warp), then remapped.
// fp32 throughout like the Fortran build; compiled with -fmad=false.
#include <mutex>

#include "common.cuh"
#include "ppm_stream.cuh"

namespace b200cs {

struct GlobalAcc {
    const float* pe1; int64_t pe_ls;
    const float* q1; int64_t q_ls;
    __device__ __forceinline__ float pe(int j) const { return __ldg(pe1 + (int64_t)(j - 1) * pe_ls); }
    __device__ __forceinline__ float q(int, int j) const { return __ldg(q1 + (int64_t)(j - 1) * q_ls); }
};

__global__ void __launch_bounds__(128)
mappm_stream_kernel(const float* __restrict__ pe1, int64_t pe1_cs, int64_t pe1_ls,
                    const float* __restrict__ q1, int64_t q1_cs, int64_t q1_ls,
                    const float* __restrict__ pe2, int64_t pe2_cs, int64_t pe2_ls,
                    float* __restrict__ q2, int64_t q2_cs, int64_t q2_ls, int64_t ncol, int km, int kn,
                    int iv, int kord) {
    for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < ncol;
         c += (int64_t)gridDim.x * blockDim.x) {
        GlobalAcc acc{pe1 + c * pe1_cs, pe1_ls, q1 + c * q1_cs, q1_ls};
        PpmStream<1> s;
        s.init(acc, km, iv, kord);
        const float* p2 = pe2 + c * pe2_cs;
        float pt = __ldg(p2);
        for (int k = 0; k < kn; k++) {
            const float pb = __ldg(p2 + (int64_t)(k + 1) * pe2_ls);
            float o[1];
            s.remap(acc, pt, pb, o);
            q2[c * q2_cs + (int64_t)k * q2_ls] = o[0];
            pt = pb;
        }
    }
}

// ---- generic two-sweep path -------------------------------------------------------------
struct Lev {  // 1-based level view into a [level][slot] workspace plane
    float* p; int64_t s;
    __device__ __forceinline__ float& operator()(int k) const { return p[(int64_t)(k - 1) * s]; }
};
struct CLev {  // 1-based read-only strided input column
    const float* p; int64_t s;
    __device__ __forceinline__ float operator()(int k) const { return __ldg(p + (int64_t)(k - 1) * s); }
};

__device__ __forceinline__ void cs_limit(bool extm, float a1, float& al, float& ar, float& a6, int iv) {
    // cs_limiters, mappm.f90:535-611
    if (iv == 0) {
        if (a1 <= 0.f) { al = a1; ar = a1; a6 = 0.f; }
        else if (fabsf(ar - al) < -a6) {
            const float d = ar - al;
            if ((a1 + 0.25f * (d * d) / a6 + a6 * (1.f / 12.f)) < 0.f) {
                if (a1 < ar && a1 < al) { ar = a1; al = a1; a6 = 0.f; }
                else if (ar > al) { a6 = 3.f * (al - a1); ar = al - a6; }
                else { a6 = 3.f * (ar - a1); al = ar - a6; }
            }
        }
        return;
    }
    const bool flat = (iv == 1) ? ((a1 - al) * (a1 - ar) >= 0.f) : extm;
    if (flat) { al = a1; ar = a1; a6 = 0.f; return; }
    const float da1 = ar - al, da2 = da1 * da1, a6da = a6 * da1;
    if (a6da < -da2) { a6 = 3.f * (al - a1); ar = al - a6; }
    else if (a6da > da2) { a6 = 3.f * (ar - a1); al = ar - a6; }
}

// ppm_profile for kord == 7 (general kord handled too), mappm.f90:614-851
__device__ void ppm_profile_ws(const CLev& A1, const Lev& AL, const Lev& AR, const Lev& A6, const Lev& DP,
                               const Lev& DC, const Lev& H2, const Lev& DELQ, const Lev& D4, int km, int iv,
                               int kord) {
    const int km1 = km - 1;
    for (int k = 2; k <= km; k++) {
        DELQ(k - 1) = A1(k) - A1(k - 1);
        D4(k) = DP(k - 1) + DP(k);
    }
    for (int k = 2; k <= km1; k++) {
        const float c1 = (DP(k - 1) + 0.5f * DP(k)) / D4(k + 1);
        const float c2 = (DP(k + 1) + 0.5f * DP(k)) / D4(k);
        const float df2 = DP(k) * (c1 * DELQ(k) + c2 * DELQ(k - 1)) / (D4(k) + DP(k + 1));
        const float a = A1(k - 1), b = A1(k), c = A1(k + 1);
        DC(k) = sign_of(fabsf(min3f(fabsf(df2), max3f(a, b, c) - b, b - min3f(a, b, c))), df2);
    }
    for (int k = 3; k <= km1; k++) {
        const float c1 = DELQ(k - 1) * DP(k - 1) / D4(k);
        const float a1 = D4(k - 1) / (D4(k) + DP(k - 1));
        const float a2 = D4(k + 1) / (D4(k) + DP(k));
        AL(k) = A1(k - 1) + c1 +
                2.f / (D4(k - 1) + D4(k + 1)) * (DP(k) * (c1 * (a1 - a2) + a2 * DC(k - 1)) - DP(k - 1) * a1 * DC(k));
    }
    {
        const float d1 = DP(1), d2 = DP(2);
        const float qm = (d2 * A1(1) + d1 * A1(2)) / (d1 + d2);
        const float dq = 2.f * (A1(2) - A1(1)) / (d1 + d2);
        const float c1 = 4.f * (AL(3) - qm - d2 * dq) / (d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1)));
        const float c3 = dq - 0.5f * c1 * (d2 * (5.f * d1 + d2) - 3.f * d1 * d1);
        float x2 = qm - 0.25f * c1 * d1 * d2 * (d2 + 3.f * d1);
        AL(1) = d1 * (2.f * c1 * (d1 * d1) - c3) + x2;
        x2 = fmaxf(x2, fminf(A1(1), A1(2)));
        x2 = fminf(x2, fmaxf(A1(1), A1(2)));
        AL(2) = x2;
        DC(1) = 0.5f * (x2 - A1(1));
    }
    if (iv == 0) { AL(1) = fmaxf(0.f, AL(1)); AL(2) = fmaxf(0.f, AL(2)); }
    else if (iv == -1) { if (AL(1) * A1(1) <= 0.f) AL(1) = 0.f; }
    else if (iv == 2 || iv == -2) { AL(1) = A1(1); }
    {
        const float d1 = DP(km), d2 = DP(km1);
        const float qm = (d2 * A1(km) + d1 * A1(km1)) / (d1 + d2);
        const float dq = 2.f * (A1(km1) - A1(km)) / (d1 + d2);
        const float c1 = (AL(km1) - qm - d2 * dq) / (d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1)));
        const float c3 = dq - 2.0f * c1 * (d2 * (5.f * d1 + d2) - 3.f * d1 * d1);
        float l = qm - c1 * d1 * d2 * (d2 + 3.f * d1);
        float r = d1 * (8.f * c1 * (d1 * d1) - c3) + l;
        l = fmaxf(l, fminf(A1(km), A1(km1)));
        l = fminf(l, fmaxf(A1(km), A1(km1)));
        DC(km) = 0.5f * (A1(km) - l);
        if (iv == 0) { l = fmaxf(0.f, l); r = fmaxf(0.f, r); }
        else if (iv < 0) { if (A1(km) * r <= 0.f) r = 0.f; }
        AL(km) = l; AR(km) = r;
    }
    for (int k = 1; k <= km1; k++) AR(k) = AL(k + 1);
    auto std_a6 = [&](int k) { return 3.f * (2.f * A1(k) - (AL(k) + AR(k))); };
    auto limit = [&](int k, int lmt) {
        float l = AL(k), r = AR(k), a = A6(k);
        ppm_limit(DC(k), A1(k), l, r, a, lmt);
        AL(k) = l; AR(k) = r; A6(k) = a;
    };
    for (int k = 1; k <= 2; k++) { A6(k) = std_a6(k); limit(k, 0); }
    if (kord >= 7) {
        for (int k = 2; k <= km1; k++)
            H2(k) = 2.f * (DC(k + 1) / DP(k + 1) - DC(k - 1) / DP(k - 1)) / (DP(k) + 0.5f * (DP(k - 1) + DP(k + 1))) *
                    (DP(k) * DP(k));
        for (int k = 3; k <= km - 2; k++) {
            const float a = A1(k), pmp = 2.f * DC(k);
            float qmp = a + pmp;
            float lac = a + 1.5f * H2(k - 1) + DC(k);
            AR(k) = fminf(fmaxf(AR(k), min3f(a, qmp, lac)), max3f(a, qmp, lac));
            qmp = a - pmp;
            lac = a + 1.5f * H2(k + 1) - DC(k);
            AL(k) = fminf(fmaxf(AL(k), min3f(a, qmp, lac)), max3f(a, qmp, lac));
            A6(k) = std_a6(k);
            if (iv == 0 && kord >= 6) limit(k, 2);
        }
    } else {
        int lmt = kord - 3;
        if (lmt < 0) lmt = 0;
        if (iv == 0 && lmt > 2) lmt = 2;
        for (int k = 3; k <= km - 2; k++) {
            if (kord != 4) A6(k) = std_a6(k); else A6(k) = 0.f;
            if (kord != 6) limit(k, lmt);
        }
    }
    for (int k = km1; k <= km; k++) { A6(k) = std_a6(k); limit(k, 0); }
}

// cs_profile, mappm.f90:132-532.  GAM, Q (km+1) and FLG (bit0 extm, bit1 ext5, bit2 ext6) planes.
__device__ void cs_profile_ws(float qs, const CLev& A1, const Lev& AL, const Lev& AR, const Lev& A6,
                              const Lev& DP, const Lev& GAM, const Lev& Q, const Lev& FLG, int km, int iv,
                              int kord) {
    const int akord = kord < 0 ? -kord : kord;
    float d4 = 0.f, bet, grat;
    if (iv == -2) {
        GAM(2) = 0.5f;
        Q(1) = 1.5f * A1(1);
        for (int k = 2; k <= km - 1; k++) {
            grat = DP(k - 1) / DP(k);
            bet = 2.f + grat + grat - GAM(k);
            Q(k) = (3.f * (A1(k - 1) + A1(k)) - Q(k - 1)) / bet;
            GAM(k + 1) = grat / bet;
        }
        grat = DP(km - 1) / DP(km);
        Q(km) = (3.f * (A1(km - 1) + A1(km)) - grat * qs - Q(km - 1)) / (2.f + grat + grat - GAM(km));
        Q(km + 1) = qs;
        for (int k = km - 1; k >= 1; k--) Q(k) = Q(k) - GAM(k + 1) * Q(k + 1);
    } else {
        grat = DP(2) / DP(1);
        bet = grat * (grat + 0.5f);
        Q(1) = ((grat + grat) * (grat + 1.f) * A1(1) + A1(2)) / bet;
        GAM(1) = (1.f + grat * (grat + 1.5f)) / bet;
        for (int k = 2; k <= km; k++) {
            d4 = DP(k - 1) / DP(k);
            bet = 2.f + d4 + d4 - GAM(k - 1);
            Q(k) = (3.f * (A1(k - 1) + d4 * A1(k)) - Q(k - 1)) / bet;
            GAM(k) = d4 / bet;
        }
        const float a_bot = 1.f + d4 * (d4 + 1.5f);
        Q(km + 1) = (2.f * d4 * (d4 + 1.f) * A1(km) + A1(km - 1) - a_bot * Q(km)) / (d4 * (d4 + 0.5f) - a_bot * GAM(km));
        for (int k = km; k >= 1; k--) Q(k) = Q(k) - GAM(k) * Q(k + 1);
    }
    auto std_a6 = [&](int k) { return 3.f * (2.f * A1(k) - (AL(k) + AR(k))); };
    if (akord > 16) {
        for (int k = 1; k <= km; k++) { AL(k) = Q(k); AR(k) = Q(k + 1); A6(k) = std_a6(k); }
        return;
    }
    Q(2) = fminf(Q(2), fmaxf(A1(1), A1(2)));
    Q(2) = fmaxf(Q(2), fminf(A1(1), A1(2)));
    for (int k = 2; k <= km; k++) GAM(k) = A1(k) - A1(k - 1);
    for (int k = 3; k <= km - 1; k++) {
        const float lo = fminf(A1(k - 1), A1(k)), hi = fmaxf(A1(k - 1), A1(k));
        if (GAM(k - 1) * GAM(k + 1) > 0.f) {
            Q(k) = fminf(Q(k), hi);
            Q(k) = fmaxf(Q(k), lo);
        } else if (GAM(k - 1) > 0.f) {
            Q(k) = fmaxf(Q(k), lo);
        } else {
            Q(k) = fminf(Q(k), hi);
            if (iv == 0) Q(k) = fmaxf(0.f, Q(k));
        }
    }
    Q(km) = fminf(Q(km), fmaxf(A1(km - 1), A1(km)));
    Q(km) = fmaxf(Q(km), fminf(A1(km - 1), A1(km)));
    for (int k = 1; k <= km; k++) { AL(k) = Q(k); AR(k) = Q(k + 1); }
    for (int k = 1; k <= km; k++) {
        int flg;
        if (k == 1 || k == km) flg = ((AL(k) - A1(k)) * (AR(k) - A1(k)) > 0.f) ? 1 : 0;
        else flg = (GAM(k) * GAM(k + 1) < 0.f) ? 1 : 0;
        if (akord > 9) {
            const float x0 = 2.f * A1(k) - (AL(k) + AR(k));
            const float x1 = fabsf(AL(k) - AR(k));
            A6(k) = 3.f * x0;
            if (fabsf(x0) > x1) flg |= 2;
            if (fabsf(A6(k)) > x1) flg |= 4;
        }
        FLG(k) = __int_as_float(flg);
    }
    auto extm = [&](int k) { return (__float_as_int(FLG(k)) & 1) != 0; };
    auto ext5 = [&](int k) { return (__float_as_int(FLG(k)) & 2) != 0; };
    auto ext6 = [&](int k) { return (__float_as_int(FLG(k)) & 4) != 0; };
    auto climit = [&](int k, int mode) {
        float l = AL(k), r = AR(k), a = A6(k);
        cs_limit(extm(k), A1(k), l, r, a, mode);
        AL(k) = l; AR(k) = r; A6(k) = a;
    };
    if (iv == 0) AL(1) = fmaxf(0.f, AL(1));
    else if (iv == -1) { if (AL(1) * A1(1) <= 0.f) AL(1) = 0.f; }
    else if (iv == 2) { AL(1) = A1(1); AR(1) = A1(1); A6(1) = 0.f; }
    if (iv != 2) { A6(1) = std_a6(1); climit(1, 1); }
    A6(2) = std_a6(2);
    climit(2, 2);
    auto huynh = [&](int k) {
        const float a = A1(k);
        const float pmp_1 = a - 2.f * GAM(k + 1);
        const float lac_1 = pmp_1 + 1.5f * GAM(k + 2);
        AL(k) = fminf(fmaxf(AL(k), min3f(a, pmp_1, lac_1)), max3f(a, pmp_1, lac_1));
        const float pmp_2 = a + 2.f * GAM(k);
        const float lac_2 = pmp_2 - 1.5f * GAM(k - 1);
        AR(k) = fminf(fmaxf(AR(k), min3f(a, pmp_2, lac_2)), max3f(a, pmp_2, lac_2));
    };
    auto flat = [&](int k) { AL(k) = A1(k); AR(k) = A1(k); };
    auto six_a6 = [&](int k) { return 6.f * A1(k) - 3.f * (AL(k) + AR(k)); };
    for (int k = 3; k <= km - 2; k++) {
        if (akord < 9) {
            huynh(k); A6(k) = std_a6(k);
        } else if (akord == 9) {
            if (extm(k) && (extm(k - 1) || extm(k + 1))) { flat(k); A6(k) = 0.f; }
            else {
                A6(k) = six_a6(k);
                if (fabsf(A6(k)) > fabsf(AL(k) - AR(k))) { huynh(k); A6(k) = six_a6(k); }
            }
        } else if (akord == 10) {
            if (ext5(k)) {
                if (ext5(k - 1) || ext5(k + 1)) flat(k);
                else if (ext6(k - 1) || ext6(k + 1)) huynh(k);
            } else if (ext6(k)) {
                if (ext5(k - 1) || ext5(k + 1)) huynh(k);
            }
            A6(k) = std_a6(k);
        } else if (akord == 12) {
            if (extm(k)) { flat(k); A6(k) = 0.f; }
            else {
                A6(k) = six_a6(k);
                if (fabsf(A6(k)) > fabsf(AL(k) - AR(k))) { huynh(k); A6(k) = six_a6(k); }
            }
        } else if (akord == 13) {
            if (ext6(k) && ext6(k - 1) && ext6(k + 1)) flat(k);
            A6(k) = std_a6(k);
        } else if (akord == 14) {
            A6(k) = std_a6(k);
        } else if (akord == 15) {
            if (ext5(k)) { if (ext5(k - 1) || ext5(k + 1)) flat(k); }
            else if (ext6(k)) huynh(k);
            A6(k) = std_a6(k);
        } else if (akord == 16) {
            if (ext5(k)) {
                if (ext5(k - 1) || ext5(k + 1)) flat(k);
                else if (ext6(k - 1) || ext6(k + 1)) huynh(k);
            }
            A6(k) = std_a6(k);
        } else {  // kord = 11
            if (ext5(k) && (ext5(k - 1) || ext5(k + 1))) { flat(k); A6(k) = 0.f; }
            else A6(k) = std_a6(k);
        }
        if (iv == 0) climit(k, 0);
    }
    if (iv == 0) AR(km) = fmaxf(0.f, AR(km));
    else if (iv == -1) { if (AR(km) * A1(km) <= 0.f) AR(km) = 0.f; }
    A6(km - 1) = std_a6(km - 1); climit(km - 1, 2);
    A6(km) = std_a6(km); climit(km, 1);
}

constexpr int kWsPlanes = 8;  // AL AR A6 DP + 4 work planes, each (km+1) rows (Q needs km+1)

__global__ void __launch_bounds__(128)
mappm_generic_kernel(const float* __restrict__ pe1, int64_t pe1_cs, int64_t pe1_ls,
                     const float* __restrict__ q1, int64_t q1_cs, int64_t q1_ls,
                     const float* __restrict__ pe2, int64_t pe2_cs, int64_t pe2_ls,
                     float* __restrict__ q2, int64_t q2_cs, int64_t q2_ls, int64_t ncol, int km, int kn,
                     int iv, int kord, float* __restrict__ ws, int64_t nslot) {
    const int64_t slot = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (slot >= nslot) return;
    const int64_t plane = (int64_t)(km + 1) * nslot;
    auto lev = [&](int p) { return Lev{ws + p * plane + slot, nslot}; };
    const Lev AL = lev(0), AR = lev(1), A6 = lev(2), DP = lev(3);
    const float r3 = 1.f / 3.f, r23 = 2.f / 3.f;
    for (int64_t c = slot; c < ncol; c += nslot) {
        const CLev PE1{pe1 + c * pe1_cs, pe1_ls}, A1{q1 + c * q1_cs, q1_ls}, PE2{pe2 + c * pe2_cs, pe2_ls};
        for (int k = 1; k <= km; k++) DP(k) = PE1(k + 1) - PE1(k);
        if (kord > 7) cs_profile_ws(0.f, A1, AL, AR, A6, DP, lev(4), lev(5), lev(6), km, iv, kord);
        else ppm_profile_ws(A1, AL, AR, A6, DP, lev(4), lev(5), lev(6), lev(7), km, iv, kord);
        // remap sweep, mappm.f90:58-124 (forward-only: k0 never decreases)
        int k0 = 1;
        bool poisoned = false;
        const float ptop = PE1(1), psurf = PE1(km + 1);
        for (int k = 1; k <= kn; k++) {
            const float pt = PE2(k), pb = PE2(k + 1);
            float res;
            if (pt <= ptop) res = A1(1);
            else if (pt >= psurf) res = A1(km);
            else if (poisoned) res = NAN;
            else {
                int L = k0;
                while (L <= km && !(pt >= PE1(L) && pt <= PE1(L + 1))) L++;
                if (L > km) { poisoned = true; res = NAN; }
                else {
                    k0 = L;
                    const float PL = (pt - PE1(L)) / DP(L);
                    if (pb <= PE1(L + 1)) {
                        const float PR = (pb - PE1(L)) / DP(L);
                        const float TT = r3 * (PR * (PR + PL) + PL * PL);
                        res = AL(L) + 0.5f * (A6(L) + AR(L) - AL(L)) * (PR + PL) - A6(L) * TT;
                    } else {
                        float delp = PE1(L + 1) - pt;
                        const float TT = r3 * (1.f + PL * (1.f + PL));
                        float qsum = delp * (AL(L) + 0.5f * (A6(L) + AR(L) - AL(L)) * (1.f + PL) - A6(L) * TT);
                        float dpsum = delp;
                        bool closed = false;
                        for (L = L + 1; L <= km; L++) {
                            if (pb > PE1(L + 1)) {
                                qsum = qsum + DP(L) * A1(L);
                                dpsum = dpsum + DP(L);
                            } else {
                                delp = pb - PE1(L);
                                const float esl = delp / DP(L);
                                qsum = qsum + delp * (AL(L) + 0.5f * esl * (AR(L) - AL(L) + A6(L) * (1.f - r23 * esl)));
                                dpsum = dpsum + delp;
                                k0 = L;
                                closed = true;
                                break;
                            }
                        }
                        if (!closed) {
                            delp = pb - psurf;
                            if (delp > 0.f) { qsum = qsum + delp * A1(km); dpsum = dpsum + delp; }
                        }
                        res = qsum / dpsum;
                    }
                }
            }
            q2[c * q2_cs + (int64_t)(k - 1) * q2_ls] = res;
        }
    }
}

constexpr int64_t kMaxSlots = 148LL * 8 * 128;

}  // namespace b200cs

using namespace b200cs;

extern "C" int64_t b200cs_mappm_workspace_bytes(int64_t ncol, int64_t km) {
    if (ncol <= 0 || km <= 0) return 0;
    const int64_t nslot = ncol < kMaxSlots ? ncol : kMaxSlots;
    return nslot * (km + 1) * kWsPlanes * (int64_t)sizeof(float);
}

extern "C" int b200cs_mappm(const float* pe1, int64_t pe1_cs, int64_t pe1_ls, const float* q1, int64_t q1_cs,
                            int64_t q1_ls, const float* pe2, int64_t pe2_cs, int64_t pe2_ls, float* q2,
                            int64_t q2_cs, int64_t q2_ls, int64_t ncol, int km, int kn, int iv, int kord,
                            void* workspace, int64_t workspace_bytes, void* stream) {
    B200CS_REQUIRE(pe1 && q1 && pe2 && q2, "mappm: NULL argument");
    B200CS_REQUIRE(ncol >= 0 && kn >= 1, "mappm: bad extents");
    B200CS_REQUIRE(km >= 4 && (kord <= 7 || km >= 5),
                   "mappm: km=%d too small (the Fortran indexes out of range below 4 levels, 5 for kord>7)", km);
    if (ncol == 0) return B200CS_OK;
    cudaStream_t st = as_stream(stream);
    if (kord <= 6) {
        int64_t blocks = (ncol + 127) / 128;
        if (blocks > 148LL * 16) blocks = 148LL * 16;
        mappm_stream_kernel<<<(unsigned)blocks, 128, 0, st>>>(pe1, pe1_cs, pe1_ls, q1, q1_cs, q1_ls, pe2, pe2_cs,
                                                              pe2_ls, q2, q2_cs, q2_ls, ncol, km, kn, iv, kord);
        return check_launch("mappm(stream)");
    }
    const int64_t need = b200cs_mappm_workspace_bytes(ncol, km);
    if (!workspace || workspace_bytes < need)
        return set_err(B200CS_EWORKSPACE, "mappm: kord=%d needs %lld workspace bytes, got %lld", kord,
                       (long long)need, (long long)workspace_bytes);
    const int64_t nslot = ncol < kMaxSlots ? ncol : kMaxSlots;
    mappm_generic_kernel<<<(unsigned)((nslot + 127) / 128), 128, 0, st>>>(
        pe1, pe1_cs, pe1_ls, q1, q1_cs, q1_ls, pe2, pe2_cs, pe2_ls, q2, q2_cs, q2_ls, ncol, km, kn, iv, kord,
        (float*)workspace, nslot);
    return check_launch("mappm(generic)");
}

// ---------------------------------------------------------------------------------------------
// b200cs_mappm_static: b200cs_mappm for callers that may not allocate (CUDA graph capture, serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The [level][slot] workspace of the kord > 6 machine lives in a file-scope
// device array of B200CS_MAPPM_STATIC_MAX_BYTES that every GPU loading the library holds until
// exit.  The kernel writes every workspace element before reading it (DP, then the profile planes,
// per column), so nothing has to be cleared between calls; all calls on a device share the array,
// so calls that overlap on the GPU (different streams) race.
// ---------------------------------------------------------------------------------------------
namespace b200cs {
static __device__ __align__(256) float g_mappm_static_ws[B200CS_MAPPM_STATIC_MAX_BYTES / sizeof(float)];

constexpr int kMaxStaticDevices = 64;
static float* g_static_ws_ptr[kMaxStaticDevices] = {};
static std::mutex g_static_ws_mutex;

// device address of the array on the CURRENT device (the name above is a host-side stand-in in host
// code, not an address a kernel can use); resolved once per device
static int mappm_static_ws(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "mappm_static: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxStaticDevices) return set_err(B200CS_EINVAL, "mappm_static: device %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_static_ws_mutex);
    if (!g_static_ws_ptr[dev]) {
        float* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_mappm_static_ws);
        if (e != cudaSuccess || !ptr)
            return set_err(B200CS_ECUDA, "mappm_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_static_ws_ptr[dev] = ptr;
    }
    *out = g_static_ws_ptr[dev];
    return B200CS_OK;
}
}  // namespace b200cs

extern "C" int b200cs_mappm_static(const float* pe1, int64_t pe1_cs, int64_t pe1_ls, const float* q1, int64_t q1_cs,
                                   int64_t q1_ls, const float* pe2, int64_t pe2_cs, int64_t pe2_ls, float* q2,
                                   int64_t q2_cs, int64_t q2_ls, int64_t ncol, int km, int kn, int iv, int kord,
                                   void* stream) {
    float* ws = nullptr;
    int64_t need = 0;
    if (kord > 6 && ncol > 0 && km > 0) {
        need = b200cs_mappm_workspace_bytes(ncol, km);
        if (need > (int64_t)B200CS_MAPPM_STATIC_MAX_BYTES)
            return set_err(B200CS_EWORKSPACE,
                           "mappm_static: ncol=%lld, km=%d needs %lld workspace bytes, the static array holds %lld "
                           "(use b200cs_mappm with a caller-provided workspace)",
                           (long long)ncol, km, (long long)need, (long long)B200CS_MAPPM_STATIC_MAX_BYTES);
        const int rc = mappm_static_ws(&ws);
        if (rc != B200CS_OK) return rc;
    }
    return b200cs_mappm(pe1, pe1_cs, pe1_ls, q1, q1_cs, q1_ls, pe2, pe2_cs, pe2_ls, q2, q2_cs, q2_ls, ncol, km, kn, iv,
                        kord, ws, need, stream);
}
