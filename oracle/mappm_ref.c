/*
 * oracle/mappm_ref.c -- TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * CPU restatement, in C99 single precision, of the reference's only native code on
 * the coarse-graining path: external/mappm/mappm/mappm.f90 (mappm, ppm_profile,
 * ppm_limiters, cs_profile, cs_limiters).  The f2py build of that file uses default
 * `real` = fp32 (external/mappm/setup.py:31, no -fdefault-real-8), so everything
 * here is `float`, constants are fp32 literals (r3 = 1./3. evaluated in fp32), and
 * the operation ORDER of every expression follows the Fortran text (left-to-right,
 * parentheses honoured).  Build with -O2 -ffp-contract=off (no FMA) -- see
 * oracle/Makefile.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.
 *
 * Parity pinning: kord=1, iv=1 is pinned by external/vcm/tests/test_mappm.py:5-44
 * (three known answers) and by the coarsen_restarts regression JSONs
 * (external/vcm/tests/_coarsen_restarts_regression_tests/reference/pressure-level-*).
 * Every other (kord, iv) combination, i.e. limiter modes 1-2, the Huynh branch and
 * all of cs_profile / cs_limiters, is PARITY UNPINNED: no reference test exercises
 * it; the Fortran text alone defines truth.
 *
 * Defined behaviour where the reference has undefined behaviour:
 *   - mappm.f90:70-97: if the level search `do 45` finds no layer (non-monotone or
 *     NaN pressures) the Fortran falls through with k1/qsum uninitialised.  Here the
 *     output layer becomes NaN and the column stays "poisoned": every later output
 *     layer of that column is NaN as well (a forward-only sweep can implement this).
 *   - cs_profile iv=-2 reads qs uninitialised when called from mappm
 *     (mappm.f90:33,49,168-170); here qs = 0.
 *   - km < 4 (kord<=7) / km < 5 (kord>7) index out of range in the Fortran; rejected
 *     with return code -1.
 *
 * MIN/MAX follow gfortran's inline expansion (ignore a NaN operand when the other is
 * a number), which is fminf/fmaxf.  SIGN(a,b) = copysignf(|a|,b), with a NaN b taken as positive.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline float min2(float a, float b) { return fminf(a, b); }
static inline float max2(float a, float b) { return fmaxf(a, b); }
static inline float min3(float a, float b, float c) { return fminf(fminf(a, b), c); }
static inline float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }
/* SIGN(a,b); a NaN b counts as positive (gfortran copies b's sign BIT, which for a NaN depends on
 * how the NaN arose and on the platform -- defined here so that CPU and GPU agree) */
static inline float sgn(float a, float b) { return (b != b) ? fabsf(a) : copysignf(fabsf(a), b); }

/* 1-based accessors onto per-column arrays of length km (or km+1) */
#define A1(k) a1_[(k)-1]
#define AL(k) al_[(k)-1]
#define AR(k) ar_[(k)-1]
#define A6(k) a6_[(k)-1]
#define DP(k) dp_[(k)-1]

/* ppm_limiters, mappm.f90:854-931, for ONE column entry */
static void ppm_limiter1(float dm, float a1, float *al, float *ar, float *a6, int lmt)
{
    const float r12 = 1.f / 12.f;
    if (lmt == 3) return;
    if (lmt == 0) { /* :875-894 */
        if (dm == 0.f) {
            *al = a1; *ar = a1; *a6 = 0.f;
        } else {
            float da1 = *ar - *al;
            float da2 = da1 * da1;
            float a6da = *a6 * da1;
            if (a6da < -da2) {
                *a6 = 3.f * (*al - a1);
                *ar = *al - *a6;
            } else if (a6da > da2) {
                *a6 = 3.f * (*ar - a1);
                *al = *ar - *a6;
            }
        }
    } else if (lmt == 1) { /* :896-905 */
        float qmp = 2.f * dm;
        *al = a1 - sgn(min2(fabsf(qmp), fabsf(*al - a1)), qmp);
        *ar = a1 + sgn(min2(fabsf(qmp), fabsf(*ar - a1)), qmp);
        *a6 = 3.f * (2.f * a1 - (*al + *ar));
    } else if (lmt == 2) { /* :907-927 */
        if (fabsf(*ar - *al) < -(*a6)) {
            float d = *ar - *al;
            float fmin_ = a1 + 0.25f * (d * d) / *a6 + *a6 * r12;
            if (fmin_ < 0.f) {
                if (a1 < *ar && a1 < *al) {
                    *ar = a1; *al = a1; *a6 = 0.f;
                } else if (*ar > *al) {
                    *a6 = 3.f * (*al - a1);
                    *ar = *al - *a6;
                } else {
                    *a6 = 3.f * (*ar - a1);
                    *al = *ar - *a6;
                }
            }
        }
    }
}

/* cs_limiters, mappm.f90:535-611, ONE column entry */
static void cs_limiter1(int extm, float a1, float *al, float *ar, float *a6, int iv)
{
    const float r12 = 1.f / 12.f;
    if (iv == 0) { /* :545-570 */
        if (a1 <= 0.f) {
            *al = a1; *ar = a1; *a6 = 0.f;
        } else if (fabsf(*ar - *al) < -(*a6)) {
            float d = *ar - *al;
            if ((a1 + 0.25f * (d * d) / *a6 + *a6 * r12) < 0.f) {
                if (a1 < *ar && a1 < *al) {
                    *ar = a1; *al = a1; *a6 = 0.f;
                } else if (*ar > *al) {
                    *a6 = 3.f * (*al - a1);
                    *ar = *al - *a6;
                } else {
                    *a6 = 3.f * (*ar - a1);
                    *al = *ar - *a6;
                }
            }
        }
    } else if (iv == 1) { /* :571-589 */
        if ((a1 - *al) * (a1 - *ar) >= 0.f) {
            *al = a1; *ar = a1; *a6 = 0.f;
        } else {
            float da1 = *ar - *al;
            float da2 = da1 * da1;
            float a6da = *a6 * da1;
            if (a6da < -da2) {
                *a6 = 3.f * (*al - a1);
                *ar = *al - *a6;
            } else if (a6da > da2) {
                *a6 = 3.f * (*ar - a1);
                *al = *ar - *a6;
            }
        }
    } else { /* :590-610 */
        if (extm) {
            *al = a1; *ar = a1; *a6 = 0.f;
        } else {
            float da1 = *ar - *al;
            float da2 = da1 * da1;
            float a6da = *a6 * da1;
            if (a6da < -da2) {
                *a6 = 3.f * (*al - a1);
                *ar = *al - *a6;
            } else if (a6da > da2) {
                *a6 = 3.f * (*ar - a1);
                *al = *ar - *a6;
            }
        }
    }
}

/* ppm_profile, mappm.f90:614-851, one column.  work = 5*km floats. */
static void ppm_profile1(const float *a1_, float *al_, float *ar_, float *a6_, const float *dp_,
                         int km, int iv, int kord, float *work)
{
    float *dc_ = work, *h2_ = work + km, *delq_ = work + 2 * km, *df2_ = work + 3 * km,
          *d4_ = work + 4 * km;
#define DC(k) dc_[(k)-1]
#define H2(k) h2_[(k)-1]
#define DELQ(k) delq_[(k)-1]
#define DF2(k) df2_[(k)-1]
#define D4(k) d4_[(k)-1]
    const int km1 = km - 1;
    int k;
    for (k = 2; k <= km; k++) { /* :651-656 */
        DELQ(k - 1) = A1(k) - A1(k - 1);
        D4(k) = DP(k - 1) + DP(k);
    }
    for (k = 2; k <= km1; k++) { /* :658-668 */
        float c1 = (DP(k - 1) + 0.5f * DP(k)) / D4(k + 1);
        float c2 = (DP(k + 1) + 0.5f * DP(k)) / D4(k);
        DF2(k) = DP(k) * (c1 * DELQ(k) + c2 * DELQ(k - 1)) / (D4(k) + DP(k + 1));
        DC(k) = sgn(min3(fabsf(DF2(k)), max3(A1(k - 1), A1(k), A1(k + 1)) - A1(k),
                         A1(k) - min3(A1(k - 1), A1(k), A1(k + 1))),
                    DF2(k));
    }
    for (k = 3; k <= km1; k++) { /* :674-683 */
        float c1 = DELQ(k - 1) * DP(k - 1) / D4(k);
        float a1 = D4(k - 1) / (D4(k) + DP(k - 1));
        float a2 = D4(k + 1) / (D4(k) + DP(k));
        AL(k) = A1(k - 1) + c1 +
                2.f / (D4(k - 1) + D4(k + 1)) *
                    (DP(k) * (c1 * (a1 - a2) + a2 * DC(k - 1)) - DP(k - 1) * a1 * DC(k));
    }
    { /* top, :689-707 */
        float d1 = DP(1), d2 = DP(2);
        float qm = (d2 * A1(1) + d1 * A1(2)) / (d1 + d2);
        float dq = 2.f * (A1(2) - A1(1)) / (d1 + d2);
        float c1 = 4.f * (AL(3) - qm - d2 * dq) / (d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1)));
        float c3 = dq - 0.5f * c1 * (d2 * (5.f * d1 + d2) - 3.f * d1 * d1);
        AL(2) = qm - 0.25f * c1 * d1 * d2 * (d2 + 3.f * d1);
        AL(1) = d1 * (2.f * c1 * (d1 * d1) - c3) + AL(2);
        AL(2) = max2(AL(2), min2(A1(1), A1(2)));
        AL(2) = min2(AL(2), max2(A1(1), A1(2)));
        DC(1) = 0.5f * (AL(2) - A1(1));
    }
    if (iv == 0) { /* :711-725 */
        AL(1) = max2(0.f, AL(1));
        AL(2) = max2(0.f, AL(2));
    } else if (iv == -1) {
        if (AL(1) * A1(1) <= 0.f) AL(1) = 0.f;
    } else if (abs(iv) == 2) {
        AL(1) = A1(1);
        AR(1) = A1(1);
    }
    { /* bottom, :729-747 */
        float d1 = DP(km), d2 = DP(km1);
        float qm = (d2 * A1(km) + d1 * A1(km1)) / (d1 + d2);
        float dq = 2.f * (A1(km1) - A1(km)) / (d1 + d2);
        float c1 = (AL(km1) - qm - d2 * dq) / (d2 * (2.f * d2 * d2 + d1 * (d2 + 3.f * d1)));
        float c3 = dq - 2.0f * c1 * (d2 * (5.f * d1 + d2) - 3.f * d1 * d1);
        AL(km) = qm - c1 * d1 * d2 * (d2 + 3.f * d1);
        AR(km) = d1 * (8.f * c1 * (d1 * d1) - c3) + AL(km);
        AL(km) = max2(AL(km), min2(A1(km), A1(km1)));
        AL(km) = min2(AL(km), max2(A1(km), A1(km1)));
        DC(km) = 0.5f * (A1(km) - AL(km));
    }
    if (iv == 0) { /* :752-761 */
        AL(km) = max2(0.f, AL(km));
        AR(km) = max2(0.f, AR(km));
    } else if (iv < 0) {
        if (A1(km) * AR(km) <= 0.f) AR(km) = 0.f;
    }
    for (k = 1; k <= km1; k++) AR(k) = AL(k + 1); /* :763-767 */

    for (k = 1; k <= 2; k++) { /* :773-778 */
        A6(k) = 3.f * (2.f * A1(k) - (AL(k) + AR(k)));
        ppm_limiter1(DC(k), A1(k), &AL(k), &AR(k), &A6(k), 0);
    }
    if (kord >= 7) { /* Huynh, :780-826 */
        for (k = 2; k <= km1; k++) {
            H2(k) = 2.f * (DC(k + 1) / DP(k + 1) - DC(k - 1) / DP(k - 1)) /
                    (DP(k) + 0.5f * (DP(k - 1) + DP(k + 1))) * (DP(k) * DP(k));
        }
        const float fac = 1.5f;
        for (k = 3; k <= km - 2; k++) {
            float pmp = 2.f * DC(k);
            float qmp = A1(k) + pmp;
            float lac = A1(k) + fac * H2(k - 1) + DC(k);
            AR(k) = min2(max2(AR(k), min3(A1(k), qmp, lac)), max3(A1(k), qmp, lac));
            qmp = A1(k) - pmp;
            lac = A1(k) + fac * H2(k + 1) - DC(k);
            AL(k) = min2(max2(AL(k), min3(A1(k), qmp, lac)), max3(A1(k), qmp, lac));
            A6(k) = 3.f * (2.f * A1(k) - (AL(k) + AR(k)));
            if (iv == 0 && kord >= 6) ppm_limiter1(DC(k), A1(k), &AL(k), &AR(k), &A6(k), 2);
        }
    } else { /* :828-842 */
        int lmt = kord - 3;
        if (lmt < 0) lmt = 0;
        if (iv == 0 && lmt > 2) lmt = 2;
        for (k = 3; k <= km - 2; k++) {
            if (kord != 4) A6(k) = 3.f * (2.f * A1(k) - (AL(k) + AR(k)));
            if (kord != 6) ppm_limiter1(DC(k), A1(k), &AL(k), &AR(k), &A6(k), lmt);
        }
    }
    for (k = km1; k <= km; k++) { /* :844-849 */
        A6(k) = 3.f * (2.f * A1(k) - (AL(k) + AR(k)));
        ppm_limiter1(DC(k), A1(k), &AL(k), &AR(k), &A6(k), 0);
    }
#undef DC
#undef H2
#undef DELQ
#undef DF2
#undef D4
}

/* cs_profile, mappm.f90:132-532, one column.  work = (2*km+1) floats + 3*km ints. */
static void cs_profile1(float qs, const float *a1_, float *al_, float *ar_, float *a6_,
                        const float *dp_, int km, int iv, int kord, float *work, int *iwork)
{
    float *gam_ = work, *q_ = work + km;
    int *extm_ = iwork, *ext5_ = iwork + km, *ext6_ = iwork + 2 * km;
#define GAM(k) gam_[(k)-1]
#define Q(k) q_[(k)-1]
#define EXTM(k) extm_[(k)-1]
#define EXT5(k) ext5_[(k)-1]
#define EXT6(k) ext6_[(k)-1]
    int k;
    const int akord = abs(kord);
    float d4 = 0.f, bet, grat;
    memset(iwork, 0, sizeof(int) * 3 * (size_t)km);
    if (iv == -2) { /* :153-176 */
        GAM(2) = 0.5f;
        Q(1) = 1.5f * A1(1);
        for (k = 2; k <= km - 1; k++) {
            grat = DP(k - 1) / DP(k);
            bet = 2.f + grat + grat - GAM(k);
            Q(k) = (3.f * (A1(k - 1) + A1(k)) - Q(k - 1)) / bet;
            GAM(k + 1) = grat / bet;
        }
        grat = DP(km - 1) / DP(km);
        Q(km) = (3.f * (A1(km - 1) + A1(km)) - grat * qs - Q(km - 1)) /
                (2.f + grat + grat - GAM(km));
        Q(km + 1) = qs;
        for (k = km - 1; k >= 1; k--) Q(k) = Q(k) - GAM(k + 1) * Q(k + 1);
    } else { /* :178-204 */
        grat = DP(2) / DP(1);
        bet = grat * (grat + 0.5f);
        Q(1) = ((grat + grat) * (grat + 1.f) * A1(1) + A1(2)) / bet;
        GAM(1) = (1.f + grat * (grat + 1.5f)) / bet;
        for (k = 2; k <= km; k++) {
            d4 = DP(k - 1) / DP(k);
            bet = 2.f + d4 + d4 - GAM(k - 1);
            Q(k) = (3.f * (A1(k - 1) + d4 * A1(k)) - Q(k - 1)) / bet;
            GAM(k) = d4 / bet;
        }
        {
            float a_bot = 1.f + d4 * (d4 + 1.5f);
            Q(km + 1) = (2.f * d4 * (d4 + 1.f) * A1(km) + A1(km - 1) - a_bot * Q(km)) /
                        (d4 * (d4 + 0.5f) - a_bot * GAM(km));
        }
        for (k = km; k >= 1; k--) Q(k) = Q(k) - GAM(k) * Q(k + 1);
    }
    if (akord > 16) { /* perfectly linear, :207-216 */
        for (k = 1; k <= km; k++) {
            AL(k) = Q(k);
            AR(k) = Q(k + 1);
            A6(k) = 3.f * (2.f * A1(k) - (AL(k) + AR(k)));
        }
        return;
    }
    /* large-scale constraints, :225-260 */
    Q(2) = min2(Q(2), max2(A1(1), A1(2)));
    Q(2) = max2(Q(2), min2(A1(1), A1(2)));
    for (k = 2; k <= km; k++) GAM(k) = A1(k) - A1(k - 1);
    for (k = 3; k <= km - 1; k++) {
        if (GAM(k - 1) * GAM(k + 1) > 0.f) {
            Q(k) = min2(Q(k), max2(A1(k - 1), A1(k)));
            Q(k) = max2(Q(k), min2(A1(k - 1), A1(k)));
        } else if (GAM(k - 1) > 0.f) {
            Q(k) = max2(Q(k), min2(A1(k - 1), A1(k)));
        } else {
            Q(k) = min2(Q(k), max2(A1(k - 1), A1(k)));
            if (iv == 0) Q(k) = max2(0.f, Q(k));
        }
    }
    Q(km) = min2(Q(km), max2(A1(km - 1), A1(km)));
    Q(km) = max2(Q(km), min2(A1(km - 1), A1(km)));
    for (k = 1; k <= km; k++) { /* :262-267 */
        AL(k) = Q(k);
        AR(k) = Q(k + 1);
    }
    for (k = 1; k <= km; k++) { /* :269-288 */
        if (k == 1 || k == km)
            EXTM(k) = (AL(k) - A1(k)) * (AR(k) - A1(k)) > 0.f;
        else
            EXTM(k) = GAM(k) * GAM(k + 1) < 0.f;
        if (akord > 9) {
            float x0 = 2.f * A1(k) - (AL(k) + AR(k));
            float x1 = fabsf(AL(k) - AR(k));
            A6(k) = 3.f * x0;
            EXT5(k) = fabsf(x0) > x1;
            EXT6(k) = fabsf(A6(k)) > x1;
        }
    }
    /* top layer, :296-317 */
    if (iv == 0) {
        AL(1) = max2(0.f, AL(1));
    } else if (iv == -1) {
        if (AL(1) * A1(1) <= 0.f) AL(1) = 0.f;
    } else if (iv == 2) {
        AL(1) = A1(1);
        AR(1) = A1(1);
        A6(1) = 0.f;
    }
    if (iv != 2) {
        A6(1) = 3.f * (2.f * A1(1) - (AL(1) + AR(1)));
        cs_limiter1(EXTM(1), A1(1), &AL(1), &AR(1), &A6(1), 1);
    }
    A6(2) = 3.f * (2.f * A1(2) - (AL(2) + AR(2))); /* :320-323 */
    cs_limiter1(EXTM(2), A1(2), &AL(2), &AR(2), &A6(2), 2);

#define HUYNH_EDGES(k)                                                                   \
    do {                                                                                 \
        float pmp_1 = A1(k) - 2.f * GAM(k + 1);                                          \
        float lac_1 = pmp_1 + 1.5f * GAM(k + 2);                                         \
        AL(k) = min2(max2(AL(k), min3(A1(k), pmp_1, lac_1)), max3(A1(k), pmp_1, lac_1)); \
        float pmp_2 = A1(k) + 2.f * GAM(k);                                              \
        float lac_2 = pmp_2 - 1.5f * GAM(k - 1);                                         \
        AR(k) = min2(max2(AR(k), min3(A1(k), pmp_2, lac_2)), max3(A1(k), pmp_2, lac_2)); \
    } while (0)
#define FLAT(k)        \
    do {               \
        AL(k) = A1(k); \
        AR(k) = A1(k); \
    } while (0)
#define STD_A6(k) (3.f * (2.f * A1(k) - (AL(k) + AR(k))))

    for (k = 3; k <= km - 2; k++) { /* :328-509 */
        if (akord < 9) {
            HUYNH_EDGES(k);
            A6(k) = STD_A6(k);
        } else if (akord == 9) {
            if (EXTM(k) && EXTM(k - 1)) {
                FLAT(k);
                A6(k) = 0.f;
            } else if (EXTM(k) && EXTM(k + 1)) {
                FLAT(k);
                A6(k) = 0.f;
            } else {
                A6(k) = 6.f * A1(k) - 3.f * (AL(k) + AR(k));
                if (fabsf(A6(k)) > fabsf(AL(k) - AR(k))) {
                    HUYNH_EDGES(k);
                    A6(k) = 6.f * A1(k) - 3.f * (AL(k) + AR(k));
                }
            }
        } else if (akord == 10) {
            if (EXT5(k)) {
                if (EXT5(k - 1) || EXT5(k + 1)) {
                    FLAT(k);
                } else if (EXT6(k - 1) || EXT6(k + 1)) {
                    HUYNH_EDGES(k);
                }
            } else if (EXT6(k)) {
                if (EXT5(k - 1) || EXT5(k + 1)) HUYNH_EDGES(k);
            }
            A6(k) = STD_A6(k);
        } else if (akord == 12) {
            if (EXTM(k)) {
                FLAT(k);
                A6(k) = 0.f;
            } else {
                A6(k) = 6.f * A1(k) - 3.f * (AL(k) + AR(k));
                if (fabsf(A6(k)) > fabsf(AL(k) - AR(k))) {
                    HUYNH_EDGES(k);
                    A6(k) = 6.f * A1(k) - 3.f * (AL(k) + AR(k));
                }
            }
        } else if (akord == 13) {
            if (EXT6(k)) {
                if (EXT6(k - 1) && EXT6(k + 1)) FLAT(k);
            }
            A6(k) = STD_A6(k);
        } else if (akord == 14) {
            A6(k) = STD_A6(k);
        } else if (akord == 15) {
            if (EXT5(k)) {
                if (EXT5(k - 1) || EXT5(k + 1)) FLAT(k);
            } else if (EXT6(k)) {
                HUYNH_EDGES(k);
            }
            A6(k) = STD_A6(k);
        } else if (akord == 16) {
            if (EXT5(k)) {
                if (EXT5(k - 1) || EXT5(k + 1)) {
                    FLAT(k);
                } else if (EXT6(k - 1) || EXT6(k + 1)) {
                    HUYNH_EDGES(k);
                }
            }
            A6(k) = STD_A6(k);
        } else { /* kord = 11 */
            if (EXT5(k) && (EXT5(k - 1) || EXT5(k + 1))) {
                FLAT(k);
                A6(k) = 0.f;
            } else {
                A6(k) = STD_A6(k);
            }
        }
        if (iv == 0) cs_limiter1(EXTM(k), A1(k), &AL(k), &AR(k), &A6(k), 0);
    }
    /* bottom, :514-530 */
    if (iv == 0) {
        AR(km) = max2(0.f, AR(km));
    } else if (iv == -1) {
        if (AR(km) * A1(km) <= 0.f) AR(km) = 0.f;
    }
    for (k = km - 1; k <= km; k++) {
        A6(k) = STD_A6(k);
        if (k == km - 1) cs_limiter1(EXTM(k), A1(k), &AL(k), &AR(k), &A6(k), 2);
        if (k == km) cs_limiter1(EXTM(k), A1(k), &AL(k), &AR(k), &A6(k), 1);
    }
#undef GAM
#undef Q
#undef EXTM
#undef EXT5
#undef EXT6
#undef HUYNH_EDGES
#undef FLAT
#undef STD_A6
}

/* mappm main sweep, mappm.f90:41-124, one column.  pe1_[km+1], pe2_[kn+1]. */
static void remap1(const float *pe1_, const float *a1_, const float *al_, const float *ar_,
                   const float *a6_, const float *dp_, int km, const float *pe2_, float *q2_,
                   int kn)
{
#define PE1(k) pe1_[(k)-1]
#define PE2(k) pe2_[(k)-1]
#define Q2(k) q2_[(k)-1]
    const float r3 = 1.f / 3.f, r23 = 2.f / 3.f;
    int k0 = 1, poisoned = 0;
    for (int k = 1; k <= kn; k++) {
        if (PE2(k) <= PE1(1)) { /* :62-64 */
            Q2(k) = A1(1);
            continue;
        }
        if (PE2(k) >= PE1(km + 1)) { /* :65-67 */
            Q2(k) = A1(km);
            continue;
        }
        if (poisoned) {
            Q2(k) = NAN;
            continue;
        }
        int L, found = 0, done = 0, k1 = 0;
        float qsum = 0.f, dpsum = 0.f;
        for (L = k0; L <= km; L++) { /* :70-95 */
            if (PE2(k) >= PE1(L) && PE2(k) <= PE1(L + 1)) {
                found = 1;
                k0 = L;
                float PL = (PE2(k) - PE1(L)) / DP(L);
                if (PE2(k + 1) <= PE1(L + 1)) {
                    float PR = (PE2(k + 1) - PE1(L)) / DP(L);
                    float TT = r3 * (PR * (PR + PL) + PL * PL);
                    Q2(k) = AL(L) + 0.5f * (A6(L) + AR(L) - AL(L)) * (PR + PL) - A6(L) * TT;
                    done = 1;
                } else {
                    float delp = PE1(L + 1) - PE2(k);
                    float TT = r3 * (1.f + PL * (1.f + PL));
                    qsum = delp * (AL(L) + 0.5f * (A6(L) + AR(L) - AL(L)) * (1.f + PL) - A6(L) * TT);
                    dpsum = delp;
                    k1 = L + 1;
                }
                break;
            }
        }
        if (done) continue;
        if (!found) { /* reference: undefined behaviour (see header) */
            poisoned = 1;
            Q2(k) = NAN;
            continue;
        }
        int closed = 0;
        for (L = k1; L <= km; L++) { /* :98-114 */
            if (PE2(k + 1) > PE1(L + 1)) {
                qsum = qsum + DP(L) * A1(L);
                dpsum = dpsum + DP(L);
            } else {
                float delp = PE2(k + 1) - PE1(L);
                float esl = delp / DP(L);
                qsum = qsum + delp * (AL(L) + 0.5f * esl * (AR(L) - AL(L) + A6(L) * (1.f - r23 * esl)));
                dpsum = dpsum + delp;
                k0 = L;
                closed = 1;
                break;
            }
        }
        if (!closed) { /* :115-120 */
            float delp = PE2(k + 1) - PE1(km + 1);
            if (delp > 0.f) {
                qsum = qsum + delp * A1(km);
                dpsum = dpsum + delp;
            }
        }
        Q2(k) = qsum / dpsum; /* :121 */
    }
#undef PE1
#undef PE2
#undef Q2
}

/*
 * C-order column-major-free entry point mirroring the f2py signature
 *   q2 = mappm.mappm(pe1[ncol,km+1], q1[ncol,km], pe2[ncol,kn+1], i1, i2, iv, kord, ptop)
 * (external/vcm/vcm/cubedsphere/regridz.py:332-334).  ptop is unused as in the Fortran.
 * nthreads <= 0 -> all OpenMP threads.  Returns 0, or -1 for unsupported km.
 */
int mappm_ref(int ncol, int km, const float *pe1, const float *q1, int kn, const float *pe2,
              float *q2, int iv, int kord, int nthreads)
{
    if (km < 4 || (kord > 7 && km < 5) || kn < 1) return -1;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
#pragma omp parallel num_threads(nthreads)
    {
        float *buf = (float *)malloc(sizeof(float) * (size_t)(10 * km + 8));
        int *ibuf = (int *)malloc(sizeof(int) * (size_t)(3 * km + 3));
        float *al_ = buf, *ar_ = buf + km, *a6_ = buf + 2 * km, *dp_ = buf + 3 * km,
              *work = buf + 4 * km;
#pragma omp for schedule(static)
        for (int i = 0; i < ncol; i++) {
            const float *pe1_ = pe1 + (size_t)i * (km + 1);
            const float *a1_ = q1 + (size_t)i * km;
            for (int k = 1; k <= km; k++) DP(k) = pe1_[k] - pe1_[k - 1]; /* :41-46 */
            if (kord > 7)
                cs_profile1(0.f, a1_, al_, ar_, a6_, dp_, km, iv, kord, work, ibuf);
            else
                ppm_profile1(a1_, al_, ar_, a6_, dp_, km, iv, kord, work);
            remap1(pe1_, a1_, al_, ar_, a6_, dp_, km, pe2 + (size_t)i * (kn + 1),
                   q2 + (size_t)i * kn, kn);
        }
        free(buf);
        free(ibuf);
    }
    return 0;
}

/* expose the profile alone (AL, AR, A6 as [ncol,km] arrays) for profile-level tests */
int mappm_profile_ref(int ncol, int km, const float *pe1, const float *q1, float *al, float *ar,
                      float *a6, int iv, int kord)
{
    if (km < 4 || (kord > 7 && km < 5)) return -1;
    float *buf = (float *)malloc(sizeof(float) * (size_t)(7 * km + 8));
    int *ibuf = (int *)malloc(sizeof(int) * (size_t)(3 * km + 3));
    float *dp_ = buf, *work = buf + km;
    for (int i = 0; i < ncol; i++) {
        const float *pe1_ = pe1 + (size_t)i * (km + 1);
        const float *a1_ = q1 + (size_t)i * km;
        float *al_ = al + (size_t)i * km, *ar_ = ar + (size_t)i * km, *a6_ = a6 + (size_t)i * km;
        for (int k = 1; k <= km; k++) DP(k) = pe1_[k] - pe1_[k - 1];
        if (kord > 7)
            cs_profile1(0.f, a1_, al_, ar_, a6_, dp_, km, iv, kord, work, ibuf);
        else
            ppm_profile1(a1_, al_, ar_, a6_, dp_, km, iv, kord, work);
    }
    free(buf);
    free(ibuf);
    return 0;
}
