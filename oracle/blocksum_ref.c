/*
 * oracle/blocksum_ref.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Explicit-order restatement of what numpy does for the window sum that xarray 0.19
 * `DataArray.coarsen({y: f, x: f}).sum()` performs (call sites
 * external/vcm/vcm/cubedsphere/coarsen.py:211-212,260-261):
 *     a.reshape(B, ny/fy, fy, nx/fx, fx).sum(axis=(-3, -1))
 * on a C-contiguous array.  numpy reduces the contiguous fx-long axis with its
 * `pairwise_sum` (n < 8: sequential from 0; 8 <= n <= 128: eight strided
 * accumulators, then ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), then the n%8 tail
 * sequentially) and adds the fy row sums sequentially, in row order, onto an
 * output initialised with the identity 0.  The CUDA kernels follow this order so
 * that block sums are bit-identical to the reference's; tests/test_oracle.py checks
 * this file against numpy itself.
 */
#include <stddef.h>

#define DEFINE_ROWSUM(T, NAME)                                                        \
    static T NAME(const T *a, int n)                                                  \
    {                                                                                 \
        if (n < 8) {                                                                  \
            T res = 0;                                                                \
            for (int i = 0; i < n; i++) res += a[i];                                  \
            return res;                                                               \
        } else if (n <= 128) {                                                        \
            T r[8];                                                                   \
            int i;                                                                    \
            for (int j = 0; j < 8; j++) r[j] = a[j];                                  \
            for (i = 8; i < n - (n % 8); i += 8)                                      \
                for (int j = 0; j < 8; j++) r[j] += a[i + j];                         \
            T res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7])); \
            for (; i < n; i++) res += a[i];                                           \
            return res;                                                               \
        } else {                                                                      \
            int n2 = n / 2;                                                           \
            n2 -= n2 % 8;                                                             \
            return NAME(a, n2) + NAME(a + n2, n - n2);                                \
        }                                                                             \
    }

DEFINE_ROWSUM(float, rowsum_f32)
DEFINE_ROWSUM(double, rowsum_f64)

#define DEFINE_BLOCKSUM(T, NAME, ROWSUM)                                                   \
    void NAME(const T *a, long nbatch, int ny, int nx, int fy, int fx, T *out)             \
    {                                                                                      \
        int cy = ny / fy, cx = nx / fx;                                                    \
        for (long b = 0; b < nbatch; b++)                                                  \
            for (int j = 0; j < cy; j++)                                                   \
                for (int i = 0; i < cx; i++) {                                             \
                    T acc = 0;                                                             \
                    for (int r = 0; r < fy; r++)                                           \
                        acc += ROWSUM(a + ((size_t)b * ny + (size_t)j * fy + r) * nx + (size_t)i * fx, fx); \
                    out[((size_t)b * cy + j) * cx + i] = acc;                              \
                }                                                                          \
    }

DEFINE_BLOCKSUM(float, blocksum_ordered_f32, rowsum_f32)
DEFINE_BLOCKSUM(double, blocksum_ordered_f64, rowsum_f64)
