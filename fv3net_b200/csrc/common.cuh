// common.cuh -- shared helpers of libb200cs (status handling, dtype promotion, numpy-order sums)
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <type_traits>

#include "../../include/b200cs.h"

namespace b200cs {

// thread-local last-error text (b200cs_last_error)
char* err_buf();
int set_err(int code, const char* fmt, ...);
int check_launch(const char* what);

#define B200CS_REQUIRE(cond, ...)                                  \
    do {                                                           \
        if (!(cond)) return b200cs::set_err(B200CS_EINVAL, __VA_ARGS__); \
    } while (0)

template <typename A, typename B>
struct Promote {
    using type = typename std::conditional<std::is_same<A, double>::value || std::is_same<B, double>::value,
                                           double, float>::type;
};

template <typename T>
__device__ __forceinline__ T nan_to_zero(T v) {
    return (v != v) ? T(0) : v;
}

// numpy pairwise_sum order for one contiguous window row of n <= 128 elements
// (numpy/core/src/umath/loops_utils.h.src; oracle/blocksum_ref.c)
template <typename T, class Load>
__device__ __forceinline__ T rowsum_numpy(int n, Load ld) {
    if (n < 8) {
        T r = T(0);
        for (int i = 0; i < n; i++) r += ld(i);
        return r;
    }
    T r0 = ld(0), r1 = ld(1), r2 = ld(2), r3 = ld(3), r4 = ld(4), r5 = ld(5), r6 = ld(6), r7 = ld(7);
    int i;
    for (i = 8; i < n - (n % 8); i += 8) {
        r0 += ld(i); r1 += ld(i + 1); r2 += ld(i + 2); r3 += ld(i + 3);
        r4 += ld(i + 4); r5 += ld(i + 5); r6 += ld(i + 6); r7 += ld(i + 7);
    }
    T res = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
    for (; i < n; i++) res += ld(i);
    return res;
}

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

}  // namespace b200cs
