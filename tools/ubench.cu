// ubench.cu -- issue-rate microbenchmarks for the instruction mix of the PPM remap kernel
// (not product code; results recorded in profiles/ubench_r01.txt and DESIGN.md).
// Each kernel runs ITER iterations of 16 independent ops per thread; reports cycles per
// warp-instruction per SM sub-partition (4 SMSPs/SM), with 8 warps per SMSP resident.
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 2048
#define NACC 16

#define KERNEL(name, DECL, BODY, FIN)                                            \
    __global__ void __launch_bounds__(256) name(float* out, float s, float t) {  \
        DECL;                                                                    \
        for (int it = 0; it < ITER; it++) {                                      \
            _Pragma("unroll") for (int j = 0; j < NACC; j++) { BODY; }           \
        }                                                                        \
        float r = 0;                                                             \
        _Pragma("unroll") for (int j = 0; j < NACC; j++) { FIN; }                \
        if (r == 12345.678f) out[threadIdx.x] = r;                               \
    }

#define DECLF float a[NACC]; _Pragma("unroll") for (int j = 0; j < NACC; j++) a[j] = threadIdx.x * 0.001f + j
#define DECLF2 float2 a[NACC]; _Pragma("unroll") for (int j = 0; j < NACC; j++) a[j] = make_float2(threadIdx.x * 0.001f + j, j * 0.5f)

KERNEL(k_ffma, DECLF, a[j] = fmaf(a[j], s, t), r += a[j])
KERNEL(k_fadd, DECLF, a[j] = a[j] + s, r += a[j])
KERNEL(k_fmul, DECLF, a[j] = a[j] * s, r += a[j])
KERNEL(k_fmnmx, DECLF, a[j] = fminf(a[j], a[(j + 1) % NACC] + 0.f) , r += a[j])
KERNEL(k_fmnmx_only, DECLF, a[j] = fminf(a[j], a[(j + 5) % NACC]), r += a[j])
KERNEL(k_fmnmx3, DECLF, a[j] = fminf(fminf(a[j], a[(j + 5) % NACC]), a[(j + 3) % NACC]), r += a[j])
KERNEL(k_rcp, DECLF, { float v; asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(v) : "f"(a[j])); a[j] = v + s; }, r += a[j])
KERNEL(k_div, DECLF, a[j] = s / a[j], r += a[j])
KERNEL(k_fdividef, DECLF, a[j] = __fdividef(s, a[j]), r += a[j])
KERNEL(k_setp_sel, DECLF, a[j] = (a[j] > s) ? t : a[j] , r += a[j])
KERNEL(k_copysign, DECLF, a[j] = copysignf(a[(j + 5) % NACC], a[j]) , r += a[j])
KERNEL(k_shfl, DECLF, a[j] = __shfl_xor_sync(0xffffffffu, a[j], 1), r += a[j])
KERNEL(k_ffma2, DECLF2,
       asm volatile("{.reg .b64 x, y, z; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %2}; mov.b64 z, {%3, %3};"
                    "fma.rn.f32x2 x, x, y, z; mov.b64 {%0, %1}, x;}" : "+f"(a[j].x), "+f"(a[j].y) : "f"(s), "f"(t)),
       r += a[j].x + a[j].y)
KERNEL(k_fadd2, DECLF2,
       asm volatile("{.reg .b64 x, y; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %2};"
                    "add.rn.f32x2 x, x, y; mov.b64 {%0, %1}, x;}" : "+f"(a[j].x), "+f"(a[j].y) : "f"(s)),
       r += a[j].x + a[j].y)
// mixes: one FMA-pipe op and one ALU-pipe op per accumulator
KERNEL(k_mix_ffma_fmnmx, DECLF, a[j] = fminf(fmaf(a[j], s, t), t), r += a[j])
KERNEL(k_mix_ffma_setp, DECLF, { float v = fmaf(a[j], s, t); a[j] = (v > s) ? t : v; }, r += a[j])
KERNEL(k_mix_ffma2_fmnmx, DECLF2,
       asm volatile("{.reg .b64 x, y, z; mov.b64 x, {%0, %1}; mov.b64 y, {%2, %2}; mov.b64 z, {%3, %3};"
                    "fma.rn.f32x2 x, x, y, z; mov.b64 {%0, %1}, x; min.f32 %0, %0, %3; min.f32 %1, %1, %3;}"
                    : "+f"(a[j].x), "+f"(a[j].y) : "f"(s), "f"(t)),
       r += a[j].x + a[j].y)
KERNEL(k_mix_ffma_rcp4, DECLF, { a[j] = fmaf(a[j], s, t); if ((j & 3) == 0) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[j])); }, r += a[j])

__global__ void __launch_bounds__(256) k_lds(float* out, float s, float t) {
    __shared__ float sm[256 * 17];
    for (int j = 0; j < 17; j++) sm[threadIdx.x + 256 * j] = threadIdx.x + j;
    __syncthreads();
    float a[NACC];
#pragma unroll
    for (int j = 0; j < NACC; j++) a[j] = 0;
    int idx = threadIdx.x;
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int j = 0; j < NACC; j++) a[j] += sm[idx + 256 * j];
        idx = (idx + (int)s) & 255;
    }
    float r = 0;
#pragma unroll
    for (int j = 0; j < NACC; j++) r += a[j];
    if (r == 12345.678f) out[threadIdx.x] = r;
}
__global__ void __launch_bounds__(256) k_sts128(float* out, float s, float t) {
    __shared__ float4 sm[256 * 4];
    float4 v = make_float4(s, t, s, t);
    int idx = threadIdx.x;
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int j = 0; j < NACC; j++) { sm[idx + 256 * (j & 3)] = v; v.x += s; }
        idx = (idx + (int)s) & 255;
    }
    __syncthreads();
    if (sm[threadIdx.x].x == 12345.678f) out[threadIdx.x] = v.x;
}

template <class K>
void run(const char* name, K kern, double ops_per_iter) {
    float* out;
    cudaMalloc(&out, 4096);
    int nsm = 148;
    dim3 grid(nsm * 4), block(256);  // 8 CTAs x 8 warps... occupancy permitting: 32 warps/SM = 8/SMSP
    kern<<<grid, block>>>(out, 1.0001f, 0.5f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    kern<<<grid, block>>>(out, 1.0001f, 0.5f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double warp_instr_per_smsp = (double)ITER * ops_per_iter * (4.0 * 8 /*warps per SM*/) / 4.0;
    double cyc = ms * 1e-3 * khz * 1e3;
    printf("%-20s %8.3f ms  %6.3f cyc/warp-op/SMSP (at %d MHz nominal)  err=%s\n", name, ms, cyc / warp_instr_per_smsp,
           khz / 1000, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    run("ffma", k_ffma, NACC); run("fadd", k_fadd, NACC); run("fmul", k_fmul, NACC);
    run("fadd+fmnmx", k_fmnmx, 2 * NACC); run("fmnmx", k_fmnmx_only, NACC); run("fmnmx3", k_fmnmx3, NACC);
    run("mufu.rcp+fadd (pair)", k_rcp, NACC); run("div.rn (per div)", k_div, NACC); run("fdividef (per div)", k_fdividef, NACC);
    run("fsetp+fsel (per pair)", k_setp_sel, NACC); run("copysign", k_copysign, NACC); run("shfl", k_shfl, NACC);
    run("ffma2 (per instr)", k_ffma2, NACC); run("fadd2 (per instr)", k_fadd2, NACC);
    run("ffma+fmnmx (pair)", k_mix_ffma_fmnmx, NACC); run("ffma+setp+sel (tri)", k_mix_ffma_setp, NACC);
    run("ffma2+2fmnmx (tri)", k_mix_ffma2_fmnmx, NACC); run("4ffma+1rcp (per 4ffma)", k_mix_ffma_rcp4, NACC / 4.0);
    run("lds.32", k_lds, NACC); run("sts.128", k_sts128, NACC);
    return 0;
}
