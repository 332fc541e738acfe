// mappm_fast.cu -- K7 in B200CS_FAST arithmetic: the standalone column remap of
// external/mappm/mappm/mappm.f90:10-126 (call site external/vcm/vcm/cubedsphere/regridz.py:329-334)
// on the layer-synchronous machine of ppm_sweep.cuh.  Same level search and the same parabolas as the
// Fortran for the profile every reference caller asks for (kord <= 3, regridz.py:227-228 passes 1);
// only the rounding of coefficients and integrals differs (geometry hoisted, rcp.approx, FMA
// contraction): within 1e-6 relative of b200cs_mappm (BASELINE.json's tolerance, stated in
// tests/test_gpu_parity.py).  One thread per column; with the restart layout [z, ncol] every
// load and store of a warp is one 128-byte line, and each input element is read once.
// Compiled with -fmad=true.
#include "common.cuh"
#include "ppm_sweep.cuh"

namespace b200cs {

// The outer loop runs over source layers and is the same for every column of the warp: every
// lane does the expensive part -- slopes, edge values, limited parabola -- exactly once per
// iteration, and the windows rotate by renaming (four instances of the layer step).  The
// data-dependent part is the inner loop over the output edges that fall into the current source
// layer (edge 0 only opens the first output layer).
//
// Level data travels through per-thread cp.async rings in shared memory, not through registers:
// a register prefetch ring needs more loads in flight than the SM has scoreboards (6), and a
// consumer that waits on a shared scoreboard waits for the youngest load on it -- ncu showed the
// prefetch distance collapsing to one layer (profiles/r01_mappm_fast_ncu.txt).  cp.async groups
// are counted per thread instead:
//   * source ring: (q[L+2], pe[L+3]) for layer L, issued kMsDepth layers ahead, one commit group
//     per layer, cp.async.wait_group kMsDepth before the layer is read;
//   * edge ring: pe2[e+1 .. e+kMsRing-1]; consuming an edge refills the freed slot.  How far the
//     ring is known to have landed (esafe) follows the per-layer groups through a small history
//     of "highest edge issued" (s_hist); a lane that outruns it drains its groups (rare: an
//     output grid much finer than the source).
constexpr int kMsDepth = 6, kMsRing = 8, kMsThreads = 128;

__device__ __forceinline__ void cp_async4(const float* sdst, const void* gsrc) {
    const unsigned sa = (unsigned)__cvta_generic_to_shared(sdst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gsrc) : "memory");
}

struct MsCtx {
    PpmSweep<1> m;
    // running pointers; the level strides are BYTE strides read straight from the constant bank
    const char* qpf;   // next source values to issue: q[L+kMsDepth+2] ...
    const char* ppf;   // ... and pe[L+kMsDepth+3] on entry to layer L
    const char* p2f;   // next output edge to issue: pe2[e+kMsRing-1]
    char* o2;          // output layer e-1
    float2* s_src;                    // + tid; slot stride kMsThreads
    float* s_edge;
    int* s_hist;
    float pt, pb;                     // previous edge, edge e (+inf beyond the last)
    float pe_top, qtop;
    int e, esafe, km, kn;
};

template <bool LAST>
__device__ __forceinline__ void ms_emit(MsCtx& c, const int64_t p2_bs, const int64_t o_bs) {
    while (LAST ? (c.e <= c.kn) : (c.pb <= c.m.pe_hi)) {
        // Edge 0 only opens the first output layer: the same evaluation leaves -D(s) in R, its
        // "output" is not stored (pt = +inf).  While pt <= pe[0] the layer copies the top source
        // value (mappm.f90:62-64); whatever such an edge leaves in R is replaced by the next one.
        float o[1];
        c.m.template edge<LAST>(c.pt, c.pb, fast_rcp(c.pb - c.pt), o);
        if (c.e > 0) *reinterpret_cast<float*>(c.o2) = (c.pt <= c.pe_top) ? c.qtop : o[0];
        c.o2 += o_bs;
        c.pt = c.pb;
        c.e++;
        if (c.e > c.esafe) {  // everything issued so far: edges <= (e-1) + kMsRing-1
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            c.esafe = c.e + kMsRing - 2;
        }
        const float nxt = c.s_edge[(c.e & (kMsRing - 1)) * kMsThreads];  // stale beyond kn, unused
        c.pb = (c.e <= c.kn) ? nxt : INFINITY;
        if (c.e + kMsRing - 1 <= c.kn) cp_async4(c.s_edge + ((c.e - 1) & (kMsRing - 1)) * kMsThreads, c.p2f);
        c.p2f += p2_bs;
    }
    if (!LAST && c.e > 0) c.m.close(c.pt);
}

template <int PH>
__device__ __forceinline__ void ms_enter(MsCtx& c, int L, int iv, const int64_t q_bs, const int64_t pe_bs) {
    const int km = c.km;
    if (L + kMsDepth + 3 <= km) {  // level data of layer L+kMsDepth
        const float2* slot = c.s_src + ((L + kMsDepth) & (kMsRing - 1)) * kMsThreads;
        cp_async4(&slot->x, c.qpf);
        cp_async4(&slot->y, c.ppf);
    }
    c.qpf += q_bs;
    c.ppf += pe_bs;
    asm volatile("cp.async.commit_group;" ::: "memory");  // also carries the edges issued by the last emit
    c.s_hist[(L & (kMsRing - 1)) * kMsThreads] = c.e + kMsRing - 1;
    asm volatile("cp.async.wait_group %0;" ::"n"(kMsDepth) : "memory");
    c.esafe = max(c.esafe, c.s_hist[((L - kMsDepth) & (kMsRing - 1)) * kMsThreads]);
    if (L <= km - 3) {
        const float2 v = c.s_src[(L & (kMsRing - 1)) * kMsThreads];
        float qn[1] = {v.x};
        c.m.template step<PH>(v.y, qn);
    } else if (L == km - 2) {
        c.m.template step_bottom<PH>(iv);
    } else {
        c.m.template step_last<PH>();
    }
}

// strides in BYTES; blockIdx.y = batch (slab) index
__global__ void __launch_bounds__(kMsThreads)
mappm_sweep_kernel(const char* __restrict__ pe1, int64_t pe1_cs, const int64_t pe1_ls, int64_t pe1_bs,
                   const char* __restrict__ q1, int64_t q1_cs, const int64_t q1_ls, int64_t q1_bs,
                   const char* __restrict__ pe2, int64_t pe2_cs, const int64_t pe2_ls, int64_t pe2_bs,
                   char* __restrict__ q2, int64_t q2_cs, const int64_t q2_ls, int64_t q2_bs, int64_t ncol, int km,
                   int kn, int iv) {
    __shared__ float2 s_src[kMsRing * kMsThreads];
    __shared__ float s_edge[kMsRing * kMsThreads];
    __shared__ int s_hist[kMsRing * kMsThreads];
    const int64_t col = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (col >= ncol) return;
    const int64_t b = blockIdx.y;
    MsCtx c;
    const char* qc = q1 + b * q1_bs + col * q1_cs;
    const char* pc = pe1 + b * pe1_bs + col * pe1_cs;
    const char* p2 = pe2 + b * pe2_bs + col * pe2_cs;
    c.o2 = q2 + b * q2_bs + col * q2_cs - q2_ls;  // edge e closes output layer e-1
    c.km = km; c.kn = kn;
    c.s_src = s_src + threadIdx.x; c.s_edge = s_edge + threadIdx.x; c.s_hist = s_hist + threadIdx.x;
    auto ld = [](const char* p) { return __ldg(reinterpret_cast<const float*>(p)); };
    // prologue: one group per layer 1..kMsDepth; the first one also carries edges 1..kMsRing-1
#pragma unroll
    for (int i = 1; i < kMsRing; i++)
        if (i <= kn) cp_async4(c.s_edge + i * kMsThreads, p2 + i * pe2_ls);
#pragma unroll
    for (int j = 1; j <= kMsDepth; j++) {
        if (j + 3 <= km) {
            const float2* slot = c.s_src + j * kMsThreads;
            cp_async4(&slot->x, qc + (j + 2) * q1_ls);
            cp_async4(&slot->y, pc + (j + 3) * pe1_ls);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
#pragma unroll
    for (int i = 0; i < kMsRing; i++) c.s_hist[i * kMsThreads] = 0;
    c.qpf = qc + (kMsDepth + 3) * q1_ls;
    c.ppf = pc + (kMsDepth + 4) * pe1_ls;
    c.p2f = p2 + kMsRing * pe2_ls;
    float p[5], qA[1], qB[1], qC[1], qD[1];
#pragma unroll
    for (int j = 0; j < 5; j++) p[j] = ld(pc + j * pe1_ls);
    qA[0] = ld(qc);
    qB[0] = ld(qc + q1_ls);
    qC[0] = ld(qc + 2 * q1_ls);
    qD[0] = ld(qc + 3 * q1_ls);
    c.pb = ld(p2);
    c.m.init(p, qA, qB, qC, qD, iv);
    c.pe_top = p[0]; c.qtop = qA[0];
    c.e = 0; c.esafe = 0; c.pt = INFINITY;
    ms_emit<false>(c, pe2_ls, q2_ls);  // layer 0 (km >= 4: never the last)
#pragma unroll 1
    for (int L = 1; L < km; L++) {
        switch (L & 3) {
            case 1: ms_enter<1>(c, L, iv, q1_ls, pe1_ls); break;
            case 2: ms_enter<2>(c, L, iv, q1_ls, pe1_ls); break;
            case 3: ms_enter<3>(c, L, iv, q1_ls, pe1_ls); break;
            default: ms_enter<0>(c, L, iv, q1_ls, pe1_ls); break;
        }
        if (L < km - 1) ms_emit<false>(c, pe2_ls, q2_ls);
    }
    ms_emit<true>(c, pe2_ls, q2_ls);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

}  // namespace b200cs

using namespace b200cs;

extern "C" int b200cs_mappm_fast(const float* pe1, const int64_t pe1_s[3], const float* q1, const int64_t q1_s[3],
                                 const float* pe2, const int64_t pe2_s[3], float* q2, const int64_t q2_s[3],
                                 int64_t nbatch, int64_t ncol, int km, int kn, int iv, int kord, void* stream) {
    B200CS_REQUIRE(pe1 && q1 && pe2 && q2 && pe1_s && q1_s && pe2_s && q2_s, "mappm_fast: NULL argument");
    B200CS_REQUIRE(ncol >= 0 && nbatch >= 0 && kn >= 1, "mappm_fast: bad extents");
    B200CS_REQUIRE(nbatch <= 65535, "mappm_fast: nbatch > 65535");
    B200CS_REQUIRE(km >= 4, "mappm_fast: km=%d too small (the Fortran indexes out of range below 4 levels)", km);
    B200CS_REQUIRE(kord <= 3, "mappm_fast: kord=%d is not a lmt=0 profile; use b200cs_mappm", kord);
    if (ncol == 0 || nbatch == 0) return B200CS_OK;
    const dim3 grid((unsigned)((ncol + kMsThreads - 1) / kMsThreads), (unsigned)nbatch);
    mappm_sweep_kernel<<<grid, kMsThreads, 0, as_stream(stream)>>>(
        (const char*)pe1, 4 * pe1_s[1], 4 * pe1_s[2], 4 * pe1_s[0], (const char*)q1, 4 * q1_s[1], 4 * q1_s[2], 4 * q1_s[0],
        (const char*)pe2, 4 * pe2_s[1], 4 * pe2_s[2], 4 * pe2_s[0], (char*)q2, 4 * q2_s[1], 4 * q2_s[2], 4 * q2_s[0], ncol,
        km, kn, iv);
    return check_launch("mappm_fast");
}
