// pressure_coarsen_tma.cuh -- K6, third design: the fused pressure-level coarse-graining kernel
// fed by TMA (B200CS_FAST arithmetic, fp32 data, 8 x 8 windows, kord <= 3).  Replaces
// regrid_to_area_weighted_pressure -> regrid_vertical/mappm -> _mask_weights ->
// weighted_block_average (external/vcm/vcm/cubedsphere/regridz.py:31-220,
// external/mappm/mappm/mappm.f90, external/vcm/vcm/cubedsphere/coarsen_restarts.py:485-534).
//
// Same dataflow and results as pressure_coarsen_sweep.cuh (coarse delp, interface pressures, masks
// and the level search bit-identical to the EXACT kernel; values within 1e-6).  What is new
// (profiles/r01_sweep_v8_ncu_summary.txt: issue-bound, 405 warp instructions per level):
//
//   * Every level of every array reaches shared memory as ONE 2-D TMA tile (cp.async.bulk.tensor,
//     8 rows x 32 columns = 1 KB, 128-byte swizzle) issued by a single lane; completion is an
//     mbarrier transaction count.  The per-thread LDGSTS, their 64-bit pointer arithmetic and the
//     cp.async group bookkeeping are gone.  A slot of the 8-level ring is refilled by the LAST warp
//     that releases it (shared-memory counter), so nobody ever waits for a free slot.
//   * A warp owns HALF A COARSE CELL: 8 rows x 4 columns (one 16-byte chunk per row, which the
//     128-byte swizzle spreads over all banks), two warps per cell.  All 32 lanes share the
//     cell's level table (uniform loads) and the weighted values of an output level are summed
//     over 32 lanes, not 8: accumulators shrink from [8 rows][km][4][NF+1] to [8 warps][km][NF+1]
//     (50 KB -> 12.6 KB), which pays for the deeper level ring.  Neighbouring columns of a
//     compact 8 x 4 patch also sit on closer output levels than a 32-column strip.
//   * The column machine works on PAIRS of fields with packed fp32 instructions and a clamp-form
//     limiter (ppm_pack.cuh).
//   * Phase A reads delp from TMA tiles too (8 levels per tile); one thread per (level, cell) forms
//     the 64 products and adds them in numpy's order straight from shared memory.
//   * NaN-skipping (xarray's coarsen().sum()) is checked once per reduced level on the warp total
//     instead of once per staged value.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "ppm_pack.cuh"
#include "pressure_coarsen_sweep.cuh"

namespace b200cs {

constexpr int kTmaSlots = 8;          // level ring depth: two halves of 4 slots (the layer loop is unrolled by 4)
constexpr int kTmaTile = 1024;        // bytes of one array-level tile (8 rows x 128 B)
constexpr int kTmaKA = 8;             // Phase A: levels per delp tile
constexpr int kTmaMaxArrays = 5;      // 4 fields + delp
constexpr int kTmaMaxChunks = 16;     // Phase A: delp tiles in flight (km <= 128)

struct TmaMaps {
    CUtensorMap lvl[kTmaMaxArrays];   // [NFL] fields, then delp: box (32, 8, 1, 1)
    CUtensorMap dchunk;               // delp: box (32, 8, kTmaKA, 1)
};

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// (the suspend-time hint, in ns, lets the hardware park the warp instead of having it re-issue the poll:
// 1.186 -> 1.181 ms at C384; 1 us and 10 us give the same)
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "LAB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra DONE;\n"
        "bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(bar), "r"(parity), "r"(1000u) : "memory");
}
// ... for the role that runs AHEAD of the other one (warp-specialised kernel: the geometry warps): it
// sleeps between polls instead of taking issue slots from the warps it is waiting for.  (A
// suspend-time hint on try_wait changed nothing: 91 polls per level and warp either way.)
__device__ __forceinline__ void mbar_wait_sleep(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra DONE;\n"
        "LAB_WAIT:\n"
        "nanosleep.u32 200;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@!P1 bra LAB_WAIT;\n"
        "DONE:\n"
        "}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load4(unsigned dst, const CUtensorMap* map, int x, int y, int z, int b, unsigned bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cta.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
        ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(z), "r"(b), "r"(bar) : "memory");
}

// shared memory carve-up (bytes from the 1 KB-aligned base).  Small fixed parts first, then the level
// ring (two HALVES of 4 slots of NA tiles; ONE register, 0 or `half`, selects the half: tiles at
// ring + hoff, the half's four "full" barriers at bars + (hoff >> 8), its four release counters at
// cnts + (hoff >> 9)), the staging ring, the accumulators, the level tables.  Phase A borrows
// ring + stage + accumulators (contiguous) for its delp tiles: 84 KB at NA = 5, km = 79 -- the whole
// column of the tile in flight at once.
template <int NA>
struct TmaSmem {
    static constexpr int w = 0;                                         // float [8][32] weights (NaN -> 0)
    static constexpr int den = w + kTileY * kTileX * 4;                 // float [4] cell weight sums
    static constexpr int bars = den + 16;                               // u64 full[2][4] of the ring, 32 B per half at stride half >> 8
    static constexpr int cnts = bars + 256;                             // unsigned released[2][4], 16 B per half at stride half >> 9
    static constexpr int bar_a = cnts + 128;                            // u64 fullA[kTmaMaxChunks] (Phase A tiles)
    static constexpr int ring = 2048;
    static constexpr int half = 4 * NA * kTmaTile;                      // tiles of 4 slots
    static constexpr int stage = ring + 2 * half;                       // float4 [kStage][256]
    static constexpr int acc = stage + kStage * kThreads * 16;          // float [8][km][NV]; then float4 [4][LS] level tables
    static_assert((half >> 8) >= 32 && 2 * (half >> 8) <= 256 && (half >> 9) >= 16 && 2 * (half >> 9) <= 128, "ring bookkeeping strides");
    static_assert(bar_a + 8 * 16 <= ring, "fixed part overflows");
};

// keeps a cheap-to-recompute value in its register (ptxas rematerialises address arithmetic inside the hot loop)
__device__ __forceinline__ unsigned opaque(unsigned v) {
    unsigned r;
    asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}
__device__ __forceinline__ float4 lds128(unsigned a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts128(unsigned a, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float lds32(unsigned a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts32(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }

// The level ring: level j lives in slot j % 4 of half (j / 4) % 2 and holds NA tiles (fields, then
// delp); its barrier completes with phase parity (j / 8) % 2.  A slot is refilled, with level j + 8,
// by the LAST warp that releases it (shared-memory counter): nobody ever waits for a free slot, and
// up to 7 levels (35 KB at NA = 5) are in flight per CTA -- the sweep of one level takes a warp a
// few hundred cycles, DRAM latency under load is a few thousand.
template <int NA>
struct TmaRing {
    unsigned sb;          // shared address of the carve-up's base
    unsigned nact;        // warps that consume the ring
    // hoff = 0 or TmaSmem<NA>::half
    __device__ __forceinline__ unsigned full(unsigned hoff, int slot) const { return sb + TmaSmem<NA>::bars + (hoff >> 8) + slot * 8; }
    // (inline on purpose: out of line -- one lane in 256 runs it per level -- the call's register
    // constraints cost 8 %, 1.15 -> 1.24 ms)
    __device__ __forceinline__ void issue(const TmaMaps& maps, unsigned hoff, int slot, int level) const {
        const unsigned bar = full(hoff, slot);
        mbar_expect_tx(bar, NA * kTmaTile);
#pragma unroll
        for (int i = 0; i < NA; i++)
            tma_load4(sb + TmaSmem<NA>::ring + hoff + (slot * NA + i) * kTmaTile, &maps.lvl[i], (int)(blockIdx.x * kTileX), (int)(blockIdx.y * kTileY),
                      level, (int)blockIdx.z, bar);
    }
    // every lane of the warp has read its values of `level` from (hoff, slot): the last warp to say
    // so refills the slot with level + 8.  (The warp is converged at the shared loads, which
    // therefore precede lane 0's atomic.)
    __device__ __forceinline__ void release(const TmaMaps& maps, unsigned hoff, int slot, int level, int km, int lane) const {
        if (lane == 0) {
            unsigned old;
            asm volatile("atom.shared.inc.u32 %0, [%1], %2;" : "=r"(old) : "r"(sb + TmaSmem<NA>::cnts + (hoff >> 9) + slot * 4), "r"(nact - 1) : "memory");
            if (old == nact - 1) {
                if (level + kTmaSlots < km) issue(maps, hoff, slot, level + kTmaSlots);
            }
        }
    }
};

template <int NP>
struct TmaCtx {
    PpmPack<NP> m;
    float pacc;           // pe[] of the last consumed level + 1 (fp32 cumsum, as Phase A)
    unsigned lvl;         // shared address of this warp's cell table: float4 [km + 1] {pb, 1/(pb-pt), threshold, coarse delp}
    unsigned acc;         // shared address of this warp's accumulators: float [km][2 NP + 1]
    unsigned buf;         // shared address of this thread's staging ring: float4 [kStage] (stride kThreads)
    int k;                // next output level of this lane
    int kdone;            // (warp-uniform) output levels below this one are reduced
    int nanwin;           // (warp-uniform) > 0: a NaN level is inside the stencil of some lane, literal limiter tests
    unsigned hoff;        // half of the ring the next levels are read from (0 or TmaSmem::half)
    unsigned par;         // phase parity of that half's barriers
    float pt, pb, rdpc, thr;
    float wv, whalf, psurf, psmin;
    int km, lane;
};

// sums 2 NP values over the 32 lanes of the warp; afterwards every lane holds the total of field
// lane >> 3 (NP = 2) or lane >> 4 (NP = 1)
template <int NP>
__device__ __forceinline__ float reduce_fields32(float4 pv, int lane) {
    const unsigned m = 0xffffffffu;
    float t;
    if (NP == 2) {
        const bool hi = lane & 16;
        const f2 p01 = f2_make(pv.x, pv.y), p23 = f2_make(pv.z, pv.w);
        f2 keep = hi ? p23 : p01;
        const f2 send = hi ? p01 : p23;
        f2 recv = send;
#if defined(__CUDA_ARCH__)
        recv.v = __shfl_xor_sync(m, send.v, 16);
#endif
        keep = f2_add(keep, recv);
        const bool b3 = lane & 8;
        const float k0 = b3 ? f2_hi(keep) : f2_lo(keep), s0 = b3 ? f2_lo(keep) : f2_hi(keep);
        t = k0 + __shfl_xor_sync(m, s0, 8);
    } else {
        const bool hi = lane & 16;
        const float k0 = hi ? pv.y : pv.x, s0 = hi ? pv.x : pv.y;
        t = k0 + __shfl_xor_sync(m, s0, 16);
        t = t + __shfl_xor_sync(m, t, 8);
    }
    t = t + __shfl_xor_sync(m, t, 4);
    t = t + __shfl_xor_sync(m, t, 2);
    return t + __shfl_xor_sync(m, t, 1);
}
__device__ __forceinline__ float reduce_max32(float t) {
    const unsigned m = 0xffffffffu;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t = fmaxf(t, __shfl_xor_sync(m, t, o));
    return t;
}
__device__ __forceinline__ float reduce_all32(float t) {
    const unsigned m = 0xffffffffu;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t = t + __shfl_xor_sync(m, t, o);
    return t;
}

// Lanes that ran more than kStage levels ahead of the slowest lane of their warp add their values
// straight into the accumulators (see sweep_accumulate_slow).  Called by exactly the lanes of
// `mask`; those on the same level take turns in lane order, so the sum is deterministic.
template <int NP>
__device__ __noinline__ void tma_accumulate_slow(unsigned acc, int k, int lane, float4 pv, unsigned mask) {
    constexpr int NV = 2 * NP + 1;
    const unsigned peers = __match_any_sync(mask, (unsigned)k);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int turns = (int)__reduce_max_sync(mask, (unsigned)rank + 1u);
    const unsigned dst = acc + k * NV * 4;
    const float p[4] = {nan_to_zero(pv.x), nan_to_zero(pv.y), nan_to_zero(pv.z), nan_to_zero(pv.w)};
    for (int t = 0; t < turns; t++) {
        if (rank == t) {
#pragma unroll
            for (int i = 0; i < 2 * NP; i++) sts32(dst + i * 4, lds32(dst + i * 4) + p[i]);
        }
        __syncwarp(mask);
    }
}

template <int NP>
__device__ __forceinline__ float4 unpack4(const f2 (&o)[NP]) {
    if (NP == 2) return make_float4(f2_lo(o[0]), f2_hi(o[0]), f2_lo(o[NP - 1]), f2_hi(o[NP - 1]));
    return make_float4(f2_lo(o[0]), f2_hi(o[0]), 0.f, 0.f);
}

// sums the four values (a.lo, a.hi, b.lo, b.hi) over the 8 lanes that share lane >> 3, transposing:
// lanes 2f and 2f + 1 of the group end up with the total of value f (NP = 1: two values, lanes
// 0..3 hold the first, 4..7 the second)
template <int NP>
__device__ __forceinline__ float reduce_fields8t(f2 a, f2 b, int lane) {
    const unsigned m = 0xffffffffu;
    float t;
    if (NP == 2) {
        const bool hi = lane & 4;
        f2 keep = hi ? b : a;
        const f2 send = hi ? a : b;
        f2 recv = send;
#if defined(__CUDA_ARCH__)
        recv.v = __shfl_xor_sync(m, send.v, 4);
#endif
        keep = f2_add(keep, recv);
        const bool b1 = lane & 2;
        const float k0 = b1 ? f2_hi(keep) : f2_lo(keep), s0 = b1 ? f2_lo(keep) : f2_hi(keep);
        t = k0 + __shfl_xor_sync(m, s0, 2);
    } else {
        const bool hi = lane & 4;
        const float k0 = hi ? f2_hi(a) : f2_lo(a), s0 = hi ? f2_lo(a) : f2_hi(a);
        t = k0 + __shfl_xor_sync(m, s0, 4);
        t = t + __shfl_xor_sync(m, t, 2);
    }
    return t + __shfl_xor_sync(m, t, 1);
}

// Reduces `nlev` (1..4, warp-uniform) staged levels starting at c.kdone, all of which every lane of
// the warp has passed.  The 8 lanes with lane >> 3 == i take level kdone + i: each reads the
// staged float4 of four source lanes (sub, sub + 8, sub + 16, sub + 24: 8 reducer lanes cover 128
// contiguous bytes per load, conflict-free), adds them with packed adds, and a 3-step transposing
// butterfly finishes the sum -- 4 LDS.128 + 4 STS.128 + ~25 arithmetic / shuffle instructions
// for four levels, where one butterfly over 32 lanes per level took 45 per level.
// (Not inlined, scalar arguments: the code exists once per kernel instead of once per unrolled phase,
// and the caller's column state stays in registers.)
struct TmaRed {
    unsigned wbuf;   // shared address of the warp's staging ring (lane 0's slot 0)
    unsigned lvl, acc;
    float psmin, psurf, wv, whalf;
    int kdone, lane;
};
template <int NP>
__device__ __noinline__ void tma_reduce_batch(const TmaRed c, int nlev) {
    const unsigned full = 0xffffffffu;
    constexpr int NV = 2 * NP + 1;
    const int i = c.lane >> 3, sub = c.lane & 7;
    const int kk = c.kdone + i;
    const bool on = i < nlev;
    const unsigned base = c.wbuf + (kk & (kStage - 1)) * (kThreads * 16) + sub * 16;
    __syncwarp();  // the staged values were written by other lanes
    float4 v[4];
#pragma unroll
    for (int t = 0; t < 4; t++) v[t] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (on) {
#pragma unroll
        for (int t = 0; t < 4; t++) v[t] = lds128(base + t * 128);
        // a slot that holds no pending level reads as zero: a lane that overflowed the ring at
        // a level went through tma_accumulate_slow and must not be counted again
#pragma unroll
        for (int t = 0; t < 4; t++) sts128(base + t * 128, make_float4(0.f, 0.f, 0.f, 0.f));
    }
    f2 a = f2_add(f2_add(f2_make(v[0].x, v[0].y), f2_make(v[1].x, v[1].y)), f2_add(f2_make(v[2].x, v[2].y), f2_make(v[3].x, v[3].y)));
    f2 b = f2_make(0.f, 0.f);
    if (NP == 2) b = f2_add(f2_add(f2_make(v[0].z, v[0].w), f2_make(v[1].z, v[1].w)), f2_add(f2_make(v[2].z, v[2].w), f2_make(v[3].z, v[3].w)));
    float tot = reduce_fields8t<NP>(a, b, c.lane);
    if (__any_sync(full, tot != tot)) {  // NaN-skipping sums (xarray's coarsen().sum()): rare
#pragma unroll
        for (int t = 0; t < 4; t++) {
            v[t].x = nan_to_zero(v[t].x); v[t].y = nan_to_zero(v[t].y); v[t].z = nan_to_zero(v[t].z); v[t].w = nan_to_zero(v[t].w);
        }
        a = f2_add(f2_add(f2_make(v[0].x, v[0].y), f2_make(v[1].x, v[1].y)), f2_add(f2_make(v[2].x, v[2].y), f2_make(v[3].x, v[3].y)));
        if (NP == 2) b = f2_add(f2_add(f2_make(v[0].z, v[0].w), f2_make(v[1].z, v[1].w)), f2_add(f2_make(v[2].z, v[2].w), f2_make(v[3].z, v[3].w)));
        tot = reduce_fields8t<NP>(a, b, c.lane);
    }
    // weight sum of the level: every lane of the warp is above the mask threshold (c.psmin = the
    // smallest surface pressure of the 32 columns) except near the surface
    const float thr = on ? lds32(c.lvl + kk * 16 + 8) : -INFINITY;
    float rd = c.whalf;
    if (__any_sync(full, !(thr < c.psmin))) {
        for (int ii = 0; ii < nlev; ii++) {
            const float thr_i = __shfl_sync(full, thr, ii * 8);
            const float r = reduce_all32(thr_i < c.psurf ? nan_to_zero(c.wv) : 0.f);
            rd = (i == ii) ? r : rd;
        }
    }
    if (on) {
        const unsigned dst = c.acc + kk * (NV * 4);
        const int f = NP == 2 ? (sub >> 1) : (sub >> 2);
        if (sub == 1) sts32(dst + 2 * NP * 4, rd);  // a level is reduced once; tma_accumulate_slow adds values only
        else if (NP == 2 ? ((sub & 1) == 0) : ((sub & 3) == 0)) sts32(dst + f * 4, lds32(dst + f * 4) + tot);
    }
    __syncwarp();  // the zeros are in place before any lane stages into these slots again
}

// Reduces the staged levels every lane of the warp has passed, [c.kdone, min over lanes of k), four
// at a time; fewer only when a lane is about to overflow its staging ring (steep terrain inside the
// warp's patch) or at the end of the column (FLUSH).
template <int NP, bool FLUSH>
__device__ __forceinline__ void tma_reduce(TmaCtx<NP>& c) {
    const unsigned full = 0xffffffffu;
    const int kend = (int)__reduce_min_sync(full, (unsigned)c.k);
    int n = kend - c.kdone;
    if (!FLUSH && n < 4) {
        if (n <= 0) return;
        if ((int)__reduce_max_sync(full, (unsigned)c.k) - c.kdone < kStage - 1) return;
    }
    TmaRed r;
    r.wbuf = c.buf - c.lane * 16; r.lvl = c.lvl; r.acc = c.acc;
    r.psmin = c.psmin; r.psurf = c.psurf; r.wv = c.wv; r.whalf = c.whalf; r.lane = c.lane;
    while (n > 0) {
        const int nl = n < 4 ? n : 4;
        r.kdone = c.kdone;
        tma_reduce_batch<NP>(r, nl);
        c.kdone += nl;
        n -= 4;
    }
}

// One pass of the edge loop: evaluates the next output edge of EVERY lane; the lanes that have none
// in this layer (`can` false) run along predicated off, so the body is straight-line code.  The
// weighted values (mask and column weight folded into the scale) go into the lane's staging ring --
// NaNs are dealt with when the level is reduced.  CHECK: some lane may be more than kStage levels
// ahead of the slowest lane of its warp (steep terrain inside the warp's patch; rare): those lanes
// add their values straight into the accumulators.
template <int NP, bool LAST, bool CHECK>
__device__ __forceinline__ void tma_edge_pass(TmaCtx<NP>& c, bool can) {
    const unsigned full = 0xffffffffu;
    unsigned overmask = 0;
    if (CHECK) overmask = __ballot_sync(full, can && (c.k - c.kdone >= kStage));
    f2 o[NP];
    // 1/thickness, the column weight and the below-surface mask enter as one factor
    // (MASK_NONE: thr = -inf; level 0: thr = +inf, see kernel)
    const float scale = (c.thr < c.psurf) ? c.rdpc * c.wv : 0.f;
    c.m.template edge<LAST>(c.pt, c.pb, scale, o, can, LAST ? c.m.Q[1] : nullptr);  // the last layer runs at phase 1
    if (CHECK && overmask != 0) {
        if (can) {
            if ((overmask >> c.lane) & 1u) tma_accumulate_slow<NP>(c.acc, c.k, c.lane, unpack4<NP>(o), overmask);
            else sts128(c.buf + (c.k & (kStage - 1)) * (kThreads * 16), unpack4<NP>(o));
        }
    } else {
        if (can) sts128(c.buf + (c.k & (kStage - 1)) * (kThreads * 16), unpack4<NP>(o));
    }
    if (can) {
        c.pt = c.pb;
        c.k++;
        const float4 lv = lds128(c.lvl + c.k * 16);
        c.pb = lv.x; c.rdpc = lv.y; c.thr = lv.z;
    }
}

// output edges inside the current source layer (all remaining ones when it is the last layer).
// Common case (no lane within two levels of overflowing its staging ring, at most two edges per
// lane in this layer): two unrolled passes without the overflow test.  general: the layer's edge
// values may be NaN next to a finite mean (boundary layers, NaN level in the stencil): close_general.
// QI = slot of the current layer's mean in the window (its phase).
template <int NP, bool LAST, int QI>
__device__ __forceinline__ void tma_emit(TmaCtx<NP>& c, bool general) {
    const unsigned full = 0xffffffffu;
    // the level table ends with a +inf sentinel: lanes that are through never qualify
    bool can = LAST ? (c.k < c.km) : (c.pb <= c.m.pe_hi);
    if (!LAST && !__any_sync(full, c.k - c.kdone >= kStage - 2)) {
        tma_edge_pass<NP, LAST, false>(c, can);
        can = c.pb <= c.m.pe_hi;
        if (__any_sync(full, can)) {
            tma_edge_pass<NP, LAST, false>(c, can);
            can = c.pb <= c.m.pe_hi;
        }
    }
    while (__any_sync(full, can)) {
        tma_edge_pass<NP, LAST, true>(c, can);
        can = LAST ? (c.k < c.km) : (c.pb <= c.m.pe_hi);
    }
    if (!LAST) {
        if (general) c.m.close_general(c.pt, c.m.Q[QI]);  // (warp-uniform)
        else c.m.close(c.pt);
    }
    tma_reduce<NP, LAST>(c);
}

// this thread's values of one level: NFL fields (an odd count repeats the last one) and delp
template <int NP, int NFL>
__device__ __forceinline__ void tma_read(unsigned src, f2 (&q)[NP], float& d) {
    float v[2 * NP];
#pragma unroll
    for (int i = 0; i < 2 * NP; i++) {
        const int t = i < NFL ? i : NFL - 1;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v[i]) : "r"(src + t * kTmaTile));
    }
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(d) : "r"(src + NFL * kTmaTile));
#pragma unroll
    for (int p = 0; p < NP; p++) q[p] = f2_make(v[2 * p], v[2 * p + 1]);
}

// Interior source layer L (PH = L%4; 1 <= L <= km-3): step to it, emit its output edges.  Level
// L+2 is already in the window (slot (PH+2)%4, loaded by the previous call); level L+3 is read
// from the ring into the free slot (PH+3)%4 right after the step was issued, so that neither the
// barrier wait nor the shared-memory latency sits in front of the arithmetic.
// c.nanwin (warp-uniform) > 0: a NaN level is inside the stencil of some lane -> literal limiter tests.
template <int NP, int NFL, int PH>
__device__ __forceinline__ void tma_layer(TmaCtx<NP>& c, const TmaRing<NFL + 1>& rg, const TmaMaps& maps, unsigned src0, int L) {
    const unsigned full = 0xffffffffu;
    const bool general = (L == 1) || (c.nanwin > 0);  // layer 1: its upper edge value comes from the top boundary formula
    c.nanwin = c.nanwin > 0 ? c.nanwin - 1 : 0;
    const float pn = c.pacc;  // pe[L+3]
    if (L + 3 < c.km) {
        // level L+3 lives in slot (PH+3)&3; slot 0 starts the next half
        constexpr int slot = (PH + 3) & 3;
        if (slot == 0) {
            c.par ^= (c.hoff != 0);  // a new round of the ring begins when half 0 comes up again
            c.hoff = TmaSmem<NFL + 1>::half - c.hoff;
        }
        float d;
        mbar_wait(rg.full(c.hoff, slot), c.par);
        tma_read<NP, NFL>(src0 + c.hoff + slot * (NFL + 1) * kTmaTile, c.m.Q[slot], d);
        rg.release(maps, c.hoff, slot, L + 3, c.km, c.lane);
        c.pacc = c.pacc + nan_to_zero(d);
        f2 s2 = c.m.Q[slot][0];
        if (NP == 2) s2 = f2_add(s2, c.m.Q[slot][NP - 1]);
        const float sv = f2_lo(s2) + f2_hi(s2);
        if (__any_sync(full, sv != sv)) c.nanwin = 4;  // NaN (or inf - inf) entering the stencil: rare
    }
    c.m.template step_inplace<PH>(pn, general);
    tma_emit<NP, false, PH>(c, general);
}

template <int NFL>
__device__ __forceinline__ void pc_gather_tail(const FieldPtrs<NFL>& fp, const float* delp_coarse, const PeerOut& peers,
                                               int first, int64_t b0, int km, int64_t cyN, int64_t cxN, int tid) {
    // The CTAs of one row of tiles (same blockIdx.y, b0) produce one complete coarse row.  The
    // last of them to finish copies that row -- contiguous cxN-float segments, read back through
    // L2 -- into this rank's slot on every peer with 16-byte stores over NVLink.
    // One release-acquire atomic by thread 0 after the CTA barrier (cumulative over the Phase C stores
    // of all warps) decides; r01 had every thread fence first (+65 us per launch at C384, 5 % --
    // tools/gather_diag.py).  (Leaving the copy to warp 0 alone so that the other warps need not wait
    // for the atomic was far worse: 1.50 ms per step at N = 2.)
    __shared__ int s_last;
    __syncthreads();
    if (tid == 0) {
        int prev;
        asm volatile("atom.add.acq_rel.gpu.global.s32 %0, [%1], 1;" : "=r"(prev) : "l"(peers.counters + (b0 * gridDim.y + blockIdx.y)) : "memory");
        s_last = (prev == (int)gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    const int64_t cy0 = blockIdx.y;  // kTileY / 8 = 1 coarse row per row of tiles
    const bool vec = (cxN % 4 == 0);
    const int64_t per_row = vec ? cxN / 4 : cxN;
    const int64_t total = (int64_t)km * per_row;
#pragma unroll
    for (int sl = 0; sl <= NFL; sl++) {
        const float* src = sl < NFL ? (const float*)fp.out[sl < NFL ? sl : 0] : delp_coarse;
        if (sl == NFL && !(first == 0 && delp_coarse)) continue;
        const int64_t slot = sl < NFL ? (first + sl) : peers.nfields;
        for (int64_t e = tid; e < total; e += kThreads) {
            const int64_t c = e % per_row;
            const int64_t k = e / per_row;
            if (cy0 >= cyN) continue;
            const int64_t o = ((b0 * km + k) * cyN + cy0) * cxN;
            if (vec) {
                const float4 v = __ldcg(reinterpret_cast<const float4*>(src + o) + c);
                for (int p = 0; p < peers.n; p++)
                    reinterpret_cast<float4*>(peers.base[p] + slot * peers.field_elems + o)[c] = v;
            } else {
                const float v = __ldcg(src + o + c);
                for (int p = 0; p < peers.n; p++) peers.base[p][slot * peers.field_elems + o + c] = v;
            }
        }
    }
}

// KR = (km - 3) % 4: the number of interior layers left over by the loop of four phases -- a template
// parameter so that the unrolled loop has ONE exit (with exits between the phases ptxas brings the
// rotating window back to canonical registers after every phase: 22 MOV per level)
// One- and two-field launches (NP = 1; the 13th field of a whole timestep) run three CTAs per SM at 80
// registers (32-108 B of spills): 0.876 -> 0.768 ms for one field, 0.909 -> 0.806 ms for two at C384
// (tools/nfield_time.py).  The four-field kernel spills 480 B at 80 registers and stays at two.
#ifndef B200CS_NP1_CTAS
#define B200CS_NP1_CTAS 3
#endif
template <int NP, int NFL, int KR>
__global__ void __launch_bounds__(kThreads, NP == 1 ? B200CS_NP1_CTAS : 2)
pressure_coarsen_tma_kernel(const __grid_constant__ TmaMaps maps, FieldPtrs<NFL> fp,
                            const float* __restrict__ w, int64_t ws0, int64_t wsy,
                            float* __restrict__ delp_coarse, int km, int64_t ny, int64_t nx, float toa,
                            int mask_mode, int iv, const __grid_constant__ PeerOut peers, int first) {
    constexpr int NA = NFL + 1;
    constexpr int NV = 2 * NP + 1;
    constexpr int CW = kTileX / 8;  // cells per tile
    using SM = TmaSmem<NA>;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int r = lane >> 2, c4 = lane & 3;
    const int xl = wid * 4 + c4;  // column inside the tile: 16-byte chunk index = wid
    const int64_t xg = (int64_t)blockIdx.x * kTileX + xl, yg = (int64_t)blockIdx.y * kTileY + r;
    const int64_t b0 = blockIdx.z;
    const int64_t cxN = nx / 8, cyN = ny / 8;
    const bool active = xg < nx;  // ny % 8 == 0: rows are always inside
    const unsigned my_off = (unsigned)(r * 128 + ((wid ^ r) << 4) + c4 * 4);

    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    // the swizzle pattern of a tile is a function of the shared ADDRESS: 1 KB alignment by hand
    unsigned char* smem_raw = smem_dyn + ((1024u - ((unsigned)__cvta_generic_to_shared(smem_dyn) & 1023u)) & 1023u);
    // (volatile: ptxas otherwise re-derives the window address -- S2R SR_CgaCtaId + LEA -- at every use)
    const unsigned sbase = opaque((unsigned)__cvta_generic_to_shared(smem_raw));
    const int LS = sweep_lvl_stride(km);
    float* s_acc = reinterpret_cast<float*>(smem_raw + SM::acc);              // [8 warps][km][NV]
    float4* s_lvl = reinterpret_cast<float4*>(s_acc + (size_t)8 * km * NV);   // [CW][LS], last = sentinel (16-byte aligned: km * NV * 32 bytes)
    float* s_w = reinterpret_cast<float*>(smem_raw + SM::w);                  // [8][32]
    float* s_den = reinterpret_cast<float*>(smem_raw + SM::den);              // [CW]
    const unsigned bar_a = sbase + SM::bar_a;
    const unsigned lvl_sh = sbase + SM::acc + 8 * km * NV * 4;

    TmaRing<NA> rg;
    rg.sb = sbase;
    {
        const int64_t left = nx - (int64_t)blockIdx.x * kTileX;
        rg.nact = left >= kTileX ? 8u : (unsigned)(left / 8) * 2u;
    }
    // Phase A streams the tile's delp in chunks of kTmaKA levels (8 KB); ring + staging ring +
    // accumulators hold nbuf of them at once (all ten at NA = 5, km = 79)
    const int nchunk = (km + kTmaKA - 1) / kTmaKA;
    int nbuf = (2 * SM::half + kStage * kThreads * 16 + 8 * km * NV * 4) / (kTmaKA * kTmaTile);
    nbuf = nbuf < nchunk ? nbuf : nchunk;
    nbuf = nbuf < kTmaMaxChunks ? nbuf : kTmaMaxChunks;
    const unsigned tiles_a = sbase + SM::ring;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            mbar_init(rg.full(0, i), 1);
            mbar_init(rg.full(SM::half, i), 1);
            sts32(sbase + SM::cnts + i * 4, 0.f);
            sts32(sbase + SM::cnts + (SM::half >> 9) + i * 4, 0.f);
        }
        for (int i = 0; i < nbuf; i++) mbar_init(bar_a + i * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int cch = 0; cch < nbuf; cch++) {
            mbar_expect_tx(bar_a + cch * 8, kTmaKA * kTmaTile);
            tma_load4(tiles_a + cch * kTmaKA * kTmaTile, &maps.dchunk, (int)(blockIdx.x * kTileX),
                      (int)(blockIdx.y * kTileY), cch * kTmaKA, (int)b0, bar_a + cch * 8);
        }
    }

    // ---------------- Phase A: coarse pressure levels of the tile's cells -------------------
    const float wv = active ? __ldg(w + b0 * ws0 + yg * wsy + xg) : 0.f;
    s_w[r * kTileX + xl] = nan_to_zero(wv);
    const float whalf = reduce_all32(nan_to_zero(wv));
    __syncthreads();  // barriers initialised, weights in shared memory
    if (tid < CW) {   // weight sum of a cell: rows pairwise (numpy), then row after row
        float d = 0.f;
        for (int rr = 0; rr < 8; rr++) {
            const float4 a = *reinterpret_cast<const float4*>(s_w + rr * kTileX + tid * 8);
            const float4 b = *reinterpret_cast<const float4*>(s_w + rr * kTileX + tid * 8 + 4);
            d = __fadd_rn(d, __fadd_rn(__fadd_rn(__fadd_rn(a.x, a.y), __fadd_rn(a.z, a.w)), __fadd_rn(__fadd_rn(b.x, b.y), __fadd_rn(b.z, b.w))));
        }
        s_den[tid] = d;
    }
    __syncthreads();
    float pacc = toa;
#pragma unroll 1
    for (int c0 = 0; c0 < nchunk; c0 += nbuf) {  // rounds of nbuf chunks (one round when everything fits)
        const unsigned rpar = (unsigned)(c0 / nbuf) & 1u;
        const int cend = c0 + nbuf < nchunk ? c0 + nbuf : nchunk;
        // coarse delp: warp (ch % 8) takes chunk ch as soon as it has landed, one lane per
        // (level, cell): 64 products delp*w in numpy's order (rows pairwise, then row after row),
        // straight from the swizzled tile; bit-identical to the EXACT kernel
#pragma unroll 1
        for (int ch = c0 + wid; ch < cend; ch += 8) {
            const unsigned tile = tiles_a + (ch - c0) * kTmaKA * kTmaTile;
            mbar_wait(bar_a + (ch - c0) * 8, rpar);
            const int kk = lane >> 2, cc = lane & 3, k = ch * kTmaKA + kk;
            const unsigned lt = tile + kk * kTmaTile;
            float num = 0.f;
#pragma unroll 2
            for (int rr = 0; rr < 8; rr++) {
                const float4 a = lds128(lt + rr * 128 + (((2 * cc) ^ rr) << 4));
                const float4 b = lds128(lt + rr * 128 + (((2 * cc + 1) ^ rr) << 4));
                const float4 wa = *reinterpret_cast<const float4*>(s_w + rr * kTileX + cc * 8);
                const float4 wb = *reinterpret_cast<const float4*>(s_w + rr * kTileX + cc * 8 + 4);
                float p0 = __fmul_rn(a.x, wa.x), p1 = __fmul_rn(a.y, wa.y), p2 = __fmul_rn(a.z, wa.z), p3 = __fmul_rn(a.w, wa.w);
                float p4 = __fmul_rn(b.x, wb.x), p5 = __fmul_rn(b.y, wb.y), p6 = __fmul_rn(b.z, wb.z), p7 = __fmul_rn(b.w, wb.w);
                float rs = __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3)), __fadd_rn(__fadd_rn(p4, p5), __fadd_rn(p6, p7)));
                if (rs != rs) {  // NaN in delp: skipped (rare)
                    p0 = nan_to_zero(p0); p1 = nan_to_zero(p1); p2 = nan_to_zero(p2); p3 = nan_to_zero(p3);
                    p4 = nan_to_zero(p4); p5 = nan_to_zero(p5); p6 = nan_to_zero(p6); p7 = nan_to_zero(p7);
                    rs = __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), __fadd_rn(p2, p3)), __fadd_rn(__fadd_rn(p4, p5), __fadd_rn(p6, p7)));
                }
                num = __fadd_rn(num, rs);
            }
            if (k < km) {
                const float dc = num / s_den[cc];
                s_lvl[cc * LS + k].w = dc;
                const int64_t cyg = blockIdx.y, cxg = (int64_t)blockIdx.x * CW + cc;
                if (delp_coarse && cxg < cxN) delp_coarse[((b0 * km + k) * cyN + cyg) * cxN + cxg] = dc;
            }
        }
        // fine surface pressure: sequential fp32 cumsum of the column (xarray cumsum == nancumsum),
        // every thread through all chunks in level order
#pragma unroll 1
        for (int ch = c0; ch < cend; ch++) {
            const unsigned tile = tiles_a + (ch - c0) * kTmaKA * kTmaTile;
            mbar_wait(bar_a + (ch - c0) * 8, rpar);
            float d[kTmaKA];
#pragma unroll
            for (int kk = 0; kk < kTmaKA; kk++) d[kk] = lds32(tile + kk * kTmaTile + my_off);
#pragma unroll
            for (int kk = 0; kk < kTmaKA; kk++) pacc = pacc + nan_to_zero(d[kk]);  // levels beyond km read as 0
        }
        __syncthreads();  // everybody is through with these tiles
        if (tid == 0 && cend < nchunk) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            for (int ch = cend; ch < cend + nbuf && ch < nchunk; ch++) {
                mbar_expect_tx(bar_a + (ch - cend) * 8, kTmaKA * kTmaTile);
                tma_load4(tiles_a + (ch - cend) * kTmaKA * kTmaTile, &maps.dchunk, (int)(blockIdx.x * kTileX), (int)(blockIdx.y * kTileY),
                          ch * kTmaKA, (int)b0, bar_a + (ch - cend) * 8);
            }
        }
    }
    // the delp tiles are done with: levels 0..7 of all arrays into the ring (delp from L2 now), while
    // the level tables are finished below
    if (tid == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int jj = 0; jj < kTmaSlots; jj++)
            if (jj < km) rg.issue(maps, jj < 4 ? 0u : (unsigned)SM::half, jj & 3, jj);
    }
    for (int t = tid; t < 8 * km * NV; t += kThreads) s_acc[t] = 0.f;
    // coarse interfaces: the cumsum is sequential (exact)
    if (tid < CW) {
        float pc = toa;
        for (int k0 = 0; k0 < km; k0 += 8) {
            float dc[8];
#pragma unroll
            for (int j = 0; j < 8; j++) dc[j] = (k0 + j < km) ? s_lvl[tid * LS + k0 + j].w : 0.f;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                if (k0 + j < km) {
                    pc = pc + nan_to_zero(dc[j]);
                    s_lvl[tid * LS + k0 + j].x = pc;
                }
            }
        }
    }
    __syncthreads();  // coarse interface pressures are in s_lvl[].x
    // per (cell, level): 1/thickness and the mask threshold, in parallel; sentinel entry; the
    // staging ring (which Phase A borrowed) starts out as zeros
    for (int t = tid; t < CW * (km + 1); t += kThreads) {
        const int cc = t / (km + 1), k = t % (km + 1);
        if (k == km) {
            s_lvl[cc * LS + km] = make_float4(INFINITY, 0.f, INFINITY, 0.f);
        } else {
            const float4 e = s_lvl[cc * LS + k];
            const float pn = e.x, dc = e.w;
            const float pc = k > 0 ? s_lvl[cc * LS + k - 1].x : toa;
            float thr = pn;  // PHALF: bottom interface of coarse layer k
            if (mask_mode == B200CS_MASK_PFULL) thr = dc / (log(pn) - log(pc));
            if (mask_mode == B200CS_MASK_NONE) thr = -INFINITY;  // always below the surface pressure
            s_lvl[cc * LS + k].y = 1.0f / (pn - pc);
            s_lvl[cc * LS + k].z = thr;
        }
    }
#pragma unroll
    for (int j = 0; j < kStage; j++) sts128(sbase + SM::stage + (j * kThreads + tid) * 16, make_float4(0.f, 0.f, 0.f, 0.f));
    __syncthreads();

    // ---------------- Phase B: layer-synchronous sweep ---------------------------------------
    if (wid < (int)rg.nact) {  // warps of cells beyond nx have nothing to do (uniform per warp)
        TmaCtx<NP> c;
        c.km = km; c.lane = lane;
        c.wv = wv; c.whalf = whalf; c.psurf = pacc;
        c.lvl = lvl_sh + (wid >> 1) * LS * 16;
        c.acc = sbase + SM::acc + wid * km * NV * 4;
        c.buf = sbase + SM::stage + tid * 16;
        c.par = 0;
        c.hoff = 0;  // levels 0..3 (init) come from half 0; level 4 flips to half 1
        c.psmin = -reduce_max32(-pacc);
        const unsigned src0 = opaque(sbase + SM::ring + my_off);
        float4 v0;  // weighted level-0 values
        {
            f2 qA[NP], qB[NP], qC[NP], qD[NP];
            float d0, d1, d2, d3, p[5];
            mbar_wait(rg.full(0, 0), 0); tma_read<NP, NFL>(src0 + 0 * NA * kTmaTile, qA, d0); rg.release(maps, 0, 0, 0, km, lane);
            mbar_wait(rg.full(0, 1), 0); tma_read<NP, NFL>(src0 + 1 * NA * kTmaTile, qB, d1); rg.release(maps, 0, 1, 1, km, lane);
            mbar_wait(rg.full(0, 2), 0); tma_read<NP, NFL>(src0 + 2 * NA * kTmaTile, qC, d2); rg.release(maps, 0, 2, 2, km, lane);
            mbar_wait(rg.full(0, 3), 0); tma_read<NP, NFL>(src0 + 3 * NA * kTmaTile, qD, d3); rg.release(maps, 0, 3, 3, km, lane);
            p[0] = toa;
            p[1] = p[0] + nan_to_zero(d0);
            p[2] = p[1] + nan_to_zero(d1);
            p[3] = p[2] + nan_to_zero(d2);
            p[4] = p[3] + nan_to_zero(d3);
            c.pacc = p[4];
            float a[2 * NP], b[2 * NP], cc2[2 * NP], dd[2 * NP];
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < NP; i++) {
                a[2 * i] = f2_lo(qA[i]); a[2 * i + 1] = f2_hi(qA[i]);
                b[2 * i] = f2_lo(qB[i]); b[2 * i + 1] = f2_hi(qB[i]);
                cc2[2 * i] = f2_lo(qC[i]); cc2[2 * i + 1] = f2_hi(qC[i]);
                dd[2 * i] = f2_lo(qD[i]); dd[2 * i + 1] = f2_hi(qD[i]);
                s += (a[2 * i] + a[2 * i + 1]) + (b[2 * i] + b[2 * i + 1]) + (cc2[2 * i] + cc2[2 * i + 1]) + (dd[2 * i] + dd[2 * i + 1]);
            }
            c.nanwin = __any_sync(0xffffffffu, s != s) ? 5 : 0;  // warp-uniform
            c.m.init(p, a, b, cc2, dd, iv);
            // output level 0 copies the top source layer (mappm.f90:62-64): its weighted sums go
            // straight into the accumulators; the edge that closes level 0 still runs (it opens
            // level 1) but stages zeros because its threshold is +inf
            const float4 lv = lds128(c.lvl);
            const float wk0 = (lv.z < c.psurf) ? wv : 0.f;
            v0 = make_float4(nan_to_zero(a[0] * wk0), nan_to_zero(a[1] * wk0), NP == 2 ? nan_to_zero(a[2 * NP - 2] * wk0) : 0.f,
                             NP == 2 ? nan_to_zero(a[2 * NP - 1] * wk0) : 0.f);
            c.k = 0; c.kdone = 0; c.pt = toa; c.pb = lv.x; c.rdpc = lv.y; c.thr = INFINITY;
        }
        {
            const float tot = reduce_fields32<NP>(v0, lane);
            const bool fwriter = NP == 2 ? ((lane & 7) == 0) : ((lane & 15) == 0);
            if (fwriter) sts32(c.acc + (NP == 2 ? (lane >> 3) : (lane >> 4)) * 4, tot);  // accumulators start at zero
        }
        tma_emit<NP, false, 0>(c, true);  // layer 0 (km >= 4: never the last)
        // interior layers: the windows rotate by renaming -- the loop is unrolled by four phases,
        // every phase with its own copy of the step AND of the edge code, so that no register is
        // moved where the phases meet
        const int Lend = km - 3;  // last interior layer; Lend % 4 == KR
        int L = 1;
#pragma unroll 1
        for (; L + 3 <= Lend; L += 4) {
            tma_layer<NP, NFL, 1>(c, rg, maps, src0, L);
            tma_layer<NP, NFL, 2>(c, rg, maps, src0, L + 1);
            tma_layer<NP, NFL, 3>(c, rg, maps, src0, L + 2);
            tma_layer<NP, NFL, 0>(c, rg, maps, src0, L + 3);
        }
        if (KR >= 1) tma_layer<NP, NFL, 1>(c, rg, maps, src0, L);
        if (KR >= 2) tma_layer<NP, NFL, 2>(c, rg, maps, src0, L + 1);
        if (KR >= 3) tma_layer<NP, NFL, 3>(c, rg, maps, src0, L + 2);
        c.m.template rotate_to0<(KR + 1) & 3>();  // layer km-2 is entered at phase 0
        // the two bottom layers, once, at phase 0 / 1
        c.m.template step_bottom<0>(iv);
        tma_emit<NP, false, 0>(c, true);
        c.m.template step_last<1>();
        tma_emit<NP, true, 1>(c, true);
    }
    __syncthreads();

    // ---------------- Phase C: the two halves of a cell, divide, store -----------------------
    // Fused all-gather (b200cs_pressure_coarsen_gather), common case: the tile's four coarse values of
    // one (level, field) are 16 contiguous bytes of the output row -- each goes to the local output
    // and straight into this rank's slot of every peer's gather buffer with one 16-byte store over
    // NVLink.  No counters, no fences, nobody waits: the stores of a CTA ride under the arithmetic of
    // the CTAs that follow (the r01 scheme -- last CTA of a tile row copies the finished row -- made every
    // CTA wait for a device-scope atomic and moved the data in bursts at the ends of the waves:
    // +65 us per launch at C384, 5 %; tools/gather_diag.py).
    // (Two ranks only: with seven peers the 16-byte NVLink packets of all CTAs cost 295 us per launch, 0.80 weak-scaling
    // efficiency at N = 8; there the finished rows are copied in 192-byte segments by the last CTA of a tile row, below.)
    bool vecg = peers.n > 0 && peers.n <= 2 && (cxN % 4 == 0) && ((int64_t)blockIdx.x * CW + CW <= cxN) && (peers.field_elems % 4 == 0);
    if (vecg) {
#pragma unroll
        for (int i = 0; i < NFL; i++) vecg = vecg && (((uintptr_t)fp.out[i] & 15) == 0);
        for (int p = 0; p < peers.n; p++) vecg = vecg && (((uintptr_t)peers.base[p] & 15) == 0);
    }
    if (vecg) {
        const bool with_delp = (first == 0 && delp_coarse != nullptr);
        const int nslot = NFL + (with_delp ? 1 : 0);
        for (int t = tid; t < km * nslot; t += kThreads) {
            const int k = t / nslot, i = t % nslot;
            float v[4];
#pragma unroll
            for (int cc = 0; cc < CW; cc++) {
                if (i < NFL) {
                    const float* a = s_acc + ((size_t)(2 * cc) * km + k) * NV;
                    const float* b = s_acc + ((size_t)(2 * cc + 1) * km + k) * NV;
                    const float den = a[2 * NP] + b[2 * NP];
                    v[cc] = (a[i] + b[i]) / den;
                } else {
                    v[cc] = s_lvl[cc * LS + k].w;
                }
            }
            const float4 v4 = make_float4(v[0], v[1], v[2], v[3]);
            const int64_t o = ((b0 * km + k) * cyN + (int64_t)blockIdx.y) * cxN + (int64_t)blockIdx.x * CW;
            if (i < NFL) *reinterpret_cast<float4*>((float*)fp.out[i] + o) = v4;
            const int64_t slot = i < NFL ? (first + i) : peers.nfields;
            for (int p = 0; p < peers.n; p++) *reinterpret_cast<float4*>(peers.base[p] + slot * peers.field_elems + o) = v4;
        }
        return;
    }
    for (int t = tid; t < km * CW; t += kThreads) {
        const int cc = t % CW, k = t / CW;
        const int64_t cyg = blockIdx.y, cxg = (int64_t)blockIdx.x * CW + cc;
        if (cxg < cxN) {
            const float* a = s_acc + ((size_t)(2 * cc) * km + k) * NV;
            const float* b = s_acc + ((size_t)(2 * cc + 1) * km + k) * NV;
            const float den = a[2 * NP] + b[2 * NP];
            const int64_t o = ((b0 * km + k) * cyN + cyg) * cxN + cxg;
#pragma unroll
            for (int i = 0; i < NFL; i++) ((float*)fp.out[i])[o] = (a[i] + b[i]) / den;
        }
    }

    // ---------------- fused all-gather, general case: the last CTA of a tile row copies the row ----
    if (peers.n > 0) pc_gather_tail<NFL>(fp, delp_coarse, peers, first, b0, km, cyN, cxN, tid);
}

template <int NP, int NFL>
size_t pressure_coarsen_tma_smem(int km) {
    constexpr int NV = 2 * NP + 1, CW = kTileX / 8;
    return (size_t)TmaSmem<NFL + 1>::acc + (size_t)CW * sweep_lvl_stride(km) * sizeof(float4) +
           sizeof(float) * (size_t)8 * km * NV + 1024;  // + alignment slack
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn tma_encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// [nb0, km, ny, nx] fp32 with element strides (s0, sl, sy, 1) as a rank-4 tensor map, box (32, 8, bz, 1)
inline bool tma_encode_field(CUtensorMap* m, const void* base, int64_t nb0, int km, int64_t ny, int64_t nx,
                             int64_t s0, int64_t sl, int64_t sy, int bz) {
    EncodeTiledFn fn = tma_encode_fn();
    if (!fn) return false;
    if (nb0 == 1) s0 = sl * km;
    if (((uintptr_t)base & 15) || sy <= 0 || sl <= 0 || s0 <= 0 || (sy & 3) || (sl & 3) || (s0 & 3)) return false;
    const cuuint64_t dims[4] = {(cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)km, (cuuint64_t)nb0};
    const cuuint64_t strides[3] = {(cuuint64_t)sy * 4, (cuuint64_t)sl * 4, (cuuint64_t)s0 * 4};
    const cuuint32_t box[4] = {(cuuint32_t)kTileX, (cuuint32_t)kTileY, (cuuint32_t)bz, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// returns B200CS_OK, an error, or -1 when the arrays cannot be described to TMA (caller falls back)
template <int NP, int NFL, int KR>
int launch_pc_tma(const PcArgs& a, int first) {
    auto kern = pressure_coarsen_tma_kernel<NP, NFL, KR>;
    const size_t smem = pressure_coarsen_tma_smem<NP, NFL>(a.km);
    if (smem > 227 * 1024) return -1;
    TmaMaps maps;
    for (int i = 0; i < NFL; i++)
        if (!tma_encode_field(&maps.lvl[i], a.fields[first + i], a.nb0, a.km, a.ny, a.nx, a.fstride[0], a.fstride[1], a.fstride[2], 1)) return -1;
    if (!tma_encode_field(&maps.lvl[NFL], a.delp, a.nb0, a.km, a.ny, a.nx, a.dstride[0], a.dstride[1], a.dstride[2], 1)) return -1;
    if (!tma_encode_field(&maps.dchunk, a.delp, a.nb0, a.km, a.ny, a.nx, a.dstride[0], a.dstride[1], a.dstride[2], kTmaKA)) return -1;
    for (int i = NFL + 1; i < kTmaMaxArrays; i++) maps.lvl[i] = maps.lvl[0];
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen(tma): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    FieldPtrs<NFL> fp;
    for (int i = 0; i < NFL; i++) { fp.in[i] = a.fields[first + i]; fp.out[i] = a.out[first + i]; }
    dim3 grid((unsigned)((a.nx + kTileX - 1) / kTileX), (unsigned)(a.ny / kTileY), (unsigned)a.nb0);
    PeerOut po{};
    po.n = a.npeer;
    po.nfields = a.nfields;
    po.field_elems = a.nb0 * a.km * (a.ny / 8) * (a.nx / 8);
    po.counters = a.counters;
    if (a.npeer > 0) {
        cudaError_t em = cudaMemsetAsync(a.counters, 0, sizeof(int) * (size_t)grid.y * (size_t)a.nb0, a.st);
        if (em != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen_gather: cudaMemsetAsync: %s", cudaGetErrorString(em));
    }
    for (int p = 0; p < a.npeer && p < 8; p++) po.base[p] = (float*)a.peer[p];
    kern<<<grid, kThreads, smem, a.st>>>(maps, fp, (const float*)a.w, a.wstride[0], a.wstride[1],
                                         first == 0 ? (float*)a.delp_coarse : nullptr, a.km, a.ny, a.nx, (float)a.toa,
                                         a.mask_mode, a.iv, po, first);
    return check_launch("pressure_coarsen(tma)");
}

// the TMA kernel covers: what the sweep kernel covers, full 8 x 8 windows, numpy's pairwise window order
inline bool pc_tma_eligible(const PcArgs& a) {
    return a.fy == 8 && a.fx == 8 && a.seq == 0 && a.ny % 8 == 0 && a.nx % 8 == 0 && a.nx < (int64_t)1 << 31 &&
           a.ny < (int64_t)1 << 31;
}

// fields [first, ...) in groups of 4 / 3 / 2 / 1; -1: not describable to TMA (nothing was launched)
template <int KR>
int launch_pc_tma_fields_kr(const PcArgs& a) {
    int first = 0;
    while (first < a.nfields) {
        const int left = a.nfields - first;
        int rc, n;
        if (left >= 4) { rc = launch_pc_tma<2, 4, KR>(a, first); n = 4; }
        else if (left == 3) { rc = launch_pc_tma<2, 3, KR>(a, first); n = 3; }
        else if (left == 2) { rc = launch_pc_tma<1, 2, KR>(a, first); n = 2; }
        else { rc = launch_pc_tma<1, 1, KR>(a, first); n = 1; }
        if (rc != B200CS_OK) return rc;
        first += n;
    }
    return B200CS_OK;
}
inline int launch_pc_tma_fields(const PcArgs& a) {
    switch ((a.km - 3) & 3) {
        case 0: return launch_pc_tma_fields_kr<0>(a);
        case 1: return launch_pc_tma_fields_kr<1>(a);
        case 2: return launch_pc_tma_fields_kr<2>(a);
        default: return launch_pc_tma_fields_kr<3>(a);
    }
}

}  // namespace b200cs
