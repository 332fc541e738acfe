// pressure_coarsen_ws.cuh -- K6, fourth design: the fused pressure-level coarse-graining kernel with
// WARP-SPECIALISED roles (B200CS_FAST arithmetic, fp32 data, 8 x 8 windows, kord <= 3).  Replaces
// regrid_to_area_weighted_pressure -> regrid_vertical/mappm -> _mask_weights ->
// weighted_block_average (external/vcm/vcm/cubedsphere/regridz.py:31-220,
// external/mappm/mappm/mappm.f90, external/vcm/vcm/cubedsphere/coarsen_restarts.py:485-534).
//
// Why (profiles/r02_tma_g_ncu_summary.txt): with one thread per column carrying the
// geometry of the column AND the PPM state of four fields the kernel needs ~120 registers, 16 warps
// per SM, and is latency-bound per warp (7.6 cycles per instruction and warp).  Here a column is
// served by TWO threads of the same CTA:
//   * a GEOMETRY thread (warps 0-3, 56 registers): pressure cumsum, the six reciprocals of the layer
//     step, the search for the output edges inside the layer, mask and weight of each edge.  Per
//     source layer it writes a 64-byte record {A, B, G1, G2, G3 | close g, t | flags | 2 x edge
//     {c1, c2, h, scale}} into a shared-memory ring;
//   * a FIELD thread (warps 4-7, 104 registers): the PPM coefficients of the fields from the record's
//     factors (ppm_pack.cuh: step_fields), the anchored integrals of the record's edges, the staging
//     ring and the 32-lane reductions -- no pressures, no level tables, no edge search.
// setmaxnreg moves registers between the two warpgroups (128 x 56 + 128 x 104 = 256 x 80), a CTA is
// 16 columns x 8 rows (two coarse cells), three CTAs per SM = 24 warps.  Results are those of
// pressure_coarsen_tma.cuh bit for bit (same ppm_pack.cuh functions on the same values, same
// reduction order).
//
// STATUS (round 2): EXPERIMENTAL, not in the product build (compiled only with -DB200CS_WITH_WS and
// selected with B200CS_DEV_WS=1).  Correct on the bring-up cases (tools/r02_ws_probe.sh: coarse delp
// bit-identical, fields 2.9e-7, partial tiles, rough terrain, the bench slab 3.8e-7) but SLOWER than
// the one-thread-per-column kernel: 1.97 ms against 1.15 ms at C384, N = 4 (profiles/r02_ws_ncu_summary.txt).
// The field warps are the critical role and their layer costs ~240 issued instructions, not the ~165
// planned (record and flag handling, predicates, the ring refill and the reduction decision stay with
// them), at 12 field warps per SM; two roles x four unrolled phases are 190 KB of SASS and the field
// side stalls 30 % of its time on instruction fetch; the geometry warps idle (they sleep between
// polls).  What would have to change is in DESIGN.md 4.
#pragma once
#include <cuda.h>

#include "../common.cuh"
#include "../ppm_pack.cuh"
#include "../pressure_coarsen_sweep.cuh"
#include "../pressure_coarsen_tma.cuh"

namespace b200cs {

constexpr int kWsTX = 16;              // tile: 16 columns x 8 rows = 128 columns = 2 coarse cells
constexpr int kWsCols = 128;
constexpr int kWsThreads = 256;        // warps 0-3 geometry, warps 4-7 fields
constexpr int kWsTile = 512;           // bytes of one array-level tile (8 rows x 64 B, 64-byte swizzle)
constexpr int kWsSlots = 4;            // level ring depth
constexpr int kWsRec = 4;              // record ring depth (rounds) per warp pair
constexpr int kWsRecBytes = 2048;      // one round of one warp: 4 float4 rows x 32 lanes
constexpr int kWsStage = 8;            // per-thread staging ring of weighted output values (levels)
constexpr int kWsKA = 8;               // Phase A: levels per delp tile
constexpr int kWsMaxChunks = 16;

// record flags (R1.w as int)
constexpr int kRfEdges = 3;            // bits 0-1: edges in this record (0..2)
constexpr int kRfFirst = 4;            // first record of the layer: step the fields
constexpr int kRfClose = 8;            // this lane leaves the layer after this record's edges
constexpr int kRfWhole = 16;           // close: no edge of this lane lay inside the layer
constexpr int kRfMore = 32;            // (warp-uniform) another round of records follows for this layer
constexpr int kRfE0Whole = 64, kRfE0Below = 128, kRfE1Whole = 256, kRfE1Below = 512;  // last layer, per edge

template <int NA>
struct WsSmem {
    static constexpr int w = 0;                                   // float [8][16] weights (NaN -> 0)
    static constexpr int den = 512;                               // float [2] cell weight sums
    static constexpr int wred = 528;                              // float [4 warps][2]: {min surface pressure, weight sum} of the warp's columns
    static constexpr int psf = 576;                               // float [128] fine surface pressure of the columns
    static constexpr int bars = 1088;                             // u64 full[kWsSlots] of the level ring
    static constexpr int cnts = bars + 8 * kWsSlots;              // unsigned released[kWsSlots]
    static constexpr int bar_a = cnts + 4 * kWsSlots;             // u64 fullA[kWsMaxChunks] (Phase A tiles)
    static constexpr int rfull = bar_a + 8 * kWsMaxChunks;        // u64 [4 warps][kWsRec]
    static constexpr int rempty = rfull + 8 * 4 * kWsRec;         // u64 [4 warps][kWsRec]
    static constexpr int ring = 2048;                             // kWsSlots x NA tiles
    static constexpr int rec = ring + kWsSlots * NA * kWsTile;    // [4 warps][kWsRec] rounds of records
    static constexpr int stage = rec + 4 * kWsRec * kWsRecBytes;  // float4 [kWsStage][128]
    static constexpr int acc = stage + kWsStage * kWsCols * 16;   // float [4 warps][km][NV]; then float4 [2][LS] level tables
    static_assert(rempty + 8 * 4 * kWsRec <= ring, "fixed part overflows");
};

struct WsMaps {
    CUtensorMap lvl[kTmaMaxArrays];   // [NFL] fields, then delp: box (16, 8, 1, 1)
    CUtensorMap dchunk;               // delp: box (16, 8, kWsKA, 1)
};

__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// the level ring, shared by the geometry warps (delp tile) and the field warps (field tiles): level j
// lives in slot j % 4, its barrier completes with phase parity (j / 4) % 2; the slot is refilled with
// level j + 4 by the LAST warp (of either role) that releases it
template <int NA>
struct WsRing {
    unsigned sb;
    unsigned nact;   // warps that consume the ring
    __device__ __forceinline__ unsigned full(int slot) const { return sb + WsSmem<NA>::bars + slot * 8; }
    __device__ __forceinline__ void issue(const WsMaps& maps, int slot, int level) const {
        const unsigned bar = full(slot);
        mbar_expect_tx(bar, NA * kWsTile);
#pragma unroll
        for (int i = 0; i < NA; i++)
            tma_load4(sb + WsSmem<NA>::ring + (slot * NA + i) * kWsTile, &maps.lvl[i], (int)(blockIdx.x * kWsTX), (int)(blockIdx.y * kTileY),
                      level, (int)blockIdx.z, bar);
    }
    __device__ __forceinline__ void release(const WsMaps& maps, int slot, int level, int km, int lane) const {
        if (lane == 0) {
            unsigned old;
            asm volatile("atom.shared.inc.u32 %0, [%1], %2;" : "=r"(old) : "r"(sb + WsSmem<NA>::cnts + slot * 4), "r"(nact - 1) : "memory");
            if (old == nact - 1) {
                if (level + kWsSlots < km) issue(maps, slot, level + kWsSlots);
            }
        }
    }
};

// one round of records of one warp: 4 float4 rows x 32 lanes (conflict-free 16-byte accesses)
struct WsRecord { float4 r0, r1, r2, r3; };
__device__ __forceinline__ void ws_rec_write(unsigned base, int lane, const WsRecord& r) {
    sts128(base + lane * 16, r.r0);
    sts128(base + 512 + lane * 16, r.r1);
    sts128(base + 1024 + lane * 16, r.r2);
    sts128(base + 1536 + lane * 16, r.r3);
}
__device__ __forceinline__ void ws_rec_read(unsigned base, int lane, WsRecord& r) {
    r.r0 = lds128(base + lane * 16);
    r.r1 = lds128(base + 512 + lane * 16);
    r.r2 = lds128(base + 1024 + lane * 16);
    r.r3 = lds128(base + 1536 + lane * 16);
}

// ------------------------------------------------------------------------------------------------
// geometry side
// ------------------------------------------------------------------------------------------------
struct WsGeo {
    // pressures of the column around the current layer L
    float pe_lo, pe_hi, p2, pacc;   // pe[L], pe[L+1], pe[L+2], pe[L+3] (cumsum of the levels read so far)
    float dpm, dp0, dp1;            // dp[L-1], dp[L], dp[L+1]
    float r12, rdp0;
    // output-edge search
    unsigned lvl;                   // shared address of the warp's cell table: float4 [km + 1] {pb, 1/(pb-pt), threshold, coarse delp}
    int k;
    float pt, pb, rdpc, thr;
    float wv, psurf;
    // record ring
    unsigned rec, rfull, rempty;    // shared addresses: the warp's rounds, its full / empty barriers
    unsigned nrec;                  // rounds produced so far
    int km, lane;
};

// All records of one source layer: up to two output edges per lane and round, more rounds while some
// lane of the warp has more (steep terrain, last layer).  r0 / r1.x carry the layer's step factors in
// the first round.
template <bool LAST>
__device__ __forceinline__ void ws_geo_layer(WsGeo& g, float4 r0, float r1x) {
    const unsigned full = 0xffffffffu;
    bool first = true, closed = LAST;
    bool more;
    do {
        WsRecord rc;
        rc.r0 = r0;
        int flags = first ? kRfFirst : 0;
        float4 e[2];
        e[0] = make_float4(0.f, 0.f, 0.f, 0.f);
        e[1] = e[0];
        float ext[2] = {0.f, 0.f};
        int ne = 0;
#pragma unroll
        for (int j = 0; j < 2; j++) {
            // the level table ends with a +inf sentinel: lanes that are through never qualify
            const bool can = LAST ? (g.k < g.km) : (g.pb <= g.pe_hi);
            const PpmEdgeGeo eg = ppm_edge_geo<LAST>(g.pt, g.pb, g.pe_lo, g.pe_hi, g.rdp0);
            // 1/thickness, the column weight and the below-surface mask enter as one factor
            // (MASK_NONE: thr = -inf; level 0: thr = +inf, see kernel)
            const float scale = (g.thr < g.psurf) ? g.rdpc * g.wv : 0.f;
            if (can) {
                e[j] = make_float4(eg.c1, eg.c2, eg.h, scale);
                ext[j] = eg.ext;
                if (LAST) flags |= (eg.whole ? (j ? kRfE1Whole : kRfE0Whole) : 0) | (eg.below ? (j ? kRfE1Below : kRfE0Below) : 0);
                ne++;
                g.pt = g.pb;
                g.k++;
                const float4 lv = lds128(g.lvl + g.k * 16);
                g.pb = lv.x; g.rdpc = lv.y; g.thr = lv.z;
            }
        }
        const bool lane_more = LAST ? (g.k < g.km) : (g.pb <= g.pe_hi);
        more = __any_sync(full, lane_more);
        flags |= ne | (more ? kRfMore : 0);
        float cg_g = 0.f, cg_t = 0.f;
        if (!lane_more && !closed) {
            closed = true;
            const float gg = g.pe_hi - fmaxf(g.pt, g.pe_lo);  // ppm_close_geo, with t passed on instead of e1, e2
            cg_g = gg;
            cg_t = gg * g.rdp0;
            flags |= kRfClose | (!(g.pt > g.pe_lo) ? kRfWhole : 0);
        }
        if (LAST) { rc.r0.x = ext[0]; rc.r0.y = ext[1]; }
        rc.r1 = make_float4(r1x, cg_g, cg_t, __int_as_float(flags));
        rc.r2 = e[0];
        rc.r3 = e[1];
        const unsigned slot = g.nrec & (kWsRec - 1);
        mbar_wait_sleep(g.rempty + slot * 8, ((g.nrec / kWsRec) & 1u) ^ 1u);  // (a fresh barrier passes the wait for the phase before its first)
        ws_rec_write(g.rec + slot * kWsRecBytes, g.lane, rc);
        __syncwarp();
        if (g.lane == 0) mbar_arrive(g.rfull + slot * 8);
        g.nrec++;
        first = false;
    } while (more);
}

// ------------------------------------------------------------------------------------------------
// field side
// ------------------------------------------------------------------------------------------------
template <int NP>
struct WsFld {
    PpmPack<NP> m;        // only the field state is used (Q, DQ, E, DC, AL, AR, DL, DR, R)
    unsigned buf;         // shared address of this thread's staging ring: float4 [kWsStage] (stride kWsCols)
    unsigned sb;          // shared address of the carve-up's base
    unsigned rec, rfull, rempty;
    unsigned nrec;        // rounds consumed so far
    int k, kdone;
    int nanwin;           // (warp-uniform) > 0: a NaN level is inside the stencil of some lane
    unsigned par;         // phase parity of the level ring slot about to wrap
    int km, lane, fw;     // fw = field warp 0..3
};

// staged levels [kdone, kdone + nlev) of the warp -> accumulators (tma_reduce_batch of the TMA kernel,
// with the context it needs read from shared memory; not inlined)
template <int NP, int NFL>
__device__ __noinline__ void ws_reduce_batch(unsigned sb, int km, int kdone, int nlev, int ftid) {
    const unsigned full = 0xffffffffu;
    constexpr int NV = 2 * NP + 1;
    using SM = WsSmem<NFL + 1>;
    const int lane = ftid & 31, fw = ftid >> 5;
    const int i = lane >> 3, sub = lane & 7;
    const int kk = kdone + i;
    const bool on = i < nlev;
    const unsigned base = sb + SM::stage + fw * 512 + (kk & (kWsStage - 1)) * (kWsCols * 16) + sub * 16;
    const unsigned acc = sb + SM::acc + fw * km * (NV * 4);
    const unsigned lvl = sb + SM::acc + 4 * km * (NV * 4) + (fw >> 1) * sweep_lvl_stride(km) * 16;
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
    float tot = reduce_fields8t<NP>(a, b, lane);
    if (__any_sync(full, tot != tot)) {  // NaN-skipping sums (xarray's coarsen().sum()): rare
#pragma unroll
        for (int t = 0; t < 4; t++) {
            v[t].x = nan_to_zero(v[t].x); v[t].y = nan_to_zero(v[t].y); v[t].z = nan_to_zero(v[t].z); v[t].w = nan_to_zero(v[t].w);
        }
        a = f2_add(f2_add(f2_make(v[0].x, v[0].y), f2_make(v[1].x, v[1].y)), f2_add(f2_make(v[2].x, v[2].y), f2_make(v[3].x, v[3].y)));
        if (NP == 2) b = f2_add(f2_add(f2_make(v[0].z, v[0].w), f2_make(v[1].z, v[1].w)), f2_add(f2_make(v[2].z, v[2].w), f2_make(v[3].z, v[3].w)));
        tot = reduce_fields8t<NP>(a, b, lane);
    }
    // weight sum of the level: every lane of the warp is above the mask threshold (psmin = the
    // smallest surface pressure of the 32 columns) except near the surface
    const float thr = on ? lds32(lvl + kk * 16 + 8) : -INFINITY;
    const float psmin = lds32(sb + SM::wred + fw * 8);
    float rd = lds32(sb + SM::wred + fw * 8 + 4);
    if (__any_sync(full, !(thr < psmin))) {
        const float psurf = lds32(sb + SM::psf + ftid * 4);
        const int r8 = lane >> 2, c4 = lane & 3;
        const float wv0 = lds32(sb + SM::w + (r8 * kWsTX + fw * 4 + c4) * 4);  // NaN -> 0 already
        for (int ii = 0; ii < nlev; ii++) {
            const float thr_i = __shfl_sync(full, thr, ii * 8);
            const float r = reduce_all32(thr_i < psurf ? wv0 : 0.f);
            rd = (i == ii) ? r : rd;
        }
    }
    if (on) {
        const unsigned dst = acc + kk * (NV * 4);
        const int f = NP == 2 ? (sub >> 1) : (sub >> 2);
        if (sub == 1) sts32(dst + 2 * NP * 4, rd);  // a level is reduced once; tma_accumulate_slow adds values only
        else if (NP == 2 ? ((sub & 1) == 0) : ((sub & 3) == 0)) sts32(dst + f * 4, lds32(dst + f * 4) + tot);
    }
    __syncwarp();  // the zeros are in place before any lane stages into these slots again
}

template <int NP, int NFL, bool FLUSH>
__device__ __forceinline__ void ws_reduce(WsFld<NP>& c) {
    const unsigned full = 0xffffffffu;
    const int kend = (int)__reduce_min_sync(full, (unsigned)c.k);
    int n = kend - c.kdone;
    if (!FLUSH && n < 4) {
        if (n <= 0) return;
        if ((int)__reduce_max_sync(full, (unsigned)c.k) - c.kdone < kWsStage - 1) return;
    }
    while (n > 0) {
        const int nl = n < 4 ? n : 4;
        ws_reduce_batch<NP, NFL>(c.sb, c.km, c.kdone, nl, c.fw * 32 + c.lane);
        c.kdone += nl;
        n -= 4;
    }
}

// one round of records of the warp: wait, read, hand the slot back
template <int NP>
__device__ __forceinline__ void ws_fld_read(WsFld<NP>& c, WsRecord& rc) {
    const unsigned slot = c.nrec & (kWsRec - 1);
    mbar_wait(c.rfull + slot * 8, (c.nrec / kWsRec) & 1u);
    ws_rec_read(c.rec + slot * kWsRecBytes, c.lane, rc);
    __syncwarp();
    if (c.lane == 0) mbar_arrive(c.rempty + slot * 8);
    c.nrec++;
}

// the (up to two) output edges of a record and, when the lane leaves the layer with it, the close.
// Common case (no lane within two levels of overflowing its staging ring): straight-line code, the
// staging store predicated on the edge count.
template <int NP, int NFL, int PH, bool LAST>
__device__ __forceinline__ void ws_fld_edges(WsFld<NP>& c, const WsRecord& rc, bool general) {
    const unsigned full = 0xffffffffu;
    constexpr int NV = 2 * NP + 1;
    const int flags = __float_as_int(rc.r1.w);
    const int ne = flags & kRfEdges;
    const bool risky = __any_sync(full, c.k - c.kdone >= kWsStage - 2);  // steep terrain inside the warp's patch, rare
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const bool on = j < ne;
        const float4 e = j ? rc.r3 : rc.r2;
        PpmEdgeGeo eg;
        eg.c1 = e.x; eg.c2 = e.y; eg.h = e.z;
        eg.ext = LAST ? (j ? rc.r0.y : rc.r0.x) : 0.f;
        eg.whole = LAST && (flags & (j ? kRfE1Whole : kRfE0Whole));
        eg.below = LAST && (flags & (j ? kRfE1Below : kRfE0Below));
        f2 o[NP];
        c.m.template edge_fields<LAST>(eg, e.w, o, on, LAST ? c.m.Q[1] : nullptr);  // the last layer runs at phase 1
        if (!risky) {
            if (on) sts128(c.buf + (c.k & (kWsStage - 1)) * (kWsCols * 16), unpack4<NP>(o));
        } else {
            const unsigned overmask = __ballot_sync(full, on && (c.k - c.kdone >= kWsStage));
            if (on) {
                if ((overmask >> c.lane) & 1u)
                    tma_accumulate_slow<NP>(c.sb + WsSmem<NFL + 1>::acc + c.fw * c.km * (NV * 4), c.k, c.lane, unpack4<NP>(o), overmask);
                else
                    sts128(c.buf + (c.k & (kWsStage - 1)) * (kWsCols * 16), unpack4<NP>(o));
            }
        }
        c.k += on ? 1 : 0;
    }
    if (!LAST && (flags & kRfClose)) {
        PpmCloseGeo cg;
        cg.g = rc.r1.y;
        const float t = rc.r1.z, sv = 1.f - t;
        cg.e2 = sv * t;
        cg.e1 = t + cg.e2;
        cg.whole = flags & kRfWhole;
        if (general) c.m.close_general_fields(cg, c.m.Q[PH]);  // (warp-uniform)
        else c.m.close_fields(cg);
    }
}

// LAYER: 0 = layer 0 (fields already initialised), 1 = interior, 2 = layer km-2, 3 = last layer.
// Consumes every round of records of one source layer: the first steps the fields, each evaluates up
// to two output edges of this lane, one closes the layer.  More than one round is rare (steep
// terrain, and the last layer, which emits all remaining levels).
template <int NP, int NFL, int PH, int LAYER>
__device__ __forceinline__ void ws_fld_layer(WsFld<NP>& c, bool general, int iv) {
    constexpr bool LAST = LAYER == 3;
    WsRecord rc;
    ws_fld_read<NP>(c, rc);
    if (LAYER == 1) {
        PpmStepGeo geo;
        geo.A = rc.r0.x; geo.B = rc.r0.y; geo.G1 = rc.r0.z; geo.G2 = rc.r0.w; geo.nG3 = rc.r1.x;
        c.m.template step_fields<PH>(geo, general);
    } else if (LAYER == 2) {
        c.m.template step_bottom_fields<PH>(rc.r0.x, rc.r0.y, iv);
    } else if (LAYER == 3) {
        c.m.template step_last_fields<PH>();
    }
    ws_fld_edges<NP, NFL, PH, LAST>(c, rc, general);
    while (__float_as_int(rc.r1.w) & kRfMore) {  // (warp-uniform)
        ws_fld_read<NP>(c, rc);
        ws_fld_edges<NP, NFL, PH, LAST>(c, rc, general);
    }
    ws_reduce<NP, NFL, LAST>(c);
}

// this thread's field values of one level (an odd count repeats the last field)
template <int NP, int NFL>
__device__ __forceinline__ void ws_read_fields(unsigned src, f2 (&q)[NP]) {
    float v[2 * NP];
#pragma unroll
    for (int i = 0; i < 2 * NP; i++) {
        const int t = i < NFL ? i : NFL - 1;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v[i]) : "r"(src + t * kWsTile));
    }
#pragma unroll
    for (int p = 0; p < NP; p++) q[p] = f2_make(v[2 * p], v[2 * p + 1]);
}

// interior layer L of the field side (PH = L % 4): level L+2 is in the window (slot (PH+2)%4); level
// L+3 is read from the ring into the free slot at the end
template <int NP, int NFL, int PH>
__device__ __forceinline__ void ws_fld_interior(WsFld<NP>& c, const WsRing<NFL + 1>& rg, const WsMaps& maps, unsigned src0, int L, int iv) {
    const unsigned full = 0xffffffffu;
    {   // level L+2 (read at the end of the previous layer): NaN (or inf - inf) entering the stencil is rare
        constexpr int i2 = (PH + 2) & 3;
        f2 s2 = c.m.Q[i2][0];
        if (NP == 2) s2 = f2_add(s2, c.m.Q[i2][NP - 1]);
        const float sv = f2_lo(s2) + f2_hi(s2);
        if (__any_sync(full, sv != sv)) c.nanwin = 4;
    }
    const bool general = (L == 1) || (c.nanwin > 0);  // layer 1: its upper edge value comes from the top boundary formula
    c.nanwin = c.nanwin > 0 ? c.nanwin - 1 : 0;
    ws_fld_layer<NP, NFL, PH, 1>(c, general, iv);
    if (L + 3 < c.km) {
        constexpr int slot = (PH + 3) & 3;  // level L+3 lives in slot (L+3) % 4
        if (slot == 0) c.par ^= 1u;
        mbar_wait(rg.full(slot), c.par);
        ws_read_fields<NP, NFL>(src0 + slot * (NFL + 1) * kWsTile, c.m.Q[slot]);
        rg.release(maps, slot, L + 3, c.km, c.lane);
    }
}

template <int NP, int NFL, int KR>
__global__ void __launch_bounds__(kWsThreads, 3)
pressure_coarsen_ws_kernel(const __grid_constant__ WsMaps maps, FieldPtrs<NFL> fp,
                           const float* __restrict__ w, int64_t ws0, int64_t wsy,
                           float* __restrict__ delp_coarse, int km, int64_t ny, int64_t nx, float toa,
                           int mask_mode, int iv, const __grid_constant__ PeerOut peers, int first) {
    constexpr int NA = NFL + 1;
    constexpr int NV = 2 * NP + 1;
    constexpr int CW = kWsTX / 8;  // cells per tile
    using SM = WsSmem<NA>;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const bool is_fld = wid >= 4;
    const int rw = wid & 3;                 // warp inside its role: columns 4 rw .. 4 rw + 3
    const int ctid = tid & (kWsCols - 1);   // the column inside the tile's thread order (both roles)
    const int r = lane >> 2, c4 = lane & 3;
    const int xl = rw * 4 + c4;
    const int64_t xg = (int64_t)blockIdx.x * kWsTX + xl, yg = (int64_t)blockIdx.y * kTileY + r;
    const int64_t b0 = blockIdx.z;
    const int64_t cxN = nx / 8, cyN = ny / 8;
    const bool active = xg < nx;  // ny % 8 == 0: rows are always inside
    // 64-byte swizzle: 16-byte chunk index (= rw) xor bits 7-8 of the offset (= row >> 1)
    const unsigned my_off = (unsigned)(r * 64 + ((rw ^ ((r >> 1) & 3)) << 4) + c4 * 4);

    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    unsigned char* smem_raw = smem_dyn + ((1024u - ((unsigned)__cvta_generic_to_shared(smem_dyn) & 1023u)) & 1023u);
    const unsigned sbase = opaque((unsigned)__cvta_generic_to_shared(smem_raw));
    const int LS = sweep_lvl_stride(km);
    float* s_acc = reinterpret_cast<float*>(smem_raw + SM::acc);              // [4 field warps][km][NV]
    float4* s_lvl = reinterpret_cast<float4*>(s_acc + (size_t)4 * km * NV);   // [CW][LS], last = sentinel
    float* s_w = reinterpret_cast<float*>(smem_raw + SM::w);                  // [8][16]
    float* s_den = reinterpret_cast<float*>(smem_raw + SM::den);              // [CW]
    const unsigned bar_a = sbase + SM::bar_a;
    const unsigned lvl_sh = sbase + SM::acc + 4 * km * NV * 4;

    WsRing<NA> rg;
    rg.sb = sbase;
    int npair;  // warp pairs (geometry + field) with columns inside the grid
    {
        const int64_t left = nx - (int64_t)blockIdx.x * kWsTX;
        npair = left >= kWsTX ? 4 : (int)(left / 8) * 2;
        rg.nact = 2u * (unsigned)npair;
    }
    const int nchunk = (km + kWsKA - 1) / kWsKA;
    int nbuf = (kWsSlots * NA * kWsTile + 4 * kWsRec * kWsRecBytes + kWsStage * kWsCols * 16 + 4 * km * NV * 4) / (kWsKA * kWsTile);
    nbuf = nbuf < nchunk ? nbuf : nchunk;
    nbuf = nbuf < kWsMaxChunks ? nbuf : kWsMaxChunks;
    const unsigned tiles_a = sbase + SM::ring;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < kWsSlots; i++) {
            mbar_init(rg.full(i), 1);
            sts32(sbase + SM::cnts + i * 4, 0.f);
        }
        for (int i = 0; i < 4 * kWsRec; i++) {
            mbar_init(sbase + SM::rfull + i * 8, 1);
            mbar_init(sbase + SM::rempty + i * 8, 1);
        }
        for (int i = 0; i < nbuf; i++) mbar_init(bar_a + i * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int cch = 0; cch < nbuf; cch++) {
            mbar_expect_tx(bar_a + cch * 8, kWsKA * kWsTile);
            tma_load4(tiles_a + cch * kWsKA * kWsTile, &maps.dchunk, (int)(blockIdx.x * kWsTX),
                      (int)(blockIdx.y * kTileY), cch * kWsKA, (int)b0, bar_a + cch * 8);
        }
    }

    // ---------------- Phase A (all eight warps): coarse pressure levels of the tile's cells ---
    const float wv = active ? __ldg(w + b0 * ws0 + yg * wsy + xg) : 0.f;
    if (!is_fld) s_w[r * kWsTX + xl] = nan_to_zero(wv);
    const float whalf = reduce_all32(nan_to_zero(wv));
    __syncthreads();  // barriers initialised, weights in shared memory
    if (tid < CW) {   // weight sum of a cell: rows pairwise (numpy), then row after row
        float d = 0.f;
        for (int rr = 0; rr < 8; rr++) {
            const float4 a = *reinterpret_cast<const float4*>(s_w + rr * kWsTX + tid * 8);
            const float4 b = *reinterpret_cast<const float4*>(s_w + rr * kWsTX + tid * 8 + 4);
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
        // coarse delp: warp (ch % 8) takes chunk ch, one lane per (level, cell) -- 16 lanes: 64 products
        // delp*w in numpy's order (rows pairwise, then row after row), straight from the swizzled tile
#pragma unroll 1
        for (int ch = c0 + wid; ch < cend; ch += 8) {
            const unsigned tile = tiles_a + (ch - c0) * kWsKA * kWsTile;
            mbar_wait(bar_a + (ch - c0) * 8, rpar);
            if (lane < kWsKA * CW) {
                const int kk = lane >> 1, cc = lane & 1, k = ch * kWsKA + kk;
                const unsigned lt = tile + kk * kWsTile;
                float num = 0.f;
#pragma unroll 2
                for (int rr = 0; rr < 8; rr++) {
                    const int sw = (rr >> 1) & 3;
                    const float4 a = lds128(lt + rr * 64 + (((2 * cc) ^ sw) << 4));
                    const float4 b = lds128(lt + rr * 64 + (((2 * cc + 1) ^ sw) << 4));
                    const float4 wa = *reinterpret_cast<const float4*>(s_w + rr * kWsTX + cc * 8);
                    const float4 wb = *reinterpret_cast<const float4*>(s_w + rr * kWsTX + cc * 8 + 4);
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
        }
        // fine surface pressure: sequential fp32 cumsum of the column (xarray cumsum == nancumsum); the
        // geometry threads need it (both roles run it: the field warps would only wait otherwise)
#pragma unroll 1
        for (int ch = c0; ch < cend; ch++) {
            const unsigned tile = tiles_a + (ch - c0) * kWsKA * kWsTile;
            mbar_wait(bar_a + (ch - c0) * 8, rpar);
            float d[kWsKA];
#pragma unroll
            for (int kk = 0; kk < kWsKA; kk++) d[kk] = lds32(tile + kk * kWsTile + my_off);
#pragma unroll
            for (int kk = 0; kk < kWsKA; kk++) pacc = pacc + nan_to_zero(d[kk]);  // levels beyond km read as 0
        }
        __syncthreads();  // everybody is through with these tiles
        if (tid == 0 && cend < nchunk) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            for (int ch = cend; ch < cend + nbuf && ch < nchunk; ch++) {
                mbar_expect_tx(bar_a + (ch - cend) * 8, kWsKA * kWsTile);
                tma_load4(tiles_a + (ch - cend) * kWsKA * kWsTile, &maps.dchunk, (int)(blockIdx.x * kWsTX), (int)(blockIdx.y * kTileY),
                          ch * kWsKA, (int)b0, bar_a + (ch - cend) * 8);
            }
        }
    }
    // the delp tiles are done with: levels 0..3 of all arrays into the ring (delp from L2 now)
    if (tid == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#pragma unroll
        for (int jj = 0; jj < kWsSlots; jj++)
            if (jj < km) rg.issue(maps, jj, jj);
    }
    for (int t = tid; t < 4 * km * NV; t += kWsThreads) s_acc[t] = 0.f;
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
    for (int t = tid; t < CW * (km + 1); t += kWsThreads) {
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
    // what the (out-of-line) reduction of the field side needs of a column / warp
    if (!is_fld) {
        sts32(sbase + SM::psf + ctid * 4, pacc);
        const float pmin = -reduce_max32(-pacc);
        if (lane == 0) { sts32(sbase + SM::wred + rw * 8, pmin); sts32(sbase + SM::wred + rw * 8 + 4, whalf); }
    }
    if (tid < kWsCols) {
#pragma unroll
        for (int j = 0; j < kWsStage; j++) sts128(sbase + SM::stage + (j * kWsCols + tid) * 16, make_float4(0.f, 0.f, 0.f, 0.f));
    }
    __syncthreads();

    // ---------------- Phase B: the two roles ---------------------------------------------------
    const unsigned src0 = opaque(sbase + SM::ring + my_off);
    const int Lend = km - 3;  // last interior layer; Lend % 4 == KR
    if (!is_fld) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
        if (rw < npair) {  // warps of cells beyond nx have nothing to do (uniform per warp)
            WsGeo g;
            g.km = km; g.lane = lane;
            g.wv = wv; g.psurf = pacc;
            g.lvl = lvl_sh + (rw >> 1) * LS * 16;
            g.rec = sbase + SM::rec + rw * kWsRec * kWsRecBytes;
            g.rfull = sbase + SM::rfull + rw * kWsRec * 8;
            g.rempty = sbase + SM::rempty + rw * kWsRec * 8;
            g.nrec = 0;
            const unsigned dsrc = src0 + NFL * kWsTile;  // the delp tile of a slot
            float d0, d1, d2, d3;
            mbar_wait(rg.full(0), 0); d0 = lds32(dsrc + 0 * NA * kWsTile); rg.release(maps, 0, 0, km, lane);
            mbar_wait(rg.full(1), 0); d1 = lds32(dsrc + 1 * NA * kWsTile); rg.release(maps, 1, 1, km, lane);
            mbar_wait(rg.full(2), 0); d2 = lds32(dsrc + 2 * NA * kWsTile); rg.release(maps, 2, 2, km, lane);
            mbar_wait(rg.full(3), 0); d3 = lds32(dsrc + 3 * NA * kWsTile); rg.release(maps, 3, 3, km, lane);
            const float p0 = toa;
            const float p1 = p0 + nan_to_zero(d0);
            const float p2 = p1 + nan_to_zero(d1);
            const float p3 = p2 + nan_to_zero(d2);
            const float p4 = p3 + nan_to_zero(d3);
            const float e1 = p1 - p0, e2 = p2 - p1, e3 = p3 - p2, e4 = p4 - p3;  // PpmPack::init's d1..d4
            {
                const float4 lv = lds128(g.lvl);
                g.k = 0; g.pt = toa; g.pb = lv.x; g.rdpc = lv.y; g.thr = INFINITY;
            }
            // layer 0
            g.pe_lo = p0; g.pe_hi = p1; g.rdp0 = fast_rcp(e1);
            ws_geo_layer<false>(g, make_float4(e1, e2, e3, e4), 0.f);
            // interior layers
            g.r12 = fast_rcp(e2 + e3);
            g.dpm = e1; g.dp0 = e2; g.dp1 = e3;
            g.pe_lo = p1; g.pe_hi = p2; g.p2 = p3; g.pacc = p4;
            unsigned par = 0;
#pragma unroll 1
            for (int L = 1; L <= Lend; L++) {
                const float pn = g.pacc;                 // pe[L+3]
                const float dp2 = pn - g.p2;
                const float r01 = g.r12;
                g.rdp0 = fast_rcp(g.dp0);
                const PpmStepGeo geo = ppm_step_geo(g.dpm, g.dp0, g.dp1, dp2, r01, g.r12);
                ws_geo_layer<false>(g, make_float4(geo.A, geo.B, geo.G1, geo.G2), geo.nG3);
                g.dpm = g.dp0; g.dp0 = g.dp1; g.dp1 = dp2;
                g.pe_lo = g.pe_hi; g.pe_hi = g.p2; g.p2 = pn;
                if (L + 3 < km) {
                    const int slot = (L + 3) & 3;
                    if (slot == 0) par ^= 1u;
                    mbar_wait_sleep(rg.full(slot), par);
                    const float d = lds32(dsrc + slot * NA * kWsTile);
                    rg.release(maps, slot, L + 3, km, lane);
                    g.pacc = g.pacc + nan_to_zero(d);
                }
            }
            // layer km-2: dp0 = dp[km-2], dp1 = dp[km-1]
            g.rdp0 = fast_rcp(g.dp0);
            ws_geo_layer<false>(g, make_float4(g.dp0, g.dp1, 0.f, 0.f), 0.f);
            // layer km-1
            g.pe_lo = g.pe_hi; g.pe_hi = g.p2;
            g.rdp0 = fast_rcp(g.dp1);
            ws_geo_layer<true>(g, make_float4(0.f, 0.f, 0.f, 0.f), 0.f);
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
        if (rw < npair) {
            WsFld<NP> c;
            c.km = km; c.lane = lane; c.fw = rw;
            c.sb = sbase;
            c.buf = sbase + SM::stage + ctid * 16;
            c.rec = sbase + SM::rec + rw * kWsRec * kWsRecBytes;
            c.rfull = sbase + SM::rfull + rw * kWsRec * 8;
            c.rempty = sbase + SM::rempty + rw * kWsRec * 8;
            c.nrec = 0;
            c.par = 0;
            c.k = 0; c.kdone = 0;
            float4 v0;  // weighted level-0 values
            {
                f2 qA[NP], qB[NP], qC[NP], qD[NP];
                mbar_wait(rg.full(0), 0); ws_read_fields<NP, NFL>(src0 + 0 * NA * kWsTile, qA); rg.release(maps, 0, 0, km, lane);
                mbar_wait(rg.full(1), 0); ws_read_fields<NP, NFL>(src0 + 1 * NA * kWsTile, qB); rg.release(maps, 1, 1, km, lane);
                mbar_wait(rg.full(2), 0); ws_read_fields<NP, NFL>(src0 + 2 * NA * kWsTile, qC); rg.release(maps, 2, 2, km, lane);
                mbar_wait(rg.full(3), 0); ws_read_fields<NP, NFL>(src0 + 3 * NA * kWsTile, qD); rg.release(maps, 3, 3, km, lane);
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
                // the first record of layer 0 carries dp[0..3]: peek at it (it is consumed below)
                mbar_wait(c.rfull, 0);
                const float4 dps = lds128(c.rec + lane * 16);
                c.m.init_fields(dps.x, dps.y, dps.z, dps.w, a, b, cc2, dd, iv);
                // output level 0 copies the top source layer (mappm.f90:62-64): its weighted sums go
                // straight into the accumulators; the edge that closes level 0 still runs (it opens
                // level 1) but stages zeros because its threshold is +inf
                const float4 lv = lds128(lvl_sh + (rw >> 1) * LS * 16);
                const float psurf = lds32(sbase + SM::psf + ctid * 4);
                const float wk0 = (lv.z < psurf) ? wv : 0.f;
                v0 = make_float4(nan_to_zero(a[0] * wk0), nan_to_zero(a[1] * wk0), NP == 2 ? nan_to_zero(a[2 * NP - 2] * wk0) : 0.f,
                                 NP == 2 ? nan_to_zero(a[2 * NP - 1] * wk0) : 0.f);
            }
            {
                const float tot = reduce_fields32<NP>(v0, lane);
                const bool fwriter = NP == 2 ? ((lane & 7) == 0) : ((lane & 15) == 0);
                if (fwriter) sts32(sbase + SM::acc + rw * km * NV * 4 + (NP == 2 ? (lane >> 3) : (lane >> 4)) * 4, tot);  // accumulators start at zero
            }
            ws_fld_layer<NP, NFL, 0, 0>(c, true, iv);  // layer 0
            int L = 1;
#pragma unroll 1
            for (; L + 3 <= Lend; L += 4) {
                ws_fld_interior<NP, NFL, 1>(c, rg, maps, src0, L, iv);
                ws_fld_interior<NP, NFL, 2>(c, rg, maps, src0, L + 1, iv);
                ws_fld_interior<NP, NFL, 3>(c, rg, maps, src0, L + 2, iv);
                ws_fld_interior<NP, NFL, 0>(c, rg, maps, src0, L + 3, iv);
            }
            if (KR >= 1) ws_fld_interior<NP, NFL, 1>(c, rg, maps, src0, L, iv);
            if (KR >= 2) ws_fld_interior<NP, NFL, 2>(c, rg, maps, src0, L + 1, iv);
            if (KR >= 3) ws_fld_interior<NP, NFL, 3>(c, rg, maps, src0, L + 2, iv);
            c.m.template rotate_to0<(KR + 1) & 3>();  // layer km-2 is entered at phase 0
            ws_fld_layer<NP, NFL, 0, 2>(c, true, iv);
            ws_fld_layer<NP, NFL, 1, 3>(c, true, iv);
        }
    }
    __syncthreads();

    // ---------------- Phase C: the two halves of a cell, divide, store -----------------------
    for (int t = tid; t < km * CW; t += kWsThreads) {
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

    // ---------------- fused all-gather (b200cs_pressure_coarsen_gather) ----------------------
    if (peers.n > 0) pc_gather_tail<NFL>(fp, delp_coarse, peers, first, b0, km, cyN, cxN, tid);
}

template <int NP, int NFL>
size_t pressure_coarsen_ws_smem(int km) {
    constexpr int NV = 2 * NP + 1, CW = kWsTX / 8;
    return (size_t)WsSmem<NFL + 1>::acc + (size_t)CW * sweep_lvl_stride(km) * sizeof(float4) +
           sizeof(float) * (size_t)4 * km * NV + 1024;  // + alignment slack
}

// [nb0, km, ny, nx] fp32 with element strides (s0, sl, sy, 1) as a rank-4 tensor map, box (16, 8, bz, 1), 64-byte swizzle
inline bool ws_encode_field(CUtensorMap* m, const void* base, int64_t nb0, int km, int64_t ny, int64_t nx,
                            int64_t s0, int64_t sl, int64_t sy, int bz) {
    EncodeTiledFn fn = tma_encode_fn();
    if (!fn) return false;
    if (nb0 == 1) s0 = sl * km;
    if (((uintptr_t)base & 15) || sy <= 0 || sl <= 0 || s0 <= 0 || (sy & 3) || (sl & 3) || (s0 & 3)) return false;
    const cuuint64_t dims[4] = {(cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)km, (cuuint64_t)nb0};
    const cuuint64_t strides[3] = {(cuuint64_t)sy * 4, (cuuint64_t)sl * 4, (cuuint64_t)s0 * 4};
    const cuuint32_t box[4] = {(cuuint32_t)kWsTX, (cuuint32_t)kTileY, (cuuint32_t)bz, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// returns B200CS_OK, an error, or -1 when the arrays cannot be described to TMA (caller falls back)
template <int NP, int NFL, int KR>
int launch_pc_ws(const PcArgs& a, int first) {
    auto kern = pressure_coarsen_ws_kernel<NP, NFL, KR>;
    const size_t smem = pressure_coarsen_ws_smem<NP, NFL>(a.km);
    if (smem > 227 * 1024) return -1;
    WsMaps maps;
    for (int i = 0; i < NFL; i++)
        if (!ws_encode_field(&maps.lvl[i], a.fields[first + i], a.nb0, a.km, a.ny, a.nx, a.fstride[0], a.fstride[1], a.fstride[2], 1)) return -1;
    if (!ws_encode_field(&maps.lvl[NFL], a.delp, a.nb0, a.km, a.ny, a.nx, a.dstride[0], a.dstride[1], a.dstride[2], 1)) return -1;
    if (!ws_encode_field(&maps.dchunk, a.delp, a.nb0, a.km, a.ny, a.nx, a.dstride[0], a.dstride[1], a.dstride[2], kWsKA)) return -1;
    for (int i = NFL + 1; i < kTmaMaxArrays; i++) maps.lvl[i] = maps.lvl[0];
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen(ws): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    FieldPtrs<NFL> fp;
    for (int i = 0; i < NFL; i++) { fp.in[i] = a.fields[first + i]; fp.out[i] = a.out[first + i]; }
    dim3 grid((unsigned)((a.nx + kWsTX - 1) / kWsTX), (unsigned)(a.ny / kTileY), (unsigned)a.nb0);
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
    kern<<<grid, kWsThreads, smem, a.st>>>(maps, fp, (const float*)a.w, a.wstride[0], a.wstride[1],
                                           first == 0 ? (float*)a.delp_coarse : nullptr, a.km, a.ny, a.nx, (float)a.toa,
                                           a.mask_mode, a.iv, po, first);
    return check_launch("pressure_coarsen(ws)");
}

// fields [first, ...) in groups of 4 / 3 / 2 / 1; -1: not describable to TMA (nothing was launched)
template <int KR>
int launch_pc_ws_fields_kr(const PcArgs& a) {
    int first = 0;
    while (first < a.nfields) {
        const int left = a.nfields - first;
        int rc, n;
        if (left >= 4) { rc = launch_pc_ws<2, 4, KR>(a, first); n = 4; }
        else if (left == 3) { rc = launch_pc_ws<2, 3, KR>(a, first); n = 3; }
        else if (left == 2) { rc = launch_pc_ws<1, 2, KR>(a, first); n = 2; }
        else { rc = launch_pc_ws<1, 1, KR>(a, first); n = 1; }
        if (rc != B200CS_OK) return rc;
        first += n;
    }
    return B200CS_OK;
}
inline int launch_pc_ws_fields(const PcArgs& a) {
    switch ((a.km - 3) & 3) {
        case 0: return launch_pc_ws_fields_kr<0>(a);
        case 1: return launch_pc_ws_fields_kr<1>(a);
        case 2: return launch_pc_ws_fields_kr<2>(a);
        default: return launch_pc_ws_fields_kr<3>(a);
    }
}

}  // namespace b200cs
