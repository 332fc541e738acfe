// pressure_coarsen_sweep.cuh -- K6-fast, second design: the fused pressure-level coarse-graining
// kernel with the SOURCE layer as the outer loop (B200CS_FAST arithmetic, fp32 data, 8-wide
// windows, kord <= 3).  Replaces regrid_to_area_weighted_pressure -> regrid_vertical/mappm ->
// _mask_weights -> weighted_block_average (external/vcm/vcm/cubedsphere/regridz.py:31-220,
// external/mappm/mappm/mappm.f90, external/vcm/vcm/cubedsphere/coarsen_restarts.py:485-534).
//
// One CTA = 8 rows x 32 columns of fine columns, one thread per column (as pressure_coarsen.cuh).
//   Phase A  streams delp once: block average of delp in numpy order (bit-exact) -> coarse
//            interface pressures, 1/thickness and mask threshold per (cell, level) in shared memory.
//   Phase B  every thread sweeps its column from the model top, ONE source layer per iteration
//            for all threads (PpmSweep, ppm_sweep.cuh): level data arrives through a 4-deep
//            cp.async ring in shared memory with compile-time slots; the output edges that fall
//            into the current layer are evaluated in an inner loop and the masked, weighted
//            values are summed over the 8 lanes of a window row (transposing butterfly when the
//            row's lanes sit on the same output level -- the common case -- else shared-memory
//            reductions per lane) into per-row accumulators [row][level][cell][NF+1].
//   Phase C  rows are added in row order, one division per coarse value, coalesced stores.
// No barrier between the start of Phase B and Phase C; HBM traffic = delp + NF fields + weights
// read once (delp a second time through L2), coarse outputs written once.
#pragma once
#include "common.cuh"
#include "ppm_sweep.cuh"
#include "pressure_coarsen_fast.cuh"

namespace b200cs {

template <int NF>
struct SweepCtx {
    PpmSweep<NF> m;
    const char* pf[NF];   // this column's field pointers at the next level to issue
    const char* pd;       // same for delp
    int64_t step_q, step_d;
    unsigned ring;        // shared address of this thread's ring element (slot 0, value 0)
    float pacc;           // pe[] of the last consumed level + 1 (fp32 cumsum, as Phase A)
    // emission state
    const float4* lvl;    // this thread's cell: [km] {pb, 1/(pb-pt), threshold, -}
    float* acc;           // this thread's row accumulators + cell offset: [km][CW][NF+1]
    int k;                // next output level of this lane
    int kdone;            // (warp-uniform) output levels below this one are reduced
    float4* buf;          // this thread's staging ring: [kStage] (stride kThreads)
    float pt, pb, rdpc, thr;
    float wv, wrow, psurf;
    int km, lane, mask_mode;
};

template <int NF>
__device__ __forceinline__ void sweep_issue(SweepCtx<NF>& c, int slot, bool valid) {
    constexpr unsigned kValStride = kThreads * 4, kSlotStride = (NF + 1) * kThreads * 4;
    if (valid) {
        const unsigned dst = c.ring + slot * kSlotStride;
#pragma unroll
        for (int i = 0; i < NF; i++) { cp_async4(dst + i * kValStride, c.pf[i]); c.pf[i] += c.step_q; }
        cp_async4(dst + NF * kValStride, c.pd);
        c.pd += c.step_d;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

template <int NF>
__device__ __forceinline__ void sweep_read(const SweepCtx<NF>& c, int slot, float (&q)[NF], float& d) {
    constexpr unsigned kValStride = kThreads * 4, kSlotStride = (NF + 1) * kThreads * 4;
    const unsigned src = c.ring + slot * kSlotStride;
#pragma unroll
    for (int i = 0; i < NF; i++) asm volatile("ld.shared.f32 %0, [%1];" : "=f"(q[i]) : "r"(src + i * kValStride));
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(d) : "r"(src + NF * kValStride));
}

constexpr int kStage = 8;  // per-thread staging ring of weighted output values (levels)

// Fallback when a lane runs more than kStage levels ahead of the slowest lane of its warp (steep
// terrain inside one 32-column row): the overflowing lanes add their values straight into the
// row's accumulators.  Two lanes collide only when they belong to the same cell AND sit on the
// same level; __match_any_sync groups those, and the members of a group take turns in lane order
// (__syncwarp orders the read-modify-writes), so the sum is deterministic and needs no atomics
// (a shared-memory float atomicAdd compiles to a CAS loop).  On rough terrain the levels of the
// lanes are spread out and one or two turns suffice; the first version ran one butterfly pass
// per distinct level (7.6 passes per edge, profiles/r01_sweep_rough_ncu.txt).
template <int NF>
__device__ __noinline__ void sweep_accumulate_slow(float* acc, int k, int lane, float4 pv, bool on) {
    const unsigned full = 0xffffffffu;
    constexpr int kLevelStride = (kTileX / 8) * (NF + 1);
    const unsigned key = on ? (((unsigned)(lane >> 3) << 16) | (unsigned)k) : (0x80000000u | (unsigned)lane);
    const unsigned peers = __match_any_sync(full, key);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    const int turns = (int)__reduce_max_sync(full, on ? (unsigned)rank + 1u : 0u);
    float* dst = acc + k * kLevelStride;  // the weight sum of every level is formed in sweep_reduce, for all lanes
    const float p[4] = {pv.x, pv.y, pv.z, pv.w};
    for (int t = 0; t < turns; t++) {
        if (on && rank == t) {
#pragma unroll
            for (int i = 0; i < NF; i++) dst[i] = dst[i] + p[i];
        }
        __syncwarp(full);
    }
}

// This lane's weighted output values of level c.k (mask and weight already applied by the edge
// evaluation): NaN -> 0, then into the staging ring.
template <int NF>
__device__ __forceinline__ void sweep_stage(SweepCtx<NF>& c, const float (&o)[NF], bool on) {
    const unsigned full = 0xffffffffu;
    float prod[4];
#pragma unroll
    for (int i = 0; i < 4; i++) prod[i] = i < NF ? o[i < NF ? i : 0] : 0.f;
    // NaN-skipping sums (xarray's coarsen().sum()) and ring overflow are both rare: one vote
    const float chk = (prod[0] + prod[1]) + (prod[2] + prod[3]);
    const bool fits = c.k - c.kdone < kStage;
    const bool over = on && !fits;
    if (__any_sync(full, (chk != chk) || over)) {
#pragma unroll
        for (int i = 0; i < 4; i++) prod[i] = nan_to_zero(prod[i]);
        if (__any_sync(full, over))
            sweep_accumulate_slow<NF>(c.acc, c.k, c.lane, make_float4(prod[0], prod[1], prod[2], prod[3]), over);
    }
    if (on && fits) c.buf[(c.k & (kStage - 1)) * kThreads] = make_float4(prod[0], prod[1], prod[2], prod[3]);
}

// Reduces the staged levels every lane of the warp has passed: [c.kdone, min over lanes of k).
template <int NF>
__device__ __forceinline__ void sweep_reduce(SweepCtx<NF>& c) {
    const unsigned full = 0xffffffffu;
    constexpr int kLevelStride = (kTileX / 8) * (NF + 1);
    const int kend = (int)__reduce_min_sync(full, (unsigned)c.k);
    const int l8 = c.lane & 7;
    const int slot = l8 < NF ? fast_owner<NF>(c.lane) : NF;
    for (int kk = c.kdone; kk < kend; kk++) {
        float4* slot_p = c.buf + (kk & (kStage - 1)) * kThreads;
        const float4 pv = *slot_p;
        // a slot that holds no pending level reads as zero: a lane that overflowed the ring at
        // level kk went through sweep_accumulate_slow and must not be counted again
        *slot_p = make_float4(0.f, 0.f, 0.f, 0.f);
        const float v[4] = {pv.x, pv.y, pv.z, pv.w};
        float vn[NF];
#pragma unroll
        for (int i = 0; i < NF; i++) vn[i] = v[i];
        const float tot = reduce_fields8<NF>(vn, c.lane);
        const bool keep = c.lvl[kk].z < c.psurf;
        float rd = c.wrow;
        if (!__all_sync(full, keep)) rd = reduce_row8(keep ? nan_to_zero(c.wv) : 0.f);
        if (l8 <= NF) {
            float* dst = c.acc + kk * kLevelStride + slot;
            *dst = *dst + (l8 < NF ? tot : rd);
        }
    }
    c.kdone = kend > c.kdone ? kend : c.kdone;
}

// output edges inside the current source layer (all remaining ones when it is the last layer)
template <int NF, bool LAST>
__device__ __forceinline__ void sweep_emit(SweepCtx<NF>& c) {
    const unsigned full = 0xffffffffu;
    while (true) {
        // the level table ends with a +inf sentinel: lanes that are through never qualify
        const bool can = LAST ? (c.k < c.km) : (c.pb <= c.m.pe_hi);
        if (!__any_sync(full, can)) break;
        float o[NF];
#pragma unroll
        for (int i = 0; i < NF; i++) o[i] = 0.f;
        // 1/thickness, the column weight and the below-surface mask enter as one factor
        // (MASK_NONE: thr = -inf; level 0: thr = +inf, see kernel)
        const float scale = (can && c.thr < c.psurf) ? c.rdpc * c.wv : 0.f;
        if (can) c.m.template edge<LAST>(c.pt, c.pb, scale, o);
        sweep_stage<NF>(c, o, can);
        if (can) {
            c.pt = c.pb;
            c.k++;
            const float4 lv = c.lvl[c.k];
            c.pb = lv.x; c.rdpc = lv.y; c.thr = lv.z;
        }
    }
    if (!LAST) c.m.close(c.pt);
    sweep_reduce<NF>(c);
}

// enter source layer L (PH = L%4; L >= 1)
template <int NF, int PH>
__device__ __forceinline__ void sweep_enter(SweepCtx<NF>& c, int L, int iv) {
    const int km = c.km;
    if (L <= km - 3) {
        float qn[NF], d;
        if (L >= 2) {
            // consume level L+2 (slot (PH+2)&3), issue level L+5 into the slot of level L+1
            sweep_issue<NF>(c, (PH + 1) & 3, L + 5 < km);
            asm volatile("cp.async.wait_group 3;" ::: "memory");
            sweep_read<NF>(c, (PH + 2) & 3, qn, d);
            c.pacc = c.pacc + nan_to_zero(d);
        } else {  // L == 1: q[3] and pe[4] are left over from init()
            sweep_issue<NF>(c, (PH + 1) & 3, L + 5 < km);
#pragma unroll
            for (int i = 0; i < NF; i++) qn[i] = c.m.Q[3][i];
        }
        c.m.template step<PH>(c.pacc, qn);
    } else if (L == km - 2) {
        c.m.template step_bottom<PH>(iv);
    } else {
        c.m.template step_last<PH>();
    }
}

// per-cell stride of the level table in float4 units: km+1 entries (the last is a sentinel), padded
// to 1 mod 8 so that the tables of the cells of a tile row start 4 banks apart (a warp reads up to
// four of them in one LDS.128)
__host__ __device__ inline int sweep_lvl_stride(int km) { return ((km + 1 + 6) / 8) * 8 + 1; }

template <int NF, int FY, bool WIDE = false>
__global__ void __launch_bounds__(kThreads, 2)
pressure_coarsen_sweep_kernel(FieldPtrs<NF> fp, int64_t fs0, int64_t fsl, int64_t fsy,
                              const float* __restrict__ delp, int64_t ds0, int64_t dsl, int64_t dsy,
                              const float* __restrict__ w, int64_t ws0, int64_t wsy,
                              float* __restrict__ delp_coarse, int km, int64_t ny, int64_t nx, float toa,
                              int mask_mode, int iv, int seq_i, const __grid_constant__ PeerOut peers, int first,
                              const __grid_constant__ WideArgs wide) {
    constexpr int FX = 8;
    const bool seq = seq_i != 0;
    constexpr int CW = kTileX / FX;   // cells per tile row
    constexpr int CH = kTileY / FY;   // cell rows per tile
    constexpr int NCELL = CW * CH;
    constexpr int NV = NF + 1;
    const int tid = threadIdx.x, lane = tid & 31, ty = tid >> 5;
    const int64_t xg = (int64_t)blockIdx.x * kTileX + lane, yg = (int64_t)blockIdx.y * kTileY + ty;
    const int64_t b0 = blockIdx.z;
    const bool active = (xg < nx) && (yg < ny);
    // out-of-range threads shadow an in-range column with zero weight (keeps the warp converged)
    const int64_t x = xg < nx ? xg : nx - 1, y = yg < ny ? yg : ny - 1;
    const int64_t cxN = nx / FX, cyN = ny / FY;
    // shadow threads also take the cell of the column they shadow (its levels are defined)
    int my_cell = (ty / FY) * CW + (int)(x - (int64_t)blockIdx.x * kTileX) / FX;
    if constexpr (WIDE) my_cell = (int)(x - (int64_t)blockIdx.x * kTileX) / (wide.f < kTileX ? wide.f : kTileX);

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* s_lvl = reinterpret_cast<float4*>(smem_raw);                    // [NCELL][km+1], last = sentinel
    const int LS = sweep_lvl_stride(km);
    float* s_acc = reinterpret_cast<float*>(s_lvl + (size_t)NCELL * LS);  // [8][km][CW][NV]
    float* s_ring = s_acc + (size_t)kTileY * km * CW * NV;                  // [4][NV][256]
    float* s_rows = s_ring + 4 * NV * kThreads;                             // [2][KA][8][CW]
    float* s_den = s_rows + 2 * kFastKA * kTileY * CW;                      // [NCELL]
    float4* s_buf = reinterpret_cast<float4*>(s_den + ((NCELL + 3) & ~3));  // [kStage][256]

    SweepCtx<NF> c;
    const float* dcol = delp + b0 * ds0 + y * dsy + x;
    {   // start the level ring before Phase A: levels 0..3
        const int64_t fbase = b0 * fs0 + y * fsy + x;
#pragma unroll
        for (int i = 0; i < NF; i++) c.pf[i] = (const char*)((const float*)fp.in[i] + fbase);
        c.pd = (const char*)dcol;
        c.step_q = fsl * 4; c.step_d = dsl * 4;
        c.ring = (unsigned)__cvta_generic_to_shared(s_ring + tid);
#pragma unroll
        for (int j = 0; j < 4; j++) sweep_issue<NF>(c, j, true);  // km >= 4
    }
    for (int t = tid; t < kTileY * km * CW * NV; t += kThreads) s_acc[t] = 0.f;

    // ---------------- Phase A: coarse pressure levels of the tile's cells -------------------
    const float wv = active ? __ldg(w + b0 * ws0 + y * wsy + x) : 0.f;
    const float wrow = reduce_row8(nan_to_zero(wv), seq, lane);  // weight sum of this window row
    float pacc = toa;
    if constexpr (!WIDE) {
        if ((lane & (FX - 1)) == 0) s_rows[ty * CW + lane / FX] = wrow;
        __syncthreads();
        if (tid < NCELL) {
            const int cyi = tid / CW, cxi = tid % CW;
            float d = 0.f;
            for (int r = 0; r < FY; r++) d = d + s_rows[(cyi * FY + r) * CW + cxi];
            s_den[tid] = d;
        }
        __syncthreads();
        // Products delp*w go to shared memory level by level (conflict-free: one 128-B line per row);
        // one thread per (level, cell) then adds its window rows from two LDS.128 each, in numpy's
        // order and entirely in registers -- no shuffles.  The staging ring of Phase B is not in use
        // yet and lends its memory.
        constexpr int kProdLevel = kTileY * kTileX + 4;  // +4 words: levels start 4 banks apart
        float* s_prod = reinterpret_cast<float*>(s_buf);  // [2][KA][kProdLevel]
        static_assert(2 * kFastKA * kProdLevel * sizeof(float) <= kStage * kThreads * sizeof(float4), "s_prod must fit in s_buf");
        float dn[kFastKA];  // next chunk of delp, in flight while the current one is reduced
    #pragma unroll
        for (int kk = 0; kk < kFastKA; kk++) dn[kk] = (kk < km) ? __ldg(dcol + (int64_t)kk * dsl) : 0.f;
        for (int k0 = 0; k0 < km; k0 += kFastKA) {
            const int buf = (k0 / kFastKA) & 1;
            float d[kFastKA], pr[kFastKA];
    #pragma unroll
            for (int kk = 0; kk < kFastKA; kk++) d[kk] = dn[kk];
    #pragma unroll
            for (int kk = 0; kk < kFastKA; kk++)
                dn[kk] = (k0 + kFastKA + kk < km) ? __ldg(dcol + (int64_t)(k0 + kFastKA + kk) * dsl) : 0.f;
            float chk = 0.f;
    #pragma unroll
            for (int kk = 0; kk < kFastKA; kk++) { pr[kk] = d[kk] * wv; chk += pr[kk]; }
            if (__any_sync(0xffffffffu, chk != chk)) {  // NaN in delp or weights: rare, exact handling
    #pragma unroll
                for (int kk = 0; kk < kFastKA; kk++) { d[kk] = nan_to_zero(d[kk]); pr[kk] = nan_to_zero(pr[kk]); }
            }
            float* sp = s_prod + (size_t)buf * kFastKA * kProdLevel + tid;
    #pragma unroll
            for (int kk = 0; kk < kFastKA; kk++) {
                pacc = pacc + d[kk];  // xarray cumsum == nancumsum (levels beyond km add 0)
                sp[kk * kProdLevel] = pr[kk];
            }
            __syncthreads();
            // KA*NCELL (level, cell) sums per chunk, handed to a different warp each chunk
            const int t = (tid - (k0 / kFastKA) * 32) & (kThreads - 1);
            if (t < kFastKA * NCELL) {
                const int kk = t / NCELL, cc = t % NCELL, k = k0 + kk;
                if (k < km) {
                    const int cyi = cc / CW, cxi = cc % CW;
                    const float* src = s_prod + (size_t)(buf * kFastKA + kk) * kProdLevel + cxi * FX;
                    float num = 0.f;
    #pragma unroll
                    for (int r = 0; r < FY; r++) {
                        const float4 a = *reinterpret_cast<const float4*>(src + (cyi * FY + r) * kTileX);
                        const float4 b = *reinterpret_cast<const float4*>(src + (cyi * FY + r) * kTileX + 4);
                        float rs;
                        if (seq) rs = ((((((((0.f + a.x) + a.y) + a.z) + a.w) + b.x) + b.y) + b.z) + b.w);
                        else rs = ((a.x + a.y) + (a.z + a.w)) + ((b.x + b.y) + (b.z + b.w));
                        num = num + rs;
                    }
                    const float dc = num / s_den[cc];
                    s_lvl[cc * LS + k].w = dc;
                    const int64_t cyg = (int64_t)blockIdx.y * CH + cyi, cxg = (int64_t)blockIdx.x * CW + cxi;
                    if (delp_coarse && cyg < cyN && cxg < cxN) delp_coarse[((b0 * km + k) * cyN + cyg) * cxN + cxg] = dc;
                }
            }
            // s_prod is double-buffered: the barrier of the next chunk orders these reads before the
            // chunk after that overwrites this half
        }
    } else {
        // coarse delp of the tile's cell(s) from the block-average pre-pass; the fine surface
        // pressure is the sequential cumsum of the column (the mask compares against it)
        const int ncx = wide.f < kTileX ? kTileX / wide.f : 1;
        const int64_t cyW = ny / wide.f, cxW = nx / wide.f;
        const int64_t cyg = ((int64_t)blockIdx.y * kTileY) / wide.f;
        for (int t = tid; t < ncx * km; t += kThreads) {
            const int cc = t / km, k = t % km;
            const int64_t cxg = ((int64_t)blockIdx.x * kTileX) / wide.f + cc;
            float dc = 0.f;
            if (cyg < cyW && cxg < cxW) dc = __ldg(wide.delp_c + ((b0 * km + k) * cyW + cyg) * cxW + cxg);
            s_lvl[cc * LS + k].w = dc;
        }
#pragma unroll 8
        for (int k = 0; k < km; k++) pacc = pacc + nan_to_zero(__ldg(dcol + (int64_t)k * dsl));
    }
    __syncthreads();
    // coarse interfaces: the cumsum is sequential (exact) and touches nothing else; meanwhile the
    // other warps set their column machines up (fine data only)
    if (tid < NCELL) {
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

    // ---------------- Phase B: layer-synchronous sweep ---------------------------------------
    c.km = km; c.lane = lane; c.mask_mode = mask_mode;
    c.wv = wv; c.wrow = wrow; c.psurf = pacc;
    c.lvl = s_lvl + my_cell * LS;
    c.acc = s_acc + (size_t)ty * km * CW * NV + (lane / FX) * NV;  // own window row (shadow rows: unused cells)
    // Rows beyond ny (edge-line launches: ny = n/f + 1 lines in tiles of 8) belong to no cell: their
    // warps take part in the barriers and skip the sweep (uniform per warp).  They used to shadow
    // the last row with zero weight -- with cells of zero weight, hence NaN levels, hence every
    // output edge on the overflow path: time only, the results were discarded.
    const bool row_ok = yg < ny;
    {
        float qA[NF], qB[NF], qC[NF], qD[NF], d0, d1, d2, d3, p[5];
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        if (row_ok) {
            sweep_read<NF>(c, 0, qA, d0);
            sweep_read<NF>(c, 1, qB, d1);
            sweep_read<NF>(c, 2, qC, d2);
            sweep_read<NF>(c, 3, qD, d3);
            p[0] = toa;
            p[1] = p[0] + nan_to_zero(d0);
            p[2] = p[1] + nan_to_zero(d1);
            p[3] = p[2] + nan_to_zero(d2);
            p[4] = p[3] + nan_to_zero(d3);
            c.pacc = p[4];
            c.m.init(p, qA, qB, qC, qD, iv);
        }
        __syncthreads();  // coarse interface pressures are in s_lvl[].x
        // per (cell, level): 1/thickness and the mask threshold, in parallel; sentinel entry; the
        // staging ring (which Phase A borrowed) starts out as zeros
        for (int t = tid; t < NCELL * (km + 1); t += kThreads) {
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
        for (int j = 0; j < kStage; j++) s_buf[j * kThreads + tid] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        if (row_ok) {
            sweep_issue<NF>(c, 0, 4 < km);
            sweep_issue<NF>(c, 1, 5 < km);
            // output level 0 copies the top source layer (mappm.f90:62-64): its weighted sums go
            // straight into the accumulators; the edge that closes level 0 still runs (it opens level
            // 1) but stages zeros because its threshold is +inf
            const float4 lv = c.lvl[0];
            c.k = 0; c.kdone = 0; c.pt = toa; c.pb = lv.x; c.rdpc = lv.y; c.thr = INFINITY;
            c.buf = s_buf + tid;
            const float wk0 = (lv.z < c.psurf) ? wv : 0.f;
            float v[NF];
#pragma unroll
            for (int i = 0; i < NF; i++) v[i] = nan_to_zero(qA[i] * wk0);
            const float tot = reduce_fields8<NF>(v, lane);
            if ((lane & 7) < NF) c.acc[fast_owner<NF>(lane)] = tot;  // accumulators start at zero
        }
    }
    if (row_ok) {
        sweep_emit<NF, false>(c);  // layer 0 (km >= 4: never the last)
        // the windows rotate by renaming (four instances of the layer step, selected by a uniform
        // branch); the edge / staging / reduction code exists once
#pragma unroll 1
        for (int L = 1; L < km; L++) {
            switch (L & 3) {
                case 1: sweep_enter<NF, 1>(c, L, iv); break;
                case 2: sweep_enter<NF, 2>(c, L, iv); break;
                case 3: sweep_enter<NF, 3>(c, L, iv); break;
                default: sweep_enter<NF, 0>(c, L, iv); break;
            }
            if (L < km - 1) sweep_emit<NF, false>(c);
        }
        sweep_emit<NF, true>(c);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---------------- Phase C: rows in row order, divide, store ------------------------------
    if constexpr (!WIDE) {
        for (int t = tid; t < km * NCELL; t += kThreads) {
            const int cc = t % NCELL, k = t / NCELL;
            const int cyi = cc / CW, cxi = cc % CW;
            const int64_t cyg = (int64_t)blockIdx.y * CH + cyi, cxg = (int64_t)blockIdx.x * CW + cxi;
            if (cyg < cyN && cxg < cxN) {
                float num[NF], den = 0.f;
    #pragma unroll
                for (int i = 0; i < NF; i++) num[i] = 0.f;
    #pragma unroll
                for (int r = 0; r < FY; r++) {
                    const float* src = s_acc + ((size_t)((cyi * FY + r) * km + k) * CW + cxi) * NV;
    #pragma unroll
                    for (int i = 0; i < NF; i++) num[i] = num[i] + src[i];
                    den = den + src[NF];
                }
                const int64_t o = ((b0 * km + k) * cyN + cyg) * cxN + cxg;
    #pragma unroll
                for (int i = 0; i < NF; i++) ((float*)fp.out[i])[o] = num[i] / den;
            }
        }
    } else {
        // partial sums of this CTA per (level, cell of the tile): rows, then the 8-lane groups of
        // the cell, in a fixed order; wide_finalize_kernel adds the CTAs of a cell and divides
        const int ncx = wide.f < kTileX ? kTileX / wide.f : 1;
        const int gpc = CW / ncx;
        const int64_t cta = ((int64_t)b0 * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        for (int t = tid; t < km * ncx * NV; t += kThreads) {
            const int v = t % NV, cc = (t / NV) % ncx, k = t / (NV * ncx);
            float sum = 0.f;
            for (int r = 0; r < kTileY; r++)
                for (int g = 0; g < gpc; g++) sum = sum + s_acc[((size_t)(r * km + k) * CW + cc * gpc + g) * NV + v];
            wide.partial[((cta * km + k) * ncx + cc) * NV + v] = sum;
        }
    }

    // ---------------- fused all-gather (b200cs_pressure_coarsen_gather) ----------------------
    // The CTAs of one row of tiles (same blockIdx.y, b0) produce CH complete coarse rows.  The
    // last of them to finish copies those rows -- contiguous cxN-float segments, read back through
    // L2 -- into this rank's slot on every peer with 16-byte stores over NVLink: full-size
    // packets instead of the 4-byte scattered stores of Phase C, no separate collective kernel.
    if (peers.n > 0) {
        __shared__ int s_last;
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const int prev = atomicAdd(&peers.counters[b0 * gridDim.y + blockIdx.y], 1);
            s_last = (prev == (int)gridDim.x - 1);
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            const int64_t rows = CH;                       // coarse rows of this row of tiles
            const int64_t cy0 = (int64_t)blockIdx.y * CH;
            const bool vec = (cxN % 4 == 0);
            const int64_t per_row = vec ? cxN / 4 : cxN;
            const int64_t total = (int64_t)km * rows * per_row;
#pragma unroll
            for (int sl = 0; sl <= NF; sl++) {             // compile-time slot index: no local copy of fp
                const float* src = sl < NF ? (const float*)fp.out[sl < NF ? sl : 0] : delp_coarse;
                if (sl == NF && !(first == 0 && delp_coarse)) continue;
                const int64_t slot = sl < NF ? (first + sl) : peers.nfields;
                for (int64_t e = tid; e < total; e += kThreads) {
                    const int64_t c = e % per_row;
                    const int64_t r = (e / per_row) % rows;
                    const int64_t k = e / (per_row * rows);
                    if (cy0 + r >= cyN) continue;
                    const int64_t o = ((b0 * km + k) * cyN + cy0 + r) * cxN;
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
    }
}

template <int NF, int FY>
size_t pressure_coarsen_sweep_smem(int km) {
    constexpr int CW = kTileX / 8, NCELL = CW * (kTileY / FY), NV = NF + 1;
    return (size_t)NCELL * sweep_lvl_stride(km) * sizeof(float4) +
           sizeof(float) * ((size_t)kTileY * km * CW * NV + 4 * NV * kThreads + 2 * kFastKA * kTileY * CW + ((NCELL + 3) & ~3)) +
           sizeof(float4) * kStage * kThreads;
}

template <int NF, int FY>
int launch_pc_sweep(const PcArgs& a, int first) {
    auto kern = pressure_coarsen_sweep_kernel<NF, FY>;
    const size_t smem = pressure_coarsen_sweep_smem<NF, FY>(a.km);
    if (smem > 227 * 1024)
        return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen(fast): km=%d needs %zu B of shared memory (> 227 KB)", a.km, smem);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen(fast): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    FieldPtrs<NF> fp;
    for (int i = 0; i < NF; i++) { fp.in[i] = a.fields[first + i]; fp.out[i] = a.out[first + i]; }
    dim3 grid((unsigned)((a.nx + kTileX - 1) / kTileX), (unsigned)((a.ny + kTileY - 1) / kTileY), (unsigned)a.nb0);
    PeerOut po{};
    po.n = a.npeer;
    po.nfields = a.nfields;
    po.field_elems = a.nb0 * a.km * (a.ny / FY) * (a.nx / 8);
    po.counters = a.counters;
    if (a.npeer > 0) {
        cudaError_t em = cudaMemsetAsync(a.counters, 0, sizeof(int) * (size_t)grid.y * (size_t)a.nb0, a.st);
        if (em != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen_gather: cudaMemsetAsync: %s", cudaGetErrorString(em));
    }
    for (int p = 0; p < a.npeer && p < 8; p++) po.base[p] = (float*)a.peer[p];
    kern<<<grid, kThreads, smem, a.st>>>(fp, a.fstride[0], a.fstride[1], a.fstride[2], (const float*)a.delp,
                                         a.dstride[0], a.dstride[1], a.dstride[2], (const float*)a.w, a.wstride[0],
                                         a.wstride[1], first == 0 ? (float*)a.delp_coarse : nullptr, a.km, a.ny, a.nx,
                                         (float)a.toa, a.mask_mode, a.iv, a.seq, po, first, WideArgs{});
    return check_launch("pressure_coarsen(fast)");
}

template <int FY>
int launch_pc_sweep_fields(const PcArgs& a) {
    int first = 0;
    while (first < a.nfields) {
        const int left = a.nfields - first;
        int rc;
        if (left >= 4) { rc = launch_pc_sweep<4, FY>(a, first); first += 4; }
        else if (left >= 2) { rc = launch_pc_sweep<2, FY>(a, first); first += 2; }
        else { rc = launch_pc_sweep<1, FY>(a, first); first += 1; }
        if (rc != B200CS_OK) return rc;
    }
    return B200CS_OK;
}

// ---- wide windows (f = 16, 32, 64) -----------------------------------------------------------

template <int NF>
__global__ void wide_finalize_kernel(const float* __restrict__ partial, FieldPtrs<NF> fp, int64_t nb0, int km,
                                     int64_t cyW, int64_t cxW, int f, int gx, int gy) {
    constexpr int NV = NF + 1;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nb0 * km * cyW * cxW) return;
    const int64_t cxg = t % cxW, cyg = (t / cxW) % cyW, k = (t / (cxW * cyW)) % km, b0 = t / (cxW * cyW * km);
    const int ncx = f < kTileX ? kTileX / f : 1, ntx = f < kTileX ? 1 : f / kTileX, nty = f / kTileY;
    const int64_t bx0 = f < kTileX ? cxg / ncx : cxg * ntx, by0 = cyg * nty;
    const int cxl = f < kTileX ? (int)(cxg % ncx) : 0;
    float num[NF], den = 0.f;
#pragma unroll
    for (int i = 0; i < NF; i++) num[i] = 0.f;
    for (int ry = 0; ry < nty; ry++)
        for (int rx = 0; rx < ntx; rx++) {
            const int64_t cta = (b0 * gy + by0 + ry) * gx + bx0 + rx;
            const float* p = partial + ((cta * km + k) * ncx + cxl) * NV;
#pragma unroll
            for (int i = 0; i < NF; i++) num[i] = num[i] + p[i];
            den = den + p[NF];
        }
#pragma unroll
    for (int i = 0; i < NF; i++) ((float*)fp.out[i])[t] = num[i] / den;
}

inline int64_t pc_wide_partial_floats(int64_t nb0, int km, int64_t ny, int64_t nx, int f) {
    const int64_t gx = (nx + kTileX - 1) / kTileX, gy = (ny + kTileY - 1) / kTileY;
    return nb0 * gy * gx * km * (f < kTileX ? kTileX / f : 1) * 5;  // NV <= 5
}

template <int NF>
int launch_pc_sweep_wide(const PcArgs& a, int first, const float* delp_c, float* partial) {
    auto kern = pressure_coarsen_sweep_kernel<NF, 8, true>;
    const size_t smem = pressure_coarsen_sweep_smem<NF, 8>(a.km);
    if (smem > 227 * 1024)
        return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen(wide): km=%d needs %zu B of shared memory (> 227 KB)", a.km, smem);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return set_err(B200CS_ECUDA, "pressure_coarsen(wide): cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    FieldPtrs<NF> fp;
    for (int i = 0; i < NF; i++) { fp.in[i] = a.fields[first + i]; fp.out[i] = a.out[first + i]; }
    dim3 grid((unsigned)((a.nx + kTileX - 1) / kTileX), (unsigned)((a.ny + kTileY - 1) / kTileY), (unsigned)a.nb0);
    WideArgs wa{a.fx, delp_c, partial};
    kern<<<grid, kThreads, smem, a.st>>>(fp, a.fstride[0], a.fstride[1], a.fstride[2], (const float*)a.delp,
                                         a.dstride[0], a.dstride[1], a.dstride[2], (const float*)a.w, a.wstride[0],
                                         a.wstride[1], nullptr, a.km, a.ny, a.nx, (float)a.toa, a.mask_mode, a.iv, 0,
                                         PeerOut{}, first, wa);
    int rc = check_launch("pressure_coarsen(wide)");
    if (rc != B200CS_OK) return rc;
    const int64_t cyW = a.ny / a.fx, cxW = a.nx / a.fx, n = a.nb0 * a.km * cyW * cxW;
    wide_finalize_kernel<NF><<<(unsigned)((n + 255) / 256), 256, 0, a.st>>>(partial, fp, a.nb0, a.km, cyW, cxW, a.fx,
                                                                           (int)grid.x, (int)grid.y);
    return check_launch("pressure_coarsen(wide finalize)");
}

inline int launch_pc_sweep_wide_fields(const PcArgs& a, const float* delp_c, float* partial) {
    int first = 0;
    while (first < a.nfields) {
        const int left = a.nfields - first;
        int rc;
        if (left >= 4) { rc = launch_pc_sweep_wide<4>(a, first, delp_c, partial); first += 4; }
        else if (left >= 2) { rc = launch_pc_sweep_wide<2>(a, first, delp_c, partial); first += 2; }
        else { rc = launch_pc_sweep_wide<1>(a, first, delp_c, partial); first += 1; }
        if (rc != B200CS_OK) return rc;
    }
    return B200CS_OK;
}

}  // namespace b200cs
