// pressure_coarsen_api.cu -- argument checking and dispatch of b200cs_pressure_coarsen
#include "pressure_coarsen.cuh"

namespace b200cs {
int pc_entry_fx2_m0(const PcArgs&, int, int);
int pc_entry_fx4_m0(const PcArgs&, int, int);
int pc_entry_fx8_m0(const PcArgs&, int, int);
int pc_entry_fx2_m1(const PcArgs&, int, int);
int pc_entry_fx4_m1(const PcArgs&, int, int);
int pc_entry_fx8_m1(const PcArgs&, int, int);
bool pc_fast_covers(const PcArgs&, int, int, int);
int pc_fast_entry(const PcArgs&);
int64_t pc_wide_workspace_floats(int64_t nb0, int km, int64_t ny, int64_t nx, int f);
int pc_wide_entry(const PcArgs&, float* ws);
}  // namespace b200cs

using namespace b200cs;

extern "C" int64_t b200cs_pressure_coarsen_gather_workspace_bytes(int64_t nb0, int64_t ny) {
    return (int64_t)sizeof(int) * nb0 * ((ny + kTileY - 1) / kTileY);
}

static int pc_dispatch(int nfields, const void* const* fields, int field_dtype, const int64_t field_stride[3],
                       const void* delp, int p_dtype, const int64_t delp_stride[3], const void* w, int w_dtype,
                       const int64_t w_stride[2], void* const* out, void* delp_coarse, int64_t nb0, int km, int64_t ny,
                       int64_t nx, int fy, int fx, double toa_pressure, int mask_mode, int iv, int kord, int mode,
                       int npeers, void* const* peer_out, void* workspace, int64_t workspace_bytes, void* stream) {
    B200CS_REQUIRE(nfields >= 0 && nfields <= B200CS_MAX_FIELDS, "pressure_coarsen: nfields must be 0..%d", B200CS_MAX_FIELDS);
    B200CS_REQUIRE(delp && w && delp_stride && w_stride, "pressure_coarsen: NULL argument");
    B200CS_REQUIRE(nfields == 0 || (fields && out && field_stride), "pressure_coarsen: NULL field arrays");
    for (int i = 0; i < nfields; i++) B200CS_REQUIRE(fields[i] && out[i], "pressure_coarsen: field %d is NULL", i);
    B200CS_REQUIRE(fx == 2 || fx == 4 || fx == 8, "pressure_coarsen: fx must be 2, 4 or 8 (got %d)", fx);
    B200CS_REQUIRE(fy == fx || fy == 1, "pressure_coarsen: fy must equal fx or be 1 (got %d)", fy);
    B200CS_REQUIRE(nb0 >= 0 && nb0 <= 65535 && ny >= 1 && nx >= 1 && ny % fy == 0 && nx % fx == 0,
                   "pressure_coarsen: extents must be multiples of the window");
    B200CS_REQUIRE((ny + kTileY - 1) / kTileY <= 65535, "pressure_coarsen: ny too large");
    B200CS_REQUIRE(km >= 4, "pressure_coarsen: km=%d; the PPM profile needs at least 4 levels", km);
    B200CS_REQUIRE(mask_mode >= B200CS_MASK_NONE && mask_mode <= B200CS_MASK_PFULL, "pressure_coarsen: bad mask_mode");
    B200CS_REQUIRE((mode & ~(B200CS_FAST | B200CS_SEQ_WINDOW)) == 0, "pressure_coarsen: bad mode");
    const int seq = (mode & B200CS_SEQ_WINDOW) ? 1 : 0;
    mode &= B200CS_FAST;
    B200CS_REQUIRE((field_dtype | 1) == 1 && (p_dtype | 1) == 1 && (w_dtype | 1) == 1,
                   "pressure_coarsen: dtypes must be F32 or F64");
    if (kord > 6)
        return set_err(B200CS_EUNSUPPORTED,
                       "pressure_coarsen: kord=%d needs a two-sweep profile; use the unfused path (b200cs_mappm)", kord);
    if (nb0 == 0) return B200CS_OK;
    PcArgs a{nfields, fields, field_dtype, field_stride, delp, delp_stride, w, w_stride, out, delp_coarse,
             nb0, km, ny, nx, fy, fx, toa_pressure, mask_mode, iv, kord, as_stream(stream), seq};
    a.npeer = npeers;
    a.peer = peer_out;
    if (npeers > 0) {
        B200CS_REQUIRE(npeers <= 8 && peer_out, "pressure_coarsen_gather: 1..8 peers");
        for (int p = 0; p < npeers; p++) B200CS_REQUIRE(peer_out[p], "pressure_coarsen_gather: peer %d is NULL", p);
        B200CS_REQUIRE(delp_coarse, "pressure_coarsen_gather: delp_coarse is part of the gathered slot");
        if (!workspace || workspace_bytes < b200cs_pressure_coarsen_gather_workspace_bytes(nb0, ny))
            return set_err(B200CS_EWORKSPACE, "pressure_coarsen_gather: workspace of %lld bytes needed",
                           (long long)b200cs_pressure_coarsen_gather_workspace_bytes(nb0, ny));
        a.counters = (int*)workspace;
        if (mode != B200CS_FAST || !pc_fast_covers(a, field_dtype, p_dtype, w_dtype) || fy != fx)
            return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen_gather: only the FAST fp32 8x8 kernel (kord <= 3) scatters to peers");
    }
    if (nfields == 0)
        return set_err(B200CS_EINVAL, "pressure_coarsen: nfields == 0 (use b200cs_weighted_block_average for delp)");
    if (mode == B200CS_EXACT) {
        if (fx == 8) return pc_entry_fx8_m0(a, p_dtype, w_dtype);
        if (fx == 4) return pc_entry_fx4_m0(a, p_dtype, w_dtype);
        return pc_entry_fx2_m0(a, p_dtype, w_dtype);
    }
    // the specialised kernel (fp32, 8-wide windows, kord <= 3); everything else: generic kernel + FMA
    if (pc_fast_covers(a, field_dtype, p_dtype, w_dtype)) return pc_fast_entry(a);
    if (fx == 8) return pc_entry_fx8_m1(a, p_dtype, w_dtype);
    if (fx == 4) return pc_entry_fx4_m1(a, p_dtype, w_dtype);
    return pc_entry_fx2_m1(a, p_dtype, w_dtype);
}

extern "C" int b200cs_pressure_coarsen(int nfields, const void* const* fields, int field_dtype,
                                       const int64_t field_stride[3], const void* delp, int p_dtype,
                                       const int64_t delp_stride[3], const void* w, int w_dtype,
                                       const int64_t w_stride[2], void* const* out, void* delp_coarse,
                                       int64_t nb0, int km, int64_t ny, int64_t nx, int fy, int fx,
                                       double toa_pressure, int mask_mode, int iv, int kord, int mode,
                                       void* stream) {
    return pc_dispatch(nfields, fields, field_dtype, field_stride, delp, p_dtype, delp_stride, w, w_dtype, w_stride, out,
                       delp_coarse, nb0, km, ny, nx, fy, fx, toa_pressure, mask_mode, iv, kord, mode, 0, nullptr, nullptr, 0, stream);
}

extern "C" int b200cs_pressure_coarsen_gather(int nfields, const void* const* fields, int field_dtype,
                                              const int64_t field_stride[3], const void* delp, int p_dtype,
                                              const int64_t delp_stride[3], const void* w, int w_dtype,
                                              const int64_t w_stride[2], void* const* out, void* delp_coarse,
                                              int64_t nb0, int km, int64_t ny, int64_t nx, int fy, int fx,
                                              double toa_pressure, int mask_mode, int iv, int kord, int mode,
                                              int npeers, void* const* peer_out, void* workspace,
                                              int64_t workspace_bytes, void* stream) {
    if (npeers < 1) return set_err(B200CS_EINVAL, "pressure_coarsen_gather: npeers must be >= 1");
    return pc_dispatch(nfields, fields, field_dtype, field_stride, delp, p_dtype, delp_stride, w, w_dtype, w_stride, out,
                       delp_coarse, nb0, km, ny, nx, fy, fx, toa_pressure, mask_mode, iv, kord, mode, npeers, peer_out,
                       workspace, workspace_bytes, stream);
}

extern "C" int64_t b200cs_pressure_coarsen_wide_workspace_bytes(int64_t nb0, int km, int64_t ny, int64_t nx, int factor) {
    if (factor < 16 || nb0 < 0 || km < 1 || ny < 1 || nx < 1) return 0;
    return (int64_t)sizeof(float) * pc_wide_workspace_floats(nb0, km, ny, nx, factor);
}

extern "C" int b200cs_pressure_coarsen_wide(int nfields, const void* const* fields, const int64_t field_stride[3],
                                            const void* delp, const int64_t delp_stride[3], const void* w,
                                            const int64_t w_stride[2], void* const* out, void* delp_coarse,
                                            int64_t nb0, int km, int64_t ny, int64_t nx, int factor,
                                            double toa_pressure, int mask_mode, int iv, int kord, void* workspace,
                                            int64_t workspace_bytes, void* stream) {
    B200CS_REQUIRE(nfields >= 1 && nfields <= B200CS_MAX_FIELDS, "pressure_coarsen_wide: nfields must be 1..%d", B200CS_MAX_FIELDS);
    B200CS_REQUIRE(fields && out && field_stride && delp && w && delp_stride && w_stride, "pressure_coarsen_wide: NULL argument");
    for (int i = 0; i < nfields; i++) B200CS_REQUIRE(fields[i] && out[i], "pressure_coarsen_wide: field %d is NULL", i);
    B200CS_REQUIRE(factor == 16 || factor == 32 || factor == 64, "pressure_coarsen_wide: factor must be 16, 32 or 64 (got %d)", factor);
    B200CS_REQUIRE(nb0 >= 0 && nb0 <= 65535 && ny >= 1 && nx >= 1 && ny % factor == 0 && nx % factor == 0,
                   "pressure_coarsen_wide: extents must be multiples of the window");
    B200CS_REQUIRE((ny + kTileY - 1) / kTileY <= 65535, "pressure_coarsen_wide: ny too large");
    B200CS_REQUIRE(km >= 4, "pressure_coarsen_wide: km=%d; the PPM profile needs at least 4 levels", km);
    B200CS_REQUIRE(mask_mode >= B200CS_MASK_NONE && mask_mode <= B200CS_MASK_PFULL, "pressure_coarsen_wide: bad mask_mode");
    if (kord > 3) return set_err(B200CS_EUNSUPPORTED, "pressure_coarsen_wide: kord <= 3 only (got %d)", kord);
    if (nb0 == 0) return B200CS_OK;
    const int64_t need = b200cs_pressure_coarsen_wide_workspace_bytes(nb0, km, ny, nx, factor);
    if (!workspace || workspace_bytes < need)
        return set_err(B200CS_EWORKSPACE, "pressure_coarsen_wide: workspace of %lld bytes needed", (long long)need);
    PcArgs a{nfields, fields, B200CS_F32, field_stride, delp, delp_stride, w, w_stride, out, delp_coarse,
             nb0, km, ny, nx, factor, factor, toa_pressure, mask_mode, iv, kord, as_stream(stream), 0};
    const int64_t span_q = (int64_t)km * field_stride[1] * 4, span_d = (int64_t)km * delp_stride[1] * 4;
    B200CS_REQUIRE(field_stride[1] >= 0 && delp_stride[1] >= 0 && span_q < (int64_t)0xffffffff && span_d < (int64_t)0xffffffff,
                   "pressure_coarsen_wide: a column must span less than 4 GiB");
    return pc_wide_entry(a, (float*)workspace);
}
