"""ctypes binding of libb200cs.so (include/b200cs.h).

There is NO CPU fallback: if the CUDA library is missing or a call fails this raises.
PyTorch is only the device-buffer carrier (``tensor.data_ptr()``) and stream provider.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200cs.so")

F32, F64 = 0, 1
OP_SUM, OP_MEAN, OP_MIN, OP_MAX, OP_MEDIAN, OP_MODE = range(6)
MASK_NONE, MASK_PHALF, MASK_PFULL = 0, 1, 2
EXACT, FAST = 0, 1
EW_MUL, EW_WHERE, EW_WHERE_NOT, EW_SELECT, EW_FILLNA, EW_MIN_SCALAR, EW_DIFF_OVER = range(7)
CMP_ISCLOSE, CMP_LT, CMP_GT = range(3)
SEQ_WINDOW = 2
MAX_FIELDS = 16

EXPORTS = [
    "b200cs_version", "b200cs_last_error", "b200cs_weighted_block_average", "b200cs_block_reduce",
    "b200cs_block_upsample", "b200cs_stagger", "b200cs_gather_columns", "b200cs_transpose_last2",
    "b200cs_pressure_at_interface", "b200cs_mappm_workspace_bytes", "b200cs_mappm",
    "b200cs_pressure_coarsen", "b200cs_mask_weights", "b200cs_surface_pressure", "b200cs_blending_weights",
    "b200cs_blend", "b200cs_hydrostatic_balance", "b200cs_elementwise", "b200cs_compare_scalar",
    "b200cs_dominant_mask", "b200cs_pressure_at_midpoint_log", "b200cs_exposed_area_weights",
    "b200cs_pressure_coarsen_gather", "b200cs_pressure_coarsen_gather_workspace_bytes",
    "b200cs_pressure_coarsen_wide", "b200cs_pressure_coarsen_wide_workspace_bytes", "b200cs_mappm_fast",
    "b200cs_host_copy", "b200cs_mappm_static",
]

_lib = None


class B200csError(RuntimeError):
    pass


def load():
    """Loads the shared library; raises (loudly) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200csError(
            f"{LIB_PATH} is missing: build it with `python fv3net_b200/csrc/build.py` "
            "(or __graft_entry__.build()).  fv3net_b200 has no CPU fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, dbl = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
    p64 = ctypes.POINTER(ctypes.c_int64)
    pvp = ctypes.POINTER(ctypes.c_void_p)
    lib.b200cs_version.restype = i32
    lib.b200cs_last_error.restype = ctypes.c_char_p
    lib.b200cs_weighted_block_average.argtypes = [vp, i32, p64, vp, i32, p64, vp, i32, p64, vp, i32,
                                                  i64, i64, i64, i64, i32, i32, i32, i32, vp]
    lib.b200cs_block_reduce.argtypes = [vp, i32, p64, vp, i64, i64, i64, i32, i32, i32, i32, i32, i32, i32, dbl, vp]
    lib.b200cs_block_upsample.argtypes = [vp, vp, i32, i64, i64, i64, i32, i32, i32, i32, vp]
    lib.b200cs_stagger.argtypes = [vp, vp, i32, i64, i64, i32, i32, vp]
    lib.b200cs_gather_columns.argtypes = [vp, p64, vp, i32, i64, i64, i64, i32, vp]
    lib.b200cs_transpose_last2.argtypes = [vp, vp, i32, i64, i64, i64, vp]
    lib.b200cs_pressure_at_interface.argtypes = [vp, i32, p64, vp, i64, i64, i64, dbl, vp]
    lib.b200cs_mappm_workspace_bytes.argtypes = [i64, i64]
    lib.b200cs_mappm_workspace_bytes.restype = i64
    lib.b200cs_mappm.argtypes = [vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, i64, i32, i32, i32, i32,
                                 vp, i64, vp]
    lib.b200cs_mappm_static.argtypes = [vp, i64, i64, vp, i64, i64, vp, i64, i64, vp, i64, i64, i64, i32, i32, i32, i32, vp]
    lib.b200cs_mappm_fast.argtypes = [vp, p64, vp, p64, vp, p64, vp, p64, i64, i64, i32, i32, i32, i32, vp]
    lib.b200cs_pressure_coarsen.argtypes = [i32, pvp, i32, p64, vp, i32, p64, vp, i32, p64, pvp, vp, i64, i32,
                                            i64, i64, i32, i32, dbl, i32, i32, i32, i32, vp]
    lib.b200cs_pressure_coarsen_gather.argtypes = [i32, pvp, i32, p64, vp, i32, p64, vp, i32, p64, pvp, vp, i64, i32,
                                                   i64, i64, i32, i32, dbl, i32, i32, i32, i32, i32, pvp, vp, i64, vp]
    lib.b200cs_pressure_coarsen_gather_workspace_bytes.argtypes = [i64, i64]
    lib.b200cs_pressure_coarsen_gather_workspace_bytes.restype = i64
    lib.b200cs_pressure_coarsen_wide.argtypes = [i32, pvp, p64, vp, p64, vp, p64, pvp, vp, i64, i32, i64, i64, i32, dbl,
                                                 i32, i32, i32, vp, i64, vp]
    lib.b200cs_pressure_coarsen_wide_workspace_bytes.argtypes = [i64, i32, i64, i64, i32]
    lib.b200cs_pressure_coarsen_wide_workspace_bytes.restype = i64
    lib.b200cs_mask_weights.argtypes = [vp, i32, vp, vp, i32, vp, i64, i64, i64, i32, vp]
    lib.b200cs_surface_pressure.argtypes = [vp, i32, p64, vp, i64, i32, i64, dbl, vp]
    lib.b200cs_blending_weights.argtypes = [vp, vp, vp, i32, i64, i32, i64, dbl, dbl, vp]
    lib.b200cs_blend.argtypes = [vp, vp, vp, vp, i32, i64, vp]
    lib.b200cs_elementwise.argtypes = [i32, vp, vp, vp, vp, i32, i64, dbl, vp]
    lib.b200cs_compare_scalar.argtypes = [i32, vp, i32, dbl, dbl, dbl, vp, vp, i64, vp]
    lib.b200cs_dominant_mask.argtypes = [vp, vp, i32, i64, i64, i64, i32, dbl, dbl, vp, vp, vp]
    lib.b200cs_pressure_at_midpoint_log.argtypes = [vp, vp, i32, i64, i32, i64, dbl, vp]
    lib.b200cs_exposed_area_weights.argtypes = [vp, vp, i32, vp, i32, vp, i64, i32, i64, i64, i32, vp]
    lib.b200cs_host_copy.argtypes = [vp, vp, i64, i32]
    lib.b200cs_hydrostatic_balance.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i64, i32, i64, dbl, dbl, dbl, dbl, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name not in ("b200cs_last_error", "b200cs_mappm_workspace_bytes", "b200cs_pressure_coarsen_gather_workspace_bytes",
                        "b200cs_pressure_coarsen_wide_workspace_bytes"):
            fn.restype = i32
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().b200cs_last_error().decode("utf-8", "replace")
        raise B200csError(f"libb200cs {what} failed with status {rc}: {msg}")


def i64arr(values):
    return (ctypes.c_int64 * len(values))(*[int(v) for v in values])


def ptrarr(ptrs):
    return (ctypes.c_void_p * len(ptrs))(*[ctypes.c_void_p(int(p)) for p in ptrs])
