"""fv3net_b200 -- B200-native (sm_100a) cubed-sphere coarse-graining: a drop-in for the
vcm.cubedsphere block-average / mappm vertical-remap hot path of VulcanClimateModeling/fv3net.

Layers:
  fv3net_b200.core      array-level operators on device buffers (torch CUDA tensors or numpy)
  fv3net_b200.restarts  restart coarse-graining recipes (fv_core / fv_tracer / fv_srf_wnd)
  fv3net_b200.cubedsphere  the reference's function names and labelled-array semantics
  libb200cs.so          hand-written CUDA kernels behind the C ABI of include/b200cs.h
"""
from . import _lib  # noqa: F401
from .core import (  # noqa: F401
    EXACT,
    FAST,
    block_reduce,
    block_upsample,
    block_upsample_like,
    edge_weighted_block_average,
    mappm,
    mask_weights,
    pressure_at_interface,
    pressure_coarsen,
    regrid_vertical,
    stagger,
    weighted_block_average,
)

__version__ = "0.1.0"
