"""Drop-in for the reference's f2py extension package ``mappm`` (external/mappm/mappm/__init__.py:
``from .mappm import interpolate_2d, mappm``), so that ``import mappm`` at
external/vcm/vcm/cubedsphere/regridz.py:7-10 resolves to the B200 library.

``mappm.mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop)`` has the f2py signature of
external/mappm/mappm/mappm.f90:10 (numpy in, float32 ``[ncol, kn]`` numpy out) and runs
``b200cs_mappm`` (fv3net_b200/csrc/mappm.cu) through ``fv3net_b200.core.mappm``; there is no CPU
path -- it raises when the CUDA library or a device is missing.

``interpolate_2d`` (external/mappm/mappm/interpolate_2d.f90, used by vcm.interpolate only) is not on
the coarse-graining path this repository replaces: calling it raises.
"""
from fv3net_b200.core import mappm as _mappm

__all__ = ["mappm", "interpolate_2d"]


def mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop):
    """f2py: q2 = mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop); i1..i2 = 1..ncol as every reference
    caller passes them (regridz.py:329-334); ptop is unused by the Fortran."""
    return _mappm(pe1, q1, pe2, i1, i2, iv, kord, ptop)


def interpolate_2d(*args, **kwargs):
    raise NotImplementedError("mappm.interpolate_2d (vcm.interpolate) is outside the coarse-graining path of fv3net_b200")
