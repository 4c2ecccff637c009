"""lsc_internship_b200 -- B200-native (sm_100a CUDA behind a C ABI) implementation of the
2-D MUSCL-Hancock + HLLC + GLM ideal-MHD step of JosephineMellor/LSC_internship
(MHD_codes/MUSCL_NEW.cpp) and of its 2-D Euler SLIC sweep (Euler_codes/SLIC_euler_2d.cpp)."""
from ._capi import LscError, MhdConfig, lib, LIB_PATH  # noqa: F401
from .hllc2d import HllcMhd2d  # noqa: F401
from .euler import SlicEuler2d, SlicEuler2dGroup, euler_kat_cfl_bound  # noqa: F401
from .mhd import MusclNew, MusclNewGroup, kat_hllc, kat_fluxdef, kat_cfl_bound, slab_range, neighbours  # noqa: F401

__all__ = ["HllcMhd2d", "SlicEuler2d", "SlicEuler2dGroup", "MusclNew", "MusclNewGroup", "slab_range", "neighbours", "LscError", "MhdConfig", "lib", "LIB_PATH", "kat_hllc", "kat_fluxdef", "kat_cfl_bound", "euler_kat_cfl_bound"]
