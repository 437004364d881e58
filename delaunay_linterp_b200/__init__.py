"""delaunay_linterp_b200 -- B200-native batched evaluation of Delaunay piecewise-linear interpolants.

Drop-in for the interp hot path of rncarpio/delaunay_linterp (locate -> barycentric -> gather),
behind a C ABI (include/dli.h).  The CUDA library is required; there is no CPU fallback.
"""
from ._capi import (DliError, LIB_PATH, STATUS_NO_MESH, STATUS_OK, STATUS_OUTSIDE_NAN, STATUS_OUTSIDE_PROJECTED,
                    STATUS_STEP_LIMIT, STATUS_ON_EDGE, STATUS_ON_VERTEX, lib)
from .mesh import Mesh, debug_orient
from .interp2 import DelaunayInterp2
from .interpd import Delaunay_incremental_interp_d, get_lifted_point

__all__ = ["Mesh", "DelaunayInterp2", "Delaunay_incremental_interp_d", "get_lifted_point", "debug_orient", "DliError", "lib", "LIB_PATH", "STATUS_OK", "STATUS_OUTSIDE_PROJECTED",
           "STATUS_OUTSIDE_NAN", "STATUS_STEP_LIMIT", "STATUS_NO_MESH", "STATUS_ON_EDGE", "STATUS_ON_VERTEX"]
