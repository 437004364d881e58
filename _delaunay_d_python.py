"""Module name of the reference's Boost.Python binding (src/delaunay_d_python.cpp:247: BOOST_PYTHON_MODULE(_delaunay_d_python)),
so that `import _delaunay_d_python as _delaunayd` (delaunay_d_test.py:4) resolves to the B200 engine.
Exposes Delaunay_incremental_interp_d(n_dims, fn) with the reference's members (:364-373): read-only triang / fn,
insert, interp, get_error_queue, get_largest_error_tuple, insert_largest_error_point, get_line_segments,
get_all_vertices; and get_lifted_point (:375).  The commented-out interp_vec of the reference (:383) is the batched hook:
interp_vec(xs) == interp(ndarray (n, d))."""
from delaunay_linterp_b200.interpd import Delaunay_incremental_interp_d, get_lifted_point

if not hasattr(Delaunay_incremental_interp_d, "interp_vec"):
    Delaunay_incremental_interp_d.interp_vec = lambda self, xs: self.interp(xs)

__all__ = ["Delaunay_incremental_interp_d", "get_lifted_point"]
