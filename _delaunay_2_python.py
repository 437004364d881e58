"""Module name of the reference's Boost.Python binding (src/delaunay_2_python.cpp:200: BOOST_PYTHON_MODULE(_delaunay_2_python)),
so that `import _delaunay_2_python as _delaunay2` (delaunay_2.py:22, delaunay_2_test.py) resolves to the B200 engine.
Exposes DelaunayInterp2 with the reference's methods (:222-231): insert, interp, get_error_queue,
get_largest_error_tuple, insert_largest_error_point, get_line_segments, get_all_vertices, gradient."""
from delaunay_linterp_b200.interp2 import DelaunayInterp2

__all__ = ["DelaunayInterp2"]
