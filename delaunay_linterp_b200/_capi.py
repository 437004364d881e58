"""ctypes binding of the C ABI in include/dli.h (libdli.so).

There is no CPU fallback: if the library is missing, or no CUDA device is present when a mesh is
created, an error is raised.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DLI_LIB") or os.path.join(_HERE, "libdli.so")  # DLI_LIB: developer override (kernel experiments)

DLI_MAX_DIM = 4
DLI_MESH_HEADER_WORDS = 16
DLI_NOALLOC_MAX_QUERIES = 1 << 20

STATUS_OK, STATUS_OUTSIDE_PROJECTED, STATUS_OUTSIDE_NAN, STATUS_STEP_LIMIT, STATUS_NO_MESH = 0, 1, 2, 3, 4
STATUS_ON_EDGE, STATUS_ON_VERTEX = 5, 6


class DliError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("dli error %d: %s" % (code, msg))
        self.code = code


class Options(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32),
        ("device", C.c_int32),
        ("hint_cells_per_simplex", C.c_float),
        ("max_walk_steps", C.c_int32),
        ("reserved0", C.c_int32),
        ("epilogue", C.c_int32),
        ("threads_per_block", C.c_int32),
        ("queries_per_warp", C.c_int32),
        ("packed_faces", C.c_int32),
        ("prefetch_ring", C.c_int32),
        ("outside_nan", C.c_int32),
        ("start_records", C.c_int32),
        ("reserved", C.c_int32 * 4),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("queries", C.c_int64),
        ("walk_steps", C.c_int64),
        ("max_walk_steps", C.c_int64),
        ("exact_fallbacks", C.c_int64),
        ("outside", C.c_int64),
        ("kernel_launches", C.c_int64),
        ("fast_kernel_launches", C.c_int64),
        ("fast_kernel_ms", C.c_double),
        ("full_kernel_ms", C.c_double),
        ("straggler_queries", C.c_int64),
        ("slow_simplices", C.c_int64),
    ]


# every symbol include/dli.h declares: (name, restype, argtypes)
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_bp = C.POINTER(C.c_uint8)
_vp = C.c_void_p
SYMBOLS = [
    ("dli_options_init", None, [C.POINTER(Options)]),
    ("dli_mesh_create", C.c_int, [C.c_int, C.c_int64, _vp, _vp, C.c_int64, _vp, _vp, C.POINTER(Options), C.POINTER(_vp)]),
    ("dli_mesh_set_values", C.c_int, [_vp, _vp]),
    ("dli_mesh_update", C.c_int, [_vp, C.c_int64, _vp, _vp, C.c_int64, _vp, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp]),
    ("dli_mesh_destroy", None, [_vp]),
    ("dli_mesh_dim", C.c_int, [_vp]),
    ("dli_mesh_num_vertices", C.c_int64, [_vp]),
    ("dli_mesh_num_simplices", C.c_int64, [_vp]),
    ("dli_mesh_device_bytes", C.c_int64, [_vp]),
    ("dli_mesh_get_simplices", C.c_int, [_vp, _vp, _vp]),
    ("dli_interp", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    ("dli_interp_noalloc", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    ("dli_locate", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp]),
    ("dli_gradient", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    ("dli_simplex_error_count", C.c_int64, [_vp]),
    ("dli_simplex_error_prepare", C.c_int, [_vp, _vp, _vp, _vp]),
    ("dli_simplex_error_finish", C.c_int, [_vp, C.c_int64, _vp, _vp]),
    ("dli_mesh_num_edges", C.c_int64, [_vp]),
    ("dli_mesh_line_segments", C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    ("dli_mesh_header", C.c_int, [_vp, C.POINTER(C.c_int64)]),
    ("dli_mesh_create_from_header", C.c_int, [C.POINTER(C.c_int64), C.POINTER(Options), C.POINTER(_vp)]),
    ("dli_mesh_num_buffers", C.c_int, [_vp]),
    ("dli_mesh_buffer", C.c_int, [_vp, C.c_int, C.POINTER(_vp), C.POINTER(C.c_int64)]),
    ("dli_mesh_buffer_read", C.c_int, [_vp, C.c_int, _vp]),
    ("dli_mesh_buffer_write", C.c_int, [_vp, C.c_int, _vp]),
    ("dli_mesh_get_grid_params", C.c_int, [_vp, _vp]),
    ("dli_mesh_set_grid_params", C.c_int, [_vp, _vp]),
    ("dli_mesh_replicate", C.c_int, [C.POINTER(_vp), _vp, C.c_int]),
    ("dli_mesh_save", C.c_int, [_vp, C.c_char_p]),
    ("dli_mesh_load", C.c_int, [C.c_char_p, C.POINTER(Options), C.POINTER(_vp)]),
    ("dli_mesh_collect_stats", C.c_int, [_vp, C.c_int]),
    ("dli_mesh_get_stats", C.c_int, [_vp, C.POINTER(Stats)]),
    ("dli_debug_orient", C.c_int, [C.c_int, C.c_int64, _vp, _vp, _vp]),
    ("dli_debug_orient_host", C.c_int, [C.c_int, C.c_int64, _vp, _vp]),
    ("dli_last_error", C.c_char_p, []),
    ("dli_device_count", C.c_int, []),
]

DLI_FN = C.CFUNCTYPE(C.c_double, C.c_int, C.POINTER(C.c_double), C.c_void_p)
DLI_FN_BATCH = C.CFUNCTYPE(None, C.c_int, C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)

# include/dli_host.h
HOST_SYMBOLS = [
    ("dli_di2_create", _vp, [DLI_FN, _vp]),
    ("dli_di2_destroy", None, [_vp]),
    ("dli_di2_insert", C.c_int32, [_vp, C.c_double, C.c_double, C.c_double]),
    ("dli_di2_insert_many", C.c_int, [_vp, C.c_int64, _vp, _vp, C.c_int, C.c_int]),
    ("dli_di2_insert_largest_error_point", C.c_int, [_vp, _dp, _dp, _dp]),
    ("dli_di2_get_largest_error_tuple", C.c_int, [_vp, _dp, _dp, _dp]),
    ("dli_di2_refine", C.c_int64, [_vp, C.c_int64, C.c_double, DLI_FN_BATCH, _vp, _vp, _vp, _vp]),
    ("dli_di2_error_queue_size", C.c_int64, [_vp]),
    ("dli_di2_get_error_queue", C.c_int, [_vp, _vp, _vp, _vp]),
    ("dli_di2_num_vertices", C.c_int64, [_vp]),
    ("dli_di2_num_faces", C.c_int64, [_vp]),
    ("dli_di2_dimension", C.c_int, [_vp]),
    ("dli_di2_export", C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    ("dli_di2_interp", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    ("dli_di2_gradient", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    ("dli_di2_sync_counts", None, [_vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ("dli_di2_device_mesh", _vp, [_vp]),
    ("dli_di2_set_options", C.c_int, [_vp, C.POINTER(Options)]),
    ("dli_di2_get_all_vertices", C.c_int64, [_vp, _vp]),
    ("dli_di2_get_line_segments", C.c_int64, [_vp, _vp]),
    ("dli_di2_line_segments", C.c_int64, [_vp, _vp, _vp, _vp, _vp]),
    ("dli_dd_create", _vp, [C.c_int, DLI_FN, _vp]),
    ("dli_dd_destroy", None, [_vp]),
    ("dli_dd_dimension", C.c_int, [_vp]),
    ("dli_dd_current_dimension", C.c_int, [_vp]),
    ("dli_dd_num_vertices", C.c_int64, [_vp]),
    ("dli_dd_num_simplices", C.c_int64, [_vp]),
    ("dli_dd_insert", C.c_int32, [_vp, _vp, C.c_double]),
    ("dli_dd_insert_many", C.c_int, [_vp, C.c_int64, _vp, _vp, C.c_int, C.c_int]),
    ("dli_dd_insert_largest_error_point", C.c_int, [_vp, _dp, _dp, _vp]),
    ("dli_dd_get_largest_error_tuple", C.c_int, [_vp, _dp, _dp, _vp]),
    ("dli_dd_error_queue_size", C.c_int64, [_vp]),
    ("dli_dd_get_error_queue", C.c_int64, [_vp, C.c_int64, _vp, _vp, _vp]),
    ("dli_dd_export", C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    ("dli_dd_interp", C.c_int, [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    ("dli_dd_device_mesh", _vp, [_vp]),
    ("dli_dd_refine", C.c_int64, [_vp, C.c_int64, C.c_double, DLI_FN_BATCH, _vp, _vp, _vp, _vp]),
    ("dli_dd_sync_counts", None, [_vp, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    ("dli_dd_set_options", C.c_int, [_vp, C.POINTER(Options)]),
    ("dli_dd_get_all_vertices", C.c_int64, [_vp, _vp]),
    ("dli_dd_get_line_segments", C.c_int64, [_vp, _vp, _vp, _vp]),
    ("dli_dd_validate", C.c_int, [_vp, C.c_int]),
]

_lib = None


def lib():
    """Loads libdli.so (building it first if the sources are newer and nvcc is available)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "delaunay_linterp_b200: %s is missing. Build it with `python -m delaunay_linterp_b200.build` "
            "(needs nvcc); there is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS + HOST_SYMBOLS:
        fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc):
    if rc < 0:
        raise DliError(rc, lib().dli_last_error().decode("utf-8", "replace"))
    return rc
