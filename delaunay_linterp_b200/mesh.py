"""Mesh: a flattened simplex mesh resident on one B200, and the batched interp entry points.

Thin Python layer over the C ABI (include/dli.h).  numpy arrays take the host-pointer path
(chunked H2D / compute / D2H inside the library); torch CUDA tensors take the device-pointer
path (asynchronous on the current torch stream).
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import DLI_MAX_DIM, DLI_MESH_HEADER_WORDS, Options, Stats, check, lib


def _is_torch_tensor(x):
    return hasattr(x, "data_ptr") and hasattr(x, "is_cuda")


def _make_options(device=-1, hint_cells_per_simplex=0.0, max_walk_steps=0, epilogue=0,
                  threads_per_block=0, queries_per_warp=0, packed_faces=0, prefetch_ring=0, outside_nan=0, start_records=0):
    o = Options()
    lib().dli_options_init(C.byref(o))
    o.device = int(device)
    o.hint_cells_per_simplex = float(hint_cells_per_simplex)
    o.max_walk_steps = int(max_walk_steps)
    o.epilogue = {"auto": 0, "closed_form": 1, "reference": 2}.get(epilogue, epilogue)
    o.threads_per_block = int(threads_per_block)
    o.queries_per_warp = int(queries_per_warp)
    o.packed_faces = int(packed_faces)
    o.prefetch_ring = int(prefetch_ring)
    o.outside_nan = int(outside_nan)
    o.start_records = int(start_records)
    return o


class Mesh:
    """Device-resident structure-of-arrays copy of a Delaunay simplex mesh (dim 2, 3 or 4).

    coords (nv, dim) float64; f (nv,) float64; simplices (ns, dim+1) int32;
    neighbors (ns, dim+1) int32 or None (k-th neighbour is opposite vertex k, -1 = hull; this is
    also scipy.spatial.Delaunay's convention).
    """

    def __init__(self, coords, f, simplices, neighbors=None, **options):
        coords = np.ascontiguousarray(coords, dtype=np.float64)
        f = np.ascontiguousarray(f, dtype=np.float64)
        if coords.ndim != 2:
            raise ValueError("coords must be (n_vertices, dim)")
        dim = coords.shape[1]
        simplices = np.ascontiguousarray(simplices, dtype=np.int32).reshape(-1, dim + 1)
        if f.shape != (coords.shape[0],):
            raise ValueError("f must be (n_vertices,)")
        nb_ptr = None
        if neighbors is not None:
            neighbors = np.ascontiguousarray(neighbors, dtype=np.int32)
            if neighbors.shape != simplices.shape:
                raise ValueError("neighbors must have the shape of simplices")
            nb_ptr = neighbors.ctypes.data
        self._h = C.c_void_p()
        opt = _make_options(**options)
        check(lib().dli_mesh_create(dim, coords.shape[0], coords.ctypes.data, f.ctypes.data, simplices.shape[0],
                                    simplices.ctypes.data if simplices.size else None, nb_ptr, C.byref(opt),
                                    C.byref(self._h)))
        self.dim = dim

    @classmethod
    def _from_handle(cls, handle, dim):
        self = cls.__new__(cls)
        self._h = handle
        self.dim = dim
        return self

    def close(self):
        if getattr(self, "_h", None):
            lib().dli_mesh_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- introspection
    @property
    def n_vertices(self):
        return lib().dli_mesh_num_vertices(self._h)

    @property
    def n_simplices(self):
        return lib().dli_mesh_num_simplices(self._h)

    @property
    def device_bytes(self):
        return lib().dli_mesh_device_bytes(self._h)

    def simplices(self):
        """canonical (positively oriented) simplices and neighbours as the engine stores them."""
        ns = self.n_simplices
        sv = np.empty((ns, self.dim + 1), np.int32)
        sn = np.empty((ns, self.dim + 1), np.int32)
        check(lib().dli_mesh_get_simplices(self._h, sv.ctypes.data, sn.ctypes.data))
        return sv, sn

    def set_values(self, f):
        f = np.ascontiguousarray(f, dtype=np.float64)
        if f.shape != (self.n_vertices,):
            raise ValueError("f must be (n_vertices,)")
        check(lib().dli_mesh_set_values(self._h, f.ctypes.data))

    def update(self, new_coords, new_f, n_simplices_total, dirty_ids, dirty_vertices, dirty_neighbors,
               value_ids=None, values=None):
        """Patch the device mesh in place after insertions (dli_mesh_update; the reference's dirty set is the star
        of the inserted vertex, src/delaunay_2_interp.h:434-448).  Returns False when the mesh cannot be patched
        (DLI_ERR_UNSUPPORTED) and has to be re-created; the mesh is unchanged in that case."""
        nc = np.ascontiguousarray(new_coords, dtype=np.float64).reshape(-1, self.dim)
        nf = np.ascontiguousarray(new_f, dtype=np.float64).reshape(-1)
        ids = np.ascontiguousarray(dirty_ids, dtype=np.int32).reshape(-1)
        sv = np.ascontiguousarray(dirty_vertices, dtype=np.int32).reshape(-1, self.dim + 1)
        sn = np.ascontiguousarray(dirty_neighbors, dtype=np.int32).reshape(-1, self.dim + 1)
        vi = np.ascontiguousarray(value_ids if value_ids is not None else [], dtype=np.int32).reshape(-1)
        vv = np.ascontiguousarray(values if values is not None else [], dtype=np.float64).reshape(-1)
        assert nc.shape[0] == nf.shape[0] and ids.shape[0] == sv.shape[0] == sn.shape[0] and vi.shape[0] == vv.shape[0]
        rc = lib().dli_mesh_update(self._h, self.n_vertices + nc.shape[0], nc.ctypes.data, nf.ctypes.data, vi.shape[0],
                                   vi.ctypes.data, vv.ctypes.data, int(n_simplices_total), ids.shape[0], ids.ctypes.data,
                                   sv.ctypes.data, sn.ctypes.data)
        if rc == -5:
            return False
        check(rc)
        return True

    # ---- the hot path
    def interp(self, xs, return_simplex=False, return_status=False, out=None, stream=None):
        """Batched interp(xs) -> ys.  xs: (n, dim) float64 numpy array or torch CUDA tensor."""
        if _is_torch_tensor(xs):
            return self._interp_torch(xs, return_simplex, return_status, out, stream, values=True)
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, self.dim)
        n = xs.shape[0]
        ys = out if out is not None else np.empty(n, np.float64)
        if ys.dtype != np.float64 or ys.shape != (n,) or not ys.flags.c_contiguous:
            raise ValueError("out must be a contiguous float64 array of shape (n,)")
        simplex = np.empty(n, np.int32) if return_simplex else None
        status = np.empty(n, np.uint8) if return_status else None
        check(lib().dli_interp(self._h, n, xs.ctypes.data, ys.ctypes.data,
                               simplex.ctypes.data if simplex is not None else None,
                               status.ctypes.data if status is not None else None, None))
        res = (ys,)
        if return_simplex:
            res += (simplex,)
        if return_status:
            res += (status,)
        return res[0] if len(res) == 1 else res

    def locate(self, xs, return_status=False, stream=None):
        if _is_torch_tensor(xs):
            return self._interp_torch(xs, True, return_status, None, stream, values=False)
        xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, self.dim)
        n = xs.shape[0]
        simplex = np.empty(n, np.int32)
        status = np.empty(n, np.uint8) if return_status else None
        check(lib().dli_locate(self._h, n, xs.ctypes.data, simplex.ctypes.data,
                               status.ctypes.data if status is not None else None, None))
        return (simplex, status) if return_status else simplex

    def gradient(self, xs, return_simplex=False, return_status=False, stream=None):
        """Batched gradient(xs) -> (n, 2) (2D meshes; src/delaunay_2_interp.h:608-642).  NaN outside the hull."""
        if self.dim != 2:
            raise ValueError("gradient: 2D meshes only (the reference's dD class has no gradient)")
        if _is_torch_tensor(xs):
            import torch
            if not xs.is_cuda or xs.dtype != torch.float64:
                raise ValueError("torch input must be a float64 CUDA tensor")
            xs = xs.contiguous().view(-1, 2)
            n = xs.shape[0]
            gs = torch.empty((n, 2), dtype=torch.float64, device=xs.device)
            simplex = torch.empty(n, dtype=torch.int32, device=xs.device) if return_simplex else None
            status = torch.empty(n, dtype=torch.uint8, device=xs.device) if return_status else None
            st = stream if stream is not None else torch.cuda.current_stream(xs.device).cuda_stream
            ptr = lambda t: t.data_ptr() if t is not None else None
            check(lib().dli_gradient(self._h, n, xs.data_ptr(), gs.data_ptr(), ptr(simplex), ptr(status), st))
        else:
            xs = np.ascontiguousarray(xs, dtype=np.float64).reshape(-1, 2)
            n = xs.shape[0]
            gs = np.empty((n, 2), np.float64)
            simplex = np.empty(n, np.int32) if return_simplex else None
            status = np.empty(n, np.uint8) if return_status else None
            ptr = lambda a: a.ctypes.data if a is not None else None
            check(lib().dli_gradient(self._h, n, xs.ctypes.data, gs.ctypes.data, ptr(simplex), ptr(status), None))
        res = tuple(t for t in (gs, simplex, status) if t is not None)
        return res[0] if len(res) == 1 else res

    def _interp_torch(self, xs, return_simplex, return_status, out, stream, values):
        import torch
        if not xs.is_cuda or xs.dtype != torch.float64:
            raise ValueError("torch input must be a float64 CUDA tensor")
        xs = xs.contiguous().view(-1, self.dim)
        n = xs.shape[0]
        ys = None
        if values:
            ys = out if out is not None else torch.empty(n, dtype=torch.float64, device=xs.device)
        simplex = torch.empty(n, dtype=torch.int32, device=xs.device) if return_simplex else None
        status = torch.empty(n, dtype=torch.uint8, device=xs.device) if return_status else None
        st = stream if stream is not None else torch.cuda.current_stream(xs.device).cuda_stream
        if values:
            check(lib().dli_interp(self._h, n, xs.data_ptr(), ys.data_ptr(),
                                   simplex.data_ptr() if simplex is not None else None,
                                   status.data_ptr() if status is not None else None, st))
        else:
            check(lib().dli_locate(self._h, n, xs.data_ptr(), simplex.data_ptr(),
                                   status.data_ptr() if status is not None else None, st))
        res = tuple(t for t in (ys, simplex if return_simplex else None, status) if t is not None)
        return res[0] if len(res) == 1 else res

    def interp_device_ptr(self, n, xs_ptr, ys_ptr, simplex_ptr=None, status_ptr=None, stream=None):
        """raw device-pointer call (bench / interop)."""
        check(lib().dli_interp(self._h, n, xs_ptr, ys_ptr, simplex_ptr, status_ptr, stream))

    # ---- per-simplex error evaluator (feeds the host adaptive-insertion queue)
    def simplex_error_items(self):
        """(centroids (n, dim), item_simplex (n,), item_vertices (n, dim+1) padded with -1)."""
        n = check(lib().dli_simplex_error_count(self._h))
        cent = np.empty((n, self.dim), np.float64)
        item_simplex = np.empty(n, np.int32)
        item_verts = np.empty((n, self.dim + 1), np.int32)
        check(lib().dli_simplex_error_prepare(self._h, cent.ctypes.data, item_simplex.ctypes.data,
                                              item_verts.ctypes.data))
        return cent, item_simplex, item_verts

    def simplex_error_centroids_device(self, out=None):
        """the centroids of every item as a float64 CUDA tensor (n, dim), written by the kernel itself (device-pointer form
        of dli_simplex_error_prepare: no host copy) -- for an f that is evaluated on the GPU."""
        import torch
        n = check(lib().dli_simplex_error_count(self._h))
        cent = out if out is not None else torch.empty((n, self.dim), dtype=torch.float64, device="cuda")
        if n:
            check(lib().dli_simplex_error_prepare(self._h, cent.data_ptr(), None, None))
        return cent

    def simplex_error(self, phi):
        """err for every item given phi[i] = f(centroid_i).  phi: numpy array, or a float64 CUDA tensor (the result is then
        a CUDA tensor too and nothing passes through the host)."""
        if _is_torch_tensor(phi):
            import torch
            if not phi.is_cuda or phi.dtype != torch.float64:
                raise ValueError("torch input must be a float64 CUDA tensor")
            phi = phi.contiguous()
            err = torch.empty_like(phi)
            check(lib().dli_simplex_error_finish(self._h, phi.shape[0], phi.data_ptr(), err.data_ptr()))
            return err
        phi = np.ascontiguousarray(phi, dtype=np.float64)
        err = np.empty_like(phi)
        check(lib().dli_simplex_error_finish(self._h, phi.shape[0], phi.ctypes.data, err.ctypes.data))
        return err

    # ---- plot data
    def line_segments(self):
        """2D: (segments (ne, 2, 3), concave (ne,), min (3,), max (3,)) -- get_line_segments + is_edge_concave
        (src/delaunay_2_interp.h:842-862, :766-803) for every finite edge, computed on the device."""
        ne = check(lib().dli_mesh_num_edges(self._h))
        seg = np.empty((ne, 2, 3), np.float64)
        cc = np.empty(ne, np.int32)
        mn, mx = np.empty(3), np.empty(3)
        check(lib().dli_mesh_line_segments(self._h, seg.ctypes.data, cc.ctypes.data, mn.ctypes.data, mx.ctypes.data))
        return seg, cc, mn, mx

    # ---- instrumentation
    def collect_stats(self, enable=True, timing=False):
        """counters add atomics to the kernels; timing records CUDA events around every interp launch."""
        check(lib().dli_mesh_collect_stats(self._h, (1 if enable else 0) | (2 if timing else 0)))

    def stats(self):
        s = Stats()
        check(lib().dli_mesh_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in Stats._fields_}

    # ---- replication across GPUs (one process per GPU; torch.distributed / NCCL moves the bytes)
    def header(self):
        h = (C.c_int64 * DLI_MESH_HEADER_WORDS)()
        check(lib().dli_mesh_header(self._h, h))
        gp = np.empty(3 * DLI_MAX_DIM, np.float64)
        check(lib().dli_mesh_get_grid_params(self._h, gp.ctypes.data))
        return np.array(list(h), dtype=np.int64), gp

    @classmethod
    def from_header(cls, header, grid_params, **options):
        h = (C.c_int64 * DLI_MESH_HEADER_WORDS)(*[int(v) for v in header])
        handle = C.c_void_p()
        opt = _make_options(**options)
        check(lib().dli_mesh_create_from_header(h, C.byref(opt), C.byref(handle)))
        gp = np.ascontiguousarray(grid_params, dtype=np.float64)
        check(lib().dli_mesh_set_grid_params(handle, gp.ctypes.data))
        return cls._from_handle(handle, int(header[1]))

    # ---- snapshot (flat file: header, grid parameters, device buffers)
    def save(self, path):
        check(lib().dli_mesh_save(self._h, str(path).encode()))

    @classmethod
    def load(cls, path, **options):
        """a replica of a saved mesh: interp / locate / gradient only (no host connectivity)."""
        handle = C.c_void_p()
        opt = _make_options(**options)
        check(lib().dli_mesh_load(str(path).encode(), C.byref(opt), C.byref(handle)))
        return cls._from_handle(handle, lib().dli_mesh_dim(handle))

    def buffers(self):
        """[(device_ptr, nbytes)] of the mesh's device buffers."""
        out = []
        for i in range(lib().dli_mesh_num_buffers(self._h)):
            p = C.c_void_p()
            b = C.c_int64()
            check(lib().dli_mesh_buffer(self._h, i, C.byref(p), C.byref(b)))
            out.append((p.value or 0, b.value))
        return out

    def buffer_read(self, index, dst_ptr):
        check(lib().dli_mesh_buffer_read(self._h, index, dst_ptr))

    def buffer_write(self, index, src_ptr):
        check(lib().dli_mesh_buffer_write(self._h, index, src_ptr))


def debug_orient(pts, on_device=True):
    """signs of the orientation determinants of pts (n, dim+1, dim) through the engine's predicate."""
    pts = np.ascontiguousarray(pts, dtype=np.float64)
    n, k, d = pts.shape
    assert k == d + 1
    sign = np.empty(n, np.int8)
    if on_device:
        used = np.empty(n, np.uint8)
        check(lib().dli_debug_orient(d, n, pts.ctypes.data, sign.ctypes.data, used.ctypes.data))
        return sign, used.astype(bool)
    check(lib().dli_debug_orient_host(d, n, pts.ctypes.data, sign.ctypes.data))
    return sign
