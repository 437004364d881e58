"""Delaunay_incremental_interp_d: Python mirror of the reference's `_delaunay_d_python.Delaunay_incremental_interp_d`
(src/delaunay_d_python.cpp:364-373; C++ class src/delaunay_d_interp.h:753-1142) for d = 2, 3, 4.

Same method names and conventions (points are lists of floats, ErrorTuple -> (err, f, [x...]), interp outside the
hull returns NaN, :1005-1008), plus batched interp(ndarray (n, d)) -> ndarray (n,) on the GPU.

Host side: the reference builds on the patched CGAL Delaunay_d (modified_cgal/Delaunay_d.h), which is unavailable;
here the d-dimensional triangulation is maintained by Qhull in incremental mode (scipy.spatial.Delaunay(incremental=True)),
which stays on the host like the reference's CGAL.  Per-insertion error tuples (a handful of simplices) are computed on
the host with the closed form of include/dli_error_math.h; rebuild_error_queue() evaluates the whole mesh on the GPU.
"""
import heapq
import itertools

import numpy as np

from .mesh import Mesh, _is_torch_tensor


def _simplex_error(w, phi):
    """closed form of compute_simplex_error (src/delaunay_d_interp.h:1078-1093): w (np, d+1) lifted vertices."""
    fbar = w[:, -1].mean()
    u = (w[1:] - w[0]).T                       # (d+1, np-1)
    q, _ = np.linalg.qr(u)
    ez = np.zeros(w.shape[1])
    ez[-1] = 1.0
    r = ez - q @ (q.T @ ez)
    r = r - q @ (q.T @ r)
    return float((phi - fbar) ** 2 * (r @ r))


class Delaunay_incremental_interp_d:
    def __init__(self, n_dims, fn=None, **options):
        if not 2 <= n_dims <= 4:
            raise ValueError("n_dims must be 2, 3 or 4")
        self.N = int(n_dims)
        self.fn = fn
        self._options = options
        self._pts = []          # inserted points (lists)
        self._f = []
        self._index = {}        # point tuple -> vertex id
        self._tri = None        # scipy Delaunay once d+1 affinely independent points exist
        self._incremental = False
        self._mesh = None       # device copy, rebuilt lazily
        self._keys = {}         # frozenset(vertex ids) of current simplices -> live entry count holder
        self._heap = []         # (-err, seq, f, x tuple, simplex key)
        self._seq = 0

    def get_dimension(self):
        return self.N

    # ---- construction (host)
    def _point(self, x):
        if len(x) != self.N:
            raise TypeError("a point is a list of %d floats" % self.N)
        return tuple(float(v) for v in x)

    def insert(self, x, f):
        p = self._point(x)
        self._mesh = None
        if p in self._index:                      # existing point: value overwritten
            self._f[self._index[p]] = float(f)
            return
        self._index[p] = len(self._pts)
        self._pts.append(p)
        self._f.append(float(f))
        if self._tri is not None and self._incremental:
            self._tri.add_points(np.array([p]))
        elif len(self._pts) >= self.N + 1:
            self._try_start()
        if self._tri is not None and self.fn is not None:
            self._update_queue()

    def _try_start(self):
        """Qhull's incremental mode cannot start from cospherical / too few sites; until it can, the (small)
        triangulation is rebuilt from scratch on every insertion (default options, which include Qz)."""
        from scipy.spatial import Delaunay
        from scipy.spatial import QhullError
        pts = np.array(self._pts)
        try:
            self._tri = Delaunay(pts, incremental=True)
            self._incremental = True
            return
        except QhullError:
            pass
        try:
            self._tri = Delaunay(pts)
            self._incremental = False
        except QhullError:
            self._tri = None                      # fewer than d+1 affinely independent points so far

    def _current(self):
        sv = self._tri.simplices.astype(np.int32)
        sn = self._tri.neighbors.astype(np.int32)
        return sv, sn

    def _items_of(self, simplex, neighbors):
        """vertex subsets that get an error tuple for this simplex: itself + every subset (size >= 2) of each hull facet
        (src/delaunay_d_interp.h:810-848); duplicates across simplices are kept, as in the reference."""
        k = self.N + 1
        items = [tuple(int(v) for v in simplex)]
        for j in range(k):
            if neighbors[j] < 0:
                facet = [int(simplex[i]) for i in range(k) if i != j]
                for size in range(2, len(facet) + 1):
                    items.extend(itertools.combinations(facet, size))
        return items

    def _push_simplex(self, key, simplex, neighbors):
        pts = np.array(self._pts)
        fv = np.array(self._f)
        n = 0
        for verts in self._items_of(simplex, neighbors):
            v = list(verts)
            c = pts[v].sum(axis=0) / len(v)
            phi = float(self.fn([float(t) for t in c]))
            w = np.c_[pts[v], fv[v]]
            heapq.heappush(self._heap, (-_simplex_error(w, phi), self._seq, phi, tuple(c.tolist()), key))
            self._seq += 1
            n += 1
        self._keys[key] = n

    def _update_queue(self):
        sv, sn = self._current()
        new = {}
        for s in range(len(sv)):
            new[frozenset(int(v) for v in sv[s])] = s
        for key in [k for k in self._keys if k not in new]:
            del self._keys[key]                   # entries become stale (lazy deletion)
        for key, s in new.items():
            if key not in self._keys:
                self._push_simplex(key, sv[s], sn[s])

    def rebuild_error_queue(self):
        """recompute every tuple of the current triangulation in one batch on the GPU (dli_simplex_error_*)."""
        assert self.fn is not None and self._tri is not None
        m = self._device_mesh()
        cent, item_simplex, item_verts = m.simplex_error_items()
        phi = np.array([float(self.fn(c.tolist())) for c in cent])
        err = m.simplex_error(phi)
        csv, _ = m.simplices()
        self._heap, self._keys = [], {}
        for i in range(len(err)):
            key = frozenset(int(v) for v in csv[item_simplex[i]])
            self._keys[key] = self._keys.get(key, 0) + 1
            self._heap.append((-float(err[i]), self._seq, float(phi[i]), tuple(cent[i].tolist()), key))
            self._seq += 1
        heapq.heapify(self._heap)

    def _prune(self):
        while self._heap and self._heap[0][4] not in self._keys:
            heapq.heappop(self._heap)

    def get_largest_error_tuple(self):
        self._prune()
        assert self.fn is not None and self._heap, "empty queue or no callback"
        e = self._heap[0]
        return (-e[0], e[2], list(e[3]))

    def insert_largest_error_point(self):
        t = self.get_largest_error_tuple()
        self.insert(t[2], t[1])
        return t

    def get_error_queue(self):
        live = [e for e in self._heap if e[4] in self._keys]
        live.sort()
        return [(-e[0], e[2], list(e[3])) for e in live]

    # ---- the hot path (GPU)
    def _device_mesh(self):
        if self._mesh is None:
            if self._tri is None:
                sv = np.zeros((0, self.N + 1), np.int32)
                pts = np.array(self._pts, dtype=np.float64).reshape(-1, self.N)
                self._mesh = Mesh(pts, np.array(self._f, dtype=np.float64), sv, None, **self._options)
            else:
                sv, sn = self._current()
                self._mesh = Mesh(np.array(self._pts), np.array(self._f), sv, sn, **self._options)
        return self._mesh

    def interp(self, x, return_status=False):
        """interp([x...]) -> float (NaN outside the hull) or interp(ndarray (n, d) / torch CUDA tensor) -> array."""
        m = self._device_mesh()
        if _is_torch_tensor(x):
            return m.interp(x, return_status=return_status)
        scalar = isinstance(x, (list, tuple)) and len(x) == self.N and not isinstance(x[0], (list, tuple, np.ndarray))
        xs = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, self.N)
        ys, st = m.interp(xs, return_status=True)
        if scalar:
            return float(ys[0])
        return (ys, st) if return_status else ys

    # ---- plot data
    def get_all_vertices(self):
        return [list(p) + [f] for p, f in zip(self._pts, self._f)]

    def get_line_segments(self):
        """(all_segments, min_coords, max_coords) as the reference returns it (src/delaunay_d_interp.h:1115-1142): the
        segments of a simplex are its consecutive vertex pairs (i, i+1 mod N+1) (:1103-1112), every pair once, flag 1."""
        n1 = self.N + 1
        lo, hi = [float(np.finfo(np.float64).max)] * n1, [-float(np.finfo(np.float64).max)] * n1
        if self._tri is None:
            return [], lo, hi
        sv, _ = self._current()
        edges = set()
        for s in sv:
            for i in range(n1):
                a, b = int(s[i]), int(s[(i + 1) % n1])
                edges.add((min(a, b), max(a, b)))
        out = []
        for a, b in sorted(edges):
            pa, pb = list(self._pts[a]) + [self._f[a]], list(self._pts[b]) + [self._f[b]]
            out.append((pa, pb, 1))
            for p in (pa, pb):
                lo = [min(u, v) for u, v in zip(lo, p)]
                hi = [max(u, v) for u, v in zip(hi, p)]
        return out, lo, hi

    def export(self):
        sv, sn = self._current() if self._tri is not None else (np.zeros((0, self.N + 1), np.int32),) * 2
        return np.array(self._pts), np.array(self._f), sv, sn


def get_lifted_point(x):
    """free function of the reference module (src/delaunay_d_interp.h:1145-1149): (x, |x|^2)."""
    x = [float(v) for v in x]
    return x + [sum(v * v for v in x)]
