"""Shared mesh/query generators for the tests (the BASELINE.json configs, scaled down)."""
import numpy as np


def cobb_douglas(x):
    """f = prod x_i^alpha_i, alpha = (0.3, 0.6, 0.2, 0.1)[:d]  (src/delaunay_2_test.cpp:32-50)."""
    x = np.asarray(x, dtype=np.float64)
    a = np.array([0.3, 0.6, 0.2, 0.1, 0.05, 0.05])[: x.shape[-1]]
    return np.prod(np.power(x, a), axis=-1)


def linspace_ref(first, last, n):
    """the reference's linspace: first + i*step (README.md:37-42)."""
    step = (last - first) / (n - 1)
    return np.array([first + i * step for i in range(n)])


def grid_mesh(lo=0.0, hi=5.0, n=10):
    """config 1: n x n grid, inserted i-outer/j-inner (README.md:55-60); every cell split by the
    diagonal (x1,y0)-(x0,y1), the tie-break CGAL's perturbed in-circle test produces
    (SURVEY.md appendix B.2 / C.1).  Returns coords, f, triangles (ccw)."""
    g = linspace_ref(lo, hi, n)
    coords = np.array([[g[i], g[j]] for i in range(n) for j in range(n)])
    vid = lambda i, j: i * n + j
    tris = []
    for i in range(n - 1):
        for j in range(n - 1):
            a, b, c, d = vid(i, j), vid(i + 1, j), vid(i, j + 1), vid(i + 1, j + 1)
            tris.append([a, b, c])   # A(x0,y0) B(x1,y0) C(x0,y1)
            tris.append([b, d, c])   # B D(x1,y1) C      -> diagonal B-C
    return coords, cobb_douglas(coords), np.array(tris, dtype=np.int32)


def grid_queries(lo=0.0, hi=5.0, n=20):
    q = linspace_ref(lo, hi, n)
    return np.array([[q[i], q[j]] for i in range(n) for j in range(n)])


def scipy_mesh(points):
    """Delaunay triangulation of `points` through Qhull: (simplices int32, neighbors int32).
    scipy's convention matches the C ABI: neighbour k is opposite vertex k, -1 = hull."""
    from scipy.spatial import Delaunay
    tri = Delaunay(points)
    return tri.simplices.astype(np.int32), tri.neighbors.astype(np.int32)


def random_mesh(dim, nv, seed):
    rng = np.random.default_rng(seed)
    pts = rng.random((nv, dim))
    sv, sn = scipy_mesh(pts)
    return pts, cobb_douglas(pts), sv, sn


def lattice_predicate_inputs(d, seed=4):
    """orientation-predicate inputs (n, d+1, d) on a regular grid i/(n-1) (full 53-bit mantissas): generic, all points
    on an axis-aligned facet hyperplane, on a Kuhn diagonal hyperplane x_1 - x_0 = c (exact zeros), one coordinate far
    below the others (outside the fixed-point bit budget), and a negated / shifted copy."""
    rng = np.random.default_rng(seed)
    n = 47
    lat = rng.integers(0, n, (400, d + 1, d)) / float(n - 1)
    lat[:100, :, 0] = lat[:100, :1, 0]
    lat[100:200, :, 1] = lat[100:200, :, 0] + (lat[100:200, :1, 1] - lat[100:200, :1, 0])
    tiny = lat.copy()[:100]
    tiny[:, 0, 0] = 1e-30
    neg = -lat[:100] * 3.0 + 1.0
    return np.concatenate([lat, tiny, neg])


def regular_grid_mesh_2d(n, lo=0.0, hi=5.0):
    """config 1 scaled up, vectorised: n x n grid on [lo, hi]^2, every cell split by the (x1,y0)-(x0,y1) diagonal."""
    g = np.linspace(lo, hi, n)
    X, Y = np.meshgrid(g, g, indexing="ij")
    coords = np.stack([X.ravel(), Y.ravel()], 1)
    I, J = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    a = (I * n + J).ravel(); b = a + n; c = a + 1; d = a + n + 1
    tri = np.concatenate([np.stack([a, b, c], 1), np.stack([b, d, c], 1)]).astype(np.int32)
    return coords, cobb_douglas(coords + 0.5), tri


def kuhn_grid_mesh(dim, n):
    """regular grid on [0,1]^dim, every cell split into dim! Kuhn simplices (vectorised)."""
    import itertools
    g = np.arange(n, dtype=np.float64) / (n - 1)
    grid = np.stack(np.meshgrid(*([g] * dim), indexing="ij"), -1).reshape(-1, dim)
    cells = np.stack(np.meshgrid(*([np.arange(n - 1)] * dim), indexing="ij"), -1).reshape(-1, dim)
    simp = []
    for perm in itertools.permutations(range(dim)):
        cur = cells.copy()
        verts = [np.ravel_multi_index(cur.T, (n,) * dim)]
        for ax in perm:
            cur[:, ax] += 1
            verts.append(np.ravel_multi_index(cur.T, (n,) * dim))
        simp.append(np.stack(verts, 1))
    return grid, 1.0 + cobb_douglas(grid + 0.5), np.concatenate(simp).astype(np.int32)
