"""Exact-rational twin of the oracle for tiny cases.  TEST INFRASTRUCTURE ONLY.

Pure Python over fractions.Fraction: exact orientation signs, brute-force closed-simplex
location and exact barycentric interpolation.  Used to check oracle/dli_oracle.c (filter +
expansion arithmetic, walk) on small adversarial inputs.  Follows the same reference lines as
dli_oracle.c (src/delaunay_2_interp.h:515-572, :223-239; src/delaunay_d_interp.h:988-1010).
"""
from fractions import Fraction as Fr


def det(m):
    n = len(m)
    if n == 1:
        return m[0][0]
    if n == 2:
        return m[0][0] * m[1][1] - m[0][1] * m[1][0]
    s = Fr(0)
    for r in range(n):
        if m[r][0] == 0:
            continue
        minor = [row[1:] for i, row in enumerate(m) if i != r]
        s += (-1) ** r * m[r][0] * det(minor)
    return s


def orient(pts):
    """sign of det[p 1]; pts = d+1 points of d coords (floats are converted exactly)."""
    d = len(pts) - 1
    m = [[Fr(float(c)) for c in p] + [Fr(1)] for p in pts]
    v = det(m)
    return (v > 0) - (v < 0)


def barycentric(simplex_pts, p):
    k = len(simplex_pts)
    whole = det([[Fr(float(c)) for c in q] + [Fr(1)] for q in simplex_pts])
    out = []
    for i in range(k):
        pts = [p if j == i else simplex_pts[j] for j in range(k)]
        out.append(det([[Fr(float(c)) for c in q] + [Fr(1)] for q in pts]) / whole)
    return out


def locate_all(coords, simplices, p):
    """ids of every simplex whose closed hull contains p (exact)."""
    hits = []
    for s, sv in enumerate(simplices):
        pts = [coords[v] for v in sv]
        b = barycentric(pts, p)
        if all(x >= 0 for x in b):
            hits.append(s)
    return hits


def interp_exact(coords, f, simplices, p):
    """(exact value as Fraction, list of containing simplices) or (None, []) outside."""
    hits = locate_all(coords, simplices, p)
    if not hits:
        return None, []
    sv = simplices[hits[0]]
    b = barycentric([coords[v] for v in sv], p)
    val = sum(bi * Fr(float(f[v])) for bi, v in zip(b, sv))
    return val, hits


def face_gradient_exact(coords, f, sv):
    """exact gradient of the linear interpolant on one triangle (src/delaunay_2_interp.h:242-263)."""
    (x0, y0), (x1, y1), (x2, y2) = [[Fr(float(c)) for c in coords[v]] for v in sv]
    f0, f1, f2 = [Fr(float(f[v])) for v in sv]
    a, b, c, d = x1 - x0, x2 - x0, y1 - y0, y2 - y0
    dt = a * d - b * c
    return ((f1 - f0) * d - (f2 - f0) * c) / dt, (a * (f2 - f0) - b * (f1 - f0)) / dt


def gradient_exact(coords, f, simplices, p):
    """Reference semantics of gradient() (src/delaunay_2_interp.h:608-642) in exact arithmetic: the mean of the
    gradients of every finite face whose closed hull contains p (1 face inside, 2 on an interior edge, the whole
    fan on a vertex); (None, []) outside the hull."""
    hits = locate_all(coords, simplices, p)
    if not hits:
        return None, []
    gs = [face_gradient_exact(coords, f, simplices[s]) for s in hits]
    n = len(gs)
    return (sum(g[0] for g in gs) / n, sum(g[1] for g in gs) / n), hits
