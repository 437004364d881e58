// dli_refmath.h -- the reference's per-query arithmetic in its own operation order, as plain functions usable on the
// host and on the device: the exact pass of the kernels (dli_device.cuh) and the host side of small batches
// (dli_hostmesh.cpp, dli_hostmesh_d.cpp) run the same code, so both return the same bits.
//
// Reference behaviour restated (paths relative to the reference tree):
//   2D epilogue src/delaunay_2_interp.h:195-239, 563-569 (Eigen dynamic determinant = partial-pivot LU)
//   2D outside-hull projection :544-545 (CGAL Cartesian Line_2 + projection)
//   dD epilogue src/delaunay_d_interp.h:601-650, 998-1004 (exact differences truncated, exact sum truncated)
//
// Must be compiled with FMA contraction off (see dli_exact.cuh).
#pragma once
#include <cmath>
#include "dli_exact.cuh"

namespace dli {

// Eigen partial-pivot LU determinant of [[a, b], [c, d]]
DLI_HDI double lu_det2(double a, double b, double c, double d) {
  double s = 1.0;
  if (fabs(c) > fabs(a)) {
    double t = a; a = c; c = t;
    t = b; b = d; d = t;
    s = -1.0;
  }
  double l = (a != 0.0) ? c / a : 0.0;
  double u = d - l * b;
  return s * (a * u);
}
// signed area: column j = P_{j+1} - P_0 (src/delaunay_2_interp.h:202-207), det / 2!
DLI_HDI double vol2(double ax, double ay, double bx, double by, double cx, double cy) {
  return lu_det2(bx - ax, cx - ax, by - ay, cy - ay) * 0.5;
}
// barycentric coordinates (:223-239) + weighted gather (:563-569), vertices A, B, C with values fa, fb, fc
DLI_HDI double epilogue2_ref_xy(double px, double py, double ax, double ay, double fa, double bx, double by, double fb,
                                double cx, double cy, double fc) {
  double whole = vol2(ax, ay, bx, by, cx, cy);
  double b0 = vol2(px, py, bx, by, cx, cy) / whole;
  double b1 = vol2(ax, ay, px, py, cx, cy) / whole;
  double b2 = vol2(ax, ay, bx, by, px, py) / whole;
  double sum = 0.0;
  sum = sum + b0 * fa;
  sum = sum + b1 * fb;
  sum = sum + b2 * fc;
  return sum;
}

// CGAL Cartesian Line_2(P,Q) + projection in plain doubles
DLI_HD inline void project_on_line(double Px, double Py, double Qx, double Qy, double &px, double &py) {
  double a, b, c;
  if (Py == Qy) {
    a = 0.0;
    if (Qx > Px) { b = 1.0; c = -Py; }
    else if (Qx == Px) { b = 0.0; c = 0.0; }
    else { b = -1.0; c = Py; }
  } else if (Qx == Px) {
    b = 0.0;
    if (Qy > Py) { a = -1.0; c = Px; }
    else { a = 1.0; c = -Px; }
  } else {
    a = Py - Qy;
    b = Qx - Px;
    c = -Px * a - Py * b;
  }
  double x, y;
  if (a == 0.0) { x = px; y = -c / b; }
  else if (b == 0.0) { x = -c / a; y = py; }
  else {
    double a2 = a * a, b2 = b * b, d = a2 + b2;
    x = (b2 * px - a * b * py - a * c) / d;
    y = (-a * b * px + a2 * py - b * c) / d;
  }
  px = x; py = y;
}

// squared distance of p from the line PQ (plain fp64): the score of a visible hull edge (see
// canonical_visible_hull_edge in dli_device.cuh)
DLI_HDI double hull_edge_score(double Px, double Py, double Qx, double Qy, double px, double py) {
  const double dx = Qx - Px, dy = Qy - Py;
  const double det = dx * (py - Py) - dy * (px - Px);
  return (det * det) / (dx * dx + dy * dy);
}

// a - b rounded toward zero (Gmpq subtraction + mpq_get_d, which truncates)
DLI_HDI double sub_rz(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dsub_rz(a, b);
#else
  double s, e;
  two_sum(a, -b, s, e);
  if (e == 0.0 || !std::isfinite(s)) return s;
  // exact difference = s + e with |e| <= ulp(s) / 2: the truncation is s when e points away from zero, else the
  // neighbour of s towards zero
  if ((e > 0.0) == (s > 0.0)) return s;
  return nextafter(s, 0.0);
#endif
}

// Eigen PartialPivLU determinant (unblocked), n = D, matrix row-major in registers (destroyed)
template <int D>
DLI_HDI double lu_det(double *a) {
  double sgn = 1.0;
#pragma unroll
  for (int c = 0; c < D; c++) {
    int piv = c;
    double best = fabs(a[c * D + c]);
#pragma unroll
    for (int r = c + 1; r < D; r++) {
      double v = fabs(a[r * D + c]);
      if (v > best) { best = v; piv = r; }
    }
    // swap rows c and piv (piv is dynamic: select-based swap keeps everything in registers)
#pragma unroll
    for (int r = c + 1; r < D; r++) {
      if (r == piv) {
#pragma unroll
        for (int j = 0; j < D; j++) { double t = a[c * D + j]; a[c * D + j] = a[r * D + j]; a[r * D + j] = t; }
      }
    }
    if (piv != c) sgn = -sgn;
    double pv = a[c * D + c];
#pragma unroll
    for (int r = c + 1; r < D; r++) {
      double l = (pv != 0.0) ? a[r * D + c] / pv : a[r * D + c];
      a[r * D + c] = l;
#pragma unroll
      for (int j = c + 1; j < D; j++) a[r * D + j] = a[r * D + j] - l * a[c * D + j];
    }
  }
  double prod = a[0];
#pragma unroll
  for (int i = 1; i < D; i++) prod = prod * a[i * D + i];
  return sgn * prod;
}

// signed volume; column j = pts[j+1] - pts[0]; exact difference rounded toward zero (Gmpq
// subtraction + mpq_get_d, src/delaunay_d_interp.h:608-613), det / D!
template <int D>
DLI_HDI double volD(const double *const *pts) {
  double a[D * D];
#pragma unroll
  for (int j = 0; j < D; j++)
#pragma unroll
    for (int i = 0; i < D; i++) a[i * D + j] = sub_rz(pts[j + 1][i], pts[0][i]);
  constexpr double fact = (D == 2 ? 2.0 : D == 3 ? 6.0 : D == 4 ? 24.0 : 120.0);
  return lu_det<D>(a) / fact;
}

// exact sum of products truncated toward zero (Gmpq accumulation + mpq_get_d, :998-1004): slow, exact path
template <int K>
DLI_HD_NOINLINE double dot_exact_trunc_slow(const double *b, const double *f) {
  double e[2 * K + 1], t[2 * K + 2];
  int n = 0;
  for (int i = 0; i < K; i++) {
    double p, q;
    two_prod(b[i], f[i], p, q);
    grow_ze(e, n, q);
    grow_ze(e, n, p);
  }
  double A = 0.0;
  for (int i = n - 1; i >= 0; i--) A += e[i];
  if (!isfinite(A)) return A;
  for (int it = 0; it < 8; it++) {
    int tn = n;
    for (int i = 0; i < n; i++) t[i] = e[i];
    grow_ze(t, tn, -A);
    int sd = expansion_sign(t, tn);
    int sa = (A > 0.0) - (A < 0.0);
    if (sd == 0) break;
    if (sa != 0 && sd == -sa) { A = nextafter(A, 0.0); continue; }
    double up = nextafter(A, sd > 0 ? (double)INFINITY : -(double)INFINITY);
    tn = n;
    for (int i = 0; i < n; i++) t[i] = e[i];
    grow_ze(t, tn, -up);
    int su = expansion_sign(t, tn);
    if (su == 0 || su == sd) { A = up; if (su == 0) break; continue; }
    break;
  }
  return A;
}

// Same value, register-only fast path: compensated (double-double) accumulation gives the total as h + l
// with |l| <= ulp(h)/2 up to an error far below |l|; the truncation toward zero is then h when l has the
// sign of h (or is an exact zero), else the neighbour of h towards zero.  When l is too small to trust its
// sign (relative to the accumulated rounding bound) the exact path decides.
template <int K>
DLI_HDI double dot_exact_trunc(const double *b, const double *f) {
  double s, lo = 0.0, pabs, q;
  two_prod(b[0], f[0], s, q);
  lo = q;
  pabs = fabs(s);
  bool all_exact = (q == 0.0);
#pragma unroll
  for (int i = 1; i < K; i++) {
    double p, qq, sn, e;
    two_prod(b[i], f[i], p, qq);
    two_sum(s, p, sn, e);
    s = sn;
    lo += e + qq;
    pabs += fabs(p);
    all_exact = all_exact && (e == 0.0) && (qq == 0.0);
  }
  if (all_exact) return s;  // every operation was exact: s is the exact total
  double h, l;
  two_sum(s, lo, h, l);  // (|lo| <= |s| is not guaranteed under cancellation, so not fast_two_sum)
  if (!isfinite(h)) return h;
  // rounding of `lo` itself is bounded by ~K^2 2^-106 sum|p_i|; demand a wide margin
  if (!(fabs(l) > 8.673617379884035e-19 * 1.1102230246251565e-16 * 1048576.0 * pabs)) return dot_exact_trunc_slow<K>(b, f);
  bool same = (l > 0.0) == (h > 0.0);
  return same ? h : nextafter(h, 0.0);
}

// barycentric coordinates + gather of the dD reference (:628-650, :998-1004) for the simplex with vertex coordinates
// x [(D+1)][D] and values f [D+1].
// COMPACT (exact pass of the kernels): the loop over the D+1 replaced vertices stays rolled -- that pass is bound by
// instruction fetch, and D+2 unrolled copies of the pivoting LU are most of its code.  Same operations, same order.
template <int D, bool COMPACT = false>
DLI_HDI double epilogueD_ref_xf(const double *p, const double *x, const double *f) {
  const double *pts[D + 1];
#pragma unroll
  for (int i = 0; i <= D; i++) pts[i] = &x[i * D];
  double b[D + 1];
  if (COMPACT) {
    double whole = 1.0;
#pragma unroll 1
    for (int i = -1; i <= D; i++) {
      const double *q[D + 1];
#pragma unroll
      for (int j = 0; j <= D; j++) q[j] = (j == i) ? p : pts[j];
      const double v = volD<D>(q);
      if (i < 0) whole = v;
      else {
#pragma unroll
        for (int j = 0; j <= D; j++) if (j == i) b[j] = v / whole;
      }
    }
  } else {
    double whole = volD<D>(pts);
#pragma unroll
    for (int i = 0; i <= D; i++) {
      const double *q[D + 1];
#pragma unroll
      for (int j = 0; j <= D; j++) q[j] = (j == i) ? p : pts[j];
      b[i] = volD<D>(q) / whole;
    }
  }
  return dot_exact_trunc<D + 1>(b, f);
}

}  // namespace dli
