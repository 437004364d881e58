// dli_exact.cuh -- orientation predicates with a static floating-point filter and an exact
// fallback in floating-point expansion arithmetic (Shewchuk-style), usable on host and device.
//
// Replaces what CGAL's Exact_predicates_inexact_constructions_kernel (2D,
// src/delaunay_2_interp.h:66) and Cartesian_d<Gmpq> (dD, src/delaunay_d_interp.h:43-50) give the
// reference's locate: the exact sign of the orientation determinant.
//
// Must be compiled with FMA contraction off (-fmad=false / -ffp-contract=off): the error-free
// transformations below rely on individually rounded operations; the one fused operation that is
// wanted (two_prod) is written as an explicit fma().
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define DLI_HD __host__ __device__
#define DLI_HDI __host__ __device__ __forceinline__
#define DLI_HD_NOINLINE __host__ __device__ __noinline__
#else
#define DLI_HD
#define DLI_HDI inline
#define DLI_HD_NOINLINE __attribute__((noinline))
#endif

#include <cstring>

namespace dli {

constexpr double kEps = 1.1102230246251565e-16;  // 2^-53
// |computed det - true det| <= kFilter[d] * permanent for the cofactor expansion over rounded
// differences used below.  d=2: Shewchuk's (3 + 16 eps) eps.  d=3,4: (number of roundings on the
// longest path: d entry roundings + (d-1) products + 1+2+..+(d-1) accumulations = 8 resp. 13),
// tripled for slack against the (1+eps)^k - 1 higher-order terms and the rounding of the
// permanent itself.
DLI_HDI double filter_const(int d) {
  return d == 2 ? (3.0 + 16.0 * kEps) * kEps : (d == 3 ? 24.0 * kEps : 64.0 * kEps);
}

DLI_HDI void two_sum(double a, double b, double &x, double &y) {
  double s = a + b;
  double bv = s - a;
  double av = s - bv;
  x = s;
  y = (a - av) + (b - bv);
}
DLI_HDI void two_prod(double a, double b, double &x, double &y) {
  double p = a * b;
  x = p;
  y = fma(a, b, -p);
}

// e[0..n) += b, e grows to n+1 components (zeros are kept; the expansion stays nonoverlapping)
DLI_HDI void grow(double *e, int &n, double b) {
  double Q = b;
  for (int i = 0; i < n; i++) {
    double Qn, q;
    two_sum(Q, e[i], Qn, q);
    e[i] = q;
    Q = Qn;
  }
  e[n++] = Q;
}
// same with zero elimination (keeps long expansions short)
DLI_HDI void grow_ze(double *e, int &n, double b) {
  double Q = b;
  int h = 0;
  for (int i = 0; i < n; i++) {
    double Qn, q;
    two_sum(Q, e[i], Qn, q);
    Q = Qn;
    if (q != 0.0) e[h++] = q;
  }
  if (Q != 0.0 || h == 0) e[h++] = Q;
  n = h;
}
DLI_HDI int expansion_sign(const double *e, int n) {
  for (int i = n - 1; i >= 0; i--) {
    if (e[i] > 0.0) return 1;
    if (e[i] < 0.0) return -1;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------
// 2D
// ---------------------------------------------------------------------------------------------

// exact sign of (ax-cx)(by-cy) - (ay-cy)(bx-cx): the six surviving products of raw coordinates,
// each split exactly with two_prod, summed exactly.
template <int D>
DLI_HD inline bool orient_exact_int(const double *pts, int &sign);   // fixed-point integer path, defined below

DLI_HD inline int orient2d_exact(double ax, double ay, double bx, double by, double cx, double cy) {
  // Common case first (Shewchuk's adaptive stage B): when the four coordinate differences are exact -- always for
  // points on a lattice, and for most near-degenerate inputs, where the operands are close -- the determinant is the
  // difference of two products of doubles: two two_prods and a 4-term expansion.
  {
    double acx, acxt, bcx, bcxt, acy, acyt, bcy, bcyt;
    two_sum(ax, -cx, acx, acxt);
    two_sum(bx, -cx, bcx, bcxt);
    two_sum(ay, -cy, acy, acyt);
    two_sum(by, -cy, bcy, bcyt);
    if (acxt == 0.0 && bcxt == 0.0 && acyt == 0.0 && bcyt == 0.0) {
      double l1, l0, r1, r0, e[4];
      int n = 0;
      two_prod(acx, bcy, l1, l0);
      two_prod(acy, bcx, r1, r0);
      grow(e, n, l0); grow(e, n, -r0); grow(e, n, l1); grow(e, n, -r1);
      return expansion_sign(e, n);
    }
  }
  {  // coordinates on a common fixed-point grid: one 128-bit 2 x 2 determinant (e.g. a point projected on a hull
    // edge line tested against that edge: differences are not exact, the filter cannot decide)
    const double pts[6] = {ax, ay, bx, by, cx, cy};
    int s;
    if (orient_exact_int<2>(pts, s)) return s;
  }
  double e[12];
  int n = 0;
  const double pa[6] = {ax, -ax, -cx, -ay, ay, cy};
  const double pb[6] = {by, cy, by, bx, cx, bx};
  for (int i = 0; i < 6; i++) {
    double p, q;
    two_prod(pa[i], pb[i], p, q);
    grow(e, n, q);
    grow(e, n, p);
  }
  return expansion_sign(e, n);
}

// filtered evaluation.  Returns the sign if certain, otherwise 2.
DLI_HDI int orient2d_filtered(double ax, double ay, double bx, double by, double cx, double cy) {
  double l = (ax - cx) * (by - cy);
  double r = (ay - cy) * (bx - cx);
  double det = l - r;
  double bound = filter_const(2) * (fabs(l) + fabs(r));
  if (det > bound) return 1;
  if (-det > bound) return -1;
  return 2;
}
DLI_HDI int orient2d(double ax, double ay, double bx, double by, double cx, double cy) {
  int s = orient2d_filtered(ax, ay, bx, by, cx, cy);
  return s == 2 ? orient2d_exact(ax, ay, bx, by, cx, cy) : s;
}

// ---------------------------------------------------------------------------------------------
// d dimensions (d = 3, 4): sign of det[p_i 1], i = 0..d  (positive = our canonical orientation)
// ---------------------------------------------------------------------------------------------

// determinant and permanent of a K x K matrix held in registers (row-major), cofactor expansion
// along column 0.
template <int K>
struct DetPerm {
  DLI_HDI static double eval(const double *m, double &perm) {
    double det = 0.0, pm = 0.0;
#pragma unroll
    for (int r = 0; r < K; r++) {
      double sub[(K - 1) * (K - 1)];
      int rr = 0;
#pragma unroll
      for (int i = 0; i < K; i++) {
        if (i == r) continue;
#pragma unroll
        for (int j = 1; j < K; j++) sub[rr * (K - 1) + (j - 1)] = m[i * K + j];
        rr++;
      }
      double sp;
      double sd = DetPerm<K - 1>::eval(sub, sp);
      double t = m[r * K] * sd;
      det = (r & 1) ? det - t : det + t;
      pm += fabs(m[r * K]) * sp;
    }
    perm = pm;
    return det;
  }
};
template <>
struct DetPerm<2> {
  DLI_HDI static double eval(const double *m, double &perm) {
    double a = m[0] * m[3], b = m[1] * m[2];
    perm = fabs(a) + fabs(b);
    return a - b;
  }
};
template <>
struct DetPerm<1> {
  DLI_HDI static double eval(const double *m, double &perm) {
    perm = fabs(m[0]);
    return m[0];
  }
};

// filtered orientation of D+1 points; q replaces vertex `slot` when slot >= 0.
// pts: (D+1) x D row-major.  Returns sign if certain else 2.
template <int D>
DLI_HDI int orient_filtered(const double *pts) {
  double m[D * D];
#pragma unroll
  for (int i = 0; i < D; i++)
#pragma unroll
    for (int j = 0; j < D; j++) m[i * D + j] = pts[i * D + j] - pts[D * D + j];
  double perm;
  double det = DetPerm<D>::eval(m, perm);
  double bound = filter_const(D) * perm;
  if (det > bound) return 1;
  if (-det > bound) return -1;
  return 2;
}

// Exact product of k doubles accumulated into an expansion: the running product is itself an
// expansion (at most 2^(k-1) components) scaled exactly term by term.
// Exact determinant sign of the (D+1)x(D+1) homogeneous matrix [1 p]: sum over permutations is
// organised as a Laplace expansion along the ones column; every minor is a D x D determinant of
// raw coordinates expanded into its D! signed products.
template <int D>
struct ExactCap {  // components of the full expansion (before zero elimination)
  static constexpr int fact = (D == 1 ? 1 : D == 2 ? 2 : D == 3 ? 6 : D == 4 ? 24 : 120);
  static constexpr int value = (D + 1) * fact * (1 << (D - 1));
};

// ---- exact sign through fixed-point integers (the common case of the exact fallback) ----
// Every finite double is an integer multiple of a power of two.  When all (D+1) x D coordinates of a predicate fit a
// common grid 2^g with at most 61 significant bits (coordinates of similar magnitude: true for meshes in a box unless
// a coordinate is within 2^-8 of zero relative to the largest), the differences are 63-bit integers and the D x D
// determinant is evaluated exactly in 128 / 256-bit two's-complement arithmetic: a few dozen integer multiplies
// instead of an expansion of up to 96 (3D) / 960 (4D) components.  Queries that sit exactly on facets of a regular
// grid -- every orientation test an exact zero, the filter can never certify them -- depend on this path.
struct I256 {
  unsigned long long w[4];  // little-endian limbs, two's complement
};
DLI_HDI I256 i256_from_i128(__int128 v) {
  I256 r;
  r.w[0] = (unsigned long long)v;
  r.w[1] = (unsigned long long)(v >> 64);
  r.w[2] = r.w[3] = (v < 0) ? ~0ull : 0ull;
  return r;
}
DLI_HDI I256 i256_neg(const I256 &a) {
  I256 r;
  unsigned long long c = 1;
  for (int i = 0; i < 4; i++) {
    unsigned long long t = ~a.w[i] + c;
    c = (c && t == 0) ? 1ull : 0ull;
    r.w[i] = t;
  }
  return r;
}
DLI_HDI I256 i256_add(const I256 &a, const I256 &b) {
  I256 r;
  unsigned long long c = 0;
  for (int i = 0; i < 4; i++) {
    unsigned __int128 t = (unsigned __int128)a.w[i] + b.w[i] + c;
    r.w[i] = (unsigned long long)t;
    c = (unsigned long long)(t >> 64);
  }
  return r;
}
DLI_HDI I256 i256_mul_i64(const I256 &a, long long b) {  // no overflow by construction (see the bit budget above)
  const bool na = (a.w[3] >> 63) != 0, nb = b < 0;
  const I256 ma = na ? i256_neg(a) : a;
  const unsigned long long mb = nb ? (unsigned long long)(-(b + 1)) + 1ull : (unsigned long long)b;
  I256 r;
  unsigned long long c = 0;
  for (int i = 0; i < 4; i++) {
    unsigned __int128 t = (unsigned __int128)ma.w[i] * mb + c;
    r.w[i] = (unsigned long long)t;
    c = (unsigned long long)(t >> 64);
  }
  return (na != nb) ? i256_neg(r) : r;
}
DLI_HDI I256 i256_mul_i128(__int128 a, __int128 b) {   // |a|, |b| < 2^126
  const bool neg = (a < 0) != (b < 0);
  const unsigned __int128 ua = a < 0 ? (unsigned __int128)0 - (unsigned __int128)a : (unsigned __int128)a;
  const unsigned __int128 ub = b < 0 ? (unsigned __int128)0 - (unsigned __int128)b : (unsigned __int128)b;
  const unsigned long long a0 = (unsigned long long)ua, a1 = (unsigned long long)(ua >> 64);
  const unsigned long long b0 = (unsigned long long)ub, b1 = (unsigned long long)(ub >> 64);
  const unsigned __int128 p00 = (unsigned __int128)a0 * b0, p01 = (unsigned __int128)a0 * b1,
                          p10 = (unsigned __int128)a1 * b0, p11 = (unsigned __int128)a1 * b1;
  I256 r;
  r.w[0] = (unsigned long long)p00;
  const unsigned __int128 mid = (p00 >> 64) + (unsigned long long)p01 + (unsigned long long)p10;
  r.w[1] = (unsigned long long)mid;
  const unsigned __int128 hi = (mid >> 64) + (p01 >> 64) + (p10 >> 64) + (unsigned long long)p11;
  r.w[2] = (unsigned long long)hi;
  r.w[3] = (unsigned long long)((hi >> 64) + (p11 >> 64));
  return neg ? i256_neg(r) : r;
}
DLI_HDI int i256_sign(const I256 &a) {
  if (a.w[3] >> 63) return -1;
  return (a.w[0] | a.w[1] | a.w[2] | a.w[3]) ? 1 : 0;
}
DLI_HDI int dli_ctz64(unsigned long long m) {
#ifdef __CUDA_ARCH__
  return __ffsll((long long)m) - 1;
#else
  return __builtin_ctzll(m);
#endif
}
DLI_HDI long long dli_double_bits(double x) {
#ifdef __CUDA_ARCH__
  return __double_as_longlong(x);
#else
  long long b;
  memcpy(&b, &x, sizeof b);
  return b;
#endif
}
DLI_HDI __int128 i128_minor2(long long a, long long b, long long c, long long d) { return (__int128)a * d - (__int128)b * c; }
// 3 x 3 determinant of 63-bit entries (row-major, stride ld)
DLI_HDI I256 i256_det3(const long long *m, int ld) {
  I256 r = i256_mul_i64(i256_from_i128(i128_minor2(m[ld + 1], m[ld + 2], m[2 * ld + 1], m[2 * ld + 2])), m[0]);
  r = i256_add(r, i256_neg(i256_mul_i64(i256_from_i128(i128_minor2(m[ld], m[ld + 2], m[2 * ld], m[2 * ld + 2])), m[1])));
  r = i256_add(r, i256_mul_i64(i256_from_i128(i128_minor2(m[ld], m[ld + 1], m[2 * ld], m[2 * ld + 1])), m[2]));
  return r;
}
// exact sign of det[p_i - p_D]; false = the coordinates do not fit the fixed-point budget (caller falls back)
// (loops deliberately NOT unrolled: the exact pass is bound by instruction fetch -- its hot code must stay inside the
// 32 KB L1.5 instruction cache -- not by the few hundred integer operations of one evaluation)
template <int D>
DLI_HD inline bool orient_exact_int(const double *pts, int &sign) {
  static_assert(D >= 2 && D <= 4, "integer path is written for D = 2, 3, 4");
  int gexp = 1 << 30, top = -(1 << 30);
#pragma unroll 1
  for (int i = 0; i < (D + 1) * D; i++) {
    const long long b = dli_double_bits(pts[i]);
    const int e = (int)((b >> 52) & 0x7ff);
    const unsigned long long frac = (unsigned long long)b & ((1ull << 52) - 1);
    if (e == 0x7ff) return false;                 // Inf / NaN
    if (e == 0) { if (frac == 0) continue; return false; }   // zero: no constraint; subnormal: fall back
    const unsigned long long m = frac | (1ull << 52);
    const int lsb = e - 1075 + dli_ctz64(m), msb = e - 1023;
    gexp = lsb < gexp ? lsb : gexp;
    top = msb > top ? msb : top;
  }
  if (top < gexp) { sign = 0; return true; }      // all coordinates zero
  if (top - gexp > 60) return false;
  long long X[(D + 1) * D];
#pragma unroll 1
  for (int i = 0; i < (D + 1) * D; i++) {
    const long long b = dli_double_bits(pts[i]);
    const int e = (int)((b >> 52) & 0x7ff);
    if (e == 0) { X[i] = 0; continue; }
    const unsigned long long m = ((unsigned long long)b & ((1ull << 52) - 1)) | (1ull << 52);
    const int sh = e - 1075 - gexp;               // >= -ctz(m)
    const long long v = (long long)(sh >= 0 ? m << sh : m >> (-sh));
    X[i] = b < 0 ? -v : v;
  }
  long long M[D * D];
#pragma unroll 1
  for (int i = 0; i < D; i++)
#pragma unroll
    for (int j = 0; j < D; j++) M[i * D + j] = X[i * D + j] - X[D * D + j];
  if (D == 2) {
    const __int128 d2 = i128_minor2(M[0], M[1], M[2], M[3]);
    sign = d2 > 0 ? 1 : (d2 < 0 ? -1 : 0);
  } else if (D == 3) {
    sign = i256_sign(i256_det3(M, 3));
  } else {
    // Laplace expansion by the first two rows: six products of complementary 2 x 2 minors (|minor| < 2^125)
    I256 acc;
    acc.w[0] = acc.w[1] = acc.w[2] = acc.w[3] = 0;
    const int ci[6] = {0, 0, 0, 1, 1, 2}, cj[6] = {1, 2, 3, 2, 3, 3}, ck[6] = {2, 1, 1, 0, 0, 0}, cl[6] = {3, 3, 2, 3, 2, 1};
#pragma unroll 1
    for (int t = 0; t < 6; t++) {
      const __int128 top = i128_minor2(M[ci[t]], M[cj[t]], M[4 + ci[t]], M[4 + cj[t]]);
      const __int128 bot = i128_minor2(M[8 + ck[t]], M[8 + cl[t]], M[12 + ck[t]], M[12 + cl[t]]);
      const I256 pr = i256_mul_i128(top, bot);
      acc = i256_add(acc, ((ci[t] + cj[t]) & 1) ? pr : i256_neg(pr));   // sign (-1)^(i + j + 1)
    }
    sign = i256_sign(acc);
  }
  return true;
}

// merge-free exact accumulation: products are generated one at a time and added with grow_ze.
// Cost O(n^2) two_sums in the worst case; n <= 96 (D=3) / 960 (D=4); it only runs when the filter fails.
// (kept out of line: it is large -- 8 KB of local memory in 4D -- and rare; inlined, it bloated the exact-pass kernels
// to 8.7k instructions / 168 registers and the pass stalled on instruction fetch)
template <int D>
DLI_HD_NOINLINE int orient_exact_expansion(const double *pts);
template <int D>
DLI_HD inline int orient_exact(const double *pts) {
  int s;
  if (orient_exact_int<D>(pts, s)) return s;
  return orient_exact_expansion<D>(pts);
}
template <int D>
DLI_HD_NOINLINE int orient_exact_expansion(const double *pts) {
  double e[ExactCap<D>::value];
  int n = 0;
  // iterate over the rows r left out (the Laplace expansion along the ones column) and over all
  // permutations of the D columns for the remaining D rows.
  int rows[D];
  for (int r = 0; r <= D; r++) {
    int c = 0;
    for (int i = 0; i <= D; i++)
      if (i != r) rows[c++] = i;
    // sign of the cofactor of entry (r, D) (ones as the LAST column): (-1)^(r + D)
    int base_sign = ((r + D) & 1) ? -1 : 1;
    // Heap-free permutation enumeration via factorial number system
    constexpr int F = ExactCap<D>::fact;
    for (int code = 0; code < F; code++) {
      int perm[D], avail[D];
      for (int i = 0; i < D; i++) avail[i] = i;
      int cc = code, parity = 0, left = D;
      for (int i = 0; i < D; i++) {
        int radix = left;
        int idx = cc % radix;
        cc /= radix;
        perm[i] = avail[idx];
        parity += idx;  // number of smaller remaining elements skipped = inversions contributed
        for (int j = idx; j < left - 1; j++) avail[j] = avail[j + 1];
        left--;
      }
      double sgn = ((parity & 1) ? -1.0 : 1.0) * (double)base_sign;
      // exact product of D factors as an expansion t[0..tn)
      double t[1 << (D - 1)];
      int tn = 1;
      t[0] = sgn * pts[rows[0] * D + perm[0]];
      for (int i = 1; i < D; i++) {
        double b = pts[rows[i] * D + perm[i]];
        double u[1 << (D - 1)];
        int un = 0;
        for (int j = 0; j < tn; j++) {
          double p, q;
          two_prod(t[j], b, p, q);
          u[un++] = q;
          u[un++] = p;
        }
        for (int j = 0; j < un; j++) t[j] = u[j];
        tn = un;
      }
      // the components of t are exact partial products; add them one by one (any order is exact)
      for (int j = 0; j < tn; j++)
        if (t[j] != 0.0) grow_ze(e, n, t[j]);
    }
  }
  return expansion_sign(e, n);
}

template <int D>
DLI_HDI int orient(const double *pts) {
  int s = orient_filtered<D>(pts);
  return s == 2 ? orient_exact<D>(pts) : s;
}

// run-time dimension front end for host code (flattener)
inline int orient_host(int d, const double *pts) {
  switch (d) {
    case 2: return orient2d(pts[0], pts[1], pts[2], pts[3], pts[4], pts[5]);
    case 3: return orient<3>(pts);
    case 4: return orient<4>(pts);
    default: return 0;
  }
}

}  // namespace dli
