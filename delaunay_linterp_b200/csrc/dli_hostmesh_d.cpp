// dli_hostmesh_d.cpp -- host-side incremental d-dimensional Delaunay triangulation (d = 2, 3, 4:
// Bowyer-Watson with an infinite vertex, exact orientation / in-sphere predicates) and the adaptive
// insertion queue of Delaunay_incremental_interp_d; C ABI `dli_dd_*` in include/dli_host.h.
//
// This is where the reference keeps its patched CGAL::Delaunay_d (modified_cgal/Delaunay_d.h on top of
// modified_cgal/Convex_hull_d.h: the lower hull of the points lifted to the paraboloid, with a history
// DAG), the vertex -> value map m_pointMap (src/delaunay_d_interp.h:757) and the boost skew_heap queue
// (:750-751, :770, :832-848, :874-916, :939-956).  Construction is host work by design (north_star); the GPU
// engine consumes the flattened result (dli_dd_export -> dli_mesh_create).  The triangulation is built
// directly in d dimensions (no lifting, no history): for points in general position the Delaunay
// triangulation is unique, so it equals the reference's; among cospherical points any valid Delaunay
// triangulation is accepted (the reference's choice there depends on its hull's facet triangulation).
#include <algorithm>
#include <climits>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <new>
#include <queue>
#include <vector>

#include "../../include/dli_error_math.h"
#include "../../include/dli_host.h"
#include "dli_exact.cuh"
#include "dli_host_small.h"
#include "dli_refmath.h"

namespace {

using dli::two_prod;
using dli::two_sum;

constexpr int MAXD = DLI_MAX_DIM;
constexpr int INF = -1;  // the infinite vertex

// ---------------------------------------------------------------------------------------------
// expansion arithmetic (Shewchuk 1997: nonoverlapping expansions, components by increasing magnitude)
// ---------------------------------------------------------------------------------------------
typedef std::vector<double> Exp;

inline void fast_two_sum(double a, double b, double &x, double &y) {
  x = a + b;
  y = b - (x - a);
}
// h = e + f, zero components eliminated (FAST-EXPANSION-SUM)
void ex_sum(const Exp &e, const Exp &f, Exp &h) {
  h.clear();
  const size_t el = e.size(), fl = f.size();
  if (el == 0) { h = f; return; }
  if (fl == 0) { h = e; return; }
  size_t ei = 0, fi = 0;
  auto take_e = [&]() { return fi >= fl || (ei < el && std::fabs(e[ei]) <= std::fabs(f[fi])); };
  double Q = take_e() ? e[ei++] : f[fi++];
  bool first = true;
  while (ei < el || fi < fl) {
    const double nx = take_e() ? e[ei++] : f[fi++];
    double Qn, hh;
    if (first) { fast_two_sum(nx, Q, Qn, hh); first = false; }
    else two_sum(Q, nx, Qn, hh);
    Q = Qn;
    if (hh != 0.0) h.push_back(hh);
  }
  if (Q != 0.0 || h.empty()) h.push_back(Q);
}
// h = b * e (SCALE-EXPANSION)
void ex_scale(const Exp &e, double b, Exp &h) {
  h.clear();
  if (e.empty()) return;
  double Q, hh;
  two_prod(e[0], b, Q, hh);
  if (hh != 0.0) h.push_back(hh);
  for (size_t i = 1; i < e.size(); i++) {
    double p1, p0, s;
    two_prod(e[i], b, p1, p0);
    two_sum(Q, p0, s, hh);
    if (hh != 0.0) h.push_back(hh);
    fast_two_sum(p1, s, Q, hh);
    if (hh != 0.0) h.push_back(hh);
  }
  if (Q != 0.0 || h.empty()) h.push_back(Q);
}
inline bool ex_zero(const Exp &e) { return e.empty() || (e.size() == 1 && e[0] == 0.0); }
// h = a * b
void ex_mul(const Exp &a, const Exp &b, Exp &h) {
  h.clear();
  if (ex_zero(a) || ex_zero(b)) { h.push_back(0.0); return; }
  const Exp &l = a.size() >= b.size() ? a : b, &s = a.size() >= b.size() ? b : a;
  Exp t, u;
  for (size_t j = 0; j < s.size(); j++) {
    ex_scale(l, s[j], t);
    if (j == 0) h = t;
    else { ex_sum(h, t, u); h.swap(u); }
  }
}
inline int ex_sign(const Exp &e) {
  if (e.empty()) return 0;
  const double t = e.back();
  return t > 0.0 ? 1 : (t < 0.0 ? -1 : 0);
}

// orientation of d+1 points in d dimensions, d = 1 .. 4 (exact sign; dli_exact.cuh for d >= 2)
int orient_any(int d, const double *pts) {
  if (d == 1) return pts[0] > pts[1] ? 1 : (pts[0] < pts[1] ? -1 : 0);
  return dli::orient_host(d, pts);
}

// In-sphere determinant of D+2 points in D dimensions: rows (v_i - p, |v_i - p|^2), i = 0 .. D.
// Filtered fp64 evaluation (cofactor expansion with the permanent as the error scale), then exact expansions.
template <int D>
int insphere_raw(const double *v /* (D+1) x D */, const double *p) {
  constexpr int K = D + 1;
  double m[K * K];
  for (int i = 0; i < K; i++) {
    double l = 0.0;
    for (int j = 0; j < D; j++) {
      const double d = v[i * D + j] - p[j];
      m[i * K + j] = d;
      l += d * d;
    }
    m[i * K + D] = l;
  }
  double perm;
  const double det = dli::DetPerm<K>::eval(m, perm);
  // roundings on the longest path: D entry differences + the lift (D + 2) + K - 1 products + K (K + 1) / 2
  // accumulations <= 34 for K = 5; tripled for the higher-order terms and the rounding of the permanent itself
  const double bound = (16.0 * K + 16.0) * dli::kEps * perm;
  if (perm > 1e-280 && perm < 1e280) {
    if (det > bound) return 1;
    if (-det > bound) return -1;
  }
  // exact: entries as expansions, determinant by dynamic programming over column subsets (rows taken in order)
  Exp a[K][K], t, u, w;
  for (int i = 0; i < K; i++) {
    Exp lift;
    for (int j = 0; j < D; j++) {
      double hi, lo;
      two_sum(v[i * D + j], -p[j], hi, lo);
      Exp &d = a[i][j];
      d.clear();
      if (lo != 0.0) d.push_back(lo);
      if (hi != 0.0 || d.empty()) d.push_back(hi);
      ex_mul(d, d, t);
      if (j == 0) lift = t;
      else { ex_sum(lift, t, u); lift.swap(u); }
    }
    a[i][D] = lift;
  }
  Exp minor[1 << K];
  minor[0].assign(1, 1.0);
  for (int mask = 1; mask < (1 << K); mask++) {
    const int r = __builtin_popcount((unsigned)mask) - 1;   // the row this minor ends with
    Exp acc(1, 0.0);
    int pos = 0;
    for (int j = 0; j < K; j++) {
      if (!(mask >> j & 1)) continue;
      ex_mul(a[r][j], minor[mask & ~(1 << j)], t);
      if (((r + pos) & 1) != 0)
        for (double &c : t) c = -c;
      ex_sum(acc, t, w);
      acc.swap(w);
      pos++;
    }
    minor[mask].swap(acc);
  }
  return ex_sign(minor[(1 << K) - 1]);
}
int insphere_raw_any(int d, const double *v, const double *p) {
  switch (d) {
    case 2: return insphere_raw<2>(v, p);
    case 3: return insphere_raw<3>(v, p);
    case 4: return insphere_raw<4>(v, p);
    default: return 0;
  }
}

// ---------------------------------------------------------------------------------------------
// triangulation
// ---------------------------------------------------------------------------------------------
struct Sx {
  int v[MAXD + 1];
  int n[MAXD + 1];
};

struct RidgeEntry {
  int key[MAXD];
  int s, slot;
  uint32_t stamp;
};

struct HostMeshD {
  int D = 0;
  std::vector<double> X;   // [nv][D]
  std::vector<double> F;
  std::vector<Sx> sx;
  std::vector<uint8_t> alive, mark;   // mark: 1 = in the cavity, 2 = tested, not in conflict
  std::vector<uint32_t> gen;
  std::vector<int> freelist;
  int cur_dim = -1;        // affine dimension of the points so far (-1: none); == D once triangulated
  std::vector<int> basis;  // cur_dim + 1 affinely independent vertices (until triangulated)
  int last = -1;           // some alive finite simplex
  int64_t n_finite = 0;
  int sphere_sign = 0;     // sign of insphere_raw for a point inside a positively oriented simplex
  uint32_t rng = 0x9e3779b9u;
  std::vector<int> new_sx, removed_sx;   // finite simplices created / removed by the last insertion
  // Stable compact ids of the finite simplices (what the device mesh and simplex_out use).  An insertion hands the ids of
  // the finite simplices it removes to the ones it creates and appends the rest, so ids change only inside the cavity
  // and the id space stays 0 .. n_finite-1 -- unless an insertion removes more finite simplices than it creates (possible
  // in 3D / 4D: a large cavity with few boundary facets): then fid_compact drops and the next flatten renumbers.
  std::vector<int32_t> fid, free_fid;
  int32_t next_fid = 0;
  bool fid_compact = true;
  void renumber() {
    int32_t c = 0;
    for (size_t s = 0; s < sx.size(); s++) fid[s] = (alive[s] && !is_inf((int)s)) ? c++ : -1;
    next_fid = c;
    free_fid.clear();
    fid_compact = true;
  }
  // scratch
  std::vector<int> cavity, stack, tested, created;
  std::vector<RidgeEntry> table;
  uint32_t stamp = 0;

  int64_t nv() const { return (int64_t)F.size(); }
  const double *pt(int v) const { return &X[(size_t)v * D]; }
  int inf_slot(int s) const {
    for (int i = 0; i <= D; i++)
      if (sx[s].v[i] == INF) return i;
    return -1;
  }
  bool is_inf(int s) const { return inf_slot(s) >= 0; }
  int new_simplex() {
    int id;
    if (!freelist.empty()) { id = freelist.back(); freelist.pop_back(); }
    else { id = (int)sx.size(); sx.push_back(Sx()); alive.push_back(0); mark.push_back(0); gen.push_back(0); fid.push_back(-1); }
    alive[id] = 1;
    fid[id] = -1;
    return id;
  }
  // orientation of the simplex `s` with the vertex in slot k replaced by the point p (k < 0: nothing replaced)
  int orient_replaced(int s, int k, const double *p) const {
    double pts[(MAXD + 1) * MAXD];
    for (int i = 0; i <= D; i++) {
      const double *src = (i == k) ? p : pt(sx[s].v[i]);
      for (int c = 0; c < D; c++) pts[i * D + c] = src[c];
    }
    return orient_any(D, pts);
  }
  // > 0: p strictly inside the circumsphere of the finite, positively oriented simplex s
  int insphere(int s, const double *p) {
    double v[(MAXD + 1) * MAXD];
    for (int i = 0; i <= D; i++)
      for (int c = 0; c < D; c++) v[i * D + c] = pt(sx[s].v[i])[c];
    if (sphere_sign == 0) {   // calibrate the sign convention on this simplex: its centroid is inside
      double c[MAXD];
      for (int k = 0; k < D; k++) {
        c[k] = 0.0;
        for (int i = 0; i <= D; i++) c[k] += v[i * D + k];
        c[k] /= (D + 1);
      }
      sphere_sign = insphere_raw_any(D, v, c);
      if (sphere_sign == 0) sphere_sign = 1;   // (cannot happen for a non-degenerate simplex)
    }
    return sphere_sign * insphere_raw_any(D, v, p);
  }
  // Is p in conflict with simplex s?  Finite: strictly inside the circumsphere.  Infinite (hull facet H, infinite vertex
  // in slot j; the invariant of the structure is that s with slot j replaced by a point beyond H is positively
  // oriented): strictly beyond H, or in the hyperplane of H and inside the circumsphere of the finite simplex behind H
  // (= inside the circumsphere of H within its hyperplane).
  bool conflict(int s, const double *p) {
    const int j = inf_slot(s);
    if (j < 0) return insphere(s, p) > 0;
    const int o = orient_replaced(s, j, p);
    if (o != 0) return o > 0;
    return insphere(sx[s].n[j], p) > 0;
  }

  // Exact visibility walk from `start` (remembering, facets tried from a random first index: terminates with
  // probability 1 on any triangulation).  Returns a finite simplex whose closed hull contains p, or the infinite simplex
  // whose hull facet p lies strictly beyond.
  int locate(const double *p, int start) {
    int s = start;
    if (s < 0 || !alive[s]) s = last;
    { const int j = inf_slot(s); if (j >= 0) s = sx[s].n[j]; }
    int from = -1;
    const int64_t cap = (int64_t)sx.size() * 4 + 64;
    for (int64_t step = 0; step < cap; step++) {
      rng = rng * 1664525u + 1013904223u;
      const int k0 = (int)((rng >> 16) % (unsigned)(D + 1));
      int moved = -1;
      for (int t = 0; t <= D; t++) {
        const int k = (k0 + t) % (D + 1);
        if (sx[s].n[k] == from) continue;
        if (orient_replaced(s, k, p) < 0) { moved = k; break; }
      }
      if (moved < 0) return s;
      const int nb = sx[s].n[moved];
      if (is_inf(nb)) return nb;
      from = s;
      s = nb;
    }
    // (not reached on a valid triangulation) exhaustive search
    for (int t = 0; t < (int)sx.size(); t++) {
      if (!alive[t]) continue;
      const int j = inf_slot(t);
      if (j >= 0) { if (orient_replaced(t, j, p) > 0) return t; continue; }
      bool in = true;
      for (int k = 0; k <= D && in; k++) in = orient_replaced(t, k, p) >= 0;
      if (in) return t;
    }
    return last;
  }

  // ---- ridge table: links the new simplices of one insertion to each other
  static uint32_t hash_key(const int *key, int n) {
    uint32_t h = 2166136261u;
    for (int i = 0; i < n; i++) { h ^= (uint32_t)key[i] + 0x9e3779b9u; h *= 16777619u; h ^= h >> 15; }
    return h;
  }
  void table_begin(size_t n_entries) {
    size_t want = 64;
    while (want < 2 * n_entries) want <<= 1;
    if (table.size() < want) { table.assign(want, RidgeEntry()); stamp = 0; }
    if (++stamp == 0) { for (RidgeEntry &e : table) e.stamp = 0; stamp = 1; }
  }
  // the facet of s opposite slot m (which contains the new vertex p): first sight is stored, second sight links
  void link_ridge(int s, int m, int p) {
    int key[MAXD], nk = 0;
    for (int i = 0; i <= D; i++) {
      const int v = sx[s].v[i];
      if (i != m && v != p) key[nk++] = v;
    }
    std::sort(key, key + nk);
    const size_t mask = table.size() - 1;
    for (size_t h = hash_key(key, nk) & mask;; h = (h + 1) & mask) {
      RidgeEntry &e = table[h];
      if (e.stamp != stamp) {
        e.stamp = stamp; e.s = s; e.slot = m;
        for (int i = 0; i < nk; i++) e.key[i] = key[i];
        return;
      }
      bool same = true;
      for (int i = 0; i < nk && same; i++) same = e.key[i] == key[i];
      if (same) {
        sx[s].n[m] = e.s;
        sx[e.s].n[e.slot] = s;
        return;
      }
    }
  }

  // insert vertex p (already appended, no duplicate) given the result of locate()
  void carve(int loc, int p) {
    const double *pp = pt(p);
    cavity.clear(); stack.clear(); tested.clear(); created.clear();
    stack.push_back(loc);
    mark[loc] = 1;
    while (!stack.empty()) {
      const int s = stack.back();
      stack.pop_back();
      cavity.push_back(s);
      for (int k = 0; k <= D; k++) {
        const int nb = sx[s].n[k];
        if (mark[nb]) continue;
        if (conflict(nb, pp)) { mark[nb] = 1; stack.push_back(nb); }
        else { mark[nb] = 2; tested.push_back(nb); }
      }
    }
    // a new simplex for every boundary facet: the cavity simplex with the vertex opposite the facet replaced by p
    // (keeps positive orientation: p and that vertex are on the same side of the facet -- the cavity is star-shaped --
    // and the invariant of the infinite simplices, see conflict())
    for (int c : cavity)
      for (int k = 0; k <= D; k++) {
        const int out = sx[c].n[k];
        if (mark[out] == 1) continue;
        const int id = new_simplex();
        Sx &S = sx[id];
        const Sx &C = sx[c];
        for (int i = 0; i <= D; i++) { S.v[i] = C.v[i]; S.n[i] = -1; }
        S.v[k] = p;
        S.n[k] = out;
        for (int i = 0; i <= D; i++)
          if (sx[out].n[i] == c) { sx[out].n[i] = id; break; }
        created.push_back(id);
      }
    table_begin(created.size() * (size_t)D);
    for (int id : created)
      for (int m = 0; m <= D; m++)
        if (sx[id].v[m] != p) link_ridge(id, m, p);
    for (int t : tested) mark[t] = 0;
    for (int c : cavity) {
      mark[c] = 0;
      alive[c] = 0;
      gen[c]++;
      if (!is_inf(c)) { removed_sx.push_back(c); n_finite--; free_fid.push_back(fid[c]); fid[c] = -1; }
      freelist.push_back(c);
    }
    for (int id : created)
      if (!is_inf(id)) {
        new_sx.push_back(id); n_finite++; last = id;
        if (free_fid.empty()) fid[id] = next_fid++;
        else { fid[id] = free_fid.back(); free_fid.pop_back(); }
      }
    if (!free_fid.empty()) fid_compact = false;   // fewer created than removed: holes in the id space
  }

  // does the point p leave the affine hull of the basis?  (exact: some (k x k) minor of the differences is non-zero,
  // evaluated as an orientation in the k coordinates it selects)
  bool extends_basis(const double *p) const {
    const int k = (int)basis.size();
    if (k == 0) return true;
    int cols[MAXD];
    for (int i = 0; i < k; i++) cols[i] = i;
    for (;;) {
      double pts[(MAXD + 1) * MAXD];
      for (int i = 0; i <= k; i++) {
        const double *src = i < k ? pt(basis[i]) : p;
        for (int c = 0; c < k; c++) pts[i * k + c] = src[cols[c]];
      }
      if (orient_any(k, pts) != 0) return true;
      int i = k - 1;   // next k-subset of the D columns
      while (i >= 0 && cols[i] == D - k + i) i--;
      if (i < 0) return false;
      cols[i]++;
      for (int j = i + 1; j < k; j++) cols[j] = cols[j - 1] + 1;
    }
  }
  // the first simplex (the basis) and the D + 1 infinite simplices around it
  void start_triangulation() {
    const int t0 = new_simplex();
    for (int i = 0; i <= D; i++) sx[t0].v[i] = basis[i];
    if (orient_replaced(t0, -1, nullptr) < 0) std::swap(sx[t0].v[0], sx[t0].v[1]);
    int inf_id[MAXD + 1];
    for (int k = 0; k <= D; k++) {
      const int id = new_simplex();
      inf_id[k] = id;
      for (int i = 0; i <= D; i++) sx[id].v[i] = sx[t0].v[i];
      sx[id].v[k] = INF;
      // t0 with slot k replaced by a point beyond facet k is negatively oriented: exchange two finite slots
      const int a = (k + 1) % (D + 1), b = (k + 2) % (D + 1);
      std::swap(sx[id].v[a], sx[id].v[b]);
    }
    for (int k = 0; k <= D; k++) {
      sx[t0].n[k] = inf_id[k];
      const int id = inf_id[k];
      for (int m = 0; m <= D; m++) {
        const int v = sx[id].v[m];
        if (v == INF) { sx[id].n[m] = t0; continue; }
        // the facet opposite finite vertex v: shared with the infinite simplex over the facet of t0 opposite v
        for (int q = 0; q <= D; q++)
          if (sx[t0].v[q] == v) sx[id].n[m] = inf_id[q];
      }
    }
    n_finite = 1;
    last = t0;
    new_sx.push_back(t0);
    fid[t0] = next_fid++;
  }

  // returns the vertex id; *is_new = false when the point already exists (value overwritten).
  // hint: a simplex to start the walk from (-1: the last one created)
  int insert(const double *p, double fv, bool *is_new, int hint = -1) {
    new_sx.clear();
    removed_sx.clear();
    *is_new = true;
    const int n = (int)nv();
    if (cur_dim < D) {
      for (int v = 0; v < n; v++)
        if (std::equal(p, p + D, pt(v))) { F[v] = fv; *is_new = false; return v; }
      X.insert(X.end(), p, p + D);
      F.push_back(fv);
      if (extends_basis(pt(n))) { basis.push_back(n); cur_dim++; }
      if (cur_dim == D) {
        // the first D + 1 affinely independent points span the first simplex; the points seen before that (affinely
        // dependent on their predecessors) follow through the normal insertion path
        start_triangulation();
        std::vector<uint8_t> in_basis((size_t)n + 1, 0);
        for (int b : basis) in_basis[b] = 1;
        for (int v = 0; v <= n; v++)
          if (!in_basis[v]) carve(locate(pt(v), last), v);
        new_sx.clear();       // everything alive and finite is new
        removed_sx.clear();
        for (int s = 0; s < (int)sx.size(); s++)
          if (alive[s] && !is_inf(s)) new_sx.push_back(s);
      }
      return n;
    }
    const int loc = locate(p, hint);
    if (!is_inf(loc))
      for (int i = 0; i <= D; i++) {
        const int v = sx[loc].v[i];
        if (std::equal(p, p + D, pt(v))) { F[v] = fv; *is_new = false; return v; }
      }
    X.insert(X.end(), p, p + D);
    F.push_back(fv);
    carve(loc, n);
    return n;
  }
};

struct QEntryD {
  double err, f;
  double x[MAXD];
  uint64_t seq;
  int s;
  uint32_t gen;
};
struct QLessD {
  bool operator()(const QEntryD &a, const QEntryD &b) const {
    if (a.err != b.err) return a.err < b.err;   // max-heap on err (ErrorTuple::operator<, src/delaunay_d_interp.h:737-739)
    return a.seq > b.seq;                       // ties: earliest pushed first (the reference leaves ties unspecified)
  }
};

double simplex_error_any(int d, int np, const double *w, double phi) {
  switch (d) {
    case 2: return dli_simplex_error<2>(np, w, phi);
    case 3: return dli_simplex_error<3>(np, w, phi);
    case 4: return dli_simplex_error<4>(np, w, phi);
    default: return 0.0;
  }
}

// Morton key of a point in the unit box (bits per axis chosen so that the key fits 63 bits)
uint64_t morton_key(int d, const double *u) {
  const int bits = 63 / d > 21 ? 21 : 63 / d;
  uint32_t q[MAXD];
  for (int c = 0; c < d; c++) {
    double t = u[c] * (double)((1u << bits) - 1);
    q[c] = t > 0.0 ? (uint32_t)std::min<double>(t, (double)((1u << bits) - 1)) : 0u;
  }
  uint64_t key = 0;
  for (int b = bits - 1; b >= 0; b--)
    for (int c = d - 1; c >= 0; c--) key = (key << 1) | ((q[c] >> b) & 1u);
  return key;
}

}  // namespace

struct dli_dd {
  HostMeshD hm;
  dli_fn fn = nullptr;
  void *user = nullptr;
  std::priority_queue<QEntryD, std::vector<QEntryD>, QLessD> pq;
  std::vector<int> nent;   // live queue entries per simplex handle
  int64_t live = 0;
  uint64_t seq = 0;
  // device side
  dli_mesh *mesh = nullptr;
  bool dirty = true;
  dli_options opt;
  bool have_opt = false;
  int64_t n_full = 0, n_patch = 0;
  // incremental device update (dli_mesh_update; the reference's dirty list is p_modified_simplices,
  // modified_cgal/Convex_hull_d.h:1172-1178, consumed at src/delaunay_d_interp.h:939-965): simplices created since the
  // last sync, vertices whose value was overwritten, vertices the device mesh already holds
  std::vector<int> pend_new, pend_values;
  int64_t nv_synced = 0;
  bool need_full = true;
  int64_t patched_since_full = 0, simplices_at_full = 0;
  // host side of small batches (dli_host_small.h): last simplex found
  int host_last = -1;

  // the dD class answers NaN outside the hull in every dimension (src/delaunay_d_interp.h:1005-1008); the 2D engine
  // projects by default (the 2D class, src/delaunay_2_interp.h:530-545)
  dli_options engine_options() const {
    dli_options o;
    if (have_opt) o = opt;
    else dli_options_init(&o);
    if (hm.D == 2) o.outside_nan = 1;
    return o;
  }
  bool entry_live(const QEntryD &e) const { return hm.alive[e.s] && hm.gen[e.s] == e.gen; }
  void push(double err, double f, const double *x, int s) {
    QEntryD e;
    e.err = err; e.f = f; e.seq = seq++; e.s = s; e.gen = hm.gen[s];
    for (int c = 0; c < MAXD; c++) e.x[c] = c < hm.D ? x[c] : 0.0;
    pq.push(e);
    if ((int)nent.size() <= s) nent.resize(hm.sx.size(), 0);
    nent[s]++;
    live++;
  }
  void drop_entries(int s) {
    if (s < (int)nent.size()) { live -= nent[s]; nent[s] = 0; }
  }
  // does the point already exist as a vertex?  (a closed simplex containing a vertex's position has it as a vertex)
  bool point_exists(const double *p, int start) {
    if (hm.cur_dim < hm.D) return false;
    const int s = hm.locate(p, start);
    if (hm.is_inf(s)) return false;
    for (int i = 0; i <= hm.D; i++)
      if (std::equal(p, p + hm.D, hm.pt(hm.sx[s].v[i]))) return true;
    return false;
  }
  // Dead entries leave the top; so do entries whose insertion point already IS a vertex.  The latter arise when the
  // ROUNDED centroid of a hull sub-simplex was inserted an ulp off the hull: the sub-simplex survives on a neighbouring
  // hull facet and keeps proposing the same point.  The reference re-inserts it (a value overwrite that changes
  // nothing) and returns the same tuple forever (insert_largest_error_point, src/delaunay_d_interp.h:1028-1034; its
  // demo never meets this: the box hull there has exactly representable face coordinates); here the tuple is dropped.
  void prune() {
    for (;;) {
      while (!pq.empty() && !entry_live(pq.top())) pq.pop();
      if (pq.empty() || !point_exists(pq.top().x, pq.top().s)) return;
      const int s = pq.top().s;
      if (s < (int)nent.size() && nent[s] > 0) nent[s]--;
      live--;
      pq.pop();
    }
  }
  // items of a simplex, in the order of the batched evaluator (include/dli.h, dli_simplex_error_*): the simplex itself,
  // then for every hull facet (ascending slot) its vertex subsets of size >= 2 by ascending bitmask
  // (set_simplex_info + add_error_of_boundary_subsimplices, src/delaunay_d_interp.h:886-916, :832-848, :810-830)
  template <class Emit>
  void for_each_item(int s, Emit emit) const {
    const int D = hm.D, k = D + 1;
    const Sx &t = hm.sx[s];
    for (int item = -1; item < k; item++) {
      int fv[MAXD + 1], nf = 0;
      if (item == -1) { for (int j = 0; j < k; j++) fv[nf++] = t.v[j]; }
      else {
        if (!hm.is_inf(t.n[item])) continue;
        for (int j = 0; j < k; j++) if (j != item) fv[nf++] = t.v[j];
      }
      const int lo = item == -1 ? (1 << nf) - 1 : 0;
      for (int mask = lo; mask < (1 << nf); mask++) {
        if (__builtin_popcount((unsigned)mask) < 2) continue;
        int sel[MAXD + 1], np = 0;
        for (int b = 0; b < nf; b++) if (mask >> b & 1) sel[np++] = fv[b];
        emit(np, sel);
      }
    }
  }
  void add_simplex_errors(int s) {
    const int D = hm.D, N = D + 1;
    for_each_item(s, [&](int np, const int *sel) {
      double w[(MAXD + 1) * (MAXD + 1)], c[MAXD];
      for (int a = 0; a < np; a++) {
        for (int q = 0; q < D; q++) w[a * N + q] = hm.pt(sel[a])[q];
        w[a * N + D] = hm.F[sel[a]];
      }
      for (int q = 0; q < D; q++) {   // get_centroid (:108-121 of the 2D header, same in dD): sequential sum / np
        double acc = w[q];
        for (int a = 1; a < np; a++) acc += w[a * N + q];
        c[q] = acc / np;
      }
      // A hull sub-simplex whose ROUNDED centroid was inserted an ulp inside the hull leaves its hull facet in place,
      // now on a sliver (facet + the new vertex), and the sliver would queue the same item again -- insertion point =
      // its own vertex, the same tuple on top of the queue forever (the reference shares this livelock: its
      // insert_largest_error_point re-inserts the rounded doubles, src/delaunay_d_interp.h:1028-1034; its own demo
      // never meets it because the box hull there has exactly representable face coordinates).  An item whose
      // insertion point already is a vertex of its simplex carries no information (the interpolant is exact there) and
      // is not queued.
      for (int a = 0; a <= D; a++)
        if (std::equal(c, c + D, hm.pt(hm.sx[s].v[a]))) return;
      const double phi = fn(D, c, user);
      push(simplex_error_any(D, np, w, phi), phi, c, s);
    });
  }
  int insert(const double *x, double f, int hint = -1) {
    bool is_new;
    const int v = hm.insert(x, f, &is_new, hint);
    dirty = true;
    if (mesh && !need_full) {
      pend_new.insert(pend_new.end(), hm.new_sx.begin(), hm.new_sx.end());
      if (!is_new) pend_values.push_back(v);
      if (!hm.fid_compact) need_full = true;
    }
    if (fn && hm.cur_dim == hm.D) {
      for (int r : hm.removed_sx) drop_entries(r);
      if (nent.size() < hm.sx.size()) nent.resize(hm.sx.size(), 0);
      if (!is_new) {
        // value change of an existing vertex: the simplices of its star are "modified" (:942-954): refresh their entries
        for (int s = 0; s < (int)hm.sx.size(); s++) {
          if (!hm.alive[s] || hm.is_inf(s)) continue;
          bool has = false;
          for (int i = 0; i <= hm.D; i++) has = has || hm.sx[s].v[i] == v;
          if (has) { drop_entries(s); hm.gen[s]++; add_simplex_errors(s); }
        }
      }
      for (int s : hm.new_sx) add_simplex_errors(s);
    }
    return v;
  }
  // compact ids of the finite simplices in slot order; any pointer may be NULL
  void flatten(double *coords, int32_t *sv, int32_t *sn, std::vector<int32_t> *handle) const {
    const int D = hm.D, k = D + 1;
    if (coords) memcpy(coords, hm.X.data(), sizeof(double) * hm.X.size());
    if (hm.cur_dim < D || !(sv || sn || handle)) return;
    // ids: the stable compact ids when they are compact (callers that own the mesh renumber first), else slot order
    std::vector<int32_t> id(hm.sx.size(), -1);
    int32_t c = 0;
    if (hm.fid_compact) { id = hm.fid; c = (int32_t)hm.n_finite; }
    else
      for (size_t s = 0; s < hm.sx.size(); s++)
        if (hm.alive[s] && !hm.is_inf((int)s)) id[s] = c++;
    if (handle) handle->assign((size_t)c, -1);
    for (size_t s = 0; s < hm.sx.size(); s++) {
      const int32_t t = id[s];
      if (t < 0) continue;
      if (handle) (*handle)[(size_t)t] = (int32_t)s;
      for (int j = 0; j < k; j++) {
        if (sv) sv[(size_t)t * k + j] = hm.sx[s].v[j];
        if (sn) sn[(size_t)t * k + j] = id[(size_t)hm.sx[s].n[j]];
      }
    }
  }
};

#define DLI_GUARD_BEGIN try {
#define DLI_GUARD_END                                      \
  } catch (const std::bad_alloc &) { return DLI_ERR_NOMEM; \
  } catch (...) { return DLI_ERR_ARG; }

// incremental path: the simplices created since the last sync plus the ring around them (their neighbour links changed)
// and the stars of overwritten vertices go to dli_mesh_update; DLI_ERR_UNSUPPORTED = the mesh cannot be patched
static int dd_patch(dli_dd *d) {
  HostMeshD &hm = d->hm;
  const int D = hm.D, k = D + 1;
  const int64_t nv = hm.nv(), nv0 = d->nv_synced, nf = hm.n_finite;
  std::vector<int> dirty;
  auto add = [&](int s) {
    if (s < 0 || !hm.alive[s] || hm.fid[s] < 0 || hm.mark[s]) return;
    hm.mark[s] = 1;
    dirty.push_back(s);
  };
  for (int s : d->pend_new) {
    if (!hm.alive[s] || hm.fid[s] < 0) continue;   // created and destroyed again before this sync
    add(s);
    for (int j = 0; j < k; j++) add(hm.sx[s].n[j]);
  }
  std::vector<int32_t> vid;
  std::vector<double> vval;
  if (!d->pend_values.empty()) {
    std::vector<uint8_t> changed((size_t)nv, 0);
    for (int v : d->pend_values) {
      if (v < nv0 && !changed[(size_t)v]) { vid.push_back(v); vval.push_back(hm.F[(size_t)v]); }
      changed[(size_t)v] = 1;
    }
    for (int s = 0; s < (int)hm.sx.size(); s++) {     // the interpolants of their stars change
      if (!hm.alive[s] || hm.fid[s] < 0) continue;
      bool has = false;
      for (int j = 0; j < k; j++) has = has || changed[(size_t)hm.sx[s].v[j]];
      if (has) add(s);
    }
  }
  for (int s : dirty) hm.mark[s] = 0;
  std::vector<int32_t> ids(dirty.size()), sv(dirty.size() * k), sn(dirty.size() * k);
  for (size_t i = 0; i < dirty.size(); i++) {
    const int s = dirty[i];
    ids[i] = hm.fid[s];
    for (int j = 0; j < k; j++) {
      sv[i * k + j] = hm.sx[s].v[j];
      sn[i * k + j] = hm.fid[hm.sx[s].n[j]];   // -1 across the hull (infinite simplices carry no id)
    }
  }
  std::vector<double> nx((size_t)(nv - nv0) * D + 1), nfv((size_t)(nv - nv0) + 1);
  for (int64_t v = nv0; v < nv; v++) {
    for (int c = 0; c < D; c++) nx[(size_t)(v - nv0) * D + c] = hm.X[(size_t)v * D + c];
    nfv[(size_t)(v - nv0)] = hm.F[(size_t)v];
  }
  d->patched_since_full += (int64_t)ids.size();
  static const bool trace = getenv("DLI_TRACE_UPDATE") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  const int rc = dli_mesh_update(d->mesh, nv, nx.data(), nfv.data(), (int64_t)vid.size(), vid.data(), vval.data(), nf,
                                 (int64_t)ids.size(), ids.data(), sv.data(), sn.data());
  if (trace) fprintf(stderr, "[dli_dd] dli_mesh_update %.3f ms, %zu dirty, %zu pending\n",
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(), ids.size(), d->pend_new.size());
  return rc;
}

static int dd_refresh(dli_dd *d) {
  if (!d->dirty && d->mesh) return DLI_OK;
  // the start-hint grid is not patched: once half of the mesh it was built for has been replaced, start over
  const bool stale_hints = d->patched_since_full > std::max<int64_t>(4096, d->simplices_at_full / 2);
  if (d->mesh && !d->need_full && !stale_hints && d->hm.fid_compact && d->hm.cur_dim == d->hm.D) {
    static const bool trace = getenv("DLI_TRACE_UPDATE") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = dd_patch(d);
    if (trace) fprintf(stderr, "[dli_dd] dd_patch %.3f ms (rc %d)\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(), rc);
    if (rc == DLI_OK) {
      d->dirty = false;
      d->n_patch++;
      d->pend_new.clear(); d->pend_values.clear();
      d->nv_synced = d->hm.nv();
      return DLI_OK;
    }
    if (rc != DLI_ERR_UNSUPPORTED) {   // a failed patch may leave the device mesh half-written: drop it
      dli_mesh_destroy(d->mesh); d->mesh = nullptr;
      return rc;
    }
  }
  if (d->mesh) { dli_mesh_destroy(d->mesh); d->mesh = nullptr; }
  if (d->hm.cur_dim == d->hm.D && !d->hm.fid_compact) d->hm.renumber();
  const HostMeshD &hm = d->hm;
  const int D = hm.D, k = D + 1;
  const int64_t nv = hm.nv(), ns = hm.cur_dim == D ? hm.n_finite : 0;
  std::vector<double> coords((size_t)nv * D + 1), fv(hm.F);
  fv.push_back(0.0);
  std::vector<int32_t> sv((size_t)ns * k + 1), sn((size_t)ns * k + 1);
  d->flatten(coords.data(), ns ? sv.data() : nullptr, ns ? sn.data() : nullptr, nullptr);
  const dli_options o = d->engine_options();
  int rc = dli_mesh_create(D, nv, coords.data(), fv.data(), ns, sv.data(), sn.data(), &o, &d->mesh);
  if (rc == DLI_OK) {
    d->dirty = false; d->n_full++;
    d->need_full = false;
    d->pend_new.clear(); d->pend_values.clear();
    d->nv_synced = nv;
    d->patched_since_full = 0;
    d->simplices_at_full = ns;
  }
  return rc;
}

// Small batches are served on the host: exact visibility walk on the triangulation itself, started at the simplex the
// previous query ended in; closed simplices; outside the hull -> NaN (src/delaunay_d_interp.h:1005-1008); epilogue in the
// reference's operation order (dli_refmath.h -- the code the exact pass of the kernels runs; D = 2 uses the 2D form,
// as the device engine does for dim-2 meshes).
template <int D>
static double dd_epilogue(const HostMeshD &hm, int s, const double *p) {
  double x[(D + 1) * D], f[D + 1];
  for (int i = 0; i <= D; i++) {
    const int v = hm.sx[s].v[i];
    for (int c = 0; c < D; c++) x[i * D + c] = hm.pt(v)[c];
    f[i] = hm.F[(size_t)v];
  }
  if (D == 2) return dli::epilogue2_ref_xy(p[0], p[1], x[0], x[1], f[0], x[2], x[3], f[1], x[4], x[5], f[2]);
  return dli::epilogueD_ref_xf<D>(p, x, f);
}
static void dd_host_interp(dli_dd *d, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto) {
  HostMeshD &hm = d->hm;
  const int D = hm.D;
  const double nan = std::nan("");
  const bool have_mesh = hm.cur_dim == D;
  // simplex ids are the stable compact ids the device mesh uses (renumbered here if an insertion left holes: the
  // device mesh then follows with a full flatten)
  if (so && have_mesh && !hm.fid_compact) { hm.renumber(); d->need_full = true; d->dirty = true; }
  int start = d->host_last;
  for (int64_t i = 0; i < n; i++) {
    const double *p = xs + i * D;
    double y = nan;
    int32_t sid = -1;
    uint8_t st = DLI_STATUS_NO_MESH;
    bool finite = true;
    for (int c = 0; c < D; c++) finite = finite && std::isfinite(p[c]);
    if (have_mesh && !finite) st = DLI_STATUS_STEP_LIMIT;
    else if (have_mesh) {
      const int s = hm.locate(p, start);
      if (hm.is_inf(s)) { st = DLI_STATUS_OUTSIDE_NAN; start = s; }
      else {
        st = DLI_STATUS_OK;
        start = s;
        y = D == 2 ? dd_epilogue<2>(hm, s, p) : (D == 3 ? dd_epilogue<3>(hm, s, p) : dd_epilogue<4>(hm, s, p));
        if (so) sid = hm.fid[(size_t)s];
      }
    }
    ys[i] = y;
    if (so) so[i] = sid;
    if (sto) sto[i] = st;
  }
  d->host_last = start;
}

extern "C" {

dli_dd *dli_dd_create(int n_dims, dli_fn fn, void *user) {
  if (n_dims < 2 || n_dims > MAXD) return nullptr;
  try {
    dli_dd *d = new dli_dd;
    d->hm.D = n_dims;
    d->fn = fn;
    d->user = user;
    return d;
  } catch (...) { return nullptr; }
}
void dli_dd_destroy(dli_dd *d) {
  if (!d) return;
  if (d->mesh) dli_mesh_destroy(d->mesh);
  delete d;
}
int dli_dd_dimension(const dli_dd *d) { return d ? d->hm.D : 0; }
int dli_dd_current_dimension(const dli_dd *d) { return d ? d->hm.cur_dim : -1; }
int64_t dli_dd_num_vertices(const dli_dd *d) { return d ? d->hm.nv() : 0; }
int64_t dli_dd_num_simplices(const dli_dd *d) { return d && d->hm.cur_dim == d->hm.D ? d->hm.n_finite : 0; }

int32_t dli_dd_insert(dli_dd *d, const double *x, double f) {
  DLI_GUARD_BEGIN
  if (!d || !x) return DLI_ERR_ARG;
  for (int c = 0; c < d->hm.D; c++)
    if (!std::isfinite(x[c])) return DLI_ERR_ARG;
  return d->insert(x, f);
  DLI_GUARD_END
}

int dli_dd_insert_many(dli_dd *d, int64_t n, const double *xs, const double *f, int spatial_sort, int use_gpu) {
  DLI_GUARD_BEGIN
  if (!d || n < 0 || (n > 0 && (!xs || !f))) return DLI_ERR_ARG;
  HostMeshD &hm = d->hm;
  const int D = hm.D;
  for (int64_t i = 0; i < n * D; i++)
    if (!std::isfinite(xs[i])) return DLI_ERR_ARG;   // validated before the first insertion: nothing to roll back
  d->need_full = true;                               // bulk insertion: the device mesh is flattened from scratch
  std::vector<int64_t> order((size_t)n);
  for (int64_t i = 0; i < n; i++) order[(size_t)i] = i;
  if (spatial_sort && n > 1) {
    // biased randomised insertion order: random rounds of doubling size, each along a Morton curve -- short walks
    // from the previous point, and no adversarial intermediate triangulations
    double lo[MAXD], hi[MAXD], sc[MAXD];
    for (int c = 0; c < D; c++) { lo[c] = INFINITY; hi[c] = -INFINITY; }
    for (int64_t i = 0; i < n; i++)
      for (int c = 0; c < D; c++) { lo[c] = std::min(lo[c], xs[i * D + c]); hi[c] = std::max(hi[c], xs[i * D + c]); }
    for (int c = 0; c < D; c++) sc[c] = hi[c] > lo[c] ? 1.0 / (hi[c] - lo[c]) : 0.0;
    std::vector<uint64_t> key((size_t)n);
    for (int64_t i = 0; i < n; i++) {
      double u[MAXD];
      for (int c = 0; c < D; c++) u[c] = (xs[i * D + c] - lo[c]) * sc[c];
      key[(size_t)i] = morton_key(D, u);
    }
    uint64_t r = 0x2545F4914F6CDD1Dull;
    for (int64_t i = n - 1; i > 0; i--) {   // Fisher-Yates with xorshift64*
      r ^= r >> 12; r ^= r << 25; r ^= r >> 27;
      const uint64_t j = (r * 0x2545F4914F6CDD1Dull) % (uint64_t)(i + 1);
      std::swap(order[(size_t)i], order[(size_t)j]);
    }
    int64_t end = n;
    while (end > 0) {   // the last half, the quarter before it, ...: each round sorted along the curve
      const int64_t begin = end > 64 ? end / 2 : 0;
      std::sort(order.begin() + begin, order.begin() + end, [&key](int64_t a, int64_t b) { return key[(size_t)a] < key[(size_t)b]; });
      end = begin;
    }
  }
  dli_fn fn = d->fn;
  d->fn = nullptr;   // triangulation first, the queue afterwards in one batch
  for (int64_t k = 0; k < n; k++) {
    const int64_t i = order[(size_t)k];
    d->insert(xs + i * D, f[i]);
  }
  d->fn = fn;
  if (!fn || hm.cur_dim < D) return DLI_OK;
  while (!d->pq.empty()) d->pq.pop();
  d->nent.assign(hm.sx.size(), 0);
  d->live = 0;
  if (!use_gpu) {
    for (int s = 0; s < (int)hm.sx.size(); s++)
      if (hm.alive[s] && !hm.is_inf(s)) d->add_simplex_errors(s);
    return DLI_OK;
  }
  // batched: flatten -> GPU centroids -> host callback -> GPU error kernel (dli_simplex_error_*)
  const int k1 = D + 1;
  const int64_t nv = hm.nv(), ns = hm.n_finite;
  std::vector<double> coords((size_t)nv * D);
  std::vector<int32_t> sv((size_t)ns * k1), sn((size_t)ns * k1), handle;
  d->flatten(coords.data(), sv.data(), sn.data(), &handle);
  if (d->mesh) { dli_mesh_destroy(d->mesh); d->mesh = nullptr; }
  const dli_options o = d->engine_options();
  int rc = dli_mesh_create(D, nv, coords.data(), hm.F.data(), ns, sv.data(), sn.data(), &o, &d->mesh);
  if (rc != DLI_OK) return rc;
  d->dirty = false;
  d->n_full++;
  const int64_t ni = dli_simplex_error_count(d->mesh);
  if (ni < 0) return (int)ni;
  std::vector<double> cent((size_t)ni * D), phi((size_t)ni), err((size_t)ni);
  std::vector<int32_t> item_simplex((size_t)ni);
  rc = dli_simplex_error_prepare(d->mesh, cent.data(), item_simplex.data(), nullptr);
  if (rc != DLI_OK) return rc;
  for (int64_t i = 0; i < ni; i++) phi[(size_t)i] = fn(D, &cent[(size_t)i * D], d->user);
  rc = dli_simplex_error_finish(d->mesh, ni, phi.data(), err.data());
  if (rc != DLI_OK) return rc;
  for (int64_t i = 0; i < ni; i++)
    d->push(err[(size_t)i], phi[(size_t)i], &cent[(size_t)i * D], handle[(size_t)item_simplex[(size_t)i]]);
  return DLI_OK;
  DLI_GUARD_END
}

int dli_dd_get_largest_error_tuple(dli_dd *d, double *err, double *f, double *x) {
  if (!d || !d->fn) return DLI_ERR_ARG;
  d->prune();
  if (d->pq.empty()) return DLI_ERR_ARG;
  const QEntryD &e = d->pq.top();
  if (err) *err = e.err;
  if (f) *f = e.f;
  if (x) for (int c = 0; c < d->hm.D; c++) x[c] = e.x[c];
  return DLI_OK;
}
int dli_dd_insert_largest_error_point(dli_dd *d, double *err, double *f, double *x) {
  DLI_GUARD_BEGIN
  if (!d || !d->fn) return DLI_ERR_ARG;
  d->prune();
  if (d->pq.empty()) return DLI_ERR_ARG;
  const QEntryD e = d->pq.top();
  if (err) *err = e.err;
  if (f) *f = e.f;
  if (x) for (int c = 0; c < d->hm.D; c++) x[c] = e.x[c];
  d->insert(e.x, e.f, e.s);   // the point is the centroid of a (sub)simplex of e.s: start the walk there
  return DLI_OK;
  DLI_GUARD_END
}
// k-at-a-time adaptive refinement in d dimensions (the dD form of dli_di2_refine, same rules): tuples
// are taken in queue order; a tuple whose simplex shares a vertex with a simplex already taken in this round stays in the
// queue (the cavities could overlap); a tuple whose simplex an earlier insertion of the round destroyed is gone with it,
// as insert() drops the entries of removed simplices; a round ends at the first tuple below rel_threshold x the round's
// largest error.  The items of every simplex created in the round -- the simplex and the vertex subsets of its hull
// facets, in the order of add_simplex_errors -- are then evaluated with ONE call of the batched callback.  In 4D an
// insertion creates ~500 simplices: that is thousands of scalar callbacks per point otherwise.
int64_t dli_dd_refine(dli_dd *d, int64_t k, double rel_threshold, dli_fn_batch fnb, void *user_b, double *err_out, double *f_out,
                      double *x_out) {
  DLI_GUARD_BEGIN
  if (!d || k < 0 || (!d->fn && !fnb)) return DLI_ERR_ARG;
  HostMeshD &hm = d->hm;
  const int D = hm.D, N = D + 1;
  if (hm.cur_dim < D) return 0;
  double top_err = -1.0;
  std::vector<QEntryD> deferred;
  std::vector<int> round_sx;
  std::vector<uint8_t> taken;            // vertices of the simplices taken in this round
  dli_fn fn = d->fn;
  int64_t done = 0;
  while (done < k) {
    d->prune();
    if (d->pq.empty()) break;
    const QEntryD e = d->pq.top();
    if (top_err < 0.0) top_err = e.err;
    else if (e.err < rel_threshold * top_err) break;
    d->pq.pop();
    if (e.s < (int)d->nent.size() && d->nent[e.s] > 0) d->nent[e.s]--;
    d->live--;
    const Sx &t = hm.sx[e.s];
    if (taken.size() < (size_t)hm.nv()) taken.resize((size_t)hm.nv() + 1, 0);
    bool clash = false;
    for (int j = 0; j <= D; j++) clash = clash || taken[(size_t)t.v[j]];
    if (clash) { deferred.push_back(e); continue; }
    for (int j = 0; j <= D; j++) taken[(size_t)t.v[j]] = 1;
    d->fn = nullptr;                     // insert without per-item callbacks; entries of removed simplices dropped below
    d->insert(e.x, e.f, e.s);
    d->fn = fn;
    for (int r : hm.removed_sx) d->drop_entries(r);
    if (d->nent.size() < hm.sx.size()) d->nent.resize(hm.sx.size(), 0);
    round_sx.insert(round_sx.end(), hm.new_sx.begin(), hm.new_sx.end());
    if (err_out) err_out[done] = e.err;
    if (f_out) f_out[done] = e.f;
    if (x_out) for (int c = 0; c < D; c++) x_out[done * D + c] = e.x[c];
    done++;
  }
  for (const QEntryD &e : deferred)      // back into the queue, unless their simplex died meanwhile
    if (d->entry_live(e)) { d->pq.push(e); d->nent[e.s]++; d->live++; }
  // ---- items of the simplices that survived the round
  std::vector<int> item_s, item_np;
  std::vector<double> w, cent;           // lifted vertices [item][N][N] (padded), centroids [item][D]
  for (int s : round_sx) {
    if (!hm.alive[s] || hm.is_inf(s) || hm.mark[s]) continue;
    hm.mark[s] = 1;
    d->for_each_item(s, [&](int np, const int *sel) {
      double tw[(MAXD + 1) * (MAXD + 1)] = {0.0}, c[MAXD];
      for (int a = 0; a < np; a++) {
        for (int q = 0; q < D; q++) tw[a * N + q] = hm.pt(sel[a])[q];
        tw[a * N + D] = hm.F[sel[a]];
      }
      for (int q = 0; q < D; q++) {
        double acc = tw[q];
        for (int a = 1; a < np; a++) acc += tw[a * N + q];
        c[q] = acc / np;
      }
      for (int a = 0; a <= D; a++)       // (an item whose insertion point is a vertex of its simplex is not queued)
        if (std::equal(c, c + D, hm.pt(hm.sx[s].v[a]))) return;
      item_s.push_back(s); item_np.push_back(np);
      w.insert(w.end(), tw, tw + (size_t)N * N);
      cent.insert(cent.end(), c, c + D);
    });
  }
  for (int s : round_sx) hm.mark[s] = 0;
  const int64_t ni = (int64_t)item_s.size();
  std::vector<double> phi((size_t)ni);
  if (fnb) { if (ni > 0) fnb(D, ni, cent.data(), phi.data(), user_b); }
  else for (int64_t i = 0; i < ni; i++) phi[(size_t)i] = fn(D, &cent[(size_t)i * D], d->user);
  for (int64_t i = 0; i < ni; i++)
    d->push(simplex_error_any(D, item_np[(size_t)i], &w[(size_t)i * N * N], phi[(size_t)i]), phi[(size_t)i], &cent[(size_t)i * D],
            item_s[(size_t)i]);
  return done;
  DLI_GUARD_END
}
int64_t dli_dd_error_queue_size(dli_dd *d) { return d ? d->live : 0; }
int64_t dli_dd_get_error_queue(dli_dd *d, int64_t cap, double *err, double *f, double *xs) {
  DLI_GUARD_BEGIN
  if (!d || cap < 0) return DLI_ERR_ARG;
  auto copy = d->pq;
  int64_t n = 0;
  const int D = d->hm.D;
  while (!copy.empty() && n < cap) {
    const QEntryD &e = copy.top();
    if (d->entry_live(e)) {
      if (err) err[n] = e.err;
      if (f) f[n] = e.f;
      if (xs) for (int c = 0; c < D; c++) xs[n * D + c] = e.x[c];
      n++;
    }
    copy.pop();
  }
  return n;
  DLI_GUARD_END
}

int dli_dd_export(const dli_dd *d, double *coords, double *f, int32_t *sv, int32_t *sn) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  if (f) memcpy(f, d->hm.F.data(), sizeof(double) * d->hm.F.size());
  d->flatten(coords, sv, sn, nullptr);
  return DLI_OK;
  DLI_GUARD_END
}

int dli_dd_set_options(dli_dd *d, const dli_options *opt) {
  if (!d) return DLI_ERR_ARG;
  d->have_opt = opt != nullptr;
  if (opt) d->opt = *opt;
  d->dirty = true;
  d->need_full = true;
  return DLI_OK;
}

int dli_dd_interp(dli_dd *d, int64_t n, const double *xs, double *ys, int32_t *simplex_out, uint8_t *status_out, void *stream) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  if (n > 0 && xs && ys && dli::host_batch_is_small(n, d->hm.cur_dim == d->hm.D ? d->hm.n_finite : 0, d->hm.D) &&
      dli::is_plain_host_pointer(xs) && dli::is_plain_host_pointer(ys) && dli::is_plain_host_pointer(simplex_out) &&
      dli::is_plain_host_pointer(status_out)) {
    dd_host_interp(d, n, xs, ys, simplex_out, status_out);
    return DLI_OK;
  }
  static const bool trace = getenv("DLI_TRACE_UPDATE") != nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  int rc = dd_refresh(d);
  if (rc != DLI_OK) return rc;
  const auto t1 = std::chrono::steady_clock::now();
  rc = dli_interp(d->mesh, n, xs, ys, simplex_out, status_out, stream);
  if (trace) fprintf(stderr, "[dli_dd] refresh %.3f ms, dli_interp call %.3f ms\n", std::chrono::duration<double, std::milli>(t1 - t0).count(),
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count());
  return rc;
  DLI_GUARD_END
}
const dli_mesh *dli_dd_device_mesh(dli_dd *d) {
  if (!d) return nullptr;
  if (dd_refresh(d) != DLI_OK) return nullptr;
  return d->mesh;
}
void dli_dd_sync_counts(const dli_dd *d, int64_t *n_full, int64_t *n_incremental) {
  if (n_full) *n_full = d ? d->n_full : 0;
  if (n_incremental) *n_incremental = d ? d->n_patch : 0;
}

int64_t dli_dd_get_all_vertices(const dli_dd *d, double *out) {
  if (!d) return 0;
  const int D = d->hm.D;
  const int64_t nv = d->hm.nv();
  if (out)
    for (int64_t v = 0; v < nv; v++) {
      for (int c = 0; c < D; c++) out[v * (D + 1) + c] = d->hm.X[(size_t)v * D + c];
      out[v * (D + 1) + D] = d->hm.F[(size_t)v];
    }
  return nv;
}

// get_line_segments (src/delaunay_d_interp.h:1103-1142): the segments of a simplex are its consecutive vertex pairs
// (i, i + 1 mod N + 1), every unordered pair once; segments [ne][2][N + 1] (coordinates then f); min / max over the end
// points ([N + 1], DBL_MAX / -DBL_MAX when there is no segment, as the reference).  Order: by (smaller id, larger id).
int64_t dli_dd_get_line_segments(const dli_dd *d, double *segments, double *min_coords, double *max_coords) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  const HostMeshD &hm = d->hm;
  const int D = hm.D, k = D + 1;
  std::vector<uint64_t> pairs;
  if (hm.cur_dim == D)
    for (int s = 0; s < (int)hm.sx.size(); s++) {
      if (!hm.alive[s] || hm.is_inf(s)) continue;
      for (int i = 0; i < k; i++) {
        const uint32_t a = (uint32_t)hm.sx[s].v[i], b = (uint32_t)hm.sx[s].v[(i + 1) % k];
        pairs.push_back(((uint64_t)std::min(a, b) << 32) | std::max(a, b));
      }
    }
  std::sort(pairs.begin(), pairs.end());
  pairs.erase(std::unique(pairs.begin(), pairs.end()), pairs.end());
  if (min_coords) for (int c = 0; c < k; c++) min_coords[c] = 1.7976931348623157e308;
  if (max_coords) for (int c = 0; c < k; c++) max_coords[c] = -1.7976931348623157e308;
  for (size_t e = 0; e < pairs.size(); e++) {
    const int ends[2] = {(int)(pairs[e] >> 32), (int)(pairs[e] & 0xffffffffu)};
    for (int t = 0; t < 2; t++)
      for (int c = 0; c < k; c++) {
        const double val = c < D ? hm.pt(ends[t])[c] : hm.F[(size_t)ends[t]];
        if (segments) segments[(e * 2 + t) * k + c] = val;
        if (min_coords) min_coords[c] = std::min(min_coords[c], val);
        if (max_coords) max_coords[c] = std::max(max_coords[c], val);
      }
  }
  return (int64_t)pairs.size();
  DLI_GUARD_END
}

// test hook: structural invariants of the triangulation (neighbour symmetry, shared facets, positive orientation, the
// invariant of the infinite simplices) and, with check_delaunay != 0, the empty-sphere property against the vertex
// opposite every interior facet (local Delaunay => global).  0 = valid, > 0 = the number of the failed check.
int dli_dd_validate(const dli_dd *dc, int check_delaunay) {
  DLI_GUARD_BEGIN
  if (!dc) return DLI_ERR_ARG;
  HostMeshD &hm = const_cast<dli_dd *>(dc)->hm;
  const int D = hm.D;
  if (hm.cur_dim < D) return 0;
  int64_t finite = 0;
  for (int s = 0; s < (int)hm.sx.size(); s++) {
    if (!hm.alive[s]) continue;
    const Sx &t = hm.sx[s];
    const int j = hm.inf_slot(s);
    if (j < 0) { finite++; if (hm.orient_replaced(s, -1, nullptr) <= 0) return 1; }
    for (int k = 0; k <= D; k++) {
      const int nb = t.n[k];
      if (nb < 0 || nb >= (int)hm.sx.size() || !hm.alive[nb]) return 2;
      const Sx &u = hm.sx[nb];
      int back = -1;
      for (int i = 0; i <= D; i++) if (u.n[i] == s) back = i;
      if (back < 0) return 3;
      // the shared facet: every vertex of s except slot k appears in nb, and not in slot `back`
      for (int i = 0; i <= D; i++) {
        if (i == k) continue;
        bool found = false;
        for (int q = 0; q <= D; q++) found = found || (q != back && u.v[q] == t.v[i]);
        if (!found) return 4;
      }
      if (j < 0 && hm.inf_slot(nb) < 0) {
        // opposite sides of the shared facet
        if (hm.orient_replaced(s, k, hm.pt(u.v[back])) >= 0) return 5;
        if (check_delaunay && hm.insphere(s, hm.pt(u.v[back])) > 0) return 6;
      }
      if (j >= 0 && k == j) {
        // infinite simplex: the vertex of the finite neighbour opposite the hull facet is strictly behind it
        if (hm.inf_slot(nb) >= 0) return 7;
        if (hm.orient_replaced(s, j, hm.pt(u.v[back])) >= 0) return 8;
      }
    }
  }
  if (finite != hm.n_finite) return 9;
  return 0;
  DLI_GUARD_END
}

}  // extern "C"
