// dli_hostmesh.cpp -- host-side incremental 2D Delaunay triangulation (Bowyer-Watson with an
// infinite vertex, exact predicates, CGAL-compatible symbolic perturbation) and the adaptive
// insertion queue of Delaunay_incremental_interp_2; C ABI in include/dli_host.h.
//
// This is where the reference keeps CGAL::Delaunay_triangulation_2 (src/delaunay_2_interp.h:66-95,
// insert :434-448) and its boost skew_heap queue (:397-400, :455-508).  Construction is host work by
// design (north_star); the GPU engine consumes the flattened result (dli_di2_export -> dli_mesh_create).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <queue>
#include <unordered_map>
#include <unordered_set>
#include <new>
#include <vector>

#include "../../include/dli_error_math.h"
#include "../../include/dli_host.h"
#include "dli_exact.cuh"
#include "dli_host_small.h"
#include "dli_refmath.h"
#include <atomic>

namespace {

using dli::grow_ze;
using dli::two_prod;

// ---------------------------------------------------------------------------------------------
// exact in-circle
// ---------------------------------------------------------------------------------------------
typedef std::vector<double> Exp;

void exp_add(Exp &e, double b) {
  if (b == 0.0 && !e.empty()) return;
  e.push_back(0.0);
  int n = (int)e.size() - 1;
  grow_ze(e.data(), n, b);
  e.resize(n);
}
// exact orient2d determinant of raw coordinates as an expansion
Exp orient_exp(double ax, double ay, double bx, double by, double cx, double cy) {
  Exp e;
  const double pa[6] = {ax, -ax, -cx, -ay, ay, cy};
  const double pb[6] = {by, cy, by, bx, cx, bx};
  for (int i = 0; i < 6; i++) {
    double p, q;
    two_prod(pa[i], pb[i], p, q);
    exp_add(e, q);
    exp_add(e, p);
  }
  return e;
}
// acc += sgn * (x*x + y*y) * m   (all exact)
void add_lift_times(Exp &acc, double x, double y, const Exp &m, double sgn) {
  double l[4];
  two_prod(x, x, l[1], l[0]);
  two_prod(y, y, l[3], l[2]);
  for (int i = 0; i < 4; i++) {
    if (l[i] == 0.0) continue;
    for (double mj : m) {
      double p, q;
      two_prod(sgn * l[i], mj, p, q);
      exp_add(acc, q);
      exp_add(acc, p);
    }
  }
}
int incircle_exact(double ax, double ay, double bx, double by, double cx, double cy, double dx, double dy) {
  // det4 = +lift_a o(b,c,d) - lift_b o(a,c,d) + lift_c o(a,b,d) - lift_d o(a,b,c)
  Exp acc;
  add_lift_times(acc, ax, ay, orient_exp(bx, by, cx, cy, dx, dy), 1.0);
  add_lift_times(acc, bx, by, orient_exp(ax, ay, cx, cy, dx, dy), -1.0);
  add_lift_times(acc, cx, cy, orient_exp(ax, ay, bx, by, dx, dy), 1.0);
  add_lift_times(acc, dx, dy, orient_exp(ax, ay, bx, by, cx, cy), -1.0);
  return dli::expansion_sign(acc.data(), (int)acc.size());
}
// > 0: d strictly inside the circle through a, b, c (ccw); Shewchuk's static filter then exact
int incircle(double ax, double ay, double bx, double by, double cx, double cy, double dx, double dy) {
  double adx = ax - dx, ady = ay - dy, bdx = bx - dx, bdy = by - dy, cdx = cx - dx, cdy = cy - dy;
  double bdxcdy = bdx * cdy, cdxbdy = cdx * bdy, alift = adx * adx + ady * ady;
  double cdxady = cdx * ady, adxcdy = adx * cdy, blift = bdx * bdx + bdy * bdy;
  double adxbdy = adx * bdy, bdxady = bdx * ady, clift = cdx * cdx + cdy * cdy;
  double det = alift * (bdxcdy - cdxbdy) + blift * (cdxady - adxcdy) + clift * (adxbdy - bdxady);
  double perm = (std::fabs(bdxcdy) + std::fabs(cdxbdy)) * alift + (std::fabs(cdxady) + std::fabs(adxcdy)) * blift +
                (std::fabs(adxbdy) + std::fabs(bdxady)) * clift;
  double bound = (10.0 + 96.0 * dli::kEps) * dli::kEps * perm;
  if (det > bound) return 1;
  if (-det > bound) return -1;
  return incircle_exact(ax, ay, bx, by, cx, cy, dx, dy);
}

inline int orient(double ax, double ay, double bx, double by, double cx, double cy) {
  return dli::orient2d(ax, ay, bx, by, cx, cy);
}

// ---------------------------------------------------------------------------------------------
// triangulation
// ---------------------------------------------------------------------------------------------
constexpr int INF = -1;  // the infinite vertex

struct Face {
  int v[3];
  int n[3];
};

struct HostMesh2 {
  std::vector<double> X, Y, F;
  std::vector<Face> fc;
  std::vector<uint8_t> alive;
  std::vector<uint32_t> gen;
  std::vector<int> freelist;
  int dim = -1;
  int last = -1;  // some alive finite face
  int64_t n_finite = 0;
  std::vector<int> new_faces, removed_faces;  // finite faces touched by the last insertion
  std::vector<int> ring_faces;                // finite faces around the cavity (one neighbour link changed)
  // Stable compact ids of the finite faces (the ids the device mesh and export() use): an insertion hands the ids of
  // the finite faces it removes to the faces it creates and appends the rest -- the number of finite faces never
  // decreases (+2 inside, +1 on a hull edge, +visible edges outside), so the id space stays 0 .. n_finite-1 and an
  // insertion changes only the ids of its own star (incremental device update, dli_mesh_update).
  std::vector<int> fid;       // face slot -> id (-1: infinite or dead)
  std::vector<int> fid_free;  // ids released by the current insertion
  bool fid_compact = true;    // false: ids must be reassigned (renumber) before the next export
  // scratch
  std::vector<int> cavity, stack;
  std::vector<uint8_t> mark;

  bool is_inf(int f) const { return fc[f].v[0] == INF || fc[f].v[1] == INF || fc[f].v[2] == INF; }
  int new_face() {
    int id;
    if (!freelist.empty()) { id = freelist.back(); freelist.pop_back(); }
    else { id = (int)fc.size(); fc.push_back(Face()); alive.push_back(0); gen.push_back(0); mark.push_back(0); fid.push_back(-1); }
    alive[id] = 1;
    fid[id] = -1;
    return id;
  }
  // ids for the finite faces created by an insertion (after the removed ones released theirs)
  void assign_ids(int64_t finite_before) {
    int64_t next_id = finite_before;  // ids in use before this insertion: 0 .. finite_before-1
    for (int f : new_faces) {
      if (!fid_free.empty()) { fid[f] = fid_free.back(); fid_free.pop_back(); }
      else fid[f] = (int)next_id++;
    }
    if (!fid_free.empty()) { fid_free.clear(); fid_compact = false; }  // cannot happen for a Delaunay insertion
  }
  void renumber() {
    int c = 0;
    for (int f = 0; f < (int)fc.size(); f++) fid[f] = (alive[f] && !is_inf(f)) ? c++ : -1;
    fid_compact = true;
  }
  int o3(int a, int b, double px, double py) const { return orient(X[a], Y[a], X[b], Y[b], px, py); }
  bool lex_less(int a, int b) const { return X[a] < X[b] || (X[a] == X[b] && Y[a] < Y[b]); }

  // CGAL Delaunay_triangulation_2::side_of_oriented_circle(p0,p1,p2,p,perturb = true)
  int side_of_circle_perturbed(int p0, int p1, int p2, int p) const {
    int s = incircle(X[p0], Y[p0], X[p1], Y[p1], X[p2], Y[p2], X[p], Y[p]);
    if (s != 0) return s;
    int pts[4] = {p0, p1, p2, p};
    std::sort(pts, pts + 4, [this](int a, int b) { return lex_less(a, b); });
    for (int i = 3; i > 0; --i) {
      if (pts[i] == p) return -1;
      int o;
      if (pts[i] == p2 && (o = orient(X[p0], Y[p0], X[p1], Y[p1], X[p], Y[p])) != 0) return o;
      if (pts[i] == p1 && (o = orient(X[p0], Y[p0], X[p], Y[p], X[p2], Y[p2])) != 0) return o;
      if (pts[i] == p0 && (o = orient(X[p], Y[p], X[p1], Y[p1], X[p2], Y[p2])) != 0) return o;
    }
    return -1;
  }

  // is vertex p (already appended to X/Y) in conflict with face f?
  bool conflict(int f, int p) const {
    const Face &t = fc[f];
    for (int i = 0; i < 3; i++)
      if (t.v[i] == INF) {
        int a = t.v[(i + 1) % 3], b = t.v[(i + 2) % 3];
        int o = o3(a, b, X[p], Y[p]);
        if (o > 0) return true;
        if (o < 0) return false;
        // collinear with the hull edge: in conflict iff strictly between its end points
        double ax = X[a], ay = Y[a], bx = X[b], by = Y[b], px = X[p], py = Y[p];
        if (ax != bx) return (ax < px && px < bx) || (bx < px && px < ax);
        return (ay < py && py < by) || (by < py && py < ay);
      }
    return side_of_circle_perturbed(t.v[0], t.v[1], t.v[2], p) > 0;
  }

  // exact visibility walk from `last`; returns a finite face whose closed hull contains p, or the
  // infinite face p is strictly visible from
  int locate(double px, double py, int start = -1) const {
    int f = (start >= 0 && start < (int)fc.size() && alive[start]) ? start : last, from = -1;
    if (is_inf(f))
      for (int i = 0; i < 3; i++)
        if (fc[f].v[i] == INF) { f = fc[f].n[i]; break; }
    for (;;) {
      const Face &t = fc[f];
      int moved = -1;
      for (int j = 0; j < 3; j++) {
        if (t.n[j] == from) continue;
        int a = t.v[(j + 1) % 3], b = t.v[(j + 2) % 3];
        if (o3(a, b, px, py) < 0) { moved = j; break; }
      }
      if (moved < 0) return f;
      int nb = t.n[moved];
      if (is_inf(nb)) return nb;
      from = f;
      f = nb;
    }
  }

  void link_all() {  // (re)build every neighbour link from directed edges; creates the infinite faces
    std::unordered_map<uint64_t, std::pair<int, int>> emap;
    auto key = [](int u, int v) { return ((uint64_t)(uint32_t)(u + 1) << 32) | (uint32_t)(v + 1); };
    int nf = (int)fc.size();
    for (int f = 0; f < nf; f++)
      if (alive[f])
        for (int j = 0; j < 3; j++) emap[key(fc[f].v[(j + 1) % 3], fc[f].v[(j + 2) % 3])] = {f, j};
    std::vector<std::pair<int, int>> hull;
    for (int f = 0; f < nf; f++)
      if (alive[f])
        for (int j = 0; j < 3; j++) {
          int u = fc[f].v[(j + 1) % 3], v = fc[f].v[(j + 2) % 3];
          if (!emap.count(key(v, u))) hull.push_back({v, u});
        }
    for (auto &e : hull) {
      int id = new_face();
      fc[id].v[0] = INF; fc[id].v[1] = e.first; fc[id].v[2] = e.second;
      for (int j = 0; j < 3; j++) emap[key(fc[id].v[(j + 1) % 3], fc[id].v[(j + 2) % 3])] = {id, j};
    }
    for (int f = 0; f < (int)fc.size(); f++)
      if (alive[f])
        for (int j = 0; j < 3; j++) {
          auto it = emap.find(key(fc[f].v[(j + 2) % 3], fc[f].v[(j + 1) % 3]));
          fc[f].n[j] = it == emap.end() ? -1 : it->second.first;
        }
  }

  // returns vertex id; *is_new = false when the point already exists (value overwritten)
  int insert(double px, double py, double fv, bool *is_new) {
    new_faces.clear();
    removed_faces.clear();
    ring_faces.clear();
    *is_new = true;
    const int64_t finite_before = n_finite;
    int nv = (int)X.size();
    if (dim < 2) {
      for (int v = 0; v < nv; v++)
        if (X[v] == px && Y[v] == py) { F[v] = fv; *is_new = false; return v; }
      X.push_back(px); Y.push_back(py); F.push_back(fv);
      int p = nv;
      if (dim == -1) { dim = 0; return p; }
      if (dim == 0) { dim = 1; return p; }
      // dim == 1: still collinear?
      int a = 0, b = 1;
      if (orient(X[a], Y[a], X[b], Y[b], px, py) == 0) return p;
      // first non-collinear point: fan over the sorted collinear points
      std::vector<int> s(nv);
      for (int i = 0; i < nv; i++) s[i] = i;
      std::sort(s.begin(), s.end(), [this](int u, int v) { return lex_less(u, v); });
      for (int i = 0; i + 1 < nv; i++) {
        int id = new_face();
        bool ccw = orient(X[s[i]], Y[s[i]], X[s[i + 1]], Y[s[i + 1]], px, py) > 0;
        fc[id].v[0] = ccw ? s[i] : s[i + 1];
        fc[id].v[1] = ccw ? s[i + 1] : s[i];
        fc[id].v[2] = p;
        new_faces.push_back(id);
        n_finite++;
      }
      link_all();
      last = new_faces[0];
      dim = 2;
      assign_ids(finite_before);
      return p;
    }
    int loc = locate(px, py);
    if (!is_inf(loc)) {
      for (int i = 0; i < 3; i++) {
        int v = fc[loc].v[i];
        if (X[v] == px && Y[v] == py) { F[v] = fv; *is_new = false; return v; }
      }
    }
    X.push_back(px); Y.push_back(py); F.push_back(fv);
    int p = nv;
    // cavity = connected set of faces in conflict with p
    cavity.clear();
    stack.clear();
    stack.push_back(loc);
    mark[loc] = 1;
    while (!stack.empty()) {
      int f = stack.back();
      stack.pop_back();
      cavity.push_back(f);
      for (int j = 0; j < 3; j++) {
        int nb = fc[f].n[j];
        if (!mark[nb] && conflict(nb, p)) { mark[nb] = 1; stack.push_back(nb); }
      }
    }
    // boundary edges -> new faces (p, a, b)
    struct BEdge { int a, b, out; };
    std::vector<BEdge> be;
    for (int f : cavity)
      for (int j = 0; j < 3; j++) {
        int nb = fc[f].n[j];
        if (!mark[nb]) be.push_back({fc[f].v[(j + 1) % 3], fc[f].v[(j + 2) % 3], nb});
      }
    std::vector<int> created(be.size());
    std::unordered_map<int, int> start_at;
    start_at.reserve(be.size() * 2);
    std::vector<int> fresh;  // allocate without reusing the cavity's own slots yet
    for (size_t i = 0; i < be.size(); i++) {
      int id = new_face();
      created[i] = id;
      fc[id].v[0] = p; fc[id].v[1] = be[i].a; fc[id].v[2] = be[i].b;
      fc[id].n[0] = be[i].out;
      Face &o = fc[be[i].out];
      for (int k = 0; k < 3; k++)
        if (o.v[(k + 1) % 3] == be[i].b && o.v[(k + 2) % 3] == be[i].a) o.n[k] = id;
      if (!is_inf(be[i].out)) ring_faces.push_back(be[i].out);
      start_at[be[i].a] = id;
    }
    for (size_t i = 0; i < be.size(); i++) {
      int id = created[i];
      int nxt = start_at[be[i].b];  // the new face whose edge starts where this one ends
      fc[id].n[1] = nxt;            // slot 1 = edge b -> p
      fc[nxt].n[2] = id;            // slot 2 of (p, b, c) = edge p -> b
    }
    for (int f : cavity) {
      mark[f] = 0;
      alive[f] = 0;
      gen[f]++;
      if (!is_inf(f)) { removed_faces.push_back(f); n_finite--; fid_free.push_back(fid[f]); }
      fid[f] = -1;
      freelist.push_back(f);
    }
    for (int id : created)
      if (!is_inf(id)) { new_faces.push_back(id); n_finite++; }
    assign_ids(finite_before);
    last = new_faces.empty() ? last : new_faces[0];
    return p;
  }
};

// Hilbert index of a point in the unit square (16 bits per axis)
uint32_t hilbert16(uint32_t x, uint32_t y) {
  uint32_t d = 0;
  for (uint32_t s = 1u << 15; s > 0; s >>= 1) {
    uint32_t rx = (x & s) ? 1 : 0, ry = (y & s) ? 1 : 0;
    d += s * s * ((3 * rx) ^ ry);
    if (ry == 0) {
      if (rx == 1) { x = 65535 - x; y = 65535 - y; }
      uint32_t t = x; x = y; y = t;
    }
  }
  return d;
}

struct QEntry {
  double err, f, x, y;
  uint64_t seq;
  int face;
  uint32_t gen;
};
struct QLess {
  bool operator()(const QEntry &a, const QEntry &b) const {
    if (a.err != b.err) return a.err < b.err;  // max-heap on err (ErrorTuple::operator<, :392-394)
    return a.seq > b.seq;                      // ties: earliest pushed first (the reference leaves ties unspecified)
  }
};

}  // namespace

struct dli_di2 {
  HostMesh2 hm;
  dli_fn fn = nullptr;
  void *user = nullptr;
  std::priority_queue<QEntry, std::vector<QEntry>, QLess> pq;
  std::vector<int> nent;  // live queue entries per face handle
  int64_t live = 0;
  uint64_t seq = 0;
  // device side
  dli_mesh *mesh = nullptr;
  bool dirty = true;
  dli_options opt;
  bool have_opt = false;
  // incremental device update (the reference's dirty set is the star returned by
  // insert_and_get_modifications, src/delaunay_2_interp.h:434-448): face slots and vertices changed since the device
  // mesh was last in step with the triangulation
  bool need_full = true;            // the next refresh re-creates the device mesh
  std::vector<int> pend_faces;      // face slots created or re-linked since the last refresh (may hold dead slots)
  std::vector<int> pend_values;     // existing vertices whose value was overwritten
  int64_t nv_synced = 0;            // vertices the device mesh holds
  int64_t n_full = 0, n_incremental = 0;
  int64_t patched_since_full = 0, faces_at_full = 0;   // re-flatten once half the mesh has been patched (hint decay)
  std::atomic<int> host_last{-1};   // face the last host-side query ended in (start of the next walk; a hint only)

  // flattened triangulation under the stable face ids (HostMesh2::fid); any pointer may be NULL
  void flatten(double *coords, int32_t *tri, int32_t *nbr, int32_t *handle) {
    if (!hm.fid_compact) { hm.renumber(); need_full = true; }
    int64_t nv = (int64_t)hm.X.size();
    if (coords) for (int64_t v = 0; v < nv; v++) { coords[2 * v] = hm.X[v]; coords[2 * v + 1] = hm.Y[v]; }
    if (!(tri || nbr || handle) || hm.dim < 2) return;
    for (int fi = 0; fi < (int)hm.fc.size(); fi++) {
      int t = hm.alive[fi] ? hm.fid[fi] : -1;
      if (t < 0) continue;
      if (handle) handle[t] = fi;
      for (int j = 0; j < 3; j++) {
        if (tri) tri[3 * t + j] = hm.fc[fi].v[j];
        if (nbr) nbr[3 * t + j] = hm.fid[hm.fc[fi].n[j]];   // -1 for an infinite neighbour
      }
    }
  }

  void synced(bool full) {
    dirty = false; need_full = false;
    pend_faces.clear(); pend_values.clear();
    nv_synced = (int64_t)hm.X.size();
    (full ? n_full : n_incremental)++;
    if (full) { patched_since_full = 0; faces_at_full = hm.n_finite; }
  }
  bool entry_live(const QEntry &e) const { return hm.alive[e.face] && hm.gen[e.face] == e.gen; }
  void push(double err, double f, double x, double y, int face) {
    pq.push(QEntry{err, f, x, y, seq++, face, hm.gen[face]});
    if ((int)nent.size() <= face) nent.resize(hm.fc.size(), 0);
    nent[face]++;
    live++;
  }
  void drop_face_entries(int face) {
    if (face < (int)nent.size()) { live -= nent[face]; nent[face] = 0; }
  }
  // Dead entries leave the top; so do entries whose insertion point already IS a vertex: the ROUNDED midpoint of a
  // hull edge that is not axis-aligned lands an ulp off the hull when it is inserted, the edge survives (on a sliver, or
  // on the hull face the point was not beyond) and keeps proposing the same point.  The reference re-inserts it -- a
  // value overwrite that changes nothing -- and returns the same tuple forever (src/delaunay_2_interp.h:498-504); its
  // README run never meets this (box hull, exactly representable coordinates: every call there adds a vertex, which
  // this rule leaves untouched).  Here such a tuple is dropped.
  bool point_exists(double x, double y, int start) const {
    if (hm.dim < 2) return false;
    const int f = hm.locate(x, y, start);
    if (hm.is_inf(f)) return false;
    for (int i = 0; i < 3; i++)
      if (hm.X[hm.fc[f].v[i]] == x && hm.Y[hm.fc[f].v[i]] == y) return true;
    return false;
  }
  void prune() {
    for (;;) {
      while (!pq.empty() && !entry_live(pq.top())) pq.pop();
      if (pq.empty() || !point_exists(pq.top().x, pq.top().y, pq.top().face)) return;
      const int face = pq.top().face;
      if (face < (int)nent.size() && nent[face] > 0) nent[face]--;
      live--;
      pq.pop();
    }
  }
  // compute_simplex_error for the face and its hull edges (src/delaunay_2_interp.h:476-491)
  void add_face_errors(int face) {
    const Face &t = hm.fc[face];
    double w[9];
    for (int i = 0; i < 3; i++) { w[i * 3] = hm.X[t.v[i]]; w[i * 3 + 1] = hm.Y[t.v[i]]; w[i * 3 + 2] = hm.F[t.v[i]]; }
    double c[2] = {(w[0] + w[3] + w[6]) / 3, (w[1] + w[4] + w[7]) / 3};
    double phi = fn(2, c, user);
    push(dli_simplex_error<2>(3, w, phi), phi, c[0], c[1], face);
    for (int j = 0; j < 3; j++)
      if (hm.is_inf(t.n[j])) {  // hull edge = the two vertices other than j, in slot order
        double e[6];
        int k = 0;
        for (int i = 0; i < 3; i++)
          if (i != j) { e[k * 3] = w[i * 3]; e[k * 3 + 1] = w[i * 3 + 1]; e[k * 3 + 2] = w[i * 3 + 2]; k++; }
        double ce[2] = {(e[0] + e[3]) / 2, (e[1] + e[4]) / 2};
        double ph = fn(2, ce, user);
        push(dli_simplex_error<2>(2, e, ph), ph, ce[0], ce[1], face);
      }
  }
  int insert(double x, double y, double f) {
    bool is_new;
    const bool had_mesh = hm.dim == 2;
    int v = hm.insert(x, y, f, &is_new);
    dirty = true;
    if (!had_mesh) need_full = true;
    if (!need_full) {
      if (!is_new) {
        pend_values.push_back(v);
        for (int fi = 0; fi < (int)hm.fc.size(); fi++)   // the star's interpolants change with the value
          if (hm.alive[fi] && !hm.is_inf(fi) && (hm.fc[fi].v[0] == v || hm.fc[fi].v[1] == v || hm.fc[fi].v[2] == v))
            pend_faces.push_back(fi);
      }
      pend_faces.insert(pend_faces.end(), hm.new_faces.begin(), hm.new_faces.end());
      pend_faces.insert(pend_faces.end(), hm.ring_faces.begin(), hm.ring_faces.end());
      if ((int64_t)pend_faces.size() > hm.n_finite) need_full = true;   // cheaper to start over
    }
    if (fn) {
      for (int r : hm.removed_faces) drop_face_entries(r);
      if (nent.size() < hm.fc.size()) nent.resize(hm.fc.size(), 0);
      if (!is_new && hm.dim == 2) {
        // value change of an existing vertex: its star is "modified" (:438-446): refresh those entries
        for (int fi = 0; fi < (int)hm.fc.size(); fi++)
          if (hm.alive[fi] && !hm.is_inf(fi) && (hm.fc[fi].v[0] == v || hm.fc[fi].v[1] == v || hm.fc[fi].v[2] == v)) {
            drop_face_entries(fi);
            hm.gen[fi]++;
            add_face_errors(fi);
          }
      }
      for (int nf : hm.new_faces) add_face_errors(nf);
    }
    return v;
  }
};


// "no exception ever crosses the ABI" (include/dli.h): host-side entry points that allocate run inside this guard
#define DLI_GUARD_BEGIN try {
#define DLI_GUARD_END                                   \
  } catch (const std::bad_alloc &) { return DLI_ERR_NOMEM; \
  } catch (...) { return DLI_ERR_ARG; }
extern "C" {

dli_di2 *dli_di2_create(dli_fn fn, void *user) {
  dli_di2 *d = new dli_di2;
  d->fn = fn;
  d->user = user;
  return d;
}
void dli_di2_destroy(dli_di2 *d) {
  if (!d) return;
  if (d->mesh) dli_mesh_destroy(d->mesh);
  delete d;
}
int32_t dli_di2_insert(dli_di2 *d, double x, double y, double f) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  if (!(std::isfinite(x) && std::isfinite(y))) return DLI_ERR_ARG;
  return d->insert(x, y, f);
  DLI_GUARD_END
}

int dli_di2_insert_many(dli_di2 *d, int64_t n, const double *xy, const double *f, int spatial_sort, int use_gpu) {
  DLI_GUARD_BEGIN
  if (!d || n < 0 || (n > 0 && (!xy || !f))) return DLI_ERR_ARG;
  // every coordinate is checked BEFORE the first insertion: a rejected call leaves the triangulation and its queue untouched
  for (int64_t i = 0; i < 2 * n; i++)
    if (!std::isfinite(xy[i])) return DLI_ERR_ARG;
  d->need_full = true;
  std::vector<int64_t> order((size_t)n);
  for (int64_t i = 0; i < n; i++) order[i] = i;
  if (spatial_sort && n > 1) {
    double lo[2] = {INFINITY, INFINITY}, hi[2] = {-INFINITY, -INFINITY};
    for (int64_t i = 0; i < n; i++)
      for (int k = 0; k < 2; k++) { lo[k] = std::min(lo[k], xy[2 * i + k]); hi[k] = std::max(hi[k], xy[2 * i + k]); }
    double sx = hi[0] > lo[0] ? 65535.0 / (hi[0] - lo[0]) : 0.0, sy = hi[1] > lo[1] ? 65535.0 / (hi[1] - lo[1]) : 0.0;
    std::vector<uint32_t> key((size_t)n);
    for (int64_t i = 0; i < n; i++)
      key[i] = hilbert16((uint32_t)((xy[2 * i] - lo[0]) * sx), (uint32_t)((xy[2 * i + 1] - lo[1]) * sy));
    std::sort(order.begin(), order.end(), [&key](int64_t a, int64_t b) { return key[a] < key[b]; });
  }
  dli_fn fn = d->fn;
  d->fn = nullptr;  // build the triangulation first, then the queue in one batch
  for (int64_t k = 0; k < n; k++) {
    int64_t i = order[k];
    d->insert(xy[2 * i], xy[2 * i + 1], f[i]);
  }
  d->fn = fn;
  if (!fn || d->hm.dim < 2) return DLI_OK;
  // rebuild the whole queue
  while (!d->pq.empty()) d->pq.pop();
  d->nent.assign(d->hm.fc.size(), 0);
  d->live = 0;
  HostMesh2 &hm = d->hm;
  if (!use_gpu) {
    for (int fi = 0; fi < (int)hm.fc.size(); fi++)
      if (hm.alive[fi] && !hm.is_inf(fi)) d->add_face_errors(fi);
    return DLI_OK;
  }
  // batched: flatten -> GPU centroids -> host callback -> GPU error kernel
  int64_t nv = (int64_t)hm.X.size(), nf = hm.n_finite;
  std::vector<double> coords((size_t)nv * 2);
  std::vector<int32_t> tri((size_t)nf * 3), nbr((size_t)nf * 3), handle((size_t)nf);
  d->flatten(coords.data(), tri.data(), nbr.data(), handle.data());
  if (d->mesh) { dli_mesh_destroy(d->mesh); d->mesh = nullptr; }
  int rc = dli_mesh_create(2, nv, coords.data(), hm.F.data(), nf, tri.data(), nbr.data(), d->have_opt ? &d->opt : nullptr,
                           &d->mesh);
  if (rc != DLI_OK) return rc;
  d->synced(true);
  int64_t ni = dli_simplex_error_count(d->mesh);
  if (ni < 0) return (int)ni;
  std::vector<double> cent((size_t)ni * 2), phi((size_t)ni), err((size_t)ni);
  std::vector<int32_t> item_simplex((size_t)ni);
  rc = dli_simplex_error_prepare(d->mesh, cent.data(), item_simplex.data(), nullptr);
  if (rc != DLI_OK) return rc;
  for (int64_t i = 0; i < ni; i++) phi[i] = fn(2, &cent[2 * i], d->user);
  rc = dli_simplex_error_finish(d->mesh, ni, phi.data(), err.data());
  if (rc != DLI_OK) return rc;
  for (int64_t i = 0; i < ni; i++) d->push(err[i], phi[i], cent[2 * i], cent[2 * i + 1], handle[item_simplex[i]]);
  return DLI_OK;
  DLI_GUARD_END
}

int dli_di2_get_largest_error_tuple(dli_di2 *d, double *err, double *f, double *xy) {
  if (!d || !d->fn) return DLI_ERR_ARG;
  d->prune();
  if (d->pq.empty()) return DLI_ERR_ARG;
  const QEntry &e = d->pq.top();
  if (err) *err = e.err;
  if (f) *f = e.f;
  if (xy) { xy[0] = e.x; xy[1] = e.y; }
  return DLI_OK;
}
int dli_di2_insert_largest_error_point(dli_di2 *d, double *err, double *f, double *xy) {
  DLI_GUARD_BEGIN
  if (!d || !d->fn) return DLI_ERR_ARG;
  d->prune();
  if (d->pq.empty()) return DLI_ERR_ARG;
  QEntry e = d->pq.top();
  if (err) *err = e.err;
  if (f) *f = e.f;
  if (xy) { xy[0] = e.x; xy[1] = e.y; }
  d->insert(e.x, e.y, e.f);
  return DLI_OK;
  DLI_GUARD_END
}
// k-at-a-time adaptive refinement.  Sequential semantics are kept wherever the k insertions do not
// interact: tuples are taken in queue order; a tuple whose face shares a vertex with a face already taken in this
// round stays in the queue (its star could overlap the earlier one); a tuple whose face a previous insertion of the
// round destroyed is dropped with its face, exactly as insert() drops the entries of removed faces (:459-474).  The
// errors of every face created in the round are then evaluated in ONE batch: centroids of the faces and of their
// hull edges -> one call of the batched callback -> the closed form of include/dli_error_math.h (the same code
// k_error_finish runs on the device; a round holds a few items per inserted point, so it is evaluated on the host).
int64_t dli_di2_refine(dli_di2 *d, int64_t k, double rel_threshold, dli_fn_batch fnb, void *user_b, double *err_out, double *f_out,
                       double *xy_out) {
  DLI_GUARD_BEGIN
  if (!d || k < 0 || (!d->fn && !fnb)) return DLI_ERR_ARG;
  double top_err = -1.0;
  HostMesh2 &hm = d->hm;
  std::vector<QEntry> deferred;
  std::vector<int> round_faces;          // faces created in this round (may die again within the round)
  std::unordered_set<int> taken;         // vertices of the faces taken in this round
  dli_fn fn = d->fn;
  int64_t done = 0;
  while (done < k) {
    d->prune();
    if (d->pq.empty()) break;
    QEntry e = d->pq.top();
    if (top_err < 0.0) top_err = e.err;
    else if (e.err < rel_threshold * top_err) break;   // the rest of the queue is not worth a point yet
    d->pq.pop();
    d->nent[e.face]--; d->live--;        // (entry is live: prune() left it on top)
    const Face &t = hm.fc[e.face];
    if (taken.count(t.v[0]) || taken.count(t.v[1]) || taken.count(t.v[2])) { deferred.push_back(e); continue; }
    for (int j = 0; j < 3; j++) taken.insert(t.v[j]);
    d->fn = nullptr;                     // insert without per-face callbacks; the queue entries of removed faces are
    d->insert(e.x, e.y, e.f);            // dropped below
    d->fn = fn;
    for (int r : hm.removed_faces) d->drop_face_entries(r);
    if (d->nent.size() < hm.fc.size()) d->nent.resize(hm.fc.size(), 0);
    round_faces.insert(round_faces.end(), hm.new_faces.begin(), hm.new_faces.end());
    if (err_out) err_out[done] = e.err;
    if (f_out) f_out[done] = e.f;
    if (xy_out) { xy_out[2 * done] = e.x; xy_out[2 * done + 1] = e.y; }
    done++;
  }
  for (const QEntry &e : deferred)       // back into the queue, unless their face died meanwhile
    if (d->entry_live(e)) { d->pq.push(e); d->nent[e.face]++; d->live++; }
  // ---- items of the faces that survived the round: the face, then its hull edges by ascending slot (the order
  //      add_face_errors pushes them in)
  std::vector<int> item_face;
  std::vector<double> w, cent;           // lifted vertices [item][<=3][3] (padded), centroids [item][2]
  std::vector<int> item_nv;
  for (int fi : round_faces) {
    if (!hm.alive[fi] || hm.is_inf(fi) || hm.mark[fi]) continue;
    hm.mark[fi] = 1;
    const Face &t = hm.fc[fi];
    double tw[9];
    for (int i = 0; i < 3; i++) { tw[i * 3] = hm.X[t.v[i]]; tw[i * 3 + 1] = hm.Y[t.v[i]]; tw[i * 3 + 2] = hm.F[t.v[i]]; }
    item_face.push_back(fi); item_nv.push_back(3);
    w.insert(w.end(), tw, tw + 9);
    cent.push_back((tw[0] + tw[3] + tw[6]) / 3); cent.push_back((tw[1] + tw[4] + tw[7]) / 3);
    for (int j = 0; j < 3; j++)
      if (hm.is_inf(t.n[j])) {
        double e[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        int c = 0;
        for (int i = 0; i < 3; i++)
          if (i != j) { e[c * 3] = tw[i * 3]; e[c * 3 + 1] = tw[i * 3 + 1]; e[c * 3 + 2] = tw[i * 3 + 2]; c++; }
        item_face.push_back(fi); item_nv.push_back(2);
        w.insert(w.end(), e, e + 9);
        cent.push_back((e[0] + e[3]) / 2); cent.push_back((e[1] + e[4]) / 2);
      }
  }
  for (int fi : round_faces) hm.mark[fi] = 0;
  const int64_t ni = (int64_t)item_face.size();
  std::vector<double> phi((size_t)ni);
  if (fnb) { if (ni > 0) fnb(2, ni, cent.data(), phi.data(), user_b); }
  else for (int64_t i = 0; i < ni; i++) phi[(size_t)i] = fn(2, &cent[2 * (size_t)i], d->user);
  for (int64_t i = 0; i < ni; i++)
    d->push(dli_simplex_error<2>(item_nv[(size_t)i], &w[9 * (size_t)i], phi[(size_t)i]), phi[(size_t)i], cent[2 * (size_t)i],
            cent[2 * (size_t)i + 1], item_face[(size_t)i]);
  return done;
  DLI_GUARD_END
}
int64_t dli_di2_error_queue_size(dli_di2 *d) { return d ? d->live : 0; }
int dli_di2_get_error_queue(dli_di2 *d, double *err, double *f, double *xy) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  std::vector<QEntry> all;
  auto copy = d->pq;
  while (!copy.empty()) {
    if (d->entry_live(copy.top())) all.push_back(copy.top());
    copy.pop();
  }
  for (size_t i = 0; i < all.size(); i++) {
    if (err) err[i] = all[i].err;
    if (f) f[i] = all[i].f;
    if (xy) { xy[2 * i] = all[i].x; xy[2 * i + 1] = all[i].y; }
  }
  return DLI_OK;
  DLI_GUARD_END
}

int64_t dli_di2_num_vertices(const dli_di2 *d) { return d ? (int64_t)d->hm.X.size() : 0; }
int64_t dli_di2_num_faces(const dli_di2 *d) { return d ? d->hm.n_finite : 0; }
int dli_di2_dimension(const dli_di2 *d) { return d ? d->hm.dim : -1; }

int dli_di2_export(const dli_di2 *d, double *coords, double *f, int32_t *tri, int32_t *nbr) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  const HostMesh2 &hm = d->hm;
  int64_t nv = (int64_t)hm.X.size();
  if (f) memcpy(f, hm.F.data(), sizeof(double) * (size_t)nv);
  const_cast<dli_di2 *>(d)->flatten(coords, tri, nbr, nullptr);   // (renumbers only after a non-compact insertion)
  return DLI_OK;
  DLI_GUARD_END
}

int dli_di2_set_options(dli_di2 *d, const dli_options *opt) {
  if (!d) return DLI_ERR_ARG;
  d->have_opt = opt != nullptr;
  if (opt) d->opt = *opt;
  d->dirty = true;
  d->need_full = true;
  return DLI_OK;
}

// ---------------------------------------------------------------------------------------------
// Host side of small batches (dli_host_small.h): interp_point of the reference (src/delaunay_2_interp.h:515-572) on
// the host triangulation -- exact visibility walk over closed faces from the face the previous query ended in,
// outside-hull projection with the rules of the device's exact pass (locate2_full in dli_device.cuh: canonical visible
// hull edge = the strictly visible one whose line is farthest, ties to the smaller (face id, slot); CGAL-style Line_2
// projection; <= 16 rounds), reference-order epilogue (dli_refmath.h).  Same code as the exact pass of the kernels, so
// the values are those of the device in `reference` mode.
// ---------------------------------------------------------------------------------------------
namespace {
struct HullEdge2 {
  int inf_face, face, slot;   // the infinite face, the finite face behind it, the slot of the hull edge in that face
  int P, Q;                   // edge end points, counter-clockwise in `face`
};
HullEdge2 hull_edge_of(const HostMesh2 &hm, int inf_face) {
  HullEdge2 h;
  h.inf_face = inf_face;
  const Face &I = hm.fc[inf_face];
  int i = I.v[0] == INF ? 0 : (I.v[1] == INF ? 1 : 2);
  h.face = I.n[i];
  const Face &F = hm.fc[h.face];
  h.slot = F.n[0] == inf_face ? 0 : (F.n[1] == inf_face ? 1 : 2);
  h.P = F.v[(h.slot + 1) % 3];
  h.Q = F.v[(h.slot + 2) % 3];
  return h;
}
// the infinite face sharing Q (dir 0: next) / P (dir 1: prev) with hull edge h
int hull_step(const HostMesh2 &hm, const HullEdge2 &h, int dir) {
  const Face &I = hm.fc[h.inf_face];
  for (int k = 0; k < 3; k++)
    if (I.v[k] == (dir ? h.Q : h.P)) return I.n[k];   // across the edge opposite the OTHER end point
  return -1;
}
HullEdge2 canonical_visible_hull_edge(const HostMesh2 &hm, double px, double py, int inf0) {
  HullEdge2 best = hull_edge_of(hm, inf0);
  double best_score = dli::hull_edge_score(hm.X[best.P], hm.Y[best.P], hm.X[best.Q], hm.Y[best.Q], px, py);
  const int64_t cap = (int64_t)hm.fc.size();
  for (int dir = 0; dir < 2; dir++) {
    HullEdge2 h = hull_edge_of(hm, inf0);
    for (int64_t it = 0; it < cap; it++) {
      const int nx = hull_step(hm, h, dir);
      if (nx < 0 || nx == inf0) break;
      h = hull_edge_of(hm, nx);
      if (hm.o3(h.P, h.Q, px, py) >= 0) break;
      const double sc = dli::hull_edge_score(hm.X[h.P], hm.Y[h.P], hm.X[h.Q], hm.Y[h.Q], px, py);
      const bool lower = hm.fid[h.face] < hm.fid[best.face] || (hm.fid[h.face] == hm.fid[best.face] && h.slot < best.slot);
      if (sc > best_score || (sc == best_score && lower)) { best = h; best_score = sc; }
    }
  }
  return best;
}
constexpr int kMaxProjectionsHost = 16;   // = kMaxProjections of dli_device.cuh

// one query; returns the status, the face in *face_out (-1: none), the value in *y.  outside_nan: the handle's
// dli_options.outside_nan -- a query outside the hull is not projected, it reports NaN like the device kernels do
uint8_t host_interp_point2(const HostMesh2 &hm, double px, double py, int start, int *face_out, double *y,
                           bool outside_nan = false) {
  *face_out = -1;
  *y = std::nan("");
  if (hm.dim < 2) return DLI_STATUS_NO_MESH;
  if (!(std::isfinite(px) && std::isfinite(py))) return DLI_STATUS_STEP_LIMIT;
  uint8_t status = DLI_STATUS_OK;
  bool stuck = false;
  int prev_face = -1, prev_slot = -1, f = start;
  for (int proj = 0;; proj++) {
    f = hm.locate(px, py, f);
    if (!hm.is_inf(f)) break;
    if (outside_nan) return DLI_STATUS_OUTSIDE_NAN;   // *y is NaN, *face_out is -1
    status = DLI_STATUS_OUTSIDE_PROJECTED;
    if (proj >= kMaxProjectionsHost || stuck) { f = hull_edge_of(hm, f).face; break; }   // accept the hull face
    const HullEdge2 h = canonical_visible_hull_edge(hm, px, py, f);
    const double opx = px, opy = py;
    dli::project_on_line(hm.X[h.P], hm.Y[h.P], hm.X[h.Q], hm.Y[h.Q], px, py);
    stuck = (h.face == prev_face && h.slot == prev_slot && px == opx && py == opy);
    prev_face = h.face; prev_slot = h.slot;
    f = h.face;
  }
  const Face &t = hm.fc[f];
  *y = dli::epilogue2_ref_xy(px, py, hm.X[t.v[0]], hm.Y[t.v[0]], hm.F[t.v[0]], hm.X[t.v[1]], hm.Y[t.v[1]], hm.F[t.v[1]],
                             hm.X[t.v[2]], hm.Y[t.v[2]], hm.F[t.v[2]]);
  *face_out = f;
  return status;
}
}  // namespace

static void di2_host_interp(dli_di2 *d, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto) {
  HostMesh2 &hm = d->hm;
  if (so && !hm.fid_compact) { hm.renumber(); d->need_full = true; d->dirty = true; }
  int start = d->host_last.load(std::memory_order_relaxed);
  const bool outside_nan = d->have_opt && d->opt.outside_nan != 0;
  for (int64_t i = 0; i < n; i++) {
    int face;
    double y;
    const uint8_t st = host_interp_point2(hm, xs[2 * i], xs[2 * i + 1], start, &face, &y, outside_nan);
    ys[i] = y;
    if (so) so[i] = face >= 0 ? hm.fid[face] : -1;
    if (sto) sto[i] = st;
    if (face >= 0) start = face;
  }
  d->host_last.store(start, std::memory_order_relaxed);
}
static bool di2_take_host(const dli_di2 *d, int64_t n, const double *xs, const double *ys, const int32_t *so, const uint8_t *sto) {
  // the host walk evaluates in the reference's operation order: a handle that asked for the closed-form epilogue
  // keeps every batch on the device, so that its values do not depend on the batch size
  if (d->have_opt && d->opt.epilogue == DLI_EPILOGUE_CLOSED_FORM) return false;
  return n > 0 && xs && ys && dli::host_batch_is_small(n, d->hm.dim == 2 ? d->hm.n_finite : 0, 2) &&
         dli::is_plain_host_pointer(xs) && dli::is_plain_host_pointer(ys) && dli::is_plain_host_pointer(so) &&
         dli::is_plain_host_pointer(sto);
}

// incremental path: hand the dirty star(s) to dli_mesh_update; DLI_ERR_UNSUPPORTED = the mesh cannot be patched
static int di2_patch(dli_di2 *d) {
  HostMesh2 &hm = d->hm;
  const int64_t nv = (int64_t)hm.X.size(), nv0 = d->nv_synced, nf = hm.n_finite;
  std::vector<double> nxy((size_t)(nv - nv0) * 2 + 2), nf_((size_t)(nv - nv0) + 1);
  for (int64_t v = nv0; v < nv; v++) { nxy[2 * (v - nv0)] = hm.X[v]; nxy[2 * (v - nv0) + 1] = hm.Y[v]; nf_[v - nv0] = hm.F[v]; }
  std::vector<int32_t> vid;
  std::vector<double> vval;
  for (int v : d->pend_values)
    if (v < nv0) { vid.push_back(v); vval.push_back(hm.F[v]); }
  // live finite faces among the pending slots, once each
  std::vector<int32_t> ids, sv, sn;
  for (int fi : d->pend_faces) {
    if (!hm.alive[fi] || hm.fid[fi] < 0 || hm.mark[fi]) continue;
    hm.mark[fi] = 1;
    ids.push_back(hm.fid[fi]);
    for (int j = 0; j < 3; j++) sv.push_back(hm.fc[fi].v[j]);
    for (int j = 0; j < 3; j++) sn.push_back(hm.fid[hm.fc[fi].n[j]]);
  }
  for (int fi : d->pend_faces) hm.mark[fi] = 0;
  d->patched_since_full += (int64_t)ids.size();
  return dli_mesh_update(d->mesh, nv, nxy.data(), nf_.data(), (int64_t)vid.size(), vid.data(), vval.data(), nf,
                         (int64_t)ids.size(), ids.data(), sv.data(), sn.data());
}

static int di2_refresh(dli_di2 *d) {
  if (!d->dirty && d->mesh) return DLI_OK;
  // the start-hint grid is not patched: once half of the mesh it was built for has been replaced, start over
  const bool stale_hints = d->patched_since_full > std::max<int64_t>(4096, d->faces_at_full / 2);
  if (d->mesh && !d->need_full && !stale_hints && d->hm.fid_compact && d->hm.dim == 2) {
    int rc = di2_patch(d);
    if (rc == DLI_OK) { d->synced(false); return DLI_OK; }
    if (rc != DLI_ERR_UNSUPPORTED) return rc;
  }
  if (d->mesh) { dli_mesh_destroy(d->mesh); d->mesh = nullptr; }
  int64_t nv = (int64_t)d->hm.X.size(), nf = d->hm.dim == 2 ? d->hm.n_finite : 0;
  std::vector<double> coords((size_t)nv * 2 + 2);
  std::vector<int32_t> tri((size_t)nf * 3 + 3), nbr((size_t)nf * 3 + 3);
  d->flatten(coords.data(), nf ? tri.data() : nullptr, nf ? nbr.data() : nullptr, nullptr);
  std::vector<double> fv(d->hm.F);
  fv.push_back(0.0);
  int rc = dli_mesh_create(2, nv, coords.data(), fv.data(), nf, tri.data(), nbr.data(), d->have_opt ? &d->opt : nullptr,
                           &d->mesh);
  if (rc == DLI_OK) d->synced(true);
  return rc;
}

void dli_di2_sync_counts(const dli_di2 *d, int64_t *n_full, int64_t *n_incremental) {
  if (n_full) *n_full = d ? d->n_full : 0;
  if (n_incremental) *n_incremental = d ? d->n_incremental : 0;
}

int dli_di2_interp(dli_di2 *d, int64_t n, const double *xs, double *ys, int32_t *simplex_out, uint8_t *status_out,
                   void *stream) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  if (di2_take_host(d, n, xs, ys, simplex_out, status_out)) { di2_host_interp(d, n, xs, ys, simplex_out, status_out); return DLI_OK; }
  int rc = di2_refresh(d);
  if (rc != DLI_OK) return rc;
  return dli_interp(d->mesh, n, xs, ys, simplex_out, status_out, stream);
  DLI_GUARD_END
}
int dli_di2_gradient(dli_di2 *d, int64_t n, const double *xs, double *grads, int32_t *simplex_out, uint8_t *status_out,
                     void *stream) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  int rc = di2_refresh(d);
  if (rc != DLI_OK) return rc;
  return dli_gradient(d->mesh, n, xs, grads, simplex_out, status_out, stream);
  DLI_GUARD_END
}
const dli_mesh *dli_di2_device_mesh(dli_di2 *d) {
  if (!d) return nullptr;
  if (di2_refresh(d) != DLI_OK) return nullptr;
  return d->mesh;
}

// get_line_segments with concavity flags and min / max (src/delaunay_2_interp.h:842-862, :766-803), batched on the GPU
int64_t dli_di2_line_segments(dli_di2 *d, double *segments, int32_t *concave, double *min_coords, double *max_coords) {
  DLI_GUARD_BEGIN
  if (!d) return DLI_ERR_ARG;
  int rc = di2_refresh(d);
  if (rc != DLI_OK) return rc;
  int64_t ne = dli_mesh_num_edges(d->mesh);
  if (ne < 0) return ne;
  if (segments || concave || min_coords || max_coords) {
    rc = dli_mesh_line_segments(d->mesh, segments, concave, min_coords, max_coords);
    if (rc != DLI_OK) return rc;
  }
  return ne;
  DLI_GUARD_END
}

int64_t dli_di2_get_all_vertices(const dli_di2 *d, double *xyz) {
  if (!d) return 0;
  int64_t nv = (int64_t)d->hm.X.size();
  if (xyz) for (int64_t v = 0; v < nv; v++) { xyz[3 * v] = d->hm.X[v]; xyz[3 * v + 1] = d->hm.Y[v]; xyz[3 * v + 2] = d->hm.F[v]; }
  return nv;
}
int64_t dli_di2_get_line_segments(const dli_di2 *d, double *seg) {
  DLI_GUARD_BEGIN
  if (!d) return 0;
  const HostMesh2 &hm = d->hm;
  int64_t ne = 0;
  for (int fi = 0; fi < (int)hm.fc.size(); fi++) {
    if (!hm.alive[fi] || hm.is_inf(fi)) continue;
    for (int j = 0; j < 3; j++) {
      int nb = hm.fc[fi].n[j];
      if (!hm.is_inf(nb) && nb < fi) continue;  // interior edges once
      int a = hm.fc[fi].v[(j + 1) % 3], b = hm.fc[fi].v[(j + 2) % 3];
      if (seg) {
        double *s = seg + 6 * ne;
        s[0] = hm.X[a]; s[1] = hm.Y[a]; s[2] = hm.F[a];
        s[3] = hm.X[b]; s[4] = hm.Y[b]; s[5] = hm.F[b];
      }
      ne++;
    }
  }
  return ne;
  DLI_GUARD_END
}

}  // extern "C"
