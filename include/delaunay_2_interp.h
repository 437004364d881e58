// delaunay_2_interp.h -- source-compatible stand-in for the reference's header of the same name
// (rncarpio/delaunay_linterp, src/delaunay_2_interp.h): class Delaunay_incremental_interp_2 with the same
// public members and semantics, plus the batched entry point interp_batch(xs -> ys) that runs on the B200.
//
//   reference member (src/delaunay_2_interp.h)            here
//   ctor() / ctor(std::function<double(int,double*)>) :410-416   same
//   insert(IterT, IterT, double)                :450-453   same (host, native triangulation: include/dli_host.h)
//   insert_point(Point const&, double)          :455-493   same
//   insert_largest_error_point() -> ErrorTuple  :498-504   same (asserts on empty queue / no callback like :499-500)
//   get_largest_error_tuple()                   :506-508   same
//   interp(IterT, IterT) / interp_point(Point)  :510-572   same result; one point at a time is answered on the host (exact walk on the
//                                                          native triangulation from the previous result, reference-order epilogue:
//                                                          no GPU call, ~0.3 us per call); batches go through the GPU engine
//   get_error_queue()                           :743-746   same (ordered by decreasing error)
//   get_all_vertices() / get_line_segments()    :806-862   same data, plain arrays
//   gradient(Point)                             :608-642   same (FACE / EDGE / VERTEX averaging, NaN outside the hull); gradient_batch for n queries
//   public data m_pTriang / m_err_pqueue / m_face_to_pqueue_handles are CGAL/Boost typed in the reference and have
//   no counterpart here (neither library is available); m_fn is kept.
// New:
//   interp_batch(xs, n, ys, status)   n queries at once; host or device pointers (see dli_interp in dli.h)
//
// Link with -ldli (delaunay_linterp_b200/libdli.so).
#ifndef DELAUNAY_2_INTERP_H_
#define DELAUNAY_2_INTERP_H_

#include <array>
#include <cassert>
#include <cstdio>
#include <functional>
#include <limits>
#include <memory>
#include <stdexcept>
#include <algorithm>
#include <string>
#include <tuple>
#include <vector>

#include "dli_host.h"

// the reference's Point is CGAL::Point_2<Epick>; this one offers the accessors the reference code uses on it
class Point {
 public:
  Point() : c_{{0.0, 0.0}} {}
  Point(double x, double y) : c_{{x, y}} {}
  double x() const { return c_[0]; }
  double y() const { return c_[1]; }
  double operator[](int i) const { return c_[i]; }
  int dimension() const { return 2; }
  bool operator==(Point const &o) const { return c_ == o.c_; }

 private:
  std::array<double, 2> c_;
};

// (error, f(x), x): src/delaunay_2_interp.h:376-395
class ErrorTuple {
 public:
  double m_err, m_f;
  Point m_x;
  ErrorTuple(double err, double f, Point const &x) : m_err(err), m_f(f), m_x(x) {}
  bool operator<(ErrorTuple const &b) const { return m_err < b.m_err; }
};

class Delaunay_incremental_interp_2 {
 public:
  static const int N = 2;
  std::function<double(int, double *)> m_fn;  // callback function (:408)

  Delaunay_incremental_interp_2() : h_(dli_di2_create(nullptr, nullptr)) {}
  explicit Delaunay_incremental_interp_2(std::function<double(int, double *)> fn) : m_fn(fn) {
    h_ = dli_di2_create(m_fn ? &Delaunay_incremental_interp_2::trampoline : nullptr, this);
  }
  virtual ~Delaunay_incremental_interp_2() { dli_di2_destroy(h_); }
  Delaunay_incremental_interp_2(Delaunay_incremental_interp_2 const &) = delete;
  Delaunay_incremental_interp_2 &operator=(Delaunay_incremental_interp_2 const &) = delete;

  template <class IterT>
  void insert(IterT x_begin, IterT x_end, double f) {
    (void)x_end;
    insert_point(Point(x_begin[0], x_begin[1]), f);
  }
  void insert_point(Point const &x, double f) {
    if (dli_di2_insert(h_, x.x(), x.y(), f) < 0) throw std::invalid_argument("insert: point is not finite");
  }
  // bulk form of insert (new): same triangulation, inserted along a Hilbert curve
  void insert_many(const double *xy, const double *f, size_t n, bool rebuild_queue_on_gpu = true) {
    check(dli_di2_insert_many(h_, (int64_t)n, xy, f, 1, rebuild_queue_on_gpu ? 1 : 0), "insert_many");
  }

  // take the tuple with the largest error and insert it; f is not evaluated again (:498-504)
  ErrorTuple insert_largest_error_point() {
    double e, f, xy[2];
    int rc = dli_di2_insert_largest_error_point(h_, &e, &f, xy);
    assert(rc == DLI_OK && "insert_largest_error_point: empty queue or no callback");
    if (rc != DLI_OK) throw std::logic_error("insert_largest_error_point: empty queue or no callback");
    return ErrorTuple(e, f, Point(xy[0], xy[1]));
  }
  // k-at-a-time form (new, dli_di2_refine): up to k largest-error tuples on pairwise non-adjacent faces are inserted,
  // then the errors of all new faces are evaluated in one batch; fn_batch(n_dims, n, xs, out, user) is called once
  // per round (nullptr: the scalar callback per item).  k = 1 is insert_largest_error_point().
  std::vector<ErrorTuple> insert_largest_error_points(size_t k, dli_fn_batch fn_batch = nullptr, void *user = nullptr,
                                                      double rel_threshold = 0.25) {
    std::vector<double> e(k + 1), f(k + 1), xy(2 * k + 2);
    int64_t n = dli_di2_refine(h_, (int64_t)k, rel_threshold, fn_batch, user, e.data(), f.data(), xy.data());
    if (n < 0) throw std::logic_error("insert_largest_error_points: no callback");
    std::vector<ErrorTuple> out;
    for (int64_t i = 0; i < n; i++) out.push_back(ErrorTuple(e[i], f[i], Point(xy[2 * i], xy[2 * i + 1])));
    return out;
  }
  ErrorTuple get_largest_error_tuple() const {
    double e, f, xy[2];
    if (dli_di2_get_largest_error_tuple(h_, &e, &f, xy) != DLI_OK)
      throw std::logic_error("get_largest_error_tuple: empty queue or no callback");
    return ErrorTuple(e, f, Point(xy[0], xy[1]));
  }

  template <class IterT>
  double interp(IterT x_begin, IterT x_end) {
    (void)x_end;
    return interp_point(Point(x_begin[0], x_begin[1]));
  }
  double interp_point(Point const &x) const {
    double xs[2] = {x.x(), x.y()}, y;
    uint8_t st;
    check(dli_di2_interp(h_, 1, xs, &y, nullptr, &st, nullptr), "interp");
    if (st == DLI_STATUS_NO_MESH || st == DLI_STATUS_STEP_LIMIT) {  // :546-551
      char buf[160];
      snprintf(buf, sizeof buf, "interp: no simplex containing the point could be found: status=%d", (int)st);
      throw std::invalid_argument(buf);
    }
    return y;
  }
  // batched interp (new): xs [n][2] -> ys [n]; host or device pointers; status may be NULL.
  // Queries outside the hull are projected like interp_point does (:530-545) and flagged in status.
  void interp_batch(const double *xs, size_t n, double *ys, unsigned char *status = nullptr, void *cuda_stream = nullptr) const {
    check(dli_di2_interp(h_, (int64_t)n, xs, ys, nullptr, status, cuda_stream), "interp_batch");
  }

  // gradient of the interpolant at x (:608-642): the face's gradient; the mean of the incident faces on an edge or
  // a vertex; quiet NaN outside the hull
  std::array<double, 2> gradient(Point const &x) const {
    double xs[2] = {x.x(), x.y()};
    std::array<double, 2> g;
    check(dli_di2_gradient(h_, 1, xs, g.data(), nullptr, nullptr, nullptr), "gradient");
    return g;
  }
  // batched gradient (new): xs [n][2] -> grads [n][2]; host or device pointers
  void gradient_batch(const double *xs, size_t n, double *grads, unsigned char *status = nullptr, void *cuda_stream = nullptr) const {
    check(dli_di2_gradient(h_, (int64_t)n, xs, grads, nullptr, status, cuda_stream), "gradient_batch");
  }

  std::vector<ErrorTuple> get_error_queue() const {
    int64_t n = dli_di2_error_queue_size(h_);
    std::vector<double> e((size_t)n), f((size_t)n), xy((size_t)n * 2);
    check(dli_di2_get_error_queue(h_, e.data(), f.data(), xy.data()), "get_error_queue");
    std::vector<ErrorTuple> out;
    out.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++) out.emplace_back(e[i], f[i], Point(xy[2 * i], xy[2 * i + 1]));
    return out;
  }

  // plot data (:806-862): vertices as (x, y, f); edges as pairs of such points
  std::vector<std::array<double, 3>> get_all_vertices() const {
    std::vector<std::array<double, 3>> v((size_t)dli_di2_get_all_vertices(h_, nullptr));
    if (!v.empty()) dli_di2_get_all_vertices(h_, &v[0][0]);
    return v;
  }
  // host-side list of the finite edges (no GPU, no flags)
  std::vector<std::array<std::array<double, 3>, 2>> get_edges() const {
    std::vector<std::array<std::array<double, 3>, 2>> s((size_t)dli_di2_get_line_segments(h_, nullptr));
    if (!s.empty()) dli_di2_get_line_segments(h_, &s[0][0][0]);
    return s;
  }
  // for plotting (:842-862): all line segments with the is_edge_concave flag (:766-803), and the min / max at each
  // coordinate; evaluated on the GPU for all edges at once (dli_mesh_line_segments)
  typedef std::tuple<std::array<double, 3>, std::array<double, 3>, int> Segment;
  std::tuple<std::vector<Segment>, std::array<double, 3>, std::array<double, 3>> get_line_segments() const {
    int64_t ne = dli_di2_line_segments(h_, nullptr, nullptr, nullptr, nullptr);
    check((int)std::min<int64_t>(ne, 0), "get_line_segments");
    std::vector<double> seg((size_t)ne * 6 + 6);
    std::vector<int32_t> cc((size_t)ne + 1);
    std::array<double, 3> mn, mx;
    check((int)std::min<int64_t>(dli_di2_line_segments(h_, seg.data(), cc.data(), mn.data(), mx.data()), 0), "get_line_segments");
    std::vector<Segment> out;
    for (int64_t i = 0; i < ne; i++) {
      std::array<double, 3> p = {{seg[6 * i], seg[6 * i + 1], seg[6 * i + 2]}}, q = {{seg[6 * i + 3], seg[6 * i + 4], seg[6 * i + 5]}};
      out.push_back(Segment(p, q, (int)cc[i]));
    }
    return std::make_tuple(out, mn, mx);
  }

  size_t number_of_vertices() const { return (size_t)dli_di2_num_vertices(h_); }
  size_t number_of_faces() const { return (size_t)dli_di2_num_faces(h_); }
  // flattened triangulation (what the engine uploads): coords [nv][2], f [nv], tri [nf][3], nbr [nf][3]
  void flatten(std::vector<double> &coords, std::vector<double> &f, std::vector<int32_t> &tri, std::vector<int32_t> &nbr) const {
    size_t nv = number_of_vertices(), nf = number_of_faces();
    coords.resize(nv * 2); f.resize(nv); tri.resize(nf * 3); nbr.resize(nf * 3);
    check(dli_di2_export(h_, coords.data(), f.data(), nf ? tri.data() : nullptr, nf ? nbr.data() : nullptr), "flatten");
  }
  dli_di2 *handle() const { return h_; }

 private:
  dli_di2 *h_;
  static double trampoline(int n, const double *x, void *user) {
    auto *self = static_cast<Delaunay_incremental_interp_2 *>(user);
    double args[2] = {x[0], x[1]};
    return self->m_fn(n, args);
  }
  static void check(int rc, const char *who) {
    if (rc != DLI_OK) throw std::runtime_error(std::string(who) + ": " + dli_last_error());
  }
};

#endif  // DELAUNAY_2_INTERP_H_
