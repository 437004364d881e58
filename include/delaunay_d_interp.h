// delaunay_d_interp.h -- source-compatible stand-in for the reference's header of the same name
// (rncarpio/delaunay_linterp, src/delaunay_d_interp.h): class Delaunay_incremental_interp_d with the same public
// members and semantics for n_dims = 2, 3, 4, plus the batched entry point interp_batch(xs -> ys) that runs on the B200.
//
//   reference member (src/delaunay_d_interp.h)                 here
//   ctor(int) / ctor(int, std::function<double(int,double*)>) :772-781   same
//   get_dimension()                                 :783        same
//   insert(IterT, IterT, double)                    :933-937    same (host, native triangulation: include/dli_host.h)
//   insert_point(Point const&, double)              :939-956    same
//   interp(IterT, IterT) / interp_point(Point)      :982-1010   same result; quiet NaN outside the hull (:1005-1008).
//                                                               One point at a time is answered on the host (exact walk
//                                                               from the previous result, reference-order epilogue);
//                                                               batches go through the CUDA engine
//   insert_largest_error_point() -> ErrorTuple      :1028-1034  same (asserts on empty queue / no callback like :1029-1030)
//   get_largest_error_tuple()                       :1036-1038  same
//   get_error_queue()                               :1040-1043  same (ordered by decreasing error)
//   get_all_vertices()                              :1097-1101  same data: rows (x_0 .. x_{N-1}, f) as doubles (the
//                                                               reference's RT is a Gmpq-backed number type)
//   get_line_segments()                             :1115-1141  same: (p, q, 1) per segment + min / max per coordinate
//   free function get_lifted_point(...)             :1145-1149  get_lifted_point(x): (x, |x|^2)
//   public data N, m_N, m_fn are kept; m_pDelaunay / m_pointMap / m_simplex_info_map / m_err_pqueue are CGAL / Boost
//   typed in the reference and have no counterpart here (neither library is available).
// New:
//   interp_batch(xs, n, ys, status)   n queries at once; host or device pointers (see dli_interp in dli.h)
//   insert_many(xs, f, n)             bulk insertion (spatially sorted)
//   insert_largest_error_points(k, fn_batch)   k-at-a-time refinement with one batched callback per round (dli_dd_refine)
//
// Point / ErrorTuple live in namespace delaunay_d and are exported to the global namespace unless the 2D header
// (which, as in the reference, defines global types of the same names) was included first.
//
// Link with -ldli (delaunay_linterp_b200/libdli.so).
#ifndef DELAUNAY_D_INTERP_H_
#define DELAUNAY_D_INTERP_H_

#include <cassert>
#include <cfloat>
#include <cstdio>
#include <functional>
#include <limits>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "dli_host.h"

namespace delaunay_d {

// the reference's Point is CGAL::Point_d<Cartesian_d<Gmpq>> (:43-55); this one offers the accessors the reference
// code uses on it: Point(d, begin, end), dimension(), operator[], cartesian_begin() / cartesian_end()
class Point {
 public:
  Point() {}
  template <class IterT>
  Point(int d, IterT first, IterT last) : c_(first, last) {
    (void)d;
    assert((int)c_.size() == d);
  }
  explicit Point(std::vector<double> const &c) : c_(c) {}
  int dimension() const { return (int)c_.size(); }
  double operator[](int i) const { return c_[(size_t)i]; }
  double cartesian(int i) const { return c_[(size_t)i]; }
  std::vector<double>::const_iterator cartesian_begin() const { return c_.begin(); }
  std::vector<double>::const_iterator cartesian_end() const { return c_.end(); }
  const double *data() const { return c_.data(); }
  bool operator==(Point const &o) const { return c_ == o.c_; }

 private:
  std::vector<double> c_;
};

// (error, f(x), x): src/delaunay_d_interp.h:724-746
class ErrorTuple {
 public:
  double m_err, m_f;
  Point m_x;
  ErrorTuple(double err, double f, Point const &x) : m_err(err), m_f(f), m_x(x) {}
  bool operator<(ErrorTuple const &b) const { return m_err < b.m_err; }
  std::string to_str() const {
    std::string s = "err: " + std::to_string(m_err) + ", f: " + std::to_string(m_f) + ", x:";
    for (int i = 0; i < m_x.dimension(); i++) s += " " + std::to_string(m_x[i]);
    return s;
  }
};

class Delaunay_incremental_interp_d {
 public:
  int N, m_N;
  std::function<double(int, double *)> m_fn;  // callback function (:758)

  explicit Delaunay_incremental_interp_d(int n_dims) : N(n_dims), m_N(n_dims), h_(dli_dd_create(n_dims, nullptr, nullptr)) {
    if (!h_) throw std::invalid_argument("Delaunay_incremental_interp_d: n_dims must be 2, 3 or 4");
  }
  Delaunay_incremental_interp_d(int n_dims, std::function<double(int, double *)> fn) : N(n_dims), m_N(n_dims), m_fn(fn) {
    h_ = dli_dd_create(n_dims, m_fn ? &Delaunay_incremental_interp_d::trampoline : nullptr, this);
    if (!h_) throw std::invalid_argument("Delaunay_incremental_interp_d: n_dims must be 2, 3 or 4");
  }
  virtual ~Delaunay_incremental_interp_d() { dli_dd_destroy(h_); }
  Delaunay_incremental_interp_d(Delaunay_incremental_interp_d const &) = delete;
  Delaunay_incremental_interp_d &operator=(Delaunay_incremental_interp_d const &) = delete;

  int get_dimension() const { return m_N; }

  template <class IterT>
  void insert(IterT x_begin, IterT x_end, double f) {
    assert(x_end - x_begin == N);
    insert_point(Point(N, x_begin, x_end), f);
  }
  void insert_point(Point const &x, double f) {
    if (x.dimension() != N || dli_dd_insert(h_, x.data(), f) < 0)
      throw std::invalid_argument("insert: point must have n_dims finite coordinates");
  }
  // bulk form of insert (new): xs [n][N] row-major; same triangulation, inserted along a space-filling curve
  void insert_many(const double *xs, const double *f, size_t n, bool rebuild_queue_on_gpu = true) {
    check(dli_dd_insert_many(h_, (int64_t)n, xs, f, 1, rebuild_queue_on_gpu ? 1 : 0), "insert_many");
  }

  template <class IterT>
  double interp(IterT x_begin, IterT x_end) {
    assert(x_end - x_begin == N);
    return interp_point(Point(N, x_begin, x_end));
  }
  // quiet NaN when no simplex contains the point (:1005-1008)
  double interp_point(Point const &p) const {
    if (p.dimension() != N) throw std::invalid_argument("interp: point must have n_dims coordinates");
    double y = std::numeric_limits<double>::quiet_NaN();
    check(dli_dd_interp(h_, 1, p.data(), &y, nullptr, nullptr, nullptr), "interp");
    return y;
  }
  // batched interp (new): xs [n][N] -> ys [n]; host or device pointers; status may be NULL
  void interp_batch(const double *xs, size_t n, double *ys, unsigned char *status = nullptr, void *cuda_stream = nullptr) const {
    check(dli_dd_interp(h_, (int64_t)n, xs, ys, nullptr, status, cuda_stream), "interp_batch");
  }

  // take the tuple with the largest error and insert it; f is not evaluated again (:1025-1034)
  ErrorTuple insert_largest_error_point() {
    double e, f;
    std::vector<double> x((size_t)N);
    int rc = dli_dd_insert_largest_error_point(h_, &e, &f, x.data());
    assert(rc == DLI_OK && "insert_largest_error_point: empty queue or no callback");
    if (rc != DLI_OK) throw std::logic_error("insert_largest_error_point: empty queue or no callback");
    return ErrorTuple(e, f, Point(x));
  }
  // k-at-a-time form (new, dli_dd_refine): up to k largest-error tuples on simplices that share no vertex are inserted,
  // then the items of every simplex created in the round are evaluated with ONE call of fn_batch (nullptr: m_fn per
  // item).  k = 1 is insert_largest_error_point.
  std::vector<ErrorTuple> insert_largest_error_points(size_t k, dli_fn_batch fn_batch = nullptr, void *user = nullptr,
                                                      double rel_threshold = 0.25) {
    std::vector<double> e(k + 1), f(k + 1), xs((k + 1) * (size_t)N);
    int64_t n = dli_dd_refine(h_, (int64_t)k, rel_threshold, fn_batch, user, e.data(), f.data(), xs.data());
    if (n < 0) throw std::logic_error("insert_largest_error_points: no callback");
    std::vector<ErrorTuple> out;
    for (int64_t i = 0; i < n; i++)
      out.push_back(ErrorTuple(e[(size_t)i], f[(size_t)i], Point(std::vector<double>(xs.begin() + i * N, xs.begin() + (i + 1) * N))));
    return out;
  }
  ErrorTuple get_largest_error_tuple() const {
    double e, f;
    std::vector<double> x((size_t)N);
    if (dli_dd_get_largest_error_tuple(h_, &e, &f, x.data()) != DLI_OK)
      throw std::logic_error("get_largest_error_tuple: empty queue or no callback");
    return ErrorTuple(e, f, Point(x));
  }
  std::vector<ErrorTuple> get_error_queue() const {
    int64_t cap = dli_dd_error_queue_size(h_);
    std::vector<double> e((size_t)cap + 1), f((size_t)cap + 1), xs((size_t)cap * N + 1);
    int64_t n = dli_dd_get_error_queue(h_, cap, e.data(), f.data(), xs.data());
    if (n < 0) throw std::runtime_error("get_error_queue failed");
    std::vector<ErrorTuple> out;
    out.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++) out.emplace_back(e[i], f[i], Point(N, xs.begin() + i * N, xs.begin() + (i + 1) * N));
    return out;
  }

  // plot data (:1097-1101): the coordinates of every vertex plus f as an additional coordinate
  std::vector<std::vector<double>> get_all_vertices() const {
    int64_t nv = dli_dd_get_all_vertices(h_, nullptr);
    std::vector<double> flat((size_t)nv * (N + 1) + 1);
    dli_dd_get_all_vertices(h_, flat.data());
    std::vector<std::vector<double>> out((size_t)nv);
    for (int64_t v = 0; v < nv; v++) out[(size_t)v].assign(flat.begin() + v * (N + 1), flat.begin() + (v + 1) * (N + 1));
    return out;
  }
  // for plotting (:1115-1141): all line segments (p, q, 1), and the min / max at each coordinate
  typedef std::tuple<std::vector<double>, std::vector<double>, int> Segment;
  std::tuple<std::vector<Segment>, std::vector<double>, std::vector<double>> get_line_segments() const {
    int64_t ne = dli_dd_get_line_segments(h_, nullptr, nullptr, nullptr);
    if (ne < 0) throw std::runtime_error("get_line_segments failed");
    std::vector<double> seg((size_t)ne * 2 * (N + 1) + 1), mn((size_t)N + 1), mx((size_t)N + 1);
    dli_dd_get_line_segments(h_, seg.data(), mn.data(), mx.data());
    std::vector<Segment> out;
    out.reserve((size_t)ne);
    for (int64_t i = 0; i < ne; i++) {
      auto b = seg.begin() + i * 2 * (N + 1);
      out.emplace_back(std::vector<double>(b, b + N + 1), std::vector<double>(b + N + 1, b + 2 * (N + 1)), 1);
    }
    return std::make_tuple(out, mn, mx);
  }

  size_t number_of_vertices() const { return (size_t)dli_dd_num_vertices(h_); }
  size_t number_of_simplices() const { return (size_t)dli_dd_num_simplices(h_); }
  int current_dimension() const { return dli_dd_current_dimension(h_); }
  // flattened triangulation (what the engine uploads): coords [nv][N], f [nv], sv / sn [ns][N + 1]
  void flatten(std::vector<double> &coords, std::vector<double> &f, std::vector<int32_t> &sv, std::vector<int32_t> &sn) const {
    size_t nv = number_of_vertices(), ns = number_of_simplices();
    coords.resize(nv * N); f.resize(nv); sv.resize(ns * (N + 1)); sn.resize(ns * (N + 1));
    check(dli_dd_export(h_, coords.data(), f.data(), ns ? sv.data() : nullptr, ns ? sn.data() : nullptr), "flatten");
  }
  dli_dd *handle() const { return h_; }

 private:
  dli_dd *h_;
  static double trampoline(int n, const double *x, void *user) {
    auto *self = static_cast<Delaunay_incremental_interp_d *>(user);
    double args[DLI_MAX_DIM];
    for (int i = 0; i < n; i++) args[i] = x[i];
    return self->m_fn(n, args);
  }
  static void check(int rc, const char *who) {
    if (rc != DLI_OK) throw std::runtime_error(std::string(who) + ": " + dli_last_error());
  }
};

// the N+1 dimensional lifted point (x, |x|^2) (:1144-1149; the reference takes the triangulation as first argument
// only to reach its lifting functor)
inline Point get_lifted_point(Point const &x) {
  std::vector<double> l(x.cartesian_begin(), x.cartesian_end());
  double s = 0.0;
  for (double v : l) s += v * v;
  l.push_back(s);
  return Point(l);
}
template <class Ignored>
inline Point get_lifted_point(Ignored const &, Point const &x) { return get_lifted_point(x); }

}  // namespace delaunay_d

using delaunay_d::Delaunay_incremental_interp_d;
using delaunay_d::get_lifted_point;
#ifndef DELAUNAY_2_INTERP_H_
using delaunay_d::ErrorTuple;
using delaunay_d::Point;
#endif

#endif  // DELAUNAY_D_INTERP_H_
