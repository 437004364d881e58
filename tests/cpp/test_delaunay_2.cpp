// C++ port of the reference's demo driver src/delaunay_2_test2.cpp (README.md:29-110), with the assertions the
// reference never had.  Uses the drop-in header include/delaunay_2_interp.h; link with libdli.so.
//   ./test_delaunay_2            full run (needs a GPU: interp goes through the CUDA engine)
//   ./test_delaunay_2 --host     host-side part only (insert / adaptive queue / flatten)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "delaunay_2_interp.h"

using std::array;
using std::vector;

static vector<double> linspace(double first, double last, int len) {
  vector<double> result(len);
  double step = (last - first) / (len - 1);
  for (int i = 0; i < len; i++) result[i] = first + i * step;
  return result;
}
static double fn(int, double *x) { return pow(x[0], 0.3) * pow(x[1], 0.6); }

#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) { printf("CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); return 1; } \
  } while (0)

static int run(double lo, const char *want_regular, const char *want_adaptive, bool host_only) {
  int n_points = 10;
  Delaunay_incremental_interp_2 triang;
  vector<double> grid = linspace(lo, 5., n_points);
  for (int i = 0; i < n_points; i++)
    for (int j = 0; j < n_points; j++) {
      array<double, 2> args = {grid[i], grid[j]};
      triang.insert(args.begin(), args.end(), fn(2, &args[0]));
    }
  CHECK(triang.number_of_vertices() == 100 && triang.number_of_faces() == 162);

  std::function<double(int, double *)> fn_obj(fn);
  Delaunay_incremental_interp_2 adaptive_triang(fn_obj);
  array<double, 2> args;
  args[0] = grid.front(); args[1] = grid.front(); adaptive_triang.insert(args.begin(), args.end(), fn(2, &args[0]));
  args[0] = grid.front(); args[1] = grid.back();  adaptive_triang.insert(args.begin(), args.end(), fn(2, &args[0]));
  args[0] = grid.back();  args[1] = grid.front(); adaptive_triang.insert(args.begin(), args.end(), fn(2, &args[0]));
  args[0] = grid.back();  args[1] = grid.back();  adaptive_triang.insert(args.begin(), args.end(), fn(2, &args[0]));
  ErrorTuple top = adaptive_triang.get_largest_error_tuple();
  ErrorTuple got = adaptive_triang.insert_largest_error_point();
  CHECK(top.m_err == got.m_err && top.m_x == got.m_x);
  for (int i = 0; i < n_points * n_points - 5; i++) adaptive_triang.insert_largest_error_point();
  CHECK(adaptive_triang.number_of_vertices() == 100);
  vector<ErrorTuple> q = adaptive_triang.get_error_queue();
  for (size_t i = 0; i + 1 < q.size(); i++) CHECK(!(q[i] < q[i + 1]));
  CHECK(adaptive_triang.get_all_vertices().size() == 100);
  CHECK(adaptive_triang.get_edges().size() == 100 + adaptive_triang.number_of_faces() - 1);  // Euler: E = V + F - 1
  if (host_only) { printf("host-only checks ok (lo=%g)\n", lo); return 0; }
  {  // get_line_segments as the reference returns it (:842-862): flags (hull edges are always 1, :778-780) + min / max
    auto ls = adaptive_triang.get_line_segments();
    CHECK(std::get<0>(ls).size() == 100 + adaptive_triang.number_of_faces() - 1);
    size_t n_concave = 0;
    for (auto &sg : std::get<0>(ls)) n_concave += std::get<2>(sg);
    CHECK(n_concave >= 4 && n_concave <= std::get<0>(ls).size());
    CHECK(std::get<1>(ls)[0] == lo && std::get<2>(ls)[0] == 5.0 && std::get<1>(ls)[1] == lo && std::get<2>(ls)[1] == 5.0);
  }

  vector<double> interp_grid = linspace(lo, 5., 2 * n_points), true_f, reg, ada, xs;
  auto t0 = std::chrono::steady_clock::now();
  for (size_t i = 0; i < interp_grid.size(); i++)
    for (size_t j = 0; j < interp_grid.size(); j++) {
      array<double, 2> a = {interp_grid[i], interp_grid[j]};
      reg.push_back(triang.interp(a.begin(), a.end()));            // scalar API, one query per call (:78-83)
    }
  const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  printf("regular grid: %zu interpolations (one call each), %.3f ms\n", reg.size(), ms);
  // README.md:128-129 publishes 2 ms for these 400 calls; they are answered on the host (dli_host_small.h), no GPU call
  CHECK(ms < 2.0);
  for (size_t i = 0; i < interp_grid.size(); i++)
    for (size_t j = 0; j < interp_grid.size(); j++) {
      array<double, 2> a = {interp_grid[i], interp_grid[j]};
      true_f.push_back(fn(2, &a[0]));
      ada.push_back(adaptive_triang.interp(a.begin(), a.end()));
      xs.push_back(a[0]); xs.push_back(a[1]);
    }
  // the batched API (GPU) gives the same values: within the engine's 1e-13 bound in its default mode, and bit for bit
  // with the reference-order epilogue (the per-point host path runs the same code as the device's exact pass)
  vector<double> reg_b(true_f.size()), ada_b(true_f.size());
  vector<unsigned char> st(true_f.size());
  triang.interp_batch(xs.data(), true_f.size(), reg_b.data(), st.data());
  adaptive_triang.interp_batch(xs.data(), true_f.size(), ada_b.data());
  double sse1 = 0, sse2 = 0;
  for (size_t i = 0; i < true_f.size(); i++) {
    // (1e-14 absolute = a few ulps of the data scale max|f| = 4.3: a query exactly on a vertex / edge where f vanishes,
    // e.g. the corner (0, 5) with lo = 0, is answered from whichever incident face the walk reaches first -- exactly 0
    // from one face, 0 * f_i + 1e-17 * f_j from another)
    const double tol_abs = 1e-14;
    if (!(std::fabs(reg[i] - reg_b[i]) <= 1e-12 * std::fabs(reg[i]) + tol_abs && std::fabs(ada[i] - ada_b[i]) <= 1e-12 * std::fabs(ada[i]) + tol_abs))
      printf("query %zu (%.17g, %.17g): regular host %.17g batch %.17g | adaptive host %.17g batch %.17g | status %d\n", i,
             xs[2 * i], xs[2 * i + 1], reg[i], reg_b[i], ada[i], ada_b[i], (int)st[i]);
    CHECK(std::fabs(reg[i] - reg_b[i]) <= 1e-12 * std::fabs(reg[i]) + tol_abs && std::fabs(ada[i] - ada_b[i]) <= 1e-12 * std::fabs(ada[i]) + tol_abs);
    CHECK(st[i] == DLI_STATUS_OK);
    sse1 += (true_f[i] - reg[i]) * (true_f[i] - reg[i]);
    sse2 += (true_f[i] - ada[i]) * (true_f[i] - ada[i]);
  }
  {
    dli_options o;
    dli_options_init(&o);
    o.epilogue = DLI_EPILOGUE_REFERENCE;
    CHECK(dli_di2_set_options(adaptive_triang.handle(), &o) == DLI_OK);
    adaptive_triang.interp_batch(xs.data(), true_f.size(), ada_b.data());
    size_t same = 0;
    for (size_t i = 0; i < true_f.size(); i++) same += ada[i] == ada_b[i];
    // (queries exactly on an edge may be answered from either incident face: equal within rounding, not bit for bit)
    CHECK(same + 40 >= true_f.size());
    for (size_t i = 0; i < true_f.size(); i++) CHECK(std::fabs(ada[i] - ada_b[i]) <= 4e-16 * std::fabs(ada[i]) + 1e-14);
    CHECK(dli_di2_set_options(adaptive_triang.handle(), nullptr) == DLI_OK);
  }
  char b1[64], b2[64];
  snprintf(b1, sizeof b1, "%f", sse1);
  snprintf(b2, sizeof b2, "%f", sse2);
  printf("regular grid: sum of squared errors: %s\nadaptive grid: sum of squared errors: %s\n", b1, b2);
  CHECK(strcmp(b1, want_regular) == 0);
  if (want_adaptive) CHECK(strcmp(b2, want_adaptive) == 0);
  // outside the hull: projected, not an error (:530-545)
  CHECK(std::fabs(triang.interp_point(Point(-1.0, 2.5)) - triang.interp_point(Point(lo, 2.5))) < 1e-14);
  // gradient (:608-642): the face's gradient = slope of interp inside the face; batch == scalar; NaN outside
  {
    array<double, 2> g = triang.gradient(Point(1.3, 2.1));
    double q[2] = {1.3, 2.1}, gb[2];
    unsigned char gst;
    triang.gradient_batch(q, 1, gb, &gst);
    CHECK(g[0] == gb[0] && g[1] == gb[1] && gst == DLI_STATUS_OK);
    const double h = 1e-3, y0 = triang.interp_point(Point(1.3, 2.1));
    CHECK(std::fabs((triang.interp_point(Point(1.3 + h, 2.1)) - y0) / h - g[0]) < 1e-9);
    CHECK(std::fabs((triang.interp_point(Point(1.3, 2.1 + h)) - y0) / h - g[1]) < 1e-9);
    CHECK(std::isnan(triang.gradient(Point(-1.0, 2.5))[0]));
    double vq[2] = {grid[3], grid[4]};   // a grid vertex: mean over its fan
    triang.gradient_batch(vq, 1, gb, &gst);
    CHECK(gst == DLI_STATUS_ON_VERTEX && std::isfinite(gb[0]) && std::isfinite(gb[1]));
  }
  // no triangulation yet: std::invalid_argument (:546-551)
  Delaunay_incremental_interp_2 empty;
  bool threw = false;
  try { empty.interp_point(Point(0.5, 0.5)); } catch (std::invalid_argument const &) { threw = true; }
  CHECK(threw);
  return 0;
}

static void fn_batch(int nd, int64_t n, const double *xs, double *out, void *) {
  for (int64_t i = 0; i < n; i++) out[i] = pow(xs[i * nd], 0.3) * pow(xs[i * nd + 1], 0.6);
}
// k-at-a-time refinement through the header: k = 1 equals the sequential call; a batched round inserts several points
static int run_refine() {
  Delaunay_incremental_interp_2 a(fn), b(fn), c(fn);
  double corners[4][2] = {{0.01, 0.01}, {0.01, 5.0}, {5.0, 0.01}, {5.0, 5.0}};
  for (auto &p : corners) { a.insert(p, p + 2, fn(2, p)); b.insert(p, p + 2, fn(2, p)); c.insert(p, p + 2, fn(2, p)); }
  for (int i = 0; i < 50; i++) {
    ErrorTuple ta = a.insert_largest_error_point();
    vector<ErrorTuple> tb = b.insert_largest_error_points(1, fn_batch);
    CHECK(tb.size() == 1 && tb[0].m_err == ta.m_err && tb[0].m_f == ta.m_f);
  }
  CHECK(a.get_largest_error_tuple().m_err == b.get_largest_error_tuple().m_err);
  size_t total = 0, rounds = 0;
  while (total < 300) { total += c.insert_largest_error_points(32, fn_batch).size(); rounds++; }
  CHECK(rounds * 4 < total);
  return 0;
}

int main(int argc, char **argv) {
  bool host_only = argc > 1 && strcmp(argv[1], "--host") == 0;
  if (run_refine()) return 1;
  // README.md:130-131 (grid on [0,5]) and the src/delaunay_2_test2.cpp variant (grid on [0.01,5])
  if (run(0.0, "5.800002", "0.044019", host_only)) return 1;
  if (run(0.01, "1.909975", nullptr, host_only)) return 1;
  printf("ok\n");
  return 0;
}
