// C++ port of the reference's demo driver src/delaunay_d_test2.cpp:33-115 (N = 2, as the reference runs it, and the
// same program for N = 3), with the assertions the reference never had.  Uses the drop-in header
// include/delaunay_d_interp.h; link with libdli.so.
//   ./test_delaunay_d            full run (needs a GPU: interp_batch goes through the CUDA engine)
//   ./test_delaunay_d --host     host-side part only (insert / adaptive queue / per-point interp, which is
//                                answered on the host)
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <vector>

#include "delaunay_d_interp.h"

using std::array;
using std::vector;

static vector<double> linspace(double first, double last, int len) {
  vector<double> result(len);
  double step = (last - first) / (len - 1);
  for (int i = 0; i < len; i++) result[i] = first + i * step;
  return result;
}
// the function to interpolate (src/delaunay_d_test2.cpp:30; the third exponent extends it to N = 3)
static double fn(int n, double *x) {
  double r = pow(x[0], 0.3) * pow(x[1], 0.6);
  if (n > 2) r *= pow(x[2], 0.2);
  return r;
}

#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) { printf("CHECK FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); return 1; } \
  } while (0)

// all points of a regular grid with n_points per axis in N dimensions, first axis outermost
static void grid_points(int N, const vector<double> &g, vector<vector<double>> &out) {
  size_t total = 1;
  for (int k = 0; k < N; k++) total *= g.size();
  for (size_t t = 0; t < total; t++) {
    vector<double> x((size_t)N);
    size_t r = t;
    for (int k = N - 1; k >= 0; k--) { x[(size_t)k] = g[r % g.size()]; r /= g.size(); }
    out.push_back(x);
  }
}

static int run(int N, bool host_only) {
  const int n_points = N == 2 ? 10 : 5;
  // first, a rectangular grid (:44-53)
  Delaunay_incremental_interp_d triang(N);
  CHECK(triang.get_dimension() == N);
  vector<double> grid = linspace(0.01, 5., n_points);
  vector<vector<double>> gp;
  grid_points(N, grid, gp);
  for (auto &x : gp) triang.insert(x.begin(), x.end(), fn(N, x.data()));
  CHECK(triang.number_of_vertices() == gp.size());
  CHECK(dli_dd_validate(triang.handle(), 1) == 0);
  // a triangulated box: the simplex volumes add up to the box (N! simplices per cell at least ... exactly N! per cell
  // when every cell is split along a diagonal; in any case >= cells)
  size_t cells = 1;
  for (int k = 0; k < N; k++) cells *= (size_t)(n_points - 1);
  CHECK(triang.number_of_simplices() >= (N == 2 ? 2 : 5) * cells);

  // second, adaptive point placement (:55-71): the corners of the box, then insert_largest_error_point
  std::function<double(int, double *)> fn_obj(fn);
  Delaunay_incremental_interp_d adaptive(N, fn_obj);
  for (int c = 0; c < (1 << N); c++) {
    vector<double> x((size_t)N);
    for (int k = 0; k < N; k++) x[(size_t)k] = (c >> k & 1) ? grid.back() : grid.front();
    adaptive.insert(x.begin(), x.end(), fn(N, x.data()));
  }
  CHECK(adaptive.current_dimension() == N);
  ErrorTuple top = adaptive.get_largest_error_tuple();
  ErrorTuple got = adaptive.insert_largest_error_point();
  CHECK(top.m_err == got.m_err && top.m_x == got.m_x && top.m_f == got.m_f);
  const size_t target = gp.size();
  while (adaptive.number_of_vertices() < target) {
    const size_t before = adaptive.number_of_vertices();
    adaptive.insert_largest_error_point();
    CHECK(adaptive.number_of_vertices() == before + 1);
  }
  CHECK(dli_dd_validate(adaptive.handle(), 1) == 0);
  vector<ErrorTuple> q = adaptive.get_error_queue();
  CHECK(!q.empty());
  for (size_t i = 0; i + 1 < q.size(); i++) CHECK(!(q[i] < q[i + 1]));
  CHECK(q[0].m_err == adaptive.get_largest_error_tuple().m_err);
  CHECK(adaptive.get_all_vertices().size() == target && adaptive.get_all_vertices()[0].size() == (size_t)N + 1);
  {
    auto ls = adaptive.get_line_segments();
    CHECK(!std::get<0>(ls).empty());
    CHECK(std::get<1>(ls)[0] == grid.front() && std::get<2>(ls)[0] == grid.back());
    for (auto &sg : std::get<0>(ls)) CHECK(std::get<2>(sg) == 1 && std::get<0>(sg).size() == (size_t)N + 1);
  }

  // compare interpolated value vs. actual function (:73-100), one interp() per point as the reference does
  vector<double> interp_grid = linspace(0.01 + 0.005, 5. - 0.005, 2 * n_points);
  vector<vector<double>> qp;
  grid_points(N, interp_grid, qp);
  vector<double> true_f, reg, ada, xs;
  // (one untimed call first: on a GPU box the library's first CUDA runtime query -- "is this a device pointer?" -- creates
  // the CUDA context, a one-time ~40 ms that is not part of the per-call cost)
  (void)triang.interp(qp[0].begin(), qp[0].end());
  auto t0 = std::chrono::steady_clock::now();
  for (auto &x : qp) {
    true_f.push_back(fn(N, x.data()));
    reg.push_back(triang.interp(x.begin(), x.end()));
    ada.push_back(adaptive.interp(x.begin(), x.end()));
    xs.insert(xs.end(), x.begin(), x.end());
  }
  const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  printf("N=%d: %zu + %zu per-point interpolations in %.3f ms\n", N, qp.size(), qp.size(), ms);
  // the reference publishes ~200 kq/s for this loop (README.md:128-129: 400 interpolations in 2 ms); the drop-in must
  // not be slower (generous bound: 10 us per call)
  CHECK(ms < 0.01 * 2 * (double)qp.size() + 5.0);
  double sse1 = 0, sse2 = 0;
  for (size_t i = 0; i < true_f.size(); i++) {
    CHECK(std::isfinite(reg[i]) && std::isfinite(ada[i]));
    sse1 += (true_f[i] - reg[i]) * (true_f[i] - reg[i]);
    sse2 += (true_f[i] - ada[i]) * (true_f[i] - ada[i]);
  }
  printf("N=%d regular grid: sum of squared errors: %f\nN=%d adaptive grid: sum of squared errors: %f\n", N, sse1, N, sse2);
  if (N == 2) {
    // the same grid and queries through the 2D class give 1.909975 for the regular grid (tests/cpp/test_delaunay_2.cpp,
    // SURVEY 8(c)) on the queries linspace(0.01, 5, 20); here the queries are the reference dD driver's (:74). The
    // regular-grid interpolant does not depend on how each grid square is split only where f is bilinear; accept the
    // published order of magnitude and require the adaptive grid to beat the regular one, as the reference shows
    CHECK(sse1 > 0.5 && sse1 < 5.0);
    // Tie the dD arithmetic (N + 2 LU determinants of Gmpq differences, src/delaunay_d_interp.h:601-650, 988-1010) to
    // the 2D class, whose branch the README's two known answers pin (src/delaunay_2_interp.h:515-572): the same points
    // in general position -- so that the Delaunay triangulation is unique and both classes hold the same triangles --
    // through both classes, compared query by query.  (The two drop-in headers both define the reference's global names
    // Point / ErrorTuple and cannot share a translation unit, as in the reference; the 2D class is driven through the
    // dli_di2_* calls its header forwards to.)
    Delaunay_incremental_interp_d dd(2);
    dli_di2 *d2 = dli_di2_create(nullptr, nullptr);
    CHECK(d2 != nullptr);
    unsigned long long lcg = 88172645463325252ull;
    auto rnd = [&lcg]() {
      lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
      return (double)(lcg >> 11) / 9007199254740992.0;
    };
    for (int i = 0; i < 400; i++) {
      array<double, 2> x = {0.01 + 4.99 * rnd(), 0.01 + 4.99 * rnd()};
      const double f = fn(2, x.data());
      dd.insert(x.begin(), x.end(), f);
      CHECK(dli_di2_insert(d2, x[0], x[1], f) >= 0);
    }
    int compared = 0;
    double worst = 0.0;
    for (int i = 0; i < 2000; i++) {
      array<double, 2> q = {0.01 + 4.99 * rnd(), 0.01 + 4.99 * rnd()};
      const double a = dd.interp(q.begin(), q.end());
      if (std::isnan(a)) continue;                       // outside the hull: the dD class answers NaN, the 2D class projects
      double b = 0.0;
      CHECK(dli_di2_interp(d2, 1, q.data(), &b, nullptr, nullptr, nullptr) == 0);
      const double rel = std::fabs(a - b) / std::fmax(std::fabs(b), 1e-300);
      if (rel > worst) worst = rel;
      if (rel > 1e-12) printf("dD(2) vs 2D class: q = (%.17g, %.17g): %.17g vs %.17g\n", q[0], q[1], a, b);
      compared++;
    }
    printf("N=2: dD class vs 2D class on %d queries over 400 random points: max rel diff %.3g\n", compared, worst);
    dli_di2_destroy(d2);
    CHECK(compared > 1500 && worst <= 1e-12);
  }
  CHECK(sse2 < sse1);
  // outside the hull: quiet NaN, no exception (:1005-1008)
  {
    vector<double> out((size_t)N, 2.5);
    out[0] = -1.0;
    CHECK(std::isnan(triang.interp(out.begin(), out.end())));
  }
  // fewer than N + 1 affinely independent points: NaN as well (locate finds no simplex)
  {
    Delaunay_incremental_interp_d flat(N);
    for (int i = 0; i < 5; i++) {
      vector<double> x((size_t)N, 0.25 * i);   // collinear
      flat.insert(x.begin(), x.end(), 1.0);
    }
    CHECK(flat.current_dimension() == 1 && flat.number_of_simplices() == 0);
    vector<double> x((size_t)N, 0.3);
    CHECK(std::isnan(flat.interp(x.begin(), x.end())));
    // the point that leaves the line (N = 2) / the plane starts the triangulation with everything seen before
    vector<double> y((size_t)N, 0.0);
    y[0] = 1.0;
    flat.insert(y.begin(), y.end(), 1.0);
    if (N == 2) {
      CHECK(flat.current_dimension() == 2 && flat.number_of_simplices() == 4);
      vector<double> z = {0.5, 0.3};
      CHECK(std::fabs(flat.interp(z.begin(), z.end()) - 1.0) < 1e-14);
    }
  }
  // duplicate insert overwrites the value
  {
    const size_t nv = triang.number_of_vertices();
    triang.insert(gp[7].begin(), gp[7].end(), 123.0);
    CHECK(triang.number_of_vertices() == nv);
    CHECK(triang.interp(gp[7].begin(), gp[7].end()) == 123.0);
    triang.insert(gp[7].begin(), gp[7].end(), fn(N, gp[7].data()));
  }
  Point lp = get_lifted_point(Point(N, gp[3].begin(), gp[3].end()));
  CHECK(lp.dimension() == N + 1);
  if (host_only) { printf("host-only checks ok (N=%d)\n", N); return 0; }

  // batched API (GPU) agrees with the per-point results within the engine's bound, statuses OK
  vector<double> reg_b(true_f.size()), ada_b(true_f.size());
  vector<unsigned char> st(true_f.size());
  triang.interp_batch(xs.data(), true_f.size(), reg_b.data(), st.data());
  adaptive.interp_batch(xs.data(), true_f.size(), ada_b.data());
  for (size_t i = 0; i < true_f.size(); i++) {
    CHECK(st[i] == DLI_STATUS_OK);
    CHECK(std::fabs(reg[i] - reg_b[i]) <= 1e-12 * std::fabs(reg[i]));
    CHECK(std::fabs(ada[i] - ada_b[i]) <= 1e-12 * std::fabs(ada[i]));
  }
  // outside the hull in a batch: NaN + DLI_STATUS_OUTSIDE_NAN in every dimension (the 2D engine projects by default)
  {
    // (repeated so that the batch is large enough to go to the GPU; small host batches are answered on the host)
    const size_t reps = 700;
    vector<double> q2((size_t)N * 3 * reps, 2.5);
    for (size_t r = 0; r < reps; r++) { q2[r * 3 * N] = -1.0; q2[r * 3 * N + (size_t)N + 1] = 7.0; }
    vector<double> y(3 * reps);
    vector<unsigned char> s3(3 * reps);
    triang.interp_batch(q2.data(), 3 * reps, y.data(), s3.data());
    for (size_t r = 0; r < reps; r++) {
      CHECK(std::isnan(y[3 * r]) && std::isnan(y[3 * r + 1]) && std::isfinite(y[3 * r + 2]));
      CHECK(s3[3 * r] == DLI_STATUS_OUTSIDE_NAN && s3[3 * r + 1] == DLI_STATUS_OUTSIDE_NAN && s3[3 * r + 2] == DLI_STATUS_OK);
    }
  }
  return 0;
}

int main(int argc, char **argv) {
  bool host_only = argc > 1 && strcmp(argv[1], "--host") == 0;
  if (run(2, host_only)) return 1;
  if (run(3, host_only)) return 1;
  printf("ok\n");
  return 0;
}
