/*
 * dli_error_math.h -- the per-(sub)simplex error metric in closed form, shared by the CUDA
 * kernel (k_error_finish) and the host-side adaptive insertion queue.
 *
 * Replaces compute_simplex_error's SVD (src/delaunay_2_interp.h:725-741 -> sqdist_point_hplane_of_points
 * :166-192 -> best_fit_subspace :125-139 -> sqdist_point_subspace :143-151; dD src/delaunay_d_interp.h:1078-1093):
 * for np lifted vertices w_i = (v_i, f_i) in R^{D+1} and the point (c, phi) above their centroid c,
 * the squared orthogonal distance to the affine hull of the w_i is
 *     err = (phi - mean f)^2 * | e_z - proj_span(e_z) |^2,   span = span{w_i - w_0}.
 */
#ifndef DLI_ERROR_MATH_H_
#define DLI_ERROR_MATH_H_

#include <math.h>

#if defined(__CUDACC__)
#define DLI_EM_HD __host__ __device__
#else
#define DLI_EM_HD
#endif

/* w: np x (D+1) row-major lifted vertices (coords then f).  Returns err. */
template <int D>
DLI_EM_HD inline double dli_simplex_error(int np, const double *w, double phi) {
  const int N = D + 1;
  double fbar = 0.0;
  for (int a = 0; a < np; a++) fbar += w[a * N + D];
  fbar /= np;
  double q[D * N];
  int nq = 0;
  for (int j = 1; j < np; j++) {
    double u[N];
    for (int k = 0; k < N; k++) u[k] = w[j * N + k] - w[k];
    for (int pass = 0; pass < 2; pass++) /* modified Gram-Schmidt, re-orthogonalised once */
      for (int t = 0; t < nq; t++) {
        double d = 0.0;
        for (int k = 0; k < N; k++) d += u[k] * q[t * N + k];
        for (int k = 0; k < N; k++) u[k] -= d * q[t * N + k];
      }
    double nn = 0.0;
    for (int k = 0; k < N; k++) nn += u[k] * u[k];
    if (nn > 0.0) {
      double inv = 1.0 / sqrt(nn);
      for (int k = 0; k < N; k++) q[nq * N + k] = u[k] * inv;
      nq++;
    }
  }
  double r[N];
  for (int k = 0; k < N; k++) r[k] = (k == D) ? 1.0 : 0.0;
  for (int pass = 0; pass < 2; pass++)
    for (int t = 0; t < nq; t++) {
      double d = 0.0;
      for (int k = 0; k < N; k++) d += r[k] * q[t * N + k];
      for (int k = 0; k < N; k++) r[k] -= d * q[t * N + k];
    }
  double fac = 0.0;
  for (int k = 0; k < N; k++) fac += r[k] * r[k];
  double delta = phi - fbar;
  return delta * delta * fac;
}

#endif /* DLI_ERROR_MATH_H_ */
