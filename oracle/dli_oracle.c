/*
 * dli_oracle.c -- CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's interp hot path (locate -> barycentric
 * -> weighted gather) and of its per-simplex error metric.  Nothing in the product
 * (delaunay_linterp_b200/, include/) includes, links or calls this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load the library built from it.
 *
 * PARITY STATUS: the reference's own implementation of this path cannot be compiled
 * here (it needs CGAL 4.2, Eigen and Boost, none vendored under /root/reference and
 * none installed).  The only pinned known answer the reference
 * publishes for this path is README.md:130 ("regular grid: sum of squared errors:
 * 5.800002"), which this oracle reproduces (tests/test_oracle_golden.py).  Beyond
 * that vector: **parity unpinned**.
 *
 * Reference lines followed (relative to /root/reference):
 *   2D driver            src/delaunay_2_interp.h:515-572   -> dlo_interp (dim == 2)
 *   2D volume / bary     src/delaunay_2_interp.h:195-219, 223-239
 *   2D gather            src/delaunay_2_interp.h:563-569
 *   dD driver            src/delaunay_d_interp.h:988-1010  -> dlo_interp (dim >= 3)
 *   dD volume / bary     src/delaunay_d_interp.h:601-624, 628-650
 *   dD locate semantics  modified_cgal/Delaunay_d.h:811-850, 793-808 (closed simplices, null outside)
 *   error metric         src/delaunay_2_interp.h:725-741, 166-192, 125-151, 108-121
 *   2D gradient          src/delaunay_2_interp.h:608-642, 574-606, 242-263  -> dlo_gradient
 *   boundary subsimplex  src/delaunay_2_interp.h:657-699, 331-342
 *   plot data            src/delaunay_2_interp.h:842-862 (get_line_segments), 766-803 (is_edge_concave)  -> dlo_line_segments
 * Third-party behaviour restated from its published algorithm (not in the tree):
 *   CGAL Triangulation_2::locate: exact-sign visibility walk, closed faces
 *   CGAL Line_2(P,Q) / projection (Cartesian, plain double constructions)
 *   Eigen MatrixXd::determinant() for dynamic sizes = PartialPivLU (unblocked, n <= 16)
 *   Eigen JacobiSVD best-fit subspace -> restated as a one-sided Jacobi SVD below
 *   GMP mpq_get_d: truncation toward zero
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -pthread).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <stdio.h>
#include <pthread.h>

#define DLO_MAXD 6               /* max supported dimension */
#define DLO_OK 0
#define DLO_OUTSIDE_PROJECTED 1  /* 2D: query was outside the hull, projected on a hull edge line */
#define DLO_OUTSIDE_NAN 2        /* dD: query outside the hull -> NaN */
#define DLO_STEP_LIMIT 3
#define DLO_NO_MESH 4

/* ------------------------------------------------------------------------- */
/* Exact arithmetic on floating-point expansions (Shewchuk 1997, restated).  */
/* An expansion is a sum of doubles, nonoverlapping, increasing magnitude.    */
/* ------------------------------------------------------------------------- */

static inline void two_sum(double a, double b, double *x, double *y) {
  double s = a + b;
  double bv = s - a;
  double av = s - bv;
  *x = s;
  *y = (a - av) + (b - bv);
}
static inline void two_prod(double a, double b, double *x, double *y) {
  double p = a * b;
  *x = p;
  *y = fma(a, b, -p);
}

/* h = e + b ; returns length of h (zero components eliminated). h may alias e. */
static int grow_expansion(int elen, const double *e, double b, double *h) {
  double Q = b, q;
  int hl = 0;
  for (int i = 0; i < elen; i++) {
    double Qn;
    two_sum(Q, e[i], &Qn, &q);
    Q = Qn;
    if (q != 0.0) h[hl++] = q;
  }
  if (Q != 0.0 || hl == 0) h[hl++] = Q;
  return hl;
}

/* h = e + f ; h must not alias f; h may alias e.  O(elen*flen), fine for an oracle. */
static int expansion_sum(int elen, const double *e, int flen, const double *f, double *h) {
  int hl = elen;
  if (h != e) memcpy(h, e, sizeof(double) * (size_t)elen);
  for (int i = 0; i < flen; i++) hl = grow_expansion(hl, h, f[i], h);
  return hl;
}

/* h = e * b ; h must not alias e.  Length <= 2*elen. */
static int scale_expansion(int elen, const double *e, double b, double *h) {
  /* every partial product is exact (two_prod); accumulate them with grow_expansion so that
     correctness does not depend on any ordering lemma */
  int hl = 0;
  h[0] = 0.0;
  hl = 1;
  for (int i = 0; i < elen; i++) {
    double p, q;
    two_prod(e[i], b, &p, &q);
    hl = grow_expansion(hl, h, q, h);
    hl = grow_expansion(hl, h, p, h);
  }
  return hl;
}

static int expansion_sign(int elen, const double *e) {
  for (int i = elen - 1; i >= 0; i--) {
    if (e[i] > 0.0) return 1;
    if (e[i] < 0.0) return -1;
  }
  return 0;
}

/* exact determinant of a k x k matrix of doubles (row-major, stride k) as an expansion.
   Laplace expansion along column 0; out must hold dlo_det_cap(k) doubles. */
static int dlo_det_cap(int k) {
  /* k! * 2^(k-1) terms is a safe upper bound for the recursive scheme below */
  int c = 1;
  for (int i = 2; i <= k; i++) c *= i;
  return c * (1 << (k > 0 ? k - 1 : 0)) + 4;
}

static int exact_det(int k, const double *m, double *out) {
  if (k == 1) {
    out[0] = m[0];
    return 1;
  }
  if (k == 2) {
    double p1, q1, p2, q2;
    two_prod(m[0], m[3], &p1, &q1);
    two_prod(m[1], m[2], &p2, &q2);
    out[0] = q1;
    int n = 1;
    n = grow_expansion(n, out, p1, out);
    n = grow_expansion(n, out, -q2, out);
    n = grow_expansion(n, out, -p2, out);
    return n;
  }
  int capm = dlo_det_cap(k - 1);
  double *minor = (double *)malloc(sizeof(double) * (size_t)((k - 1) * (k - 1)));
  double *md = (double *)malloc(sizeof(double) * (size_t)capm);
  double *sc = (double *)malloc(sizeof(double) * (size_t)(2 * capm + 4));
  int n = 1;
  out[0] = 0.0;
  for (int r = 0; r < k; r++) {
    double a = m[r * k + 0];
    if (a == 0.0) continue;
    int rr = 0;
    for (int i = 0; i < k; i++) {
      if (i == r) continue;
      for (int j = 1; j < k; j++) minor[rr * (k - 1) + (j - 1)] = m[i * k + j];
      rr++;
    }
    int ml = exact_det(k - 1, minor, md);
    int sl = scale_expansion(ml, md, (r & 1) ? -a : a, sc);
    n = expansion_sum(n, out, sl, sc, out);
  }
  free(minor);
  free(md);
  free(sc);
  return n;
}

/* Exact sign of the orientation determinant of d+1 points in d dims:
 *   | p0 1 |
 *   | .. 1 |   (homogeneous form: entries are the raw input doubles, hence exact)
 *   | pd 1 |
 * normalised so that in 2D a counter-clockwise triangle is positive.
 * pts: array of d+1 pointers to d doubles. */
static int orient_exact(int d, const double *const *pts) {
  int k = d + 1;
  double m[(DLO_MAXD + 1) * (DLO_MAXD + 1)];
  /* put the ones in column 0 so that Laplace expansion along it needs no multiplications */
  for (int i = 0; i < k; i++) {
    m[i * k] = 1.0;
    for (int j = 0; j < d; j++) m[i * k + 1 + j] = pts[i][j];
  }
  int cap = dlo_det_cap(k);
  double *e = (double *)malloc(sizeof(double) * (size_t)cap);
  int n = exact_det(k, m, e);
  int s = expansion_sign(n, e);
  free(e);
  /* det[1 p] (ones first) relates to det[p 1] (ones last) by d column swaps: sign (-1)^d.
     det[p 1] for a ccw triangle (d=2) is positive.  We normalise to det[p 1]. */
  return (d & 1) ? -s : s;
}

/* Semi-static floating-point filter in front of orient_exact.
 * Computes det of the d x d matrix of differences (p_i - p_d) by cofactor expansion in
 * doubles together with the same expression over absolute values (the "permanent").
 * |computed - true| <= c_d * eps * permanent with a generous c_d; if |det| is above that
 * bound its sign is certain.  (d=2: Shewchuk's (3+16eps)eps; larger d: see DESIGN.md.) */
static double det_and_perm(int k, const double *m, int stride, double *perm) {
  if (k == 1) {
    *perm = fabs(m[0]);
    return m[0];
  }
  if (k == 2) {
    double a = m[0] * m[stride + 1], b = m[1] * m[stride];
    *perm = fabs(a) + fabs(b);
    return a - b;
  }
  double sub[(DLO_MAXD) * (DLO_MAXD)];
  double det = 0.0, pm = 0.0;
  for (int r = 0; r < k; r++) {
    int rr = 0;
    for (int i = 0; i < k; i++) {
      if (i == r) continue;
      for (int j = 1; j < k; j++) sub[rr * (k - 1) + (j - 1)] = m[i * stride + j];
      rr++;
    }
    double sp;
    double sd = det_and_perm(k - 1, sub, k - 1, &sp);
    double t = m[r * stride] * sd;
    det += (r & 1) ? -t : t;
    pm += fabs(m[r * stride]) * sp;
  }
  *perm = pm;
  return det;
}

static int orient_sign(int d, const double *const *pts) {
  double m[DLO_MAXD * DLO_MAXD];
  for (int i = 0; i < d; i++)
    for (int j = 0; j < d; j++) m[i * d + j] = pts[i][j] - pts[d][j];
  double perm;
  double det = det_and_perm(d, m, d, &perm);
  /* det(p_i - p_d) = det[p 1]: subtracting the last row from the others and expanding
     along the ones column (entry (d,d), sign +) */
  static const double eps = 1.1102230246251565e-16; /* 2^-53 */
  /* generous constants: number of roundings on any path, doubled */
  static const double cd[DLO_MAXD + 1] = {0, 0, 8.0, 24.0, 64.0, 160.0, 384.0};
  double bound = cd[d] * eps * perm;
  if (det > bound) return 1;
  if (-det > bound) return -1;
  return orient_exact(d, pts);
}

/* test hooks */
int dlo_orient_exact(int d, const double *pts /* (d+1) x d row-major */) {
  const double *pp[DLO_MAXD + 1];
  for (int i = 0; i <= d; i++) pp[i] = pts + (size_t)i * d;
  return orient_exact(d, pp);
}
int dlo_orient_filtered(int d, const double *pts) {
  const double *pp[DLO_MAXD + 1];
  for (int i = 0; i <= d; i++) pp[i] = pts + (size_t)i * d;
  return orient_sign(d, pp);
}

/* ------------------------------------------------------------------------- */
/* tiny pthread parallel-for (static contiguous chunks)                      */
/* ------------------------------------------------------------------------- */
typedef void (*dlo_range_fn)(int64_t lo, int64_t hi, int tid, void *ctx);
typedef struct { dlo_range_fn fn; int64_t lo, hi; int tid; void *ctx; } dlo_task;
static void *dlo_task_main(void *a) {
  dlo_task *t = (dlo_task *)a;
  t->fn(t->lo, t->hi, t->tid, t->ctx);
  return NULL;
}
static void dlo_parallel_for(int64_t n, int nthreads, dlo_range_fn fn, void *ctx) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  if ((int64_t)nthreads > n) nthreads = n > 0 ? (int)n : 1;
  if (nthreads == 1) { fn(0, n, 0, ctx); return; }
  pthread_t th[256];
  dlo_task tk[256];
  for (int t = 0; t < nthreads; t++) {
    tk[t].fn = fn; tk[t].ctx = ctx; tk[t].tid = t;
    tk[t].lo = n * t / nthreads; tk[t].hi = n * (t + 1) / nthreads;
    if (t > 0) pthread_create(&th[t], NULL, dlo_task_main, &tk[t]);
  }
  dlo_task_main(&tk[0]);
  for (int t = 1; t < nthreads; t++) pthread_join(th[t], NULL);
}

/* ------------------------------------------------------------------------- */
/* Mesh                                                                      */
/* ------------------------------------------------------------------------- */

typedef struct dlo_mesh {
  int dim;
  int64_t nv, ns;
  double *x;      /* [nv][dim] */
  double *f;      /* [nv] */
  int32_t *sv;    /* [ns][dim+1], positively oriented */
  int32_t *sn;    /* [ns][dim+1], neighbour k is opposite vertex k, -1 = hull */
  int32_t start;  /* lowest-numbered simplex that has a hull facet (reference-faithful start) */
  /* 2D hull edge table: edge h = (face ht[h], slot hs[h]); P = v[slot+1], Q = v[slot+2] */
  int32_t nh;
  int32_t *ht, *hs, *hnext, *hprev;
  int32_t *hull_of; /* [ns][3] -> hull edge index or -1 (2D only) */
} dlo_mesh;

typedef struct {
  int32_t key[DLO_MAXD];
  int32_t s, slot;
} facet_rec;

static int g_fk = 0;
static int facet_cmp(const void *a, const void *b) {
  const facet_rec *x = (const facet_rec *)a, *y = (const facet_rec *)b;
  for (int i = 0; i < g_fk; i++) {
    if (x->key[i] < y->key[i]) return -1;
    if (x->key[i] > y->key[i]) return 1;
  }
  return 0;
}

static int derive_neighbors(dlo_mesh *m) {
  int d = m->dim, k = d + 1;
  int64_t nf = m->ns * k;
  facet_rec *fr = (facet_rec *)malloc(sizeof(facet_rec) * (size_t)nf);
  if (!fr) return -1;
  for (int64_t s = 0; s < m->ns; s++)
    for (int slot = 0; slot < k; slot++) {
      facet_rec *r = &fr[s * k + slot];
      int c = 0;
      for (int j = 0; j < k; j++)
        if (j != slot) r->key[c++] = m->sv[s * k + j];
      /* insertion sort of d ids */
      for (int a = 1; a < d; a++) {
        int32_t v = r->key[a];
        int b = a - 1;
        while (b >= 0 && r->key[b] > v) {
          r->key[b + 1] = r->key[b];
          b--;
        }
        r->key[b + 1] = v;
      }
      r->s = (int32_t)s;
      r->slot = slot;
    }
  g_fk = d;
  qsort(fr, (size_t)nf, sizeof(facet_rec), facet_cmp);
  for (int64_t i = 0; i < nf; i++) m->sn[(int64_t)fr[i].s * k + fr[i].slot] = -1;
  int rc = 0;
  for (int64_t i = 0; i + 1 < nf;) {
    if (facet_cmp(&fr[i], &fr[i + 1]) == 0) {
      if (i + 2 < nf && facet_cmp(&fr[i], &fr[i + 2]) == 0) rc = -2; /* non-manifold */
      m->sn[(int64_t)fr[i].s * k + fr[i].slot] = fr[i + 1].s;
      m->sn[(int64_t)fr[i + 1].s * k + fr[i + 1].slot] = fr[i].s;
      i += 2;
    } else
      i += 1;
  }
  free(fr);
  return rc;
}

void dlo_mesh_destroy(dlo_mesh *m) {
  if (!m) return;
  free(m->x); free(m->f); free(m->sv); free(m->sn);
  free(m->ht); free(m->hs); free(m->hnext); free(m->hprev); free(m->hull_of);
  free(m);
}

/* returns NULL on failure (degenerate simplex, bad index, non-manifold facet) */
dlo_mesh *dlo_mesh_create(int dim, int64_t nv, const double *coords, const double *f, int64_t ns,
                          const int32_t *sv, const int32_t *sn /* may be NULL */) {
  if (dim < 2 || dim > DLO_MAXD || nv < 0 || ns < 0) return NULL;
  int k = dim + 1;
  dlo_mesh *m = (dlo_mesh *)calloc(1, sizeof(dlo_mesh));
  m->dim = dim; m->nv = nv; m->ns = ns;
  m->x = (double *)malloc(sizeof(double) * (size_t)(nv * dim + 1));
  m->f = (double *)malloc(sizeof(double) * (size_t)(nv + 1));
  m->sv = (int32_t *)malloc(sizeof(int32_t) * (size_t)(ns * k + 1));
  m->sn = (int32_t *)malloc(sizeof(int32_t) * (size_t)(ns * k + 1));
  memcpy(m->x, coords, sizeof(double) * (size_t)(nv * dim));
  memcpy(m->f, f, sizeof(double) * (size_t)nv);
  memcpy(m->sv, sv, sizeof(int32_t) * (size_t)(ns * k));
  for (int64_t i = 0; i < ns * k; i++)
    if (m->sv[i] < 0 || m->sv[i] >= nv) { dlo_mesh_destroy(m); return NULL; }
  if (sn) memcpy(m->sn, sn, sizeof(int32_t) * (size_t)(ns * k));
  else if (derive_neighbors(m) != 0) { dlo_mesh_destroy(m); return NULL; }
  /* canonical orientation: positive; a negative simplex gets slots 0 and 1 exchanged
     (vertices and the neighbours opposite them). */
  for (int64_t s = 0; s < ns; s++) {
    const double *pp[DLO_MAXD + 1];
    for (int j = 0; j < k; j++) pp[j] = m->x + (size_t)m->sv[s * k + j] * dim;
    int o = orient_sign(dim, pp);
    if (o == 0) { dlo_mesh_destroy(m); return NULL; }
    if (o < 0) {
      int32_t t = m->sv[s * k]; m->sv[s * k] = m->sv[s * k + 1]; m->sv[s * k + 1] = t;
      t = m->sn[s * k]; m->sn[s * k] = m->sn[s * k + 1]; m->sn[s * k + 1] = t;
    }
  }
  m->start = -1;
  for (int64_t s = 0; s < ns && m->start < 0; s++)
    for (int j = 0; j < k; j++)
      if (m->sn[s * k + j] < 0) { m->start = (int32_t)s; break; }
  if (m->start < 0 && ns > 0) m->start = 0;
  if (dim == 2) {
    int32_t nh = 0;
    for (int64_t i = 0; i < ns * 3; i++) nh += (m->sn[i] < 0);
    m->nh = nh;
    m->ht = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nh + 1));
    m->hs = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nh + 1));
    m->hnext = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nh + 1));
    m->hprev = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nh + 1));
    m->hull_of = (int32_t *)malloc(sizeof(int32_t) * (size_t)(ns * 3 + 1));
    int32_t *from = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nv + 1));
    int32_t *to = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nv + 1));
    for (int64_t i = 0; i < nv; i++) from[i] = to[i] = -1;
    int32_t h = 0;
    for (int64_t s = 0; s < ns; s++)
      for (int j = 0; j < 3; j++) {
        m->hull_of[s * 3 + j] = -1;
        if (m->sn[s * 3 + j] < 0) {
          m->ht[h] = (int32_t)s; m->hs[h] = j; m->hull_of[s * 3 + j] = h;
          from[m->sv[s * 3 + (j + 1) % 3]] = h; /* edge h starts at P */
          to[m->sv[s * 3 + (j + 2) % 3]] = h;   /* edge h ends at Q */
          h++;
        }
      }
    for (h = 0; h < nh; h++) {
      int32_t P = m->sv[m->ht[h] * 3 + (m->hs[h] + 1) % 3];
      int32_t Q = m->sv[m->ht[h] * 3 + (m->hs[h] + 2) % 3];
      m->hnext[h] = from[Q];
      m->hprev[h] = to[P];
    }
    free(from); free(to);
  }
  return m;
}

int dlo_mesh_set_values(dlo_mesh *m, const double *f) {
  memcpy(m->f, f, sizeof(double) * (size_t)m->nv);
  return 0;
}
int32_t dlo_mesh_start(const dlo_mesh *m) { return m->start; }
void dlo_mesh_get_simplices(const dlo_mesh *m, int32_t *sv, int32_t *sn) {
  int k = m->dim + 1;
  if (sv) memcpy(sv, m->sv, sizeof(int32_t) * (size_t)(m->ns * k));
  if (sn) memcpy(sn, m->sn, sizeof(int32_t) * (size_t)(m->ns * k));
}

/* ------------------------------------------------------------------------- */
/* Locate: exact-sign visibility walk over closed simplices                  */
/* ------------------------------------------------------------------------- */

/* Walk from simplex `s` towards p.  Returns the simplex whose closed hull contains p
 * (*exit_slot = -1), or the hull simplex through whose hull facet *exit_slot the point is
 * strictly visible (outside the convex hull).  *steps counts visited simplices. */
static int32_t walk(const dlo_mesh *m, const double *p, int32_t s, int *exit_slot, int64_t *steps,
                    int64_t max_steps) {
  int d = m->dim, k = d + 1;
  int32_t from = -1;
  const double *pp[DLO_MAXD + 1];
  for (int64_t it = 0; it < max_steps; it++) {
    (*steps)++;
    const int32_t *v = m->sv + (size_t)s * k;
    const int32_t *nb = m->sn + (size_t)s * k;
    int moved = 0;
    for (int j = 0; j < k; j++) {
      if (nb[j] >= 0 && nb[j] == from) continue; /* we entered through this facet */
      for (int i = 0; i < k; i++) pp[i] = (i == j) ? p : m->x + (size_t)v[i] * d;
      if (orient_sign(d, pp) < 0) {
        if (nb[j] < 0) { *exit_slot = j; return s; }
        from = s; s = nb[j]; moved = 1;
        break;
      }
    }
    if (!moved) { *exit_slot = -1; return s; }
  }
  *exit_slot = -2;
  return s;
}

/* Eigen dynamic-size determinant: PartialPivLU (unblocked), det = sign * prod(diag), sequential. */
static double lu_det(int n, double *a /* n x n row-major, destroyed */) {
  int sgn = 1;
  for (int c = 0; c < n; c++) {
    int piv = c;
    double best = fabs(a[c * n + c]);
    for (int r = c + 1; r < n; r++) {
      double v = fabs(a[r * n + c]);
      if (v > best) { best = v; piv = r; }
    }
    if (a[piv * n + c] != 0.0) {
      if (piv != c) {
        for (int j = 0; j < n; j++) { double t = a[c * n + j]; a[c * n + j] = a[piv * n + j]; a[piv * n + j] = t; }
        sgn = -sgn;
      }
      for (int r = c + 1; r < n; r++) a[r * n + c] = a[r * n + c] / a[c * n + c];
    }
    for (int r = c + 1; r < n; r++)
      for (int j = c + 1; j < n; j++) a[r * n + j] = a[r * n + j] - a[r * n + c] * a[c * n + j];
  }
  double prod = a[0];
  for (int i = 1; i < n; i++) prod = prod * a[i * n + i];
  return (double)sgn * prod;
}

/* exact difference a-b rounded toward zero (what Gmpq subtraction + mpq_get_d gives) */
static double sub_trunc(double a, double b) {
  double s, e;
  two_sum(a, -b, &s, &e);
  if ((s > 0.0 && e < 0.0) || (s < 0.0 && e > 0.0)) s = nextafter(s, 0.0);
  return s;
}

/* signed volume of the simplex pts[0..n] in n dims; column j = pts[j+1] - pts[0]; / n! */
static double volume(int n, const double *const *pts, int trunc_diff) {
  double a[DLO_MAXD * DLO_MAXD];
  for (int j = 0; j < n; j++)
    for (int i = 0; i < n; i++)
      a[i * n + j] = trunc_diff ? sub_trunc(pts[j + 1][i], pts[0][i]) : (pts[j + 1][i] - pts[0][i]);
  double det = lu_det(n, a);
  int fact = 1;
  for (int i = 1; i <= n; i++) fact *= i;
  return det / (double)fact;
}

static void barycentric(int n, const double *const *pts, const double *p, int trunc_diff, double *b) {
  double whole = volume(n, pts, trunc_diff);
  const double *q[DLO_MAXD + 1];
  for (int i = 0; i <= n; i++) {
    for (int j = 0; j <= n; j++) q[j] = (j == i) ? p : pts[j];
    b[i] = volume(n, q, trunc_diff) / whole;
  }
}

/* exact sum of products b[i]*f[i], truncated toward zero to a double (Gmpq sum + mpq_get_d) */
static double dot_exact_trunc(int n, const double *b, const double *f) {
  double e[4 * (DLO_MAXD + 2)], t[4 * (DLO_MAXD + 2)];
  int el = 1;
  e[0] = 0.0;
  for (int i = 0; i < n; i++) {
    double p, q;
    two_prod(b[i], f[i], &p, &q);
    el = grow_expansion(el, e, q, e);
    el = grow_expansion(el, e, p, e);
  }
  double A = 0.0;
  for (int i = el - 1; i >= 0; i--) A += e[i];
  if (!isfinite(A)) return A;
  /* fix up: largest-magnitude double not exceeding |total| */
  for (int it = 0; it < 8; it++) {
    int tl = grow_expansion(el, e, -A, t);           /* total - A */
    int sd = expansion_sign(tl, t);
    int sa = (A > 0) - (A < 0);
    if (sd == 0) break;
    if (sa != 0 && sd == -sa) { A = nextafter(A, 0.0); continue; }   /* |A| > |total| */
    /* |A| < |total| (or A == 0): can we move away from zero without passing total? */
    double up = nextafter(A, sd > 0 ? INFINITY : -INFINITY);
    tl = grow_expansion(el, e, -up, t);
    int su = expansion_sign(tl, t);
    if (su == 0 || su == sd) { A = up; if (su == 0) break; continue; }
    break;
  }
  return A;
}

/* CGAL Cartesian Line_2(P,Q) + projection, plain double constructions */
static void project_on_line(const double *P, const double *Q, double *p) {
  double a, b, c;
  if (P[1] == Q[1]) {
    a = 0.0;
    if (Q[0] > P[0]) { b = 1.0; c = -P[1]; }
    else if (Q[0] == P[0]) { b = 0.0; c = 0.0; }
    else { b = -1.0; c = P[1]; }
  } else if (Q[0] == P[0]) {
    b = 0.0;
    if (Q[1] > P[1]) { a = -1.0; c = P[0]; }
    else { a = 1.0; c = -P[0]; }
  } else {
    a = P[1] - Q[1];
    b = Q[0] - P[0];
    c = -P[0] * a - P[1] * b;
  }
  double x, y;
  if (a == 0.0) { x = p[0]; y = -c / b; }
  else if (b == 0.0) { x = -c / a; y = p[1]; }
  else {
    double a2 = a * a, b2 = b * b, d = a2 + b2;
    x = (b2 * p[0] - a * b * p[1] - a * c) / d;
    y = (-a * b * p[0] + a2 * p[1] - b * c) / d;
  }
  p[0] = x; p[1] = y;
}

/* Among the hull edges strictly visible from p (a contiguous chain containing h0) pick the one whose supporting
   line is farthest from p (largest det^2 / |PQ|^2 in plain fp64, ties: smallest index): a canonical stand-in for
   "whichever infinite face CGAL's walk ended in" (src/delaunay_2_interp.h:530-545), which is walk-dependent in the
   reference.  Farthest-line-first guarantees progress: after a projection p sits on the line of the edge it was
   projected on, within rounding, and that edge may still test "strictly visible" by an ulp; preferring it again
   (e.g. a smallest-index rule) would re-project p on the same line forever while the edge that p is really
   outside of is never chosen -- the reference escapes that loop through CGAL's stochastic walk. */
static double hull_edge_score(const dlo_mesh *m, const double *p, int32_t h) {
  int32_t t = m->ht[h], s = m->hs[h];
  const double *P = m->x + (size_t)m->sv[t * 3 + (s + 1) % 3] * 2;
  const double *Q = m->x + (size_t)m->sv[t * 3 + (s + 2) % 3] * 2;
  double dx = Q[0] - P[0], dy = Q[1] - P[1];
  double det = dx * (p[1] - P[1]) - dy * (p[0] - P[0]);
  return (det * det) / (dx * dx + dy * dy);
}
static int32_t canonical_visible_hull_edge(const dlo_mesh *m, const double *p, int32_t h0) {
  int32_t best = h0;
  double best_score = hull_edge_score(m, p, h0);
  const double *pp[3];
  for (int dir = 0; dir < 2; dir++) {
    int32_t h = h0;
    for (int32_t it = 0; it < m->nh; it++) {
      h = dir ? m->hprev[h] : m->hnext[h];
      if (h < 0 || h == h0) break;
      int32_t t = m->ht[h], s = m->hs[h];
      pp[0] = m->x + (size_t)m->sv[t * 3 + (s + 1) % 3] * 2;
      pp[1] = m->x + (size_t)m->sv[t * 3 + (s + 2) % 3] * 2;
      pp[2] = p;
      if (orient_sign(2, pp) >= 0) break;
      double sc = hull_edge_score(m, p, h);
      if (sc > best_score || (sc == best_score && h < best)) { best = h; best_score = sc; }
    }
  }
  return best;
}

#define DLO_MAX_PROJECTIONS 16

/* one query.  start: simplex to start the walk from. */
static double interp_one(const dlo_mesh *m, const double *x, int32_t start, int32_t *simplex_out,
                         uint8_t *status_out, int64_t *steps) {
  int d = m->dim, k = d + 1;
  double p[DLO_MAXD];
  for (int i = 0; i < d; i++) p[i] = x[i];
  uint8_t status = DLO_OK;
  int32_t s = start;
  int64_t max_steps = m->ns + 16;
  if (m->ns <= 0) { *simplex_out = -1; *status_out = DLO_NO_MESH; return NAN; }
  for (int proj = 0;; proj++) {
    int slot;
    s = walk(m, p, s, &slot, steps, max_steps);
    if (slot == -1) break;
    if (slot == -2) { *simplex_out = s; *status_out = DLO_STEP_LIMIT; return NAN; }
    if (d != 2) { /* src/delaunay_d_interp.h:1005-1008 */
      *simplex_out = -1; *status_out = DLO_OUTSIDE_NAN; return NAN;
    }
    status = DLO_OUTSIDE_PROJECTED;
    if (proj >= DLO_MAX_PROJECTIONS) break; /* accept the hull face (ulp-level extrapolation) */
    int32_t h = canonical_visible_hull_edge(m, p, m->hull_of[s * 3 + slot]);
    int32_t t = m->ht[h], hs = m->hs[h];
    project_on_line(m->x + (size_t)m->sv[t * 3 + (hs + 1) % 3] * 2,
                    m->x + (size_t)m->sv[t * 3 + (hs + 2) % 3] * 2, p);
    s = t;
  }
  const double *pts[DLO_MAXD + 1];
  double b[DLO_MAXD + 1], fv[DLO_MAXD + 1];
  for (int i = 0; i < k; i++) {
    pts[i] = m->x + (size_t)m->sv[(size_t)s * k + i] * d;
    fv[i] = m->f[m->sv[(size_t)s * k + i]];
  }
  double result;
  if (d == 2) {
    barycentric(2, pts, p, 0, b);
    double sum = 0.0;
    for (int i = 0; i < 3; i++) sum = sum + b[i] * fv[i];
    result = sum;
  } else {
    barycentric(d, pts, p, 1, b);
    result = dot_exact_trunc(k, b, fv);
  }
  *simplex_out = s;
  *status_out = status;
  return result;
}

/* mode 0: reference-faithful, every query walks from the fixed hull-adjacent start simplex
 *         (what locate(p, lt, li) without a hint does, src/delaunay_2_interp.h:525)
 * mode 1: walk from the simplex the previous query of the same thread ended in (same result for
 *         interior queries, much shorter walks; used by tests to keep run time down)
 * nthreads <= 1: serial.  total_steps (may be NULL) receives the number of simplices visited. */
typedef struct {
  const dlo_mesh *m; const double *xs; double *ys; int32_t *so; uint8_t *st; int mode;
  int64_t steps[256];
} interp_ctx;
static void interp_range(int64_t lo, int64_t hi, int tid, void *vctx) {
  interp_ctx *c = (interp_ctx *)vctx;
  const dlo_mesh *m = c->m;
  int32_t last = m->start;
  int64_t steps = 0;
  for (int64_t i = lo; i < hi; i++) {
    int32_t so; uint8_t st;
    double y = interp_one(m, c->xs + (size_t)i * m->dim, c->mode == 1 ? last : m->start, &so, &st, &steps);
    if (c->ys) c->ys[i] = y;
    if (c->so) c->so[i] = so;
    if (c->st) c->st[i] = st;
    if (so >= 0) last = so;
  }
  c->steps[tid] = steps;
}
int dlo_interp(const dlo_mesh *m, int64_t n, const double *xs, double *ys, int32_t *simplex_out,
               uint8_t *status_out, int mode, int nthreads, int64_t *total_steps) {
  interp_ctx c;
  memset(&c, 0, sizeof(c));
  c.m = m; c.xs = xs; c.ys = ys; c.so = simplex_out; c.st = status_out; c.mode = mode;
  dlo_parallel_for(n, nthreads, interp_range, &c);
  int64_t s = 0;
  for (int t = 0; t < 256; t++) s += c.steps[t];
  if (total_steps) *total_steps = s;
  return 0;
}

/* ------------------------------------------------------------------------- */
/* 2D gradient (src/delaunay_2_interp.h:608-642)                              */
/* ------------------------------------------------------------------------- */
#define DLO_ON_EDGE 5
#define DLO_ON_VERTEX 6

/* simplex_gradient (:242-263) as called by set_simplex_gradient (:574-588): V(j,i) = P[i+1][j] - P[0][j],
 * df(i) = f[i+1] - f[0], gradient = transpose(inverse(V)) * df.  CGAL's Linear_algebraCd<double>::inverse is
 * not in the tree; restated through the adjugate: inverse(V) = adj(V) / det(V). */
static void face_gradient(const dlo_mesh *m, int32_t s, double *g) {
  const int32_t *v = m->sv + (size_t)s * 3;
  const double *A = m->x + (size_t)v[0] * 2, *B = m->x + (size_t)v[1] * 2, *C = m->x + (size_t)v[2] * 2;
  double fa = m->f[v[0]], fb = m->f[v[1]], fc = m->f[v[2]];
  double a = B[0] - A[0], c = B[1] - A[1], b = C[0] - A[0], d = C[1] - A[1]; /* V = [[a, b], [c, d]] */
  double df0 = fb - fa, df1 = fc - fa;
  double det = a * d - c * b;
  g[0] = (df0 * d - df1 * c) / det;
  g[1] = (a * df1 - b * df0) / det;
}
static int slot_of(const dlo_mesh *m, int32_t s, int32_t v) {
  const int32_t *sv = m->sv + (size_t)s * 3;
  return sv[0] == v ? 0 : (sv[1] == v ? 1 : 2);
}
/* VERTEX case (:619-632): get_avg_simplex_gradient (:594-606) over the incident finite faces = get_centroid
 * (:108-121): sequential sum from 0, then / count.  CGAL's circulator starts at vertex->face(), which is
 * arbitrary; canonical order used here (and by the CUDA path): counter-clockwise, a closed fan starting from
 * its lowest-numbered face, an open fan (hull vertex) from its clockwise-most face. */
static void vertex_avg_gradient(const dlo_mesh *m, int32_t f0, int32_t v, double *g) {
  int32_t start = f0, best = f0, f = f0;
  int open = 0;
  for (int64_t it = 0; it < m->ns + 16; it++) {
    int32_t nb = m->sn[(size_t)f * 3 + (slot_of(m, f, v) + 2) % 3]; /* clockwise neighbour */
    if (nb < 0) { open = 1; start = f; break; }
    if (nb == f0) break;
    f = nb;
    if (f < best) { best = f; start = f; }
  }
  double sx = 0.0, sy = 0.0;
  int cnt = 0;
  f = start;
  for (int64_t it = 0; it < m->ns + 16; it++) {
    double gf[2];
    face_gradient(m, f, gf);
    sx += gf[0]; sy += gf[1]; cnt++;
    int32_t nb = m->sn[(size_t)f * 3 + (slot_of(m, f, v) + 1) % 3]; /* counter-clockwise neighbour */
    if (nb < 0 || (!open && nb == start)) break;
    f = nb;
  }
  g[0] = sx / cnt; g[1] = sy / cnt;
}
static void gradient_one(const dlo_mesh *m, const double *x, int32_t start, double *g, int32_t *simplex_out,
                         uint8_t *status_out, int64_t *steps) {
  g[0] = g[1] = NAN;
  if (m->ns <= 0) { *simplex_out = -1; *status_out = DLO_NO_MESH; return; }
  int slot;
  int32_t s = walk(m, x, start, &slot, steps, m->ns + 16);
  if (slot == -2) { *simplex_out = s; *status_out = DLO_STEP_LIMIT; return; }
  if (slot != -1) { *simplex_out = -1; *status_out = DLO_OUTSIDE_NAN; return; } /* :618: not found -> NaN */
  const int32_t *v = m->sv + (size_t)s * 3;
  int zero[3], nz = 0;
  const double *pp[3];
  for (int j = 0; j < 3; j++) {
    for (int i = 0; i < 3; i++) pp[i] = (i == j) ? x : m->x + (size_t)v[i] * 2;
    zero[j] = (orient_sign(2, pp) == 0);
    nz += zero[j];
  }
  *simplex_out = s;
  if (nz == 0) { face_gradient(m, s, g); *status_out = DLO_OK; return; }
  if (nz == 1) { /* EDGE (:633-636) */
    int li = zero[0] ? 0 : (zero[1] ? 1 : 2);
    double g1[2], g2[2];
    face_gradient(m, s, g1);
    int32_t nb = m->sn[(size_t)s * 3 + li];
    if (nb >= 0) {
      face_gradient(m, nb, g2);
      g[0] = (g1[0] + g2[0]) / 2; g[1] = (g1[1] + g2[1]) / 2;
    } else { g[0] = g1[0]; g[1] = g1[1]; }
    *status_out = DLO_ON_EDGE;
    return;
  }
  int k = !zero[0] ? 0 : (!zero[1] ? 1 : 2);
  vertex_avg_gradient(m, s, v[k], g);
  *status_out = DLO_ON_VERTEX;
}
/* serial; mode as dlo_interp */
int dlo_gradient(const dlo_mesh *m, int64_t n, const double *xs, double *gs, int32_t *simplex_out,
                 uint8_t *status_out, int mode) {
  if (m->dim != 2) return -1;
  int32_t last = m->start;
  int64_t steps = 0;
  for (int64_t i = 0; i < n; i++) {
    int32_t so; uint8_t st;
    gradient_one(m, xs + 2 * (size_t)i, mode == 1 ? last : m->start, gs + 2 * (size_t)i, &so, &st, &steps);
    if (simplex_out) simplex_out[i] = so;
    if (status_out) status_out[i] = st;
    if (so >= 0) last = so;
  }
  return 0;
}

/* is p inside the closed simplex s?  exact. returns 1 inside (incl. boundary), 0 outside */
int dlo_contains(const dlo_mesh *m, int32_t s, const double *p) {
  int d = m->dim, k = d + 1;
  const double *pp[DLO_MAXD + 1];
  for (int j = 0; j < k; j++) {
    for (int i = 0; i < k; i++) pp[i] = (i == j) ? p : m->x + (size_t)m->sv[(size_t)s * k + i] * d;
    if (orient_sign(d, pp) < 0) return 0;
  }
  return 1;
}
/* is p strictly interior to simplex s? */
int dlo_strictly_inside(const dlo_mesh *m, int32_t s, const double *p) {
  int d = m->dim, k = d + 1;
  const double *pp[DLO_MAXD + 1];
  for (int j = 0; j < k; j++) {
    for (int i = 0; i < k; i++) pp[i] = (i == j) ? p : m->x + (size_t)m->sv[(size_t)s * k + i] * d;
    if (orient_sign(d, pp) <= 0) return 0;
  }
  return 1;
}
void dlo_classify(const dlo_mesh *m, int64_t n, const double *xs, const int32_t *simplex,
                  uint8_t *contains, uint8_t *strict) {
  for (int64_t i = 0; i < n; i++) {
    if (simplex[i] < 0) { contains[i] = 0; strict[i] = 0; continue; }
    contains[i] = (uint8_t)dlo_contains(m, simplex[i], xs + (size_t)i * m->dim);
    strict[i] = (uint8_t)dlo_strictly_inside(m, simplex[i], xs + (size_t)i * m->dim);
  }
}

/* ------------------------------------------------------------------------- */
/* Error metric (src/delaunay_2_interp.h:725-741 and helpers)                 */
/* ------------------------------------------------------------------------- */

/* one-sided Jacobi SVD: right singular vectors of A (r x c, row-major) sorted by decreasing
   singular value, returned as columns of V (c x c).  Restates what Eigen::JacobiSVD with
   ComputeFullV delivers (src/delaunay_2_interp.h:136-138) up to sign/rotation inside
   equal-singular-value subspaces, which the projector V_k V_k^T does not see when the
   subspace is well separated. */
static void svd_V(int r, int c, const double *A, double *V, double *sv) {
  /* Hestenes one-sided Jacobi: rotate column pairs of U = A V until mutually orthogonal */
  double U[(DLO_MAXD + 1) * (DLO_MAXD + 1)];
  memcpy(U, A, sizeof(double) * (size_t)(r * c));
  for (int i = 0; i < c; i++)
    for (int j = 0; j < c; j++) V[i * c + j] = (i == j);
  for (int sweep = 0; sweep < 60; sweep++) {
    int rotated = 0;
    for (int p = 0; p < c; p++)
      for (int q = p + 1; q < c; q++) {
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int t = 0; t < r; t++) {
          al += U[t * c + p] * U[t * c + p];
          be += U[t * c + q] * U[t * c + q];
          ga += U[t * c + p] * U[t * c + q];
        }
        if (ga == 0.0 || fabs(ga) <= 1e-17 * sqrt(al * be)) continue;
        rotated = 1;
        double ze = (be - al) / (2.0 * ga);
        double t = (ze >= 0 ? 1.0 : -1.0) / (fabs(ze) + sqrt(1.0 + ze * ze));
        double cs = 1.0 / sqrt(1.0 + t * t), sn = cs * t;
        for (int k = 0; k < r; k++) {
          double up = U[k * c + p], uq = U[k * c + q];
          U[k * c + p] = cs * up - sn * uq;
          U[k * c + q] = sn * up + cs * uq;
        }
        for (int k = 0; k < c; k++) {
          double vp = V[k * c + p], vq = V[k * c + q];
          V[k * c + p] = cs * vp - sn * vq;
          V[k * c + q] = sn * vp + cs * vq;
        }
      }
    if (!rotated) break;
  }
  double nrm[DLO_MAXD + 1];
  int idx[DLO_MAXD + 1];
  for (int j = 0; j < c; j++) {
    double s = 0.0;
    for (int t = 0; t < r; t++) s += U[t * c + j] * U[t * c + j];
    nrm[j] = sqrt(s);
    idx[j] = j;
  }
  for (int i = 0; i < c; i++)
    for (int j = i + 1; j < c; j++)
      if (nrm[idx[j]] > nrm[idx[i]]) { int t = idx[i]; idx[i] = idx[j]; idx[j] = t; }
  double Vs[(DLO_MAXD + 1) * (DLO_MAXD + 1)];
  for (int j = 0; j < c; j++) {
    sv[j] = nrm[idx[j]];
    for (int i = 0; i < c; i++) Vs[i * c + j] = V[i * c + idx[j]];
  }
  memcpy(V, Vs, sizeof(double) * (size_t)(c * c));
}

/* squared distance in R^{N} from point q to the best-fit (np-1)-dim hyperplane through the np
   points w (each N doubles): src/delaunay_2_interp.h:166-192 */
static double sqdist_point_hplane(int N, const double *q, int np, const double *w) {
  double cen[DLO_MAXD + 1];
  for (int j = 0; j < N; j++) cen[j] = 0.0;
  for (int i = 0; i < np; i++)
    for (int j = 0; j < N; j++) cen[j] += w[i * N + j];
  for (int j = 0; j < N; j++) cen[j] /= np;
  double A[(DLO_MAXD + 1) * (DLO_MAXD + 1)], V[(DLO_MAXD + 1) * (DLO_MAXD + 1)], sv[DLO_MAXD + 1];
  for (int i = 0; i < np; i++)
    for (int j = 0; j < N; j++) A[i * N + j] = w[i * N + j] - cen[j];
  double tp[DLO_MAXD + 1];
  for (int j = 0; j < N; j++) tp[j] = q[j] - cen[j];
  svd_V(np, N, A, V, sv);
  int kd = np - 1;
  /* p2 = V_k V_k^T tp */
  double coef[DLO_MAXD + 1], sq = 0.0;
  for (int c = 0; c < kd; c++) {
    double s = 0.0;
    for (int j = 0; j < N; j++) s += V[j * N + c] * tp[j];
    coef[c] = s;
  }
  for (int j = 0; j < N; j++) {
    double pj = 0.0;
    for (int c = 0; c < kd; c++) pj += V[j * N + c] * coef[c];
    double dd = tp[j] - pj;
    sq += dd * dd;
  }
  return sq;
}

/* Sub-simplex enumeration.  For every simplex: the simplex itself; for every hull facet of it,
 * every vertex subset of size >= 2 of that facet (duplicates across simplices are NOT removed,
 * src/delaunay_2_interp.h:683-699).  Item order: simplices ascending; within a simplex the
 * full simplex first, then hull facets by ascending slot, subsets by ascending bitmask over
 * the facet's vertices in slot order.
 * Pass 1 (verts == NULL): returns the item count.
 * Pass 2: fills item_simplex[n], item_nv[n], item_verts[n][dim+1] (-1 padded), centroids[n][dim]. */
int64_t dlo_error_items(const dlo_mesh *m, int32_t *item_simplex, int32_t *item_nv, int32_t *item_verts,
                        double *centroids) {
  int d = m->dim, k = d + 1;
  int64_t n = 0;
  for (int64_t s = 0; s < m->ns; s++) {
    for (int item = -1; item < k; item++) {
      int32_t fv[DLO_MAXD + 1];
      int nf = 0;
      if (item == -1) {
        for (int j = 0; j < k; j++) fv[nf++] = m->sv[s * k + j];
      } else {
        if (m->sn[s * k + item] >= 0) continue;
        for (int j = 0; j < k; j++)
          if (j != item) fv[nf++] = m->sv[s * k + j];
      }
      int lo = (item == -1) ? (1 << nf) - 1 : 0;
      for (int mask = lo; mask < (1 << nf); mask++) {
        int cnt = __builtin_popcount((unsigned)mask);
        if (cnt < 2) continue;
        if (item_verts) {
          int c = 0;
          double cen[DLO_MAXD];
          for (int j = 0; j < d; j++) cen[j] = 0.0;
          for (int b = 0; b < nf; b++)
            if (mask >> b & 1) {
              item_verts[n * k + c++] = fv[b];
              for (int j = 0; j < d; j++) cen[j] += m->x[(size_t)fv[b] * d + j];
            }
          for (; c < k; c++) item_verts[n * k + c] = -1;
          for (int j = 0; j < d; j++) centroids[n * d + j] = cen[j] / cnt;
          item_simplex[n] = (int32_t)s;
          item_nv[n] = cnt;
        }
        n++;
      }
    }
  }
  return n;
}

/* err[i] for items produced by dlo_error_items, given phi[i] = f(centroid_i).
   src/delaunay_2_interp.h:725-741 (SVD form, as the reference computes it). */
void dlo_error_eval(const dlo_mesh *m, int64_t n, const int32_t *item_nv, const int32_t *item_verts,
                    const double *centroids, const double *phi, double *err) {
  int d = m->dim, k = d + 1, N = d + 1;
  for (int64_t i = 0; i < n; i++) {
    double w[(DLO_MAXD + 1) * (DLO_MAXD + 1)], q[DLO_MAXD + 1];
    int np = item_nv[i];
    for (int a = 0; a < np; a++) {
      int32_t v = item_verts[i * k + a];
      for (int j = 0; j < d; j++) w[a * N + j] = m->x[(size_t)v * d + j];
      w[a * N + d] = m->f[v];
    }
    for (int j = 0; j < d; j++) q[j] = centroids[i * d + j];
    q[d] = phi[i];
    err[i] = sqdist_point_hplane(N, q, np, w);
  }
}

/* ------------------------------------------------------------------------- */
/* Plot data: get_line_segments (src/delaunay_2_interp.h:842-862) with the    */
/* concavity flag of is_edge_concave (:766-803).  2D only.                    */
/* ------------------------------------------------------------------------- */

/* CGAL Cartesian Line_2(P, Q) coefficients in plain doubles (the construction project_on_line uses) */
static void line_from_points(const double *P, const double *Q, double *a, double *b, double *c) {
  if (P[1] == Q[1]) {
    *a = 0.0;
    if (Q[0] > P[0]) { *b = 1.0; *c = -P[1]; }
    else if (Q[0] == P[0]) { *b = 0.0; *c = 0.0; }
    else { *b = -1.0; *c = P[1]; }
  } else if (Q[0] == P[0]) {
    *b = 0.0;
    if (Q[1] > P[1]) { *a = -1.0; *c = P[0]; }
    else { *a = 1.0; *c = -P[0]; }
  } else {
    *a = P[1] - Q[1];
    *b = Q[0] - P[0];
    *c = -P[0] * *a - P[1] * *b;
  }
}
/* Line_2::oriented_side: exact sign of a x + b y + c for the double coefficients (filtered-exact in Epick) */
static int line_side(double a, double b, double c, double x, double y) {
  double e[8], p, q;
  int n = 0;
  two_prod(a, x, &p, &q);
  n = grow_expansion(n, e, q, e);
  n = grow_expansion(n, e, p, e);
  two_prod(b, y, &p, &q);
  n = grow_expansion(n, e, q, e);
  n = grow_expansion(n, e, p, e);
  n = grow_expansion(n, e, c, e);
  return expansion_sign(n, e);
}
/* is_edge_concave for the edge opposite slot j of face s (face1 = s, face2 = its neighbour) */
static int edge_concave(const dlo_mesh *m, int32_t s, int j) {
  int32_t nb = m->sn[(size_t)s * 3 + j];
  if (nb < 0) return 1; /* an infinite face on either side (:778-780) */
  const int32_t *v = m->sv + (size_t)s * 3;
  const double *P = m->x + (size_t)v[(j + 1) % 3] * 2, *Q = m->x + (size_t)v[(j + 2) % 3] * 2;
  const double *R = m->x + (size_t)v[j] * 2;
  double g1[2], g2[2];
  face_gradient(m, s, g1);
  face_gradient(m, nb, g2);
  double dgx = g2[0] - g1[0], dgy = g2[1] - g1[1];
  double a, b, c;
  line_from_points(P, Q, &a, &b, &c);
  int side1 = line_side(a, b, c, R[0], R[1]);
  int side_left = line_side(a, b, c, P[0] - 1.0, P[1]);
  int side_down = line_side(a, b, c, P[0], P[1] - 1.0);
  if (side_left == 0) dgx = 0.0; else if (side_left != side1) dgx = -dgx;
  if (side_down == 0) dgy = 0.0; else if (side_down != side1) dgy = -dgy;
  return (dgx <= 0.0 && dgy <= 0.0) ? 1 : 0;
}
/* Every finite edge once, from its lower-numbered face (CGAL's edge iterator order is arbitrary): segments
 * [ne][2][3] = (P.x, P.y, f_P, Q.x, Q.y, f_Q) with P = vertex ccw(j), Q = vertex cw(j) of that face; concave [ne];
 * min/max [3] over the segment end points (:853-858).  Returns ne; arrays may be NULL. */
int64_t dlo_line_segments(const dlo_mesh *m, double *seg, int32_t *concave, double *mn, double *mx) {
  if (m->dim != 2) return -1;
  int64_t ne = 0;
  if (mn) for (int k = 0; k < 3; k++) { mn[k] = DBL_MAX; mx[k] = -DBL_MAX; }
  for (int64_t s = 0; s < m->ns; s++)
    for (int j = 0; j < 3; j++) {
      int32_t nb = m->sn[s * 3 + j];
      if (nb >= 0 && nb < s) continue;
      int32_t p = m->sv[s * 3 + (j + 1) % 3], q = m->sv[s * 3 + (j + 2) % 3];
      double row[6] = {m->x[(size_t)p * 2], m->x[(size_t)p * 2 + 1], m->f[p], m->x[(size_t)q * 2], m->x[(size_t)q * 2 + 1], m->f[q]};
      if (seg) memcpy(seg + 6 * ne, row, sizeof row);
      if (concave) concave[ne] = edge_concave(m, (int32_t)s, j);
      if (mn) for (int k = 0; k < 6; k++) { if (row[k] < mn[k % 3]) mn[k % 3] = row[k]; if (row[k] > mx[k % 3]) mx[k % 3] = row[k]; }
      ne++;
    }
  return ne;
}
