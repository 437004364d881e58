// dli_kernels.cuh -- __global__ kernels (sm_100a): batched interp (fast + straggler/exact pass),
// hint-grid construction, per-simplex error evaluator, predicate test hook.
#pragma once
#include "dli_device.cuh"
#include "../../include/dli_error_math.h"

namespace dli {

// stats words in device memory
enum { ST_STEPS = 0, ST_MAXSTEPS = 1, ST_EXACT = 2, ST_OUTSIDE = 3, ST_HARD = 4, ST_WORDS = 8 };

__device__ __forceinline__ void flush_stats(unsigned long long *g, const StatsAcc &st) {
  // warp-aggregate, one atomic per warp per counter
  unsigned long long s = st.steps, e = st.exact, o = st.outside;
  unsigned int mx = st.max_steps;
  unsigned mask = __activemask();
  for (int off = 16; off > 0; off >>= 1) {
    s += __shfl_down_sync(mask, s, off);
    e += __shfl_down_sync(mask, e, off);
    o += __shfl_down_sync(mask, o, off);
    mx = max(mx, __shfl_down_sync(mask, mx, off));
  }
  // with a partial mask shfl_down from an inactive lane returns the caller's own value: only use the
  // aggregated value when the whole warp is here, otherwise fall back to per-thread atomics
  if (mask == 0xffffffffu) {
    if ((threadIdx.x & 31) == 0) {
      if (s) atomicAdd(g + ST_STEPS, s);
      if (e) atomicAdd(g + ST_EXACT, e);
      if (o) atomicAdd(g + ST_OUTSIDE, o);
      atomicMax(g + ST_MAXSTEPS, (unsigned long long)mx);
    }
  } else {
    if (st.steps) atomicAdd(g + ST_STEPS, st.steps);
    if (st.exact) atomicAdd(g + ST_EXACT, st.exact);
    if (st.outside) atomicAdd(g + ST_OUTSIDE, st.outside);
    atomicMax(g + ST_MAXSTEPS, (unsigned long long)st.max_steps);
  }
}

// ---------------------------------------------------------------------------------------------
// 2D interp
// ---------------------------------------------------------------------------------------------

// Fast pass.  Each warp owns a contiguous range of `per_warp` queries and runs a lane-level state
// machine: every loop iteration performs exactly ONE walk step for every lane; a lane whose query
// finished (epilogue done, or handed to the straggler queue) is refilled with the next unclaimed
// query of the warp's range at the top of the next iteration (ballot/popc rank, no atomics, no
// shared memory).  Walk-length variance therefore costs no SIMT efficiency in the step body.
// Predicates are filter-only here; queries it cannot certify (filter failure, hull exit, long walk)
// are appended to `queue` for k_interp2_full.  If the queue overflows (`queue_cap`), the counter
// keeps growing and k_interp2_full re-runs the whole batch (degenerate inputs only).
template <bool STATS, int EPI, bool PACKED>
__global__ void __launch_bounds__(256, 4)
k_interp2_fast(Mesh2View m, int64_t n, int per_warp, const double2 *__restrict__ xs, double *__restrict__ ys,
               int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out, int32_t *__restrict__ queue,
               unsigned int *__restrict__ queue_count, unsigned int queue_cap, unsigned long long *__restrict__ stats) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * per_warp;
  if (base >= n) return;  // warp-uniform
  const int64_t end = min(n, base + (int64_t)per_warp);
  int64_t next = base;
  bool active = false;
  double px = 0.0, py = 0.0;
  int f = 0, from = -1;
  int64_t qi = 0;
  unsigned int steps = 0;
  const double fc = filter_const(2);
  StatsAcc st;
  for (;;) {
    unsigned need = __ballot_sync(FULL, !active);
    if (need) {
      int64_t idx = next + __popc(need & ((1u << lane) - 1u));
      if (!active && idx < end) {
        double2 q = __ldcs(xs + idx);
        px = q.x; py = q.y;
        qi = idx;
        f = hint2(m, px, py);
        from = -1;
        steps = 0;
        active = true;
      }
      next += __popc(need);
    }
    if (!__any_sync(FULL, active)) break;
    if (active) {
      steps++;
      Face2 F = PACKED ? load_face_packed(m, f) : load_face(m, f);
      double4 A = load_vert(m, F.v0), B = load_vert(m, F.v1), C = load_vert(m, F.v2);
      double ax = A.x - px, ay = A.y - py, bx = B.x - px, by = B.y - py, cx = C.x - px, cy = C.y - py;
      // edge opposite v0 = (B, C); opposite v1 = (C, A); opposite v2 = (A, B)
      double l0 = bx * cy, r0 = by * cx, d0 = l0 - r0, e0 = fc * (fabs(l0) + fabs(r0));
      double l1 = cx * ay, r1 = cy * ax, d1 = l1 - r1, e1 = fc * (fabs(l1) + fabs(r1));
      double l2 = ax * by, r2 = ay * bx, d2 = l2 - r2, e2 = fc * (fabs(l2) + fabs(r2));
      bool c0 = (F.n0 >= 0) & (F.n0 == from), c1 = (F.n1 >= 0) & (F.n1 == from), c2 = (F.n2 >= 0) & (F.n2 == from);
      bool neg0 = (-d0 > e0) & !c0, neg1 = (-d1 > e1) & !c1, neg2 = (-d2 > e2) & !c2;
      bool give_up = false;
      if (neg0 | neg1 | neg2) {
        int nb = neg0 ? F.n0 : (neg1 ? F.n1 : F.n2);
        if (nb < 0 || (int)steps >= m.fast_steps) give_up = true;  // hull exit / straggler
        else { from = f; f = nb; }
      } else {
        bool pos = ((d0 > e0) | c0) & ((d1 > e1) | c1) & ((d2 > e2) | c2);
        if (pos) {
          if (ys) {
            double y = (EPI == 0) ? epilogue2_ref(px, py, A, B, C) : epilogue2_fast(px, py, A, B, C);
            __stcs(ys + qi, y);
          }
          if (simplex_out) __stcs(simplex_out + qi, f);
          if (status_out) status_out[qi] = DLI_STATUS_OK;
          if (STATS) { st.steps += steps; st.max_steps = max(st.max_steps, steps); }
          active = false;
        } else {
          give_up = true;  // a decision the static filter cannot certify
        }
      }
      if (give_up) {
        unsigned int slot = atomicAdd(queue_count, 1u);
        if (slot < queue_cap) queue[slot] = (int32_t)qi;
        active = false;
      }
    }
  }
  if (STATS) flush_stats(stats, st);
}

// Straggler / exact pass over the queue written by the fast pass (count read on the device: no
// host synchronisation between the two launches).
template <bool STATS, int EPI>
__global__ void __launch_bounds__(128)
k_interp2_full(Mesh2View m, const double2 *__restrict__ xs, double *__restrict__ ys,
               int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out,
               const int32_t *__restrict__ queue, const unsigned int *__restrict__ queue_count,
               unsigned int queue_cap, int64_t n /* batch size: processed directly if the queue overflowed */,
               unsigned long long *__restrict__ stats) {
  int64_t cnt = (int64_t)*queue_count;
  const bool direct = cnt > (int64_t)queue_cap;
  if (direct) cnt = n;
  StatsAcc st;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < cnt; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = direct ? k : (int64_t)queue[k];
    double2 q = xs[i];
    double px = q.x, py = q.y;
    int f = hint2(m, px, py);
    double4 A, B, C;
    uint8_t status = locate2_full(m, px, py, f, A, B, C, st);
    double y;
    if (status == DLI_STATUS_STEP_LIMIT) {
      y = __longlong_as_double(0x7ff8000000000000LL);
    } else {
      y = (EPI == 0) ? epilogue2_ref(px, py, A, B, C) : epilogue2_fast(px, py, A, B, C);
    }
    if (ys) ys[i] = y;
    if (simplex_out) simplex_out[i] = f;
    if (status_out) status_out[i] = status;
  }
  if (STATS) {
    flush_stats(stats, st);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(stats + ST_HARD, (unsigned long long)cnt);
  }
}

// ---------------------------------------------------------------------------------------------
// dD interp
// ---------------------------------------------------------------------------------------------

// dD fast pass: the same warp-persistent lane state machine as k_interp2_fast (one walk step per lane per
// iteration, ballot/popc refill), all D+1 facet orientations of a step from shared minors (OrientAll<D>).
template <int D, bool STATS, int EPI>
__global__ void __launch_bounds__(128)
k_interpD_fast(MeshDView m, int64_t n, int per_warp, const double *__restrict__ xs, double *__restrict__ ys,
               int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out, int32_t *__restrict__ queue,
               unsigned int *__restrict__ queue_count, unsigned int queue_cap, unsigned long long *__restrict__ stats) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * per_warp;
  if (base >= n) return;  // warp-uniform
  const int64_t end = min(n, base + (int64_t)per_warp);
  int64_t next = base;
  bool active = false;
  double p[D];
#pragma unroll
  for (int k = 0; k < D; k++) p[k] = 0.0;
  int s = 0, from = -1;
  int64_t qi = 0;
  unsigned int steps = 0;
  const double fc = filter_const(D);
  StatsAcc st;
  for (;;) {
    unsigned need = __ballot_sync(FULL, !active);
    if (need) {
      int64_t idx = next + __popc(need & ((1u << lane) - 1u));
      if (!active && idx < end) {
#pragma unroll
        for (int k = 0; k < D; k++) p[k] = __ldcs(xs + idx * D + k);
        qi = idx;
        s = hintD<D>(m, p);
        from = -1;
        steps = 0;
        active = true;
      }
      next += __popc(need);
    }
    if (!__any_sync(FULL, active)) break;
    if (active) {
      steps++;
      SimplexD<D> S;
      load_simplex_ids<D>(m, s, S);
      load_simplex_verts<D>(m, S);
      double R[(D + 1) * D], o[D + 1], pm[D + 1];
#pragma unroll
      for (int i = 0; i <= D; i++)
#pragma unroll
        for (int k = 0; k < D; k++) R[i * D + k] = S.x[i * D + k] - p[k];
      OrientAll<D>::eval(R, o, pm);
      int mv = -1;
      bool all_pos = true;
#pragma unroll
      for (int j = D; j >= 0; j--) {
        bool came = (S.n[j] >= 0) & (S.n[j] == from);
        double e = fc * pm[j];
        if (!came) {
          if (-o[j] > e) mv = j;               // lowest such j wins (loop runs downwards)
          else if (!(o[j] > e)) all_pos = false;
        }
      }
      bool give_up = false;
      if (mv >= 0) {
        int nb = S.n[0];
#pragma unroll
        for (int j = 1; j <= D; j++) nb = (mv == j) ? S.n[j] : nb;
        if (nb < 0 || (int)steps >= m.fast_steps) give_up = true;
        else { from = s; s = nb; }
      } else if (all_pos) {
        if (ys) {
          double y = (EPI == 0) ? epilogueD_ref<D>(p, S) : epilogueD_fast<D>(p, S);
          __stcs(ys + qi, y);
        }
        if (simplex_out) __stcs(simplex_out + qi, s);
        if (status_out) status_out[qi] = DLI_STATUS_OK;
        if (STATS) { st.steps += steps; st.max_steps = max(st.max_steps, steps); }
        active = false;
      } else {
        give_up = true;
      }
      if (give_up) {
        unsigned int slot = atomicAdd(queue_count, 1u);
        if (slot < queue_cap) queue[slot] = (int32_t)qi;
        active = false;
      }
    }
  }
  if (STATS) flush_stats(stats, st);
}

template <int D, bool STATS, int EPI>
__global__ void __launch_bounds__(64)
k_interpD_full(MeshDView m, const double *__restrict__ xs, double *__restrict__ ys,
               int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out,
               const int32_t *__restrict__ queue, const unsigned int *__restrict__ queue_count,
               unsigned int queue_cap, int64_t n, unsigned long long *__restrict__ stats) {
  int64_t cnt = (int64_t)*queue_count;
  const bool direct = cnt > (int64_t)queue_cap;
  if (direct) cnt = n;
  StatsAcc st;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < cnt; k += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = direct ? k : (int64_t)queue[k];
    double p[D];
#pragma unroll
    for (int a = 0; a < D; a++) p[a] = xs[i * D + a];
    int s = hintD<D>(m, p);
    SimplexD<D> S;
    uint8_t status = locateD_full<D>(m, p, s, S, st);
    double y;
    if (status != DLI_STATUS_OK) {
      y = __longlong_as_double(0x7ff8000000000000LL);
      if (status == DLI_STATUS_OUTSIDE_NAN) s = -1;
    } else {
      y = (EPI == 0) ? epilogueD_ref<D>(p, S) : epilogueD_fast<D>(p, S);
    }
    if (ys) ys[i] = y;
    if (simplex_out) simplex_out[i] = s;
    if (status_out) status_out[i] = status;
  }
  if (STATS) {
    flush_stats(stats, st);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(stats + ST_HARD, (unsigned long long)cnt);
  }
}

// 32-byte face records -> 16-byte packed records (see Mesh2View::fpack)
__global__ void k_pack_faces(int ns, const int4 *__restrict__ frec, ulonglong2 *__restrict__ fpack) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= ns) return;
  int4 a = frec[2 * (size_t)f], b = frec[2 * (size_t)f + 1];
  unsigned long long n0 = a.w < 0 ? kPackMask : (unsigned)a.w, n1 = b.x < 0 ? kPackMask : (unsigned)b.x,
                     n2 = b.y < 0 ? kPackMask : (unsigned)b.y;
  ulonglong2 r;
  r.x = (unsigned long long)(unsigned)a.x | ((unsigned long long)(unsigned)a.y << 21) | ((unsigned long long)(unsigned)a.z << 42);
  r.y = n0 | (n1 << 21) | (n2 << 42);
  fpack[f] = r;
}

__global__ void k_fill_nomesh(int64_t n, double *ys, int32_t *simplex_out, uint8_t *status_out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (ys) ys[i] = __longlong_as_double(0x7ff8000000000000LL);
  if (simplex_out) simplex_out[i] = -1;
  if (status_out) status_out[i] = DLI_STATUS_NO_MESH;
}

// ---------------------------------------------------------------------------------------------
// Hint grid: cell -> a simplex containing (or near) the cell centre.  Hints need not be exact:
// any simplex is a valid walk start, so plain floating-point tests are used here.
// ---------------------------------------------------------------------------------------------

struct GridView {
  double o[DLI_MAX_DIM], cell[DLI_MAX_DIM], ic[DLI_MAX_DIM];
  int g[DLI_MAX_DIM];
  int64_t cells;
};

// one thread per simplex: visit the cells of its bounding box whose centre lies inside it
template <int D>
__global__ void k_hint_raster(GridView gv, int ns, const double *__restrict__ vrec, int vs, const int32_t *__restrict__ srec,
                              int ss, int32_t *__restrict__ hint) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  double x[(D + 1) * D];
  int lo[D], hi[D];
  for (int i = 0; i <= D; i++) {
    int v = srec[(size_t)s * ss + i];
    for (int k = 0; k < D; k++) x[i * D + k] = vrec[(size_t)v * vs + k];
  }
  int64_t total = 1;
  for (int k = 0; k < D; k++) {
    double mn = x[k], mx = x[k];
    for (int i = 1; i <= D; i++) { mn = fmin(mn, x[i * D + k]); mx = fmax(mx, x[i * D + k]); }
    // cells whose centre o + (c + 0.5) * cell lies in [mn, mx]
    int a = (int)ceil((mn - gv.o[k]) * gv.ic[k] - 0.5);
    int b = (int)floor((mx - gv.o[k]) * gv.ic[k] - 0.5);
    lo[k] = max(a, 0);
    hi[k] = min(b, gv.g[k] - 1);
    if (hi[k] < lo[k]) return;
    total *= (hi[k] - lo[k] + 1);
  }
  // whole-simplex orientation (positive by construction); tolerance relative to it
  double m0[D * D];
  for (int i = 0; i < D; i++)
    for (int k = 0; k < D; k++) m0[i * D + k] = x[i * D + k] - x[D * D + k];
  double perm;
  double whole = DetPerm<D>::eval(m0, perm);
  double tol = -1e-9 * fabs(whole);
  int c[D];
  for (int k = 0; k < D; k++) c[k] = lo[k];
  for (int64_t it = 0; it < total; it++) {
    double p[D];
    for (int k = 0; k < D; k++) p[k] = gv.o[k] + (c[k] + 0.5) * gv.cell[k];
    bool inside = true;
    for (int j = 0; j <= D && inside; j++) {
      double pts[(D + 1) * D];
      for (int i = 0; i <= D; i++)
        for (int k = 0; k < D; k++) pts[i * D + k] = (i == j) ? p[k] : x[i * D + k];
      double mm[D * D];
      for (int i = 0; i < D; i++)
        for (int k = 0; k < D; k++) mm[i * D + k] = pts[i * D + k] - pts[D * D + k];
      double pm;
      if (DetPerm<D>::eval(mm, pm) < tol) inside = false;
    }
    if (inside) {
      size_t idx = 0;
      for (int k = D - 1; k >= 0; k--) idx = idx * gv.g[k] + c[k];
      hint[idx] = s;
    }
    for (int k = 0; k < D; k++) {  // odometer
      if (++c[k] <= hi[k]) break;
      c[k] = lo[k];
    }
  }
}

// empty cells (-1) adopt a neighbour's hint; in place (any neighbour's value is a valid hint)
template <int D>
__global__ void k_hint_dilate(GridView gv, int32_t *hint, unsigned int *remaining) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= gv.cells) return;
  if (hint[idx] >= 0) return;
  int64_t stride = 1, rem = idx;
  int found = -1;
  for (int k = 0; k < D; k++) {
    int ck = (int)(rem % gv.g[k]);
    rem /= gv.g[k];
    if (found < 0 && ck > 0) { int v = ((volatile int32_t *)hint)[idx - stride]; if (v >= 0) found = v; }
    if (found < 0 && ck < gv.g[k] - 1) { int v = ((volatile int32_t *)hint)[idx + stride]; if (v >= 0) found = v; }
    stride *= gv.g[k];
  }
  if (found >= 0) hint[idx] = found;
  else atomicAdd(remaining, 1u);
}
__global__ void k_hint_fill_rest(int64_t cells, int32_t *hint, int32_t value) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < cells && hint[idx] < 0) hint[idx] = value;
}

// ---------------------------------------------------------------------------------------------
// Per-simplex error evaluator (src/delaunay_2_interp.h:725-741 and helpers; closed form of the
// reference's SVD: err = (phi - mean f)^2 * |e_z - proj_span(e_z)|^2, span = span{w_i - w_0},
// w_i = (v_i, f_i) the lifted vertices of the (sub-)simplex).
// ---------------------------------------------------------------------------------------------

// items: item_verts [n][D+1] padded with -1 (device); centroid accumulated in slot order then divided
template <int D>
__global__ void k_error_prepare(int64_t n, const int32_t *__restrict__ item_verts, const double *__restrict__ vrec,
                                int vs, double *__restrict__ centroids) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double c[D];
  for (int k = 0; k < D; k++) c[k] = 0.0;
  int cnt = 0;
  for (int a = 0; a <= D; a++) {
    int v = item_verts[i * (D + 1) + a];
    if (v < 0) break;
    for (int k = 0; k < D; k++) c[k] += vrec[(size_t)v * vs + k];
    cnt++;
  }
  for (int k = 0; k < D; k++) centroids[i * D + k] = c[k] / cnt;
}

template <int D>
__global__ void k_error_finish(int64_t n, const int32_t *__restrict__ item_verts, const double *__restrict__ vrec,
                               int vs, const double *__restrict__ phi, double *__restrict__ err) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  constexpr int N = D + 1;  // lifted dimension
  double w[(D + 1) * N];
  int np = 0;
  for (int a = 0; a <= D; a++) {
    int v = item_verts[i * (D + 1) + a];
    if (v < 0) break;
    for (int k = 0; k < D; k++) w[a * N + k] = vrec[(size_t)v * vs + k];
    w[a * N + D] = vrec[(size_t)v * vs + D];
    np++;
  }
  err[i] = dli_simplex_error<D>(np, w, phi[i]);
}

// ---------------------------------------------------------------------------------------------
// Test hook: the device predicate on explicit point sets
// ---------------------------------------------------------------------------------------------
template <int D>
__global__ void k_debug_orient(int64_t n, const double *__restrict__ pts, int8_t *__restrict__ sign_out,
                               uint8_t *__restrict__ exact_used) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double p[(D + 1) * D];
  for (int k = 0; k < (D + 1) * D; k++) p[k] = pts[i * (D + 1) * D + k];
  int s;
  bool ex;
  if constexpr (D == 2) {
    s = orient2d_filtered(p[0], p[1], p[2], p[3], p[4], p[5]);
    ex = (s == 2);
    if (ex) s = orient2d_exact(p[0], p[1], p[2], p[3], p[4], p[5]);
  } else {
    s = orient_filtered<D>(p);
    ex = (s == 2);
    if (ex) s = orient_exact<D>(p);
  }
  sign_out[i] = (int8_t)s;
  if (exact_used) exact_used[i] = ex;
}

}  // namespace dli
