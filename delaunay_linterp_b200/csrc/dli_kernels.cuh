// dli_kernels.cuh -- __global__ kernels (sm_100a): batched interp (fast + straggler/exact pass),
// hint-grid construction, per-simplex error evaluator, predicate test hook.
#pragma once
#include <cfloat>
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

// Straggler queue of the fast passes.  The queue holds a fraction of the launch (run_device); a query that does not
// fit is marked instead -- a reserved NaN bit pattern in its result slot (the simplex slot for dli_locate) -- and the
// exact pass, when the counter shows an overflow, also scans the results for marks and serves those queries from their
// start hint.  Every query is written by the fast pass (value, queue slot or mark) before the exact pass starts, and a
// stale slot that happens to hold the mark is merely answered twice, so this is exact; the scan (8 B per query) only
// runs on batches where more than 1/16 of the queries need the exact pass.
constexpr long long kOverflowBits = 0x7ff8d11d11d11d11LL;
constexpr int kOverflowSimplex = (int)0x80000001;
template <int OUT>
__device__ __forceinline__ void enqueue_straggler(int32_t *__restrict__ queue, unsigned int *__restrict__ queue_count,
                                                  unsigned int queue_cap, int64_t qi, int where, double *__restrict__ ys,
                                                  int32_t *__restrict__ simplex_out) {
  const unsigned int slot = atomicAdd(queue_count, 1u);
  if (slot < queue_cap) { queue[slot] = (int32_t)qi; queue[(size_t)queue_cap + slot] = where; }   // exact pass resumes there
  else if (ys) ys[OUT == 1 ? 2 * qi : qi] = __longlong_as_double(kOverflowBits);
  else if (simplex_out) simplex_out[qi] = kOverflowSimplex;
}
template <int OUT>
__device__ __forceinline__ bool is_overflow_mark(const double *__restrict__ ys, const int32_t *__restrict__ simplex_out, int64_t i) {
  if (ys) return __double_as_longlong(ys[OUT == 1 ? 2 * i : i]) == kOverflowBits;
  return simplex_out && simplex_out[i] == kOverflowSimplex;
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
// keeps growing, the query's result slot receives a reserved mark instead (kOverflowBits / kOverflowSimplex) and
// k_interp2_full, seeing count > cap, scans the results for marks and serves those queries from their start hint.
// OUT = 0: ys[n] receives the interpolated value; OUT = 1 (dli_gradient): ys is a double2 array and receives the
// gradient of the located face (strictly interior queries only; everything else goes to the exact pass).
template <bool STATS, int EPI, bool PACKED, int OUT = 0>
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
            if (OUT == 1) {
              __stcs(reinterpret_cast<double2 *>(ys) + qi, face_gradient2(A, B, C));
            } else {
              double y = (EPI == 0) ? epilogue2_ref(px, py, A, B, C) : epilogue2_fast(px, py, A, B, C);
              __stcs(ys + qi, y);
            }
          }
          if (simplex_out) __stcs(simplex_out + qi, __ldg(m.orig + DLI_IDX(f, m.ns)));
          if (status_out) status_out[qi] = DLI_STATUS_OK;
          if (STATS) { st.steps += steps; st.max_steps = max(st.max_steps, steps); }
          active = false;
        } else {
          give_up = true;  // a decision the static filter cannot certify
        }
      }
      if (give_up) {
        enqueue_straggler<OUT>(queue, queue_count, queue_cap, qi, f, ys, simplex_out);
        active = false;
      }
    }
  }
  if (STATS) flush_stats(stats, st);
}

// ---------------------------------------------------------------------------------------------
// 2D fast pass on 64-byte self-contained walk records (see dli_device.cuh, "FatFace2")
// ---------------------------------------------------------------------------------------------

// Builds the walk records from the canonical vertex / face records.  One thread per face.
// The slow flag is set when (a) an edge vector does not survive the fp32 conversion within 2^-24
// relative (overflow, underflow, non-finite input), or (b) the a-priori bound on the error of the
// gradient-form interpolant over the face exceeds 1e-13 * min|f_i| (slivers, faces on which f changes
// sign or vanishes at a vertex): those faces are evaluated by k_interp2_full in the reference's own
// operation order.
// numeric core shared by the two record builders: writes the 64-byte record of the face (V0, V1, V2) with the packed
// neighbour word `nbr` (slow flag added here) to `o`; returns the slow flag
__device__ __forceinline__ bool make_fat2(const double4 &V0, const double4 &V1, const double4 &V2, unsigned long long nbr,
                                          uint4 *__restrict__ o) {
  const double Ax = V1.x - V0.x, Ay = V1.y - V0.y, Bx = V2.x - V0.x, By = V2.y - V0.y;
  const float ax = (float)Ax, ay = (float)Ay, bx = (float)Bx, by = (float)By;
  const double u = 5.9604644775390625e-08;  // 2^-24
  bool slow = false;
  slow |= !(fabs((double)ax - Ax) <= u * fabs(Ax));
  slow |= !(fabs((double)ay - Ay) <= u * fabs(Ay));
  slow |= !(fabs((double)bx - Bx) <= u * fabs(Bx));
  slow |= !(fabs((double)by - By) <= u * fabs(By));
  // affine interpolant  y = f0 + gx (x - x0) + gy (y - y0)
  const double eps = 1.1102230246251565e-16;  // 2^-53
  const double d1 = V1.z - V0.z, d2 = V2.z - V0.z;
  const double dl = Ax * By, dr = Ay * Bx, det = dl - dr, Sd = fabs(dl) + fabs(dr);
  const double xl = d1 * By, xr = d2 * Ay, nx = xl - xr, Sx = fabs(xl) + fabs(xr);
  const double yl = Ax * d2, yr = Bx * d1, ny = yl - yr, Sy = fabs(yl) + fabs(yr);
  const double gx = nx / det, gy = ny / det;
  // first-order rounding analysis with a factor 2 of slack (every input difference carries one rounding,
  // each product one, each 2-term sum one): |d det| <= 4 eps Sd, |d nx| <= 4 eps Sx
  const double ad = fabs(det);
  const double dgx = 8.0 * eps * (Sx + fabs(gx) * Sd) / ad + 2.0 * eps * fabs(gx);
  const double dgy = 8.0 * eps * (Sy + fabs(gy) * Sd) / ad + 2.0 * eps * fabs(gy);
  const double wx = fmax(fabs(Ax), fabs(Bx)), wy = fmax(fabs(Ay), fabs(By));
  const double E = dgx * wx + dgy * wy + 3.0 * eps * (fabs(V0.z) + fabs(gx) * wx + fabs(gy) * wy);
  double mn = fmin(fabs(V0.z), fmin(fabs(V1.z), fabs(V2.z)));
  const bool same_sign = (V0.z > 0.0 && V1.z > 0.0 && V2.z > 0.0) || (V0.z < 0.0 && V1.z < 0.0 && V2.z < 0.0);
  const bool all_zero = (V0.z == 0.0 && V1.z == 0.0 && V2.z == 0.0);
  if (!same_sign) mn = 0.0;
  if (!all_zero) slow |= !(E <= 1e-13 * mn);   // (NaN compares false -> slow)
  else slow |= !(gx == 0.0 && gy == 0.0);
  if (slow) nbr |= kFatSlow;
  uint4 h0, h1, h2, h3;
  h0.x = (unsigned)__double2loint(V0.x); h0.y = (unsigned)__double2hiint(V0.x);
  h0.z = (unsigned)__double2loint(V0.y); h0.w = (unsigned)__double2hiint(V0.y);
  h1.x = __float_as_uint(ax); h1.y = __float_as_uint(ay); h1.z = __float_as_uint(bx); h1.w = __float_as_uint(by);
  h2.x = (unsigned)__double2loint(V0.z); h2.y = (unsigned)__double2hiint(V0.z);
  h2.z = (unsigned)__double2loint(gx); h2.w = (unsigned)__double2hiint(gx);
  h3.x = (unsigned)__double2loint(gy); h3.y = (unsigned)__double2hiint(gy);
  h3.z = (unsigned)(nbr & 0xffffffffull); h3.w = (unsigned)(nbr >> 32);
  o[0] = h0; o[1] = h1; o[2] = h2; o[3] = h3;
  return slow;
}

template <bool DELTA>
__global__ void k_build_fat2(int ns, const int4 *__restrict__ frec, const double4 *__restrict__ vrec,
                             uint4 *__restrict__ fat, unsigned int *__restrict__ n_slow) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= ns) return;
  int4 r0 = frec[2 * (size_t)f], r1 = frec[2 * (size_t)f + 1];
  // neighbour field (21 bits): kPackMask = hull; ABS: the neighbour's index; DELTA (meshes with >= 2^21-2 faces):
  // index difference + kFatBias, kFatFar when it does not fit (the kernel then reads the canonical face record;
  // rare because faces are stored in Morton order)
  auto enc = [f](int nb) -> unsigned long long {
    if (nb < 0) return (unsigned long long)kPackMask;
    if (!DELTA) return (unsigned long long)(unsigned)nb;
    long long dlt = (long long)nb - f + kFatBias;
    return (dlt < 0 || dlt >= (long long)kFatFar) ? (unsigned long long)kFatFar : (unsigned long long)dlt;
  };
  unsigned long long nbr = enc(r0.w) | (enc(r1.x) << 21) | (enc(r1.y) << 42);
  if (make_fat2(vrec[r0.x], vrec[r0.y], vrec[r0.z], nbr, fat + 4 * (size_t)f)) atomicAdd(n_slow, 1u);
}

// ---- cell-indexed walk records ("implicit hint" layout for meshes far larger than L2) ----
// The record array is laid out by hint-grid rows: row j owns the slots rowstart[j] .. rowstart[j+1]-1 = gx cell
// slots (slot rowstart[j] + i is a copy of the record of the face the hint grid assigns to cell (i, j)) followed by
// the records of the faces that own no cell and whose centroid lies in row j.  A query therefore starts at
// slot = rowstart[cj] + ci without reading a hint, and every neighbour of a record lives at most about one row
// away, so its slot delta fits the 21-bit DELTA field.  Neighbour fields point to a copy of the neighbour face in
// the 3 x 3 cell neighbourhood if there is one, else to the neighbour's canonical (lowest) slot.
constexpr int kNoSlot = 0x7f7f7f7f;  // canon[] initial value (memset 0x7f): no slot yet
struct SlotGrid {
  double ox, oy, icx, icy;
  int gx, gy;
};
__device__ __forceinline__ int face_row(const SlotGrid &g, const int4 *__restrict__ frec, const double4 *__restrict__ vrec, int f,
                                        int *col = nullptr) {
  int4 r0 = frec[2 * (size_t)f];
  double4 A = vrec[r0.x], B = vrec[r0.y], C = vrec[r0.z];
  int j = (int)(((A.y + B.y + C.y) / 3.0 - g.oy) * g.icy);
  if (col) { int i = (int)(((A.x + B.x + C.x) / 3.0 - g.ox) * g.icx); *col = max(0, min(g.gx - 1, i)); }
  return max(0, min(g.gy - 1, j));
}
// owner[f] = lowest cell whose hint is f (kNoSlot: the face owns no cell)
__global__ void k_slot_owner(int64_t cells, const int32_t *__restrict__ hint, int *__restrict__ owner) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < cells) atomicMin(owner + hint[c], (int)c);
}
__global__ void k_slot_rowcount(int ns, SlotGrid g, const int *__restrict__ owner, const int4 *__restrict__ frec,
                                const double4 *__restrict__ vrec, int *__restrict__ rowcount) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f < ns && owner[f] == kNoSlot) atomicAdd(rowcount + face_row(g, frec, vrec, f), 1);
}
// slot -> face table: cell slots from the hint grid, extra slots handed out per row
__global__ void k_slot_fill_cells(int64_t cells, int gx, const int32_t *__restrict__ hint, const int *__restrict__ rowstart,
                                  int32_t *__restrict__ slot_face) {
  int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c < cells) slot_face[rowstart[c / gx] + (int)(c % gx)] = hint[c];
}
__global__ void k_slot_fill_extra(int ns, SlotGrid g, const int *__restrict__ owner, const int4 *__restrict__ frec,
                                  const double4 *__restrict__ vrec, const int *__restrict__ rowstart, int *__restrict__ cursor,
                                  int32_t *__restrict__ slot_face) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= ns || owner[f] != kNoSlot) return;
  int j = face_row(g, frec, vrec, f);
  slot_face[rowstart[j] + g.gx + atomicAdd(cursor + j, 1)] = f;
}
// canon[f] = lowest slot holding a copy of face f
__global__ void k_slot_canon(int64_t nslots, const int32_t *__restrict__ slot_face, int *__restrict__ canon) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < nslots) atomicMin(canon + slot_face[s], (int)s);
}
__global__ void k_build_fat2_slots(int64_t nslots, SlotGrid g, const int32_t *__restrict__ hint, const int *__restrict__ rowstart,
                                   const int32_t *__restrict__ slot_face, const int *__restrict__ canon,
                                   const int4 *__restrict__ frec, const double4 *__restrict__ vrec,
                                   uint4 *__restrict__ fat, unsigned int *__restrict__ n_slow) {
  int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots) return;
  const int f = slot_face[s];
  int4 r0 = frec[2 * (size_t)f], r1 = frec[2 * (size_t)f + 1];
  // row of this slot: the last j with rowstart[j] <= s
  int lo = 0, hi = g.gy - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (rowstart[mid] <= s) lo = mid; else hi = mid - 1;
  }
  int cj = lo, ci = (int)(s - rowstart[cj]);
  if (ci >= g.gx) face_row(g, frec, vrec, f, &ci);   // an extra slot: search around the face's centroid cell
  auto enc = [&](int nb) -> unsigned long long {
    if (nb < 0) return (unsigned long long)kPackMask;
    long long target = canon[nb];
    long long best = -1;
    for (int dj = -1; dj <= 1; dj++)
      for (int di = -1; di <= 1; di++) {
        int i = ci + di, j = cj + dj;
        if (i < 0 || j < 0 || i >= g.gx || j >= g.gy) continue;
        if (hint[(size_t)j * g.gx + i] != nb) continue;
        long long c = (long long)rowstart[j] + i;
        if (best < 0 || llabs(c - s) < llabs(best - s)) best = c;
      }
    if (best >= 0) target = best;
    long long dlt = target - s + kFatBias;
    return (dlt < 0 || dlt >= (long long)kFatFar) ? (unsigned long long)kFatFar : (unsigned long long)dlt;
  };
  unsigned long long nbr = enc(r0.w) | (enc(r1.x) << 21) | (enc(r1.y) << 42);
  bool slow = make_fat2(vrec[r0.x], vrec[r0.y], vrec[r0.z], nbr, fat + 4 * (size_t)s);
  if (slow && s == canon[f]) atomicAdd(n_slow, 1u);   // count faces, not copies
}

// Fast pass on the walk records.  Same warp-persistent lane state machine as k_interp2_fast (one walk step
// per lane per iteration, ballot/popc refill), but a step costs ONE L1TEX wavefront per query instead of
// four: the two 32-byte halves of a record are fetched by the two lanes of a lane pair in one LDG.E.256
// (both addresses in the same 128-byte line), once for the even lane's face and once for the odd lane's,
// and the halves are exchanged with 8 SHFL.BFLY ("pair-transposed load").
#ifndef DLI_X_FAT_THREADS   // developer experiment knobs (scripts/build_variants.py); defaults are the shipped values
#define DLI_X_FAT_THREADS 128
#endif
#ifndef DLI_X_FAT_BLOCKS
#define DLI_X_FAT_BLOCKS 9
#endif
// RING = true adds a per-warp shared-memory prefetch ring (cp.async): the queries of the block of 32 three blocks
// ahead and the start hints of the block two ahead are in flight while the current block walks, so a loop
// iteration waits for ONE dependent memory latency (the walk record) instead of three (query -> hint -> record).
// Ring = 128 query slots (16 B) + 128 hint slots (4 B) per warp = 2.5 KB; slot = query index & 127.
constexpr int kRingSlots = 128;
constexpr int kRingBytesPerWarp = kRingSlots * (16 + 4);
// CELL = true: cell-indexed records (k_build_fat2_slots): the start slot IS the query's hint cell, no hint is read
// (the ring issues an L2 prefetch of the start record instead); a far neighbour hands the query to the exact pass.
template <bool STATS, bool RING, bool DELTA, int OUT = 0, bool CELL = false>
__global__ void __launch_bounds__(DLI_X_FAT_THREADS, DLI_X_FAT_BLOCKS)
k_interp2_fat(Mesh2View m, int64_t n, int per_warp, const double2 *__restrict__ xs, double *__restrict__ ys,
              int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out, int32_t *__restrict__ queue,
              unsigned int *__restrict__ queue_count, unsigned int queue_cap, unsigned long long *__restrict__ stats) {
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const bool odd = lane & 1;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * per_warp;
  if (base >= n) return;  // warp-uniform
  const int64_t end = min(n, base + (int64_t)per_warp);
  const char *rec = reinterpret_cast<const char *>(m.fat) + (odd ? 32 : 0);
  int64_t next = base;
  bool active = false;
  double px = 0.0, py = 0.0;
  int f = 0, from = -1;
  int64_t qi = 0;
  unsigned int steps = 0;
  StatsAcc st;
  // ---- prefetch ring (RING only)
  extern __shared__ __align__(16) unsigned char ring_smem[];
  double2 *ringq = nullptr;
  int *ringh = nullptr;
  unsigned long long pol_last = 0, pol_first = 0;
  int blk = 0;
  auto issue_q = [&](int64_t j) {  // query j -> ring (streamed: L2 evict_first)
    if (j < end)
      asm volatile("cp.async.cg.shared.global.L2::cache_hint [%0], [%1], 16, %2;" ::"r"(
                       (unsigned)__cvta_generic_to_shared(ringq + (j & (kRingSlots - 1)))),
                   "l"(xs + j), "l"(pol_first)
                   : "memory");
  };
  auto issue_h = [&](int64_t j) {  // start hint of query j (already in the ring) -> ring (hint grid: L2 evict_last)
    // (CELL: the start slot is computed from the query at refill time, nothing to fetch.  Prefetching the start
    // record into L2 here was measured and is harmful on DRAM-resident meshes: 4.2 -> 7.0 ms on cfg5, the
    // prefetched lines are evicted before use and fetched twice)
    if (!CELL && j < end) {
      double2 q = ringq[j & (kRingSlots - 1)];
      int ci = (int)((q.x - m.ox) * m.icx), cj = (int)((q.y - m.oy) * m.icy);
      ci = max(0, min(m.gx - 1, ci));
      cj = max(0, min(m.gy - 1, cj));
      asm volatile("cp.async.ca.shared.global.L2::cache_hint" DLI_L2_64B " [%0], [%1], 4, %2;" ::"r"(
                       (unsigned)__cvta_generic_to_shared(ringh + (j & (kRingSlots - 1)))),
                   "l"(m.hint + DLI_IDX((size_t)cj * m.gx + ci, (size_t)m.gx * m.gy)), "l"(pol_last)
                   : "memory");
    }
  };
  auto commit_wait = [&](bool commit) {
    if (commit) asm volatile("cp.async.commit_group;" ::: "memory");
    else { asm volatile("cp.async.wait_group 0;" ::: "memory"); __syncwarp(); }
  };
  if (RING) {
    const int wib = threadIdx.x >> 5, nw = blockDim.x >> 5;
    ringq = reinterpret_cast<double2 *>(ring_smem) + wib * kRingSlots;
    ringh = reinterpret_cast<int *>(ring_smem + (size_t)nw * kRingSlots * 16) + wib * kRingSlots;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_last));
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_first));
    issue_q(base + lane); issue_q(base + 32 + lane); issue_q(base + 64 + lane);
    commit_wait(true); commit_wait(false);
    issue_h(base + lane); issue_h(base + 32 + lane);
    commit_wait(true); commit_wait(false);
    issue_h(base + 64 + lane); issue_q(base + 96 + lane);
    commit_wait(true);
  }
  for (;;) {
    unsigned need = __ballot_sync(FULL, !active);
    if (need) {
      int64_t idx = next + __popc(need & ((1u << lane) - 1u));
      if (!active && idx < end) {
        if (RING) {
          double2 q = ringq[idx & (kRingSlots - 1)];
          px = q.x; py = q.y;
          f = CELL ? cell2(m, px, py) : ringh[idx & (kRingSlots - 1)];
        } else {
          double2 q = __ldcs(xs + idx);
          px = q.x; py = q.y;
          f = CELL ? cell2(m, px, py) : hint2(m, px, py);
        }
        qi = idx;
        from = -1;
        steps = 0;
        active = true;
      }
      next += __popc(need);
      if (RING) {
        const int nb = (int)((next - base) >> 5);
        if (nb != blk) {  // warp-uniform: `next` entered a new block of 32 -> advance the pipeline by one block
          blk = nb;
          commit_wait(false);  // the previous group (hints of blk+1, queries of blk+2) has landed
          issue_h(base + (int64_t)(blk + 2) * 32 + lane);
          issue_q(base + (int64_t)(blk + 3) * 32 + lane);
          commit_wait(true);
        }
      }
    }
    if (!__any_sync(FULL, active)) break;
    // pair-transposed load of the two records of this lane pair
    const int fs = active ? f : -1;
    const int fp = __shfl_xor_sync(FULL, fs, 1);
    const int fe = odd ? fp : fs, fo = odd ? fs : fp;
    unsigned R1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, R2[8] = {0, 0, 0, 0, 0, 0, 0, 0}, X[8];
    if (fe >= 0) ldg256(R1, rec + (size_t)DLI_IDX(fe, m.nslots) * 64);
    if (fo >= 0) ldg256(R2, rec + (size_t)DLI_IDX(fo, m.nslots) * 64);
#pragma unroll
    for (int i = 0; i < 8; i++) X[i] = __shfl_xor_sync(FULL, odd ? R1[i] : R2[i], 1);
    if (active) {
      steps++;
      unsigned G[8], V[8];  // geometry half, values half of this lane's own record
#pragma unroll
      for (int i = 0; i < 8; i++) { G[i] = odd ? X[i] : R1[i]; V[i] = odd ? R2[i] : X[i]; }
      const double dx = px - __hiloint2double((int)G[1], (int)G[0]);
      const double dy = py - __hiloint2double((int)G[3], (int)G[2]);
      const double ax = (double)__uint_as_float(G[4]), ay = (double)__uint_as_float(G[5]);
      const double bx = (double)__uint_as_float(G[6]), by = (double)__uint_as_float(G[7]);
      // o2: edge v0->v1 (opposite v2), o1: edge v2->v0 (opposite v1), o0 = w - o2 - o1: edge v1->v2
      const double l2 = ax * dy, r2 = ay * dx, o2 = l2 - r2, e2 = kFatK * (fabs(l2) + fabs(r2));
      const double l1 = dx * by, r1 = dy * bx, o1 = l1 - r1, e1 = kFatK * (fabs(l1) + fabs(r1));
      const double lw = ax * by, rw = ay * bx, w = lw - rw, ew = (2.0 * kFatK) * (fabs(lw) + fabs(rw));
      const double o0 = (w - o2) - o1, e0 = (ew + e2) + e1;
      const unsigned long long nbr = ((unsigned long long)V[7] << 32) | V[6];
      const unsigned k0 = (unsigned)(nbr & kPackMask), k1 = (unsigned)((nbr >> 21) & kPackMask), k2 = (unsigned)((nbr >> 42) & kPackMask);
      // neighbour indices; -2 = hull or (DELTA) far neighbour, never equal to `from`
      const int n0 = DELTA ? (k0 >= kFatFar ? -2 : f + (int)k0 - kFatBias) : (k0 == kPackMask ? -2 : (int)k0);
      const int n1 = DELTA ? (k1 >= kFatFar ? -2 : f + (int)k1 - kFatBias) : (k1 == kPackMask ? -2 : (int)k1);
      const int n2 = DELTA ? (k2 >= kFatFar ? -2 : f + (int)k2 - kFatBias) : (k2 == kPackMask ? -2 : (int)k2);
      const bool c0 = n0 == from, c1 = n1 == from, c2 = n2 == from;
      const bool neg0 = (-o0 > e0) & !c0, neg1 = (-o1 > e1) & !c1, neg2 = (-o2 > e2) & !c2;
      // (gradient queries honour the slow flag too: it also marks faces whose fp32 edge vectors are unusable)
      bool give_up = (nbr & kFatSlow) != 0;
      if (!give_up) {
        if (neg0 | neg1 | neg2) {
          int nb = neg0 ? n0 : (neg1 ? n1 : n2);
          if (DELTA) {
            const unsigned kc = neg0 ? k0 : (neg1 ? k1 : k2);
            if (kc == kFatFar)  // neighbour too far away in index space for the 21-bit field: canonical record
              nb = CELL ? -1 : __ldg(reinterpret_cast<const int *>(m.frec) + 8 * (size_t)DLI_IDX(f, m.ns) + 3 + (neg0 ? 0 : (neg1 ? 1 : 2)));
          }
          if (nb < 0 || (int)steps >= m.fast_steps) give_up = true;  // hull exit / straggler
          else { from = f; f = nb; }
        } else {
          // closed face: o >= e certifies "not negative".  e carries a factor 2 of slack over the rounding analysis, so
          // o == e > 0 is still strictly positive; o == e == 0 happens only when both products of the edge's cross
          // product are exact zeros (an axis-aligned edge through vertex 0 with the query's coordinate equal to the
          // vertex's: fp32(A) = 0 iff A = 0, p - p0 = 0 iff p = p0; barring underflow), i.e. the query is exactly ON
          // that edge -- CGAL's EDGE / VERTEX locate types, which the reference answers from the face at hand.
          // Gradient queries (OUT = 1) keep the strict test: on an edge / vertex they need gradient2_full's averaging.
          bool pos = OUT == 1 ? (((o0 > e0) | c0) & ((o1 > e1) | c1) & ((o2 > e2) | c2))
                              : (((o0 >= e0) | c0) & ((o1 >= e1) | c1) & ((o2 >= e2) | c2));
          if (pos) {
            if (ys) {
              const double f0 = __hiloint2double((int)V[1], (int)V[0]);
              const double gx = __hiloint2double((int)V[3], (int)V[2]);
              const double gy = __hiloint2double((int)V[5], (int)V[4]);
              if (OUT == 1) __stcs(reinterpret_cast<double2 *>(ys) + qi, make_double2(gx, gy));
              else __stcs(ys + qi, f0 + fma(gx, dx, gy * dy));
            }
            if (simplex_out) {
              const int face = CELL ? __ldg(m.slot_face + DLI_IDX(f, m.nslots)) : f;
              __stcs(simplex_out + qi, __ldg(m.orig + DLI_IDX(face, m.ns)));
            }
            if (status_out) status_out[qi] = DLI_STATUS_OK;
            if (STATS) { st.steps += steps; st.max_steps = max(st.max_steps, steps); }
            active = false;
          } else {
            give_up = true;  // a decision the static filter cannot certify
          }
        }
      }
      if (give_up) {
        enqueue_straggler<OUT>(queue, queue_count, queue_cap, qi, CELL ? __ldg(m.slot_face + DLI_IDX(f, m.nslots)) : f, ys, simplex_out);
        active = false;
      }
    }
  }
  if (RING) asm volatile("cp.async.wait_all;" ::: "memory");
  if (STATS) flush_stats(stats, st);
}

// Straggler / exact pass over the queue written by the fast pass (count read on the device: no
// host synchronisation between the two launches).
template <bool STATS, int EPI, int OUT = 0>
__global__ void __launch_bounds__(128)
k_interp2_full(Mesh2View m, const double2 *__restrict__ xs, double *__restrict__ ys,
               int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out,
               const int32_t *__restrict__ queue, const unsigned int *__restrict__ queue_count,
               unsigned int queue_cap, int64_t n /* batch size: processed directly if the queue overflowed */,
               unsigned long long *__restrict__ stats) {
  const int64_t total = (int64_t)*queue_count;
  const bool overflow = total > (int64_t)queue_cap;
  const int64_t cnt = overflow ? (int64_t)queue_cap : total;
  StatsAcc st;
  // k < cnt: queue entries; k >= cnt (overflow only): scan of the batch for marked queries
  const int64_t span = overflow ? cnt + n : cnt;
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < span; k += (int64_t)gridDim.x * blockDim.x) {
    const bool queued = k < cnt;
    const int64_t i = queued ? (int64_t)queue[k] : k - cnt;
    if (!queued && !is_overflow_mark<OUT>(ys, simplex_out, i)) continue;
    double2 q = xs[i];
    double px = q.x, py = q.y;
    if (!(isfinite(px) && isfinite(py))) {  // NaN / Inf coordinates: no face can be certified, report instead of walking
      const double nan = __longlong_as_double(0x7ff8000000000000LL);
      if (ys) { if (OUT == 1) reinterpret_cast<double2 *>(ys)[i] = make_double2(nan, nan); else ys[i] = nan; }
      if (simplex_out) simplex_out[i] = -1;
      if (status_out) status_out[i] = DLI_STATUS_STEP_LIMIT;
      continue;
    }
    // resume the walk in the face where the fast pass stopped (hull exit, step cap, uncertain sign): the walk up to
    // there is not repeated.  Marked queries (queue overflow) start from the hint.
    int f = queued ? queue[(size_t)queue_cap + k] : hint2(m, px, py);
    if (OUT == 1) {
      double2 g;
      uint8_t status = gradient2_full(m, px, py, f, g, st);
      if (ys) reinterpret_cast<double2 *>(ys)[i] = g;
      if (simplex_out) simplex_out[i] = (status == DLI_STATUS_OUTSIDE_NAN) ? -1 : __ldg(m.orig + f);
      if (status_out) status_out[i] = status;
      continue;
    }
    double4 A, B, C;
    uint8_t status = m.outside_nan ? locate2_full<false>(m, px, py, f, A, B, C, st) : locate2_full(m, px, py, f, A, B, C, st);
    double y;
    if (status == DLI_STATUS_STEP_LIMIT || status == DLI_STATUS_OUTSIDE_NAN) {
      y = __longlong_as_double(0x7ff8000000000000LL);
    } else {
      y = (EPI == 0) ? epilogue2_ref(px, py, A, B, C) : epilogue2_fast(px, py, A, B, C);
    }
    if (ys) ys[i] = y;
    if (simplex_out) simplex_out[i] = (status == DLI_STATUS_OUTSIDE_NAN || status == DLI_STATUS_STEP_LIMIT) ? -1 : __ldg(m.orig + f);
    if (status_out) status_out[i] = status;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && m.straggler_frac && n > 0) *m.straggler_frac = (float)total / (float)n;
  if (STATS) {
    flush_stats(stats, st);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(stats + ST_HARD, (unsigned long long)total);
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
        if (simplex_out) __stcs(simplex_out + qi, __ldg(m.orig + DLI_IDX(s, m.ns)));
        if (status_out) status_out[qi] = DLI_STATUS_OK;
        if (STATS) { st.steps += steps; st.max_steps = max(st.max_steps, steps); }
        active = false;
      } else {
        give_up = true;
      }
      if (give_up) {
        enqueue_straggler<0>(queue, queue_count, queue_cap, qi, s, ys, simplex_out);
        active = false;
      }
    }
  }
  if (STATS) flush_stats(stats, st);
}

#ifndef DLI_X_FULLD_BLOCKS
#define DLI_X_FULLD_BLOCKS 5   // measured on all-degenerate batches: 3 / 4 / 5 / 6 / 8 CTAs per SM -> 8.4 / 7.9 / 7.5 / 7.8 / 9.1 ms (4D)
#endif
template <int D, bool STATS, int EPI>
__global__ void __launch_bounds__(128, DLI_X_FULLD_BLOCKS)
k_interpD_full(MeshDView m, const double *__restrict__ xs, double *__restrict__ ys,
               int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out,
               const int32_t *__restrict__ queue, const unsigned int *__restrict__ queue_count,
               unsigned int queue_cap, int64_t n, unsigned long long *__restrict__ stats) {
  const int64_t total = (int64_t)*queue_count;
  const bool overflow = total > (int64_t)queue_cap;
  const int64_t cnt = overflow ? (int64_t)queue_cap : total;
  StatsAcc st;
  const int64_t span = overflow ? cnt + n : cnt;   // queue entries, then (overflow only) a scan for marked queries
  for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < span; k += (int64_t)gridDim.x * blockDim.x) {
    const bool queued = k < cnt;
    const int64_t i = queued ? (int64_t)queue[k] : k - cnt;
    if (!queued && !is_overflow_mark<0>(ys, simplex_out, i)) continue;
    double p[D];
#pragma unroll
    for (int a = 0; a < D; a++) p[a] = xs[i * D + a];
    bool finite = true;
#pragma unroll
    for (int a = 0; a < D; a++) finite = finite && isfinite(p[a]);
    if (!finite) {  // NaN / Inf coordinates: report instead of walking
      if (ys) ys[i] = __longlong_as_double(0x7ff8000000000000LL);
      if (simplex_out) simplex_out[i] = -1;
      if (status_out) status_out[i] = DLI_STATUS_STEP_LIMIT;
      continue;
    }
    int s = queued ? queue[(size_t)queue_cap + k] : hintD<D>(m, p);   // resume where the fast pass stopped
    SimplexD<D> S;
    uint8_t status = locateD_full<D>(m, p, s, S, st);
    double y;
    if (status != DLI_STATUS_OK) {
      y = __longlong_as_double(0x7ff8000000000000LL);
      s = -1;
    } else {
      y = (EPI == 0) ? epilogueD_ref<D, true>(p, S) : epilogueD_fast<D>(p, S);
    }
    if (ys) ys[i] = y;
    if (simplex_out) simplex_out[i] = s >= 0 ? __ldg(m.orig + DLI_IDX(s, m.ns)) : -1;
    if (status_out) status_out[i] = status;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && m.straggler_frac && n > 0) *m.straggler_frac = (float)total / (float)n;
  if (STATS) {
    flush_stats(stats, st);
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(stats + ST_HARD, (unsigned long long)total);
  }
}

// ---------------------------------------------------------------------------------------------
// dD fast pass on walk records (see dli_device.cuh, FatD<D>)
// ---------------------------------------------------------------------------------------------

// double-double helpers for the one-time record build (relative error of one operation < 2^-100)
struct dd { double hi, lo; };
__device__ __forceinline__ dd dd_make(double a, double b) { dd r; two_sum(a, b, r.hi, r.lo); return r; }
__device__ __forceinline__ dd dd_neg(dd a) { dd r; r.hi = -a.hi; r.lo = -a.lo; return r; }
__device__ __forceinline__ dd dd_add(dd a, dd b) {
  double s, e, t, f2;
  two_sum(a.hi, b.hi, s, e);
  two_sum(a.lo, b.lo, t, f2);
  e += t;
  double s2, e2;
  two_sum(s, e, s2, e2);   // (fast renormalisation is enough: |e| << |s| after the first pass or both tiny)
  e2 += f2;
  dd r;
  two_sum(s2, e2, r.hi, r.lo);
  return r;
}
__device__ __forceinline__ dd dd_mul(dd a, dd b) {
  double p, e;
  two_prod(a.hi, b.hi, p, e);
  e += a.hi * b.lo + a.lo * b.hi;
  dd r;
  two_sum(p, e, r.hi, r.lo);
  return r;
}
__device__ __forceinline__ dd dd_div(dd a, dd b) {
  double q1 = a.hi / b.hi;
  dd q; q.hi = q1; q.lo = 0.0;
  dd r = dd_add(a, dd_neg(dd_mul(b, q)));
  double q2 = r.hi / b.hi;
  q.hi = q2;
  r = dd_add(r, dd_neg(dd_mul(b, q)));
  double q3 = r.hi / b.hi;
  dd res = dd_make(q1, q2);
  dd t; t.hi = q3; t.lo = 0.0;
  return dd_add(res, t);
}
// determinant of the K x K dd matrix m (row-major) by cofactor expansion along column 0
template <int K>
struct DetDD {
  __device__ static dd eval(const dd *m) {
    dd det; det.hi = 0.0; det.lo = 0.0;
    for (int r = 0; r < K; r++) {
      dd sub[(K - 1) * (K - 1)];
      int rr = 0;
      for (int i = 0; i < K; i++) {
        if (i == r) continue;
        for (int c = 1; c < K; c++) sub[rr * (K - 1) + (c - 1)] = m[i * K + c];
        rr++;
      }
      dd t = dd_mul(m[r * K], DetDD<K - 1>::eval(sub));
      det = dd_add(det, (r & 1) ? dd_neg(t) : t);
    }
    return det;
  }
};
template <>
struct DetDD<1> {
  __device__ static dd eval(const dd *m) { return m[0]; }
};

// Builds the dD walk records.  One thread per simplex.  G = (E^-1)^T (E rows = v_i - v_0, formed exactly as
// double-doubles) by cofactors in double-double arithmetic, so the stored fp64 gradient g = sum_i (f_i - f_0) G_i is
// the correctly rounded exact gradient unless the simplex is so flat that even 2^-94 * permanent / |W| matters (the
// bound below accounts for it).  The slow flag is set when an entry of the fp32 G is not within 1.5 fp32
// ulps-of-the-row-maximum of the exact value, or when the a-priori error bound of the affine interpolant exceeds
// 1e-13 * min|f_i| (see k_build_fat2): k_interpD_full handles those simplices in the reference's operation order.
// `list` (dli_mesh_update): the device indices of the simplices to (re)build, n of them; NULL = all ns simplices.
// `epi_base`: first record index of the interpolant array of the split 4D layout (= the number of walk records).
template <int D, bool PACK = false>
__global__ void k_build_fatD(int ns, const int32_t *__restrict__ srec, int ss, const double *__restrict__ vrec, int vs,
                             uint4 *__restrict__ fat, unsigned int *__restrict__ n_slow, const int32_t *__restrict__ list = nullptr,
                             int n_list = 0) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (list) {
    if (s >= n_list) return;
    s = list[s];
  } else if (s >= ns) return;
  const double eps = 1.1102230246251565e-16, u = 5.9604644775390625e-08;
  const double edd = 5.048709793414476e-29;  // 2^-94: bound on the accumulated double-double error per unit of permanent
  int v[D + 1], nb[D + 1];
  for (int i = 0; i <= D; i++) { v[i] = srec[(size_t)s * ss + i]; nb[i] = srec[(size_t)s * ss + D + 1 + i]; }
  double x0[D], f[D + 1], E[D * D];
  dd Ed[D * D];
  for (int k = 0; k < D; k++) x0[k] = vrec[(size_t)v[0] * vs + k];
  f[0] = vrec[(size_t)v[0] * vs + D];
  for (int i = 0; i < D; i++) {
    for (int k = 0; k < D; k++) {
      Ed[i * D + k] = dd_make(vrec[(size_t)v[i + 1] * vs + k], -x0[k]);   // exact difference
      E[i * D + k] = Ed[i * D + k].hi;
    }
    f[i + 1] = vrec[(size_t)v[i + 1] * vs + D];
  }
  // cofactors (double-double) and their permanents (double)
  dd Cd[D * D];
  double P[D * D];
  for (int i = 0; i < D; i++)
    for (int k = 0; k < D; k++) {
      dd subd[(D - 1) * (D - 1)];
      double sub[(D - 1) * (D - 1)];
      int rr = 0;
      for (int a = 0; a < D; a++) {
        if (a == i) continue;
        int cc = 0;
        for (int b = 0; b < D; b++) {
          if (b == k) continue;
          subd[rr * (D - 1) + cc] = Ed[a * D + b];
          sub[rr * (D - 1) + cc] = E[a * D + b];
          cc++;
        }
        rr++;
      }
      double sp;
      DetPerm<D - 1>::eval(sub, sp);
      dd sd = DetDD<D - 1>::eval(subd);
      Cd[i * D + k] = ((i + k) & 1) ? dd_neg(sd) : sd;
      P[i * D + k] = sp;
    }
  dd Wd; Wd.hi = 0.0; Wd.lo = 0.0;
  double PW = 0.0;
  for (int k = 0; k < D; k++) { Wd = dd_add(Wd, dd_mul(Ed[k], Cd[k])); PW += fabs(E[k]) * P[k]; }
  const double aW = fabs(Wd.hi);
  bool slow = !(aW > 0.0);
  dd Gd[D * D];
  double G[D * D], dG[D * D];
  float g32[D * D];
  for (int i = 0; i < D * D; i++) {
    Gd[i] = dd_div(Cd[i], Wd);
    G[i] = Gd[i].hi;
    // error of the fp64 G[i] against the exact entry: one rounding + the double-double residue
    dG[i] = eps * fabs(G[i]) + edd * (P[i] + fabs(G[i]) * PW) / aW;
    g32[i] = (float)G[i];
  }
  unsigned long long W[5] = {0ull, 0ull, 0ull, 0ull, 0ull};   // PACK: block-floating-point G (FatD<4, true>)
  if constexpr (PACK) {
    for (int i = 0; i < D; i++) {
      double gmax = 0.0;
      for (int k = 0; k < D; k++) gmax = fmax(gmax, fabs(G[i * D + k]));
      bool ok = gmax > 0.0 && gmax < 1e300;
      int e = ok ? ilogb(gmax) - 16 : 0;                      // gmax / 2^e in [2^16, 2^17)
      if (e < -128 || e > 127) { ok = false; e = 0; }
      double dmax = 0.0, dec[D];
      unsigned __int128 row = 0;   // 80 bits: four 18-bit mantissas (q + 2^17), then the exponent byte (e + 128)
      for (int k = 0; k < D; k++) {
        double q = ok ? rint(ldexp(G[i * D + k], -e)) : 0.0;
        q = fmin(131071.0, fmax(-131071.0, q));
        dec[k] = ldexp(q, e);
        dmax = fmax(dmax, fabs(dec[k]));
        row |= (unsigned __int128)(unsigned long long)((long long)q + 131072ll) << (18 * k);
      }
      row |= (unsigned __int128)(unsigned long long)(e + 128) << 72;
      // every decoded entry within 2^-17 (half a mantissa unit: gmax / 2^e >= 2^16) of the decoded row maximum, the
      // error of computing G included; the 2^-10 relative allowance covers dmax < gmax by a unit and dG
      for (int k = 0; k < D; k++) ok = ok && (fabs(dec[k] - G[i * D + k]) + dG[i * D + k] <= 7.63685e-06 * dmax);
      slow |= !ok;
      // row i occupies bits [80 i, 80 i + 80) of W: byte-aligned, so bytes never straddle a 64-bit word
      for (int t = 0; t < 10; t++) {
        const int pos = 80 * i + 8 * t;
        W[pos / 64] |= (unsigned long long)((unsigned)(row >> (8 * t)) & 0xffu) << (pos % 64);
      }
    }
  } else {
  for (int i = 0; i < D; i++) {
    double gmax = 0.0;
    for (int k = 0; k < D; k++) gmax = fmax(gmax, fabs((double)g32[i * D + k]));
    slow |= !(gmax > 0.0 && gmax < 1e300);
    for (int k = 0; k < D; k++) slow |= !(fabs((double)g32[i * D + k] - G[i * D + k]) + dG[i * D + k] <= 1.5 * u * gmax);
  }
  }
  // affine interpolant (gradient accumulated in double-double) and its a-priori error bound over the simplex
  double g[D], E_eval = (D + 2) * eps * fabs(f[0]);
  double mn = fabs(f[0]);
  bool pos = f[0] > 0.0, neg = f[0] < 0.0, zero = f[0] == 0.0;
  for (int i = 1; i <= D; i++) {
    mn = fmin(mn, fabs(f[i]));
    pos &= f[i] > 0.0; neg &= f[i] < 0.0; zero &= f[i] == 0.0;
  }
  for (int k = 0; k < D; k++) {
    dd gk; gk.hi = 0.0; gk.lo = 0.0;
    double dk = 0.0, wk = 0.0;
    for (int i = 0; i < D; i++) {
      const dd df = dd_make(f[i + 1], -f[0]);
      gk = dd_add(gk, dd_mul(df, Gd[i * D + k]));
      dk += fabs(df.hi) * (edd * (P[i * D + k] + fabs(G[i * D + k]) * PW) / aW + 8.0 * edd * fabs(G[i * D + k]));
      wk = fmax(wk, fabs(E[i * D + k]));
    }
    g[k] = gk.hi;
    // rounding of g[k] to fp64 + propagated double-double residue, then the (D + 2) roundings of the evaluation
    E_eval += (eps * fabs(g[k]) + dk) * wk + (D + 2) * eps * fabs(g[k]) * wk;
  }
  if (zero) {
    for (int k = 0; k < D; k++) slow |= !(g[k] == 0.0);
  } else {
    if (!(pos || neg)) mn = 0.0;
    slow |= !(E_eval <= 1e-13 * mn);
  }
  if (slow) atomicAdd(n_slow, 1u);
  // ---- pack
  auto d2 = [](double a, double b) {
    uint4 r;
    r.x = (unsigned)__double2loint(a); r.y = (unsigned)__double2hiint(a);
    r.z = (unsigned)__double2loint(b); r.w = (unsigned)__double2hiint(b);
    return r;
  };
  auto f4 = [](float a, float b, float c, float d) {
    return make_uint4(__float_as_uint(a), __float_as_uint(b), __float_as_uint(c), __float_as_uint(d));
  };
  const double flags = __longlong_as_double(slow ? 1ll : 0ll);
  uint4 *o = fat + (size_t)s * FatD<D, PACK>::CHUNKS;
  if constexpr (D == 3) {
    o[0] = d2(x0[0], x0[1]);
    o[1] = d2(x0[2], f[0]);
    o[2] = d2(g[0], g[1]);
    o[3] = d2(g[2], flags);
    o[4] = f4(g32[0], g32[1], g32[2], g32[3]);
    o[5] = f4(g32[4], g32[5], g32[6], g32[7]);
    o[6] = f4(g32[8], 0.f, 0.f, 0.f);
    o[7] = make_uint4((unsigned)nb[0], (unsigned)nb[1], (unsigned)nb[2], (unsigned)nb[3]);
  } else if constexpr (PACK) {
    auto n24 = [&](int j) -> unsigned { return nb[j] < 0 ? kPackNb24 : (unsigned)nb[j]; };   // (ns < 2^24 - 1: checked by the host)
    o[0] = d2(x0[0], x0[1]);
    o[1] = d2(x0[2], x0[3]);
    o[2] = d2(f[0], g[0]);
    o[3] = d2(g[1], g[2]);
    uint4 c4 = d2(g[3], 0.0);
    c4.z = (unsigned)W[0]; c4.w = (unsigned)(W[0] >> 32);
    o[4] = c4;
    o[5] = make_uint4((unsigned)W[1], (unsigned)(W[1] >> 32), (unsigned)W[2], (unsigned)(W[2] >> 32));
    o[6] = make_uint4((unsigned)W[3], (unsigned)(W[3] >> 32), (unsigned)W[4], (unsigned)(W[4] >> 32));
    // 128-bit little-endian string: n_j in bits [24 j, 24 j + 24), flags in bits [120, 128)
    const unsigned a0 = n24(0), a1 = n24(1), a2 = n24(2), a3 = n24(3), a4 = n24(4);
    o[7] = make_uint4(a0 | (a1 << 24), (a1 >> 8) | (a2 << 16), (a2 >> 16) | (a3 << 8), a4 | ((slow ? 1u : 0u) << 24));
  } else {
    const unsigned fl = slow ? 1u : 0u;
    o[0] = d2(x0[0], x0[1]);
    o[1] = d2(x0[2], x0[3]);
    o[2] = f4(g32[0], g32[1], g32[2], g32[3]);
    o[3] = f4(g32[4], g32[5], g32[6], g32[7]);
    o[4] = f4(g32[8], g32[9], g32[10], g32[11]);
    o[5] = f4(g32[12], g32[13], g32[14], g32[15]);
    o[6] = make_uint4((unsigned)nb[0], (unsigned)nb[1], (unsigned)nb[2], (unsigned)nb[3]);
    o[7] = make_uint4((unsigned)nb[4], fl, 0u, 0u);
    uint4 *e = fat + (size_t)ns * FatD<D>::CHUNKS + (size_t)s * FatD<D>::EPI_CHUNKS;   // interpolant records follow
    e[0] = d2(f[0], g[0]);
    e[1] = d2(g[1], g[2]);
    e[2] = d2(g[3], 0.0);
    e[3] = make_uint4(0u, 0u, 0u, 0u);
  }
}

// start copies: record ns + c = a copy of the record of hint[c], 16-byte chunk per thread (MeshDView::start_copies)
__global__ void k_start_copies(long long cells, int ns, int chunks, const int32_t *__restrict__ hint, uint4 *__restrict__ fat) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= cells * chunks) return;
  const long long c = t / chunks;
  const int k = (int)(t % chunks);
  fat[((size_t)ns + (size_t)c) * chunks + k] = fat[(size_t)hint[c] * chunks + k];
}

// dD fast pass on walk records.  Lane state machine as k_interpD_fast; per iteration the warp stages the 32 records
// of this step into shared memory with coalesced cp.async (FatD<D>::LPR lanes x 16 B per record, one or two
// 128-byte lines per record instead of D+2 scattered gathers), then every lane reads its own record with LDS.128.
// A step is D dot products against the fp32 barycentric gradients with the bound kFatKD * max|G_i| * |d|_1.
// RING = true: per-warp prefetch ring for queries and start hints, as in k_interp2_fat, but fed through registers
// (the record staging owns the cp.async groups, which retire in order): at every block-of-32 transition a lane
// stores the query / hint it loaded one transition ago into the ring and issues the loads for the next block, so a
// refill is D+1 LDS and the only dependent global latency of an iteration is the record itself.
// Ring per warp: 3 blocks of queries (block b, b+1 complete, b+2 just stored) + 2 blocks of hints.
template <int D>
struct RingD {
  static constexpr int QBYTES = 96 * D * 8, HBYTES = 64 * 4;
  static constexpr int BYTES = QBYTES + HBYTES;
};
// min CTAs per SM of the packed 4D walk-record kernel: ptxas's own choice is 107 registers = 4 CTAs, and ncu showed the
// kernel latency-bound there (DRAM 61 % busy, profiles/r02_k_interpD_fat4_packed_4cta_ncu.txt); measured on cfg4-4D with
// 1 / 6 / 7 / 8: 14.6 / 16.3 / 12.5 / 11.1 Gq/s (80 registers at 6; 72 and 64 spill into the staging loop)
#ifndef DLI_X_FATD4_BLOCKS
#define DLI_X_FATD4_BLOCKS 6
#endif
// 3D: 6 CTAs per SM (80 registers) is the measured optimum (round 1: ptxas chose 79 by itself; 7 / 8 CTAs spill and
// lose).  Pinned here because the round-2 additions (start copies, record index checks) let ptxas drift to 96
// registers = 5 CTAs, which cost 10 % (32.7 -> 29.5 Gq/s on cfg4-3D, profiles/r02_k_interpD_fat3_ncu.txt).
#ifndef DLI_X_FATD3_BLOCKS
#define DLI_X_FATD3_BLOCKS 6
#endif
template <int D, bool STATS, bool RING = false, bool PACK = false>
__global__ void __launch_bounds__(128, (D == 4 && PACK) ? DLI_X_FATD4_BLOCKS : (D == 3 ? DLI_X_FATD3_BLOCKS : 1))
k_interpD_fat(MeshDView m, int64_t n, int per_warp, const double *__restrict__ xs, double *__restrict__ ys,
              int32_t *__restrict__ simplex_out, uint8_t *__restrict__ status_out, int32_t *__restrict__ queue,
              unsigned int *__restrict__ queue_count, unsigned int queue_cap, unsigned long long *__restrict__ stats) {
  using F = FatD<D, PACK>;
  constexpr unsigned FULL = 0xffffffffu;
  extern __shared__ __align__(16) unsigned char stage_smem[];
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * per_warp;
  if (base >= n) return;  // warp-uniform
  const int64_t end = min(n, base + (int64_t)per_warp);
  uint4 *stage = reinterpret_cast<uint4 *>(stage_smem) + (size_t)(threadIdx.x >> 5) * (F::SMEM_PER_WARP / 16);
  const uint4 *mine = stage + lane * F::SSTRIDE;                    // this lane's staged record
  // SPLIT (D = 4): the interpolant record of the simplex a lane's previous query ended in is fetched into the lane's
  // own slot together with the next step's walk records, so its latency hides behind that staging
  uint4 *eslot = stage + 32 * F::SSTRIDE + lane * F::ESTRIDE;
  const unsigned edst = (unsigned)__cvta_generic_to_shared(eslot);
  int64_t pqi = -1;   // query whose value is pending on its interpolant record
  int ps = 0;         // ... and the simplex it ended in
  const int rsub = lane / F::LPR, csub = lane % F::LPR;             // staging role: record within the group, chunk
  const unsigned sdst = (unsigned)__cvta_generic_to_shared(stage + csub);
  int64_t next = base;
  bool active = false;
  double p[D];
#pragma unroll
  for (int k = 0; k < D; k++) p[k] = 0.0;
  int s = 0, from = -1;
  int64_t qi = 0;
  unsigned int steps = 0;
  StatsAcc st;
  // ---- prefetch ring (RING only)
  double *ringq = nullptr;
  int *ringh = nullptr;
  double hq[D];      // query (blk + 2) * 32 + lane, in flight since the previous transition
  int hh = 0;        // hint of query (blk + 1) * 32 + lane, in flight since the previous transition
  int blk = 0, b3 = 0;  // block of 32 that `next` is in; blk % 3
#pragma unroll
  for (int k = 0; k < D; k++) hq[k] = 0.0;
  auto load_q = [&](int b, double *q) {   // query b * 32 + lane of this warp's range (zeros past the end)
    const int64_t j = base + (int64_t)b * 32 + lane;
#pragma unroll
    for (int k = 0; k < D; k++) q[k] = j < end ? __ldcs(xs + j * D + k) : 0.0;
  };
  auto store_q = [&](int bm3, const double *q) {
#pragma unroll
    for (int k = 0; k < D; k++) ringq[(bm3 * 32 + lane) * D + k] = q[k];
  };
  if (RING) {
    const int nw = blockDim.x >> 5, wib = threadIdx.x >> 5;
    unsigned char *rb = stage_smem + (size_t)nw * F::SMEM_PER_WARP + (size_t)wib * RingD<D>::BYTES;
    ringq = reinterpret_cast<double *>(rb);
    ringh = reinterpret_cast<int *>(rb + RingD<D>::QBYTES);
    // state "just entered block 0": queries of blocks 0..2 and hints of blocks 0, 1 in the ring; the hint of
    // block 2 and the queries of block 3 in flight
    double q0[D], q1[D];
    load_q(0, q0); load_q(1, q1); load_q(2, hq);
    store_q(0, q0); store_q(1, q1); store_q(2, hq);
    ringh[lane] = startD<D>(m, q0);
    ringh[32 + lane] = startD<D>(m, q1);
    hh = startD<D>(m, hq);
    load_q(3, hq);
    __syncwarp();
  }
  for (;;) {
    if (RING) {
      const int nb = (int)((next - base) >> 5);
      if (nb != blk) {  // warp-uniform: `next` entered block nb = blk + 1 (a refill consumes at most 32 queries)
        blk = nb;
        b3 = b3 == 2 ? 0 : b3 + 1;
        // (all ring reads of block blk - 1 were consumed by the previous iterations' arithmetic)
        ringh[((blk + 1) & 1) * 32 + lane] = hh;                 // hints of block blk + 1
        store_q(b3 == 0 ? 2 : b3 - 1, hq);                       // queries of block blk + 2  ((blk + 2) % 3 == (b3 + 2) % 3)
        hh = startD<D>(m, hq);                                    // hint of block blk + 2: in flight until the next transition
        load_q(blk + 3, hq);                                     // queries of block blk + 3: likewise
        __syncwarp();
      }
    }
    unsigned need = __ballot_sync(FULL, !active);
    if (need) {
      int64_t idx = next + __popc(need & ((1u << lane) - 1u));
      if (!active && idx < end) {
        if (RING) {
          const int lb = (int)((idx - base) >> 5);               // blk or blk + 1
          const int m3 = lb == blk ? b3 : (b3 == 2 ? 0 : b3 + 1);
          const int sl = (int)((idx - base) & 31);
#pragma unroll
          for (int k = 0; k < D; k++) p[k] = ringq[(m3 * 32 + sl) * D + k];
          s = ringh[(lb & 1) * 32 + sl];
        } else {
#pragma unroll
          for (int k = 0; k < D; k++) p[k] = __ldcs(xs + idx * D + k);
          s = startD<D>(m, p);
        }
        qi = idx;
        from = -1;
        steps = 0;
        active = true;
      }
      next += __popc(need);
    }
    if (!__any_sync(FULL, active || (F::SPLIT && pqi >= 0))) break;
    // ---- stage the records of this step
    const int fs = active ? s : -1;
    __syncwarp();  // every lane has finished reading the previous step's records
#pragma unroll
    for (int j = 0; j < F::LPR; j++) {     // 32 / (32 / LPR) instructions, 32 / LPR records each
      const int r = j * (32 / F::LPR) + rsub;
      const int fr = __shfl_sync(FULL, fs, r);
      if (fr >= 0 && csub < F::LOADED)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst + (unsigned)(r * F::SSTRIDE * 16)),
                     "l"(m.fat + (size_t)DLI_IDX(fr, m.nrec) * F::CHUNKS + csub)
                     : "memory");
    }
    if (F::SPLIT && pqi >= 0) {
      const uint4 *er = m.epi + (size_t)DLI_IDX(ps, m.ns) * F::EPI_CHUNKS;
#pragma unroll
      for (int c = 0; c < F::EPI_LOADED; c++)
        asm volatile("cp.async.cg.shared.global" DLI_L2_64B " [%0], [%1], 16;" ::"r"(edst + (unsigned)(c * 16)), "l"(er + c) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    if (F::SPLIT && pqi >= 0) {   // finish the pending query: y = f0 + g . d, same operations as the in-record form
      auto lo = [](const uint4 &c) { return __hiloint2double((int)c.y, (int)c.x); };
      auto hi = [](const uint4 &c) { return __hiloint2double((int)c.w, (int)c.z); };
      const uint4 e0 = eslot[0], e1 = eslot[1], e2 = eslot[2], d01 = eslot[3], d23 = eslot[4];
      double y = 0.0;
      y = fma(lo(e2), hi(d23), y);
      y = fma(hi(e1), lo(d23), y);
      y = fma(lo(e1), hi(d01), y);
      y = fma(hi(e0), lo(d01), y);
      __stcs(ys + pqi, lo(e0) + y);
      pqi = -1;
    }
    if (active) {
      steps++;
      double x0[D], f0, g[D];
      float G[D * D];
      double rowscale[D];   // PACK: 2^e_i of the block-floating-point rows (unused otherwise)
      int nb[D + 1];
      unsigned long long flags;
      {
        auto lo = [](const uint4 &c) { return __hiloint2double((int)c.y, (int)c.x); };
        auto hi = [](const uint4 &c) { return __hiloint2double((int)c.w, (int)c.z); };
        if constexpr (D == 3) {
          uint4 c0 = mine[0], c1 = mine[1], c2 = mine[2], c3 = mine[3], c4 = mine[4], c5 = mine[5], c6 = mine[6], c7 = mine[7];
          x0[0] = lo(c0); x0[1] = hi(c0); x0[2] = lo(c1); f0 = hi(c1);
          g[0] = lo(c2); g[1] = hi(c2); g[2] = lo(c3);
          flags = ((unsigned long long)c3.w << 32) | c3.z;
          G[0] = __uint_as_float(c4.x); G[1] = __uint_as_float(c4.y); G[2] = __uint_as_float(c4.z); G[3] = __uint_as_float(c4.w);
          G[4] = __uint_as_float(c5.x); G[5] = __uint_as_float(c5.y); G[6] = __uint_as_float(c5.z); G[7] = __uint_as_float(c5.w);
          G[8] = __uint_as_float(c6.x);
          nb[0] = (int)c7.x; nb[1] = (int)c7.y; nb[2] = (int)c7.z; nb[3] = (int)c7.w;
        } else if constexpr (PACK) {
          const uint4 c0 = mine[0], c1 = mine[1], c2 = mine[2], c3 = mine[3], c4 = mine[4], c5 = mine[5], c6 = mine[6], c7 = mine[7];
          x0[0] = lo(c0); x0[1] = hi(c0); x0[2] = lo(c1); x0[3] = hi(c1);
          f0 = lo(c2); g[0] = hi(c2); g[1] = lo(c3); g[2] = hi(c3); g[3] = lo(c4);
          const unsigned long long W[5] = {((unsigned long long)c4.w << 32) | c4.z, ((unsigned long long)c5.y << 32) | c5.x,
                                           ((unsigned long long)c5.w << 32) | c5.z, ((unsigned long long)c6.y << 32) | c6.x,
                                           ((unsigned long long)c6.w << 32) | c6.z};
          // mantissas as exact small integers in fp32 (magic-number conversion: 2^23 + field, minus 2^23 + 2^17)
          auto mant = [](unsigned fld) { return __uint_as_float(0x4B000000u | fld) - 8519680.0f; };
          G[0] = mant(pack_field<0, 18>(W));    G[1] = mant(pack_field<18, 18>(W));   G[2] = mant(pack_field<36, 18>(W));   G[3] = mant(pack_field<54, 18>(W));
          G[4] = mant(pack_field<80, 18>(W));   G[5] = mant(pack_field<98, 18>(W));   G[6] = mant(pack_field<116, 18>(W));  G[7] = mant(pack_field<134, 18>(W));
          G[8] = mant(pack_field<160, 18>(W));  G[9] = mant(pack_field<178, 18>(W));  G[10] = mant(pack_field<196, 18>(W)); G[11] = mant(pack_field<214, 18>(W));
          G[12] = mant(pack_field<240, 18>(W)); G[13] = mant(pack_field<258, 18>(W)); G[14] = mant(pack_field<276, 18>(W)); G[15] = mant(pack_field<294, 18>(W));
          // row scales 2^e as doubles (exponent byte e + 128 -> biased fp64 exponent e + 1023)
          rowscale[0] = __hiloint2double((int)((pack_field<72, 8>(W) + 895u) << 20), 0);
          rowscale[1] = __hiloint2double((int)((pack_field<152, 8>(W) + 895u) << 20), 0);
          rowscale[2] = __hiloint2double((int)((pack_field<232, 8>(W) + 895u) << 20), 0);
          rowscale[3] = __hiloint2double((int)((pack_field<312, 8>(W) + 895u) << 20), 0);
          const unsigned a0 = c7.x & kPackNb24, a1 = (c7.x >> 24) | ((c7.y & 0xffffu) << 8), a2 = (c7.y >> 16) | ((c7.z & 0xffu) << 16),
                         a3 = c7.z >> 8, a4 = c7.w & kPackNb24;
          nb[0] = a0 == kPackNb24 ? -1 : (int)a0; nb[1] = a1 == kPackNb24 ? -1 : (int)a1; nb[2] = a2 == kPackNb24 ? -1 : (int)a2;
          nb[3] = a3 == kPackNb24 ? -1 : (int)a3; nb[4] = a4 == kPackNb24 ? -1 : (int)a4;
          flags = c7.w >> 24;
        } else {
          uint4 c0 = mine[0], c1 = mine[1], c6 = mine[6], c7 = mine[7];
          x0[0] = lo(c0); x0[1] = hi(c0); x0[2] = lo(c1); x0[3] = hi(c1);
          f0 = 0.0;
#pragma unroll
          for (int k = 0; k < D; k++) g[k] = 0.0;
#pragma unroll
          for (int q = 0; q < 4; q++) {
            uint4 c = mine[2 + q];
            G[4 * q] = __uint_as_float(c.x); G[4 * q + 1] = __uint_as_float(c.y);
            G[4 * q + 2] = __uint_as_float(c.z); G[4 * q + 3] = __uint_as_float(c.w);
          }
          nb[0] = (int)c6.x; nb[1] = (int)c6.y; nb[2] = (int)c6.z; nb[3] = (int)c6.w;
          nb[4] = (int)c7.x;
          flags = c7.y;
        }
      }
      double d[D], d1 = 0.0;
#pragma unroll
      for (int k = 0; k < D; k++) { d[k] = p[k] - x0[k]; d1 += fabs(d[k]); }
      // barycentric coordinates b[1..D] = G_i . d, b[0] = 1 - sum; bounds e[j]
      double b[D + 1], e[D + 1], sb = 0.0, se = 0.0;
#pragma unroll
      for (int i = 0; i < D; i++) {
        double acc = 0.0;
        float gm = 0.f;
#pragma unroll
        for (int k = 0; k < D; k++) { acc = fma((double)G[i * D + k], d[k], acc); gm = fmaxf(gm, fabsf(G[i * D + k])); }
        if constexpr (PACK) {   // block floating point: exact power-of-two scaling of the row
          acc *= rowscale[i];
          e[i + 1] = (kFatKP * ((double)gm * rowscale[i])) * d1;
        } else {
          e[i + 1] = (kFatKD * (double)gm) * d1;
        }
        b[i + 1] = acc;
        sb += acc;
        se += e[i + 1];
      }
      b[0] = 1.0 - sb;
      e[0] = se * 1.0000001 + 8.881784197001252e-16;  // + 2^-50: roundings of the sum and of 1 - sum
      int mv = -1;
      bool all_pos = true;
#pragma unroll
      for (int j = D; j >= 0; j--) {
        const bool came = nb[j] == from;   // from = -1 only matches a hull slot, which is never "came from"... guard below
        if (!(came && nb[j] >= 0)) {
          if (-b[j] > e[j]) mv = j;        // lowest such j wins (loop runs downwards)
          else if (!(b[j] > e[j])) all_pos = false;
        }
      }
      bool give_up = (flags & 1ull) != 0;
      if (!give_up) {
        if (mv >= 0) {
          int t = nb[0];
#pragma unroll
          for (int j = 1; j <= D; j++) t = (mv == j) ? nb[j] : t;
          if (t < 0 || (int)steps >= m.fast_steps) give_up = true;
          else { from = s; s = t; }
        } else if (all_pos) {
          if (ys) {
            if constexpr (F::SPLIT) {   // value written once the interpolant record has arrived (next iteration)
              auto pk = [](double a, double b) {
                return make_uint4((unsigned)__double2loint(a), (unsigned)__double2hiint(a), (unsigned)__double2loint(b),
                                  (unsigned)__double2hiint(b));
              };
              eslot[3] = pk(d[0], d[1]);
              eslot[4] = pk(d[D - 2], d[D - 1]);
              pqi = qi;
              ps = s;
            } else {
              double y = 0.0;
#pragma unroll
              for (int k = D - 1; k >= 0; k--) y = fma(g[k], d[k], y);
              __stcs(ys + qi, f0 + y);
            }
          }
          if (simplex_out) __stcs(simplex_out + qi, __ldg(m.orig + DLI_IDX(record_simplex(m, s), m.ns)));
          if (status_out) status_out[qi] = DLI_STATUS_OK;
          if (STATS) { st.steps += steps; st.max_steps = max(st.max_steps, steps); }
          active = false;
        } else {
          give_up = true;
        }
      }
      if (give_up) {
        enqueue_straggler<0>(queue, queue_count, queue_cap, qi, record_simplex(m, s), ys, simplex_out);
        active = false;
      }
    }
  }
  if (STATS) flush_stats(stats, st);
}

// ---------------------------------------------------------------------------------------------
// Plot data (2D): get_line_segments (src/delaunay_2_interp.h:842-862) with the concavity flag of is_edge_concave
// (:766-803), every finite edge once, from the face with the lower caller id (CGAL's edge order is arbitrary).
// ---------------------------------------------------------------------------------------------
// exact sign of a x + b y + c for double coefficients (Line_2::oriented_side, filtered-exact in Epick)
__device__ inline int line_side_exact(double a, double b, double c, double x, double y) {
  double e[5], p, q;
  int n = 0;
  two_prod(a, x, p, q);
  grow(e, n, q); grow(e, n, p);
  two_prod(b, y, p, q);
  grow(e, n, q); grow(e, n, p);
  grow(e, n, c);
  return expansion_sign(e, n);
}
// CGAL Cartesian Line_2(P, Q) in plain doubles (same construction as project_on_line)
__device__ inline void line_from_points(double Px, double Py, double Qx, double Qy, double &a, double &b, double &c) {
  if (Py == Qy) {
    a = 0.0;
    if (Qx > Px) { b = 1.0; c = -Py; }
    else if (Qx == Px) { b = 0.0; c = 0.0; }
    else { b = -1.0; c = Py; }
  } else if (Qx == Px) {
    b = 0.0;
    if (Qy > Py) { a = -1.0; c = Px; }
    else { a = 1.0; c = -Px; }
  } else {
    a = Py - Qy;
    b = Qx - Px;
    c = -Px * a - Py * b;
  }
}
__device__ __forceinline__ bool edge_emitted(const int32_t *__restrict__ orig, int f, int nb) {
  return nb < 0 || __ldg(orig + f) < __ldg(orig + nb);
}
// per face: number of edges it emits
__global__ void k_edge_count(int ns, const int4 *__restrict__ frec, const int32_t *__restrict__ orig, int *__restrict__ cnt) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= ns) return;
  int4 r0 = frec[2 * (size_t)f], r1 = frec[2 * (size_t)f + 1];
  cnt[f] = (int)edge_emitted(orig, f, r0.w) + (int)edge_emitted(orig, f, r1.x) + (int)edge_emitted(orig, f, r1.y);
}
// segments [ne][2][3] = (P.x, P.y, f_P, Q.x, Q.y, f_Q) with P = vertex ccw(j), Q = vertex cw(j) of the emitting face
__global__ void k_edge_emit(int ns, const int4 *__restrict__ frec, const double4 *__restrict__ vrec,
                            const int32_t *__restrict__ orig, const int *__restrict__ offs, double *__restrict__ seg,
                            int32_t *__restrict__ concave) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= ns) return;
  int4 r0 = frec[2 * (size_t)f], r1 = frec[2 * (size_t)f + 1];
  const int v[3] = {r0.x, r0.y, r0.z}, nbs[3] = {r0.w, r1.x, r1.y};
  const double4 V[3] = {vrec[v[0]], vrec[v[1]], vrec[v[2]]};
  int o = offs[f];
  double2 g1 = make_double2(0.0, 0.0);
  bool have_g1 = false;
  for (int j = 0; j < 3; j++) {
    const int nb = nbs[j];
    if (!edge_emitted(orig, f, nb)) continue;
    const double4 P = V[(j + 1) % 3], Q = V[(j + 2) % 3], R = V[j];
    double *row = seg + 6 * (size_t)o;
    row[0] = P.x; row[1] = P.y; row[2] = P.z; row[3] = Q.x; row[4] = Q.y; row[5] = Q.z;
    int flag = 1;                                   // an infinite face on either side (:778-780)
    if (nb >= 0) {
      if (!have_g1) { g1 = face_gradient2(V[0], V[1], V[2]); have_g1 = true; }
      int4 n0 = frec[2 * (size_t)nb];
      const double2 g2 = face_gradient2(vrec[n0.x], vrec[n0.y], vrec[n0.z]);
      double dgx = g2.x - g1.x, dgy = g2.y - g1.y;
      double a, b, c;
      line_from_points(P.x, P.y, Q.x, Q.y, a, b, c);
      const int side1 = line_side_exact(a, b, c, R.x, R.y);
      const int sl = line_side_exact(a, b, c, P.x - 1.0, P.y), sd = line_side_exact(a, b, c, P.x, P.y - 1.0);
      if (sl == 0) dgx = 0.0; else if (sl != side1) dgx = -dgx;
      if (sd == 0) dgy = 0.0; else if (sd != side1) dgy = -dgy;
      flag = (dgx <= 0.0 && dgy <= 0.0) ? 1 : 0;
    }
    concave[o] = flag;
    o++;
  }
}
// min / max of (x, y, f) over the n end points ([n][3]) of the segments: per-block partials [gridDim.x][6]
__global__ void k_point_minmax(int64_t n, const double *__restrict__ pts, double *__restrict__ part) {
  double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
    const double3 a = make_double3(pts[3 * v], pts[3 * v + 1], pts[3 * v + 2]);
    mn[0] = fmin(mn[0], a.x); mn[1] = fmin(mn[1], a.y); mn[2] = fmin(mn[2], a.z);
    mx[0] = fmax(mx[0], a.x); mx[1] = fmax(mx[1], a.y); mx[2] = fmax(mx[2], a.z);
  }
  __shared__ double sh[6][8];
  for (int k = 0; k < 3; k++)
    for (int o = 16; o > 0; o >>= 1) {
      mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) for (int k = 0; k < 3; k++) { sh[k][w] = mn[k]; sh[3 + k][w] = mx[k]; }
  __syncthreads();
  if (threadIdx.x < 6) {
    double r = sh[threadIdx.x][0];
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) r = threadIdx.x < 3 ? fmin(r, sh[threadIdx.x][i]) : fmax(r, sh[threadIdx.x][i]);
    part[(size_t)blockIdx.x * 6 + threadIdx.x] = r;
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

// ---- incremental update (dli_mesh_update): the same record builders over a list of device face indices ----
// canonical face records + id map of the listed faces
__global__ void k_update_faces(int n, const int32_t *__restrict__ idx, const int4 *__restrict__ rows,
                               const int32_t *__restrict__ ids, int4 *__restrict__ frec, int32_t *__restrict__ orig) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int f = idx[i];
  frec[2 * (size_t)f] = rows[2 * (size_t)i];
  frec[2 * (size_t)f + 1] = rows[2 * (size_t)i + 1];
  orig[f] = ids[i];
}
__global__ void k_update_values(int n, const int32_t *__restrict__ vid, const double *__restrict__ val, double4 *__restrict__ vrec) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) vrec[vid[i]].z = val[i];
}
__global__ void k_pack_faces_list(int n, const int32_t *__restrict__ idx, const int4 *__restrict__ frec,
                                  ulonglong2 *__restrict__ fpack) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int f = idx[i];
  int4 a = frec[2 * (size_t)f], b = frec[2 * (size_t)f + 1];
  unsigned long long n0 = a.w < 0 ? kPackMask : (unsigned)a.w, n1 = b.x < 0 ? kPackMask : (unsigned)b.x,
                     n2 = b.y < 0 ? kPackMask : (unsigned)b.y;
  ulonglong2 r;
  r.x = (unsigned long long)(unsigned)a.x | ((unsigned long long)(unsigned)a.y << 21) | ((unsigned long long)(unsigned)a.z << 42);
  r.y = n0 | (n1 << 21) | (n2 << 42);
  fpack[f] = r;
}
// walk records of the listed faces (absolute neighbour ids); n_slow_delta accumulates (new slow flag - old slow flag),
// faces at index >= old_ns had no record
__global__ void k_build_fat2_list(int n, const int32_t *__restrict__ idx, int old_ns, const int4 *__restrict__ frec,
                                  const double4 *__restrict__ vrec, uint4 *__restrict__ fat, int *__restrict__ n_slow_delta) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int f = idx[i];
  int4 r0 = frec[2 * (size_t)f], r1 = frec[2 * (size_t)f + 1];
  auto enc = [](int nb) -> unsigned long long { return nb < 0 ? (unsigned long long)kPackMask : (unsigned long long)(unsigned)nb; };
  unsigned long long nbr = enc(r0.w) | (enc(r1.x) << 21) | (enc(r1.y) << 42);
  int was = 0;
  if (f < old_ns) was = (fat[4 * (size_t)f + 3].w >> 31) & 1u;   // kFatSlow = bit 63 of the neighbour word
  const int is = make_fat2(vrec[r0.x], vrec[r0.y], vrec[r0.z], nbr, fat + 4 * (size_t)f) ? 1 : 0;
  if (is != was) atomicAdd(n_slow_delta, is - was);
}

// dD incremental update: canonical simplex records + id map of the listed simplices (rows: ss ints each, already in
// device indices), new values of existing vertices
__global__ void k_update_simplices_d(int n, const int32_t *__restrict__ idx, const int32_t *__restrict__ rows, int ss,
                                     const int32_t *__restrict__ ids, int32_t *__restrict__ srec, int32_t *__restrict__ orig) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = idx[i];
  for (int k = 0; k < ss; k++) srec[(size_t)s * ss + k] = rows[(size_t)i * ss + k];
  orig[s] = ids[i];
}
__global__ void k_update_values_d(int n, const int32_t *__restrict__ vid, const double *__restrict__ val, double *__restrict__ vrec,
                                  int vs, int dim) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) vrec[(size_t)vid[i] * vs + dim] = val[i];
}
// slow flags of the dD walk records (recounted after a patch: a handful of bytes per simplex)
template <int D, bool PACK>
__global__ void k_count_slow_d(int ns, const uint4 *__restrict__ fat, unsigned int *__restrict__ n_slow) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= ns) return;
  const uint4 *r = fat + (size_t)s * FatD<D, PACK>::CHUNKS;
  unsigned fl;
  if constexpr (D == 3) fl = r[3].z & 1u;          // flags: low word of the second double of chunk 3
  else if constexpr (PACK) fl = (r[7].w >> 24) & 1u;
  else fl = r[7].y & 1u;
  if (fl) atomicAdd(n_slow, 1u);
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
// Snapshot validation (dli_mesh_load): every index of the canonical records, of the hint grid and of the hull table in
// range.  The derived buffers (walk records, packed faces, slot tables) are NOT decoded here; the per-buffer checksums
// of the snapshot file (snap_checksum in dli_capi.cu) are what protects them against a damaged file.
// ---------------------------------------------------------------------------------------------
__global__ void k_validate_ids(long long ns, long long nv, long long nh, int dim, int ss, const int32_t *__restrict__ srec,
                               long long cells, const int32_t *__restrict__ hint, const int4 *__restrict__ hull,
                               unsigned int *__restrict__ bad) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool wrong = false;
  if (t < ns) {
    for (int k = 0; k <= dim; k++) {
      const int v = srec[t * ss + k], nb = srec[t * ss + dim + 1 + k];
      wrong |= v < 0 || v >= nv;
      wrong |= nb >= ns || (dim == 2 ? nb < -nh : nb < -1);      // 2D: ~hull edge index; dD: -1
    }
  }
  if (t < cells) wrong |= hint[t] < 0 || hint[t] >= ns;
  if (hull && t < nh) {
    const int4 h = hull[t];
    wrong |= h.x < 0 || h.x >= ns || h.y < 0 || h.y > 2 || h.z < -1 || h.z >= nh || h.w < -1 || h.w >= nh;
  }
  if (wrong) atomicAdd(bad, 1u);
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
