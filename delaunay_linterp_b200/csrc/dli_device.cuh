// dli_device.cuh -- device-side mesh views and the per-query device functions
// (jump-and-walk locate, outside-hull handling, barycentric + gather epilogue).
//
// Reference behaviour restated (paths relative to the reference tree):
//   2D driver   src/delaunay_2_interp.h:515-572      dD driver  src/delaunay_d_interp.h:988-1010
//   2D epilogue src/delaunay_2_interp.h:195-239, 563-569 (Eigen dynamic determinant = partial-pivot LU)
//   dD epilogue src/delaunay_d_interp.h:601-650, 998-1004 (exact differences truncated, exact sum truncated)
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "dli_exact.cuh"
#include "dli_refmath.h"
#include "../../include/dli.h"

namespace dli {

// -DDLI_BOUNDS_CHECK: every index the kernels form into a mesh buffer is checked on the device (printf + trap, so the
// next CUDA call fails loudly).  compute-sanitizer is not available on the build pool; scripts/bounds_check.sh builds
// this variant and runs the small parity tests against it.  Off (the shipped build): the macro is the bare index.
#ifdef DLI_BOUNDS_CHECK
__device__ __noinline__ void dli_bounds_fail(long long i, long long n, int line) {
  printf("DLI_BOUNDS_CHECK: index %lld outside [0, %lld) at dli source line %d\n", i, n, line);
  __trap();
}
#define DLI_IDX(i, n) (((long long)(i) < 0 || (long long)(i) >= (long long)(n)) ? (::dli::dli_bounds_fail((long long)(i), (long long)(n), __LINE__), (i)) : (i))
#else
#define DLI_IDX(i, n) (i)
#endif

// DRAM fetch size of scattered record loads.  On B200 a global load that misses L2 brings in the whole 128-byte line
// whatever it asked for (126.7 B of DRAM traffic per random 16 / 32 / 64-byte request, profiles/r02_dram_granularity_ncu.txt;
// cudaLimitMaxL2FetchGranularity changes nothing); the PTX prefetch-size qualifier .L2::64B limits the fetch to the
// 64-byte half line (63.7 B per request).  MEASURED AND NOT USED: with the qualifier on every record gather and hint
// load, cfg5 moves 41 % fewer DRAM bytes (12.8 vs 21.7 GB per 125M queries) in the SAME time (3.87 vs 3.90 ms) -- the
// walk kernels are bound by the rate of random DRAM accesses (~47 G line fetches/s in the probe, whatever their size),
// not by bytes -- and cfg2 gets 4 % slower (52.4 vs 54.4 Gq/s: the other half of a fetched line is often the next
// face of the walk, Morton order).  -DDLI_X_L2_64B turns it on (scripts/build_variants.py) for experiments.
#ifdef DLI_X_L2_64B
#define DLI_L2_64B ".L2::64B"
#else
#define DLI_L2_64B ""
#endif

constexpr int kMaxProjections = 16;  // 2D outside-hull project/re-locate rounds before accepting the hull face

// ---------------------------------------------------------------------------------------------
// Device mesh views (plain structs passed by value to kernels)
// ---------------------------------------------------------------------------------------------

// 2D: vertex record double4 {x, y, f, 0} (32 B = one DRAM sector);
//     face record 2 x int4 {v0, v1, v2, n0 | n1, n2, 0, 0} (32 B); neighbour >= 0 = face id,
//     < 0 = ~(hull edge index); hull record int4 {face, slot, next, prev}.
struct Mesh2View {
  const double4 *vrec;
  const int4 *frec;
  // optional packed face records for the fast pass (meshes with <= 2^21 vertices and < 2^21-1 faces):
  // 16 B = {v0 | v1 << 21 | v2 << 42, n0 | n1 << 21 | n2 << 42}, neighbour 0x1FFFFF = hull.  Halves the
  // L2 footprint of the face array (two faces per sector).
  const ulonglong2 *fpack;
  // optional 64-byte self-contained walk records (FatFace2) for the pair-transposed fast pass k_interp2_fat
  const uint4 *fat;
  const int4 *hull;
  const int32_t *hint;
  const int32_t *orig;  // device simplex index (Morton order of centroids) -> the caller's simplex id
  // cell-indexed walk records (optional, big meshes): `fat` is laid out by hint-grid rows, the start slot of a query
  // in cell (i, j) is rowstart[j] + i (a copy of the record of face hint[j * gx + i]); slot_face maps slots to faces
  const int32_t *slot_face;
  const int *rowstart;
  double ox, oy, icx, icy;  // hint grid origin and 1/cell size
  int gx, gy;
  int ns, nh;
  int nv;          // vertices; nslots: records in `fat` (== ns unless cell-indexed)  [index checks of DLI_BOUNDS_CHECK builds]
  long long nslots;
  int max_steps;   // hard cap on simplices visited per query
  int fast_steps;  // fast-path cap; longer walks are re-run by the straggler/exact pass
  int outside_nan; // != 0: queries outside the hull report DLI_STATUS_OUTSIDE_NAN instead of being projected
  float *straggler_frac;  // mapped pinned host word: the exact pass reports the straggler fraction of its launch (queue sizing)
};

// dD (D = 3, 4): vertex record VS doubles {x_0..x_{D-1}, f, pad}; simplex record SS ints
// {v_0..v_D, n_0..n_D, pad}; neighbour < 0 = hull.
template <int D>
struct RecD {
  static constexpr int VS = (D + 1 <= 4) ? 4 : 8;        // doubles per vertex record (32 B / 64 B)
  static constexpr int SS = (2 * (D + 1) <= 8) ? 8 : 16;  // ints per simplex record (32 B / 64 B)
};
struct MeshDView {
  const double *vrec;
  const int32_t *srec;
  const int32_t *hint;
  const int32_t *orig;  // device simplex index -> the caller's simplex id
  const uint4 *fat;     // optional walk records (FatD<D>), NULL when unused
  const uint4 *epi;     // D = 4: the 64-byte affine-interpolant records that follow the walk records
  double o[DLI_MAX_DIM], ic[DLI_MAX_DIM];
  int g[DLI_MAX_DIM];
  int ns;
  int nv;            // vertices, hint cells  [index checks of DLI_BOUNDS_CHECK builds]
  long long cells;
  // start copies (4D packed records by default): `fat` holds ns canonical records followed by one COPY of the record of
  // hint[c] per hint cell c, so a query's first record is fat[ns + cell(p)] -- one DRAM access instead of two dependent
  // ones (the 4-byte hint, then the record it names).  The walk-record kernels run at the DRAM random-access rate on
  // meshes larger than L2 (~47 G line fetches/s on B200, whatever their size: profiles/r02_dram_granularity_ncu.txt),
  // so what counts is accesses per query.  A record index >= ns is a copy; its simplex is hint[index - ns].
  int start_copies;
  long long nrec;    // records in `fat`: ns, or ns + cells with start copies
  int max_steps;
  int fast_steps;
  float *straggler_frac;  // as Mesh2View::straggler_frac
};

// ---------------------------------------------------------------------------------------------
// dD walk records (fast pass k_interpD_fat).  Self-contained per simplex, in 16-byte chunks so that a warp can
// stage the 32 records of a walk step into shared memory with coalesced cp.async (LPR lanes per record):
//   p0 = vertex 0 (fp64 x D) | f0, grad f (fp64 x (D+1)): affine interpolant y = f0 + g.(p - p0) | flags (u64)
//   G (fp32, D x D): row i = gradient of the barycentric coordinate of vertex i+1, so b_{i+1} = G_i . (p - p0)
//   n_0..n_D (int32, device simplex indices, < 0 = hull)
// D = 3: 8 chunks (128 B):  c0 p0x p0y | c1 p0z f0 | c2 gx gy | c3 gz flags | c4 G0..3 | c5 G4..7 | c6 G8 - - - | c7 n0..n3
// D = 4: the 192 bytes per simplex are split so that a walk step touches ONE 128-byte line (a 192-byte record straddles
//        two): 8 chunks of geometry (128 B): c0 p0x p0y | c1 p0z p0w | c2..c5 G0..15 | c6 n0..n3 | c7 n4 flags - -
//        and, in a second array behind the ns geometry records, 4 chunks (64 B, 3 loaded) of affine interpolant:
//        e0 f0 gx | e1 gy gz | e2 gw - | e3 unused.  Only the simplex a query ends in reads its interpolant record.
// The fp32 rounding of G (and the fp64 rounding error of computing it, bounded per entry at build time to at most
// half an fp32 ulp) is covered by the filter constant kFatKD, so a certified barycentric sign is the sign of the exact
// orientation predicate on the fp64 mesh.
// ---------------------------------------------------------------------------------------------
// D = 4, PACK = true (the default for 4D meshes with < 2^24 - 1 simplices): the whole simplex in ONE 128-byte line, so a
//        query touches one line per walk step and none for the interpolant (the split layout reads a second, random
//        line per query for 40 useful bytes: 450 B per query moved against 260 B algorithmic, profiles/r01_k_interpD_fat4_split_ncu.txt):
//        c0 p0x p0y | c1 p0z p0w | c2 f0 gx | c3 gy gz | c4 gw, W0 | c5 W1 W2 | c6 W3 W4 | c7 neighbours
//        W0..W4 = 320 bits: row i of G in bits [80 i, 80 i + 80) as BLOCK FLOATING POINT -- four 18-bit mantissas
//        (q + 2^17, |q| < 2^17) and one 8-bit exponent (e + 128): G_ik ~ q_ik 2^e_i with e_i chosen from the row
//        maximum, so every entry is within 2^-17 of the ROW MAXIMUM (checked per entry at build time, together with the
//        error of computing G itself; else the slow flag).  The step's bound is relative to the row maximum anyway
//        (kFatKD * max|G_i| * |d|_1), so block scaling loses nothing but 6 bits of band: kFatKP = 2^-16.
//        c7 = five 24-bit neighbour ids (0xffffff = hull) + 8 flag bits (bit 0 = slow).
template <int D, bool PACK = false>
struct FatD {
  static_assert(!PACK || D == 4, "packed walk records exist for D = 4 only");
  static constexpr int CHUNKS = 8;                        // walk record size in 16-byte chunks (one 128-byte line)
  static constexpr int LOADED = 8;                        // chunks actually staged
  static constexpr int LPR = 8;                           // lanes cooperating on one record
  static constexpr int SSTRIDE = 9;                       // shared-memory record stride in chunks (odd: conflict-free)
  static constexpr bool SPLIT = (D == 4) && !PACK;        // affine interpolant in a separate record array
  static constexpr int EPI_CHUNKS = SPLIT ? 4 : 0;        // interpolant record stride in chunks (64 B)
  static constexpr int EPI_LOADED = SPLIT ? 3 : 0;
  static constexpr int ESTRIDE = 5;                       // per-lane shared slot: 3 interpolant chunks + d[4] (conflict-free)
  static constexpr int BYTES = (CHUNKS + EPI_CHUNKS) * 16;  // bytes per simplex in d_fat
  static constexpr int SMEM_PER_WARP = 32 * SSTRIDE * 16 + (SPLIT ? 32 * ESTRIDE * 16 : 0);
};
constexpr double kFatKD = 2.384185791015625e-07;  // 2^-22 = 4 fp32 unit roundoffs
constexpr double kFatKP = 1.52587890625e-05;      // 2^-16 = twice the block-floating-point entry error 2^-17 (packed 4D records)
constexpr unsigned kPackNb24 = 0xFFFFFFu;         // packed 4D records: 24-bit neighbour field, all ones = hull
// field [OFF, OFF + LEN) of the 320-bit string W[0..4] (compile-time position: one or two shifts)
template <int OFF, int LEN>
__host__ __device__ __forceinline__ unsigned pack_field(const unsigned long long (&W)[5]) {
  constexpr int w = OFF / 64, b = OFF % 64;
  unsigned long long v = W[w] >> b;
  if constexpr (b + LEN > 64) v |= W[w + 1] << (64 - b);
  return (unsigned)v & ((1u << LEN) - 1u);
}

struct StatsAcc {  // per-thread accumulators, flushed by the kernels when stats are enabled
  unsigned long long steps = 0, exact = 0, outside = 0;
  unsigned int max_steps = 0;
};

// ---------------------------------------------------------------------------------------------
// 2D epilogue (reference operation order)
// ---------------------------------------------------------------------------------------------

// reference operation order (dli_refmath.h: lu_det2 / vol2 / epilogue2_ref_xy, shared with the host side of small batches)
__device__ __forceinline__ double epilogue2_ref(double px, double py, const double4 &A, const double4 &B,
                                                 const double4 &C) {
  return epilogue2_ref_xy(px, py, A.x, A.y, A.z, B.x, B.y, B.z, C.x, C.y, C.z);
}
// closed form (option epilogue = 1): one reciprocal, cross products
__device__ __forceinline__ double epilogue2_fast(double px, double py, const double4 &A, const double4 &B,
                                                  const double4 &C) {
  double w = (B.x - A.x) * (C.y - A.y) - (C.x - A.x) * (B.y - A.y);
  double w0 = (B.x - px) * (C.y - py) - (C.x - px) * (B.y - py);
  double w1 = (C.x - px) * (A.y - py) - (A.x - px) * (C.y - py);
  double w2 = (A.x - px) * (B.y - py) - (B.x - px) * (A.y - py);
  double r = 1.0 / w;
  return (w0 * A.z + w1 * B.z + w2 * C.z) * r;
}

// (project_on_line, hull_edge_score: dli_refmath.h)

// a start hint (one 4-byte entry of the hint grid)
__device__ __forceinline__ int ldg_hint(const int32_t *p) {
  int v;
  asm volatile("ld.global.nc" DLI_L2_64B ".b32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int hint2(const Mesh2View &m, double x, double y) {
  int i = (int)((x - m.ox) * m.icx);
  int j = (int)((y - m.oy) * m.icy);
  i = max(0, min(m.gx - 1, i));
  j = max(0, min(m.gy - 1, j));
  return ldg_hint(m.hint + DLI_IDX((size_t)j * m.gx + i, (size_t)m.gx * m.gy));
}

// hint-grid cell index of a point (the start slot of cell-indexed walk records)
__device__ __forceinline__ int cell2(const Mesh2View &m, double x, double y) {
  int i = (int)((x - m.ox) * m.icx);
  int j = (int)((y - m.oy) * m.icy);
  i = max(0, min(m.gx - 1, i));
  j = max(0, min(m.gy - 1, j));
  return __ldg(m.rowstart + DLI_IDX(j, m.gy + 1)) + i;
}

struct Face2 {
  int v0, v1, v2, n0, n1, n2;
};
// one 32-byte record = one sector = one 256-bit load (LDG.E.256, sm_100+)
__device__ __forceinline__ Face2 load_face(const Mesh2View &m, int f) {
  Face2 r;
  int p0, p1;
  asm volatile("ld.global.nc" DLI_L2_64B ".v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.v0), "=r"(r.v1), "=r"(r.v2), "=r"(r.n0), "=r"(r.n1), "=r"(r.n2), "=r"(p0), "=r"(p1)
               : "l"(m.frec + 2 * (size_t)DLI_IDX(f, m.ns)));
  return r;
}
constexpr unsigned kPackMask = 0x1FFFFFu;
__device__ __forceinline__ Face2 load_face_packed(const Mesh2View &m, int f) {
  unsigned long long w0, w1;
  asm volatile("ld.global.nc" DLI_L2_64B ".v2.u64 {%0,%1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(m.fpack + DLI_IDX(f, m.ns)));
  Face2 r;
  r.v0 = (int)(w0 & kPackMask); r.v1 = (int)((w0 >> 21) & kPackMask); r.v2 = (int)(w0 >> 42);
  int n0 = (int)(w1 & kPackMask), n1 = (int)((w1 >> 21) & kPackMask), n2 = (int)(w1 >> 42);
  r.n0 = n0 == (int)kPackMask ? -1 : n0;
  r.n1 = n1 == (int)kPackMask ? -1 : n1;
  r.n2 = n2 == (int)kPackMask ? -1 : n2;
  return r;
}
__device__ __forceinline__ double4 load_vert(const Mesh2View &m, int v) {
  double4 r;
  asm volatile("ld.global.nc" DLI_L2_64B ".v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(m.vrec + DLI_IDX(v, m.nv)));
  return r;
}

// ---------------------------------------------------------------------------------------------
// 64-byte self-contained walk record of a 2D face (fast pass k_interp2_fat).  One walk step = ONE
// 128-byte line instead of a face record plus three vertex records:
//   half 0 (32 B, geometry): p0 = vertex 0 in fp64; a = fp32(v1 - v0), b = fp32(v2 - v0)
//   half 1 (32 B, values + topology): f0, grad f (the face's affine interpolant
//          y = f0 + g . (p - p0)); three 21-bit neighbour ids (kPackMask = hull) and, in bit 63, the
//          "slow" flag (this face must be handled by the exact / reference-order pass)
// The orientation tests of the fast pass run on the fp32-rounded edge vectors with an error bound that
// covers the rounding (kFatK below), so a certified sign is the sign of the exact predicate on the
// fp64 mesh; everything uncertain goes to k_interp2_full.  The gradient form is used only on faces
// whose a-priori error bound (k_build_fat2) keeps it within 1e-13 of the exact interpolant relative
// to the smallest |f| on the face; every other face carries the slow flag.
// ---------------------------------------------------------------------------------------------
constexpr unsigned long long kFatSlow = 1ull << 63;
constexpr unsigned kFatFar = 0x1FFFFEu;  // DELTA encoding: neighbour index difference does not fit
constexpr int kFatBias = 1 << 20;        // DELTA encoding: field = neighbour - face + kFatBias
constexpr double kFatK = 1.1920928955078125e-07;  // 2^-23: twice the fp32 unit roundoff 2^-24

#define DLI_LDG256 "ld.global.nc" DLI_L2_64B ".v8.b32"
__device__ __forceinline__ void ldg256(unsigned (&r)[8], const void *p) {
  asm volatile(DLI_LDG256 " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "l"(p));
}

// Fast path: static filter only.  Returns true with the containing face's vertices loaded when every
// decision was certain, the point stayed inside the hull and the walk was short; false = hand the
// query to the full path.
__device__ __forceinline__ bool locate2_fast(const Mesh2View &m, double px, double py, int &f, double4 &A,
                                             double4 &B, double4 &C, unsigned int &steps) {
  int from = -1;
  const double fc = filter_const(2);
  for (int it = 0; it < m.fast_steps; it++) {
    steps++;
    Face2 F = load_face(m, f);
    A = load_vert(m, F.v0);
    B = load_vert(m, F.v1);
    C = load_vert(m, F.v2);
    double ax = A.x - px, ay = A.y - py, bx = B.x - px, by = B.y - py, cx = C.x - px, cy = C.y - py;
    // edge opposite v0 = (B, C); opposite v1 = (C, A); opposite v2 = (A, B); orient(X, Y, p) with
    // differences taken against p
    double l0 = bx * cy, r0 = by * cx, d0 = l0 - r0, e0 = fc * (fabs(l0) + fabs(r0));
    double l1 = cx * ay, r1 = cy * ax, d1 = l1 - r1, e1 = fc * (fabs(l1) + fabs(r1));
    double l2 = ax * by, r2 = ay * bx, d2 = l2 - r2, e2 = fc * (fabs(l2) + fabs(r2));
    bool neg0 = (-d0 > e0) && !(F.n0 >= 0 && F.n0 == from);
    bool neg1 = (-d1 > e1) && !(F.n1 >= 0 && F.n1 == from);
    bool neg2 = (-d2 > e2) && !(F.n2 >= 0 && F.n2 == from);
    if (neg0 | neg1 | neg2) {
      int nb = neg0 ? F.n0 : (neg1 ? F.n1 : F.n2);
      if (nb < 0) return false;  // leaving the hull: full path
      from = f;
      f = nb;
      continue;
    }
    // no certainly-negative edge: inside iff the others are certainly positive (the edge we came
    // through is known positive)
    bool pos0 = (d0 > e0) || (F.n0 >= 0 && F.n0 == from);
    bool pos1 = (d1 > e1) || (F.n1 >= 0 && F.n1 == from);
    bool pos2 = (d2 > e2) || (F.n2 >= 0 && F.n2 == from);
    return pos0 & pos1 & pos2;
  }
  return false;
}

// Among the hull edges strictly visible from p (a contiguous chain containing h0) the one whose supporting line is
// farthest from p (largest det^2 / |PQ|^2 in plain fp64, ties: smallest index); canonical stand-in for "whichever
// infinite face CGAL's walk ended in".  Farthest-line-first guarantees progress: after a projection p sits on the
// line of its edge within rounding and that edge may still test strictly visible by an ulp; a fixed preference for
// it would re-project p on the same line until the round cap.
__device__ inline int canonical_visible_hull_edge(const Mesh2View &m, double px, double py, int h0,
                                                  StatsAcc &st) {
  int best = h0;
  double best_score;
  {
    int4 r = __ldg(m.hull + DLI_IDX(h0, m.nh));
    Face2 F = load_face(m, r.x);
    int vs[3] = {F.v0, F.v1, F.v2};
    double4 P = load_vert(m, vs[(r.y + 1) % 3]), Q = load_vert(m, vs[(r.y + 2) % 3]);
    best_score = hull_edge_score(P.x, P.y, Q.x, Q.y, px, py);
  }
  for (int dir = 0; dir < 2; dir++) {
    int h = h0;
    for (int it = 0; it < m.nh; it++) {
      int4 hr = __ldg(m.hull + DLI_IDX(h, m.nh));
      h = dir ? hr.w : hr.z;
      if (h < 0 || h == h0) break;
      int4 r = __ldg(m.hull + DLI_IDX(h, m.nh));
      Face2 F = load_face(m, r.x);
      int vs[3] = {F.v0, F.v1, F.v2};
      double4 P = load_vert(m, vs[(r.y + 1) % 3]), Q = load_vert(m, vs[(r.y + 2) % 3]);
      int s = orient2d_filtered(P.x, P.y, Q.x, Q.y, px, py);
      if (s == 2) { st.exact++; s = orient2d_exact(P.x, P.y, Q.x, Q.y, px, py); }
      if (s >= 0) break;
      const double sc = hull_edge_score(P.x, P.y, Q.x, Q.y, px, py);
      if (sc > best_score || (sc == best_score && h < best)) { best = h; best_score = sc; }
    }
  }
  return best;
}

// Full path: exact-sign walk over closed faces, outside-hull projection loop, status codes.
// Returns the face in f with its vertices loaded; p is updated to the projected point.
// PROJECT = false (gradient queries, src/delaunay_2_interp.h:618): leaving the hull ends the query with
// DLI_STATUS_OUTSIDE_NAN instead of projecting.
template <bool PROJECT = true>
__device__ inline uint8_t locate2_full(const Mesh2View &m, double &px, double &py, int &f, double4 &A, double4 &B,
                                       double4 &C, StatsAcc &st) {
  uint8_t status = DLI_STATUS_OK;
  int from = -1;
  unsigned int steps = 0;
  // A projected point that still tests strictly outside its own hull edge (by an ulp: the projection is an inexact
  // construction) is projected on the same line again and usually lands on itself: from then on every round repeats
  // the previous one exactly (same point, same edge, same start face), so the remaining rounds up to the cap are
  // skipped -- the result is the one the capped loop returns.
  bool stuck = false;
  int prev_h = -1;
  for (int proj = 0;; proj++) {
    int exit_h = -1;
    bool found = false;
    for (;;) {
      if ((int)steps >= m.max_steps) { st.steps += steps; return DLI_STATUS_STEP_LIMIT; }
      steps++;
      Face2 F = load_face(m, f);
      A = load_vert(m, F.v0);
      B = load_vert(m, F.v1);
      C = load_vert(m, F.v2);
      int nbs[3] = {F.n0, F.n1, F.n2};
      // The decision "first strictly negative edge in the order 0, 1, 2", taken in two convergent stages (as
      // locateD_full does): the fp64 filter on all three edges, branch-free; then the exact predicate on the uncertain
      // edges below the first certainly-negative one, lowest first, through ONE call site -- lanes resolve their k-th
      // uncertain edge together whatever its index (with the predicate inlined per edge, a warp of on-edge queries
      // ran 13 lanes wide: profiles/r02_k_interp2_full_lattice_before_ncu.txt).
      int mv = 3;
      unsigned unc = 0;
#pragma unroll
      for (int j = 0; j < 3; j++) {
        if (nbs[j] >= 0 && nbs[j] == from) continue;
        const double4 &X = (j == 0) ? B : (j == 1 ? C : A);
        const double4 &Y = (j == 0) ? C : (j == 1 ? A : B);
        const int s = orient2d_filtered(X.x, X.y, Y.x, Y.y, px, py);
        if (s == 2) unc |= 1u << j;
        else if (s < 0 && j < mv) mv = j;
      }
      unc &= (1u << mv) - 1u;
      while (unc) {
        const int j = __ffs((int)unc) - 1;
        unc &= unc - 1u;
        const double Xx = (j == 0) ? B.x : (j == 1 ? C.x : A.x), Xy = (j == 0) ? B.y : (j == 1 ? C.y : A.y);
        const double Yx = (j == 0) ? C.x : (j == 1 ? A.x : B.x), Yy = (j == 0) ? C.y : (j == 1 ? A.y : B.y);
        st.exact++;
        if (orient2d_exact(Xx, Xy, Yx, Yy, px, py) < 0) { mv = j; unc = 0; }
      }
      if (mv > 2) { found = true; break; }
      int nb = nbs[mv];
      if (nb < 0) { exit_h = ~nb; break; }
      from = f;
      f = nb;
    }
    if (found) break;
    if (status == DLI_STATUS_OK) st.outside++;
    if (!PROJECT) {
      st.steps += steps;
      if (steps > st.max_steps) st.max_steps = steps;
      return DLI_STATUS_OUTSIDE_NAN;
    }
    status = DLI_STATUS_OUTSIDE_PROJECTED;
    if (proj >= kMaxProjections || stuck) break;  // accept the hull face (ulp-level extrapolation)
    int h = canonical_visible_hull_edge(m, px, py, exit_h, st);
    int4 r = __ldg(m.hull + DLI_IDX(h, m.nh));
    Face2 F = load_face(m, r.x);
    int vs[3] = {F.v0, F.v1, F.v2};
    double4 P = load_vert(m, vs[(r.y + 1) % 3]), Q = load_vert(m, vs[(r.y + 2) % 3]);
    const double opx = px, opy = py;
    project_on_line(P.x, P.y, Q.x, Q.y, px, py);
    stuck = (h == prev_h && px == opx && py == opy);
    prev_h = h;
    f = r.x;
    from = -1;
  }
  st.steps += steps;
  if (steps > st.max_steps) st.max_steps = steps;
  return status;
}

// ---------------------------------------------------------------------------------------------
// 2D gradient queries (Delaunay_incremental_interp_2::gradient, src/delaunay_2_interp.h:608-642)
// ---------------------------------------------------------------------------------------------

// Gradient of the linear interpolant on face (A, B, C): (V^-1)^T df with the columns of V = B - A, C - A and
// df = (f_B - f_A, f_C - f_A) (simplex_gradient :242-263 via set_simplex_gradient :574-588), written out
// through the adjugate.  Same operations, in the same order, as the gradient stored in the walk records
// (k_build_fat2), so the fast pass and the exact pass return the same bits.
__device__ __forceinline__ double2 face_gradient2(const double4 &A, const double4 &B, const double4 &C) {
  const double Ax = B.x - A.x, Ay = B.y - A.y, Bx = C.x - A.x, By = C.y - A.y;
  const double d1 = B.z - A.z, d2 = C.z - A.z;
  const double dl = Ax * By, dr = Ay * Bx, det = dl - dr;
  const double xl = d1 * By, xr = d2 * Ay, yl = Ax * d2, yr = Bx * d1;
  return make_double2((xl - xr) / det, (yl - yr) / det);
}
__device__ inline double2 face_gradient2_of(const Mesh2View &m, int f) {
  Face2 F = load_face(m, f);
  return face_gradient2(load_vert(m, F.v0), load_vert(m, F.v1), load_vert(m, F.v2));
}
__device__ __forceinline__ int face_slot_of(const Face2 &F, int v) { return F.v0 == v ? 0 : (F.v1 == v ? 1 : 2); }
__device__ __forceinline__ int face_nb(const Face2 &F, int j) { return j == 0 ? F.n0 : (j == 1 ? F.n1 : F.n2); }

// Average of the gradients of the finite faces around vertex v, starting the circulation in face f0
// (:619-632 + get_avg_simplex_gradient :594-606 + get_centroid :108-121: sequential sum, then / count).
// CGAL's circulator starts at an arbitrary incident face, so the summation order is canonicalised: a closed
// fan is summed counter-clockwise from its face with the smallest caller id, an open fan (hull vertex) from
// its clockwise-most face (the convention is part of the ABI contract, include/dli.h).
__device__ inline double2 vertex_avg_gradient2(const Mesh2View &m, int f0, int v) {
  // clockwise sweep: find the start face
  int start = f0, best_id = __ldg(m.orig + DLI_IDX(f0, m.ns)), f = f0;
  bool open = false;
  for (int it = 0; it < m.max_steps; it++) {
    Face2 F = load_face(m, f);
    int nb = face_nb(F, (face_slot_of(F, v) + 2) % 3);  // across the edge (v, next vertex): clockwise neighbour
    if (nb < 0) { open = true; start = f; break; }
    if (nb == f0) break;
    f = nb;
    int id = __ldg(m.orig + DLI_IDX(f, m.ns));
    if (id < best_id) { best_id = id; start = f; }
  }
  // counter-clockwise sum from `start`
  double sx = 0.0, sy = 0.0;
  int cnt = 0;
  f = start;
  for (int it = 0; it < m.max_steps; it++) {
    double2 g = face_gradient2_of(m, f);
    sx += g.x; sy += g.y; cnt++;
    Face2 F = load_face(m, f);
    int nb = face_nb(F, (face_slot_of(F, v) + 1) % 3);  // across the edge (v, previous vertex): counter-clockwise
    if (nb < 0 || (!open && nb == start)) break;
    f = nb;
  }
  return make_double2(sx / cnt, sy / cnt);
}

// One gradient query on the exact path.  Status: DLI_STATUS_OK (strictly inside a face), DLI_STATUS_ON_EDGE,
// DLI_STATUS_ON_VERTEX, DLI_STATUS_OUTSIDE_NAN (no projection for gradients: NaN, :610-611), STEP_LIMIT.
__device__ inline uint8_t gradient2_full(const Mesh2View &m, double px, double py, int &f, double2 &g, StatsAcc &st) {
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  g = make_double2(nan, nan);
  double4 A, B, C;
  uint8_t status = locate2_full<false>(m, px, py, f, A, B, C, st);
  if (status != DLI_STATUS_OK) return status;
  Face2 F = load_face(m, f);
  int zero[3], nz = 0;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const double4 &X = (j == 0) ? B : (j == 1 ? C : A);
    const double4 &Y = (j == 0) ? C : (j == 1 ? A : B);
    int s = orient2d_filtered(X.x, X.y, Y.x, Y.y, px, py);
    if (s == 2) { st.exact++; s = orient2d_exact(X.x, X.y, Y.x, Y.y, px, py); }
    zero[j] = (s == 0);
    nz += zero[j];
  }
  if (nz == 0) { g = face_gradient2(A, B, C); return DLI_STATUS_OK; }
  if (nz == 1) {  // EDGE: this face and the one across the edge (:633-636); an infinite neighbour is skipped
    int li = zero[0] ? 0 : (zero[1] ? 1 : 2);
    double2 g1 = face_gradient2(A, B, C);
    int nb = face_nb(F, li);
    if (nb >= 0) {
      double2 g2 = face_gradient2_of(m, nb);
      g = make_double2((g1.x + g2.x) / 2, (g1.y + g2.y) / 2);
    } else {
      g = g1;
    }
    return DLI_STATUS_ON_EDGE;
  }
  // VERTEX: on two edges = the vertex opposite the third
  int k = !zero[0] ? 0 : (!zero[1] ? 1 : 2);
  int v = k == 0 ? F.v0 : (k == 1 ? F.v1 : F.v2);
  g = vertex_avg_gradient2(m, f, v);
  return DLI_STATUS_ON_VERTEX;
}

// ---------------------------------------------------------------------------------------------
// dD
// ---------------------------------------------------------------------------------------------

template <int D>
__device__ __forceinline__ int hintD(const MeshDView &m, const double *p) {
  size_t idx = 0;
#pragma unroll
  for (int a = D - 1; a >= 0; a--) {
    int i = (int)((p[a] - m.o[a]) * m.ic[a]);
    i = max(0, min(m.g[a] - 1, i));
    idx = idx * m.g[a] + i;
  }
  return ldg_hint(m.hint + DLI_IDX(idx, m.cells));
}

// first record of a query's walk: the hinted simplex, or (start copies) the copy of its record that lives in the hint cell
template <int D>
__device__ __forceinline__ int startD(const MeshDView &m, const double *p) {
  size_t idx = 0;
#pragma unroll
  for (int a = D - 1; a >= 0; a--) {
    int i = (int)((p[a] - m.o[a]) * m.ic[a]);
    i = max(0, min(m.g[a] - 1, i));
    idx = idx * m.g[a] + i;
  }
  if (m.start_copies) return m.ns + (int)idx;
  return ldg_hint(m.hint + DLI_IDX(idx, m.cells));
}
// simplex of a record index
__device__ __forceinline__ int record_simplex(const MeshDView &m, int r) {
  return r >= m.ns ? ldg_hint(m.hint + DLI_IDX(r - m.ns, m.cells)) : r;
}

template <int D>
struct SimplexD {
  int v[D + 1], n[D + 1];
  double x[(D + 1) * D];
  double f[D + 1];
};

template <int D>
__device__ __forceinline__ void load_simplex_ids(const MeshDView &m, int s, SimplexD<D> &S) {
  constexpr int SS = RecD<D>::SS;
  const int4 *r = reinterpret_cast<const int4 *>(m.srec + (size_t)DLI_IDX(s, m.ns) * SS);
  int tmp[SS];
#pragma unroll
  for (int i = 0; i < SS / 4; i++) {
    if (i * 4 < 2 * (D + 1)) {
      int4 t = __ldg(r + i);
      tmp[4 * i] = t.x; tmp[4 * i + 1] = t.y; tmp[4 * i + 2] = t.z; tmp[4 * i + 3] = t.w;
    }
  }
#pragma unroll
  for (int i = 0; i <= D; i++) { S.v[i] = tmp[i]; S.n[i] = tmp[D + 1 + i]; }
}
template <int D>
__device__ __forceinline__ void load_simplex_verts(const MeshDView &m, SimplexD<D> &S) {
  constexpr int VS = RecD<D>::VS;
#pragma unroll
  for (int i = 0; i <= D; i++) {
    const double2 *r = reinterpret_cast<const double2 *>(m.vrec + (size_t)DLI_IDX(S.v[i], m.nv) * VS);
    double tmp[VS];
#pragma unroll
    for (int k = 0; k < VS / 2; k++) {
      if (2 * k < D + 1) {
        double2 t = __ldg(r + k);
        tmp[2 * k] = t.x; tmp[2 * k + 1] = t.y;
      }
    }
#pragma unroll
    for (int k = 0; k < D; k++) S.x[i * D + k] = tmp[k];
    S.f[i] = tmp[D];
  }
}

// sign of the orientation of the simplex with vertex j replaced by p; EXACT selects the fallback
template <int D, bool EXACT>
__device__ __forceinline__ int orient_replaced(const SimplexD<D> &S, int j, const double *p, StatsAcc &st) {
  double pts[(D + 1) * D];
#pragma unroll
  for (int i = 0; i <= D; i++)
#pragma unroll
    for (int k = 0; k < D; k++) pts[i * D + k] = (i == j) ? p[k] : S.x[i * D + k];
  int s = orient_filtered<D>(pts);
  if (EXACT && s == 2) { st.exact++; s = orient_exact<D>(pts); }
  return s;
}

// All D+1 facet orientations of one walk step at once.  With R_i = v_i - q (rows of a (D+1) x D matrix),
// the orientation of the simplex with vertex j replaced by q is (-1)^(j+D) * det(R without row j); the D+1
// determinants share their lower-order minors (cofactor expansion along column 0, minors along their first
// column), so the rounding structure -- and therefore the filter constant -- is that of DetPerm<D>.
// Outputs o[j] (signed so that o[j] < 0 means "q is beyond facet j") and the permanent bound scale pm[j].
template <int D>
struct OrientAll;
template <>
struct OrientAll<3> {
  __device__ __forceinline__ static void eval(const double *R, double *o, double *pm) {
    // 2x2 minors over columns (1,2) for the 6 row pairs
    double M[4][4], P[4][4];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = a + 1; b < 4; b++) {
        double l = R[a * 3 + 1] * R[b * 3 + 2], r = R[a * 3 + 2] * R[b * 3 + 1];
        M[a][b] = l - r;
        P[a][b] = fabs(l) + fabs(r);
      }
    auto det3 = [&](int a, int b, int c, double &perm) {
      double t0 = R[a * 3] * M[b][c], t1 = R[b * 3] * M[a][c], t2 = R[c * 3] * M[a][b];
      perm = fabs(R[a * 3]) * P[b][c] + fabs(R[b * 3]) * P[a][c] + fabs(R[c * 3]) * P[a][b];
      return (t0 - t1) + t2;
    };
    o[0] = -det3(1, 2, 3, pm[0]);
    o[1] = det3(0, 2, 3, pm[1]);
    o[2] = -det3(0, 1, 3, pm[2]);
    o[3] = det3(0, 1, 2, pm[3]);
  }
};
template <>
struct OrientAll<4> {
  __device__ __forceinline__ static void eval(const double *R, double *o, double *pm) {
    double M2[5][5], P2[5][5];
#pragma unroll
    for (int a = 0; a < 5; a++)
#pragma unroll
      for (int b = a + 1; b < 5; b++) {
        double l = R[a * 4 + 2] * R[b * 4 + 3], r = R[a * 4 + 3] * R[b * 4 + 2];
        M2[a][b] = l - r;
        P2[a][b] = fabs(l) + fabs(r);
      }
    double M3[5][5][5], P3[5][5][5];
#pragma unroll
    for (int a = 0; a < 5; a++)
#pragma unroll
      for (int b = a + 1; b < 5; b++)
#pragma unroll
        for (int c = b + 1; c < 5; c++) {
          double t0 = R[a * 4 + 1] * M2[b][c], t1 = R[b * 4 + 1] * M2[a][c], t2 = R[c * 4 + 1] * M2[a][b];
          M3[a][b][c] = (t0 - t1) + t2;
          P3[a][b][c] = fabs(R[a * 4 + 1]) * P2[b][c] + fabs(R[b * 4 + 1]) * P2[a][c] + fabs(R[c * 4 + 1]) * P2[a][b];
        }
    auto det4 = [&](int a, int b, int c, int d, double &perm) {
      double t0 = R[a * 4] * M3[b][c][d], t1 = R[b * 4] * M3[a][c][d], t2 = R[c * 4] * M3[a][b][d], t3 = R[d * 4] * M3[a][b][c];
      perm = fabs(R[a * 4]) * P3[b][c][d] + fabs(R[b * 4]) * P3[a][c][d] + fabs(R[c * 4]) * P3[a][b][d] +
             fabs(R[d * 4]) * P3[a][b][c];
      return ((t0 - t1) + t2) - t3;
    };
    o[0] = det4(1, 2, 3, 4, pm[0]);
    o[1] = -det4(0, 2, 3, 4, pm[1]);
    o[2] = det4(0, 1, 3, 4, pm[2]);
    o[3] = -det4(0, 1, 2, 4, pm[3]);
    o[4] = det4(0, 1, 2, 3, pm[4]);
  }
};

// fast path (filter only).  true = located with certainty inside the hull.
template <int D>
__device__ __forceinline__ bool locateD_fast(const MeshDView &m, const double *p, int &s, SimplexD<D> &S,
                                             unsigned int &steps) {
  int from = -1;
  StatsAcc dummy;
  for (int it = 0; it < m.fast_steps; it++) {
    steps++;
    load_simplex_ids<D>(m, s, S);
    load_simplex_verts<D>(m, S);
    int mv = -1;
    bool uncertain = false;
#pragma unroll
    for (int j = 0; j <= D; j++) {
      if (mv >= 0) continue;
      if (S.n[j] >= 0 && S.n[j] == from) continue;
      int sg = orient_replaced<D, false>(S, j, p, dummy);
      if (sg == -1) mv = j;
      else if (sg != 1) uncertain = true;
    }
    if (mv < 0) return !uncertain;
    int nb = S.n[mv];
    if (nb < 0) return false;
    from = s;
    s = nb;
  }
  return false;
}

// full path: exact signs; outside the hull -> DLI_STATUS_OUTSIDE_NAN (src/delaunay_d_interp.h:1005-1008)
template <int D>
__device__ inline uint8_t locateD_full(const MeshDView &m, const double *p, int &s, SimplexD<D> &S, StatsAcc &st) {
  int from = -1;
  unsigned int steps = 0;
  for (;;) {
    if ((int)steps >= m.max_steps) { st.steps += steps; return DLI_STATUS_STEP_LIMIT; }
    steps++;
    load_simplex_ids<D>(m, s, S);
    load_simplex_verts<D>(m, S);
    // First the filter on every facet (cheap, the whole warp together), then the exact predicate only on the
    // uncertain facets below the first certainly-negative one, lowest first: the same decision as testing the facets
    // in order and leaving at the first negative one, but lanes resolve their k-th uncertain facet together whatever
    // its index, instead of serialising by facet index (exact evaluation is ~10^3 instructions; on batches with
    // many exact zeros the warp ran 3 lanes wide).
    int mv = D + 1;
    unsigned unc = 0;
    unsigned skip = 0;
#pragma unroll
    for (int j = 0; j <= D; j++) if (S.n[j] >= 0 && S.n[j] == from) skip |= 1u << j;
#pragma unroll 1
    for (int j = 0; j <= D; j++) {
      if (skip >> j & 1u) continue;
      const int sg = orient_replaced<D, false>(S, j, p, st);
      if (sg == 2) unc |= 1u << j;
      else if (sg < 0 && j < mv) mv = j;
    }
    unc &= (1u << mv) - 1u;
    while (unc) {
      const int j = __ffs((int)unc) - 1;
      unc &= unc - 1u;
      double pts[(D + 1) * D];
#pragma unroll
      for (int i = 0; i <= D; i++)
#pragma unroll
        for (int k = 0; k < D; k++) pts[i * D + k] = (i == j) ? p[k] : S.x[i * D + k];
      st.exact++;
      if (orient_exact<D>(pts) < 0) { mv = j; unc = 0; }
    }
    if (mv > D) break;
    int nb = S.n[mv];
    if (nb < 0) {
      st.steps += steps; st.outside++;
      if (steps > st.max_steps) st.max_steps = steps;
      return DLI_STATUS_OUTSIDE_NAN;
    }
    from = s;
    s = nb;
  }
  st.steps += steps;
  if (steps > st.max_steps) st.max_steps = steps;
  return DLI_STATUS_OK;
}

// (lu_det, volD, dot_exact_trunc, epilogueD_ref_xf: dli_refmath.h)

template <int D, bool COMPACT = false>
__device__ __forceinline__ double epilogueD_ref(const double *p, const SimplexD<D> &S) {
  return epilogueD_ref_xf<D, COMPACT>(p, S.x, S.f);
}

// closed form (option epilogue = 1): one LU solve of [P1-P0 .. PD-P0] lam = p - P0
template <int D>
__device__ __forceinline__ double epilogueD_fast(const double *p, const SimplexD<D> &S) {
  double a[D * (D + 1)];
#pragma unroll
  for (int i = 0; i < D; i++) {
#pragma unroll
    for (int j = 0; j < D; j++) a[i * (D + 1) + j] = S.x[(j + 1) * D + i] - S.x[i];
    a[i * (D + 1) + D] = p[i] - S.x[i];
  }
#pragma unroll
  for (int c = 0; c < D; c++) {
    int piv = c;
    double best = fabs(a[c * (D + 1) + c]);
#pragma unroll
    for (int r = c + 1; r < D; r++) {
      double v = fabs(a[r * (D + 1) + c]);
      if (v > best) { best = v; piv = r; }
    }
#pragma unroll
    for (int r = c + 1; r < D; r++)
      if (r == piv) {
#pragma unroll
        for (int j = 0; j <= D; j++) { double t = a[c * (D + 1) + j]; a[c * (D + 1) + j] = a[r * (D + 1) + j]; a[r * (D + 1) + j] = t; }
      }
    double inv = 1.0 / a[c * (D + 1) + c];
#pragma unroll
    for (int r = c + 1; r < D; r++) {
      double l = a[r * (D + 1) + c] * inv;
#pragma unroll
      for (int j = c + 1; j <= D; j++) a[r * (D + 1) + j] -= l * a[c * (D + 1) + j];
    }
  }
  double lam[D];
#pragma unroll
  for (int i = D - 1; i >= 0; i--) {
    double s = a[i * (D + 1) + D];
#pragma unroll
    for (int j = i + 1; j < D; j++) s -= a[i * (D + 1) + j] * lam[j];
    lam[i] = s / a[i * (D + 1) + i];
  }
  double l0 = 1.0, y = 0.0;
#pragma unroll
  for (int i = 0; i < D; i++) { l0 -= lam[i]; y += lam[i] * S.f[i + 1]; }
  return y + l0 * S.f[0];
}

}  // namespace dli
