// dli_capi.cu -- the C ABI (include/dli.h): flattener, hint-grid build, batched interp / locate
// dispatch (device-pointer and host-pointer paths), error evaluator, replication support.
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <cub/device/device_scan.cuh>

#include "dli_kernels.cuh"

using namespace dli;

// ---------------------------------------------------------------------------------------------
// error handling
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}
#define CU(call)                                                                                      \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess) {                                                                          \
      char b_[512];                                                                                   \
      snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return fail(DLI_ERR_CUDA, b_);                                                                  \
    }                                                                                                 \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (dev >= 0 && dev != prev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// ---------------------------------------------------------------------------------------------
// mesh object
// ---------------------------------------------------------------------------------------------
constexpr int64_t kHostChunk = 1ll << 22;  // queries per staged chunk of the host-pointer path (64 MB of 2D queries)
struct HostPipe {
  static constexpr int kSlots = 3;
  bool ready = false;
  cudaStream_t st[kSlots] = {nullptr, nullptr, nullptr};
  cudaEvent_t done[kSlots] = {nullptr, nullptr, nullptr};
  double *x[kSlots] = {nullptr, nullptr, nullptr}, *y[kSlots] = {nullptr, nullptr, nullptr};
  int32_t *s[kSlots] = {nullptr, nullptr, nullptr};
  uint8_t *t[kSlots] = {nullptr, nullptr, nullptr};
};
// Pageable host buffers (numpy arrays, std::vector): a cudaMemcpyAsync from pageable memory is staged by the driver
// through one thread at 6-10 GB/s.  Buffers that are not page-locked therefore go through slots with their own pinned
// staging memory, one worker thread per slot: every worker copies a chunk in, runs it on its stream, copies the
// results out, so the host copies of several chunks proceed in parallel and overlap the DMA of the others.
constexpr int kMaxStage = 16;
constexpr int64_t kStageChunk = 1ll << 19;        // queries per staged chunk at most (8 MB of 2D queries)
constexpr int64_t kStageMinChunk = 1ll << 16;     // ... and at least (medium batches are cut so that every worker gets two)
constexpr int64_t kStageMinQueries = 1ll << 18;   // smaller batches: the driver's own pageable path is as fast
struct StageSlot {
  cudaStream_t st = nullptr;
  double *dx = nullptr, *dy = nullptr, *hx = nullptr, *hy = nullptr;
  int32_t *ds = nullptr, *hs = nullptr;
  uint8_t *dt = nullptr, *ht = nullptr;
};
struct StagePipe {
  int nslots = 0;
  StageSlot slot[kMaxStage];
};
static void stage_free(StagePipe &P) {
  for (int i = 0; i < kMaxStage; i++) {
    StageSlot &S = P.slot[i];
    if (S.st) { cudaStreamSynchronize(S.st); cudaStreamDestroy(S.st); }
    cudaFree(S.dx); cudaFree(S.dy); cudaFree(S.ds); cudaFree(S.dt);
    cudaFreeHost(S.hx); cudaFreeHost(S.hy); cudaFreeHost(S.hs); cudaFreeHost(S.ht);
    S = StageSlot();
  }
  P.nslots = 0;
}
static void pipe_free(HostPipe &P) {
  for (int i = 0; i < HostPipe::kSlots; i++) {
    if (P.st[i]) { cudaStreamSynchronize(P.st[i]); cudaStreamDestroy(P.st[i]); }
    if (P.done[i]) cudaEventDestroy(P.done[i]);
    cudaFree(P.x[i]); cudaFree(P.y[i]); cudaFree(P.s[i]); cudaFree(P.t[i]);
    P.st[i] = nullptr; P.done[i] = nullptr; P.x[i] = P.y[i] = nullptr; P.s[i] = nullptr; P.t[i] = nullptr;
  }
  P.ready = false;
}

struct dli_mesh {
  int dim = 0;
  int device = 0;
  int64_t nv = 0, ns = 0;
  int32_t nh = 0;
  dli_options opt{};
  int vs = 0, ss = 0;  // record strides (doubles / ints)
  // host copies of the canonical connectivity (error items, get_simplices)
  std::vector<int32_t> h_sv, h_sn;
  // dli_mesh_update (2D meshes created here, not replicas): caller id -> device index, vertex coordinates,
  // capacities of the growable device buffers
  std::vector<int32_t> h_inv;
  std::vector<double> h_xy;
  int64_t cap_v = 0, cap_s = 0;
  int64_t cap_fat = 0;               // dD: walk records d_fat has room for (>= cap_s; the start copies leave room behind ns)
  int32_t cap_h = 0;
  int64_t n_updates = 0;
  unsigned char *d_upd = nullptr;   // grow-only device scratch of dli_mesh_update
  size_t upd_cap = 0;
  // device buffers
  double *d_vrec = nullptr;
  int32_t *d_srec = nullptr;
  int4 *d_hull = nullptr;
  int32_t *d_hint = nullptr;
  int32_t *d_orig = nullptr;      // device simplex index -> caller's simplex id (simplices are stored in Morton order)
  ulonglong2 *d_fpack = nullptr;  // 2D packed face records (fast pass), NULL when ids do not fit 21 bits
  uint4 *d_fat = nullptr;         // 2D 64-byte walk records (k_interp2_fat), NULL when unused
  int64_t n_slow_faces = 0;       // faces carrying the slow flag in d_fat
  int fat_enc = -1;               // neighbour fields of the 2D walk records: 0 = absolute ids, 1 = delta-encoded; -1 = not
                                  // decided yet (creating rank: from the mesh size / options; replicas and snapshots: from
                                  // header word 13, never re-derived from their own options)
  // dD start copies (MeshDView::start_copies): d_fat has nslots = ns + cells records
  bool startcopy = false;
  // cell-indexed walk records (2D meshes far larger than L2): d_fat has nslots = cells + extra records
  bool cellmode = false;
  int64_t nslots = 0;             // records in d_fat (== ns unless cellmode)
  int32_t *d_slot_face = nullptr; // [nslots] slot -> face
  int *d_rowstart = nullptr;      // [gy + 1] first slot of every hint-grid row
  GridView grid{};
  int max_steps = 0, fast_steps = 0;
  // error items (lazy)
  std::mutex mu;
  bool items_ready = false;
  int64_t n_items = 0;
  std::vector<int32_t> h_item_simplex, h_item_verts;
  int32_t *d_item_verts = nullptr;
  // stats
  bool collect_stats = false;   // counters (atomics in the kernels)
  bool collect_timing = false;  // CUDA events around the interp kernels
  unsigned long long *d_stats = nullptr;
  float *h_straggler_frac = nullptr;   // mapped pinned word written by the exact passes: straggler fraction of the last launch
  // most recent call's launch / query count; written by concurrent callers of one mesh (dli.h allows them), so atomic:
  // the reader sees the numbers of SOME recent call, never a torn value
  mutable std::atomic<int64_t> last_launches{0}, last_queries{0};
  mutable std::mutex ev_mu;
  mutable std::vector<cudaEvent_t> ev;  // triples (before fast, after fast, after full)
  // host-pointer path staging ring (created on first use)
  mutable std::mutex pipe_mu;
  mutable HostPipe pipe;
  mutable StagePipe stage;
};

// 2D meshes with at least this many faces (512 MB of walk records, 4 x L2) use cell-indexed walk records
constexpr int64_t kCellModeFaces = 8ll << 20;

// bytes per walk record (2D: 64, 3D: 128, 4D: 192)
// 4D walk records: one packed 128-byte record per simplex (FatD<4, true>) unless the mesh has too many simplices for the
// 24-bit neighbour fields or opt.packed_faces == 5 asks for the split layout (tests); replicas / snapshots follow header word 13
static bool fat_pack4(const dli_mesh *m);
static int fat_record_bytes(const dli_mesh *m) {
  return m->dim == 2 ? 64 : (m->dim == 3 ? FatD<3>::BYTES : (fat_pack4(m) ? FatD<4, true>::BYTES : FatD<4>::BYTES));
}

static void mesh_free(dli_mesh *m) {
  if (!m) return;
  DeviceGuard g(m->device);
  cudaFree(m->d_vrec);
  cudaFree(m->d_srec);
  cudaFree(m->d_hull);
  cudaFree(m->d_hint);
  cudaFree(m->d_orig);
  cudaFree(m->d_fpack);
  cudaFree(m->d_fat);
  cudaFree(m->d_slot_face);
  cudaFree(m->d_rowstart);
  cudaFree(m->d_item_verts);
  cudaFree(m->d_upd);
  cudaFree(m->d_stats);
  if (m->h_straggler_frac) cudaFreeHost(m->h_straggler_frac);
  pipe_free(m->pipe);
  stage_free(m->stage);
  for (cudaEvent_t e : m->ev) cudaEventDestroy(e);
  delete m;
}

extern "C" void dli_options_init(dli_options *o) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->struct_size = (int32_t)sizeof(dli_options);
  o->device = -1;
  o->hint_cells_per_simplex = 0.f;
  o->max_walk_steps = 0;
  o->reserved0 = 0;
  o->epilogue = 0;
  o->threads_per_block = 0;
  o->queries_per_warp = 0;
  o->packed_faces = 0;
  o->prefetch_ring = 0;
  o->outside_nan = 0;
  o->start_records = 0;
}

extern "C" const char *dli_last_error(void) { return g_err.c_str(); }

// "no exception ever crosses the ABI" (include/dli.h): entry points that allocate host memory run inside this guard
#define DLI_GUARD_BEGIN try {
#define DLI_GUARD_END(who)                                                                                   \
  } catch (const std::bad_alloc &) { return fail(DLI_ERR_NOMEM, who ": out of host memory");                 \
  } catch (const std::exception &e) { return fail(DLI_ERR_ARG, std::string(who ": ") + e.what());            \
  } catch (...) { return fail(DLI_ERR_ARG, who ": unknown exception"); }

extern "C" int dli_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// ---------------------------------------------------------------------------------------------
// flattener (host part)
// ---------------------------------------------------------------------------------------------
namespace {

struct FacetRec {
  int32_t key[DLI_MAX_DIM];
  int32_t s, slot;
};

int derive_neighbors(int d, int64_t ns, const int32_t *sv, int32_t *sn) {
  const int k = d + 1;
  std::vector<FacetRec> fr((size_t)ns * k);
  for (int64_t s = 0; s < ns; s++)
    for (int slot = 0; slot < k; slot++) {
      FacetRec &r = fr[(size_t)s * k + slot];
      int c = 0;
      for (int j = 0; j < k; j++)
        if (j != slot) r.key[c++] = sv[s * k + j];
      std::sort(r.key, r.key + d);
      r.s = (int32_t)s;
      r.slot = slot;
    }
  auto less = [d](const FacetRec &a, const FacetRec &b) {
    for (int i = 0; i < d; i++)
      if (a.key[i] != b.key[i]) return a.key[i] < b.key[i];
    return false;
  };
  auto same = [d](const FacetRec &a, const FacetRec &b) {
    for (int i = 0; i < d; i++)
      if (a.key[i] != b.key[i]) return false;
    return true;
  };
  std::sort(fr.begin(), fr.end(), less);
  for (size_t i = 0; i < fr.size(); i++) sn[(size_t)fr[i].s * k + fr[i].slot] = -1;
  for (size_t i = 0; i + 1 < fr.size();) {
    if (same(fr[i], fr[i + 1])) {
      if (i + 2 < fr.size() && same(fr[i], fr[i + 2])) return -1;  // facet shared by 3 simplices
      sn[(size_t)fr[i].s * k + fr[i].slot] = fr[i + 1].s;
      sn[(size_t)fr[i + 1].s * k + fr[i + 1].slot] = fr[i].s;
      i += 2;
    } else {
      i += 1;
    }
  }
  return 0;
}

// Device order of the simplices: Morton (Z-curve) order of their centroids, so that simplices that are close in
// space are close in memory (walk-record locality for coherent query batches; short neighbour deltas).
// perm[i] = caller's id of the simplex stored at device index i.
void morton_order(int d, int64_t nv, const double *coords, int64_t ns, const int32_t *sv, std::vector<int32_t> &perm) {
  const int k = d + 1;
  double lo[DLI_MAX_DIM], hi[DLI_MAX_DIM];
  for (int c = 0; c < d; c++) { lo[c] = INFINITY; hi[c] = -INFINITY; }
  for (int64_t v = 0; v < nv; v++)
    for (int c = 0; c < d; c++) {
      double x = coords[v * d + c];
      if (x < lo[c]) lo[c] = x;
      if (x > hi[c]) hi[c] = x;
    }
  const int bits = 63 / d > 21 ? 21 : 63 / d;  // 2D/3D: 21 bits per axis, 4D: 15
  double scale[DLI_MAX_DIM];
  for (int c = 0; c < d; c++) {
    double ext = hi[c] - lo[c];
    scale[c] = (ext > 0.0 && std::isfinite(ext)) ? (double)((1u << bits) - 1) / ext : 0.0;
  }
  std::vector<std::pair<uint64_t, int32_t>> keys((size_t)ns);
  for (int64_t s = 0; s < ns; s++) {
    uint64_t key = 0;
    uint32_t q[DLI_MAX_DIM];
    for (int c = 0; c < d; c++) {
      double acc = 0.0;
      for (int j = 0; j < k; j++) acc += coords[(size_t)sv[s * k + j] * d + c];
      double t = (acc / k - lo[c]) * scale[c];
      q[c] = (t > 0.0) ? (uint32_t)std::min<double>(t, (double)((1u << bits) - 1)) : 0u;  // NaN -> 0
    }
    for (int b = bits - 1; b >= 0; b--)
      for (int c = d - 1; c >= 0; c--) key = (key << 1) | ((q[c] >> b) & 1u);
    keys[(size_t)s] = {key, (int32_t)s};
  }
  std::sort(keys.begin(), keys.end());
  perm.resize((size_t)ns);
  for (int64_t i = 0; i < ns; i++) perm[(size_t)i] = keys[(size_t)i].second;
}

template <int D>
void launch_hint_build(dli_mesh *m, cudaStream_t st) {
  int threads = 128;
  k_hint_raster<D><<<(unsigned)((m->ns + threads - 1) / threads), threads, 0, st>>>(
      m->grid, (int)m->ns, m->d_vrec, m->vs, m->d_srec, m->ss, m->d_hint);
}
template <int D>
void launch_hint_dilate(dli_mesh *m, unsigned int *d_rem, cudaStream_t st) {
  int threads = 256;
  k_hint_dilate<D><<<(unsigned)((m->grid.cells + threads - 1) / threads), threads, 0, st>>>(m->grid, m->d_hint, d_rem);
}

int build_hint_grid(dli_mesh *m, const double *coords) {
  const int d = m->dim;
  // bounding box of the vertices
  double lo[DLI_MAX_DIM], hi[DLI_MAX_DIM];
  for (int k = 0; k < d; k++) { lo[k] = INFINITY; hi[k] = -INFINITY; }
  for (int64_t v = 0; v < m->nv; v++)
    for (int k = 0; k < d; k++) {
      double x = coords[v * d + k];
      if (x < lo[k]) lo[k] = x;
      if (x > hi[k]) hi[k] = x;
    }
  float cps = m->opt.hint_cells_per_simplex;
  // measured optima; cell-indexed records (3.1b) have no hint table to keep small: 2 / 3 / 4 / 6 cells per
  // face give 30.0 / 31.4 / 32.3 / 33.3 Gq/s on cfg5 for 6.7 / 8.9 / 11.1 / 15.6 GB
  if (!(cps > 0.f)) cps = (d == 2) ? (m->cellmode ? 4.0f : 2.0f) : (d == 3 ? 4.0f : 8.0f);
  double target = std::max(1.0, (double)m->ns * cps);
  int g = (int)std::ceil(std::pow(target, 1.0 / d));
  g = std::max(1, std::min(g, d == 2 ? 16384 : (d == 3 ? 1024 : 128)));
  m->grid.cells = 1;
  for (int k = 0; k < d; k++) {
    double ext = hi[k] - lo[k];
    if (!(ext > 0.0) || !std::isfinite(ext)) ext = 1.0;
    m->grid.g[k] = g;
    m->grid.o[k] = lo[k];
    m->grid.cell[k] = ext / g;
    m->grid.ic[k] = g / ext;
    m->grid.cells *= g;
  }
  CU(cudaMalloc(&m->d_hint, sizeof(int32_t) * (size_t)m->grid.cells));
  CU(cudaMemset(m->d_hint, 0xff, sizeof(int32_t) * (size_t)m->grid.cells));
  unsigned int *d_rem = nullptr;
  CU(cudaMalloc(&d_rem, sizeof(unsigned int)));
  cudaStream_t st = 0;
  switch (d) {
    case 2: launch_hint_build<2>(m, st); break;
    case 3: launch_hint_build<3>(m, st); break;
    case 4: launch_hint_build<4>(m, st); break;
  }
  CU(cudaGetLastError());
  // dilation rounds until no empty cell is left (or no progress)
  unsigned int prev = 0xffffffffu;
  for (int round = 0; round < 4 * g + 8; round++) {
    CU(cudaMemsetAsync(d_rem, 0, sizeof(unsigned int), st));
    switch (d) {
      case 2: launch_hint_dilate<2>(m, d_rem, st); break;
      case 3: launch_hint_dilate<3>(m, d_rem, st); break;
      case 4: launch_hint_dilate<4>(m, d_rem, st); break;
    }
    unsigned int rem = 0;
    CU(cudaMemcpy(&rem, d_rem, sizeof rem, cudaMemcpyDeviceToHost));
    if (rem == 0) break;
    if (rem == prev && round > 2) {  // nothing rasterised at all (tiny mesh): start everywhere from simplex 0
      break;
    }
    prev = rem;
  }
  k_hint_fill_rest<<<(unsigned)((m->grid.cells + 255) / 256), 256, 0, st>>>(m->grid.cells, m->d_hint, 0);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(st));
  cudaFree(d_rem);
  return DLI_OK;
}

}  // namespace

static int mesh_alloc_common(dli_mesh *m) {
  const int d = m->dim;
  if (d == 2) { m->vs = 4; m->ss = 8; }
  else if (d == 3) { m->vs = RecD<3>::VS; m->ss = RecD<3>::SS; }
  else { m->vs = RecD<4>::VS; m->ss = RecD<4>::SS; }
  m->max_steps = m->opt.max_walk_steps > 0 ? m->opt.max_walk_steps : (int)std::min<int64_t>(m->ns + 16, 0x7fffffff);
  m->fast_steps = std::min(m->max_steps, d == 2 ? 24 : 32);
  CU(cudaMalloc(&m->d_vrec, sizeof(double) * (size_t)std::max<int64_t>(m->nv, 1) * m->vs));
  CU(cudaMalloc(&m->d_srec, sizeof(int32_t) * (size_t)std::max<int64_t>(m->ns, 1) * m->ss));
  CU(cudaMalloc(&m->d_orig, sizeof(int32_t) * (size_t)std::max<int64_t>(m->ns, 1)));
  if (d == 2) CU(cudaMalloc(&m->d_hull, sizeof(int4) * (size_t)std::max<int32_t>(m->nh, 1)));
  if (d == 2 && m->opt.packed_faces >= 0 && m->ns > 0 && m->nv <= (1 << 21) && m->ns < (1 << 21) - 1)
    CU(cudaMalloc(&m->d_fpack, sizeof(ulonglong2) * (size_t)m->ns));
  m->nslots = m->ns;
  m->cap_v = std::max<int64_t>(m->nv, 1); m->cap_s = std::max<int64_t>(m->ns, 1); m->cap_h = std::max<int32_t>(m->nh, 1);
  if (m->opt.epilogue == DLI_EPILOGUE_AUTO && m->opt.packed_faces >= 0 && m->ns > 0) {
    // 2D meshes whose records are far larger than L2 get cell-indexed records (allocated once the hint grid exists)
    m->cellmode = d == 2 && (m->opt.packed_faces == 3 || (m->opt.packed_faces == 0 && m->ns >= kCellModeFaces));
    if (!m->cellmode) CU(cudaMalloc(&m->d_fat, (size_t)fat_record_bytes(m) * (size_t)m->ns));
  }
  CU(cudaMalloc(&m->d_stats, sizeof(unsigned long long) * ST_WORDS));
  CU(cudaMemset(m->d_stats, 0, sizeof(unsigned long long) * ST_WORDS));
  // (a failed allocation only loses the hint: the queue then keeps its default size)
  if (cudaHostAlloc(&m->h_straggler_frac, sizeof(float), cudaHostAllocMapped) == cudaSuccess) *m->h_straggler_frac = 0.f;
  else { cudaGetLastError(); m->h_straggler_frac = nullptr; }
  // keep stream-ordered allocations cached between calls
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, m->device) == cudaSuccess) {
    unsigned long long thr = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
  }
  return DLI_OK;
}

static bool fat_pack4(const dli_mesh *m) {
  if (m->dim != 4) return false;
  if (m->fat_enc >= 0) return m->fat_enc == 4;
  return m->ns < (1 << 24) - 1 && m->opt.packed_faces != 5;
}
// start copies on the creating rank (replicas / snapshots: header word 13): 4D packed records by default while the
// copies stay below 24 GB, 3D on request (its hint table and most of its records are L2-resident: no gain there)
constexpr int64_t kStartCopyMaxBytes = 24ll << 30;
static bool want_startcopy(const dli_mesh *m) {
  if (m->dim < 3 || !m->d_fat || m->ns == 0 || m->opt.start_records < 0) return false;
  if (m->dim == 4 && !fat_pack4(m)) return false;
  if ((m->ns + m->grid.cells) >= 0x7fffff00ll) return false;
  if (m->opt.start_records > 0) return true;
  return m->dim == 4 && m->grid.cells * (int64_t)FatD<4, true>::BYTES <= kStartCopyMaxBytes;
}
// walk records of meshes with >= 2^21-2 faces (or opt.packed_faces == 2: testing) use delta-encoded neighbours
static bool fat_delta(const dli_mesh *m) {
  if (m->fat_enc >= 0) return m->fat_enc == 1;
  return m->ns >= (1 << 21) - 2 || m->opt.packed_faces == 2;
}

static SlotGrid slot_grid(const dli_mesh *m) {
  SlotGrid g;
  g.ox = m->grid.o[0]; g.oy = m->grid.o[1]; g.icx = m->grid.ic[0]; g.icy = m->grid.ic[1];
  g.gx = m->grid.g[0]; g.gy = m->grid.g[1];
  return g;
}
// Lays out the cell-indexed record array (creating rank): row starts, slot -> face table, record buffer.
static int build_slots(dli_mesh *m) {
  const int64_t cells = m->grid.cells;
  const int gx = m->grid.g[0], gy = m->grid.g[1];
  const SlotGrid g = slot_grid(m);
  const int4 *frec = reinterpret_cast<const int4 *>(m->d_srec);
  const double4 *vrec = reinterpret_cast<const double4 *>(m->d_vrec);
  const unsigned fb = (unsigned)((m->ns + 255) / 256), cb = (unsigned)((cells + 255) / 256);
  int *d_owner = nullptr, *d_rowcount = nullptr;
  std::vector<int> rowcount((size_t)gy), rowstart((size_t)gy + 1);
  cudaError_t e = cudaMalloc(&d_owner, sizeof(int) * (size_t)m->ns);
  if (e == cudaSuccess) e = cudaMalloc(&d_rowcount, sizeof(int) * (size_t)gy);
  if (e == cudaSuccess) e = cudaMemset(d_owner, 0x7f, sizeof(int) * (size_t)m->ns);
  if (e == cudaSuccess) e = cudaMemset(d_rowcount, 0, sizeof(int) * (size_t)gy);
  if (e == cudaSuccess) {
    k_slot_owner<<<cb, 256>>>(cells, m->d_hint, d_owner);
    k_slot_rowcount<<<fb, 256>>>((int)m->ns, g, d_owner, frec, vrec, d_rowcount);
    e = cudaMemcpy(rowcount.data(), d_rowcount, sizeof(int) * (size_t)gy, cudaMemcpyDeviceToHost);
  }
  int64_t acc = 0;
  for (int j = 0; j < gy; j++) { rowstart[(size_t)j] = (int)acc; acc += gx + rowcount[(size_t)j]; }
  rowstart[(size_t)gy] = (int)acc;
  if (e == cudaSuccess && acc >= 0x7fffff00ll) e = cudaErrorInvalidValue;
  m->nslots = acc;
  if (e == cudaSuccess) e = cudaMalloc(&m->d_rowstart, sizeof(int) * ((size_t)gy + 1));
  if (e == cudaSuccess) e = cudaMemcpy(m->d_rowstart, rowstart.data(), sizeof(int) * ((size_t)gy + 1), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&m->d_slot_face, sizeof(int32_t) * (size_t)m->nslots);
  if (e == cudaSuccess) e = cudaMalloc(&m->d_fat, 64 * (size_t)m->nslots);
  if (e == cudaSuccess) e = cudaMemset(d_rowcount, 0, sizeof(int) * (size_t)gy);   // reused as the per-row cursor
  if (e == cudaSuccess) {
    k_slot_fill_cells<<<cb, 256>>>(cells, gx, m->d_hint, m->d_rowstart, m->d_slot_face);
    k_slot_fill_extra<<<fb, 256>>>((int)m->ns, g, d_owner, frec, vrec, m->d_rowstart, d_rowcount, m->d_slot_face);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaFree(d_owner); cudaFree(d_rowcount);
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("cell-indexed record layout failed: ") + cudaGetErrorString(e));
  return DLI_OK;
}

// (re)builds the 64-byte walk records from the canonical records (values and geometry)
static int build_fat_records(dli_mesh *m) {
  if (m->cellmode) {
    if (!m->d_fat) {            // first build on the creating rank: lay the slots out, allocate
      int rc = build_slots(m);
      if (rc != DLI_OK) return rc;
    }
    // face -> lowest slot holding a copy of it (build-time table, ns ints)
    int *d_canon = nullptr;
    CU(cudaMalloc(&d_canon, sizeof(int) * (size_t)m->ns));
    cudaError_t e = cudaMemset(d_canon, 0x7f, sizeof(int) * (size_t)m->ns);
    unsigned int *d_cnt = reinterpret_cast<unsigned int *>(m->d_stats + ST_WORDS - 1);
    if (e == cudaSuccess) e = cudaMemset(d_cnt, 0, sizeof(unsigned int));
    unsigned int cnt = 0;
    if (e == cudaSuccess) {
      const unsigned sb = (unsigned)((m->nslots + 255) / 256);
      k_slot_canon<<<sb, 256>>>(m->nslots, m->d_slot_face, d_canon);
      k_build_fat2_slots<<<sb, 256>>>(m->nslots, slot_grid(m), m->d_hint, m->d_rowstart, m->d_slot_face, d_canon,
                                      reinterpret_cast<const int4 *>(m->d_srec), reinterpret_cast<const double4 *>(m->d_vrec),
                                      m->d_fat, d_cnt);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(&cnt, d_cnt, sizeof cnt, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemset(d_cnt, 0, sizeof(unsigned int));
    cudaFree(d_canon);
    if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("cell-indexed record build failed: ") + cudaGetErrorString(e));
    m->n_slow_faces = cnt;
    return DLI_OK;
  }
  if (!m->d_fat) return DLI_OK;
  if (!m->startcopy && m->fat_enc < 0 && want_startcopy(m)) {   // creating rank, first build: room for the copies
    uint4 *big = nullptr;
    const size_t bytes = (size_t)fat_record_bytes(m) * (size_t)(m->ns + m->grid.cells);
    if (cudaMalloc(&big, bytes) == cudaSuccess) {
      cudaFree(m->d_fat);
      m->d_fat = big;
      m->startcopy = true;
      m->nslots = m->ns + m->grid.cells;
      m->cap_fat = m->nslots;
    } else {
      cudaGetLastError();   // not enough memory for the copies: the mesh works without them
    }
  }
  unsigned int *d_cnt = reinterpret_cast<unsigned int *>(m->d_stats + ST_WORDS - 1);
  CU(cudaMemset(d_cnt, 0, sizeof(unsigned int)));
  if (m->dim == 2) {
    auto kb = fat_delta(m) ? k_build_fat2<true> : k_build_fat2<false>;
    kb<<<(unsigned)((m->ns + 255) / 256), 256>>>((int)m->ns, reinterpret_cast<const int4 *>(m->d_srec),
                                               reinterpret_cast<const double4 *>(m->d_vrec), m->d_fat, d_cnt);
  } else if (m->dim == 3) {
    k_build_fatD<3><<<(unsigned)((m->ns + 127) / 128), 128>>>((int)m->ns, m->d_srec, m->ss, m->d_vrec, m->vs, m->d_fat, d_cnt);
  } else if (fat_pack4(m)) {
    k_build_fatD<4, true><<<(unsigned)((m->ns + 127) / 128), 128>>>((int)m->ns, m->d_srec, m->ss, m->d_vrec, m->vs, m->d_fat, d_cnt);
  } else {
    k_build_fatD<4><<<(unsigned)((m->ns + 127) / 128), 128>>>((int)m->ns, m->d_srec, m->ss, m->d_vrec, m->vs, m->d_fat, d_cnt);
  }
  CU(cudaGetLastError());
  if (m->startcopy) {
    const int chunks = fat_record_bytes(m) / 16;
    const long long work = (long long)m->grid.cells * chunks;
    k_start_copies<<<(unsigned)((work + 255) / 256), 256>>>((long long)m->grid.cells, (int)m->ns, chunks, m->d_hint, m->d_fat);
    CU(cudaGetLastError());
  }
  unsigned int cnt = 0;
  CU(cudaMemcpy(&cnt, d_cnt, sizeof cnt, cudaMemcpyDeviceToHost));
  CU(cudaMemset(d_cnt, 0, sizeof(unsigned int)));
  m->n_slow_faces = cnt;
  return DLI_OK;
}

static int resolve_options(const dli_options *opt, dli_options *out) {
  dli_options_init(out);
  if (opt) {
    size_t sz = opt->struct_size > 0 ? std::min<size_t>(opt->struct_size, sizeof(dli_options)) : sizeof(dli_options);
    memcpy(out, opt, sz);
    out->struct_size = (int32_t)sizeof(dli_options);
  }
  return DLI_OK;
}

// 2D hull-edge table {face (caller id), slot, next, prev} in the order (face id, slot) ascending; enc_sn gets ~h in
// the hull slots (the device encoding of a hull neighbour)
static void hull_table2(int64_t nv, int64_t ns, const std::vector<int32_t> &sv, const std::vector<int32_t> &sn,
                        std::vector<int4> &hull, std::vector<int32_t> *enc_sn) {
  hull.clear();
  std::vector<int32_t> from((size_t)nv, -1), to((size_t)nv, -1);
  for (int64_t s = 0; s < ns; s++)
    for (int j = 0; j < 3; j++)
      if (sn[s * 3 + j] < 0) {
        int32_t h = (int32_t)hull.size();
        hull.push_back(make_int4((int)s, j, -1, -1));
        from[sv[s * 3 + (j + 1) % 3]] = h;
        to[sv[s * 3 + (j + 2) % 3]] = h;
        if (enc_sn) (*enc_sn)[s * 3 + j] = ~h;
      }
  for (auto &h : hull) {
    int32_t P = sv[(size_t)h.x * 3 + (h.y + 1) % 3], Q = sv[(size_t)h.x * 3 + (h.y + 2) % 3];
    h.z = from[Q];
    h.w = to[P];
  }
}

extern "C" int dli_mesh_create(int dim, int64_t nv, const double *coords, const double *f, int64_t ns,
                               const int32_t *sv_in, const int32_t *sn_in, const dli_options *opt, dli_mesh **out) {
  DLI_GUARD_BEGIN
  if (!out) return fail(DLI_ERR_ARG, "dli_mesh_create: out is NULL");
  *out = nullptr;
  if (dim < 2 || dim > 4) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_create: dim must be 2, 3 or 4");
  if (nv < 0 || ns < 0 || nv > 0x7fffffff || ns > 0x7ffffff0) return fail(DLI_ERR_ARG, "dli_mesh_create: bad sizes");
  if ((nv > 0 && (!coords || !f)) || (ns > 0 && !sv_in)) return fail(DLI_ERR_ARG, "dli_mesh_create: NULL array");
  const int d = dim, k = dim + 1;
  dli_mesh *m = new dli_mesh;
  m->dim = d; m->nv = nv; m->ns = ns;
  resolve_options(opt, &m->opt);
  int ndev = dli_device_count();
  if (ndev <= 0) { delete m; return fail(DLI_ERR_CUDA, "dli_mesh_create: no CUDA device available (this library has no CPU fallback)"); }
  if (m->opt.device >= 0) m->device = m->opt.device;
  else if (cudaGetDevice(&m->device) != cudaSuccess) { delete m; return fail(DLI_ERR_CUDA, "cudaGetDevice failed"); }
  DeviceGuard guard(m->device);
  if (!guard.ok) { delete m; return fail(DLI_ERR_CUDA, "dli_mesh_create: cannot select device"); }

  // ---- connectivity: copy, derive neighbours, canonical orientation
  m->h_sv.assign(sv_in, sv_in + (size_t)ns * k);
  m->h_sn.resize((size_t)ns * k);
  for (size_t i = 0; i < m->h_sv.size(); i++)
    if (m->h_sv[i] < 0 || m->h_sv[i] >= nv) { delete m; return fail(DLI_ERR_MESH, "dli_mesh_create: vertex index out of range"); }
  if (sn_in) {
    memcpy(m->h_sn.data(), sn_in, sizeof(int32_t) * (size_t)ns * k);
    for (size_t i = 0; i < m->h_sn.size(); i++)
      if (m->h_sn[i] >= ns) { delete m; return fail(DLI_ERR_MESH, "dli_mesh_create: neighbour index out of range"); }
      else if (m->h_sn[i] < 0) m->h_sn[i] = -1;
  } else if (derive_neighbors(d, ns, m->h_sv.data(), m->h_sn.data()) != 0) {
    delete m;
    return fail(DLI_ERR_MESH, "dli_mesh_create: a facet is shared by more than two simplices");
  }
  for (int64_t s = 0; s < ns; s++) {
    double pts[(DLI_MAX_DIM + 1) * DLI_MAX_DIM];
    for (int j = 0; j < k; j++)
      for (int c = 0; c < d; c++) pts[j * d + c] = coords[(size_t)m->h_sv[s * k + j] * d + c];
    int o = orient_host(d, pts);
    if (o == 0) {
      char b[160];
      snprintf(b, sizeof b, "dli_mesh_create: simplex %lld is degenerate (zero volume)", (long long)s);
      delete m;
      return fail(DLI_ERR_MESH, b);
    }
    if (o < 0) {  // exchange slots 0 and 1 (vertices and the neighbours opposite them)
      std::swap(m->h_sv[s * k], m->h_sv[s * k + 1]);
      std::swap(m->h_sn[s * k], m->h_sn[s * k + 1]);
    }
  }

  // ---- 2D hull edge table
  std::vector<int4> hull;
  std::vector<int32_t> enc_sn(m->h_sn);  // device encoding of neighbours
  if (d == 2) {
    hull_table2(nv, ns, m->h_sv, m->h_sn, hull, &enc_sn);
    m->nh = (int32_t)hull.size();
  }

  int rc = mesh_alloc_common(m);
  if (rc != DLI_OK) { mesh_free(m); return rc; }

  // ---- pack records
  {
    std::vector<double> vrec((size_t)std::max<int64_t>(nv, 1) * m->vs, 0.0);
    for (int64_t v = 0; v < nv; v++) {
      for (int c = 0; c < d; c++) vrec[(size_t)v * m->vs + c] = coords[v * d + c];
      vrec[(size_t)v * m->vs + d] = f[v];
    }
    // simplices in Morton order of their centroids; ids the caller sees go through d_orig
    std::vector<int32_t> perm, inv((size_t)ns);
    morton_order(d, nv, coords, ns, m->h_sv.data(), perm);
    for (int64_t i = 0; i < ns; i++) inv[(size_t)perm[(size_t)i]] = (int32_t)i;
    std::vector<int32_t> srec((size_t)std::max<int64_t>(ns, 1) * m->ss, 0);
    for (int64_t i = 0; i < ns; i++) {
      const int64_t s = perm[(size_t)i];
      for (int j = 0; j < k; j++) {
        int32_t nb = enc_sn[s * k + j];
        srec[(size_t)i * m->ss + j] = m->h_sv[s * k + j];
        srec[(size_t)i * m->ss + k + j] = nb >= 0 ? inv[(size_t)nb] : nb;
      }
    }
    for (auto &h : hull) h.x = inv[(size_t)h.x];
    if ((d == 2 && ns > 0 && ns < (1 << 21) - 2) || (d > 2 && ns > 0)) {   // what dli_mesh_update needs later
      m->h_inv = inv;
      m->h_xy.assign(coords, coords + (size_t)nv * d);
    }
    cudaError_t e = cudaMemcpy(m->d_vrec, vrec.data(), sizeof(double) * vrec.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(m->d_srec, srec.data(), sizeof(int32_t) * srec.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && ns > 0) e = cudaMemcpy(m->d_orig, perm.data(), sizeof(int32_t) * perm.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && d == 2 && !hull.empty())
      e = cudaMemcpy(m->d_hull, hull.data(), sizeof(int4) * hull.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { mesh_free(m); return fail(DLI_ERR_CUDA, std::string("mesh upload failed: ") + cudaGetErrorString(e)); }
  }
  if (ns > 0) {
    rc = build_hint_grid(m, coords);
    if (rc != DLI_OK) { mesh_free(m); return rc; }
  }
  if (m->d_fpack) {
    k_pack_faces<<<(unsigned)((ns + 255) / 256), 256>>>((int)ns, reinterpret_cast<const int4 *>(m->d_srec), m->d_fpack);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { mesh_free(m); return fail(DLI_ERR_CUDA, std::string("face packing failed: ") + cudaGetErrorString(e)); }
  }
  rc = build_fat_records(m);
  if (rc != DLI_OK) { mesh_free(m); return rc; }
  *out = m;
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_create")
}

extern "C" int dli_mesh_set_values(dli_mesh *m, const double *f) {
  if (!m || !f) return fail(DLI_ERR_ARG, "dli_mesh_set_values: NULL argument");
  DeviceGuard guard(m->device);
  // strided copy into the f slot of every vertex record
  CU(cudaMemcpy2D(m->d_vrec + m->dim, sizeof(double) * m->vs, f, sizeof(double), sizeof(double), (size_t)m->nv,
                  cudaMemcpyHostToDevice));
  return build_fat_records(m);
}

// grows a device buffer to new_bytes keeping the first keep_bytes
template <class T>
static cudaError_t grow_buffer(T *&p, size_t keep_bytes, size_t new_bytes) {
  T *q = nullptr;
  cudaError_t e = cudaMalloc(&q, new_bytes);
  if (e != cudaSuccess) return e;
  if (p && keep_bytes) e = cudaMemcpy(q, p, keep_bytes, cudaMemcpyDeviceToDevice);
  if (e != cudaSuccess) { cudaFree(q); return e; }
  cudaFree(p);
  p = q;
  return cudaSuccess;
}

// dD (3D, 4D packed records) form of dli_mesh_update: same contract, no hull table, no packed faces.  The start copies
// (MeshDView::start_copies) are dropped by the first patch -- their region starts at record ns, which moves -- and come
// back with the next full flatten; the start-hint grid is kept, as in 2D.
static int mesh_update_dd(dli_mesh *m, int64_t nv_total, const double *new_coords, const double *new_f, int64_t n_fchg,
                          const int32_t *fchg_ids, const double *fchg_vals, int64_t ns_total, int64_t n_dirty,
                          const int32_t *dirty_ids, const int32_t *dirty_sv, const int32_t *dirty_sn) {
  const int d = m->dim, k = d + 1;
  const int64_t nv0 = m->nv, ns0 = m->ns;
  static const bool trace = getenv("DLI_TRACE_UPDATE") != nullptr;   // developer aid: phase times on stderr
  auto now = [] { return std::chrono::steady_clock::now(); };
  auto t_begin = now();
  auto lap = [&](const char *what) {
    if (!trace) return;
    cudaDeviceSynchronize();
    auto t = now();
    fprintf(stderr, "[dli_mesh_update dD] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t_begin).count());
    t_begin = t;
  };
  if (ns0 == 0 || m->h_inv.empty() || m->h_sv.empty() || m->h_xy.empty())
    return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: mesh cannot be patched (replica or empty)");
  if (d == 4 && m->d_fat && !fat_pack4(m))
    return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: 4D split walk records cannot be patched");
  if (d == 4 && ns_total >= (1 << 24) - 1)
    return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: ids no longer fit the packed records");
  for (int64_t v = 0; v < nv_total - nv0; v++)
    for (int c = 0; c < d; c++) {
      const double x = new_coords[d * v + c], lo = m->grid.o[c], hi = lo + m->grid.cell[c] * m->grid.g[c];
      if (!(x >= lo && x <= hi)) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: new vertex outside the start-hint grid");
    }
  {
    int64_t appended = 0;
    std::vector<uint8_t> seen((size_t)(ns_total - ns0), 0);
    for (int64_t i = 0; i < n_dirty; i++) {
      const int32_t id = dirty_ids[i];
      if (id < 0 || id >= ns_total) return fail(DLI_ERR_MESH, "dli_mesh_update: simplex id out of range");
      if (id >= ns0) {
        if (seen[(size_t)(id - ns0)]) return fail(DLI_ERR_MESH, "dli_mesh_update: simplex id listed twice");
        seen[(size_t)(id - ns0)] = 1;
        appended++;
      }
      for (int j = 0; j < k; j++) {
        if (dirty_sv[k * i + j] < 0 || dirty_sv[k * i + j] >= nv_total) return fail(DLI_ERR_MESH, "dli_mesh_update: vertex index out of range");
        if (dirty_sn[k * i + j] >= ns_total) return fail(DLI_ERR_MESH, "dli_mesh_update: neighbour index out of range");
      }
    }
    if (appended != ns_total - ns0) return fail(DLI_ERR_MESH, "dli_mesh_update: every new simplex id must be listed");
    for (int64_t i = 0; i < n_fchg; i++)
      if (fchg_ids[i] < 0 || fchg_ids[i] >= nv0) return fail(DLI_ERR_MESH, "dli_mesh_update: value-change vertex out of range");
  }
  DeviceGuard guard(m->device);
  if (!guard.ok) return fail(DLI_ERR_CUDA, "dli_mesh_update: cannot select device");
  // ---- orientation-canonical rows (checked first, then applied)
  auto X = [&](int32_t v, int c) { return v < nv0 ? m->h_xy[(size_t)d * v + c] : new_coords[(size_t)d * (v - nv0) + c]; };
  std::vector<uint8_t> flip((size_t)n_dirty, 0);
  for (int64_t i = 0; i < n_dirty; i++) {
    double pts[(DLI_MAX_DIM + 1) * DLI_MAX_DIM];
    for (int j = 0; j < k; j++)
      for (int c = 0; c < d; c++) pts[j * d + c] = X(dirty_sv[k * i + j], c);
    const int o = orient_host(d, pts);
    if (o == 0) return fail(DLI_ERR_MESH, "dli_mesh_update: degenerate simplex (zero volume)");
    flip[(size_t)i] = o < 0;
  }
  m->h_xy.insert(m->h_xy.end(), new_coords, new_coords + (size_t)(nv_total - nv0) * d);
  m->h_sv.resize((size_t)ns_total * k, 0);
  m->h_sn.resize((size_t)ns_total * k, -1);
  auto dev_index = [&](int64_t s) -> int32_t { return s < ns0 ? m->h_inv[(size_t)s] : (int32_t)s; };
  const int ss = m->ss, vs = m->vs;
  std::vector<int32_t> idx((size_t)n_dirty), rows((size_t)n_dirty * ss, 0);
  for (int64_t i = 0; i < n_dirty; i++) {
    const int64_t sid = dirty_ids[i];
    int32_t v[DLI_MAX_DIM + 1], nb[DLI_MAX_DIM + 1];
    for (int j = 0; j < k; j++) { v[j] = dirty_sv[k * i + j]; nb[j] = dirty_sn[k * i + j] < 0 ? -1 : dirty_sn[k * i + j]; }
    if (flip[(size_t)i]) { std::swap(v[0], v[1]); std::swap(nb[0], nb[1]); }
    idx[(size_t)i] = dev_index(sid);
    for (int j = 0; j < k; j++) {
      m->h_sv[(size_t)sid * k + j] = v[j];
      m->h_sn[(size_t)sid * k + j] = nb[j];
      rows[(size_t)i * ss + j] = v[j];
      rows[(size_t)i * ss + k + j] = nb[j] >= 0 ? dev_index(nb[j]) : -1;
    }
  }
  lap("validate + host rows");
  // ---- capacity
  cudaError_t e = cudaDeviceSynchronize();
  const size_t rec = m->d_fat ? (size_t)fat_record_bytes(m) : 0;
  if (m->cap_fat == 0) m->cap_fat = m->startcopy ? m->nslots : m->cap_s;
  if (m->startcopy) { m->startcopy = false; m->nslots = ns0; }   // (the region behind the records becomes spare capacity)
  if (e == cudaSuccess && nv_total > m->cap_v) {
    const int64_t cap = nv_total + nv_total / 2 + 1024;
    e = grow_buffer(m->d_vrec, sizeof(double) * (size_t)nv0 * vs, sizeof(double) * (size_t)cap * vs);
    if (e == cudaSuccess) m->cap_v = cap;
  }
  if (e == cudaSuccess && ns_total > m->cap_s) {
    const int64_t cap = ns_total + ns_total / 2 + 1024;
    e = grow_buffer(m->d_srec, sizeof(int32_t) * (size_t)ns0 * ss, sizeof(int32_t) * (size_t)cap * ss);
    if (e == cudaSuccess) e = grow_buffer(m->d_orig, sizeof(int32_t) * (size_t)ns0, sizeof(int32_t) * (size_t)cap);
    if (e == cudaSuccess) m->cap_s = cap;
  }
  if (e == cudaSuccess && m->d_fat && ns_total > m->cap_fat) {
    const int64_t cap = ns_total + ns_total / 2 + 1024;
    e = grow_buffer(m->d_fat, rec * (size_t)ns0, rec * (size_t)cap);
    if (e == cudaSuccess) m->cap_fat = cap;
  }
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_update: growing the device buffers failed (the mesh is now inconsistent; destroy it): ") + cudaGetErrorString(e));
  lap("capacity");
  // ---- upload (one scratch buffer) + rebuild the walk records of the listed simplices
  const int64_t nnew = nv_total - nv0;
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t o_idx = 0, o_ids = al(o_idx + sizeof(int32_t) * (size_t)n_dirty), o_rows = al(o_ids + sizeof(int32_t) * (size_t)n_dirty),
               o_vid = al(o_rows + sizeof(int32_t) * (size_t)ss * (size_t)n_dirty), o_val = al(o_vid + sizeof(int32_t) * (size_t)n_fchg),
               o_new = al(o_val + sizeof(double) * (size_t)n_fchg), o_cnt = al(o_new + sizeof(double) * (size_t)vs * (size_t)nnew),
               total = o_cnt + 256;
  if (total > m->upd_cap) {
    cudaFree(m->d_upd); m->d_upd = nullptr; m->upd_cap = 0;
    e = cudaMalloc(&m->d_upd, total * 2);
    if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_update: scratch allocation failed (the mesh is now inconsistent; destroy it): ") + cudaGetErrorString(e));
    m->upd_cap = total * 2;
  }
  std::vector<unsigned char> stage(total, 0);
  memcpy(stage.data() + o_idx, idx.data(), sizeof(int32_t) * (size_t)n_dirty);
  memcpy(stage.data() + o_ids, dirty_ids, sizeof(int32_t) * (size_t)n_dirty);
  memcpy(stage.data() + o_rows, rows.data(), sizeof(int32_t) * (size_t)ss * (size_t)n_dirty);
  if (n_fchg > 0) {
    memcpy(stage.data() + o_vid, fchg_ids, sizeof(int32_t) * (size_t)n_fchg);
    memcpy(stage.data() + o_val, fchg_vals, sizeof(double) * (size_t)n_fchg);
  }
  {
    double *vr = reinterpret_cast<double *>(stage.data() + o_new);
    for (int64_t v = 0; v < nnew; v++) {
      for (int c = 0; c < d; c++) vr[(size_t)vs * v + c] = new_coords[(size_t)d * v + c];
      vr[(size_t)vs * v + d] = new_f[v];
    }
  }
  unsigned int cnt = 0;
  e = cudaMemcpy(m->d_upd, stage.data(), total, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && nnew > 0)
    e = cudaMemcpyAsync(m->d_vrec + (size_t)nv0 * vs, m->d_upd + o_new, sizeof(double) * (size_t)vs * (size_t)nnew, cudaMemcpyDeviceToDevice, 0);
  if (e == cudaSuccess && n_fchg > 0)
    k_update_values_d<<<(unsigned)((n_fchg + 127) / 128), 128>>>((int)n_fchg, reinterpret_cast<const int32_t *>(m->d_upd + o_vid),
                                                                 reinterpret_cast<const double *>(m->d_upd + o_val), m->d_vrec, vs, d);
  if (e == cudaSuccess && n_dirty > 0) {
    const unsigned gb = (unsigned)((n_dirty + 127) / 128);
    const int32_t *d_idx = reinterpret_cast<const int32_t *>(m->d_upd + o_idx);
    unsigned int *d_cnt = reinterpret_cast<unsigned int *>(m->d_upd + o_cnt);
    k_update_simplices_d<<<gb, 128>>>((int)n_dirty, d_idx, reinterpret_cast<const int32_t *>(m->d_upd + o_rows), ss,
                                      reinterpret_cast<const int32_t *>(m->d_upd + o_ids), m->d_srec, m->d_orig);
    lap("upload + canonical records");
    if (m->d_fat) {
      const unsigned cb = (unsigned)((ns_total + 255) / 256);
      if (d == 3) {
        k_build_fatD<3><<<gb, 128>>>((int)ns_total, m->d_srec, ss, m->d_vrec, vs, m->d_fat, d_cnt + 1, d_idx, (int)n_dirty);
        k_count_slow_d<3, false><<<cb, 256>>>((int)ns_total, m->d_fat, d_cnt);
      } else {
        k_build_fatD<4, true><<<gb, 128>>>((int)ns_total, m->d_srec, ss, m->d_vrec, vs, m->d_fat, d_cnt + 1, d_idx, (int)n_dirty);
        k_count_slow_d<4, true><<<cb, 256>>>((int)ns_total, m->d_fat, d_cnt);
      }
    }
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(&cnt, d_cnt, sizeof cnt, cudaMemcpyDeviceToHost);
    lap("walk records + slow recount");
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_update failed (the mesh is now inconsistent; destroy it): ") + cudaGetErrorString(e));
  m->h_inv.resize((size_t)ns_total);
  for (int64_t s2 = ns0; s2 < ns_total; s2++) m->h_inv[(size_t)s2] = (int32_t)s2;
  m->nv = nv_total; m->ns = ns_total; m->nslots = ns_total;
  if (m->d_fat && n_dirty > 0) m->n_slow_faces = cnt;
  if (!(m->opt.max_walk_steps > 0)) m->max_steps = (int)std::min<int64_t>(m->ns + 16, 0x7fffffff);
  {
    std::lock_guard<std::mutex> lk(m->mu);
    if (m->items_ready) { cudaFree(m->d_item_verts); m->d_item_verts = nullptr; m->items_ready = false; }
  }
  m->n_updates++;
  return DLI_OK;
}

extern "C" int dli_mesh_update(dli_mesh *m, int64_t nv_total, const double *new_coords, const double *new_f,
                               int64_t n_fchg, const int32_t *fchg_ids, const double *fchg_vals, int64_t ns_total,
                               int64_t n_dirty, const int32_t *dirty_ids, const int32_t *dirty_sv, const int32_t *dirty_sn) {
  DLI_GUARD_BEGIN
  if (!m) return fail(DLI_ERR_ARG, "dli_mesh_update: NULL mesh");
  const int64_t nv0 = m->nv, ns0 = m->ns;
  if (nv_total < nv0 || ns_total < ns0 || n_fchg < 0 || n_dirty < 0 || nv_total > 0x7fffffff)
    return fail(DLI_ERR_ARG, "dli_mesh_update: sizes must not shrink");
  if ((nv_total > nv0 && (!new_coords || !new_f)) || (n_fchg > 0 && (!fchg_ids || !fchg_vals)) ||
      (n_dirty > 0 && (!dirty_ids || !dirty_sv || !dirty_sn)))
    return fail(DLI_ERR_ARG, "dli_mesh_update: NULL array");
  if (m->dim > 2)
    return mesh_update_dd(m, nv_total, new_coords, new_f, n_fchg, fchg_ids, fchg_vals, ns_total, n_dirty, dirty_ids, dirty_sv, dirty_sn);
  // ---- can this mesh be patched?
  if (m->dim != 2 || m->cellmode || ns0 == 0 || m->h_inv.empty() || m->h_sv.empty())
    return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: mesh cannot be patched (not 2D, replica, empty or cell-indexed records)");
  if (ns_total >= (1 << 21) - 2 || (m->d_fpack && nv_total > (1 << 21)) || fat_delta(m))
    return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: ids no longer fit the packed records");
  for (int64_t v = 0; v < nv_total - nv0; v++)
    for (int c = 0; c < 2; c++) {
      const double x = new_coords[2 * v + c], lo = m->grid.o[c], hi = lo + m->grid.cell[c] * m->grid.g[c];
      if (!(x >= lo && x <= hi)) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_update: new vertex outside the start-hint grid");
    }
  // ---- validate the dirty set
  {
    int64_t appended = 0;
    std::vector<uint8_t> seen((size_t)(ns_total - ns0), 0);
    for (int64_t i = 0; i < n_dirty; i++) {
      const int32_t id = dirty_ids[i];
      if (id < 0 || id >= ns_total) return fail(DLI_ERR_MESH, "dli_mesh_update: simplex id out of range");
      if (id >= ns0) {
        if (seen[(size_t)(id - ns0)]) return fail(DLI_ERR_MESH, "dli_mesh_update: simplex id listed twice");
        seen[(size_t)(id - ns0)] = 1;
        appended++;
      }
      for (int j = 0; j < 3; j++) {
        if (dirty_sv[3 * i + j] < 0 || dirty_sv[3 * i + j] >= nv_total) return fail(DLI_ERR_MESH, "dli_mesh_update: vertex index out of range");
        if (dirty_sn[3 * i + j] >= ns_total) return fail(DLI_ERR_MESH, "dli_mesh_update: neighbour index out of range");
      }
    }
    if (appended != ns_total - ns0) return fail(DLI_ERR_MESH, "dli_mesh_update: every new simplex id must be listed");
    for (int64_t i = 0; i < n_fchg; i++)
      if (fchg_ids[i] < 0 || fchg_ids[i] >= nv0) return fail(DLI_ERR_MESH, "dli_mesh_update: value-change vertex out of range");
  }
  DeviceGuard guard(m->device);
  if (!guard.ok) return fail(DLI_ERR_CUDA, "dli_mesh_update: cannot select device");
  // ---- host connectivity: orientation-canonical rows of the dirty simplices (checked first, then applied in place)
  auto X = [&](int32_t v, int c) { return v < nv0 ? m->h_xy[2 * (size_t)v + c] : new_coords[2 * (size_t)(v - nv0) + c]; };
  std::vector<uint8_t> flip((size_t)n_dirty, 0);
  bool hull_touched = false;
  for (int64_t i = 0; i < n_dirty; i++) {
    const int64_t s = dirty_ids[i];
    const int32_t *v = dirty_sv + 3 * i;
    const int o = orient2d(X(v[0], 0), X(v[0], 1), X(v[1], 0), X(v[1], 1), X(v[2], 0), X(v[2], 1));
    if (o == 0) return fail(DLI_ERR_MESH, "dli_mesh_update: degenerate simplex (zero area)");
    flip[(size_t)i] = o < 0;
    for (int j = 0; j < 3; j++)
      if (dirty_sn[3 * i + j] < 0 || (s < ns0 && m->h_sn[(size_t)s * 3 + j] < 0)) hull_touched = true;
  }
  m->h_xy.insert(m->h_xy.end(), new_coords, new_coords + (size_t)(nv_total - nv0) * 2);
  m->h_sv.resize((size_t)ns_total * 3, 0);
  m->h_sn.resize((size_t)ns_total * 3, -1);
  std::vector<int32_t> &sv = m->h_sv, &sn = m->h_sn;
  for (int64_t i = 0; i < n_dirty; i++) {
    const int64_t s = dirty_ids[i];
    int32_t v[3] = {dirty_sv[3 * i], dirty_sv[3 * i + 1], dirty_sv[3 * i + 2]};
    int32_t nb[3] = {dirty_sn[3 * i], dirty_sn[3 * i + 1], dirty_sn[3 * i + 2]};
    if (flip[(size_t)i]) { std::swap(v[0], v[1]); std::swap(nb[0], nb[1]); }
    for (int j = 0; j < 3; j++) {
      sv[(size_t)s * 3 + j] = v[j];
      sn[(size_t)s * 3 + j] = nb[j] < 0 ? -1 : nb[j];
    }
  }
  // ---- the set of simplices to re-send: the dirty ones, plus every hull-adjacent one when the hull table is rebuilt
  std::vector<int32_t> list(dirty_ids, dirty_ids + n_dirty);
  std::vector<int4> hull;
  std::unordered_map<int64_t, int32_t> hull_of;   // face * 3 + slot -> hull edge (only filled when the hull is rebuilt)
  if (hull_touched) {
    hull_table2(nv_total, ns_total, sv, sn, hull, nullptr);
    std::vector<uint8_t> in_list((size_t)ns_total, 0);
    for (int32_t s : list) in_list[(size_t)s] = 1;
    hull_of.reserve(hull.size() * 2);
    for (size_t h = 0; h < hull.size(); h++) {
      hull_of[(int64_t)hull[h].x * 3 + hull[h].y] = (int32_t)h;
      if (!in_list[(size_t)hull[h].x]) { in_list[(size_t)hull[h].x] = 1; list.push_back(hull[h].x); }
    }
  }
  // (without a hull change the ~h encodings of untouched simplices stay valid and no dirty simplex has a hull slot)
  auto dev_index = [&](int64_t s) -> int32_t { return s < ns0 ? m->h_inv[(size_t)s] : (int32_t)s; };
  const int64_t nl = (int64_t)list.size();
  std::vector<int32_t> idx((size_t)nl), rows((size_t)nl * 8, 0);
  for (int64_t i = 0; i < nl; i++) {
    const int64_t s = list[(size_t)i];
    idx[(size_t)i] = dev_index(s);
    for (int j = 0; j < 3; j++) {
      const int32_t nb = sn[(size_t)s * 3 + j];
      rows[(size_t)i * 8 + j] = sv[(size_t)s * 3 + j];
      rows[(size_t)i * 8 + 3 + j] = nb >= 0 ? dev_index(nb) : ~hull_of.at(s * 3 + j);
    }
  }
  for (auto &h : hull) h.x = dev_index(h.x);
  // ---- capacity
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess && nv_total > m->cap_v) {
    const int64_t cap = nv_total + nv_total / 2 + 1024;
    e = grow_buffer(m->d_vrec, sizeof(double) * (size_t)nv0 * m->vs, sizeof(double) * (size_t)cap * m->vs);
    if (e == cudaSuccess) m->cap_v = cap;
  }
  if (e == cudaSuccess && ns_total > m->cap_s) {
    const int64_t cap = std::min<int64_t>(ns_total + ns_total / 2 + 1024, (1 << 21) - 2);
    e = grow_buffer(m->d_srec, sizeof(int32_t) * (size_t)ns0 * m->ss, sizeof(int32_t) * (size_t)cap * m->ss);
    if (e == cudaSuccess) e = grow_buffer(m->d_orig, sizeof(int32_t) * (size_t)ns0, sizeof(int32_t) * (size_t)cap);
    if (e == cudaSuccess && m->d_fpack) e = grow_buffer(m->d_fpack, sizeof(ulonglong2) * (size_t)ns0, sizeof(ulonglong2) * (size_t)cap);
    if (e == cudaSuccess && m->d_fat) e = grow_buffer(m->d_fat, (size_t)64 * (size_t)ns0, (size_t)64 * (size_t)cap);
    if (e == cudaSuccess) m->cap_s = cap;
  }
  if (e == cudaSuccess && hull_touched && (int64_t)hull.size() > m->cap_h) {
    const int32_t cap = (int32_t)hull.size() * 2 + 64;
    e = grow_buffer(m->d_hull, 0, sizeof(int4) * (size_t)cap);
    if (e == cudaSuccess) m->cap_h = cap;
  }
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_update: growing the device buffers failed (the mesh is now inconsistent; destroy it): ") + cudaGetErrorString(e));
  // ---- upload + rebuild the derived records of the listed simplices (one scratch buffer, one upload)
  const int64_t nnew = nv_total - nv0;
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t o_idx = 0, o_ids = al(o_idx + sizeof(int32_t) * (size_t)nl), o_rows = al(o_ids + sizeof(int32_t) * (size_t)nl),
               o_vid = al(o_rows + sizeof(int32_t) * 8 * (size_t)nl), o_val = al(o_vid + sizeof(int32_t) * (size_t)n_fchg),
               o_new = al(o_val + sizeof(double) * (size_t)n_fchg), o_delta = al(o_new + sizeof(double) * 4 * (size_t)nnew),
               total = o_delta + 256;
  if (total > m->upd_cap) {
    cudaFree(m->d_upd); m->d_upd = nullptr; m->upd_cap = 0;
    e = cudaMalloc(&m->d_upd, total * 2);
    if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_update: scratch allocation failed (the mesh is now inconsistent; destroy it): ") + cudaGetErrorString(e));
    m->upd_cap = total * 2;
  }
  std::vector<unsigned char> stage(total, 0);
  memcpy(stage.data() + o_idx, idx.data(), sizeof(int32_t) * (size_t)nl);
  memcpy(stage.data() + o_ids, list.data(), sizeof(int32_t) * (size_t)nl);
  memcpy(stage.data() + o_rows, rows.data(), sizeof(int32_t) * 8 * (size_t)nl);
  if (n_fchg > 0) {
    memcpy(stage.data() + o_vid, fchg_ids, sizeof(int32_t) * (size_t)n_fchg);
    memcpy(stage.data() + o_val, fchg_vals, sizeof(double) * (size_t)n_fchg);
  }
  {
    double *vr = reinterpret_cast<double *>(stage.data() + o_new);
    for (int64_t v = 0; v < nnew; v++) { vr[4 * v] = new_coords[2 * v]; vr[4 * v + 1] = new_coords[2 * v + 1]; vr[4 * v + 2] = new_f[v]; }
  }
  int delta = 0;
  e = cudaMemcpy(m->d_upd, stage.data(), total, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && nnew > 0)
    e = cudaMemcpyAsync(m->d_vrec + (size_t)nv0 * 4, m->d_upd + o_new, sizeof(double) * 4 * (size_t)nnew, cudaMemcpyDeviceToDevice, 0);
  if (e == cudaSuccess && n_fchg > 0)
    k_update_values<<<(unsigned)((n_fchg + 127) / 128), 128>>>((int)n_fchg, reinterpret_cast<const int32_t *>(m->d_upd + o_vid),
                                                               reinterpret_cast<const double *>(m->d_upd + o_val),
                                                               reinterpret_cast<double4 *>(m->d_vrec));
  if (e == cudaSuccess && nl > 0) {
    const unsigned gb = (unsigned)((nl + 127) / 128);
    int4 *frec = reinterpret_cast<int4 *>(m->d_srec);
    const int32_t *d_idx = reinterpret_cast<const int32_t *>(m->d_upd + o_idx);
    k_update_faces<<<gb, 128>>>((int)nl, d_idx, reinterpret_cast<const int4 *>(m->d_upd + o_rows),
                                reinterpret_cast<const int32_t *>(m->d_upd + o_ids), frec, m->d_orig);
    if (m->d_fpack) k_pack_faces_list<<<gb, 128>>>((int)nl, d_idx, frec, m->d_fpack);
    if (m->d_fat)
      k_build_fat2_list<<<gb, 128>>>((int)nl, d_idx, (int)ns0, frec, reinterpret_cast<const double4 *>(m->d_vrec), m->d_fat,
                                     reinterpret_cast<int *>(m->d_upd + o_delta));
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpy(&delta, m->d_upd + o_delta, sizeof delta, cudaMemcpyDeviceToHost);
  }
  if (e == cudaSuccess && hull_touched && !hull.empty())
    e = cudaMemcpy(m->d_hull, hull.data(), sizeof(int4) * hull.size(), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_update failed (the mesh is now inconsistent; destroy it): ") + cudaGetErrorString(e));
  // ---- commit the host side
  m->h_inv.resize((size_t)ns_total);
  for (int64_t s = ns0; s < ns_total; s++) m->h_inv[(size_t)s] = (int32_t)s;
  m->nv = nv_total; m->ns = ns_total; m->nslots = ns_total;
  if (hull_touched) m->nh = (int32_t)hull.size();
  m->n_slow_faces += delta;
  if (!(m->opt.max_walk_steps > 0)) m->max_steps = (int)std::min<int64_t>(m->ns + 16, 0x7fffffff);
  {
    std::lock_guard<std::mutex> lk(m->mu);
    if (m->items_ready) { cudaFree(m->d_item_verts); m->d_item_verts = nullptr; m->items_ready = false; }
  }
  m->n_updates++;
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_update")
}

extern "C" void dli_mesh_destroy(dli_mesh *m) { mesh_free(m); }
extern "C" int dli_mesh_dim(const dli_mesh *m) { return m ? m->dim : 0; }
extern "C" int64_t dli_mesh_num_vertices(const dli_mesh *m) { return m ? m->nv : 0; }
extern "C" int64_t dli_mesh_num_simplices(const dli_mesh *m) { return m ? m->ns : 0; }
extern "C" int64_t dli_mesh_device_bytes(const dli_mesh *m) {
  if (!m) return 0;
  return (int64_t)(sizeof(double) * m->nv * m->vs + sizeof(int32_t) * m->ns * m->ss + sizeof(int4) * m->nh +
                   sizeof(int32_t) * m->grid.cells + sizeof(int32_t) * m->ns + (m->d_fpack ? sizeof(ulonglong2) * m->ns : 0) +
                   (m->d_fat ? (int64_t)fat_record_bytes(m) * m->nslots : 0) +
                   (m->cellmode ? (int64_t)sizeof(int32_t) * (m->nslots + m->grid.g[1] + 1) : 0));
}
extern "C" int dli_mesh_get_simplices(const dli_mesh *m, int32_t *sv, int32_t *sn) {
  DLI_GUARD_BEGIN
  if (!m) return fail(DLI_ERR_ARG, "dli_mesh_get_simplices: NULL mesh");
  if (m->h_sv.empty() && m->ns > 0) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_get_simplices: mesh was created from a header (no host copy)");
  if (sv) memcpy(sv, m->h_sv.data(), sizeof(int32_t) * m->h_sv.size());
  if (sn) memcpy(sn, m->h_sn.data(), sizeof(int32_t) * m->h_sn.size());
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_get_simplices")
}

// ---------------------------------------------------------------------------------------------
// interp / locate dispatch
// ---------------------------------------------------------------------------------------------
#ifndef DLI_X_FULL_CTAS
#define DLI_X_FULL_CTAS 6     // CTAs per SM of the 2D exact pass (grid-stride over the straggler queue)
#endif
#ifndef DLI_X_FULL_THREADS
#define DLI_X_FULL_THREADS 128
#endif
namespace {

Mesh2View view2(const dli_mesh *m) {
  Mesh2View v;
  v.vrec = reinterpret_cast<const double4 *>(m->d_vrec);
  v.frec = reinterpret_cast<const int4 *>(m->d_srec);
  v.fpack = m->d_fpack;
  v.fat = m->d_fat;
  v.hull = m->d_hull;
  v.hint = m->d_hint;
  v.orig = m->d_orig;
  v.slot_face = m->d_slot_face;
  v.rowstart = m->d_rowstart;
  v.ox = m->grid.o[0]; v.oy = m->grid.o[1];
  v.icx = m->grid.ic[0]; v.icy = m->grid.ic[1];
  v.gx = m->grid.g[0]; v.gy = m->grid.g[1];
  v.ns = (int)m->ns; v.nh = m->nh;
  v.nv = (int)m->nv; v.nslots = m->nslots;
  v.max_steps = m->max_steps; v.fast_steps = m->fast_steps;
  v.outside_nan = m->opt.outside_nan;
  v.straggler_frac = m->h_straggler_frac;
  return v;
}
MeshDView viewD(const dli_mesh *m) {
  MeshDView v;
  v.vrec = m->d_vrec; v.srec = m->d_srec; v.hint = m->d_hint; v.orig = m->d_orig;
  v.fat = m->d_fat;
  v.epi = (m->d_fat && m->dim == 4 && !fat_pack4(m)) ? m->d_fat + (size_t)m->ns * FatD<4>::CHUNKS : nullptr;
  for (int k = 0; k < DLI_MAX_DIM; k++) { v.o[k] = m->grid.o[k]; v.ic[k] = m->grid.ic[k]; v.g[k] = m->grid.g[k]; }
  v.ns = (int)m->ns; v.max_steps = m->max_steps; v.fast_steps = m->fast_steps;
  v.nv = (int)m->nv; v.cells = m->grid.cells;
  v.start_copies = m->startcopy ? 1 : 0;
  v.nrec = m->startcopy ? m->ns + m->grid.cells : m->ns;
  v.straggler_frac = m->h_straggler_frac;
  return v;
}

constexpr int64_t kChunk = 1ll << 28;      // queries per launch pair (int32 indices in the straggler queue)
// Straggler queue entries per launch pair: 1/16 of the launch (at least 2^16), 8 B each from the stream-ordered pool --
// 50 MB for a 100M-query launch (it used to be one entry per query: 0.8 GB).  Queries that do not fit are marked in
// their result slot and found again by the exact pass (enqueue_straggler in dli_kernels.cuh): an overflow costs one
// scan of the results, not a re-run of the batch.
// The exact pass reports the straggler fraction of its launch through a mapped host word (no copy, no synchronisation;
// a stale or concurrent value is only a worse hint): a mesh whose batches keep sending more than 1/16 of their queries
// to the exact pass (lattice queries on a regular grid: 30-60 %) gets a queue that holds them from its second call on.
static int64_t queue_capacity(const dli_mesh *m, int64_t n) {
  double frac = 1.0 / 16;
  if (m->h_straggler_frac) {
    const float seen = *(volatile float *)m->h_straggler_frac;
    if (seen > 0.05f && seen <= 1.0f) frac = std::min(1.0, 1.25 * (double)seen + 0.01);
  }
  return std::min<int64_t>(n, std::max<int64_t>(1 << 16, (int64_t)(frac * (double)n)));
}

struct Launch {
  const dli_mesh *m;
  cudaStream_t st;
  int32_t *queue;
  unsigned int *qcount;
  unsigned int qcap = 0;
  unsigned long long *stats;
  int launches = 0;
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  void mark(int i) {
    if (!m->collect_timing) return;
    if (cudaEventCreate(&ev[i]) == cudaSuccess) cudaEventRecord(ev[i], st);
    if (i == 2) {
      std::lock_guard<std::mutex> lk(m->ev_mu);
      for (int k = 0; k < 3; k++) m->ev.push_back(ev[k]);
    }
  }
};

template <bool STATS, int EPI, int OUT = 0>
void launch2(Launch &L, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto) {
  auto fast = L.m->d_fpack ? k_interp2_fast<STATS, EPI, true, OUT> : k_interp2_fast<STATS, EPI, false, OUT>;
  Mesh2View v = view2(L.m);
  const bool use_fat = (EPI == 0 && L.m->d_fat);
  int tpb = L.m->opt.threads_per_block > 0 ? L.m->opt.threads_per_block : (use_fat ? 128 : 256);
  tpb = std::min(use_fat ? DLI_X_FAT_THREADS : 256, std::max(32, tpb / 32 * 32));  // never above the kernel's launch bound
  // queries per warp: large enough to amortise the drain at the end of a warp's range, small enough to
  // give every SM many CTAs
  int per_warp = L.m->opt.queries_per_warp > 0 ? L.m->opt.queries_per_warp : 1024;
  while (per_warp > 32 && n / per_warp < 148 * 64) per_warp >>= 1;
  int64_t warps = (n + per_warp - 1) / per_warp;
  int wpb = tpb / 32;
  L.mark(0);
  if (use_fat) {
    const bool ring = L.m->opt.prefetch_ring >= 0, delta = fat_delta(L.m);
    auto kf = ring ? (delta ? k_interp2_fat<STATS, true, true, OUT> : k_interp2_fat<STATS, true, false, OUT>)
                   : (delta ? k_interp2_fat<STATS, false, true, OUT> : k_interp2_fat<STATS, false, false, OUT>);
    // cell-indexed records: no hint to prefetch, and the query prefetch alone does not pay
    if (L.m->cellmode) kf = k_interp2_fat<STATS, false, true, OUT, true>;
    kf<<<(unsigned)((warps + wpb - 1) / wpb), tpb, (ring && !L.m->cellmode) ? (size_t)wpb * kRingBytesPerWarp : 0, L.st>>>(
        v, n, per_warp, reinterpret_cast<const double2 *>(xs), ys, so, sto, L.queue, L.qcount, L.qcap, L.stats);
  }
  else
    fast<<<(unsigned)((warps + wpb - 1) / wpb), tpb, 0, L.st>>>(
        v, n, per_warp, reinterpret_cast<const double2 *>(xs), ys, so, sto, L.queue, L.qcount, L.qcap, L.stats);
  L.mark(1);
  k_interp2_full<STATS, EPI, OUT><<<148 * DLI_X_FULL_CTAS, DLI_X_FULL_THREADS, 0, L.st>>>(v, reinterpret_cast<const double2 *>(xs), ys, so, sto, L.queue,
                                                         L.qcount, L.qcap, n, L.stats);
  L.mark(2);
  L.launches += 2;
}
template <int D, bool STATS, int EPI>
void launchD(Launch &L, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto) {
  MeshDView v = viewD(L.m);
  int tpb = 128;
  int per_warp = L.m->opt.queries_per_warp > 0 ? L.m->opt.queries_per_warp : 512;
  while (per_warp > 32 && n / per_warp < 148 * 64) per_warp >>= 1;
  int64_t warps = (n + per_warp - 1) / per_warp;
  int wpb = tpb / 32;
  L.mark(0);
  if (EPI == 0 && L.m->d_fat) {
    const unsigned grid = (unsigned)((warps + wpb - 1) / wpb);
    const bool ring = L.m->opt.prefetch_ring >= 0;
    if constexpr (D == 4) {
      if (fat_pack4(L.m)) {
        using FP = FatD<4, true>;
        if (ring)
          k_interpD_fat<4, STATS, true, true><<<grid, tpb, (size_t)wpb * (FP::SMEM_PER_WARP + RingD<4>::BYTES), L.st>>>(
              v, n, per_warp, xs, ys, so, sto, L.queue, L.qcount, L.qcap, L.stats);
        else
          k_interpD_fat<4, STATS, false, true><<<grid, tpb, (size_t)wpb * FP::SMEM_PER_WARP, L.st>>>(
              v, n, per_warp, xs, ys, so, sto, L.queue, L.qcount, L.qcap, L.stats);
        L.mark(1);
        k_interpD_full<D, STATS, EPI><<<148 * DLI_X_FULLD_BLOCKS, 128, 0, L.st>>>(v, xs, ys, so, sto, L.queue, L.qcount, L.qcap, n, L.stats);
        L.mark(2);
        L.launches += 2;
        return;
      }
    }
    if (ring)
      k_interpD_fat<D, STATS, true><<<grid, tpb, (size_t)wpb * (FatD<D>::SMEM_PER_WARP + RingD<D>::BYTES), L.st>>>(
          v, n, per_warp, xs, ys, so, sto, L.queue, L.qcount, L.qcap, L.stats);
    else
      k_interpD_fat<D, STATS, false><<<grid, tpb, (size_t)wpb * FatD<D>::SMEM_PER_WARP, L.st>>>(
          v, n, per_warp, xs, ys, so, sto, L.queue, L.qcount, L.qcap, L.stats);
  }
  else
    k_interpD_fast<D, STATS, EPI><<<(unsigned)((warps + wpb - 1) / wpb), tpb, 0, L.st>>>(v, n, per_warp, xs, ys, so, sto, L.queue,
                                                                                        L.qcount, L.qcap, L.stats);
  L.mark(1);
  k_interpD_full<D, STATS, EPI><<<148 * DLI_X_FULLD_BLOCKS, 128, 0, L.st>>>(v, xs, ys, so, sto, L.queue, L.qcount, L.qcap, n, L.stats);
  L.mark(2);
  L.launches += 2;
}

template <bool STATS, int EPI>
void launch_any(Launch &L, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto) {
  switch (L.m->dim) {
    case 2: launch2<STATS, EPI>(L, n, xs, ys, so, sto); break;
    case 3: launchD<3, STATS, EPI>(L, n, xs, ys, so, sto); break;
    case 4: launchD<4, STATS, EPI>(L, n, xs, ys, so, sto); break;
  }
}

// all pointers are device pointers; enqueues on `st`.  out = 0: ys[n] values; out = 1: ys[n][2] gradients (2D)
int run_device(const dli_mesh *m, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto,
               cudaStream_t st, int *launches, int out = 0) {
  if (n == 0) return DLI_OK;
  if (m->ns == 0) {
    const int64_t cnt = out == 1 ? 2 * n : n;   // one thread per element: gradients are two doubles per query
    k_fill_nomesh<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(cnt, ys, nullptr, nullptr);
    if (so || sto) k_fill_nomesh<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, nullptr, so, sto);
    CU(cudaGetLastError());
    if (launches) *launches += 1;
    return DLI_OK;
  }
  // (A single-launch form for small batches -- the exact pass serving every query from its start hint, no queue, no
  // counter reset -- was measured and dropped: 36 us instead of 42 us per 1024-query call from Python, because the
  // per-call floor is host-side (ctypes call, pointer queries, stream synchronisation), not the two launches; and it
  // breaks the batch-partition invariance of the results, since the exact pass evaluates in the reference's operation
  // order where the fast pass uses the face's affine form.)
  const int64_t cap = queue_capacity(m, std::min(n, kChunk));
  int32_t *queue = nullptr;
  // [cap] query indices, [cap] faces the fast pass stopped in, then the counter
  CU(cudaMallocAsync(&queue, sizeof(int32_t) * 2 * (size_t)cap + 16, st));
  unsigned int *qcount = reinterpret_cast<unsigned int *>(queue + 2 * cap);
  // (queue + 2 * cap is 4-byte aligned, which is all an unsigned int needs)
  Launch L;
  L.m = m; L.st = st; L.queue = queue; L.qcount = qcount; L.qcap = (unsigned int)cap; L.stats = m->d_stats;
  const int d = m->dim;
  for (int64_t off = 0; off < n; off += kChunk) {
    int64_t cn = std::min(kChunk, n - off);
    CU(cudaMemsetAsync(qcount, 0, sizeof(unsigned int), st));
    const double *cx = xs + off * d;
    double *cy = ys ? ys + off * (out == 1 ? 2 : 1) : nullptr;
    int32_t *cs = so ? so + off : nullptr;
    uint8_t *ct = sto ? sto + off : nullptr;
    if (out == 1) {
      if (m->collect_stats) launch2<true, 0, 1>(L, cn, cx, cy, cs, ct);
      else launch2<false, 0, 1>(L, cn, cx, cy, cs, ct);
    } else if (m->collect_stats) {
      if (m->opt.epilogue != DLI_EPILOGUE_CLOSED_FORM) launch_any<true, 0>(L, cn, cx, cy, cs, ct);
      else launch_any<true, 1>(L, cn, cx, cy, cs, ct);
    } else {
      if (m->opt.epilogue != DLI_EPILOGUE_CLOSED_FORM) launch_any<false, 0>(L, cn, cx, cy, cs, ct);
      else launch_any<false, 1>(L, cn, cx, cy, cs, ct);
    }
    CU(cudaGetLastError());
  }
  CU(cudaFreeAsync(queue, st));
  if (launches) *launches += L.launches;
  return DLI_OK;
}

bool is_device_ptr(const void *p, bool *known) {
  cudaPointerAttributes a;
  *known = true;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); *known = false; return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Host copy between the caller's pageable buffer and a pinned staging buffer.  The destination is not read again by
// this core (the DMA engine or the caller reads it later), so it is written with non-temporal stores: no
// read-for-ownership of the destination lines (a third less host-memory traffic, which is the bound of this path).
#if defined(__SSE2__)
static void stage_copy(void *dst, const void *src, size_t bytes) {
  unsigned char *d = static_cast<unsigned char *>(dst);
  const unsigned char *s = static_cast<const unsigned char *>(src);
  size_t head = (16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15;
  if (head > bytes) head = bytes;
  memcpy(d, s, head);
  d += head; s += head; bytes -= head;
  for (size_t i = bytes / 64; i > 0; i--, s += 64, d += 64) {
    const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s));
    const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + 16));
    const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + 32));
    const __m128i e = _mm_loadu_si128(reinterpret_cast<const __m128i *>(s + 48));
    _mm_stream_si128(reinterpret_cast<__m128i *>(d), a);
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + 16), b);
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + 32), c);
    _mm_stream_si128(reinterpret_cast<__m128i *>(d + 48), e);
  }
  _mm_sfence();
  memcpy(d, s, bytes & 63);
}
#else
static void stage_copy(void *dst, const void *src, size_t bytes) { memcpy(dst, src, bytes); }
#endif
static bool is_pageable(const void *p) {
  if (!p) return false;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
  return a.type == cudaMemoryTypeUnregistered;
}
static int stage_threads() {
  static const int t = [] {
    int v = 0;
    if (const char *e = getenv("DLI_HOST_COPY_THREADS")) v = atoi(e);
    if (v <= 0) v = std::min(8, (int)std::thread::hardware_concurrency() / 2);
    return std::max(1, std::min(kMaxStage, v));
  }();
  return t;
}
// pageable host buffers: see StagePipe.  Called with m->pipe_mu held.
int run_host_staged(const dli_mesh *m, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto, int *launches,
                    int out) {
  const int d = m->dim, yw = out == 1 ? 2 : 1;
  StagePipe &P = m->stage;
  // statistics / per-launch timing are accumulated per call: keep those calls on one worker
  int T = (m->collect_stats || m->collect_timing) ? 1 : stage_threads();
  int64_t chunk = (n + 2 * T - 1) / (2 * T);
  chunk = std::min<int64_t>(kStageChunk, std::max<int64_t>(kStageMinChunk, (chunk + 4095) & ~(int64_t)4095));
  T = (int)std::min<int64_t>(T, (n + chunk - 1) / chunk);
  const int64_t cap = kStageChunk;       // slot buffers are sized for the largest chunk
  for (int i = 0; i < T; i++) {          // slots are created on first use and kept
    StageSlot &S = P.slot[i];
    cudaError_t e = cudaSuccess;
    if (!S.st) e = cudaStreamCreateWithFlags(&S.st, cudaStreamNonBlocking);
    if (e == cudaSuccess && !S.dx) e = cudaMalloc(&S.dx, sizeof(double) * (size_t)cap * d);
    if (e == cudaSuccess && !S.hx) e = cudaMallocHost(&S.hx, sizeof(double) * (size_t)cap * d);
    if (e == cudaSuccess && ys && !S.dy) e = cudaMalloc(&S.dy, sizeof(double) * (size_t)cap * 2);   // room for gradients
    if (e == cudaSuccess && ys && !S.hy) e = cudaMallocHost(&S.hy, sizeof(double) * (size_t)cap * 2);
    if (e == cudaSuccess && so && !S.ds) e = cudaMalloc(&S.ds, sizeof(int32_t) * (size_t)cap);
    if (e == cudaSuccess && so && !S.hs) e = cudaMallocHost(&S.hs, sizeof(int32_t) * (size_t)cap);
    if (e == cudaSuccess && sto && !S.dt) e = cudaMalloc(&S.dt, (size_t)cap);
    if (e == cudaSuccess && sto && !S.ht) e = cudaMallocHost(&S.ht, (size_t)cap);
    if (e != cudaSuccess) {
      stage_free(P);
      return fail(DLI_ERR_CUDA, std::string("dli_interp: staging allocation failed: ") + cudaGetErrorString(e));
    }
    P.nslots = std::max(P.nslots, i + 1);
  }
  std::atomic<int64_t> next{0};
  std::atomic<int> rc_all{DLI_OK}, launch_all{0};
  std::mutex err_mu;
  std::string err_msg;
  auto work = [&](int w) {
    StageSlot &S = P.slot[w];
    int rc = DLI_OK, nl = 0;
    if (cudaSetDevice(m->device) != cudaSuccess) rc = fail(DLI_ERR_CUDA, "dli_interp: cannot select device");
    while (rc == DLI_OK && rc_all.load(std::memory_order_relaxed) == DLI_OK) {
      const int64_t off = next.fetch_add(1) * chunk;
      if (off >= n) break;
      const int64_t cn = std::min(chunk, n - off);
      stage_copy(S.hx, xs + off * d, sizeof(double) * (size_t)cn * d);
      cudaError_t e = cudaMemcpyAsync(S.dx, S.hx, sizeof(double) * (size_t)cn * d, cudaMemcpyHostToDevice, S.st);
      if (e != cudaSuccess) { rc = fail(DLI_ERR_CUDA, std::string("H2D copy failed: ") + cudaGetErrorString(e)); break; }
      rc = run_device(m, cn, S.dx, ys ? S.dy : nullptr, so ? S.ds : nullptr, sto ? S.dt : nullptr, S.st, &nl, out);
      if (rc != DLI_OK) break;
      if (ys) e = cudaMemcpyAsync(S.hy, S.dy, sizeof(double) * (size_t)cn * yw, cudaMemcpyDeviceToHost, S.st);
      if (e == cudaSuccess && so) e = cudaMemcpyAsync(S.hs, S.ds, sizeof(int32_t) * (size_t)cn, cudaMemcpyDeviceToHost, S.st);
      if (e == cudaSuccess && sto) e = cudaMemcpyAsync(S.ht, S.dt, (size_t)cn, cudaMemcpyDeviceToHost, S.st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(S.st);
      if (e != cudaSuccess) { rc = fail(DLI_ERR_CUDA, std::string("dli_interp: ") + cudaGetErrorString(e)); break; }
      if (ys) stage_copy(ys + off * yw, S.hy, sizeof(double) * (size_t)cn * yw);
      if (so) stage_copy(so + off, S.hs, sizeof(int32_t) * (size_t)cn);
      if (sto) stage_copy(sto + off, S.ht, (size_t)cn);
    }
    launch_all.fetch_add(nl);
    if (rc != DLI_OK) {
      std::lock_guard<std::mutex> lk(err_mu);
      if (rc_all.load() == DLI_OK) { rc_all.store(rc); err_msg = g_err; }
    }
  };
  std::vector<std::thread> pool;
  try {
    for (int w = 1; w < T; w++) pool.emplace_back(work, w);
  } catch (...) {}                       // fewer helpers than planned: the remaining workers take all chunks
  work(0);
  for (std::thread &t : pool) t.join();
  if (launches) *launches += launch_all.load();
  return rc_all.load() == DLI_OK ? DLI_OK : fail(rc_all.load(), err_msg);
}

// host-pointer path: chunked H2D / compute / D2H pipeline over a ring of device slots.  The ring (streams +
// staging buffers + straggler queues) lives in the mesh and is created once: a call costs no cudaMalloc /
// cudaFree / stream creation.  H2D of chunk k+1, the kernels of chunk k and the D2H of chunk k-1 run on
// different streams, so both copy engines stay busy (PCIe is the bound: 24 B per 2D query).
int run_host(const dli_mesh *m, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto, int *launches,
             int out = 0) {
  if (n == 0) return DLI_OK;
  const int d = m->dim;
  std::lock_guard<std::mutex> lk(m->pipe_mu);  // one host-path call at a time per mesh
  if (n >= kStageMinQueries && (is_pageable(xs) || is_pageable(ys) || is_pageable(so) || is_pageable(sto)))
    return run_host_staged(m, n, xs, ys, so, sto, launches, out);
  HostPipe &P = m->pipe;
  const int64_t chunk = kHostChunk;
  if (!P.ready) {
    for (int i = 0; i < HostPipe::kSlots; i++) {
      cudaError_t e = cudaStreamCreateWithFlags(&P.st[i], cudaStreamNonBlocking);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&P.done[i], cudaEventDisableTiming);
      if (e == cudaSuccess) e = cudaMalloc(&P.x[i], sizeof(double) * (size_t)chunk * d);
      if (e == cudaSuccess) e = cudaMalloc(&P.y[i], sizeof(double) * (size_t)chunk * (d == 2 ? 2 : 1));  // 2D: room for gradients
      if (e == cudaSuccess) e = cudaMalloc(&P.s[i], sizeof(int32_t) * (size_t)chunk);
      if (e == cudaSuccess) e = cudaMalloc(&P.t[i], (size_t)chunk);
      if (e != cudaSuccess) {
        pipe_free(P);
        return fail(DLI_ERR_CUDA, std::string("dli_interp: staging allocation failed: ") + cudaGetErrorString(e));
      }
    }
    P.ready = true;
  }
  int rc = DLI_OK;
  int64_t c = 0;
  for (int64_t off = 0; off < n && rc == DLI_OK; off += chunk, c++) {
    const int i = (int)(c % HostPipe::kSlots);
    cudaStream_t st = P.st[i];
    int64_t cn = std::min(chunk, n - off);
    cudaError_t e = cudaMemcpyAsync(P.x[i], xs + off * d, sizeof(double) * (size_t)cn * d, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) { rc = fail(DLI_ERR_CUDA, std::string("H2D copy failed: ") + cudaGetErrorString(e)); break; }
    rc = run_device(m, cn, P.x[i], ys ? P.y[i] : nullptr, so ? P.s[i] : nullptr, sto ? P.t[i] : nullptr, st, launches, out);
    if (rc != DLI_OK) break;
    const int yw = out == 1 ? 2 : 1;
    if (ys) e = cudaMemcpyAsync(ys + off * yw, P.y[i], sizeof(double) * (size_t)cn * yw, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && so) e = cudaMemcpyAsync(so + off, P.s[i], sizeof(int32_t) * (size_t)cn, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && sto) e = cudaMemcpyAsync(sto + off, P.t[i], (size_t)cn, cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess) rc = fail(DLI_ERR_CUDA, std::string("D2H copy failed: ") + cudaGetErrorString(e));
  }
  for (int i = 0; i < HostPipe::kSlots; i++) {
    cudaError_t e = cudaStreamSynchronize(P.st[i]);
    if (e != cudaSuccess && rc == DLI_OK) rc = fail(DLI_ERR_CUDA, std::string("dli_interp: ") + cudaGetErrorString(e));
  }
  return rc;
}

int interp_common(const dli_mesh *m, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto, void *stream,
                  const char *who, int out = 0) {
  if (!m) return fail(DLI_ERR_ARG, std::string(who) + ": NULL mesh");
  if (n < 0) return fail(DLI_ERR_ARG, std::string(who) + ": negative n");
  if (n == 0) return DLI_OK;
  if (!xs) return fail(DLI_ERR_ARG, std::string(who) + ": xs is NULL");
  DeviceGuard guard(m->device);
  if (!guard.ok) return fail(DLI_ERR_CUDA, std::string(who) + ": cannot select device");
  bool known;
  bool dev = is_device_ptr(xs, &known);
  const void *outs[3] = {ys, so, sto};
  for (const void *p : outs)
    if (p) {
      bool k2;
      if (is_device_ptr(p, &k2) != dev) return fail(DLI_ERR_ARG, std::string(who) + ": xs and outputs must all be host or all be device pointers");
    }
  if (dev && m->dim == 2 && (((uintptr_t)xs & 15) || (out == 1 && ((uintptr_t)ys & 15))))
    return fail(DLI_ERR_ARG, std::string(who) + ": 2D device arrays of coordinate pairs must be 16-byte aligned");
  if (dev && (((uintptr_t)xs & 7) || ((uintptr_t)ys & 7) || ((uintptr_t)so & 3)))
    return fail(DLI_ERR_ARG, std::string(who) + ": misaligned device pointer");
  int launches = 0;
  int rc;
  if (m->collect_stats) {     // counters cover the whole call, however many chunks the host paths cut it into
    if (dev) CU(cudaMemsetAsync(m->d_stats, 0, sizeof(unsigned long long) * ST_WORDS, (cudaStream_t)stream));
    else { CU(cudaMemset(m->d_stats, 0, sizeof(unsigned long long) * ST_WORDS)); CU(cudaDeviceSynchronize()); }
  }
  if (dev) rc = run_device(m, n, xs, ys, so, sto, (cudaStream_t)stream, &launches, out);
  else rc = run_host(m, n, xs, ys, so, sto, &launches, out);
  m->last_launches.store(launches, std::memory_order_relaxed);
  m->last_queries.store(n, std::memory_order_relaxed);
  return rc;
}

}  // namespace

extern "C" int dli_interp(const dli_mesh *m, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto,
                          void *stream) {
  if (n > 0 && !ys) return fail(DLI_ERR_ARG, "dli_interp: ys is NULL");
  return interp_common(m, n, xs, ys, so, sto, stream, "dli_interp");
}
extern "C" int dli_locate(const dli_mesh *m, int64_t n, const double *xs, int32_t *so, uint8_t *sto, void *stream) {
  if (n > 0 && !so) return fail(DLI_ERR_ARG, "dli_locate: simplex_out is NULL");
  return interp_common(m, n, xs, nullptr, so, sto, stream, "dli_locate");
}

extern "C" int dli_gradient(const dli_mesh *m, int64_t n, const double *xs, double *grads, int32_t *so, uint8_t *sto,
                            void *stream) {
  if (m && m->dim != 2) return fail(DLI_ERR_UNSUPPORTED, "dli_gradient: 2D meshes only (the reference's dD class has no gradient)");
  if (n > 0 && !grads) return fail(DLI_ERR_ARG, "dli_gradient: grads is NULL");
  return interp_common(m, n, xs, grads, so, sto, stream, "dli_gradient", 1);
}

// ---------------------------------------------------------------------------------------------
// dli_interp_noalloc: batched interp for callers that must not allocate device memory.
//
// dli_interp takes its straggler queue (the hand-over list between the fast walk kernel and the exact-arithmetic
// kernel: two int32 per queued query plus a counter) from the stream-ordered pool on every call.  Some callers
// cannot have that: a host program that handed all of the GPU's memory to its own pool at start-up and left the
// driver nothing to give out; a solver inside a section where an allocation may take a driver lock or synchronise
// the device and stall its other streams; a driver configuration without stream-ordered pools.  For them the queue
// lives in the array below, part of the library image: every GPU that loads libdli.so holds these 8 MiB + 256 B
// until the process exits, whether or not the entry point is ever called.
//
// g_noalloc_scratch is a device array.  In host code its name stands for a host-side shadow object, not for memory the
// GPU can use, so the host never passes the name to a kernel: noalloc_scratch() asks the runtime for the array's
// address on the current device (cudaGetSymbolAddress) once per device and hands that out.
// ---------------------------------------------------------------------------------------------
static_assert(DLI_NOALLOC_MAX_QUERIES > 0 && DLI_NOALLOC_MAX_QUERIES <= (1ll << 22), "keep the resident scratch modest");
constexpr size_t kNoAllocHead = 256;   // the counter has the first 256 bytes to itself
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocHead + sizeof(int32_t) * 2 * (size_t)DLI_NOALLOC_MAX_QUERIES];

constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static unsigned char *g_noalloc_addr[kNoAllocMaxDevices];   // per device ordinal; nullptr until the set-up below ran

// device `device` must be current
static int noalloc_scratch(int device, unsigned char **out) {
  if (device < 0 || device >= kNoAllocMaxDevices)
    return fail(DLI_ERR_UNSUPPORTED, "dli_interp_noalloc: device ordinal beyond the 64 the address table holds");
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_addr[device]) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(DLI_ERR_CUDA, std::string("dli_interp_noalloc: cudaGetSymbolAddress(scratch) failed: ") + cudaGetErrorString(e));
    }
    g_noalloc_addr[device] = static_cast<unsigned char *>(p);
  }
  *out = g_noalloc_addr[device];
  return DLI_OK;
}

extern "C" int dli_interp_noalloc(const dli_mesh *m, int64_t n, const double *xs, double *ys, int32_t *so, uint8_t *sto,
                                  void *stream) {
  const char *who = "dli_interp_noalloc";
  if (!m) return fail(DLI_ERR_ARG, std::string(who) + ": NULL mesh");
  if (n < 0) return fail(DLI_ERR_ARG, std::string(who) + ": negative n");
  if (n > DLI_NOALLOC_MAX_QUERIES) {
    char b[200];
    snprintf(b, sizeof b, "%s: n = %lld exceeds DLI_NOALLOC_MAX_QUERIES = %lld (the resident scratch holds no more; split the batch or call dli_interp)",
             who, (long long)n, (long long)DLI_NOALLOC_MAX_QUERIES);
    return fail(DLI_ERR_ARG, b);
  }
  if (n == 0) return DLI_OK;
  if (!xs || !ys) return fail(DLI_ERR_ARG, std::string(who) + ": xs / ys is NULL");
  if (m->collect_timing)
    return fail(DLI_ERR_UNSUPPORTED, std::string(who) + ": per-launch timing creates CUDA events; switch it off (dli_mesh_collect_stats(mesh, 0 or 1))");
  DeviceGuard guard(m->device);
  if (!guard.ok) return fail(DLI_ERR_CUDA, std::string(who) + ": cannot select device");
  const void *ptrs[4] = {xs, ys, so, sto};
  for (const void *p : ptrs)
    if (p) {
      bool known;
      if (!is_device_ptr(p, &known)) return fail(DLI_ERR_ARG, std::string(who) + ": device pointers only (the host-pointer pipeline of dli_interp owns staging buffers)");
    }
  if (m->dim == 2 && ((uintptr_t)xs & 15)) return fail(DLI_ERR_ARG, std::string(who) + ": 2D device arrays of coordinate pairs must be 16-byte aligned");
  if (((uintptr_t)xs & 7) || ((uintptr_t)ys & 7) || ((uintptr_t)so & 3)) return fail(DLI_ERR_ARG, std::string(who) + ": misaligned device pointer");
  cudaStream_t st = (cudaStream_t)stream;
  int launches = 0;
  if (m->ns == 0) {
    k_fill_nomesh<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, ys, so, sto);
    CU(cudaGetLastError());
    launches = 1;
  } else {
    unsigned char *scratch = nullptr;
    const int rc = noalloc_scratch(m->device, &scratch);
    if (rc != DLI_OK) return rc;
    Launch L;
    L.m = m; L.st = st; L.stats = m->d_stats;
    L.qcount = reinterpret_cast<unsigned int *>(scratch);
    L.queue = reinterpret_cast<int32_t *>(scratch + kNoAllocHead);
    L.qcap = (unsigned int)n;   // room for every query of the batch: the queue cannot overflow
    if (m->collect_stats) CU(cudaMemsetAsync(m->d_stats, 0, sizeof(unsigned long long) * ST_WORDS, st));
    // the array keeps what the previous call left in it: the counter must start at zero, the slots are written
    // before they are read
    CU(cudaMemsetAsync(L.qcount, 0, sizeof(unsigned int), st));
    if (m->collect_stats) {
      if (m->opt.epilogue != DLI_EPILOGUE_CLOSED_FORM) launch_any<true, 0>(L, n, xs, ys, so, sto);
      else launch_any<true, 1>(L, n, xs, ys, so, sto);
    } else {
      if (m->opt.epilogue != DLI_EPILOGUE_CLOSED_FORM) launch_any<false, 0>(L, n, xs, ys, so, sto);
      else launch_any<false, 1>(L, n, xs, ys, so, sto);
    }
    CU(cudaGetLastError());
    launches = L.launches;
  }
  m->last_launches.store(launches, std::memory_order_relaxed);
  m->last_queries.store(n, std::memory_order_relaxed);
  return DLI_OK;
}

extern "C" int dli_mesh_collect_stats(dli_mesh *m, int enable) {
  if (!m) return fail(DLI_ERR_ARG, "NULL mesh");
  m->collect_stats = (enable & 1) != 0;
  m->collect_timing = (enable & 2) != 0;
  {
    std::lock_guard<std::mutex> lk(m->ev_mu);
    for (cudaEvent_t e : m->ev) cudaEventDestroy(e);
    m->ev.clear();
  }
  return DLI_OK;
}
extern "C" int dli_mesh_get_stats(const dli_mesh *m, dli_stats *out) {
  if (!m || !out) return fail(DLI_ERR_ARG, "NULL argument");
  DeviceGuard guard(m->device);
  unsigned long long w[ST_WORDS];
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(w, m->d_stats, sizeof w, cudaMemcpyDeviceToHost));
  out->queries = m->last_queries.load(std::memory_order_relaxed);
  out->walk_steps = (int64_t)w[ST_STEPS];
  out->max_walk_steps = (int64_t)w[ST_MAXSTEPS];
  out->exact_fallbacks = (int64_t)w[ST_EXACT];
  out->outside = (int64_t)w[ST_OUTSIDE];
  out->kernel_launches = m->last_launches.load(std::memory_order_relaxed);
  out->straggler_queries = (int64_t)w[ST_HARD];
  out->slow_simplices = m->n_slow_faces;
  out->fast_kernel_launches = 0;
  out->fast_kernel_ms = out->full_kernel_ms = 0.0;
  {
    std::lock_guard<std::mutex> lk(m->ev_mu);
    for (size_t i = 0; i + 2 < m->ev.size(); i += 3) {
      float a = 0.f, b = 0.f;
      if (cudaEventElapsedTime(&a, m->ev[i], m->ev[i + 1]) == cudaSuccess &&
          cudaEventElapsedTime(&b, m->ev[i + 1], m->ev[i + 2]) == cudaSuccess) {
        out->fast_kernel_ms += a;
        out->full_kernel_ms += b;
        out->fast_kernel_launches++;
      }
    }
  }
  return DLI_OK;
}

// ---------------------------------------------------------------------------------------------
// plot data (2D)
// ---------------------------------------------------------------------------------------------
extern "C" int64_t dli_mesh_num_edges(const dli_mesh *m) {
  if (!m) return fail(DLI_ERR_ARG, "dli_mesh_num_edges: NULL mesh");
  if (m->dim != 2) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_num_edges: 2D meshes only");
  return (3 * m->ns + m->nh) / 2;   // every interior edge is shared by two faces, every hull edge by one
}

extern "C" int dli_mesh_line_segments(const dli_mesh *m, double *segments, int32_t *concave, double *min_coords,
                                      double *max_coords) {
  DLI_GUARD_BEGIN
  if (!m) return fail(DLI_ERR_ARG, "dli_mesh_line_segments: NULL mesh");
  if (m->dim != 2) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_line_segments: 2D meshes only (as is_edge_concave)");
  DeviceGuard guard(m->device);
  if (!guard.ok) return fail(DLI_ERR_CUDA, "dli_mesh_line_segments: cannot select device");
  const int64_t ne = (3 * m->ns + m->nh) / 2;
  const int ns = (int)m->ns;
  if (ne == 0) {   // no finite edge: nothing visited (:845-846)
    for (int k = 0; k < 3; k++) { if (min_coords) min_coords[k] = DBL_MAX; if (max_coords) max_coords[k] = -DBL_MAX; }
    return DLI_OK;
  }
  int *d_cnt = nullptr, *d_off = nullptr;
  double *d_seg = nullptr;
  int32_t *d_cc = nullptr;
  void *d_tmp = nullptr;
  size_t tmp_bytes = 0;
  cudaError_t e = cudaMalloc(&d_cnt, sizeof(int) * (size_t)ns);
  if (e == cudaSuccess) e = cudaMalloc(&d_off, sizeof(int) * (size_t)ns);
  if (e == cudaSuccess) e = cudaMalloc(&d_seg, sizeof(double) * 6 * (size_t)ne);
  if (e == cudaSuccess) e = cudaMalloc(&d_cc, sizeof(int32_t) * (size_t)ne);
  if (e == cudaSuccess) e = cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_cnt, d_off, ns);
  if (e == cudaSuccess) e = cudaMalloc(&d_tmp, std::max<size_t>(tmp_bytes, 16));
  if (e == cudaSuccess) {
    const unsigned gb = (unsigned)((ns + 127) / 128);
    const int4 *frec = reinterpret_cast<const int4 *>(m->d_srec);
    k_edge_count<<<gb, 128>>>(ns, frec, m->d_orig, d_cnt);
    e = cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_cnt, d_off, ns);
    if (e == cudaSuccess) {
      k_edge_emit<<<gb, 128>>>(ns, frec, reinterpret_cast<const double4 *>(m->d_vrec), m->d_orig, d_off, d_seg, d_cc);
      e = cudaGetLastError();
    }
  }
  if (e == cudaSuccess && (min_coords || max_coords)) {
    const int nb = 296;
    double *d_part = nullptr;
    std::vector<double> part((size_t)nb * 6);
    e = cudaMalloc(&d_part, sizeof(double) * part.size());
    if (e == cudaSuccess) {
      k_point_minmax<<<nb, 256>>>(2 * ne, d_seg, d_part);
      e = cudaMemcpy(part.data(), d_part, sizeof(double) * part.size(), cudaMemcpyDeviceToHost);
    }
    cudaFree(d_part);
    for (int k = 0; k < 3 && e == cudaSuccess; k++) {
      double lo = DBL_MAX, hi = -DBL_MAX;     // the reference's initial values (:845-846)
      for (int b = 0; b < nb; b++) { lo = std::min(lo, part[(size_t)b * 6 + k]); hi = std::max(hi, part[(size_t)b * 6 + 3 + k]); }
      if (min_coords) min_coords[k] = lo;
      if (max_coords) max_coords[k] = hi;
    }
  }
  if (e == cudaSuccess && segments) e = cudaMemcpy(segments, d_seg, sizeof(double) * 6 * (size_t)ne, cudaMemcpyDefault);
  if (e == cudaSuccess && concave) e = cudaMemcpy(concave, d_cc, sizeof(int32_t) * (size_t)ne, cudaMemcpyDefault);
  cudaFree(d_cnt); cudaFree(d_off); cudaFree(d_seg); cudaFree(d_cc); cudaFree(d_tmp);
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_mesh_line_segments: ") + cudaGetErrorString(e));
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_line_segments")
}

// ---------------------------------------------------------------------------------------------
// error evaluator
// ---------------------------------------------------------------------------------------------
static int ensure_items(dli_mesh *m) {
  std::lock_guard<std::mutex> lk(m->mu);
  if (m->items_ready) return DLI_OK;
  if (m->h_sv.empty() && m->ns > 0) return fail(DLI_ERR_UNSUPPORTED, "error evaluator needs a mesh created with dli_mesh_create");
  const int d = m->dim, k = d + 1;
  m->h_item_simplex.clear();
  m->h_item_verts.clear();
  for (int64_t s = 0; s < m->ns; s++) {
    for (int item = -1; item < k; item++) {
      int32_t fv[DLI_MAX_DIM + 1];
      int nf = 0;
      if (item == -1) {
        for (int j = 0; j < k; j++) fv[nf++] = m->h_sv[s * k + j];
      } else {
        if (m->h_sn[s * k + item] >= 0) continue;
        for (int j = 0; j < k; j++)
          if (j != item) fv[nf++] = m->h_sv[s * k + j];
      }
      int lo = (item == -1) ? (1 << nf) - 1 : 0;
      for (int mask = lo; mask < (1 << nf); mask++) {
        if (__builtin_popcount((unsigned)mask) < 2) continue;
        int c = 0;
        for (int b = 0; b < nf; b++)
          if (mask >> b & 1) { m->h_item_verts.push_back(fv[b]); c++; }
        for (; c < k; c++) m->h_item_verts.push_back(-1);
        m->h_item_simplex.push_back((int32_t)s);
      }
    }
  }
  m->n_items = (int64_t)m->h_item_simplex.size();
  DeviceGuard guard(m->device);
  CU(cudaMalloc(&m->d_item_verts, sizeof(int32_t) * std::max<size_t>(m->h_item_verts.size(), 1)));
  CU(cudaMemcpy(m->d_item_verts, m->h_item_verts.data(), sizeof(int32_t) * m->h_item_verts.size(), cudaMemcpyHostToDevice));
  m->items_ready = true;
  return DLI_OK;
}

extern "C" int64_t dli_simplex_error_count(const dli_mesh *m) {
  if (!m) return fail(DLI_ERR_ARG, "NULL mesh");
  int rc = ensure_items(const_cast<dli_mesh *>(m));
  return rc != DLI_OK ? rc : m->n_items;
}

extern "C" int dli_simplex_error_prepare(const dli_mesh *mc, double *centroids, int32_t *item_simplex,
                                         int32_t *item_vertices) {
  DLI_GUARD_BEGIN
  dli_mesh *m = const_cast<dli_mesh *>(mc);
  if (!m || !centroids) return fail(DLI_ERR_ARG, "dli_simplex_error_prepare: NULL argument");
  int rc = ensure_items(m);
  if (rc != DLI_OK) return rc;
  const int64_t n = m->n_items;
  if (item_simplex) memcpy(item_simplex, m->h_item_simplex.data(), sizeof(int32_t) * (size_t)n);
  if (item_vertices) memcpy(item_vertices, m->h_item_verts.data(), sizeof(int32_t) * m->h_item_verts.size());
  if (n == 0) return DLI_OK;
  DeviceGuard guard(m->device);
  bool known;
  const bool on_device = is_device_ptr(centroids, &known);
  double *d_c = centroids;   // device pointer: the kernel writes the caller's buffer (default stream, no copy, no allocation)
  if (!on_device) CU(cudaMalloc(&d_c, sizeof(double) * (size_t)n * m->dim));
  unsigned blocks = (unsigned)((n + 255) / 256);
  switch (m->dim) {
    case 2: k_error_prepare<2><<<blocks, 256>>>(n, m->d_item_verts, m->d_vrec, m->vs, d_c); break;
    case 3: k_error_prepare<3><<<blocks, 256>>>(n, m->d_item_verts, m->d_vrec, m->vs, d_c); break;
    case 4: k_error_prepare<4><<<blocks, 256>>>(n, m->d_item_verts, m->d_vrec, m->vs, d_c); break;
  }
  cudaError_t e = cudaGetLastError();
  if (!on_device) {
    if (e == cudaSuccess) e = cudaMemcpy(centroids, d_c, sizeof(double) * (size_t)n * m->dim, cudaMemcpyDeviceToHost);
    cudaFree(d_c);
  }
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_simplex_error_prepare: ") + cudaGetErrorString(e));
  return DLI_OK;
  DLI_GUARD_END("dli_simplex_error_prepare")
}

extern "C" int dli_simplex_error_finish(const dli_mesh *mc, int64_t n, const double *phi, double *err) {
  DLI_GUARD_BEGIN
  dli_mesh *m = const_cast<dli_mesh *>(mc);
  if (!m || (n > 0 && (!phi || !err))) return fail(DLI_ERR_ARG, "dli_simplex_error_finish: NULL argument");
  int rc = ensure_items(m);
  if (rc != DLI_OK) return rc;
  if (n != m->n_items) return fail(DLI_ERR_ARG, "dli_simplex_error_finish: n does not match dli_simplex_error_count");
  if (n == 0) return DLI_OK;
  DeviceGuard guard(m->device);
  bool known;
  const bool on_device = is_device_ptr(phi, &known);
  if (is_device_ptr(err, &known) != on_device)
    return fail(DLI_ERR_ARG, "dli_simplex_error_finish: f_at_centroids and err must both be host or both be device pointers");
  if (on_device) {   // the caller's f lives on the GPU: one kernel on the default stream, no staging, no allocation
    unsigned blocks = (unsigned)((n + 127) / 128);
    switch (m->dim) {
      case 2: k_error_finish<2><<<blocks, 128>>>(n, m->d_item_verts, m->d_vrec, m->vs, phi, err); break;
      case 3: k_error_finish<3><<<blocks, 128>>>(n, m->d_item_verts, m->d_vrec, m->vs, phi, err); break;
      case 4: k_error_finish<4><<<blocks, 128>>>(n, m->d_item_verts, m->d_vrec, m->vs, phi, err); break;
    }
    CU(cudaGetLastError());
    return DLI_OK;
  }
  double *d_phi = nullptr, *d_err = nullptr;
  CU(cudaMalloc(&d_phi, sizeof(double) * (size_t)n));
  cudaError_t e = cudaMalloc(&d_err, sizeof(double) * (size_t)n);
  if (e == cudaSuccess) e = cudaMemcpy(d_phi, phi, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    unsigned blocks = (unsigned)((n + 127) / 128);
    switch (m->dim) {
      case 2: k_error_finish<2><<<blocks, 128>>>(n, m->d_item_verts, m->d_vrec, m->vs, d_phi, d_err); break;
      case 3: k_error_finish<3><<<blocks, 128>>>(n, m->d_item_verts, m->d_vrec, m->vs, d_phi, d_err); break;
      case 4: k_error_finish<4><<<blocks, 128>>>(n, m->d_item_verts, m->d_vrec, m->vs, d_phi, d_err); break;
    }
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(err, d_err, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost);
  cudaFree(d_phi);
  cudaFree(d_err);
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_simplex_error_finish: ") + cudaGetErrorString(e));
  return DLI_OK;
  DLI_GUARD_END("dli_simplex_error_finish")
}

// ---------------------------------------------------------------------------------------------
// replication support
// ---------------------------------------------------------------------------------------------
// header words: 0 version, 1 dim, 2 nv, 3 ns, 4 nh, 5 cells, 6..9 grid g[k], 10 max_steps, 11 fast_steps,
// 12 has packed faces, 13 has walk records, 14..15 reserved.  The floating-point grid parameters travel in buffer 0 (a small device block).
extern "C" int dli_mesh_header(const dli_mesh *m, int64_t h[DLI_MESH_HEADER_WORDS]) {
  if (!m || !h) return fail(DLI_ERR_ARG, "NULL argument");
  memset(h, 0, sizeof(int64_t) * DLI_MESH_HEADER_WORDS);
  h[0] = DLI_VERSION; h[1] = m->dim; h[2] = m->nv; h[3] = m->ns; h[4] = m->nh; h[5] = m->grid.cells;
  for (int k = 0; k < 4; k++) h[6 + k] = m->grid.g[k];
  h[10] = m->max_steps; h[11] = m->fast_steps;
  // optional buffers the replica must mirror, and how the walk records encode their neighbours:
  // word 13: 0 = no walk records, 1 = absolute neighbour ids (4D: split records), 2 = cell-indexed records,
  // 3 = delta-encoded neighbours, 4 = 4D packed one-line records, 5 = 4 + start copies, 6 = 1 + start copies (dD);
  // word 15 = records in the walk-record buffer
  h[12] = m->d_fpack ? 1 : 0; h[13] = m->d_fat ? (m->cellmode ? 2 : ((m->dim == 2 && fat_delta(m)) ? 3 : (fat_pack4(m) ? (m->startcopy ? 5 : 4) : (m->startcopy ? 6 : 1)))) : 0;
  h[15] = m->nslots;
  // word 12, bits 8-11: the options that change what a query returns (the walk-step cap is word 10).  A replica or a
  // loaded snapshot must answer like its source, so they travel with the mesh: bit 8 = the field is valid, bit 9 =
  // outside_nan, bits 10-11 = epilogue.  (Word 14 is free here; the snapshot file keeps its slow-face count there.)
  h[12] |= (1 << 8) | ((m->opt.outside_nan ? 1 : 0) << 9) | ((int64_t)(m->opt.epilogue & 3) << 10);
  // grid origin / inverse cell: bit patterns of doubles do not fit 16 words for 4D; pack o and ic for
  // up to 2 dims here and the rest in the reserved words -> instead expose them through buffer index
  // (see dli_mesh_buffer): index 0 = 96-byte parameter block.
  return DLI_OK;
}

struct GridParams {
  double o[DLI_MAX_DIM], cell[DLI_MAX_DIM], ic[DLI_MAX_DIM];
};

extern "C" int dli_mesh_create_from_header(const int64_t h[DLI_MESH_HEADER_WORDS], const dli_options *opt,
                                           dli_mesh **out) {
  if (!h || !out) return fail(DLI_ERR_ARG, "NULL argument");
  *out = nullptr;
  if (h[0] != DLI_VERSION) return fail(DLI_ERR_ARG, "dli_mesh_create_from_header: version mismatch");
  if (dli_device_count() <= 0) return fail(DLI_ERR_CUDA, "no CUDA device available");
  dli_mesh *m = new dli_mesh;
  m->dim = (int)h[1]; m->nv = h[2]; m->ns = h[3]; m->nh = (int32_t)h[4];
  if (m->dim < 2 || m->dim > 4) { delete m; return fail(DLI_ERR_UNSUPPORTED, "bad dim in header"); }
  resolve_options(opt, &m->opt);
  if (h[12] & (1 << 8)) {   // result-affecting options of the source win over this side's (older headers: field absent)
    m->opt.outside_nan = (int32_t)((h[12] >> 9) & 1);
    m->opt.epilogue = (int32_t)((h[12] >> 10) & 3);
  }
  if (m->opt.device >= 0) m->device = m->opt.device;
  else cudaGetDevice(&m->device);
  DeviceGuard guard(m->device);
  m->grid.cells = h[5];
  for (int k = 0; k < 4; k++) m->grid.g[k] = (int)h[6 + k];
  int rc = mesh_alloc_common(m);
  // mirror the source's optional buffers exactly (every rank must see the same buffer list)
  if (rc == DLI_OK && !(h[12] & 1) && m->d_fpack) { cudaFree(m->d_fpack); m->d_fpack = nullptr; }
  // (mesh_alloc_common sized d_fat from this side's options: drop it and allocate for the source's layout below)
  if (rc == DLI_OK && m->d_fat) { cudaFree(m->d_fat); m->d_fat = nullptr; }
  m->cellmode = h[13] == 2;
  m->fat_enc = h[13] == 3 ? 1 : ((h[13] == 4 || h[13] == 5) ? 4 : 0);
  m->startcopy = h[13] == 5 || h[13] == 6;
  m->nslots = (h[13] == 2 || m->startcopy) ? h[15] : m->ns;
  if (m->startcopy && m->nslots != m->ns + m->grid.cells) rc = fail(DLI_ERR_ARG, "header: start-copy record count does not match ns + cells");
  if (rc == DLI_OK && m->cellmode &&
      (cudaMalloc(&m->d_slot_face, sizeof(int32_t) * (size_t)std::max<int64_t>(m->nslots, 1)) != cudaSuccess ||
       cudaMalloc(&m->d_rowstart, sizeof(int) * ((size_t)m->grid.g[1] + 1)) != cudaSuccess))
    rc = fail(DLI_ERR_CUDA, "slot table allocation failed");
  if (rc == DLI_OK && (h[12] & 1) && !m->d_fpack && cudaMalloc(&m->d_fpack, sizeof(ulonglong2) * (size_t)std::max<int64_t>(m->ns, 1)) != cudaSuccess)
    rc = fail(DLI_ERR_CUDA, "packed face allocation failed");
  if (rc == DLI_OK && h[13] && !m->d_fat &&
      cudaMalloc(&m->d_fat, (size_t)fat_record_bytes(m) * (size_t)std::max<int64_t>(m->nslots, 1)) != cudaSuccess)
    rc = fail(DLI_ERR_CUDA, "walk record allocation failed");
  m->max_steps = (int)h[10];
  m->fast_steps = (int)h[11];
  if (rc != DLI_OK) { mesh_free(m); return rc; }
  cudaError_t e = cudaMalloc(&m->d_hint, sizeof(int32_t) * (size_t)std::max<int64_t>(m->grid.cells, 1));
  if (e != cudaSuccess) { mesh_free(m); return fail(DLI_ERR_CUDA, "hint allocation failed"); }
  *out = m;
  return DLI_OK;
}

extern "C" int dli_mesh_num_buffers(const dli_mesh *m) { return m ? 9 : 0; }

// buffers 0..5: vertex records, simplex records, hull table, hint grid, packed 2D face records (may be empty),
// 64-byte 2D walk records (may be empty), device-order -> caller simplex ids
extern "C" int dli_mesh_buffer(const dli_mesh *m, int index, void **ptr, int64_t *bytes) {
  if (!m || !ptr || !bytes) return fail(DLI_ERR_ARG, "NULL argument");
  switch (index) {
    case 0: *ptr = m->d_vrec; *bytes = (int64_t)sizeof(double) * std::max<int64_t>(m->nv, 1) * m->vs; break;
    case 1: *ptr = m->d_srec; *bytes = (int64_t)sizeof(int32_t) * std::max<int64_t>(m->ns, 1) * m->ss; break;
    case 2: *ptr = m->d_hull; *bytes = m->d_hull ? (int64_t)sizeof(int4) * std::max<int32_t>(m->nh, 1) : 0; break;
    case 3: *ptr = m->d_hint; *bytes = (int64_t)sizeof(int32_t) * m->grid.cells; break;
    case 4: *ptr = m->d_fpack; *bytes = m->d_fpack ? (int64_t)sizeof(ulonglong2) * m->ns : 0; break;
    case 5: *ptr = m->d_fat; *bytes = m->d_fat ? (int64_t)fat_record_bytes(m) * m->nslots : 0; break;
    case 6: *ptr = m->d_orig; *bytes = (int64_t)sizeof(int32_t) * std::max<int64_t>(m->ns, 1); break;
    case 7: *ptr = m->d_slot_face; *bytes = m->cellmode ? (int64_t)sizeof(int32_t) * m->nslots : 0; break;
    case 8: *ptr = m->d_rowstart; *bytes = m->cellmode ? (int64_t)sizeof(int) * (m->grid.g[1] + 1) : 0; break;
    default: return fail(DLI_ERR_ARG, "dli_mesh_buffer: bad index");
  }
  return DLI_OK;
}

// copy a whole mesh buffer out of / into the mesh; `ptr` may be a host or a device pointer
extern "C" int dli_mesh_buffer_read(const dli_mesh *m, int index, void *dst) {
  void *p; int64_t b;
  int rc = dli_mesh_buffer(m, index, &p, &b);
  if (rc != DLI_OK) return rc;
  if (b == 0) return DLI_OK;
  if (!dst) return fail(DLI_ERR_ARG, "dli_mesh_buffer_read: NULL dst");
  DeviceGuard guard(m->device);
  CU(cudaMemcpy(dst, p, (size_t)b, cudaMemcpyDefault));
  return DLI_OK;
}
extern "C" int dli_mesh_buffer_write(dli_mesh *m, int index, const void *src) {
  void *p; int64_t b;
  int rc = dli_mesh_buffer(m, index, &p, &b);
  if (rc != DLI_OK) return rc;
  if (b == 0) return DLI_OK;
  if (!src) return fail(DLI_ERR_ARG, "dli_mesh_buffer_write: NULL src");
  DeviceGuard guard(m->device);
  CU(cudaMemcpy(p, src, (size_t)b, cudaMemcpyDefault));
  return DLI_OK;
}

// grid parameters (doubles) for replication: host-side get/set
extern "C" int dli_mesh_get_grid_params(const dli_mesh *m, double *out /* [3 * DLI_MAX_DIM] */) {
  if (!m || !out) return fail(DLI_ERR_ARG, "NULL argument");
  for (int k = 0; k < DLI_MAX_DIM; k++) { out[k] = m->grid.o[k]; out[DLI_MAX_DIM + k] = m->grid.cell[k]; out[2 * DLI_MAX_DIM + k] = m->grid.ic[k]; }
  return DLI_OK;
}
extern "C" int dli_mesh_set_grid_params(dli_mesh *m, const double *in) {
  if (!m || !in) return fail(DLI_ERR_ARG, "NULL argument");
  for (int k = 0; k < DLI_MAX_DIM; k++) { m->grid.o[k] = in[k]; m->grid.cell[k] = in[DLI_MAX_DIM + k]; m->grid.ic[k] = in[2 * DLI_MAX_DIM + k]; }
  return DLI_OK;
}

// ---------------------------------------------------------------------------------------------
// dli_mesh_replicate: NCCL broadcast of the whole mesh.  NCCL is resolved at run time -- the
// process already holds the NCCL its communicator came from (the caller links it, or torch loaded its own copy), and
// a second copy linked into this library could not use that communicator.
// ---------------------------------------------------------------------------------------------
#include <dlfcn.h>
#include <nccl.h>   // types and enums only; no NCCL symbol is linked
namespace {
struct NcclApi {
  ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommUserRank)(const ncclComm_t, int *) = nullptr;
  ncclResult_t (*CommCuDevice)(const ncclComm_t, int *) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
  std::string why;
};
const NcclApi &nccl_api() {
  static const NcclApi api = [] {
    NcclApi a;
    void *h = nullptr;
    if (const char *e = getenv("DLI_NCCL_LIB")) h = dlopen(e, RTLD_NOW | RTLD_GLOBAL);
    auto sym = [&](const char *name) -> void * {
      void *p = h ? dlsym(h, name) : nullptr;
      if (!p) p = dlsym(RTLD_DEFAULT, name);   // the NCCL the process already has in its global scope
      return p;
    };
    if (!sym("ncclBroadcast")) {
      h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // loaded privately by someone else (torch): reuse that copy
      if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
      if (!h) h = dlopen("libnccl.so", RTLD_NOW);
    }
    a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(sym("ncclBroadcast"));
    a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
    a.CommUserRank = reinterpret_cast<decltype(a.CommUserRank)>(sym("ncclCommUserRank"));
    a.CommCuDevice = reinterpret_cast<decltype(a.CommCuDevice)>(sym("ncclCommCuDevice"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
    a.ok = a.Broadcast && a.CommUserRank && a.CommCuDevice;
    if (!a.ok) a.why = "NCCL not found (looked in the process, then libnccl.so.2; set DLI_NCCL_LIB)";
    return a;
  }();
  return api;
}
constexpr int kRepWords = DLI_MESH_HEADER_WORDS + 3 * DLI_MAX_DIM + 2;   // header, grid parameters (bit patterns), n_slow, ok flag
}  // namespace

extern "C" int dli_mesh_replicate(dli_mesh **inout, void *nccl_comm, int root) {
  DLI_GUARD_BEGIN
  if (!inout || !nccl_comm) return fail(DLI_ERR_ARG, "dli_mesh_replicate: NULL argument");
  const NcclApi &N = nccl_api();
  if (!N.ok) return fail(DLI_ERR_UNSUPPORTED, "dli_mesh_replicate: " + N.why);
  ncclComm_t comm = static_cast<ncclComm_t>(nccl_comm);
  auto nerr = [&](ncclResult_t r, const char *what) {
    return fail(DLI_ERR_CUDA, std::string("dli_mesh_replicate: ") + what + ": " + (N.GetErrorString ? N.GetErrorString(r) : "NCCL error"));
  };
  int rank = -1, dev = -1;
  ncclResult_t r = N.CommUserRank(comm, &rank);
  if (r == ncclSuccess) r = N.CommCuDevice(comm, &dev);
  if (r != ncclSuccess) return nerr(r, "communicator query");
  const bool is_root = rank == root;
  // A root that cannot send its mesh still takes part in the header broadcast, with an empty header (ok word 0): the
  // other ranks then return an error instead of waiting in ncclBroadcast for ever.
  const bool root_no_mesh = is_root && !*inout;
  const bool root_wrong_dev = is_root && *inout && (*inout)->device != dev;
  const bool root_bad = root_no_mesh || root_wrong_dev;
  DeviceGuard guard(dev);
  if (!guard.ok) return fail(DLI_ERR_CUDA, "dli_mesh_replicate: cannot select the communicator's device");
  cudaStream_t st = nullptr;
  int64_t *d_hdr = nullptr;
  CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  int rc = DLI_OK;
  int64_t w[kRepWords];
  memset(w, 0, sizeof w);
  cudaError_t e = cudaMalloc(&d_hdr, sizeof w);
  if (e == cudaSuccess && root_bad) e = cudaMemcpyAsync(d_hdr, w, sizeof w, cudaMemcpyHostToDevice, st);   // all zero
  if (e == cudaSuccess && is_root && !root_bad) {
    double gp[3 * DLI_MAX_DIM];
    dli_mesh_header(*inout, w);
    dli_mesh_get_grid_params(*inout, gp);
    memcpy(w + DLI_MESH_HEADER_WORDS, gp, sizeof gp);
    w[DLI_MESH_HEADER_WORDS + 3 * DLI_MAX_DIM] = (*inout)->n_slow_faces;
    w[kRepWords - 1] = 1;
    e = cudaMemcpyAsync(d_hdr, w, sizeof w, cudaMemcpyHostToDevice, st);
  }
  if (e == cudaSuccess) {
    r = N.Broadcast(d_hdr, d_hdr, kRepWords, ncclInt64, root, comm, st);
    if (r != ncclSuccess) rc = nerr(r, "header broadcast");
  }
  if (e == cudaSuccess && rc == DLI_OK) e = cudaMemcpyAsync(w, d_hdr, sizeof w, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess && rc == DLI_OK) e = cudaStreamSynchronize(st);
  dli_mesh *m = is_root ? *inout : nullptr;
  if (root_no_mesh) rc = fail(DLI_ERR_ARG, "dli_mesh_replicate: the root rank must pass its mesh");
  else if (root_wrong_dev) rc = fail(DLI_ERR_ARG, "dli_mesh_replicate: the root's mesh is not on the communicator's device");
  if (e == cudaSuccess && rc == DLI_OK && !is_root) {
    if (w[kRepWords - 1] != 1) rc = fail(DLI_ERR_ARG, "dli_mesh_replicate: no header received from the root (the root rank reports why)");
    dli_options o;
    dli_options_init(&o);
    o.device = dev;
    if (rc == DLI_OK) rc = dli_mesh_create_from_header(w, &o, &m);
    if (rc == DLI_OK) {
      double gp[3 * DLI_MAX_DIM];
      memcpy(gp, w + DLI_MESH_HEADER_WORDS, sizeof gp);
      dli_mesh_set_grid_params(m, gp);
      m->n_slow_faces = w[DLI_MESH_HEADER_WORDS + 3 * DLI_MAX_DIM];
    }
  }
  // Agree on success before the buffer broadcasts: a rank that could not build its replica (allocation failure, bad
  // header) cannot receive, and the others must not wait for it.  One int64 min-reduction; every rank takes part,
  // failed or not, as long as it got this far with a working stream and scratch word.
  if (N.AllReduce && d_hdr) {
    int64_t ok = (e == cudaSuccess && rc == DLI_OK) ? 1 : 0;
    cudaError_t e2 = cudaMemcpyAsync(d_hdr, &ok, sizeof ok, cudaMemcpyHostToDevice, st);
    if (e2 == cudaSuccess) {
      r = N.AllReduce(d_hdr, d_hdr, 1, ncclInt64, ncclMin, comm, st);
      if (r != ncclSuccess && rc == DLI_OK) rc = nerr(r, "status all-reduce");
    }
    int64_t all_ok = 0;
    if (e2 == cudaSuccess) e2 = cudaMemcpyAsync(&all_ok, d_hdr, sizeof all_ok, cudaMemcpyDeviceToHost, st);
    if (e2 == cudaSuccess) e2 = cudaStreamSynchronize(st);
    if (e2 != cudaSuccess && e == cudaSuccess) e = e2;
    if (e == cudaSuccess && rc == DLI_OK && all_ok != 1)
      rc = fail(DLI_ERR_CUDA, "dli_mesh_replicate: another rank could not take part (its own return value says why); nothing was broadcast");
  }
  // every rank walks the same buffer list (the header fixes which optional buffers exist and how large they are)
  if (e == cudaSuccess && rc == DLI_OK) {
    for (int i = 0; i < dli_mesh_num_buffers(m) && rc == DLI_OK; i++) {
      void *p; int64_t b;
      rc = dli_mesh_buffer(m, i, &p, &b);
      if (rc != DLI_OK || b == 0) continue;
      r = N.Broadcast(p, p, (size_t)b, ncclChar, root, comm, st);
      if (r != ncclSuccess) rc = nerr(r, "buffer broadcast");
    }
    if (rc == DLI_OK) e = cudaStreamSynchronize(st);
  }
  cudaFree(d_hdr);
  cudaStreamDestroy(st);
  if (e != cudaSuccess && rc == DLI_OK) rc = fail(DLI_ERR_CUDA, std::string("dli_mesh_replicate: ") + cudaGetErrorString(e));
  if (rc != DLI_OK) { if (!is_root && m) mesh_free(m); return rc; }
  if (!is_root) {
    if (*inout) mesh_free(*inout);   // a previous replica is replaced
    *inout = m;
  }
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_replicate")
}

// ---------------------------------------------------------------------------------------------
// mesh snapshot: one flat file = magic, header words, grid parameters, then every device buffer
// (length-prefixed).  The same bytes the replication path broadcasts; a loaded mesh is a replica
// (no host copy of the connectivity: dli_mesh_get_simplices / the error evaluator are unavailable).
// ---------------------------------------------------------------------------------------------
static const char kSnapMagic[8] = {'D', 'L', 'I', 'M', 'E', 'S', 'H', '3'};

// 64-bit checksum of a buffer as stored in the file (four interleaved FNV-1a lanes over 8-byte words, folded, then the
// tail bytes): the canonical records, the hint grid and the hull table are range-checked on the device after a load
// (k_validate_ids), but the DERIVED buffers -- walk records in six encodings, packed faces, slot tables -- carry
// neighbour ids the fast pass follows first and are not decoded again; the checksum is what keeps a damaged file from
// becoming out-of-range device reads.  It detects damage, not forgery.
static uint64_t snap_checksum(const void *data, size_t bytes) {
  const unsigned char *p = static_cast<const unsigned char *>(data);
  uint64_t h[4] = {0xcbf29ce484222325ull, 0x84222325cbf29ce4ull, 0x9e3779b97f4a7c15ull, 0xc2b2ae3d27d4eb4full};
  size_t i = 0;
  for (; i + 32 <= bytes; i += 32)
    for (int k = 0; k < 4; k++) {
      uint64_t w;
      memcpy(&w, p + i + 8 * k, 8);
      h[k] = (h[k] ^ w) * 0x100000001b3ull;
    }
  uint64_t r = bytes;
  for (int k = 0; k < 4; k++) r = (r ^ h[k]) * 0x100000001b3ull + (r >> 29);
  for (; i < bytes; i++) r = (r ^ p[i]) * 0x100000001b3ull;
  return r;
}

extern "C" int dli_mesh_save(const dli_mesh *m, const char *path) {
  DLI_GUARD_BEGIN
  if (!m || !path) return fail(DLI_ERR_ARG, "dli_mesh_save: NULL argument");
  FILE *fp = fopen(path, "wb");
  if (!fp) return fail(DLI_ERR_ARG, std::string("dli_mesh_save: cannot open ") + path);
  int64_t h[DLI_MESH_HEADER_WORDS];
  double gp[3 * DLI_MAX_DIM];
  dli_mesh_header(m, h);
  h[14] = m->n_slow_faces;
  dli_mesh_get_grid_params(m, gp);
  bool ok = fwrite(kSnapMagic, 1, 8, fp) == 8 && fwrite(h, sizeof(int64_t), DLI_MESH_HEADER_WORDS, fp) == DLI_MESH_HEADER_WORDS &&
            fwrite(gp, sizeof(double), 3 * DLI_MAX_DIM, fp) == 3 * DLI_MAX_DIM;
  std::vector<char> host;
  int rc = DLI_OK;
  for (int i = 0; ok && rc == DLI_OK && i < dli_mesh_num_buffers(m); i++) {
    void *p; int64_t b;
    rc = dli_mesh_buffer(m, i, &p, &b);
    if (rc != DLI_OK) break;
    ok = fwrite(&b, sizeof b, 1, fp) == 1;
    if (!ok || b == 0) continue;
    host.resize((size_t)b);
    rc = dli_mesh_buffer_read(m, i, host.data());
    if (rc == DLI_OK) ok = fwrite(host.data(), 1, (size_t)b, fp) == (size_t)b;
    const uint64_t sum = snap_checksum(host.data(), (size_t)b);
    if (rc == DLI_OK && ok) ok = fwrite(&sum, sizeof sum, 1, fp) == 1;
  }
  if (fclose(fp) != 0) ok = false;
  if (rc != DLI_OK) return rc;
  if (!ok) return fail(DLI_ERR_ARG, std::string("dli_mesh_save: write failed: ") + path);
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_save")
}

extern "C" int dli_mesh_load(const char *path, const dli_options *opt, dli_mesh **out) {
  DLI_GUARD_BEGIN
  if (!path || !out) return fail(DLI_ERR_ARG, "dli_mesh_load: NULL argument");
  *out = nullptr;
  FILE *fp = fopen(path, "rb");
  if (!fp) return fail(DLI_ERR_ARG, std::string("dli_mesh_load: cannot open ") + path);
  char magic[8];
  int64_t h[DLI_MESH_HEADER_WORDS];
  double gp[3 * DLI_MAX_DIM];
  bool ok = fread(magic, 1, 8, fp) == 8 && memcmp(magic, kSnapMagic, 8) == 0 &&
            fread(h, sizeof(int64_t), DLI_MESH_HEADER_WORDS, fp) == DLI_MESH_HEADER_WORDS &&
            fread(gp, sizeof(double), 3 * DLI_MAX_DIM, fp) == 3 * DLI_MAX_DIM;
  if (!ok) { fclose(fp); return fail(DLI_ERR_ARG, std::string("dli_mesh_load: not a mesh snapshot: ") + path); }
  dli_mesh *m = nullptr;
  int rc = dli_mesh_create_from_header(h, opt, &m);
  if (rc != DLI_OK) { fclose(fp); return rc; }
  dli_mesh_set_grid_params(m, gp);
  m->n_slow_faces = h[14];
  std::vector<char> host;
  for (int i = 0; rc == DLI_OK && i < dli_mesh_num_buffers(m); i++) {
    void *p; int64_t b, want;
    rc = dli_mesh_buffer(m, i, &p, &want);
    if (rc != DLI_OK) break;
    if (fread(&b, sizeof b, 1, fp) != 1 || b != want) { rc = fail(DLI_ERR_ARG, "dli_mesh_load: truncated or inconsistent snapshot"); break; }
    if (b == 0) continue;
    host.resize((size_t)b);
    uint64_t sum = 0;
    if (fread(host.data(), 1, (size_t)b, fp) != (size_t)b || fread(&sum, sizeof sum, 1, fp) != 1) { rc = fail(DLI_ERR_ARG, "dli_mesh_load: truncated snapshot"); break; }
    if (sum != snap_checksum(host.data(), (size_t)b)) { rc = fail(DLI_ERR_ARG, "dli_mesh_load: checksum mismatch (damaged snapshot)"); break; }
    rc = dli_mesh_buffer_write(m, i, host.data());
  }
  fclose(fp);
  if (rc == DLI_OK && m->ns > 0) {   // second line of defence behind the checksums: canonical records, hints and hull table in range
    DeviceGuard guard(m->device);
    unsigned int *d_bad = reinterpret_cast<unsigned int *>(m->d_stats + ST_WORDS - 1);
    unsigned int bad = 0;
    const long long span = std::max<long long>(std::max<long long>(m->ns, m->grid.cells), m->nh);
    cudaError_t e = cudaMemset(d_bad, 0, sizeof(unsigned int));
    if (e == cudaSuccess) {
      k_validate_ids<<<(unsigned)((span + 255) / 256), 256>>>(m->ns, m->nv, m->nh, m->dim, m->ss, m->d_srec, m->grid.cells, m->d_hint,
                                                            m->dim == 2 ? m->d_hull : nullptr, d_bad);
      e = cudaMemcpy(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost);
    }
    if (e == cudaSuccess) e = cudaMemset(d_bad, 0, sizeof(unsigned int));
    if (e != cudaSuccess) rc = fail(DLI_ERR_CUDA, std::string("dli_mesh_load: validation failed: ") + cudaGetErrorString(e));
    else if (bad) rc = fail(DLI_ERR_ARG, "dli_mesh_load: snapshot holds out-of-range indices (damaged file)");
  }
  if (rc != DLI_OK) { mesh_free(m); return rc; }
  *out = m;
  return DLI_OK;
  DLI_GUARD_END("dli_mesh_load")
}

// ---------------------------------------------------------------------------------------------
// predicate test hook
// ---------------------------------------------------------------------------------------------
extern "C" int dli_debug_orient(int dim, int64_t n, const double *pts, int8_t *sign_out, uint8_t *exact_used) {
  if (dim < 2 || dim > 4) return fail(DLI_ERR_UNSUPPORTED, "dli_debug_orient: dim must be 2..4");
  if (n < 0 || (n > 0 && (!pts || !sign_out))) return fail(DLI_ERR_ARG, "dli_debug_orient: bad argument");
  if (n == 0) return DLI_OK;
  if (dli_device_count() <= 0) return fail(DLI_ERR_CUDA, "no CUDA device available");
  size_t per = (size_t)(dim + 1) * dim;
  double *d_p = nullptr;
  int8_t *d_s = nullptr;
  uint8_t *d_e = nullptr;
  CU(cudaMalloc(&d_p, sizeof(double) * per * (size_t)n));
  cudaError_t e = cudaMalloc(&d_s, (size_t)n);
  if (e == cudaSuccess) e = cudaMalloc(&d_e, (size_t)n);
  if (e == cudaSuccess) e = cudaMemcpy(d_p, pts, sizeof(double) * per * (size_t)n, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    unsigned blocks = (unsigned)((n + 63) / 64);
    switch (dim) {
      case 2: k_debug_orient<2><<<blocks, 64>>>(n, d_p, d_s, d_e); break;
      case 3: k_debug_orient<3><<<blocks, 64>>>(n, d_p, d_s, d_e); break;
      case 4: k_debug_orient<4><<<blocks, 64>>>(n, d_p, d_s, d_e); break;
    }
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(sign_out, d_s, (size_t)n, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && exact_used) e = cudaMemcpy(exact_used, d_e, (size_t)n, cudaMemcpyDeviceToHost);
  cudaFree(d_p); cudaFree(d_s); cudaFree(d_e);
  if (e != cudaSuccess) return fail(DLI_ERR_CUDA, std::string("dli_debug_orient: ") + cudaGetErrorString(e));
  return DLI_OK;
}

extern "C" int dli_debug_orient_host(int dim, int64_t n, const double *pts, int8_t *sign_out) {
  if (dim < 2 || dim > 4) return fail(DLI_ERR_UNSUPPORTED, "dli_debug_orient_host: dim must be 2..4");
  if (n < 0 || (n > 0 && (!pts || !sign_out))) return fail(DLI_ERR_ARG, "dli_debug_orient_host: bad argument");
  size_t per = (size_t)(dim + 1) * dim;
  for (int64_t i = 0; i < n; i++) sign_out[i] = (int8_t)orient_host(dim, pts + per * (size_t)i);
  return DLI_OK;
}
