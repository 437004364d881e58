/*
 * dli.h -- C ABI of the B200 batched Delaunay linear-interpolation engine.
 *
 * This is the drop-in boundary for the reference's interp hot path.  The reference
 * (rncarpio/delaunay_linterp) has no FFI layer: the path is a header-only C++ member function
 * called once per query.  Each entry point below names the reference interface it replaces
 * (paths relative to the reference tree).  All pointers are plain C arrays, row-major; no C++
 * or torch types cross this boundary.  See INTEGRATION.md for the reference-side bindings.
 *
 * Conventions
 *   - return value: 0 = success, negative = error (dli_last_error() has the message).  No
 *     exception ever crosses the ABI.
 *   - query / result pointers may be HOST or DEVICE pointers (detected with
 *     cudaPointerGetAttributes).  With device pointers the call only enqueues work on `stream`
 *     and returns; with host pointers it runs a chunked H2D / compute / D2H pipeline and returns
 *     when the results are in the host buffers.  Page-locked host buffers (cudaMallocHost /
 *     cudaHostRegister) are copied by DMA directly and run at the PCIe rate; pageable buffers
 *     (malloc, std::vector, numpy) of >= 2^18 queries are staged through the library's own pinned
 *     buffers by up to 8 worker threads (environment variable DLI_HOST_COPY_THREADS overrides the
 *     default of half the host threads) -- about 4 x the driver's own pageable path.
 *   - per-query conditions are reported in status_out, never as an error return.
 *   - a mesh handle is immutable apart from dli_mesh_set_values() and dli_mesh_update(), which must
 *     not run while queries on the mesh are in flight; interp/locate calls on the same mesh may be
 *     issued concurrently from several host threads / streams.
 *   - the library needs a CUDA device (sm_100a).  There is no CPU fallback: without a device
 *     every call that touches the GPU fails with DLI_ERR_CUDA.
 */
#ifndef DLI_H_
#define DLI_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLI_VERSION 2
#define DLI_MAX_DIM 4 /* dimensions 2, 3 and 4 are supported */

/* error returns */
#define DLI_OK 0
#define DLI_ERR_ARG (-1)     /* bad argument */
#define DLI_ERR_MESH (-2)    /* invalid mesh: degenerate simplex, bad index, non-manifold facet */
#define DLI_ERR_CUDA (-3)    /* CUDA runtime error or no device */
#define DLI_ERR_NOMEM (-4)
#define DLI_ERR_UNSUPPORTED (-5)

/* per-query status (status_out) */
#define DLI_STATUS_OK 0                /* located in a closed simplex (FACE / EDGE / VERTEX in CGAL terms) */
#define DLI_STATUS_OUTSIDE_PROJECTED 1 /* 2D: outside the hull; projected on the hull edge line and re-located
                                          (src/delaunay_2_interp.h:530-545) */
#define DLI_STATUS_OUTSIDE_NAN 2       /* dD: outside the hull; value is quiet NaN, simplex -1
                                          (src/delaunay_d_interp.h:1005-1008) */
#define DLI_STATUS_STEP_LIMIT 3        /* no simplex found: walk exceeded max_walk_steps (invalid / non-Delaunay mesh), or
                                          the query has a NaN / Inf coordinate; value NaN, simplex -1.
                                          The 2D header wrapper turns this into std::invalid_argument
                                          (src/delaunay_2_interp.h:546-551) */
#define DLI_STATUS_NO_MESH 4           /* mesh has no simplex (fewer than dim+1 affinely independent points) */
#define DLI_STATUS_ON_EDGE 5           /* dli_gradient only: query exactly on an edge (CGAL Locate_type EDGE): average of
                                          the two incident faces (src/delaunay_2_interp.h:633-636) */
#define DLI_STATUS_ON_VERTEX 6         /* dli_gradient only: query exactly on a vertex (Locate_type VERTEX): average over
                                          the incident finite faces (src/delaunay_2_interp.h:619-632) */

#define DLI_EPILOGUE_AUTO 0
#define DLI_EPILOGUE_CLOSED_FORM 1
#define DLI_EPILOGUE_REFERENCE 2

typedef struct dli_mesh dli_mesh; /* opaque: owns the device structure-of-arrays buffers and the hint grid */

typedef struct dli_options {
  int32_t struct_size;            /* sizeof(dli_options); lets the struct grow compatibly */
  int32_t device;                 /* CUDA device ordinal, -1 = current device */
  float hint_cells_per_simplex;   /* start-hint grid resolution; <= 0 = default (2D: 2, or 4 with cell-indexed records; 3D: 4, 4D: 8) */
  int32_t max_walk_steps;         /* <= 0 = default (n_simplices + 16) */
  int32_t reserved0;              /* must be 0 (was: device-side query binning; measured not to pay) */
  int32_t epilogue;               /* DLI_EPILOGUE_*: 0 = AUTO (default): 2D meshes with < 2^21-1 faces use the 64-byte
                                     walk records and evaluate each face's affine interpolant f0 + g.(p - p0) on faces
                                     whose a-priori error bound keeps it within 1e-13 relative of the exact
                                     interpolant, and the reference operation order everywhere else (other faces,
                                     other dims, stragglers);  1 = closed-form barycentric solve everywhere
                                     (differs in the last ulps on slivers);  2 = reference operation order
                                     everywhere (partial-pivot LU determinants; bit-identical to a scalar restatement of the reference) */
  int32_t threads_per_block;      /* <= 0 = default */
  int32_t queries_per_warp;       /* 2D fast pass: contiguous queries owned by one warp; <= 0 = default (1024) */
  int32_t packed_faces;           /* 2D: 0 = compact fast-pass records (16-byte packed faces / 64-byte walk records; default),
                                     -1 = never, 2 = walk records with delta-encoded neighbours even on small meshes
                                     (the encoding large meshes get; for tests), 3 = cell-indexed walk records even on
                                     small meshes (the layout meshes with >= 8M faces get: a query's start record is
                                     addressed by its hint cell, no hint read; for tests);
                                     4D: 0 = one packed 128-byte walk record per simplex (block-floating-point barycentric
                                     gradients, 24-bit neighbour ids; default below 2^24 - 1 simplices), 5 = the split layout
                                     larger meshes get (128-byte geometry record + 64-byte interpolant record; for tests) */
  int32_t prefetch_ring;          /* 2D walk-record fast pass: 0 = per-warp cp.async prefetch ring for queries and start
                                     hints (default), -1 = plain loads */
  int32_t outside_nan;            /* 2D: 0 = queries outside the hull are projected on the hull (Delaunay_incremental_interp_2,
                                     src/delaunay_2_interp.h:530-545; default), 1 = they report DLI_STATUS_OUTSIDE_NAN with a
                                     quiet NaN like the dD class (src/delaunay_d_interp.h:1005-1008; what
                                     Delaunay_incremental_interp_d(2) needs).  dD meshes always report NaN. */
  int32_t start_records;          /* dD walk records: 0 = default: 4D meshes on packed records keep, behind the records, one COPY of the
                                     hinted simplex's record per hint cell, so that a query starts with ONE memory access
                                     (the record in its cell) instead of two dependent ones (hint, then record) -- costs
                                     128 B per hint cell (auto only while that stays below 24 GB); -1 = never; 1 = also in 3D (tests) */
  int32_t reserved[4];
} dli_options;

/* fills *opt with the defaults */
void dli_options_init(dli_options *opt);

/*
 * Flattener.  Replaces the pointer-based CGAL containers the path reads:
 *   2D  Delaunay_2 / Triangulation_data_structure_2 (src/delaunay_2_interp.h:66-95, member m_pTriang :405;
 *       accessors face->vertex(i), face->neighbor(i), vertex->point(), vertex->info())
 *   dD  Delaunay_d + m_pointMap (src/delaunay_d_interp.h:303,757; modified_cgal/Delaunay_d.h:389-396,
 *       426-439, 1015-1028)
 * coords            [n_vertices][dim]
 * f                 [n_vertices]
 * simplex_vertices  [n_simplices][dim+1]   finite simplices only, any orientation
 * simplex_neighbors [n_simplices][dim+1]   k-th entry is the simplex opposite vertex k, -1 across the hull;
 *                                          NULL = derive from shared facets
 * The arrays are copied; the caller keeps ownership.  n_simplices == 0 is allowed (every query then
 * reports DLI_STATUS_NO_MESH).
 */
int dli_mesh_create(int dim, int64_t n_vertices, const double *coords, const double *f, int64_t n_simplices,
                    const int32_t *simplex_vertices, const int32_t *simplex_neighbors, const dli_options *opt,
                    dli_mesh **out);

/* same geometry, new vertex values (vertex->info() = f, src/delaunay_2_interp.h:446) */
int dli_mesh_set_values(dli_mesh *mesh, const double *f /* [n_vertices], host */);

/*
 * Incremental update after insertions (2D, 3D, 4D).  Replaces the reference's per-insert bookkeeping: the dirty set of
 * an insertion is the star of the new vertex (2D: insert_and_get_modifications, src/delaunay_2_interp.h:434-448, value
 * overwrite of an existing vertex :446; dD: p_modified_simplices, modified_cgal/Convex_hull_d.h:1172-1178, consumed at
 * src/delaunay_d_interp.h:939-965) -- only those simplices are re-sent and re-derived on the device (canonical
 * records, packed records, walk records, hull-edge table when the hull changed) instead of flattening the whole mesh.
 *   n_vertices_total   vertices after the update; the (n_vertices_total - old n_vertices) new ones are appended:
 *                      new_coords [.][dim], new_f [.]
 *   n_value_changes    existing vertices whose value changed: value_ids [.], values [.]
 *   n_simplices_total  simplices after the update.  The caller's id space stays compact: a simplex that disappeared
 *                      hands its id to a new one, further new simplices are appended (ids >= old n_simplices)
 *   n_dirty            simplices whose vertex list, neighbour list or vertex values differ from what the mesh holds:
 *                      dirty_ids [.] (each id once), dirty_vertices [.][dim+1] (any orientation),
 *                      dirty_neighbors [.][dim+1] (-1 across the hull).  Every id >= old n_simplices must be listed.
 * Start hints are kept (an id that is reused stays in the neighbourhood of its old simplex), so hint quality decays
 * slowly with many updates; re-create the mesh now and then (dli_host.h does, see di2_refresh).
 * Returns DLI_ERR_UNSUPPORTED -- mesh unchanged -- when the mesh cannot be patched and must be re-created: a replica
 * or loaded snapshot (no host connectivity), cell-indexed or delta-encoded 2D walk records, split 4D walk records, ids
 * that no longer fit the packed records, or a new vertex outside the bounding box of the start-hint grid.
 * dD: the start copies of a 4D mesh (dli_options.start_records) are dropped by the first patch and come back with the
 * next full flatten.
 * No interp / locate call may be in flight on the mesh.
 */
int dli_mesh_update(dli_mesh *mesh, int64_t n_vertices_total, const double *new_coords, const double *new_f,
                    int64_t n_value_changes, const int32_t *value_ids, const double *values, int64_t n_simplices_total,
                    int64_t n_dirty, const int32_t *dirty_ids, const int32_t *dirty_vertices,
                    const int32_t *dirty_neighbors);

void dli_mesh_destroy(dli_mesh *mesh);

int dli_mesh_dim(const dli_mesh *mesh);
int64_t dli_mesh_num_vertices(const dli_mesh *mesh);
int64_t dli_mesh_num_simplices(const dli_mesh *mesh);
/* bytes of device memory held by the mesh (records + hint grid) */
int64_t dli_mesh_device_bytes(const dli_mesh *mesh);
/* canonical (positively oriented) simplices as the engine stores them, in the caller's numbering; host arrays,
   either may be NULL.  Vertex order inside a simplex is the order the barycentric epilogue uses. */
int dli_mesh_get_simplices(const dli_mesh *mesh, int32_t *simplex_vertices, int32_t *simplex_neighbors);

/*
 * Batched interp.  Replaces, for n queries at once,
 *   Delaunay_incremental_interp_2::interp / interp_point   (src/delaunay_2_interp.h:510-513, 515-572)
 *   Delaunay_incremental_interp_d::interp / interp_point   (src/delaunay_d_interp.h:982-986, 988-1010)
 * i.e. locate (CGAL Triangulation_2::locate / Delaunay_d::locate, modified_cgal/Delaunay_d.h:811-850)
 * + compute_barycentric_coords (:223-239 / :628-650) + weighted gather (:563-569 / :998-1004).
 * xs          [n][dim]  host or device
 * ys          [n]       same memory space as xs
 * simplex_out [n] or NULL: id (caller's numbering) of the simplex used, -1 if none
 * status_out  [n] or NULL: DLI_STATUS_*
 * stream      cudaStream_t (as void*), NULL = default stream
 */
int dli_interp(const dli_mesh *mesh, int64_t n, const double *xs, double *ys, int32_t *simplex_out,
               uint8_t *status_out, void *stream);

/*
 * dli_interp for callers that cannot allocate device memory.  No counterpart in the reference (it has no GPU
 * path and no allocation contract); it exists because of how this engine is built: dli_interp draws its straggler
 * queue -- the hand-over list between the fast walk kernel and the exact-arithmetic kernel -- from the CUDA
 * stream-ordered pool on every call.  That is a problem for
 *   - a host program that gave ALL device memory to its own pool at start-up, so the driver has nothing left for us;
 *   - a loop that may not risk the driver lock or the device-wide synchronisation an allocation can imply while
 *     its other streams are busy;
 *   - a driver set-up without stream-ordered memory pools.
 * This entry point makes no cudaMalloc / cudaMallocAsync / cudaFree call and calls no allocating library.  Its queue
 * is a fixed array inside libdli.so's device image: 8 MiB + 256 B on EVERY GPU that loads the library, held until
 * the process exits, used or not.  Same kernels, same arithmetic, same results as dli_interp.
 *   limit    n <= DLI_NOALLOC_MAX_QUERIES; a larger n returns DLI_ERR_ARG and touches nothing (split the batch)
 *   pointers DEVICE pointers only (xs, ys, and simplex_out / status_out when given); host pointers return
 *            DLI_ERR_ARG -- the host pipeline of dli_interp owns staging buffers, which is what this call avoids
 *   sharing  all calls on one GPU -- any mesh, any stream, any thread -- use the same array.  Two calls whose
 *            kernels overlap on the GPU race on it and return garbage: keep them on one stream, or order them
 *            with events.  The array keeps the previous call's contents; the call re-zeroes the one word
 *            (the queue counter) that has to start at zero, with a cudaMemsetAsync on `stream`.
 *   set-up   the FIRST call on each device looks up the array's device address (cudaGetSymbolAddress) and keeps
 *            it; make that first call before you begin any stream capture.
 *   timing   returns DLI_ERR_UNSUPPORTED while per-launch timing (dli_mesh_collect_stats bit 1) is on, because that
 *            mode creates CUDA events per call; the counters (bit 0) work.
 */
#define DLI_NOALLOC_MAX_QUERIES (1 << 20)
int dli_interp_noalloc(const dli_mesh *mesh, int64_t n, const double *xs, double *ys, int32_t *simplex_out,
                       uint8_t *status_out, void *stream);

/* Point location only (CGAL locate as used at src/delaunay_2_interp.h:525, src/delaunay_d_interp.h:990).
   2D queries outside the hull report the hull face behind the canonical visible hull edge. */
int dli_locate(const dli_mesh *mesh, int64_t n, const double *xs, int32_t *simplex_out, uint8_t *status_out,
               void *stream);

/*
 * Batched gradient of the piecewise-linear interpolant (2D only, as in the reference).  Replaces, for n queries,
 *   Delaunay_incremental_interp_2::gradient              (src/delaunay_2_interp.h:608-642)
 * with the per-face gradients of set_simplex_gradient / simplex_gradient (:574-588, :242-263).
 *   strictly inside a face  -> that face's gradient                                   status DLI_STATUS_OK
 *   exactly on an edge      -> mean of the two incident faces (a hull edge: the one)  status DLI_STATUS_ON_EDGE
 *   exactly on a vertex     -> mean over the incident finite faces                    status DLI_STATUS_ON_VERTEX
 *   outside the hull        -> (NaN, NaN), simplex -1; gradients are NOT projected    status DLI_STATUS_OUTSIDE_NAN
 * EDGE / VERTEX are decided by the exact orientation predicate.  The vertex mean is summed counter-clockwise
 * starting from the incident face with the smallest id (a hull vertex: from the clockwise-most face); CGAL's
 * circulator starts at an arbitrary incident face, so the reference's own sum order is not reproducible.
 * xs    [n][2]  host or device
 * grads [n][2]  same memory space as xs
 */
int dli_gradient(const dli_mesh *mesh, int64_t n, const double *xs, double *grads, int32_t *simplex_out,
                 uint8_t *status_out, void *stream);

/*
 * Batched per-simplex error evaluator.  Replaces compute_simplex_error + add_error_of_boundary_subsimplices
 * (src/delaunay_2_interp.h:725-741, 683-699, 657-680, 331-342; dD src/delaunay_d_interp.h:1078-1093, 810-848)
 * for every simplex of the mesh at once.  The host callback m_fn (src/delaunay_2_interp.h:408, 421-431)
 * stays on the host, between the two calls:
 *     n = dli_simplex_error_count(mesh);
 *     dli_simplex_error_prepare(mesh, centroids, item_simplex, item_vertices);   // device kernel
 *     phi[i] = fn(centroids[i]);                                                 // caller
 *     dli_simplex_error_finish(mesh, n, phi, err);                               // device kernel
 * Items: every simplex, then for each hull facet of it every vertex subset of size >= 2 (duplicates across
 * simplices kept, as in the reference).  Order: simplices ascending; inside a simplex the simplex itself,
 * then hull facets by ascending slot, subsets by ascending bitmask.
 * item_vertices is [n][dim+1], padded with -1; item_simplex / item_vertices are host arrays.
 * centroids (prepare) and f_at_centroids / err (finish) may be HOST pointers (the call copies and returns when the data
 * is there) or DEVICE pointers (a caller whose f is evaluated on the GPU: the kernel reads / writes the caller's buffers
 * directly, enqueued on the default stream, nothing is staged through the host and the call does not wait).
 */
int64_t dli_simplex_error_count(const dli_mesh *mesh);
int dli_simplex_error_prepare(const dli_mesh *mesh, double *centroids /* [n][dim] */,
                              int32_t *item_simplex /* [n] or NULL */, int32_t *item_vertices /* [n][dim+1] or NULL */);
int dli_simplex_error_finish(const dli_mesh *mesh, int64_t n, const double *f_at_centroids /* [n] */,
                             double *err /* [n] */);

/*
 * Multi-GPU replication support: the mesh is a set of device buffers.  Rank 0 flattens; the other ranks
 * allocate an empty mesh of the same shape from the 16-word header and receive the buffers through any
 * broadcast (NCCL / torch.distributed) on the raw device pointers.  The header carries the options that decide what a
 * query returns (outside_nan, epilogue, the walk-step cap): a replica, and a mesh loaded from a snapshot, answers
 * like its source whatever `opt` says about those; `opt` chooses the device and the launch shape.
 */
#define DLI_MESH_HEADER_WORDS 16
int dli_mesh_header(const dli_mesh *mesh, int64_t header[DLI_MESH_HEADER_WORDS]);
int dli_mesh_create_from_header(const int64_t header[DLI_MESH_HEADER_WORDS], const dli_options *opt, dli_mesh **out);
int dli_mesh_num_buffers(const dli_mesh *mesh);
int dli_mesh_buffer(const dli_mesh *mesh, int index, void **device_ptr, int64_t *bytes);
/* whole-buffer copies out of / into the mesh; the other side may be a host or a device pointer */
int dli_mesh_buffer_read(const dli_mesh *mesh, int index, void *dst);
int dli_mesh_buffer_write(dli_mesh *mesh, int index, const void *src);
/* hint-grid geometry (origin, cell size, 1/cell size; 3 x DLI_MAX_DIM doubles), host side of the replication */
int dli_mesh_get_grid_params(const dli_mesh *mesh, double *out /* [3 * DLI_MAX_DIM] */);
int dli_mesh_set_grid_params(dli_mesh *mesh, const double *in /* [3 * DLI_MAX_DIM] */);

/*
 * The same in one call for C / C++ callers that hold an NCCL communicator: the root rank passes its
 * mesh in *inout; every other rank passes *inout == NULL (or a replica to be replaced) and receives a replica on the
 * communicator's device -- one ncclBroadcast of the header block, then one per buffer, over NVLink / NVSwitch.  Collective:
 * every rank of the communicator must call it.  nccl_comm is an ncclComm_t.  NCCL itself is resolved at run time from
 * the calling process (link your program against NCCL as usual, or set DLI_NCCL_LIB to the library's path); the call
 * returns DLI_ERR_UNSUPPORTED when no NCCL can be found.  A replica answers interp / locate / gradient; it has no host
 * connectivity (dli_mesh_get_simplices, dli_mesh_update and the error evaluator need the creating rank).
 */
int dli_mesh_replicate(dli_mesh **inout, void *nccl_comm, int root);

/*
 * Mesh snapshot (no counterpart in the reference, which has no serialisation): the flattened device mesh as one
 * flat file -- magic "DLIMESH3", the 16 header words, the grid parameters, then every device buffer
 * length-prefixed and followed by a 64-bit checksum -- so that a large triangulation is built and flattened once.
 * dli_mesh_load rejects a file whose checksums do not match (damage anywhere, derived walk records included) and
 * range-checks the canonical records, hints and hull table on the device.  A loaded mesh is a replica:
 * interp / locate / gradient work, dli_mesh_get_simplices and the error evaluator need the creating process.
 */
int dli_mesh_save(const dli_mesh *mesh, const char *path);
int dli_mesh_load(const char *path, const dli_options *opt, dli_mesh **out);

/* instrumentation: counters of the most recent interp/locate on this mesh that had `collect_stats` enabled */
typedef struct dli_stats {
  int64_t queries;
  int64_t walk_steps;       /* simplices visited, all queries */
  int64_t max_walk_steps;   /* longest single walk */
  int64_t exact_fallbacks;  /* orientation tests that fell through the static filter */
  int64_t outside;          /* queries that left the hull */
  int64_t kernel_launches;  /* launches issued by the most recent call */
  /* timing mode (enable & 2): CUDA-event durations accumulated over every call since timing was enabled */
  int64_t fast_kernel_launches;
  double fast_kernel_ms;    /* sum of the fast-pass kernel durations */
  double full_kernel_ms;    /* sum of the straggler/exact-pass kernel durations */
  int64_t straggler_queries; /* queries the fast pass handed to the straggler/exact pass (counters mode) */
  int64_t slow_simplices;    /* simplices whose walk record carries the slow flag (always served by the exact pass) */
} dli_stats;
/* enable: 0 = off, bit 0 = counters (adds atomics to the kernels), bit 1 = per-launch CUDA-event timing of the
   two interp kernels on the launching stream.  Enabling resets the accumulated timing. */
int dli_mesh_collect_stats(dli_mesh *mesh, int enable);
int dli_mesh_get_stats(const dli_mesh *mesh, dli_stats *out);

/* test hook: sign of the orientation determinant of n simplices given as [n][dim+1][dim] device-side
   filtered-exact predicate (the one the walk uses); host pointers; sign_out [n] in {-1,0,1};
   exact_used_out [n] or NULL tells whether the exact fallback ran. */
int dli_debug_orient(int dim, int64_t n, const double *pts, int8_t *sign_out, uint8_t *exact_used_out);

/* the same predicate evaluated on the host (the flattener uses it to orient simplices); needs no GPU */
int dli_debug_orient_host(int dim, int64_t n, const double *pts, int8_t *sign_out);

/* thread-local message for the last non-zero return on this thread */
const char *dli_last_error(void);
/*
 * Plot data (2D), batched.  Replaces get_line_segments (src/delaunay_2_interp.h:842-862) and, per edge,
 * is_edge_concave (:766-803): every finite edge once as (P.x, P.y, f_P, Q.x, Q.y, f_Q) with its concavity flag
 * (1 on the hull, :778-780; else both components of the gradient jump across the edge, oriented by the side the
 * first face lies on, are <= 0), and the min / max of (x, y, f) over the edge end points (:853-858).
 * CGAL's edge iterator order is arbitrary; here: the order of the engine's face storage, an interior edge emitted
 * by its face with the lower caller id (P = vertex ccw(slot), Q = vertex cw(slot) of that face).
 * segments [n_edges][2][3] and concave [n_edges]: host or device pointers, either may be NULL;
 * min_coords / max_coords [3]: host, may be NULL (DBL_MAX / -DBL_MAX when there is no edge, as the reference).
 */
int64_t dli_mesh_num_edges(const dli_mesh *mesh);
int dli_mesh_line_segments(const dli_mesh *mesh, double *segments, int32_t *concave, double *min_coords,
                           double *max_coords);

/* number of CUDA devices visible (0 if none / no driver) */
int dli_device_count(void);

#ifdef __cplusplus
}
#endif
#endif /* DLI_H_ */
