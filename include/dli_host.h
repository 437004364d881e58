/*
 * dli_host.h -- C ABI of the host-side objects that stand where the reference keeps CGAL:
 * incremental Delaunay triangulations (2D: dli_di2, d dimensions: dli_dd) with vertex values and the
 * adaptive-insertion error queue, feeding the GPU engine (include/dli.h) through the flattener.
 *
 * The reference builds on CGAL::Delaunay_triangulation_2 (src/delaunay_2_interp.h:66-95), which is
 * not available here (no CGAL in the image), so construction is native: Bowyer-Watson insertion with
 * exact orientation / in-circle predicates and CGAL's symbolic perturbation for cocircular points
 * (Delaunay_triangulation_2::side_of_oriented_circle(p0,p1,p2,p,perturb)), which makes the
 * triangulation independent of insertion order and reproduces the README known answer.
 *
 * dli_di2 mirrors class Delaunay_incremental_interp_2 (src/delaunay_2_interp.h:402-864):
 *   insert                         :450-493   -> dli_di2_insert
 *   insert_largest_error_point     :498-504   -> dli_di2_insert_largest_error_point
 *   get_largest_error_tuple        :506-508   -> dli_di2_get_largest_error_tuple
 *   interp / interp_point          :510-572   -> dli_di2_interp (batched, on the GPU)
 *   gradient                       :608-642   -> dli_di2_gradient (batched, on the GPU)
 *   get_error_queue                :743-746   -> dli_di2_get_error_queue
 *   get_all_vertices               :806-840   -> dli_di2_get_all_vertices
 *   get_line_segments              :842-862   -> dli_di2_get_line_segments
 *   compute_simplex_error + add_error_of_boundary_subsimplices :725-741, :683-699 (host, per inserted star;
 *   the whole-mesh batched form is dli_simplex_error_* in dli.h)
 *
 * Threads: a dli_di2 / dli_dd handle is, like the reference's classes, a single-threaded object -- one thread at a
 * time may call into a given handle (interp included: it may re-flatten or patch the device mesh, and one-point calls
 * walk the host triangulation from the previous result).  Concurrency belongs one level down: several threads may
 * call dli_interp / dli_locate / dli_gradient on ONE dli_mesh (dli_di2_device_mesh / dli_dd_device_mesh) at once,
 * each on its own stream.
 *
 * Deviation from the reference, both classes: a queue tuple whose insertion point already is a vertex is dropped when it
 * reaches the top.  On a hull that is not an axis-aligned box the ROUNDED centroid of a hull sub-simplex is inserted an
 * ulp off the hull, the sub-simplex survives and proposes the same point again; the reference re-inserts it (a value
 * overwrite that changes nothing) and returns the same tuple from insert_largest_error_point forever
 * (src/delaunay_2_interp.h:498-504, src/delaunay_d_interp.h:1028-1034).  Sequences in which every call adds a vertex --
 * the README's -- are unaffected.
 */
#ifndef DLI_HOST_H_
#define DLI_HOST_H_

#include <stdint.h>
#include "dli.h"

#ifdef __cplusplus
extern "C" {
#endif

/* host callback: f(x), the reference's std::function<double(int, double*)> m_fn (src/delaunay_2_interp.h:408) */
typedef double (*dli_fn)(int n_dims, const double *x, void *user);

typedef struct dli_di2 dli_di2;

/* fn may be NULL: no error queue is kept (ctor without callback, src/delaunay_2_interp.h:410-412) */
dli_di2 *dli_di2_create(dli_fn fn, void *user);
void dli_di2_destroy(dli_di2 *d);

/* insert(x, f): a duplicate point only overwrites its value (CGAL insert returns the existing vertex).
   Returns the vertex id (>= 0) or a negative error. */
int32_t dli_di2_insert(dli_di2 *d, double x, double y, double f);
/* bulk insertion; spatial_sort != 0 inserts along a Hilbert curve (same triangulation, faster).
   With a callback the error queue is rebuilt afterwards (batched on the GPU when use_gpu != 0). */
int dli_di2_insert_many(dli_di2 *d, int64_t n, const double *xy, const double *f, int spatial_sort, int use_gpu);

/* pops the largest-error tuple, inserts its point with the stored f (no re-evaluation), returns it.
   DLI_ERR_ARG if the queue is empty or there is no callback (the reference asserts, :499-500). */
int dli_di2_insert_largest_error_point(dli_di2 *d, double *err, double *f, double *xy);
int dli_di2_get_largest_error_tuple(dli_di2 *d, double *err, double *f, double *xy);
/* batched callback: out[i] = f(xs[i]), xs [n][n_dims] */
typedef void (*dli_fn_batch)(int n_dims, int64_t n, const double *xs, double *out, void *user);
/* k-at-a-time adaptive refinement: takes up to k largest-error tuples in queue order -- skipping, and leaving queued,
   tuples whose face shares a vertex with a face already taken in this round -- inserts their points with the stored
   f (insert_largest_error_point :498-504 for each), then evaluates the errors of all faces created in the round
   (compute_simplex_error :725-741 + add_error_of_boundary_subsimplices :683-699) in one batch with ONE call of fn_batch
   (NULL: the scalar callback of dli_di2_create is called per item).  A round ends early at the first tuple whose error
   is below rel_threshold x (the round's largest error), so that a large k does not spend points on faces that are
   already good (0 = take everything).  k = 1 is insert_largest_error_point.  Returns the
   number of points inserted (< k when the queue runs out) or a negative error; err/f/xy [k] receive the tuples (may
   be NULL). */
int64_t dli_di2_refine(dli_di2 *d, int64_t k, double rel_threshold, dli_fn_batch fn_batch, void *user, double *err,
                       double *f, double *xy);
int64_t dli_di2_error_queue_size(dli_di2 *d);
/* all live tuples ordered by decreasing error; arrays of dli_di2_error_queue_size() entries */
int dli_di2_get_error_queue(dli_di2 *d, double *err, double *f, double *xy);

int64_t dli_di2_num_vertices(const dli_di2 *d);
int64_t dli_di2_num_faces(const dli_di2 *d); /* finite faces */
/* triangulation dimension: -1 empty, 0 one point, 1 all collinear, 2 triangulated */
int dli_di2_dimension(const dli_di2 *d);

/* flattened triangulation: coords [nv][2], f [nv], tri [nf][3] (ccw), nbr [nf][3] (-1 = hull); any may be NULL */
int dli_di2_export(const dli_di2 *d, double *coords, double *f, int32_t *tri, int32_t *nbr);

/* batched interp on the GPU; after insertions the device mesh is brought in step lazily: the stars of the inserted
   points are patched in place (dli_mesh_update, ids of export() are stable across insertions), a full re-flatten
   happens after bulk insertion, when more faces changed than exist, or when the mesh cannot be patched.  Host or device pointers as dli_interp.
   With dimension < 2 every query reports DLI_STATUS_NO_MESH (the reference throws, :546-551).
   Very small batches in plain host memory (the reference's one-point-per-call usage) are answered by an exact walk on
   the host triangulation; that path follows the handle's options: outside_nan = 1 gives NaN / DLI_STATUS_OUTSIDE_NAN
   outside the hull there too, and a handle set to DLI_EPILOGUE_CLOSED_FORM never takes it. */
int dli_di2_interp(dli_di2 *d, int64_t n, const double *xs, double *ys, int32_t *simplex_out, uint8_t *status_out,
                   void *stream);
/* batched gradient on the GPU (dli_gradient): grads [n][2]; NaN outside the hull (:610-611, :618) */
int dli_di2_gradient(dli_di2 *d, int64_t n, const double *xs, double *grads, int32_t *simplex_out, uint8_t *status_out,
                     void *stream);
/* how often the device mesh was re-created / patched in place (dli_mesh_update) so far */
void dli_di2_sync_counts(const dli_di2 *d, int64_t *n_full, int64_t *n_incremental);
/* the device mesh currently backing dli_di2_interp (NULL before the first interp); owned by d */
const dli_mesh *dli_di2_device_mesh(dli_di2 *d);
int dli_di2_set_options(dli_di2 *d, const dli_options *opt);

/* plot data: vertices as (x, y, f) triples [nv][3]; finite edges as segments [ne][2][3]; count query with NULL */
int64_t dli_di2_get_all_vertices(const dli_di2 *d, double *xyz);
int64_t dli_di2_get_line_segments(const dli_di2 *d, double *segments);
/* get_line_segments as the reference returns it (:842-862): segments [ne][2][3], the is_edge_concave flag of every
   edge (:766-803) and the min / max of (x, y, f); evaluated on the GPU (dli_mesh_line_segments).  Returns ne (call
   with NULL arrays first), or a negative error. */
int64_t dli_di2_line_segments(dli_di2 *d, double *segments, int32_t *concave, double *min_coords, double *max_coords);

/*
 * dli_dd mirrors class Delaunay_incremental_interp_d (src/delaunay_d_interp.h:753-1142) for n_dims = 2, 3, 4.
 * The reference builds on its patched CGAL Delaunay_d (modified_cgal/Delaunay_d.h over modified_cgal/Convex_hull_d.h:
 * lower hull of the lifted points, Gmpq coordinates, history DAG); here the triangulation is native: Bowyer-Watson
 * insertion directly in d dimensions with an infinite vertex, exact orientation and in-sphere predicates (fp64 filter,
 * then floating-point expansions), vertex -> value stored with the vertex (m_pointMap :757).
 *   ctor(int) / ctor(int, fn)        :772-781   -> dli_dd_create
 *   get_dimension                    :783       -> dli_dd_dimension
 *   insert / insert_point            :933-956   -> dli_dd_insert (dirty set = simplices created / removed, the
 *                                                  "modified simplices" of modified_cgal/Convex_hull_d.h:1172-1178)
 *   interp / interp_point            :982-1010  -> dli_dd_interp (batched, on the GPU; NaN outside the hull)
 *   insert_largest_error_point       :1028-1034 -> dli_dd_insert_largest_error_point
 *   get_largest_error_tuple          :1036-1038 -> dli_dd_get_largest_error_tuple
 *   get_error_queue                  :1040-1043 -> dli_dd_get_error_queue
 *   get_all_vertices                 :1097-1101 -> dli_dd_get_all_vertices
 *   get_line_segments                :1103-1141 -> dli_dd_get_line_segments
 *   set_simplex_info / add_error_of_boundary_subsimplices / clear_simplex_info :832-916 (host, per inserted star)
 */
typedef struct dli_dd dli_dd;

/* n_dims in 2 .. DLI_MAX_DIM (NULL otherwise); fn may be NULL: no error queue is kept (:772-775) */
dli_dd *dli_dd_create(int n_dims, dli_fn fn, void *user);
void dli_dd_destroy(dli_dd *d);
int dli_dd_dimension(const dli_dd *d);
/* affine dimension of the points inserted so far (Delaunay_d::current_dimension): -1 none, n_dims once simplices exist */
int dli_dd_current_dimension(const dli_dd *d);
int64_t dli_dd_num_vertices(const dli_dd *d);
int64_t dli_dd_num_simplices(const dli_dd *d); /* finite simplices */

/* insert(x [n_dims], f): a duplicate point only overwrites its value.  Returns the vertex id or a negative error. */
int32_t dli_dd_insert(dli_dd *d, const double *x, double f);
/* bulk insertion; spatial_sort != 0 inserts in a biased randomised order along a Morton curve (same triangulation for
   points in general position, much faster).  With a callback the error queue is rebuilt afterwards (batched on the GPU
   through dli_simplex_error_* when use_gpu != 0). */
int dli_dd_insert_many(dli_dd *d, int64_t n, const double *xs, const double *f, int spatial_sort, int use_gpu);

/* DLI_ERR_ARG if the queue is empty or there is no callback (the reference asserts, :1029-1030) */
int dli_dd_insert_largest_error_point(dli_dd *d, double *err, double *f, double *x /* [n_dims] */);
int dli_dd_get_largest_error_tuple(dli_dd *d, double *err, double *f, double *x /* [n_dims] */);
int64_t dli_dd_error_queue_size(dli_dd *d);
/* up to cap live tuples ordered by decreasing error; xs [cap][n_dims]; returns the number written (or a negative error) */
int64_t dli_dd_get_error_queue(dli_dd *d, int64_t cap, double *err, double *f, double *xs);

/* flattened triangulation: coords [nv][n_dims], f [nv], sv [ns][n_dims + 1] (positively oriented), sn [ns][n_dims + 1]
   (k-th entry opposite vertex k, -1 = hull); any may be NULL */
int dli_dd_export(const dli_dd *d, double *coords, double *f, int32_t *sv, int32_t *sn);

/* batched interp.  After insertions the device mesh is brought in step lazily: patched in place through dli_mesh_update
   (stable compact simplex ids), re-created when it cannot be patched (dli_dd_sync_counts tells which happened how
   often).  Host or device pointers as dli_interp;
   queries outside the hull (or before n_dims + 1 affinely independent points exist) get a quiet NaN (:1005-1008) and
   DLI_STATUS_OUTSIDE_NAN / DLI_STATUS_NO_MESH.  Small batches of host pointers (one point at a time, as the reference's
   own loops call it) are answered on the host by an exact walk on the triangulation -- no GPU call; see dli_host_small.h. */
int dli_dd_interp(dli_dd *d, int64_t n, const double *xs, double *ys, int32_t *simplex_out, uint8_t *status_out,
                  void *stream);
const dli_mesh *dli_dd_device_mesh(dli_dd *d);
/* k-at-a-time adaptive refinement, the dD form of dli_di2_refine (same rules): up to k tuples on simplices
 * that share no vertex are inserted, then the items of every simplex created in the round are evaluated with ONE call of
 * fn_batch (NULL: the scalar callback of dli_dd_create per item).  err / f [k], xs [k][n_dims] receive the inserted
 * tuples (any may be NULL).  Returns the number of points inserted (0: empty queue), < 0 on error. */
int64_t dli_dd_refine(dli_dd *d, int64_t k, double rel_threshold, dli_fn_batch fn_batch, void *user, double *err, double *f,
                      double *xs);
/* (full re-flattens, in-place patches) of the device mesh so far: after insert / insert_largest_error_point the next
 * interp re-sends only the simplices the insertions created, the ring around them and the stars of overwritten
 * vertices (dli_mesh_update; the reference's dirty list: modified_cgal/Convex_hull_d.h:1172-1178 consumed at
 * src/delaunay_d_interp.h:939-965) */
void dli_dd_sync_counts(const dli_dd *d, int64_t *n_full, int64_t *n_incremental);
int dli_dd_set_options(dli_dd *d, const dli_options *opt);

/* plot data: vertices as (x..., f) rows [nv][n_dims + 1]; count query with NULL */
int64_t dli_dd_get_all_vertices(const dli_dd *d, double *out);
/* get_line_segments (:1103-1141): the consecutive vertex pairs (i, i + 1 mod n_dims + 1) of every simplex, each
   unordered pair once, as segments [ne][2][n_dims + 1]; min / max over the end points [n_dims + 1] (DBL_MAX / -DBL_MAX
   without segments).  Returns ne (call with NULL arrays first) or a negative error. */
int64_t dli_dd_get_line_segments(const dli_dd *d, double *segments, double *min_coords, double *max_coords);
/* test hook: 0 when the triangulation is structurally valid (and Delaunay, with check_delaunay != 0) */
int dli_dd_validate(const dli_dd *d, int check_delaunay);

#ifdef __cplusplus
}
#endif
#endif /* DLI_HOST_H_ */
