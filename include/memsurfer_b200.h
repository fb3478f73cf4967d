/* memsurfer_b200.h -- C ABI of libmemsurfer_b200.so (hand-written sm_100a CUDA).
 *
 * Drop-in replacement for the native half of MemSurfer's per-frame hot path.
 * Every entry point names the reference interface it replaces (paths under
 * the LLNL/MemSurfer tree).  Conventions:
 *   - all array arguments are DEVICE pointers (HBM) unless marked "host";
 *   - the caller owns every buffer, including the workspace `ws` whose size is
 *     returned by the matching *_workspace_bytes(); nothing is allocated or
 *     kept by the library (no global mutable state; re-entrant per stream);
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on
 *     it except where "syncs" is noted (a small device->host read-back); the
 *     per-frame hot path (point normals, mesh properties, KDE) never syncs;
 *   - return value: 0 = ok, non-zero = error (see ms_last_error(), a
 *     thread-local string).  Nothing throws across this boundary (the
 *     reference aborts the interpreter on C++ exceptions: memsurfer.i has no
 *     %exception);
 *   - arithmetic is fp64 on the float32-valued inputs; *_f32 outputs are the
 *     reference's dtypes (may be NULL), *_f64 the full-precision twins.
 */
#ifndef MEMSURFER_B200_H
#define MEMSURFER_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int         ms_version(void);
const char *ms_last_error(void);

/* ---------------------------------------------------------------- point normals
 * Replaces memsurfer_cmod.PointSet: set_periodic (PointSet.cpp:94-147),
 * need_normals -> CGAL::pca_estimate_normals + CGAL::mst_orient_normals
 * (PointSet.cpp:156-200), get_normals (PointSet.cpp:70-88) and the global flip
 * of Membrane.compute_pnormals (membrane.py:145-152).                          */

/* Nothing in stages 1-4 is read back by the host: the number of points incl. ghost images stays on the device in
 * an 8 x int64 status block `info` that the caller allocates and threads through the chain (arrays are sized by
 * ms_pointset_capacity, kernels are launched over the capacity and leave early).  The caller reads the block
 * whenever it reads results.                                                                        */
#define MS_INFO_M           0   /* points incl. ghost images (written by ms_ghosts)                  */
#define MS_INFO_UNORIENTED  1   /* points not connected to the source (ms_orient_normals)            */
#define MS_INFO_KNN_TIES    2   /* diagnostic: k-NN queries redone in exact (distance, index) order  */
#define MS_INFO_MST_ROUNDS  3   /* diagnostic: Boruvka rounds                                         */
#define MS_INFO_ERROR       4   /* bit flags, 0 = ok                                                  */
#define MS_INFO_KNN_CANDS   5   /* diagnostic: candidate distances the k-NN evaluated                */
#define MS_INFO_WORDS       8
#define MS_ERR_NONFINITE    1   /* non-finite coordinates: every neighbour list is -1                */

/* Replaces the host scan of Membrane.fit_points_to_box_xy (membrane.py:100-121): *count_dev (device int64) = number
 * of x / y coordinates outside [box0, box1] (box6 on the host), i.e. how many coordinates the wrap would move.  The
 * usual answer is 0, and then the host never has to walk the points.                                            */
int ms_points_outside_box(const float *points, int64_t n, const float *box6, int64_t *count_dev, void *stream);

/* Upper bound of points incl. ghost images (each point gets <= 3 images). */
int64_t ms_pointset_capacity(int64_t n, int periodic);

/* Stage 1: ghost images.  points (n,3) f32 -> points_all (cap,3) f32 = points ++ ghosts in the
 * reference's order; ghost_src (cap,) i32 = original index of every row.  box6 (host) =
 * {b0x,b0y,b0z,b1x,b1y,b1z} as float32; blayer in (0,1).  periodic == 0 copies the points only.
 * info (device, MS_INFO_WORDS x int64): [MS_INFO_M] = n + #ghosts, the other words are cleared.       */
size_t ms_ghosts_workspace_bytes(int64_t n);
int ms_ghosts(const float *points, int64_t n, int periodic, const float *box6, float blayer,
              float *points_all, int32_t *ghost_src, int64_t *info,
              void *ws, size_t ws_bytes, void *stream);

/* Stage 2: exact (k1 = k+1)-nearest neighbours incl. self for the first m points of points_all (m_cap rows
 * allocated; m = info[MS_INFO_M] read on the device, or m_cap when info == NULL), sorted by (squared distance
 * in fp64, index), via a uniform xy cell list whose geometry is worked out on the device.  nbr (m_cap,k1) i32. */
size_t ms_knn_workspace_bytes(int64_t m_cap, int k1);
int ms_knn(const float *points_all, int64_t m_cap, int64_t *info, int k1, int32_t *nbr,
           void *ws, size_t ws_bytes, void *stream);

/* Stage 3: PCA normal of every point from its k1 neighbours (unit eigenvector of the smallest
 * eigenvalue of the 3x3 covariance; sign arbitrary).  normals (m_cap,3) f64.                      */
int ms_pca_normals(const float *points_all, int64_t m_cap, const int64_t *info, const int32_t *nbr, int k1,
                   double *normals, void *stream);

/* Stage 4: minimum-spanning-tree orientation over the Riemannian graph (edge {a<b} iff b in
 * NN(a), weight float(max(0,1-|na.nb|)), ties by (w,a,slot)); source = first max-z point flipped
 * to n_z >= 0.  normals updated in place.  info[MS_INFO_UNORIENTED] = points not connected to the
 * source.  One cooperative launch; the number of Boruvka rounds is decided on the device.         */
size_t ms_orient_workspace_bytes(int64_t m_cap, int k1);
int ms_orient_normals(const float *points_all, int64_t m_cap, int64_t *info, const int32_t *nbr, int k1,
                      double *normals, void *ws, size_t ws_bytes, void *stream);

/* Diagnostic (no reference counterpart): nanoseconds spent in every phase of the last ms_orient_normals call on
 * workspace `ws` (init, round 0 edges, round 0 hooks, then edges / hooks of every later round, orientation):
 * ns_out (host, cap doubles), *nphases = entries written.  Syncs the stream.                          */
int ms_orient_phase_times(const void *ws, size_t ws_bytes, int64_t m_cap, int k1, double *ns_out, int cap, int *nphases, void *stream);

/* Stage 5: first n normals -> float32, flipped globally if the z of the normal with the largest
 * |n_z| (float32) times ndir_hint is negative (membrane.py:148-152).                            */
size_t ms_finish_workspace_bytes(int64_t n);
int ms_finish_pnormals(const double *normals, int64_t n, int ndir_hint,
                       float *pnormals_f32, double *pnormals_f64,
                       void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- mesh properties
 * Replaces memsurfer_cmod.TriMesh.need_normals / need_pointareas (TriMesh.hpp:346-407,
 * TriMesh.cpp:99-190) and the vtkCurvatures calls of TriMesh.compute_curvatures
 * (trimesh.py:284-316, mesh assembly + half-box face filter trimesh.py:441-500).
 *
 * verts_all (nv_all,3) f32 = vertices ++ duplicated verts; faces_all (nf,3) u32 = faces ++
 * trimmed faces (the concatenation of TriMesh.hpp:360-364); outputs cover the first nv rows.   */

/* Vertex -> incident (face,corner) lists in ascending order (CSR) + the per-face keep mask of
 * the half-box filter (half_xw/half_yw < 0 disables it).  csr_off (nv+1) i32,
 * csr_adj (3*nf) i32 packed face*4+corner, keep (nf) u8.                                        */
size_t ms_mesh_topology_workspace_bytes(int64_t nv, int64_t nf);
int ms_mesh_topology(const float *verts_all, int64_t nv_all, int64_t nv,
                     const uint32_t *faces_all, int64_t nf, double half_xw, double half_yw,
                     int32_t *csr_off, int32_t *csr_adj, uint8_t *keep,
                     void *ws, size_t ws_bytes, void *stream);

/* Area-weighted vertex normals, mixed Voronoi point areas, mean (dihedral) and Gaussian (angle-deficit)
 * curvature of the first nv vertices: one pass per face (everything that depends on a single face, one
 * 128-byte record each), one pass per face edge (mean-curvature edge terms) and one deterministic
 * gather per vertex.  Any output may be NULL.                                                     */
size_t ms_mesh_properties_workspace_bytes(int64_t nv, int64_t nf);
int ms_mesh_properties(const float *verts_all, int64_t nv_all, int64_t nv,
                       const uint32_t *faces_all, int64_t nf,
                       const int32_t *csr_off, const int32_t *csr_adj, const uint8_t *keep,
                       double *vnormals_f64, float *vnormals_f32,
                       double *pareas_f64, float *pareas_f32,
                       double *mean_curv, double *gaus_curv,
                       void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- KDE densities
 * Replaces memsurfer_cmod.TriMesh.kde types 2 and 3 (TriMesh_kde.cpp:344-392, 443-532) with
 * GaussianKernel2D/3D (DensityKernels.hpp:44-84) and DistanceSquared /
 * DistancePeriodicXYSquared (DistanceKernels.hpp:75-136) fused in.
 * verts (nv,3) f32 = the mesh's own vertices (targets: all of them); ids (nids) i32 = source
 * vertex ids, nids == 0 -> all.  type 2|3.  periodic: box6 (host) = bboxl.  cutoff_sigmas <= 0:
 * all pairs (the reference's sum); > 0: only pairs inside a window of cells that covers every pair
 * closer than cutoff_sigmas*sigma are evaluated (cell list; dropped relative tail mass
 * <= exp(-cutoff_sigmas^2/2); needs box6 -- for a non-periodic mesh pass the vertex extent -- and
 * falls back to all pairs when the window would not fit the period).  get_counts as
 * TriMesh_kde.cpp:515-524.                                                                     */
size_t ms_kde_workspace_bytes(int64_t nv, int64_t nids);
int ms_kde(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
           int type, int periodic, const float *box6, float sigma, int get_counts,
           double cutoff_sigmas, double *density_f64, float *density_f32,
           void *ws, size_t ws_bytes, void *stream);

/* Label-fused KDE.  Membrane.compute_densities (membrane.py:289-297) asks for one field per lipid
 * type plus 'all' (examples/ex_3lipid.py:169-175); the reference runs kde() once per field, i.e.
 * every vertex pair is evaluated twice.  ms_kde_labels evaluates each pair once and keeps the sums
 * apart by the SOURCE's label: labels (nv) i32 in [0, nlabels) (negative: not a source),
 * raw (nlabels, nv) f64 with raw[l*nv + j] = sfactor * sum_{i: labels[i]==l} exp(efactor*dist2(v_i, v_j))
 * (TriMesh_kde.cpp:344-392 before its normalisations).  nlabels <= 64.                           */
size_t ms_kde_labels_workspace_bytes(int64_t nv, int nlabels);
int ms_kde_labels(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                  int type, int periodic, const float *box6, float sigma, double cutoff_sigmas,
                  double *raw, void *ws, size_t ws_bytes, void *stream);
/* The normalisations of TriMesh::kde (TriMesh_kde.cpp:504-524) for the union of the labels with
 * select[l] != 0 (host, nlabels bytes): density = sum_l raw[l] / nv; get_counts: * n_selected /
 * sum(density), n_selected = number of source vertices of the selected labels.                  */
size_t ms_kde_select_workspace_bytes(int64_t nv);
int ms_kde_select(const double *raw, int64_t nv, int nlabels, const uint8_t *select, double n_selected,
                  int get_counts, double *density_f64, float *density_f32,
                  void *ws, size_t ws_bytes, void *stream);

/* ms_kde_select without a workspace and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, kernels and results as ms_kde_select; the scratch is a
 * static device array inside the library, held by every GPU that loads it until exit (8 MiB + 32 KiB), so
 * nv <= MS_KDE_SELECT_STATIC_MAX_NV -- larger sizes are rejected (status 1, ms_last_error says so).  All calls on a
 * device share that array: calls that overlap on the GPU (two streams, two threads) race; calls on one stream, or
 * otherwise ordered, do not.  Nothing has to be cleared between calls.  The device of `stream` must be current.  */
#define MS_KDE_SELECT_STATIC_MAX_NV 1048576
int ms_kde_select_static(const double *raw, int64_t nv, int nlabels, const uint8_t *select, double n_selected,
                         int get_counts, double *density_f64, float *density_f32, void *stream);

/* Any number of fields of one call in two launches (instead of ms_kde_select's two per field).  raw = nrows rows of nv
 * raw sums (e.g. the (nsigma, nlabels, nv) block of ms_kde_labels_multi: nrows = nsigma * nlabels).
 * ms_kde_label_totals: partials (ms_kde_label_totals_bytes(nrows) bytes) = fixed-order partial sums of every row.
 * ms_kde_fields: field f = the union of the labels in masks[f] (bit l = label l) of the sigma whose first row is
 * row0[f], normalised like TriMesh::kde (TriMesh_kde.cpp:504-524) with n_selected[f] source vertices; row0, masks,
 * n_selected are HOST arrays of nfields (<= 64) entries; outputs (nfields, nv), either may be NULL.                */
size_t ms_kde_label_totals_bytes(int64_t nrows);
int ms_kde_label_totals(const double *raw, int64_t nrows, int64_t nv, double *partials, void *stream);
int ms_kde_fields(const double *raw, const double *partials, int64_t nv, int nlabels, int nfields,
                  const int32_t *row0, const uint64_t *masks, const double *n_selected, int get_counts,
                  double *out_f64, float *out_f32, void *stream);

/* Several sigmas in one call: raw (nsigma, nlabels, nv) = the ms_kde_labels result of every sigma (host array
 * `sigmas`).  With a cut-off whose window plan fits every sigma the vertices are binned ONCE (on the grid of the
 * smallest sigma) and only re-staged per sigma; otherwise a loop over ms_kde_labels.  This is the loop
 * `for s in sigmas` of Membrane.compute_densities (membrane.py:289-297) hoisted below the ABI.            */
size_t ms_kde_labels_multi_workspace_bytes(int64_t nv, int nlabels);
int ms_kde_labels_multi(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                        int type, int periodic, const float *box6, const float *sigmas, int nsigma, double cutoff_sigmas,
                        double *raw, void *ws, size_t ws_bytes, void *stream);
/* Diagnostic twin: pairs (host, 2 * nsigma doubles): pairs[i] = pairs the launch of sigma i evaluates, pairs[nsigma + i] =
 * those of them summed in fp64 (the rest, beyond 5 sigma, is summed in fp32).  Syncs the stream. */
int ms_kde_labels_multi_pairs(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                              int type, int periodic, const float *box6, const float *sigmas, int nsigma, double cutoff_sigmas,
                              double *pairs, void *ws, size_t ws_bytes, void *stream);

/* ms_kde_labels_multi that also emits, sigma by sigma, the normalised fields of `nsets` label sets (masks[f]: bit l =
 * label l; n_selected[f] source vertices; HOST arrays) -- what Membrane.compute_densities (membrane.py:289-297) asks
 * for in its loop over 'all' and every lipid type, normalised like TriMesh::kde (TriMesh_kde.cpp:504-524).  The fields
 * of a sigma are written right after that sigma's pair sums: out_f32 / out_f64 (nsigma, nsets, nv), out_f64 may be
 * NULL; partials = ms_kde_label_totals_bytes(nsigma * nlabels) bytes.  If host_f32 (page-locked, same shape) is
 * given, every sigma's block is also copied there on `copy_stream` (a stream other than `stream`), ordered after its
 * fields kernel: the downloads run under the next sigma's pair kernel.  The caller synchronises with copy_stream.   */
int ms_kde_labels_multi_fields(const float *verts, int64_t nv, const int32_t *labels, int nlabels,
                               int type, int periodic, const float *box6, const float *sigmas, int nsigma,
                               double cutoff_sigmas, double *raw, int nsets, const uint64_t *masks,
                               const double *n_selected, int get_counts, double *partials, double *out_f64,
                               float *out_f32, float *host_f32, void *copy_stream,
                               void *ws, size_t ws_bytes, void *stream);

/* Diagnostic (roofline accounting): the number of pairs ms_kde (labels == NULL) or ms_kde_labels
 * evaluates for the same arguments -- nv*n_sources for the all-pairs kernel, the sum over cells of
 * targets x window sources for the cell list.  pairs = TWO host doubles: all evaluated pairs, and those of them summed
 * in fp64 (the cell-list kernel sums pairs beyond 5 sigma in fp32); ws sized by the matching *_workspace_bytes.
 * Syncs the stream.                                                          */
int ms_kde_evaluated_pairs(const float *verts, int64_t nv, const int32_t *ids, int64_t nids,
                           const int32_t *labels, int nlabels,
                           int type, int periodic, const float *box6, float sigma, double cutoff_sigmas,
                           double *pairs, void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- overlapping points
 * Replaces the detection part of utils.test_overlapping_points (memsurfer/utils.py:91-159: scipy pdist
 * over all pairs < dthreshold).  points (m,3) f32 (2-D callers pad z = 0); pairs (max_pairs,2) i32
 * receives index pairs i<j in no particular order; *npairs (host) = number of pairs closer than
 * dthreshold (may exceed max_pairs); *min_above (host, may be NULL) = smallest distance > dthreshold
 * among the scanned neighbours (+inf if none).  Syncs the stream.                                        */
size_t ms_overlap_workspace_bytes(int64_t m);
int ms_overlapping_pairs(const float *points, int64_t m, double dthreshold, int32_t *pairs, int64_t max_pairs,
                         int64_t *npairs, double *min_above, void *ws, size_t ws_bytes, void *stream);

/* Same search without any read-back: box4 (host, may be NULL = measure on the device, which syncs) = xy extent for
 * the cell list {xmin, ymin, xmax, ymax}; result_dev (device, 2 x uint64) receives {pair count, IEEE bits of
 * the smallest distance above the threshold (+inf if none)}.  Asynchronous when box4 is given.          */
int ms_overlapping_pairs_async(const float *points, int64_t m, double dthreshold, const float *box4, int32_t *pairs,
                               int64_t max_pairs, unsigned long long *result_dev, void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- closest point / thickness
 * Replaces TriMesh::project_on_surface (memsurfer/src/TriMesh_cgal.cpp:295-363: CGAL AABB tree over the
 * mesh's own faces, closest_point_and_primitive per query, barycentric coordinates by TriMesh::Point2Bary
 * TriMesh.cpp:28-42) and TriMesh::distance_to_other_mesh (TriMesh_cgal.cpp:390-425; Euclidean, no
 * periodic wrap).  verts (nv,3) f32, faces (nf,3) u32 (the mesh's own faces only), box4 (host) = xy
 * extent of the vertices {xmin, ymin, xmax, ymax}; queries (nq,3) f32.  Outputs (each may be NULL):
 * face_id (nq) i32 = closest face (ties: smallest id), bary (nq,3) f64 = barycentric coordinates of the
 * closest point, dist_f64 / dist_f32 (nq) = distance.  Syncs the stream once (list size check).        */
size_t ms_closest_point_workspace_bytes(int64_t nv, int64_t nf);
int ms_closest_point(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf, const float *box4,
                     const float *queries, int64_t nq, int32_t *face_id, double *bary,
                     double *dist_f64, float *dist_f32, void *ws, size_t ws_bytes, void *stream);

/* fp64 FMA peak probe used for the KDE roofline: runs `iters` dependent-chain DFMA rounds on
 * every SM and returns the number of FLOP issued in *flop (host); time it with events.         */
int ms_dfma_probe(double *out, int64_t iters, double *flop, void *stream);

/* ---------------------------------------------------------------- per-frame summaries
 * No reference counterpart: the reference analyses one frame per process (examples/ex_3lipid.py:78)
 * and writes whole fields to .vtp (membrane.py:319-344).  When trajectory frames are sharded over
 * GPUs, ranks exchange only this record per field: out4 (device, nfields x 4 doubles) = {sum, sum|x|,
 * min, max} of each of the nfields (<= 64) fp64 fields of n values, fields = HOST array of device
 * pointers; fixed reduction order (bit-reproducible); two launches for the whole batch.           */
size_t ms_field_summary_workspace_bytes(int nfields);
int ms_field_summary(const double *const *fields, int nfields, int64_t n, double *out4,
                     void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- vertex neighbours / shells
 * ms_mesh_neighbors replaces TriMesh::need_neighbors (memsurfer/src/TriMesh.cpp:197-255): per vertex the
 * distinct vertices it shares a face with, in the order the reference's sequential face loop appends them
 * (ascending face, corner; next then next-next vertex; first appearance wins).  faces (nf,3) u32 = the mesh's
 * own faces (corners >= nv are ignored), nbr_off (nv+1) i32 offsets, nbr i32 with room for 6*nf entries (may be
 * NULL: offsets only); *total_out = nbr_off[nv].  Synchronises the stream once.
 * ms_mesh_shells replaces the breadth-first loop of TriMesh.compute_shells (memsurfer/trimesh.py:364-416):
 * shells[v] = hop distance to the nearest of the nref reference vertices (-1: not connected to any);
 * ref (nref) i32 on the device.  *n_unreached = vertices left at -1, *max_shell = largest shell.        */
size_t ms_mesh_neighbors_workspace_bytes(int64_t nv, int64_t nf);
int ms_mesh_neighbors(const uint32_t *faces, int64_t nf, int64_t nv, int32_t *nbr_off, int32_t *nbr, int64_t *total_out,
                      void *ws, size_t ws_bytes, void *stream);
size_t ms_mesh_shells_workspace_bytes(int64_t nv);
int ms_mesh_shells(const int32_t *nbr_off, const int32_t *nbr, int64_t nv, const int32_t *ref, int64_t nref,
                   int32_t *shells, int64_t *n_unreached, int32_t *max_shell, void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- periodic Delaunay (first slice)
 * Replaces TriMesh::periodicDelaunay (memsurfer/src/TriMesh_cgal.cpp:160-281; CGAL periodic triangulation,
 * 9-sheeted covering): the faces with at least one vertex in the original sheet, counter-clockwise, as
 * dfaces (cap,3) i32 original vertex ids and doffs (cap,3,2) i8 periodic offsets in {-1,0,1} (the
 * PeriodicVertex triples of :223-233).  Inputs: points_all (m,3) f32 = the n planar points (z ignored) followed
 * by their ghost images and ghost_src (m) i32 from ms_ghosts; nbr (m,k1) i32 from ms_knn on points_all with
 * z = 0; box widths; ghost_layer = width of the ghost layer (blayer * the smaller box width).  Faces are
 * ordered by their smallest vertex id, then counter-clockwise around it.  *n_incomplete = points whose Voronoi
 * cell cannot be closed inside the ghost layer, or -1 when the stars do not add up to a triangulation of the
 * torus (6n corners: exactly cocircular points, e.g. an unperturbed lattice); nothing is written then.
 * *nfaces = faces found.  dfaces == NULL: count only.  Syncs the stream once.     */
size_t ms_delaunay_periodic_workspace_bytes(int64_t n);
int ms_delaunay_periodic(const float *points_all, int64_t m, int64_t n, const int32_t *ghost_src,
                         const int32_t *nbr, int k1, double box_wx, double box_wy, double ghost_layer,
                         int32_t *dfaces, int8_t *doffs, int64_t cap, int64_t *nfaces, int64_t *n_incomplete,
                         void *ws, size_t ws_bytes, void *stream);

/* First step of TriMesh::trim_periodicDelaunay (memsurfer/src/TriMesh.cpp:493-575): the faces whose three corners
 * are in the original sheet (all offsets zero) compacted in order into full (nf,3) u32 = mFaces; the others,
 * also in order, into pfaces (nf,3) i32 / poffs (nf,3,2) i8 for the duplicated-vertex bookkeeping.  All three
 * outputs sized for nf faces; *nfull = faces in `full` (the rest, nf - *nfull, in pfaces).  Syncs once.    */
size_t ms_delaunay_split_workspace_bytes(int64_t nf);
int ms_delaunay_split(const int32_t *dfaces, const int8_t *doffs, int64_t nf, uint32_t *full, int32_t *pfaces, int8_t *poffs,
                      int64_t *nfull, void *ws, size_t ws_bytes, void *stream);

/* Second step of TriMesh::trim_periodicDelaunay (TriMesh.cpp:513-560): the crossing faces (pfaces / poffs from
 * ms_delaunay_split) re-indexed onto duplicated vertices numbered in order of first appearance in the (face, corner)
 * scan, as the reference's std::find / push_back loop numbers them.  corner_verts (3 nf, dim) f32 = this mesh's coordinates of
 * the original vertex of every corner, dim 2|3; nv = number of original vertices; box widths of THIS mesh as float32.  tfaces (nf,3) u32, dup_ids (3 nf) u32 = original vertex of every duplicate, dup_verts (3 nf, dim) f32 =
 * float(double(v) + double(float(off) * boxw)) (TriMesh.cpp:547-549); *ndup (host).  Syncs the stream once.       */
size_t ms_delaunay_duplicates_workspace_bytes(int64_t nfaces);
int ms_delaunay_duplicates(const float *corner_verts, int64_t nv, int dim, const int32_t *pfaces, const int8_t *poffs, int64_t nf,
                           float box_wx, float box_wy, uint32_t *tfaces, uint32_t *dup_ids, float *dup_verts, int64_t *ndup,
                           void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- geodesic (type 1) KDE
 * ms_geodesics replaces TriMesh::compute_geodesics_fw (memsurfer/src/TriMesh_kde.cpp:196-332): all-pairs
 * shortest paths along the edges of `faces` (for a periodic mesh: its faces followed by its periodic faces,
 * :487-489), edge length = sqrt of the periodic-xy 3-D squared distance (DistanceKernels.hpp:120-135).
 * D: pitch x pitch fp64 on the device, pitch = ms_geodesic_pitch(nv) (rows/columns >= nv are padding);
 * unreachable pairs hold DBL_MAX (FLT_MAX in the reference).  nv <= 46 000 (17 GB).
 * ms_kde_geodesic replaces kde_2m (:394-437) + the normalisation of TriMesh::kde (:504-524) with
 * GaussianKernel2D(sigma): all vertices (nids == 0): sum_i k(D[i][j]^2); selected ids: sum_i k(D[ids[i]][j])
 * -- the reference does not square the distance in that branch (:432) and neither does this.           */
int64_t ms_geodesic_pitch(int64_t nv);
int ms_geodesics(const float *verts, int64_t nv, const uint32_t *faces, int64_t nf, int periodic, const float *box6,
                 double *D, void *stream);
size_t ms_kde_geodesic_workspace_bytes(int64_t nv);
int ms_kde_geodesic(const double *D, int64_t nv, const int32_t *ids, int64_t nids, float sigma, int get_counts,
                    double *density_f64, float *density_f32, void *ws, size_t ws_bytes, void *stream);

/* ---------------------------------------------------------------- diagnostics
 * No reference counterpart (the reference times whole steps with utils.Timer, utils.py:209-241).
 * ms_launch_count: kernels launched by the library since the last reset (process-wide).
 * ms_profile_enable(1): from now on this thread records a CUDA event before every kernel it
 * launches; ms_profile_report writes {"kernel": {"launches": n, "ms": t}, ...} (device time
 * between consecutive marks, summed per kernel name) into buf (host) and clears the marks.   */
int64_t ms_launch_count(int reset);
int     ms_profile_enable(int on);
int     ms_profile_report(char *buf, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
