/* mundy_b200.h — C-ABI of the B200-native collision hot path (sm_100a).
 *
 * This header is the drop-in boundary: each entry point replaces one seam of the reference
 * (MundyRepo/MuNDy @ 797a7f48; paths below are relative to the reference root). The reference has no FFI for this path
 * (header-only C++20 templates over Kokkos), so the boundary is "what a maintainer would bind": plain pointers and
 * sizes, no C++/torch types. INTEGRATION.md shows the reference-side stubs; include/mundy_b200/ holds the C++ veneer
 * that keeps the reference's call signatures on top of these functions.
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer unless its name ends in _h (host) or the doc says otherwise.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream). Calls are asynchronous with respect
 *     to the host except where a host-visible result is returned (documented per function).
 *   - Layouts are the reference's: vectors are xyz-interleaved (3 doubles per item, `View<double*[3]>` LayoutRight /
 *     STK field layout), quaternions are (w,x,y,z) (mundy/math/src/mundy_math/Quaternion.hpp), AABBs are
 *     (min xyz, max xyz) (mundy/geom/src/mundy_geom/primitives/AABB.hpp:199-202), pairs are int32 (i,j).
 *   - All arithmetic is fp64 with the reference's operation order and no FMA contraction.
 *   - Return value: 0 = MB200_OK, negative = error (see enum); mb200_last_error_string() describes the last error of
 *     the calling thread. Nothing throws or aborts across this boundary.
 *   - Thread-safety: a handle is single-threaded; distinct handles / handle-free calls may be used concurrently.
 *   - There is no CPU fallback: without a CUDA device every compute call returns MB200_ERR_CUDA.
 */
#ifndef MUNDY_B200_H_
#define MUNDY_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define MB200_ABI_VERSION 1

enum mb200_status {
  MB200_OK = 0,
  MB200_ERR_INVALID_ARGUMENT = -1,
  MB200_ERR_CUDA = -2,
  MB200_ERR_CAPACITY = -3,       /* an output buffer supplied by the caller is too small */
  MB200_ERR_NOT_CONVERGED = -4,  /* only returned where the reference would throw (see mb200_collision_step) */
  MB200_ERR_COMM = -5
};

/* Body kinds. Sphere: centre + radius. Spherocylinder: centre + quaternion + radius + length, axis = q (x) z_hat
 * (mundy/geom/src/mundy_geom/compute_aabb.hpp:115-117). */
enum mb200_body_kind { MB200_SPHERE = 0, MB200_SPHEROCYLINDER = 1 };

typedef void* mb200_stream;

int mb200_abi_version(void);
const char* mb200_last_error_string(void);
/* Number of visible CUDA devices (0 when none / no driver). Host-only query. */
int mb200_device_count(void);

/* Device-memory helpers for hosts that do not link the CUDA runtime themselves (FFI callers, the C++ veneer's
 * single-object convenience overloads). Thin wrappers over cudaMalloc / cudaFree / cudaMemcpyAsync / stream sync. */
int mb200_device_alloc(void** out, int64_t bytes);
int mb200_device_free(void* ptr);
int mb200_copy_to_device(mb200_stream stream, void* dst, const void* src_h, int64_t bytes);
int mb200_copy_to_host(mb200_stream stream, void* dst_h, const void* src, int64_t bytes);
int mb200_stream_synchronize(mb200_stream stream);

/* ---------------------------------------------------------------------------------------------------------------
 * A. geom::compute_aabb, batched.   mundy/geom/src/mundy_geom/compute_aabb.hpp
 * ------------------------------------------------------------------------------------------------------------- */
/* compute_aabb(Sphere) :72-80.  center[3n], radius[n] -> aabb[6n] */
int mb200_compute_aabb_sphere(mb200_stream stream, int64_t n, const double* center, const double* radius,
                              double* aabb);
/* compute_aabb(Spherocylinder) :106-127.  quat[4n] = (w,x,y,z), not assumed unit-norm. */
int mb200_compute_aabb_spherocylinder(mb200_stream stream, int64_t n, const double* center, const double* quat,
                                      const double* radius, const double* length, double* aabb);
/* compute_aabb(SpherocylinderSegment) :130-143.  start[3n], end[3n]. */
int mb200_compute_aabb_segment(mb200_stream stream, int64_t n, const double* start, const double* end,
                               const double* radius, double* aabb);

/* ---------------------------------------------------------------------------------------------------------------
 * C. geom::distance (SharedNormalSigned), batched: item k of every array is one call of the reference function.
 *    Output pointers may be NULL (that output is skipped).
 * ------------------------------------------------------------------------------------------------------------- */
/* distance(Sphere, Sphere, sep)  mundy/geom/src/mundy_geom/distance/SphereSphere.hpp:54-76 */
int mb200_distance_sphere_sphere(mb200_stream stream, int64_t n, const double* c1, const double* r1, const double* c2,
                                 const double* r2, double* dist, double* sep);
/* distance(LineSegment, LineSegment, cp1, cp2, s, t, sep)  distance/LineSegmentLineSegment.hpp:189-318 */
int mb200_distance_segment_segment(mb200_stream stream, int64_t n, const double* l0, const double* l1,
                                   const double* m0, const double* m1, double* dist, double* s, double* t,
                                   double* cp1, double* cp2, double* sep);
/* distance(Point, LineSegment, cp, t, sep)  distance/PointLineSegment.hpp:128-172 (t unclamped, cp clamped) */
int mb200_distance_point_segment(mb200_stream stream, int64_t n, const double* x, const double* p1, const double* p2,
                                 double* dist, double* t, double* cp, double* sep);
/* distance(LineSegment, Sphere, cp, t, sep)  distance/LineSegmentSphere.hpp:86-99 */
int mb200_distance_segment_sphere(mb200_stream stream, int64_t n, const double* p1, const double* p2, const double* c,
                                  const double* r, double* dist, double* t, double* cp, double* sep);

/* ---------------------------------------------------------------------------------------------------------------
 * A'/C'. The remaining primitives of mundy_geom, batched like the sections above.
 *    quat[4n] = (w,x,y,z), not assumed unit-norm; radii[3n] = the three semi-axes in the body frame; a Line is
 *    (center[3n], direction[3n]). Output pointers other than aabb may be NULL.
 * ------------------------------------------------------------------------------------------------------------- */
/* compute_aabb(Point)  mundy/geom/src/mundy_geom/compute_aabb.hpp:47-54 */
int mb200_compute_aabb_point(mb200_stream stream, int64_t n, const double* point, double* aabb);
/* compute_aabb(LineSegment)  compute_aabb.hpp:57-69 */
int mb200_compute_aabb_line_segment(mb200_stream stream, int64_t n, const double* start, const double* end,
                                    double* aabb);
/* compute_aabb(Ellipsoid)  compute_aabb.hpp:83-103: corners center -/+ (orientation * radii), per-axis min / max */
int mb200_compute_aabb_ellipsoid(mb200_stream stream, int64_t n, const double* center, const double* quat,
                                 const double* radii, double* aabb);
/* distance(Point, Sphere, sep)  distance/PointSphere.hpp:68-76 */
int mb200_distance_point_sphere(mb200_stream stream, int64_t n, const double* x, const double* c, const double* r,
                                double* dist, double* sep);
/* distance(Point, Line, cp, arch_length, sep)  distance/PointLine.hpp:92-103 (the reference leaves sep unwritten) */
int mb200_distance_point_line(mb200_stream stream, int64_t n, const double* x, const double* line_center,
                              const double* line_dir, double* dist, double* cp, double* arch);
/* distance(Line, Line, cp1, cp2, arch1, arch2, sep)  distance/LineLine.hpp:131-157. Parallel lines (D < 1e-15): the
 * reference returns 0 with sep = 0 and leaves cp/arch untouched; here those outputs are written as 0. */
int mb200_distance_line_line(mb200_stream stream, int64_t n, const double* c1, const double* d1, const double* c2,
                             const double* d2, double* dist, double* cp1, double* cp2, double* arch1, double* arch2,
                             double* sep);
/* distance(Line, Sphere, cp, arch_length, sep)  distance/LineSphere.hpp:93-102 */
int mb200_distance_line_sphere(mb200_stream stream, int64_t n, const double* line_center, const double* line_dir,
                               const double* c, const double* r, double* dist, double* cp, double* arch);
/* distance(SharedNormalSigned, Point, Ellipsoid, closest_point, ellipsoid_normal)  distance/PointEllipsoid.hpp:97-135.
 * Found by math::find_min_using_approximate_derivatives (mundy/math/src/mundy_math/minimize.hpp:41-50: L-BFGS memory
 * 10, central differences, Wolfe line search) over the normal's two angles from nine starts, like the next two. */
int mb200_distance_point_ellipsoid(mb200_stream stream, int64_t n, const double* x, const double* center,
                                   const double* quat, const double* radii, double* dist, double* cp, double* normal);
/* distance(SharedNormalSigned, Ellipsoid, Ellipsoid, cp1, cp2, normal1, normal2)  distance/EllipsoidEllipsoid.hpp:109-151
 * (normal2 = -normal1 is not stored) */
int mb200_distance_ellipsoid_ellipsoid(mb200_stream stream, int64_t n, const double* c1, const double* q1,
                                       const double* radii1, const double* c2, const double* q2,
                                       const double* radii2, double* dist, double* cp1, double* cp2, double* normal1);
/* distance(Euclidean, Circle3D, Circle3D, cp1, cp2, normal1, normal2)  distance/Circle3DCircle3D.hpp:88-151 */
int mb200_distance_circle3d_circle3d(mb200_stream stream, int64_t n, const double* c1, const double* q1,
                                     const double* r1, const double* c2, const double* q2, const double* r2,
                                     double* dist, double* cp1, double* cp2, double* normal1);

/* ---------------------------------------------------------------------------------------------------------------
 * Body set in the caller's (reference) layout. kind/quat/length may be NULL when every body is a sphere.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct mb200_bodies {
  int64_t n;
  double* center;        /* [3n]  in/out for the step entry points, read-only elsewhere */
  double* quat;          /* [4n]  (w,x,y,z) */
  const double* radius;  /* [n] */
  const double* length;  /* [n]   spherocylinder centre-line length */
  const int32_t* kind;   /* [n]   mb200_body_kind */
} mb200_bodies;

/* ---------------------------------------------------------------------------------------------------------------
 * C5/K6. Narrow phase on a pair list: signed separation phi0, unit normal (from i to j), lever arms (centre-line
 * closest point minus centre; zero for spheres). Restates the constraint initialisation of
 * scrap/lcp_spheres/KokkosLCP.cpp:449-491 (spheres) and the composition of
 * scrap/parameter_interface/linkers/.../SpherocylinderSpherocylinderLinker.cpp:207-247 with geom::distance (rods).
 * pairs[2*npairs]; phi[npairs]; normal[3*npairs]; arm_i/arm_j[3*npairs] may be NULL.
 * ------------------------------------------------------------------------------------------------------------- */
int mb200_narrow_phase(mb200_stream stream, const mb200_bodies* bodies, int64_t npairs, const int32_t* pairs,
                       double* phi, double* normal, double* arm_i, double* arm_j);

/* ---------------------------------------------------------------------------------------------------------------
 * B. Broad phase = mesh::GenNeighborLinks (mundy/mesh/src/mundy_mesh/GenNeighborLinkers.hpp:294-543), whose search
 * call stk::search::coarse_search(MORTON_LBVH) (:650-664) is replaced by a Morton-sorted cell list on the device.
 * Candidate predicate (fp64, exact): i<j and |c_j - c_i|^2 <= (S_i + S_j)^2, S = bounding radius + search buffer
 * (:575-584); bounding radius = r (sphere) or L/2 + r (spherocylinder; geom/compute_bounding_radius.hpp:74-82).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct mb200_neighbor_links mb200_neighbor_links; /* opaque; owns the pair list and the rebuild state */

/* Search filters (GenNeighborLinkers.hpp:116-200 SearchFilterBase; applied by scan+scatter compaction :140-183). */
enum mb200_search_filter {
  MB200_FILTER_NONE = 0,           /* every bounding-sphere candidate becomes a link (the reference's default) */
  MB200_FILTER_SHAPE_DISTANCE = 1  /* keep (i,j) iff surface separation <= 2*buffer: shapes inflated by the buffer
                                      intersect. Identical to NONE for spheres; prunes rod pairs. */
};

int mb200_neighbor_links_create(mb200_neighbor_links** out);
int mb200_neighbor_links_destroy(mb200_neighbor_links* h);
/* Builder setters (set_search_buffer :455, set_search_filter :441, set_sort_search_results :427). */
int mb200_neighbor_links_set_search_buffer(mb200_neighbor_links* h, double buffer);
int mb200_neighbor_links_set_search_filter(mb200_neighbor_links* h, int filter);
/* Orthorhombic periodic box (BASELINE configs[1] "periodic box"). cell_size_h = HOST array of the
 * three cell edge lengths; an entry <= 0 leaves that axis free; NULL switches periodicity off (the default).
 * All three axes periodic = geom::PeriodicScaledMetric (mundy/geom/src/mundy_geom/periodicity.hpp:743-845), a subset =
 * PeriodicMetricX / Y / XY / YZ (:335-741). With a periodic box
 *   - the candidate predicate measures the centre separation with metric.sep (minimum image, :814-818):
 *     |sep(c_i, c_j)|^2 <= (S_i + S_j)^2; each unordered pair is one link (its nearest image);
 *   - the narrow phase and the shape filter evaluate geom::distance(body_i, unwrap_points_to_ref(body_j, metric, c_i))
 *     (:1487-1516), i = the lower caller index; arms of j are relative to its unwrapped centre;
 *   - the rebuild trigger measures metric.sep(point at build time, point now);
 *   - centres are NOT wrapped by the step (trajectories stay continuous; wrap with mb200_periodic_wrap when wanted).
 * The cell list tiles each periodic axis with whole cells and wraps its neighbour walk, so no ghost images are made. */
int mb200_neighbor_links_set_periodic(mb200_neighbor_links* h, const double* cell_size_h);
/* generate() :510-543. Rebuilds when never built, when n changed, when force != 0, or when some body moved more than
 * half the buffer since the last build (:603-615). *rebuilt_h (host) receives 1/0. Synchronises the stream. */
int mb200_neighbor_links_generate(mb200_neighbor_links* h, mb200_stream stream, const mb200_bodies* bodies, int force,
                                  int* rebuilt_h);
/* Number of links (host value; valid after generate). */
int64_t mb200_neighbor_links_count(const mb200_neighbor_links* h);
/* Copy the link list, in the caller's body indices, sorted by (i,j) with i<j, to pairs_out[2*capacity] (device). */
int mb200_neighbor_links_get_pairs(mb200_neighbor_links* h, mb200_stream stream, int32_t* pairs_out, int64_t capacity);

/* One-shot search on explicit search spheres (centre[3n], search radius S[n]): pairs sorted by (i,j).
 * *count_h (host) receives the number found even when it exceeds capacity (then MB200_ERR_CAPACITY). */
int mb200_search_pairs(mb200_stream stream, int64_t n, const double* center, const double* search_radius,
                       int32_t* pairs_out, int64_t capacity, int64_t* count_h);

/* The same with a periodic box (cell_size_h as in mb200_neighbor_links_set_periodic). */
int mb200_search_pairs_periodic(mb200_stream stream, int64_t n, const double* center, const double* search_radius,
                                const double* cell_size_h, int32_t* pairs_out, int64_t capacity, int64_t* count_h);
/* mb200_narrow_phase with a periodic box: body j of every pair is unwrapped to the centre of body i first. */
int mb200_narrow_phase_periodic(mb200_stream stream, const mb200_bodies* bodies, int64_t npairs, const int32_t* pairs,
                                const double* cell_size_h, double* phi, double* normal, double* arm_i, double* arm_j);

/* mb200_narrow_phase for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same arguments, same
 * kernels and bit-identical results; the packed body records that mb200_narrow_phase keeps in stream-ordered scratch
 * live in a static device array instead (4 MB, held by every GPU that loads the library until exit). The call makes no
 * cudaMalloc / cudaMallocAsync / cudaFree and no allocating library call. bodies->n > MB200_NO_ALLOC_MAX_BODIES is
 * refused with MB200_ERR_INVALID_ARGUMENT. Every call on a device shares the array: calls that overlap on the GPU
 * (different streams) race; calls on one stream are ordered and safe. */
#define MB200_NO_ALLOC_MAX_BODIES 65536
int mb200_narrow_phase_no_alloc(mb200_stream stream, const mb200_bodies* bodies, int64_t npairs, const int32_t* pairs,
                                double* phi, double* normal, double* arm_i, double* arm_j);

/* Periodic metric leaves, batched (mundy/geom/src/mundy_geom/periodicity.hpp; every array [3n]; cell_size_h HOST[3]):
 *   sep[k]  = metric.sep(p1[k], p2[k])                          :814-818  minimum-image separation vector
 *   out[k]  = metric.wrap(p[k])                                 :821-824  into [0, L) per periodic axis
 *   out[k]  = unwrap_points_to_ref(p[k], metric, ref[k])        :1333-1339 (points, sphere / spherocylinder centres)
 *   (start_out, end_out)[k] = unwrap_points_to_ref(LineSegment(start[k], end[k]), metric, ref[k])   :1368-1382 */
int mb200_periodic_sep(mb200_stream stream, int64_t n, const double* cell_size_h, const double* p1, const double* p2,
                       double* sep);
int mb200_periodic_wrap(mb200_stream stream, int64_t n, const double* cell_size_h, const double* p, double* out);
int mb200_periodic_unwrap_points(mb200_stream stream, int64_t n, const double* cell_size_h, const double* p,
                                 const double* ref, double* out);
int mb200_periodic_unwrap_segments(mb200_stream stream, int64_t n, const double* cell_size_h, const double* start,
                                   const double* end, const double* ref, double* start_out, double* end_out);

/* ---------------------------------------------------------------------------------------------------------------
 * Link data COO -> CRS. The reference keeps links in COO form (link -> linked entities,
 * mundy/mesh/src/mundy_mesh/LinkCOOData.hpp:146-260) and rebuilds the CRS form (entity -> connected links) with
 * NgpCOOToCRSSynchronizer (impl/NgpCOOToCRSSynchronizer.hpp:253-266: reset, gather_part_1_count with one atomic_add per
 * linked entity :405-468, partial sum, scatter fill). Here, for links of dimensionality 2 (neighbour pairs):
 *   linked[2 * n_links]   entity index in slot 0 / slot 1 of every link; a negative value = empty slot (skipped, :449)
 *   offsets[n_entities+1] CRS row offsets (int64)
 *   crs_links / crs_slots [offsets[n_entities] valid entries, capacity 2 * n_links] the links connected to each entity
 *                         and the slot the entity occupies in them, ascending (link, slot) per entity: deterministic,
 *                         unlike the atomic fill of the reference. crs_slots may be NULL.
 * One stable radix sort + one scan, no floating point. Synchronises (validates the entity indices).
 * ------------------------------------------------------------------------------------------------------------- */
int mb200_links_coo_to_crs(mb200_stream stream, int64_t n_entities, int64_t n_links, const int32_t* linked,
                           int64_t* offsets, int32_t* crs_links, int32_t* crs_slots);

/* ---------------------------------------------------------------------------------------------------------------
 * D. Matrix-free contact operator (there is no assembled Jacobian in the reference).
 *   D   : scrap/lcp_spheres/KokkosLCP.cpp:559-591   F_i -= gamma n, F_j += gamma n ; 6-DOF scrap/UnitTestMundy.cpp:652-674
 *   M   : KokkosLCP.cpp:593-614, scrap/parameter_interface/NGPSpheresLCP.cpp:586-618 (U = F/(6 pi mu r));
 *         rods scrap/UnitTestMundy.cpp:704-753
 *   D^T : KokkosLCP.cpp:616-647 ; 6-DOF UnitTestMundy.cpp:783-824
 * The scatter is done without atomics: body-major incidence lists + per-body ordered sums (deterministic).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct mb200_contact_op {
  int64_t n;              /* bodies */
  int64_t npairs;         /* contacts */
  const int32_t* pairs;   /* [2*npairs] */
  const double* normal;   /* [3*npairs] */
  const double* arm_i;    /* [3*npairs] or NULL => 3-DOF (spheres: force only) */
  const double* arm_j;    /* [3*npairs] or NULL */
  const double* quat;     /* [4n] or NULL => isotropic translational mobility */
  const double* radius;   /* [n] */
  const int32_t* kind;    /* [n] or NULL (all spheres) */
  double viscosity;       /* mu */
  double dt;
} mb200_contact_op;

/* F[3n] (and T[3n] when op->arm_i != NULL) = D gamma. */
int mb200_apply_D(mb200_stream stream, const mb200_contact_op* op, const double* gamma, double* F, double* T);
/* U = M F (and W = M_r T when T,W != NULL). Uses op->n, quat, radius, kind, viscosity only. */
int mb200_apply_M(mb200_stream stream, const mb200_contact_op* op, const double* F, const double* T, double* U,
                  double* W);
/* gdot[npairs] = D^T (U, W). */
int mb200_apply_DT(mb200_stream stream, const mb200_contact_op* op, const double* U, const double* W, double* gdot);
/* y = dt * D^T M D x : the LinearOp `apply(x, y)` of math::convex (mundy/math/src/mundy_math/convex.hpp:186-191). */
int mb200_apply_A(mb200_stream stream, const mb200_contact_op* op, const double* x, double* y);

/* ---------------------------------------------------------------------------------------------------------------
 * E1. Driver BBPGD  resolve_collisions  scrap/lcp_spheres/KokkosLCP.cpp:675-876 (+ external-velocity correction of
 * scrap/parameter_interface/NGPSpheresLCP.cpp:751-759 when U_ext != NULL). SURVEY.md Appendix A is the statement.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct mb200_collision_result {  /* KokkosLCP.cpp:667-671 CollisionResult */
  double max_abs_projected_sep;
  int32_t ite_count;
  int32_t converged;  /* max_abs_projected_sep < tol */
  double max_displacement;
} mb200_collision_result;

typedef struct mb200_lcp_params {
  double tol;
  int32_t max_iters;
  int32_t fixed_iters; /* >= 0: ignore tol, run exactly this many iterations (parity checks); < 0: off */
} mb200_lcp_params;

/* phi0[npairs] in/out (corrected in place by dt*D^T U_ext when U_ext != NULL); gamma[npairs] in/out;
 * F,U[3n] out; T,W[3n] out when op->arm_i != NULL (else may be NULL). result_h is a HOST struct; synchronises. */
int mb200_resolve_collisions(mb200_stream stream, const mb200_contact_op* op, const mb200_lcp_params* params,
                             const double* U_ext, const double* W_ext, double* phi0, double* gamma, double* F,
                             double* T, double* U, double* W, mb200_collision_result* result_h);

/* The same solve with the lever arms in AXIAL form, the form mb200_collision_step runs for spherocylinders: the contact
 * point of a rod lies on its centre line, so arm = sigma * h with h = (L/2) (q (x) z_hat) the half-axis vector of the body
 * (compute_aabb.hpp:115-117) and sigma = 2 clamp(s, 0, 1) - 1, s = the segment parameter geom::distance returns
 * (LineSegmentLineSegment.hpp:313-316); sigma = 0 for a sphere. sigma[2*npairs] = (sigma_i, sigma_j) per contact,
 * length[n]; op->quat and op->kind are required, op->arm_i / arm_j must be NULL. 16 bytes per contact instead of 48. */
int mb200_resolve_collisions_axial(mb200_stream stream, const mb200_contact_op* op, const double* sigma,
                                   const double* length, const mb200_lcp_params* params, const double* U_ext,
                                   const double* W_ext, double* phi0, double* gamma, double* F, double* T, double* U,
                                   double* W, mb200_collision_result* result_h);

/* Shape of the solver's deterministic global reductions (residual max and the three Barzilai-Borwein sums,
 * KokkosLCP.cpp:493-557) for a system of npairs contacts: `grid` CTAs of `threads` threads walk the contacts with a grid
 * stride, sum per thread, then through a shuffle tree per warp and per CTA; the CTA partials are reduced the same way.
 * A CPU checker that adds in this order reproduces the device's iterates bit for bit at any iteration count.
 * arm_mode: 0 no arms (3-DOF), 1 axial arms (the collision step's spherocylinders), 2 caller-supplied arm vectors. */
int mb200_lcp_reduction_shape(int64_t npairs, int arm_mode, int32_t* grid_h, int32_t* threads_h);

/* ---------------------------------------------------------------------------------------------------------------
 * E2. math::convex projected-gradient solver  mundy/math/src/mundy_math/convex.hpp
 * ------------------------------------------------------------------------------------------------------------- */
enum mb200_space_kind { MB200_SPACE_UNCONSTRAINED = 0, MB200_SPACE_LOWER = 1, MB200_SPACE_UPPER = 2, MB200_SPACE_BOX = 3 };
enum mb200_residual_kind {
  MB200_RESID_PROJECTED_DIFF = 0, /* LinfNormProjectedDiffResidual :468-496 (default) */
  MB200_RESID_PROJECTED_GRAD = 1  /* LinfNormProjectedGradientResidual :434-466 */
};
typedef struct mb200_pgd_config { /* PGDConfig :519-525 */
  uint32_t max_iters;             /* default 1000 */
  double tol;                     /* default 1e-8 */
  int32_t space_kind;
  double lo, hi;
  int32_t residual_kind;
} mb200_pgd_config;
typedef struct mb200_solve_result { /* SolveResult :527-541 */
  uint32_t num_iters;
  double residual;
  int32_t converged;
} mb200_solve_result;

/* KokkosBackend pieces (:141-285) on device vectors: the static interface of the B200 backend in the C++ veneer. */
int mb200_vec_axpby(mb200_stream stream, int64_t n, double alpha, const double* x, double beta, double* y);
int mb200_vec_wrapped_axpbyz(mb200_stream stream, int64_t n, double alpha, const double* x, double beta,
                             const double* y, double* z, int space_kind, double lo, double hi);
/* sum (x - x_old)^2 -> *out_h ; sum (x - x_old)(g - g_old) -> *out_h   (both synchronise) */
int mb200_vec_diff_dot(mb200_stream stream, int64_t n, const double* x, const double* x_old, double* out_h);
int mb200_vec_diff_dot2(mb200_stream stream, int64_t n, const double* x, const double* x_old, const double* g,
                        const double* g_old, double* out_h);
int mb200_vec_residual(mb200_stream stream, int64_t n, const double* x, const double* g, int residual_kind,
                       int space_kind, double lo, double hi, double* out_h);
/* y = A x for a dense row-major n x n matrix (the KokkosBlas::gemv path, :166-184). */
int mb200_dense_gemv(mb200_stream stream, int64_t n, const double* A, const double* x, double* y);

/* solve_cqpp / solve_lcp (:789-845) with a dense A (row-major n x n) ... */
int mb200_solve_cqpp_dense(mb200_stream stream, int64_t n, const double* A, const double* q, double* x,
                           const mb200_pgd_config* cfg, mb200_solve_result* result_h);
/* ... and with the collision operator A = dt D^T M D, q = phi0, x = gamma >= 0. */
int mb200_solve_lcp_contact(mb200_stream stream, const mb200_contact_op* op, const double* phi0, double* gamma,
                            const mb200_pgd_config* cfg, mb200_solve_result* result_h);

/* ---------------------------------------------------------------------------------------------------------------
 * G. Non-local mobility operators behind `ApplyMobilityOp`. Spheres, translational part.
 *   pair kernels  scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp: apply_stokes_kernel :686-747,
 *                 apply_rpy_kernel :847-942 (quake_inv_sqrt :72-82), apply_rpyc_kernel :954-1085, apply_local_drag
 *                 :1316-1354; scrap/lcp_spheres/KokkosLCP.cpp: RPYKernel :219-280 (one radius, 1/sqrt),
 *                 apply_rpy_kernel_neighbors :305-333, compute_the_mobility_problem :593-614
 *   operator      scrap/parameter_interface/alens/tests/performance_tests/NgpHP1.cpp: compute_rpy_mobility_spheres :677-701,
 *                 SphereLocalDragMobilityOp / SphereRpyMobilityOp :804-925
 *   solve         NgpHP1.cpp resolve_collisions<ApplyMobilityOp> :1487-1645
 * The sum over the sources of one target is taken in ascending source (CSR row) order from zero, then added to the
 * target velocity: deterministic, and bit-identical to the reference's kernels on the Kokkos Serial backend.
 * ------------------------------------------------------------------------------------------------------------- */
enum mb200_mobility_kernel {
  MB200_KERNEL_STOKES = 0,    /* Periphery.hpp apply_stokes_kernel (radii unused, may be NULL) */
  MB200_KERNEL_RPY = 1,       /* Periphery.hpp apply_rpy_kernel: a = source radius, b = target radius, quake_inv_sqrt */
  MB200_KERNEL_RPYC = 2,      /* Periphery.hpp apply_rpyc_kernel: overlap-corrected RPY */
  MB200_KERNEL_RPY_EXACT = 3  /* KokkosLCP.cpp RPYKernel: 1/sqrt, radius taken from the source, self term 1/(6 pi r mu) */
};
enum mb200_mobility_mode {
  MB200_MOBILITY_LOCAL_DRAG = 0, /* U = F / (6 pi mu r)                                  (HydroType DRY)           */
  MB200_MOBILITY_NEIGHBORS = 1,  /* + pair kernel over the sources of a CSR neighbour list (HYDRO_NEAREST / DISTANT) */
  MB200_MOBILITY_ALL = 2         /* + pair kernel over all spheres                         (HYDRO_ALL)               */
};
/* target_vel[3 nt] += sum over ALL ns sources of K(target, source) source_force. Self pairs (r^2 < 1e-12) add nothing.
 * source_radii / target_radii may be NULL for MB200_KERNEL_STOKES. */
int mb200_mobility_kernel_apply(mb200_stream stream, int kernel, int64_t nt, int64_t ns, double viscosity,
                                const double* target_pos, const double* source_pos, const double* source_radii,
                                const double* target_radii, const double* source_force, double* target_vel);
/* vel[3n] += sum over the sources indices[offsets[t] .. offsets[t+1]) of each target t (sources == targets). */
int mb200_mobility_kernel_apply_neighbors(mb200_stream stream, int kernel, int64_t n, double viscosity,
                                          const double* pos, const double* radii, const int32_t* offsets,
                                          const int32_t* indices, const double* force, double* vel);
/* vel[3n] += force / (6 pi mu r): apply_local_drag (kernel selects which of the two reference forms of the coefficient). */
int mb200_apply_local_drag(mb200_stream stream, int kernel, int64_t n, double viscosity, const double* radii,
                           const double* force, double* vel);

typedef struct mb200_mobility_op { /* the state an `ApplyMobilityOp` holds (NgpHP1.cpp:833-925) */
  int64_t n;
  int32_t mode;               /* mb200_mobility_mode */
  int32_t kernel;             /* mb200_mobility_kernel */
  double viscosity;
  const double* pos;          /* [3n] (unused for LOCAL_DRAG) */
  const double* radius;       /* [n] */
  const int32_t* nbr_offsets; /* [n+1] (NEIGHBORS) */
  const int32_t* nbr_indices;
} mb200_mobility_op;
/* U[3n] = M F: zero, non-local kernel, local drag (`apply(force, velocity)` of the operator). */
int mb200_mobility_apply(mb200_stream stream, const mb200_mobility_op* mob, const double* F, double* U);

/* BBPGD with U = M F delegated to the operator. op: 3-DOF contact operator (arm_i = arm_j = NULL; its radius / viscosity
 * are not used). F_ext[3n] != NULL: U_ext = M F_ext is written to U_ext_out (nullable) and phi0 += dt D^T U_ext.
 * phi0, gamma in/out; F, U out (collision force and velocity). result_h is a HOST struct; synchronises. */
int mb200_resolve_collisions_mobility(mb200_stream stream, const mb200_contact_op* op, const mb200_mobility_op* mob,
                                      const mb200_lcp_params* params, const double* F_ext, double* U_ext_out,
                                      double* phi0, double* gamma, double* F, double* U,
                                      mb200_collision_result* result_h);

/* ---------------------------------------------------------------------------------------------------------------
 * E3. Euler update  x += dt U (KokkosLCP.cpp:1015); rods: q <- math::rotate_quaternion(q, W, dt)
 * (mundy/math/src/mundy_math/Quaternion.hpp:1359-1385). quat/kind/W may be NULL.
 * ------------------------------------------------------------------------------------------------------------- */
int mb200_euler_update(mb200_stream stream, int64_t n, double* center, double* quat, const int32_t* kind,
                       const double* U, const double* W, double dt);

/* External forcing of the multi-step driver: the "attractive ABC flow"
 * F += (a sin z + c cos y, b sin x + a cos z, c sin y + b cos x) - 2 x/|x|, a = 1, b = sqrt 2, c = sqrt 3
 * (scrap/parameter_interface/NGPSpheresLCP.cpp:971-998). center[3n] in, F[3n] in/out (accumulated into). */
int mb200_abc_flow_force(mb200_stream stream, int64_t n, const double* center, double* F);

/* ---------------------------------------------------------------------------------------------------------------
 * Whole collision timestep (the loop body of scrap/parameter_interface/NGPSpheresLCP.cpp:1135-1292):
 *   generate links if needed -> narrow phase -> (phi0 += dt D^T U_ext) -> BBPGD -> U += U_ext -> Euler.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct mb200_step_params {
  double viscosity;
  double dt;
  double tol;
  int32_t max_iters;
  int32_t fixed_iters;       /* < 0 off */
  int32_t throw_on_max_iters; /* !=0: return MB200_ERR_NOT_CONVERGED when ite_count == max_iters (the STK drivers
                                 throw, scrap/lcp_spheres/NGPSpheresLCP.cpp:745-747) */
  int32_t force_rebuild;
} mb200_step_params;

typedef struct mb200_step_stats {
  mb200_collision_result lcp;
  int64_t num_links;       /* constraints in this step */
  int32_t rebuilt;
  int32_t kernel_launches; /* kernels of this library launched by the call */
} mb200_step_stats;

/* Optional per-body arrays (device): F_ext -> U_ext = M F_ext is added (NULL: none); outputs may be NULL.
 * bodies->center / quat are advanced in place. stats_h is a HOST struct; synchronises the stream. */
int mb200_collision_step(mb200_neighbor_links* links, mb200_stream stream, const mb200_bodies* bodies,
                         const mb200_step_params* params, const double* F_ext, const double* T_ext, double* F_col,
                         double* T_col, double* U_out, double* W_out, mb200_step_stats* stats_h);

/* Same step through HOST buffers (every pointer in bodies_h and the optional arrays is a host pointer): copies the
 * inputs to the device, runs mb200_collision_step, copies centres/quaternions (and optional outputs) back.
 * This is the call a host application that keeps its fields on the CPU would make. */
int mb200_collision_step_host(mb200_neighbor_links* links, mb200_stream stream, const mb200_bodies* bodies_h,
                              const mb200_step_params* params, const double* F_ext_h, const double* T_ext_h,
                              double* F_col_h, double* T_col_h, double* U_out_h, double* W_out_h,
                              mb200_step_stats* stats_h);

/* ---------------------------------------------------------------------------------------------------------------
 * Multi-GPU (one process per GPU, NCCL over NVLink / NVSwitch). Replaces the MPI usage of the STK drivers on this path
 * (scrap/lcp_spheres/NGPSpheresLCP.cpp:371,450-452 all_reduce of the residual and the BB sums; :256-283 ghosting).
 *
 * Decomposition: every rank OWNS a set of bodies (any number; mundy_b200/dist.py orders a global set along the reference's
 * Hilbert curve of the centres and hands rank r the r-th contiguous block) and passes ONLY those to the library; nothing is
 * replicated. At every link rebuild the ranks all-gather their owned bounding boxes (7 doubles each), every rank sends
 * the owned bodies whose search sphere reaches a peer's box to that peer (nearest periodic image when the links carry a
 * periodic box), and a rank works on the ranks' blocks in rank order: received ghosts and owned bodies in global id order,
 * hence the same pair orientation everywhere. A link with one owned end is kept by both owners and its gamma is computed
 * redundantly (bit-identically) on both; it is counted in the global sums by the owner of its lower-id body
 * (scrap/UnitTestMundy.cpp:553-555). Between rebuilds the owners refresh the ghosts' centres / quaternions once per step.
 * Per solver iteration the owners' velocity records go to the ghost rows and the ranks' four scalars to every rank: through
 * peer memory from inside the kernels (default; the body pass handles the ghosted bodies first so that the halo travels
 * while the interior is processed) or with NCCL send/recv + all-gather (see mb200_comm_exchange_mode).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct mb200_comm mb200_comm;
#define MB200_COMM_ID_BYTES 128
/* Rank 0 creates the id (host buffer of MB200_COMM_ID_BYTES) and hands it to the other ranks by any means. */
int mb200_comm_get_unique_id(char* id_out_h);
/* Collective over all ranks. The calling thread's current CUDA device is the rank's GPU. */
int mb200_comm_create(mb200_comm** out, int world, int rank, const char* id_h);
int mb200_comm_destroy(mb200_comm* comm);
int mb200_comm_world(const mb200_comm* comm);
int mb200_comm_rank(const mb200_comm* comm);

/* How the solver's per-iteration exchange runs: 1 = peer memory (CUDA IPC mappings over NVLink: the body / contact kernels
 * store velocity records and reduction totals straight into the peers' memory and signal with system-scope flags; chosen
 * automatically at the first distributed build when every rank could map every other rank), 0 = NCCL send/recv +
 * all-gather (fallback; forced with the environment variable MB200_DIST_EXCHANGE=nccl). Valid after the first step.
 * MB200_DIST_EXCHANGE=peer additionally takes the peer-memory kernels when the communicator has a single rank (no peer
 * to exchange with: a way to test and time those kernels on one GPU). */
int mb200_comm_exchange_mode(const mb200_comm* comm);

/* One collision timestep of the bodies distributed over the ranks. owned_bodies: device arrays of the bodies THIS rank owns
 * (n >= 0, may differ between ranks); on return they hold the advanced centres / quaternions. F_ext / T_ext: per owned body
 * [3n] or NULL (the same choice on every rank). F_col / U_out: optional [3n], per owned body.
 * stats_h: links and launches of THIS rank; lcp fields are global. Collective; synchronises the stream.
 * A rank that returns an error before the first collective of the step leaves its peers waiting: arguments must be valid on
 * every rank (they are checked before anything is exchanged). */
int mb200_collision_step_dist(mb200_neighbor_links* links, mb200_comm* comm, mb200_stream stream,
                              const mb200_bodies* owned_bodies, const mb200_step_params* params,
                              const double* F_ext, const double* T_ext, double* F_col, double* U_out,
                              mb200_step_stats* stats_h);
/* Bodies (owned + ghosts) and ghosts alone in the last distributed build of this rank. */
int mb200_dist_local_counts(const mb200_neighbor_links* links, int64_t* n_local_h, int64_t* n_ghost_rows_h);

/* Re-partitioning of the ranks' owned blocks on the device (splitters + all-to-all-v of the migrating bodies).
 * The reference re-balances with stk::balance::balanceStkMesh (RCB; scrap/UnitTestMundy.cpp:1448, third party) and
 * migrates entities through STK; here: 30-bit keys = position of the centre along the reference's Hilbert curve
 * (math::hilbert_3d, mundy/math/src/mundy_math/Hilbert.hpp:48-85) on a 2^10 lattice over the global bounding box, a
 * histogram of the keys' top 18 bits summed over the ranks, W - 1 splitters at the bucket boundaries nearest to equal
 * counts, NCCL send / receive of the bodies to the rank whose key range holds them. Collective; synchronises.
 *   owned_in        the bodies this rank owns now (n >= 0)
 *   extra_in/out    optional per-body payload of extra_dim doubles (external forces, ...) that travels with the bodies
 *   capacity        bodies the output arrays can hold; more arriving -> MB200_ERR_CAPACITY on this rank (owned_out->n is
 *                   still set; every collective of the call has completed, the other ranks are not left waiting)
 *   owned_out       the caller's OUTPUT arrays (written through the struct's pointers, the const members included;
 *                   quat / length / kind may be NULL for spheres); owned_out->n = bodies this rank owns afterwards,
 *                   in ascending key order (ties: by source rank, then the source's order)
 *   origin_out      optional [capacity]: (source rank << 32) | index in that rank's owned_in
 *   keys_out        optional [capacity]: the bodies' keys
 *   links           optional: the rank's link handle is marked for a rebuild at its next distributed step
 * Bodies of one bucket (16^3 lattice cells) stay together, so blocks are equal up to one bucket's population. */
int mb200_dist_repartition(mb200_comm* comm, mb200_stream stream, mb200_neighbor_links* links,
                           const mb200_bodies* owned_in, const double* extra_in, int extra_dim, int64_t capacity,
                           mb200_bodies* owned_out, double* extra_out, int64_t* origin_out, uint32_t* keys_out);
/* The keys alone: keys_out[k] = position of cell min(trunc((c_k - lo) / span * 2^10), 2^10 - 1) (per axis) along
 * math::hilbert_3d's curve (Hilbert.hpp:48-85; visiting order pinned against UnitTestHilbert.cpp's goldens in
 * tests/test_dist_cpu.py through mundy_b200/dist.py). lo_h: HOST double[3]. */
int mb200_hilbert_keys(mb200_stream stream, int64_t n, const double* center, const double* lo_h, double span,
                       uint32_t* keys_out);

/* Total kernels launched by this library in the calling process (a monotone counter; for bench accounting). */
int64_t mb200_kernel_launch_count(void);

/* ---------------------------------------------------------------------------------------------------------------
 * Stage timers (the reference drivers print Kokkos::Timer wall times per phase, scrap/lcp_spheres/NGPSpheresLCP.cpp:
 * 999-1053; here: CUDA events recorded on the stream the kernels are launched on). Off by default; while enabled every
 * launch of the listed stages is bracketed by an event pair. mb200_profile_read synchronises the recorded events,
 * returns the accumulated device time [ms] and the number of bracketed regions of one stage since the last
 * mb200_profile_enable(1), and keeps accumulating. mb200_profile_enable(k), k > 1: of the per-launch stages (LCP_BODY,
 * LCP_RATE, LCP_COMM) only every k-th region is bracketed (an event pair between back-to-back kernels costs about a
 * launch gap: 4.5 % of a 1M-rod step when every launch is bracketed); mb200_profile_calls returns the number of regions
 * entered, bracketed or not.
 * ------------------------------------------------------------------------------------------------------------- */
enum mb200_stage {
  MB200_STAGE_TRIGGER = 0,    /* rebuild trigger (moved more than buffer/2)        one region per step   */
  MB200_STAGE_BROADPHASE = 1, /* link rebuild: keys, sort, cell table, pair search, incidence lists      */
  MB200_STAGE_PACK = 2,       /* bodies -> Morton-order records and solver inputs   one region per step   */
  MB200_STAGE_NARROW = 3,     /* narrow phase                                        one region per step   */
  MB200_STAGE_LCP_SETUP = 4,  /* BBPGD: mobility records, initial state, phi0 += dt D^T U_ext   per solve */
  MB200_STAGE_LCP_BODY = 5,   /* BBPGD: gamma update, F = D gamma, U = M F           one region per launch */
  MB200_STAGE_LCP_RATE = 6,   /* BBPGD: gamma update, g = D^T U, residual, BB dots   one region per launch */
  MB200_STAGE_EULER = 7,      /* velocity totals, Euler update, scatter to the caller's order            */
  MB200_STAGE_LCP_COMM = 8,   /* multi-GPU: velocity halo (pack, NCCL send/recv, unpack), all-gather of the sums */
  MB200_STAGE_STATE_COMM = 9, /* multi-GPU: owned block in / out of the local set, per-step refresh of the ghost state   */
  MB200_STAGE_COUNT = 10
};
int mb200_profile_enable(int on);
int mb200_profile_read(int stage, double* ms_total_h, int64_t* regions_h);
int mb200_profile_calls(int stage, int64_t* calls_h);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* MUNDY_B200_H_ */
