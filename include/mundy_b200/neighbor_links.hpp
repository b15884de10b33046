// mundy_b200/neighbor_links.hpp — the mesh::GenNeighborLinks seam (mundy/mesh/src/mundy_mesh/GenNeighborLinkers.hpp:294-543)
// and the driver-level collision step on top of the C-ABI.
//
// Kept from the reference: the builder protocol (setters chain, throw std::runtime_error after concretize(), :401-462),
// `bool generate()` returning whether a rebuild happened (:510-543), the ExcludeSelfInteractions semantics (:185-200, always
// on), the rebuild rule "some body moved more than half the buffer" (:603-615).
// Replaced: STK selectors / link parts / generators (the caller-side data model, out of scope) by a flat `Bodies` view of
// device arrays in the reference's field layouts; stk::search::coarse_search by the device cell-list search.
#ifndef MUNDY_B200_NEIGHBOR_LINKS_HPP_
#define MUNDY_B200_NEIGHBOR_LINKS_HPP_

#include "core.hpp"
#include "geom.hpp"

namespace mundy_b200 {
namespace mesh {

/// Non-owning view of the body fields (device pointers). kind/quat/length may be null when every body is a sphere.
struct Bodies {
  int64_t n = 0;
  double* center = nullptr;        // [3n] NODE_COORDS
  double* quat = nullptr;          // [4n] (w,x,y,z)
  const double* radius = nullptr;  // [n]
  const double* length = nullptr;  // [n]
  const int32_t* kind = nullptr;   // [n] MB200_SPHERE / MB200_SPHEROCYLINDER
  mb200_bodies c_struct() const { return mb200_bodies{n, center, quat, radius, length, kind}; }
};

namespace search_filters {
enum class Kind { None = MB200_FILTER_NONE, ShapeDistance = MB200_FILTER_SHAPE_DISTANCE };
}

// Stand-ins for the two STK enumerations the reference's setters take (stk::topology::rank_t, stk::search::SearchMethod),
// with the enumerators a caller of this path uses; the values are not STK's.
enum class EntityRank { NODE_RANK, EDGE_RANK, FACE_RANK, ELEM_RANK };
enum class SearchMethod { KDTREE, MORTON_LBVH, ARBORX };

struct CollisionResult {  // scrap/lcp_spheres/KokkosLCP.cpp:667-671
  double max_abs_projected_sep;
  int ite_count;
  double max_displacement;
};

class GenNeighborLinks {
 public:
  explicit GenNeighborLinks(mb200_stream stream = nullptr) : stream_(stream) {
    check(mb200_neighbor_links_create(&h_), "mb200_neighbor_links_create");
  }
  ~GenNeighborLinks() { mb200_neighbor_links_destroy(h_); }
  GenNeighborLinks(const GenNeighborLinks&) = delete;
  GenNeighborLinks& operator=(const GenNeighborLinks&) = delete;

  bool is_concretized() const { return concretized_; }

  // ---- setters (GenNeighborLinkers.hpp:401-462) ----
  /// :409-418. Like the reference, only element-to-element searches are accepted.
  GenNeighborLinks& set_source_target_rank(EntityRank source_rank, EntityRank target_rank) {
    require_open("Cannot set source and target ranks after concretization.");
    if (source_rank != EntityRank::ELEM_RANK || target_rank != EntityRank::ELEM_RANK)
      throw std::runtime_error("To start, we will only support element to element neighbor searches.");
    return *this;
  }
  /// :444-448. Stored; every method names the same candidate set (the fp64 predicate of SURVEY 8a-B2), which the device
  /// finds with its own cell list whatever is asked for.
  GenNeighborLinks& set_search_method(SearchMethod method) {
    require_open("Cannot set search method after concretization.");
    search_method_ = method;
    return *this;
  }
  SearchMethod search_method() const { return search_method_; }
  GenNeighborLinks& set_enforce_source_target_symmetry(bool v) {
    require_open("Cannot set enforce source-target symmetry after concretization.");
    enforce_symmetry_ = v;  // stored, as in the reference (never consulted there either)
    return *this;
  }
  GenNeighborLinks& set_allow_duplicate_links(bool v) {
    require_open("Cannot set allow duplicate links after concretization.");
    allow_duplicates_ = v;
    return *this;
  }
  GenNeighborLinks& set_sort_search_results(bool v) {
    require_open("Cannot set sort search results after concretization.");
    sort_results_ = v;  // results are always sorted by (i, j) here
    return *this;
  }
  GenNeighborLinks& set_search_filter(search_filters::Kind k) {
    require_open("Cannot set search filter after concretization.");
    check(mb200_neighbor_links_set_search_filter(h_, static_cast<int>(k)), "mb200_neighbor_links_set_search_filter");
    return *this;
  }
  GenNeighborLinks& set_search_buffer(double buffer) {
    require_open("Cannot set search buffer after concretization.");
    check(mb200_neighbor_links_set_search_buffer(h_, buffer), "mb200_neighbor_links_set_search_buffer");
    return *this;
  }
  /// Periodic box: the search, the filter and the narrow phase of collision_step() use this metric
  /// (geom::PeriodicScaledMetric or one of the partially periodic metrics; include/mundy_b200.h
  /// mb200_neighbor_links_set_periodic). Not a reference setter: the reference has no periodic search.
  GenNeighborLinks& set_periodic_metric(const geom::PeriodicScaledMetric& metric) {
    require_open("Cannot set the periodic metric after concretization.");
    check(mb200_neighbor_links_set_periodic(h_, metric.cell_size()), "mb200_neighbor_links_set_periodic");
    return *this;
  }
  /// acts_on(source, target, generators, link parts) of the reference (:464-497) with a self-search over one body set.
  GenNeighborLinks& acts_on(const Bodies& bodies) {
    require_open("Cannot set source/targets after concretization.");
    bodies_ = bodies;
    has_bodies_ = true;
    return *this;
  }
  GenNeighborLinks& concretize() {
    if (!has_bodies_) throw std::runtime_error("GenNeighborLinks: exactly one acts_on() specialization is required.");
    concretized_ = true;
    return *this;
  }

  /// generate() :510-543. Returns true when the link list was rebuilt.
  bool generate(bool force = false) {
    if (!concretized_) throw std::runtime_error("GenNeighborLinks::generate called before concretize().");
    const mb200_bodies b = bodies_.c_struct();
    int rebuilt = 0;
    check(mb200_neighbor_links_generate(h_, stream_, &b, force ? 1 : 0, &rebuilt), "mb200_neighbor_links_generate");
    return rebuilt != 0;
  }
  int64_t num_links() const { return mb200_neighbor_links_count(h_); }
  /// The links as (i, j), i < j, in the caller's body indices, sorted; device array of 2 * num_links() int32.
  DeviceVector<int32_t> links() {
    DeviceVector<int32_t> out(static_cast<std::size_t>(2 * num_links()));
    check(mb200_neighbor_links_get_pairs(h_, stream_, out.data(), num_links()), "mb200_neighbor_links_get_pairs");
    return out;
  }

  /// One whole collision timestep of the multi-step driver (scrap/parameter_interface/NGPSpheresLCP.cpp:1135-1292):
  /// generate-if-needed -> narrow phase -> BBPGD resolve_collisions -> Euler. Throws like the STK drivers when the solver
  /// hits max_col_iterations (scrap/lcp_spheres/NGPSpheresLCP.cpp:745-747).
  CollisionResult collision_step(double viscosity, double dt, double tol, int max_col_iterations,
                                 const double* F_ext = nullptr, double* F_col = nullptr, double* U_out = nullptr,
                                 const double* T_ext = nullptr, double* T_col = nullptr, double* W_out = nullptr) {
    if (!concretized_) throw std::runtime_error("GenNeighborLinks::collision_step called before concretize().");
    const mb200_bodies b = bodies_.c_struct();
    const mb200_step_params p{viscosity, dt, tol, max_col_iterations, -1, 1, 0};
    mb200_step_stats stats{};
    const int status = mb200_collision_step(h_, stream_, &b, &p, F_ext, T_ext, F_col, T_col, U_out, W_out, &stats);
    if (status == MB200_ERR_NOT_CONVERGED) throw std::runtime_error("Collision resolution did not converge!");
    check(status, "mb200_collision_step");
    last_stats_ = stats;
    return CollisionResult{stats.lcp.max_abs_projected_sep, stats.lcp.ite_count, stats.lcp.max_displacement};
  }
  const mb200_step_stats& last_step_stats() const { return last_stats_; }
  mb200_neighbor_links* handle() { return h_; }

 private:
  void require_open(const char* msg) const {
    if (concretized_) throw std::runtime_error(msg);
  }
  mb200_neighbor_links* h_ = nullptr;
  mb200_stream stream_;
  Bodies bodies_;
  bool has_bodies_ = false, concretized_ = false;
  bool enforce_symmetry_ = false, allow_duplicates_ = false, sort_results_ = true;
  SearchMethod search_method_ = SearchMethod::MORTON_LBVH;
  mb200_step_stats last_stats_{};
};

}  // namespace mesh
}  // namespace mundy_b200
#endif  // MUNDY_B200_NEIGHBOR_LINKS_HPP_
