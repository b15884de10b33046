// mundy_b200/mobility.hpp — the `ApplyMobilityOp` seam of the templated collision solve
// (scrap/parameter_interface/alens/tests/performance_tests/NgpHP1.cpp:804-925 SphereLocalDragMobilityOp /
// SphereRpyMobilityOp, :1487-1645 resolve_collisions<ApplyMobilityOp>) and the pair kernels underneath it
// (scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp :686-1085, :1316-1354;
// scrap/lcp_spheres/KokkosLCP.cpp :219-333, :593-614) on top of the C-ABI.
//
// Kept from the reference: free functions apply_stokes_kernel / apply_rpy_kernel / apply_rpyc_kernel /
// apply_local_drag with the reference's argument order (space, viscosity, source_positions, target_positions,
// source_radii, target_radii, source_forces, target_velocities; velocities are accumulated into), the size checks that
// throw std::invalid_argument, operator classes with `update()` and `apply(force, velocity)`, and a
// resolve_collisions(…, apply_mobility_op, …) that returns {max_abs_projected_sep, ite_count, max_displacement} and throws
// std::runtime_error("Collision resolution did not converge!") at the iteration cap (:1632-1634).
// Replaced: STK fields / selectors by flat device arrays; `apply` ASSIGNS velocity = M force (KokkosLCP.cpp:603-613) —
// SphereRpyMobilityOp::apply adds into the field and relies on the caller to have zeroed it.
#ifndef MUNDY_B200_MOBILITY_HPP_
#define MUNDY_B200_MOBILITY_HPP_

#include "core.hpp"
#include "neighbor_links.hpp"

namespace mundy_b200 {
namespace alens {
namespace periphery {

struct DeviceSpace {  // stands where the reference passes an execution space
  mb200_stream stream = nullptr;
};

namespace impl {
inline void require(bool ok, const char* msg) {
  if (!ok) throw std::invalid_argument(msg);
}
inline void apply_kernel(const DeviceSpace& space, int kernel, double viscosity, const DeviceVector<double>& source_positions,
                         const DeviceVector<double>& target_positions, const DeviceVector<double>* source_radii,
                         const DeviceVector<double>* target_radii, const DeviceVector<double>& source_forces,
                         DeviceVector<double>& target_velocities, const char* who) {
  const std::size_t ns = source_positions.extent(0) / 3, nt = target_positions.extent(0) / 3;
  require(source_forces.extent(0) == 3 * ns, "source_forces must have size 3 * num_source_points.");
  require(target_velocities.extent(0) == 3 * nt, "target_velocities must have size 3 * num_target_points.");
  if (source_radii) require(source_radii->extent(0) == ns, "source_radii must have size num_source_points.");
  if (target_radii) require(target_radii->extent(0) == nt, "target_radii must have size num_target_points.");
  check(mb200_mobility_kernel_apply(space.stream, kernel, static_cast<int64_t>(nt), static_cast<int64_t>(ns), viscosity,
                                    target_positions.data(), source_positions.data(),
                                    source_radii ? source_radii->data() : nullptr,
                                    target_radii ? target_radii->data() : nullptr, source_forces.data(),
                                    target_velocities.data()),
        who);
}
}  // namespace impl

/// Periphery.hpp:686-747: u_target += M f_source, Stokeslet.
inline void apply_stokes_kernel(const DeviceSpace& space, double viscosity, const DeviceVector<double>& source_positions,
                                const DeviceVector<double>& target_positions, const DeviceVector<double>& source_forces,
                                DeviceVector<double>& target_velocities) {
  impl::apply_kernel(space, MB200_KERNEL_STOKES, viscosity, source_positions, target_positions, nullptr, nullptr,
                     source_forces, target_velocities, "apply_stokes_kernel");
}
/// Periphery.hpp:847-942: RPY, no self interaction ("simply add 1/(6 pi mu) * f to u").
inline void apply_rpy_kernel(const DeviceSpace& space, double viscosity, const DeviceVector<double>& source_positions,
                             const DeviceVector<double>& target_positions, const DeviceVector<double>& source_radii,
                             const DeviceVector<double>& target_radii, const DeviceVector<double>& source_forces,
                             DeviceVector<double>& target_velocities) {
  impl::apply_kernel(space, MB200_KERNEL_RPY, viscosity, source_positions, target_positions, &source_radii, &target_radii,
                     source_forces, target_velocities, "apply_rpy_kernel");
}
/// Periphery.hpp:954-1085: overlap-corrected RPY.
inline void apply_rpyc_kernel(const DeviceSpace& space, double viscosity, const DeviceVector<double>& source_positions,
                              const DeviceVector<double>& target_positions, const DeviceVector<double>& source_radii,
                              const DeviceVector<double>& target_radii, const DeviceVector<double>& source_forces,
                              DeviceVector<double>& target_velocities) {
  impl::apply_kernel(space, MB200_KERNEL_RPYC, viscosity, source_positions, target_positions, &source_radii, &target_radii,
                     source_forces, target_velocities, "apply_rpyc_kernel");
}
/// Periphery.hpp:1316-1354 (argument order: velocities, forces, radii).
inline void apply_local_drag(const DeviceSpace& space, double viscosity, DeviceVector<double>& sphere_velocities,
                             const DeviceVector<double>& sphere_forces, const DeviceVector<double>& sphere_radii) {
  impl::require(sphere_velocities.extent(0) == 3 * sphere_radii.extent(0),
                "apply_local_drag: sphere_velocities must have size 3 * num_spheres.");
  impl::require(sphere_forces.extent(0) == 3 * sphere_radii.extent(0),
                "apply_local_drag: sphere_forces must have size 3 * num_spheres.");
  check(mb200_apply_local_drag(space.stream, MB200_KERNEL_RPY, static_cast<int64_t>(sphere_radii.extent(0)), viscosity,
                               sphere_radii.data(), sphere_forces.data(), sphere_velocities.data()),
        "apply_local_drag");
}

}  // namespace periphery

/// Common state of the mobility operators: what `ApplyMobilityOp` must provide is update() and apply(force, velocity).
class SphereMobilityOpBase {
 public:
  void update() {}  // positions / radii are read through the pointers at every apply
  void apply(const DeviceVector<double>& force, DeviceVector<double>& velocity) {
    periphery::impl::require(force.extent(0) == 3 * static_cast<std::size_t>(op_.n) && velocity.extent(0) == force.extent(0),
                             "mobility op: force / velocity must have size 3 * num_spheres.");
    check(mb200_mobility_apply(stream_, &op_, force.data(), velocity.data()), "mb200_mobility_apply");
  }
  const mb200_mobility_op& c_struct() const { return op_; }
  mb200_stream stream() const { return stream_; }

 protected:
  SphereMobilityOpBase(mb200_stream stream, int mode, int kernel, double viscosity, const DeviceVector<double>* positions,
                       const DeviceVector<double>& radii, const DeviceVector<int32_t>* offsets,
                       const DeviceVector<int32_t>* indices)
      : stream_(stream) {
    if (positions)
      periphery::impl::require(positions->extent(0) == 3 * radii.extent(0), "mobility op: positions must have size 3 * num_spheres.");
    if (offsets) periphery::impl::require(offsets->extent(0) == radii.extent(0) + 1, "mobility op: offsets must have size num_spheres + 1.");
    op_ = mb200_mobility_op{static_cast<int64_t>(radii.extent(0)), mode, kernel, viscosity,
                            positions ? positions->data() : nullptr, radii.data(), offsets ? offsets->data() : nullptr,
                            indices ? indices->data() : nullptr};
  }
  mb200_stream stream_;
  mb200_mobility_op op_;
};

/// NgpHP1.cpp:804-831.
class SphereLocalDragMobilityOp : public SphereMobilityOpBase {
 public:
  SphereLocalDragMobilityOp(double viscosity, const DeviceVector<double>& radii, mb200_stream stream = nullptr)
      : SphereMobilityOpBase(stream, MB200_MOBILITY_LOCAL_DRAG, MB200_KERNEL_RPY, viscosity, nullptr, radii, nullptr, nullptr) {}
};
/// NgpHP1.cpp:833-925: RPY over all spheres + local drag. `kernel` selects RPY / RPYC / the driver's exact RPY.
class SphereRpyMobilityOp : public SphereMobilityOpBase {
 public:
  SphereRpyMobilityOp(double viscosity, const DeviceVector<double>& positions, const DeviceVector<double>& radii,
                      int kernel = MB200_KERNEL_RPY, mb200_stream stream = nullptr)
      : SphereMobilityOpBase(stream, MB200_MOBILITY_ALL, kernel, viscosity, &positions, radii, nullptr, nullptr) {}
};
/// KokkosLCP.cpp:305-333, :593-614 with only_use_neighbors: pair kernel over a CSR list of sources per target.
class SphereNeighborRpyMobilityOp : public SphereMobilityOpBase {
 public:
  SphereNeighborRpyMobilityOp(double viscosity, const DeviceVector<double>& positions, const DeviceVector<double>& radii,
                              const DeviceVector<int32_t>& neighbor_offsets, const DeviceVector<int32_t>& neighbor_indices,
                              int kernel = MB200_KERNEL_RPY_EXACT, mb200_stream stream = nullptr)
      : SphereMobilityOpBase(stream, MB200_MOBILITY_NEIGHBORS, kernel, viscosity, &positions, radii, &neighbor_offsets,
                             &neighbor_indices) {}
};

/// NgpHP1.cpp:1487-1645. neighbor_ids [2 C], signed_sep_dist [C] (corrected in place), con_normal_ij [3 C],
/// lagrange_multipliers [C] (initial guess in, solution out). node_force (external, may be null) enters through
/// phi0 += dt D^T M F_ext and its velocity is written to node_velocity when that is non-null.
template <typename ApplyMobilityOp>
inline mesh::CollisionResult resolve_collisions(double dt, double max_allowable_overlap, int max_col_iterations,
                                                ApplyMobilityOp& apply_mobility_op, const DeviceVector<int32_t>& neighbor_ids,
                                                const DeviceVector<double>* node_force, DeviceVector<double>* node_velocity,
                                                DeviceVector<double>& node_collision_force,
                                                DeviceVector<double>& node_collision_velocity,
                                                DeviceVector<double>& signed_sep_dist, const DeviceVector<double>& con_normal_ij,
                                                DeviceVector<double>& lagrange_multipliers) {
  const mb200_mobility_op& mob = apply_mobility_op.c_struct();
  const std::size_t c = signed_sep_dist.extent(0), n = static_cast<std::size_t>(mob.n);
  periphery::impl::require(neighbor_ids.extent(0) == 2 * c && con_normal_ij.extent(0) == 3 * c && lagrange_multipliers.extent(0) == c,
                           "resolve_collisions: constraint arrays disagree on the number of collisions.");
  periphery::impl::require(node_collision_force.extent(0) == 3 * n && node_collision_velocity.extent(0) == 3 * n,
                           "resolve_collisions: node fields must have size 3 * num_spheres.");
  mb200_contact_op op{};
  op.n = mob.n, op.npairs = static_cast<int64_t>(c), op.pairs = neighbor_ids.data(), op.normal = con_normal_ij.data();
  op.radius = mob.radius, op.viscosity = mob.viscosity, op.dt = dt;
  const mb200_lcp_params params{max_allowable_overlap, max_col_iterations, -1};
  mb200_collision_result r{};
  check(mb200_resolve_collisions_mobility(apply_mobility_op.stream(), &op, &mob, &params, node_force ? node_force->data() : nullptr,
                                          (node_force && node_velocity) ? node_velocity->data() : nullptr, signed_sep_dist.data(),
                                          lagrange_multipliers.data(), node_collision_force.data(),
                                          node_collision_velocity.data(), &r),
        "mb200_resolve_collisions_mobility");
  if (!r.converged && r.ite_count == max_col_iterations) throw std::runtime_error("Collision resolution did not converge!");
  return mesh::CollisionResult{r.max_abs_projected_sep, r.ite_count, r.max_displacement};
}

}  // namespace alens
}  // namespace mundy_b200
#endif  // MUNDY_B200_MOBILITY_HPP_
