"""ctypes binding of the C-ABI in include/mundy_b200.h (libmundy_b200.so, built in-tree by `make`).

This module is plumbing only: it declares the argument types of every exported entry point and raises on a
non-zero status. There is no CPU fallback: if the library is missing, importing this module fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmundy_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not found: build it with `make` at the repo root (or __graft_entry__.build()). "
        "mundy_b200 has no CPU fallback."
    )

lib = C.CDLL(LIB_PATH)

c_i64 = C.c_int64
c_i32 = C.c_int32
c_dbl = C.c_double
vp = C.c_void_p


class Bodies(C.Structure):
    _fields_ = [("n", c_i64), ("center", vp), ("quat", vp), ("radius", vp), ("length", vp), ("kind", vp)]


class ContactOp(C.Structure):
    _fields_ = [
        ("n", c_i64), ("npairs", c_i64), ("pairs", vp), ("normal", vp), ("arm_i", vp), ("arm_j", vp),
        ("quat", vp), ("radius", vp), ("kind", vp), ("viscosity", c_dbl), ("dt", c_dbl),
    ]


class CollisionResult(C.Structure):
    _fields_ = [("max_abs_projected_sep", c_dbl), ("ite_count", c_i32), ("converged", c_i32),
                ("max_displacement", c_dbl)]


class LcpParams(C.Structure):
    _fields_ = [("tol", c_dbl), ("max_iters", c_i32), ("fixed_iters", c_i32)]


class PgdConfig(C.Structure):
    _fields_ = [("max_iters", C.c_uint32), ("tol", c_dbl), ("space_kind", c_i32), ("lo", c_dbl), ("hi", c_dbl),
                ("residual_kind", c_i32)]


class SolveResult(C.Structure):
    _fields_ = [("num_iters", C.c_uint32), ("residual", c_dbl), ("converged", c_i32)]


class MobilityOp(C.Structure):
    _fields_ = [("n", c_i64), ("mode", c_i32), ("kernel", c_i32), ("viscosity", c_dbl), ("pos", vp), ("radius", vp),
                ("nbr_offsets", vp), ("nbr_indices", vp)]


class StepParams(C.Structure):
    _fields_ = [("viscosity", c_dbl), ("dt", c_dbl), ("tol", c_dbl), ("max_iters", c_i32), ("fixed_iters", c_i32),
                ("throw_on_max_iters", c_i32), ("force_rebuild", c_i32)]


class StepStats(C.Structure):
    _fields_ = [("lcp", CollisionResult), ("num_links", c_i64), ("rebuilt", c_i32), ("kernel_launches", c_i32)]


P = C.POINTER

# name -> (restype, argtypes). Every symbol declared in include/mundy_b200.h must appear here
# (tests/test_abi.py checks the header against this table and against the .so's export list).
SIGNATURES = {
    "mb200_abi_version": (C.c_int, []),
    "mb200_last_error_string": (C.c_char_p, []),
    "mb200_device_count": (C.c_int, []),
    "mb200_kernel_launch_count": (c_i64, []),
    "mb200_profile_enable": (C.c_int, [C.c_int]),
    "mb200_profile_read": (C.c_int, [C.c_int, P(c_dbl), P(c_i64)]),
    "mb200_profile_calls": (C.c_int, [C.c_int, P(c_i64)]),
    "mb200_device_alloc": (C.c_int, [P(vp), c_i64]),
    "mb200_device_free": (C.c_int, [vp]),
    "mb200_copy_to_device": (C.c_int, [vp, vp, vp, c_i64]),
    "mb200_copy_to_host": (C.c_int, [vp, vp, vp, c_i64]),
    "mb200_stream_synchronize": (C.c_int, [vp]),
    "mb200_compute_aabb_sphere": (C.c_int, [vp, c_i64, vp, vp, vp]),
    "mb200_compute_aabb_spherocylinder": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp]),
    "mb200_compute_aabb_segment": (C.c_int, [vp, c_i64, vp, vp, vp, vp]),
    "mb200_distance_sphere_sphere": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_segment_segment": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_point_segment": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_segment_sphere": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_compute_aabb_point": (C.c_int, [vp, c_i64, vp, vp]),
    "mb200_compute_aabb_line_segment": (C.c_int, [vp, c_i64, vp, vp, vp]),
    "mb200_compute_aabb_ellipsoid": (C.c_int, [vp, c_i64, vp, vp, vp, vp]),
    "mb200_distance_point_sphere": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp]),
    "mb200_distance_point_line": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_line_line": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_line_sphere": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_point_ellipsoid": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_ellipsoid_ellipsoid": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_distance_circle3d_circle3d": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "mb200_narrow_phase": (C.c_int, [vp, P(Bodies), c_i64, vp, vp, vp, vp, vp]),
    "mb200_neighbor_links_create": (C.c_int, [P(vp)]),
    "mb200_neighbor_links_destroy": (C.c_int, [vp]),
    "mb200_neighbor_links_set_search_buffer": (C.c_int, [vp, c_dbl]),
    "mb200_neighbor_links_set_search_filter": (C.c_int, [vp, C.c_int]),
    "mb200_neighbor_links_generate": (C.c_int, [vp, vp, P(Bodies), C.c_int, P(C.c_int)]),
    "mb200_neighbor_links_count": (c_i64, [vp]),
    "mb200_neighbor_links_get_pairs": (C.c_int, [vp, vp, vp, c_i64]),
    "mb200_search_pairs": (C.c_int, [vp, c_i64, vp, vp, vp, c_i64, P(c_i64)]),
    "mb200_neighbor_links_set_periodic": (C.c_int, [vp, P(c_dbl)]),
    "mb200_search_pairs_periodic": (C.c_int, [vp, c_i64, vp, vp, P(c_dbl), vp, c_i64, P(c_i64)]),
    "mb200_narrow_phase_periodic": (C.c_int, [vp, P(Bodies), c_i64, vp, P(c_dbl), vp, vp, vp, vp]),
    "mb200_narrow_phase_no_alloc": (C.c_int, [vp, P(Bodies), c_i64, vp, vp, vp, vp, vp]),
    "mb200_dist_repartition": (C.c_int, [vp, vp, vp, P(Bodies), vp, C.c_int, c_i64, P(Bodies), vp, vp, vp]),
    "mb200_hilbert_keys": (C.c_int, [vp, c_i64, vp, P(c_dbl), c_dbl, vp]),
    "mb200_periodic_sep": (C.c_int, [vp, c_i64, P(c_dbl), vp, vp, vp]),
    "mb200_periodic_wrap": (C.c_int, [vp, c_i64, P(c_dbl), vp, vp]),
    "mb200_periodic_unwrap_points": (C.c_int, [vp, c_i64, P(c_dbl), vp, vp, vp]),
    "mb200_periodic_unwrap_segments": (C.c_int, [vp, c_i64, P(c_dbl), vp, vp, vp, vp, vp]),
    "mb200_links_coo_to_crs": (C.c_int, [vp, c_i64, c_i64, vp, vp, vp, vp]),
    "mb200_apply_D": (C.c_int, [vp, P(ContactOp), vp, vp, vp]),
    "mb200_apply_M": (C.c_int, [vp, P(ContactOp), vp, vp, vp, vp]),
    "mb200_apply_DT": (C.c_int, [vp, P(ContactOp), vp, vp, vp]),
    "mb200_apply_A": (C.c_int, [vp, P(ContactOp), vp, vp]),
    "mb200_resolve_collisions_axial": (C.c_int, [vp, P(ContactOp), vp, vp, P(LcpParams), vp, vp, vp, vp, vp, vp, vp, vp,
                                                P(CollisionResult)]),
    "mb200_lcp_reduction_shape": (C.c_int, [c_i64, C.c_int, P(c_i32), P(c_i32)]),
    "mb200_resolve_collisions": (C.c_int, [vp, P(ContactOp), P(LcpParams), vp, vp, vp, vp, vp, vp, vp, vp,
                                           P(CollisionResult)]),
    "mb200_mobility_kernel_apply": (C.c_int, [vp, C.c_int, c_i64, c_i64, c_dbl, vp, vp, vp, vp, vp, vp]),
    "mb200_mobility_kernel_apply_neighbors": (C.c_int, [vp, C.c_int, c_i64, c_dbl, vp, vp, vp, vp, vp, vp]),
    "mb200_apply_local_drag": (C.c_int, [vp, C.c_int, c_i64, c_dbl, vp, vp, vp]),
    "mb200_mobility_apply": (C.c_int, [vp, P(MobilityOp), vp, vp]),
    "mb200_resolve_collisions_mobility": (C.c_int, [vp, P(ContactOp), P(MobilityOp), P(LcpParams), vp, vp, vp, vp, vp,
                                                    vp, P(CollisionResult)]),
    "mb200_vec_axpby": (C.c_int, [vp, c_i64, c_dbl, vp, c_dbl, vp]),
    "mb200_vec_wrapped_axpbyz": (C.c_int, [vp, c_i64, c_dbl, vp, c_dbl, vp, vp, C.c_int, c_dbl, c_dbl]),
    "mb200_vec_diff_dot": (C.c_int, [vp, c_i64, vp, vp, P(c_dbl)]),
    "mb200_vec_diff_dot2": (C.c_int, [vp, c_i64, vp, vp, vp, vp, P(c_dbl)]),
    "mb200_vec_residual": (C.c_int, [vp, c_i64, vp, vp, C.c_int, C.c_int, c_dbl, c_dbl, P(c_dbl)]),
    "mb200_dense_gemv": (C.c_int, [vp, c_i64, vp, vp, vp]),
    "mb200_solve_cqpp_dense": (C.c_int, [vp, c_i64, vp, vp, vp, P(PgdConfig), P(SolveResult)]),
    "mb200_solve_lcp_contact": (C.c_int, [vp, P(ContactOp), vp, vp, P(PgdConfig), P(SolveResult)]),
    "mb200_euler_update": (C.c_int, [vp, c_i64, vp, vp, vp, vp, vp, c_dbl]),
    "mb200_abc_flow_force": (C.c_int, [vp, c_i64, vp, vp]),
    "mb200_collision_step": (C.c_int, [vp, vp, P(Bodies), P(StepParams), vp, vp, vp, vp, vp, vp, P(StepStats)]),
    "mb200_comm_get_unique_id": (C.c_int, [vp]),
    "mb200_comm_create": (C.c_int, [P(vp), C.c_int, C.c_int, vp]),
    "mb200_comm_destroy": (C.c_int, [vp]),
    "mb200_comm_world": (C.c_int, [vp]),
    "mb200_comm_rank": (C.c_int, [vp]),
    "mb200_comm_exchange_mode": (C.c_int, [vp]),
    "mb200_collision_step_dist": (C.c_int, [vp, vp, vp, P(Bodies), P(StepParams), vp, vp, vp, vp, P(StepStats)]),
    "mb200_dist_local_counts": (C.c_int, [vp, P(c_i64), P(c_i64)]),
    "mb200_collision_step_host": (C.c_int, [vp, vp, P(Bodies), P(StepParams), vp, vp, vp, vp, vp, vp, P(StepStats)]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here == the library does not export a declared symbol
    _fn.restype = _res
    _fn.argtypes = _args


class Mb200Error(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        msg = lib.mb200_last_error_string()
        super().__init__(f"{where} failed with status {status}: {msg.decode() if msg else ''}")


def check(status, where):
    if status != 0:
        raise Mb200Error(status, where)
