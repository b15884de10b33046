"""Thin torch-facing wrappers over the C-ABI (torch supplies device memory and streams; nothing is computed here).

Every function takes CUDA float64/int32 tensors in the reference's layouts and forwards their data pointers to
libmundy_b200.so. They fail loudly (Mb200Error) when no GPU is present: there is no CPU path.
"""
import ctypes as C

import torch

from . import _capi as capi
from ._capi import lib, check


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "expected a contiguous CUDA tensor"
    return C.c_void_p(t.data_ptr())


def _f64(t):
    assert t is None or t.dtype == torch.float64
    return _p(t)


def _i32(t):
    assert t is None or t.dtype == torch.int32
    return _p(t)


def _cell(cell_size):
    """Host double[3] of the periodic cell edges (entry <= 0: free axis); None -> NULL (free space)."""
    if cell_size is None:
        return None
    return (C.c_double * 3)(*[float(x) for x in cell_size])


def _new(n, like, dtype=torch.float64):
    return torch.empty(n, dtype=dtype, device=like.device)


# ---------------------------------------------------------------------------------------------------------------
def compute_aabb_sphere(center, radius):
    n = radius.numel()
    out = _new((n, 6), center)
    check(lib.mb200_compute_aabb_sphere(_stream(), n, _f64(center), _f64(radius), _f64(out)), "compute_aabb_sphere")
    return out


def compute_aabb_spherocylinder(center, quat, radius, length):
    n = radius.numel()
    out = _new((n, 6), center)
    check(lib.mb200_compute_aabb_spherocylinder(_stream(), n, _f64(center), _f64(quat), _f64(radius), _f64(length),
                                                _f64(out)), "compute_aabb_spherocylinder")
    return out


def compute_aabb_segment(start, end, radius):
    n = radius.numel()
    out = _new((n, 6), start)
    check(lib.mb200_compute_aabb_segment(_stream(), n, _f64(start), _f64(end), _f64(radius), _f64(out)),
          "compute_aabb_segment")
    return out


def distance_sphere_sphere(c1, r1, c2, r2):
    n = r1.numel()
    dist, sep = _new(n, c1), _new((n, 3), c1)
    check(lib.mb200_distance_sphere_sphere(_stream(), n, _f64(c1), _f64(r1), _f64(c2), _f64(r2), _f64(dist),
                                           _f64(sep)), "distance_sphere_sphere")
    return dist, sep


def distance_segment_segment(l0, l1, m0, m1):
    n = l0.shape[0]
    dist, s, t = _new(n, l0), _new(n, l0), _new(n, l0)
    cp1, cp2, sep = _new((n, 3), l0), _new((n, 3), l0), _new((n, 3), l0)
    check(lib.mb200_distance_segment_segment(_stream(), n, _f64(l0), _f64(l1), _f64(m0), _f64(m1), _f64(dist), _f64(s),
                                             _f64(t), _f64(cp1), _f64(cp2), _f64(sep)), "distance_segment_segment")
    return dist, s, t, cp1, cp2, sep


def distance_point_segment(x, p1, p2):
    n = x.shape[0]
    dist, t, cp, sep = _new(n, x), _new(n, x), _new((n, 3), x), _new((n, 3), x)
    check(lib.mb200_distance_point_segment(_stream(), n, _f64(x), _f64(p1), _f64(p2), _f64(dist), _f64(t), _f64(cp),
                                           _f64(sep)), "distance_point_segment")
    return dist, t, cp, sep


def distance_segment_sphere(p1, p2, c, r):
    n = r.numel()
    dist, t, cp, sep = _new(n, c), _new(n, c), _new((n, 3), c), _new((n, 3), c)
    check(lib.mb200_distance_segment_sphere(_stream(), n, _f64(p1), _f64(p2), _f64(c), _f64(r), _f64(dist), _f64(t),
                                            _f64(cp), _f64(sep)), "distance_segment_sphere")
    return dist, t, cp, sep


# ---- the remaining primitives; argument order = the C-ABI's ------------------------------------
def _prim(name, n, like, inputs, widths):
    outs = [_new(n if w == 1 else (n, w), like) for w in widths]
    check(getattr(lib, "mb200_" + name)(_stream(), n, *[_f64(x) for x in inputs], *[_f64(o) for o in outs]), name)
    return outs[0] if len(outs) == 1 else tuple(outs)


def compute_aabb_point(point):
    return _prim("compute_aabb_point", point.shape[0], point, (point,), (6,))


def compute_aabb_line_segment(start, end):
    return _prim("compute_aabb_line_segment", start.shape[0], start, (start, end), (6,))


def compute_aabb_ellipsoid(center, quat, radii):
    return _prim("compute_aabb_ellipsoid", center.shape[0], center, (center, quat, radii), (6,))


def distance_point_sphere(x, c, r):
    """-> dist, sep"""
    return _prim("distance_point_sphere", x.shape[0], x, (x, c, r), (1, 3))


def distance_point_line(x, line_center, line_dir):
    """-> dist, closest point, arch length"""
    return _prim("distance_point_line", x.shape[0], x, (x, line_center, line_dir), (1, 3, 1))


def distance_line_line(c1, d1, c2, d2):
    """-> dist, cp1, cp2, arch1, arch2, sep"""
    return _prim("distance_line_line", c1.shape[0], c1, (c1, d1, c2, d2), (1, 3, 3, 1, 1, 3))


def distance_line_sphere(line_center, line_dir, c, r):
    """-> dist, closest point on the line, arch length"""
    return _prim("distance_line_sphere", c.shape[0], c, (line_center, line_dir, c, r), (1, 3, 1))


def distance_point_ellipsoid(x, center, quat, radii):
    """-> signed distance, closest point on the ellipsoid, ellipsoid normal there"""
    return _prim("distance_point_ellipsoid", x.shape[0], x, (x, center, quat, radii), (1, 3, 3))


def distance_ellipsoid_ellipsoid(c1, q1, radii1, c2, q2, radii2):
    """-> shared-normal signed distance, cp1, cp2, shared normal of the first"""
    return _prim("distance_ellipsoid_ellipsoid", c1.shape[0], c1, (c1, q1, radii1, c2, q2, radii2), (1, 3, 3, 3))


def distance_circle3d_circle3d(c1, q1, r1, c2, q2, r2):
    """-> Euclidean distance, cp1, cp2, unit vector from cp1 to cp2"""
    return _prim("distance_circle3d_circle3d", c1.shape[0], c1, (c1, q1, r1, c2, q2, r2), (1, 3, 3, 3))


# ---------------------------------------------------------------------------------------------------------------
def make_bodies(center, radius, quat=None, length=None, kind=None):
    b = capi.Bodies(radius.numel(), _f64(center), _f64(quat), _f64(radius), _f64(length), _i32(kind))
    b._keep = (center, radius, quat, length, kind)
    return b


def periodic_sep(cell_size, p1, p2):
    out = torch.empty_like(p1)
    check(lib.mb200_periodic_sep(_stream(), p1.shape[0], _cell(cell_size), _f64(p1), _f64(p2), _f64(out)),
          "periodic_sep")
    return out


def periodic_wrap(cell_size, p):
    out = torch.empty_like(p)
    check(lib.mb200_periodic_wrap(_stream(), p.shape[0], _cell(cell_size), _f64(p), _f64(out)), "periodic_wrap")
    return out


def periodic_unwrap_points(cell_size, p, ref):
    out = torch.empty_like(p)
    check(lib.mb200_periodic_unwrap_points(_stream(), p.shape[0], _cell(cell_size), _f64(p), _f64(ref), _f64(out)),
          "periodic_unwrap_points")
    return out


def periodic_unwrap_segments(cell_size, start, end, ref):
    o0, o1 = torch.empty_like(start), torch.empty_like(end)
    check(lib.mb200_periodic_unwrap_segments(_stream(), start.shape[0], _cell(cell_size), _f64(start), _f64(end),
                                             _f64(ref), _f64(o0), _f64(o1)), "periodic_unwrap_segments")
    return o0, o1


def narrow_phase(bodies, pairs, arms=False, cell_size=None):
    npairs = pairs.shape[0]
    ref = bodies._keep[0]
    phi, normal = _new(npairs, ref), _new((npairs, 3), ref)
    ai = _new((npairs, 3), ref) if arms else None
    aj = _new((npairs, 3), ref) if arms else None
    check(lib.mb200_narrow_phase_periodic(_stream(), C.byref(bodies), npairs, _i32(pairs), _cell(cell_size), _f64(phi),
                                          _f64(normal), _f64(ai), _f64(aj)), "narrow_phase")
    return phi, normal, ai, aj


NO_ALLOC_MAX_BODIES = 65536  # MB200_NO_ALLOC_MAX_BODIES


def narrow_phase_no_alloc(bodies, pairs, phi, normal, arm_i=None, arm_j=None):
    """mb200_narrow_phase_no_alloc: the caller passes every output; the call allocates nothing (include/mundy_b200.h)."""
    check(lib.mb200_narrow_phase_no_alloc(_stream(), C.byref(bodies), pairs.shape[0], _i32(pairs), _f64(phi),
                                          _f64(normal), _f64(arm_i), _f64(arm_j)), "narrow_phase_no_alloc")
    return phi, normal, arm_i, arm_j


def search_pairs(center, search_radius, capacity=None, cell_size=None):
    n = search_radius.numel()
    count = C.c_int64(0)
    if capacity is None:
        capacity = max(64 * n, 1024)
    out = torch.empty((capacity, 2), dtype=torch.int32, device=center.device)
    cell = _cell(cell_size)
    st = lib.mb200_search_pairs_periodic(_stream(), n, _f64(center), _f64(search_radius), cell, _i32(out), capacity,
                                         C.byref(count))
    if st == -3:  # capacity: retry with the exact size
        out = torch.empty((count.value, 2), dtype=torch.int32, device=center.device)
        st = lib.mb200_search_pairs_periodic(_stream(), n, _f64(center), _f64(search_radius), cell, _i32(out),
                                             count.value, C.byref(count))
    check(st, "search_pairs")
    return out[: count.value]


def links_coo_to_crs(n_entities, linked):
    """COO (n_links x 2 int32 entity indices, negative = empty slot) -> CRS: (offsets int64[n+1], links, slots)."""
    n_links = linked.shape[0]
    dev = linked.device
    offsets = torch.empty(n_entities + 1, dtype=torch.int64, device=dev)
    links = torch.empty(2 * n_links, dtype=torch.int32, device=dev)
    slots = torch.empty(2 * n_links, dtype=torch.int32, device=dev)
    check(lib.mb200_links_coo_to_crs(_stream(), n_entities, n_links, _i32(linked), _p(offsets), _i32(links),
                                     _i32(slots)), "links_coo_to_crs")
    total = int(offsets[-1].item())
    return offsets, links[:total], slots[:total]


class NeighborLinks:
    """mesh::GenNeighborLinks-like builder (mundy/mesh/src/mundy_mesh/GenNeighborLinkers.hpp:294-543)."""

    FILTER_NONE = 0
    FILTER_SHAPE_DISTANCE = 1

    def __init__(self):
        self._h = C.c_void_p()
        check(lib.mb200_neighbor_links_create(C.byref(self._h)), "neighbor_links_create")

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # lib is None during interpreter shutdown
            lib.mb200_neighbor_links_destroy(self._h)
            self._h = None

    def set_search_buffer(self, buffer):
        check(lib.mb200_neighbor_links_set_search_buffer(self._h, float(buffer)), "set_search_buffer")
        return self

    def set_search_filter(self, filt):
        check(lib.mb200_neighbor_links_set_search_filter(self._h, int(filt)), "set_search_filter")
        return self

    def set_periodic(self, cell_size):
        """Orthorhombic periodic box (cell edge per axis, <= 0: free axis); None switches it off."""
        check(lib.mb200_neighbor_links_set_periodic(self._h, _cell(cell_size)), "set_periodic")
        return self

    def generate(self, bodies, force=False):
        rebuilt = C.c_int(0)
        check(lib.mb200_neighbor_links_generate(self._h, _stream(), C.byref(bodies), int(force), C.byref(rebuilt)),
              "neighbor_links_generate")
        return bool(rebuilt.value)

    def count(self):
        return lib.mb200_neighbor_links_count(self._h)

    def pairs(self, device="cuda"):
        n = self.count()
        out = torch.empty((n, 2), dtype=torch.int32, device=device)
        check(lib.mb200_neighbor_links_get_pairs(self._h, _stream(), _i32(out), n), "neighbor_links_get_pairs")
        return out

    def step(self, bodies, viscosity, dt, tol, max_iters, fixed_iters=-1, throw_on_max_iters=False,
             force_rebuild=False, F_ext=None, T_ext=None, F_col=None, T_col=None, U_out=None, W_out=None):
        params = capi.StepParams(viscosity, dt, tol, max_iters, fixed_iters, int(throw_on_max_iters),
                                 int(force_rebuild))
        stats = capi.StepStats()
        check(lib.mb200_collision_step(self._h, _stream(), C.byref(bodies), C.byref(params), _f64(F_ext), _f64(T_ext),
                                       _f64(F_col), _f64(T_col), _f64(U_out), _f64(W_out), C.byref(stats)),
              "collision_step")
        return stats

    def step_host(self, center, radius, quat, length, kind, viscosity, dt, tol, max_iters, fixed_iters=-1,
                  F_ext=None, T_ext=None, F_col=None, T_col=None, U_out=None, W_out=None, force_rebuild=False):
        """Host (CPU, ideally pinned) tensors in, advanced in place. H2D + step + D2H inside the call."""
        def hp(t):
            if t is None:
                return None
            assert (not t.is_cuda) and t.is_contiguous()
            return C.c_void_p(t.data_ptr())
        b = capi.Bodies(radius.numel(), hp(center), hp(quat), hp(radius), hp(length), hp(kind))
        params = capi.StepParams(viscosity, dt, tol, max_iters, fixed_iters, 0, int(force_rebuild))
        stats = capi.StepStats()
        check(lib.mb200_collision_step_host(self._h, _stream(), C.byref(b), C.byref(params), hp(F_ext), hp(T_ext),
                                            hp(F_col), hp(T_col), hp(U_out), hp(W_out), C.byref(stats)),
              "collision_step_host")
        return stats


# ---------------------------------------------------------------------------------------------------------------
def make_contact_op(n, pairs, normal, radius, viscosity, dt, arm_i=None, arm_j=None, quat=None, kind=None):
    op = capi.ContactOp(n, pairs.shape[0], _i32(pairs), _f64(normal), _f64(arm_i), _f64(arm_j), _f64(quat),
                        _f64(radius), _i32(kind), viscosity, dt)
    op._keep = (pairs, normal, arm_i, arm_j, quat, radius, kind)
    return op


def apply_D(op, gamma):
    dev = gamma.device
    six = bool(op.arm_i)
    F = torch.empty((op.n, 3), dtype=torch.float64, device=dev)
    T = torch.empty((op.n, 3), dtype=torch.float64, device=dev) if six else None
    check(lib.mb200_apply_D(_stream(), C.byref(op), _f64(gamma), _f64(F), _f64(T)), "apply_D")
    return F, T


def apply_M(op, F, T=None):
    U = torch.empty_like(F)
    W = torch.empty_like(T) if T is not None else None
    check(lib.mb200_apply_M(_stream(), C.byref(op), _f64(F), _f64(T), _f64(U), _f64(W)), "apply_M")
    return U, W


def apply_DT(op, U, W=None):
    g = torch.empty(op.npairs, dtype=torch.float64, device=U.device)
    check(lib.mb200_apply_DT(_stream(), C.byref(op), _f64(U), _f64(W), _f64(g)), "apply_DT")
    return g


def apply_A(op, x):
    y = torch.empty_like(x)
    check(lib.mb200_apply_A(_stream(), C.byref(op), _f64(x), _f64(y)), "apply_A")
    return y


def resolve_collisions(op, phi0, gamma, tol, max_iters, fixed_iters=-1, U_ext=None, W_ext=None):
    """phi0 and gamma are updated in place. Returns (result, F, T, U, W)."""
    dev = phi0.device
    six = bool(op.arm_i)
    F = torch.empty((op.n, 3), dtype=torch.float64, device=dev)
    U = torch.empty((op.n, 3), dtype=torch.float64, device=dev)
    T = torch.empty((op.n, 3), dtype=torch.float64, device=dev) if six else None
    W = torch.empty((op.n, 3), dtype=torch.float64, device=dev) if six else None
    params = capi.LcpParams(tol, max_iters, fixed_iters)
    res = capi.CollisionResult()
    check(lib.mb200_resolve_collisions(_stream(), C.byref(op), C.byref(params), _f64(U_ext), _f64(W_ext), _f64(phi0),
                                       _f64(gamma), _f64(F), _f64(T), _f64(U), _f64(W), C.byref(res)),
          "resolve_collisions")
    return res, F, T, U, W


ARM_NONE, ARM_AXIAL, ARM_GENERAL = 0, 1, 2


def resolve_collisions_axial(op, sigma, length, phi0, gamma, tol, max_iters, fixed_iters=-1, U_ext=None, W_ext=None):
    """Arms in axial form (sigma (npairs, 2), length (n,)); phi0 and gamma are updated in place. Returns (result, F, T, U, W)."""
    dev = phi0.device
    F, U, T, W = (torch.empty((op.n, 3), dtype=torch.float64, device=dev) for _ in range(4))
    params = capi.LcpParams(tol, max_iters, fixed_iters)
    res = capi.CollisionResult()
    check(lib.mb200_resolve_collisions_axial(_stream(), C.byref(op), _f64(sigma), _f64(length), C.byref(params),
                                             _f64(U_ext), _f64(W_ext), _f64(phi0), _f64(gamma), _f64(F), _f64(T),
                                             _f64(U), _f64(W), C.byref(res)), "resolve_collisions_axial")
    return res, F, T, U, W


def lcp_reduction_shape(npairs, arm_mode):
    """(grid, threads) of the solver's deterministic global reductions for npairs contacts (see the header)."""
    g, t = C.c_int32(), C.c_int32()
    check(lib.mb200_lcp_reduction_shape(int(npairs), int(arm_mode), C.byref(g), C.byref(t)), "lcp_reduction_shape")
    return g.value, t.value


# G. non-local mobility operators ---------------------------------------------------------------
KERNEL_STOKES, KERNEL_RPY, KERNEL_RPYC, KERNEL_RPY_EXACT = 0, 1, 2, 3
MOBILITY_LOCAL_DRAG, MOBILITY_NEIGHBORS, MOBILITY_ALL = 0, 1, 2


def mobility_kernel_apply(kernel, viscosity, target_pos, source_pos, source_radii, target_radii, source_force,
                          target_vel):
    """target_vel += sum over all sources (in place)."""
    check(lib.mb200_mobility_kernel_apply(_stream(), kernel, target_pos.shape[0], source_pos.shape[0], viscosity,
                                          _f64(target_pos), _f64(source_pos), _f64(source_radii), _f64(target_radii),
                                          _f64(source_force), _f64(target_vel)), "mobility_kernel_apply")
    return target_vel


def mobility_kernel_apply_neighbors(kernel, viscosity, pos, radii, offsets, indices, force, vel):
    check(lib.mb200_mobility_kernel_apply_neighbors(_stream(), kernel, pos.shape[0], viscosity, _f64(pos), _f64(radii),
                                                    _i32(offsets), _i32(indices), _f64(force), _f64(vel)),
          "mobility_kernel_apply_neighbors")
    return vel


def apply_local_drag(kernel, viscosity, radii, force, vel):
    check(lib.mb200_apply_local_drag(_stream(), kernel, radii.numel(), viscosity, _f64(radii), _f64(force), _f64(vel)),
          "apply_local_drag")
    return vel


def make_mobility_op(mode, kernel, viscosity, radius, pos=None, offsets=None, indices=None):
    op = capi.MobilityOp(radius.numel(), mode, kernel, viscosity, _f64(pos), _f64(radius), _i32(offsets), _i32(indices))
    op._keep = (pos, radius, offsets, indices)
    return op


def mobility_apply(mob, F):
    U = torch.empty_like(F)
    check(lib.mb200_mobility_apply(_stream(), C.byref(mob), _f64(F), _f64(U)), "mobility_apply")
    return U


def resolve_collisions_mobility(op, mob, phi0, gamma, tol, max_iters, fixed_iters=-1, F_ext=None):
    """phi0 and gamma are updated in place. Returns (result, F, U, U_ext)."""
    dev = phi0.device
    F = torch.empty((op.n, 3), dtype=torch.float64, device=dev)
    U = torch.empty((op.n, 3), dtype=torch.float64, device=dev)
    Ue = torch.empty((op.n, 3), dtype=torch.float64, device=dev) if F_ext is not None else None
    params = capi.LcpParams(tol, max_iters, fixed_iters)
    res = capi.CollisionResult()
    check(lib.mb200_resolve_collisions_mobility(_stream(), C.byref(op), C.byref(mob), C.byref(params), _f64(F_ext),
                                                _f64(Ue), _f64(phi0), _f64(gamma), _f64(F), _f64(U), C.byref(res)),
          "resolve_collisions_mobility")
    return res, F, U, Ue


def solve_cqpp_dense(A, q, x, max_iters=1000, tol=1e-8, space_kind=1, lo=0.0, hi=0.0, residual_kind=0):
    cfg = capi.PgdConfig(max_iters, tol, space_kind, lo, hi, residual_kind)
    res = capi.SolveResult()
    check(lib.mb200_solve_cqpp_dense(_stream(), q.numel(), _f64(A), _f64(q), _f64(x), C.byref(cfg), C.byref(res)),
          "solve_cqpp_dense")
    return res


def solve_lcp_contact(op, phi0, gamma, max_iters=1000, tol=1e-8, residual_kind=0):
    cfg = capi.PgdConfig(max_iters, tol, 1, 0.0, 0.0, residual_kind)
    res = capi.SolveResult()
    check(lib.mb200_solve_lcp_contact(_stream(), C.byref(op), _f64(phi0), _f64(gamma), C.byref(cfg), C.byref(res)),
          "solve_lcp_contact")
    return res


def euler_update(center, U, dt, quat=None, kind=None, W=None):
    check(lib.mb200_euler_update(_stream(), center.shape[0], _f64(center), _f64(quat), _i32(kind), _f64(U), _f64(W),
                                 dt), "euler_update")


def abc_flow_force(center, F):
    """F += attractive ABC flow force at `center` (in place)."""
    check(lib.mb200_abc_flow_force(_stream(), center.shape[0], _f64(center), _f64(F)), "abc_flow_force")
    return F


def kernel_launch_count():
    return lib.mb200_kernel_launch_count()


STAGES = ("trigger", "broadphase", "pack", "narrow", "lcp_setup", "lcp_body", "lcp_rate", "euler", "lcp_comm",
          "state_comm")


def profile_enable(on=True):
    """on = True / 1: every region is bracketed; on = k > 1: every k-th region of the per-launch LCP stages."""
    check(lib.mb200_profile_enable(int(on)), "profile_enable")


def profile_calls():
    """{stage: regions entered since profile_enable, bracketed or not}."""
    out = {}
    for k, name in enumerate(STAGES):
        cnt = C.c_int64(0)
        check(lib.mb200_profile_calls(k, C.byref(cnt)), "profile_calls")
        out[name] = cnt.value
    return out


def profile_read():
    """{stage: (device ms accumulated since profile_enable(True), number of bracketed regions)}."""
    out = {}
    for k, name in enumerate(STAGES):
        ms, cnt = C.c_double(0.0), C.c_int64(0)
        check(lib.mb200_profile_read(k, C.byref(ms), C.byref(cnt)), "profile_read")
        out[name] = (ms.value, cnt.value)
    return out
