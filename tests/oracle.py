"""numpy/ctypes front-end to the CPU oracle (oracle/liboracle.so) and, when present, to the reference's own headers
compiled behind the Kokkos shim (oracle/_ref/libmundy_ref.so). TEST INFRASTRUCTURE ONLY: only tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs may import this."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libmundy_ref.so")


def _build():
    if not os.path.exists(ORACLE_SO) or (os.path.isdir("/root/reference/mundy") and not os.path.exists(REF_SO)):
        subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True, capture_output=True)


_build()
orc = C.CDLL(ORACLE_SO)
ref = C.CDLL(REF_SO) if os.path.exists(REF_SO) else None

_d = np.float64
_i = np.int32


def _a(x, dtype=_d):
    return None if x is None else np.ascontiguousarray(x, dtype=dtype)


def _p(x):
    return None if x is None else x.ctypes.data_as(C.c_void_p)


class CollisionResult(C.Structure):
    _fields_ = [("max_abs_projected_sep", C.c_double), ("ite_count", C.c_int), ("max_displacement", C.c_double)]


class ContactOpCtx(C.Structure):
    _fields_ = [("n", C.c_long), ("npairs", C.c_long), ("pairs", C.c_void_p), ("normal", C.c_void_p),
                ("arm_i", C.c_void_p), ("arm_j", C.c_void_p), ("quat", C.c_void_p), ("r", C.c_void_p),
                ("kind", C.c_void_p), ("mu", C.c_double), ("dt", C.c_double)]


def _call(lib, name, *args, restype=None):
    fn = getattr(lib, name)
    fn.restype = restype
    conv = []
    for a in args:
        if isinstance(a, np.ndarray) or a is None:
            conv.append(_p(a))
        elif isinstance(a, float):
            conv.append(C.c_double(a))
        elif isinstance(a, (int, np.integer)):
            conv.append(C.c_long(int(a)))
        else:
            conv.append(a)
    return fn(*conv)


# ---- leaves (available from both libraries: prefix orc_ / ref_) ---------------------------------------------------
def aabb_sphere(pos, r, lib=None, prefix="orc_"):
    lib = lib or orc
    pos, r = _a(pos), _a(r)
    out = np.empty((len(r), 6))
    _call(lib, prefix + "aabb_sphere", len(r), pos, r, out)
    return out


def aabb_spherocylinder(pos, quat, r, length, lib=None, prefix="orc_"):
    lib = lib or orc
    pos, quat, r, length = _a(pos), _a(quat), _a(r), _a(length)
    out = np.empty((len(r), 6))
    _call(lib, prefix + "aabb_spherocylinder", len(r), pos, quat, r, length, out)
    return out


def aabb_segment(p0, p1, r, lib=None, prefix="orc_"):
    lib = lib or orc
    p0, p1, r = _a(p0), _a(p1), _a(r)
    out = np.empty((len(r), 6))
    _call(lib, prefix + "aabb_segment", len(r), p0, p1, r, out)
    return out


def quat_rotate(quat, v, lib=None, prefix="orc_"):
    lib = lib or orc
    quat, v = _a(quat), _a(v)
    out = np.empty_like(v)
    _call(lib, prefix + "quat_rotate", len(v), quat, v, out)
    return out


def rotate_quaternion(quat, omega, dt, lib=None, prefix="orc_"):
    lib = lib or orc
    q = _a(quat).copy()
    _call(lib, prefix + "rotate_quaternion", len(q), q, _a(omega), float(dt))
    return q


# ---- space-filling-curve specifications, reference library only -----------------------------------
def ref_hilbert_3d(s):
    """Lattice points of the reference's hilbert_3d(s) in visiting order, (s^3, 3) ints (needs oracle/_ref)."""
    out = np.zeros((s ** 3, 3))
    _call(ref, "ref_hilbert_3d", int(s), out)
    return out.astype(np.int64)


def ref_zorder_less(p, q, which=0):
    """which = 0: zorder_knn::Less<Point, 3>; 1: mundy::math::zmorton_less (zmort.hpp). Returns a bool array."""
    p, q = _a(p), _a(q)
    out = np.zeros(len(p), dtype=np.int32)
    fn = ref.ref_zorder_less
    fn.restype = None
    fn(C.c_long(len(p)), C.c_int(which), _p(p), _p(q), _p(out))
    return out.astype(bool)


# ---- remaining primitives: one generic front-end, outputs shaped by PRIMITIVE_OUTPUTS ------------
# name -> per-item widths of the output arrays, in call order
PRIMITIVE_OUTPUTS = {
    "aabb_point": (6,), "aabb_line_segment": (6,), "aabb_ellipsoid": (6,),
    "point_sphere": (1, 3),                    # dist, sep
    "point_line": (1, 3, 1),                   # dist, closest point, arch length
    "line_line": (1, 3, 3, 1, 1, 3),           # dist, cp1, cp2, arch1, arch2, sep
    "line_sphere": (1, 3, 1),                  # dist, closest point, arch length
    "point_ellipsoid": (1, 3, 3),              # dist, closest point, ellipsoid normal
    "ellipsoid_ellipsoid": (1, 3, 3, 3),       # dist, cp1, cp2, shared normal of the first
    "circle_circle": (1, 3, 3, 3),             # dist, cp1, cp2, unit vector from cp1 to cp2
}


def primitive(name, *inputs, lib=None, prefix="orc_"):
    lib = lib or orc
    ins = [_a(x) for x in inputs]
    n = len(ins[0])
    outs = [np.zeros(n) if w == 1 else np.zeros((n, w)) for w in PRIMITIVE_OUTPUTS[name]]
    _call(lib, prefix + name, n, *ins, *outs)
    return outs[0] if len(outs) == 1 else tuple(outs)


def segseg(l0, l1, m0, m1, lib=None, prefix="orc_"):
    lib = lib or orc
    l0, l1, m0, m1 = _a(l0), _a(l1), _a(m0), _a(m1)
    n = len(l0)
    dist, s, t = np.empty(n), np.empty(n), np.empty(n)
    cp1, cp2, sep = np.empty((n, 3)), np.empty((n, 3)), np.empty((n, 3))
    _call(lib, prefix + "segseg", n, l0, l1, m0, m1, dist, s, t, cp1, cp2, sep)
    return dist, s, t, cp1, cp2, sep


def point_seg(x, p1, p2, lib=None, prefix="orc_"):
    lib = lib or orc
    x, p1, p2 = _a(x), _a(p1), _a(p2)
    n = len(x)
    dist, t, cp, sep = np.empty(n), np.empty(n), np.empty((n, 3)), np.empty((n, 3))
    _call(lib, prefix + "point_seg", n, x, p1, p2, dist, t, cp, sep)
    return dist, t, cp, sep


def sphere_sphere(c1, r1, c2, r2, lib=None, prefix="orc_"):
    lib = lib or orc
    c1, r1, c2, r2 = _a(c1), _a(r1), _a(c2), _a(r2)
    n = len(r1)
    dist, sep = np.empty(n), np.empty((n, 3))
    _call(lib, prefix + "sphere_sphere", n, c1, r1, c2, r2, dist, sep)
    return dist, sep


def seg_sphere(p1, p2, c, r, lib=None, prefix="orc_"):
    lib = lib or orc
    p1, p2, c, r = _a(p1), _a(p2), _a(c), _a(r)
    n = len(r)
    dist, t, cp, sep = np.empty(n), np.empty(n), np.empty((n, 3)), np.empty((n, 3))
    _call(lib, prefix + "seg_sphere", n, p1, p2, c, r, dist, t, cp, sep)
    return dist, t, cp, sep


# ---- collision path (oracle only) -----------------------------------------------------------------------------------
def search_radii(r, buffer, kind=None, length=None):
    r = _a(r)
    S = np.empty(len(r))
    _call(orc, "orc_search_radii", len(r), _a(kind, _i), r, _a(length), float(buffer), S)
    return S


def _cell(cell):
    return None if cell is None else np.ascontiguousarray(cell, dtype=_d)


PERIODIC_MODES = {7: "PeriodicScaledMetric", 1: "PeriodicMetricX", 2: "PeriodicMetricY", 3: "PeriodicMetricXY",
                  6: "PeriodicMetricYZ"}


def periodic(op, cell, a, b=None, c=None, lib=None):
    """Periodic leaves. op: 0 sep(a,b) | 1 wrap(a) | 2 unwrap point a to ref b | 3 unwrap segment (a,c) to ref b
    (| 4, 5: centre of the unwrapped Sphere / Spherocylinder, reference library only).
    lib=None: the oracle restatement; lib=ref: the reference's own classes (mode = mask of the axes with cell > 0)."""
    cell = _cell(cell)
    a, b, c = _a(a), _a(b), _a(c)
    o0 = np.empty_like(a)
    o1 = np.empty_like(a) if op == 3 else None
    if lib is None:
        fn = orc.orc_periodic
        fn.restype = None
        fn(C.c_int(op), C.c_long(len(a)), _p(cell), _p(a), _p(b), _p(c), _p(o0), _p(o1))
    else:
        mode = sum(1 << d for d in range(3) if cell[d] > 0)
        fn = lib.ref_periodic
        fn.restype = C.c_int
        st = fn(C.c_int(mode), C.c_int(op), C.c_long(len(a)), _p(cell), _p(a), _p(b), _p(c), _p(o0), _p(o1))
        assert st == 0, f"no reference metric for periodic axes mask {mode}"
    return (o0, o1) if op == 3 else o0


def search(pos, S, brute=False, cell=None):
    pos, S = _a(pos), _a(S)
    n = len(S)
    fn = "orc_search_brute_periodic" if brute else "orc_search_cells_periodic"
    cell = _cell(cell)
    cap = max(1024, 64 * n)
    while True:
        pairs = np.empty((cap, 2), dtype=_i)
        cnt = _call(orc, fn, n, pos, S, cell, pairs, cap, restype=C.c_long)
        if cnt <= cap:
            return pairs[:cnt].copy()
        cap = cnt


def filter_shape_distance(pairs, pos, quat, r, length, kind, cutoff, cell=None):
    pairs = _a(pairs, _i)
    out = np.empty_like(pairs)
    cnt = _call(orc, "orc_filter_shape_distance_periodic", len(pairs), pairs, _a(pos), _a(quat), _a(r), _a(length),
                _a(kind, _i), float(cutoff), _cell(cell), out, restype=C.c_long)
    return out[:cnt].copy()


def moved_too_much(pos, pos_at_build, buffer, cell=None):
    return bool(_call(orc, "orc_moved_too_much_periodic", len(pos), _a(pos), _a(pos_at_build), float(buffer),
                      _cell(cell), restype=C.c_int))


def moved_too_much_shapes(pos, quat, length, kind, pos_at_build, quat_at_build, buffer, cell=None):
    return bool(_call(orc, "orc_moved_too_much_shapes_periodic", len(pos), _a(pos), _a(quat), _a(length),
                      _a(kind, _i), _a(pos_at_build), _a(quat_at_build), float(buffer), _cell(cell),
                      restype=C.c_int))


def narrow_phase(pairs, pos, r, quat=None, length=None, kind=None, arms=False, cell=None, sigma=False):
    """sigma=True additionally returns the axial form of the arms, (m, 2): arm_side = sigma_side * half-axis of the body."""
    pairs = _a(pairs, _i)
    m = len(pairs)
    phi, normal = np.empty(m), np.empty((m, 3))
    ai = np.empty((m, 3)) if arms else None
    aj = np.empty((m, 3)) if arms else None
    sg = np.zeros((m, 2)) if sigma else None
    _call(orc, "orc_narrow_phase_full", m, pairs, _a(pos), _a(quat), _a(r), _a(length), _a(kind, _i), _cell(cell),
          phi, normal, ai, aj, sg)
    return (phi, normal, ai, aj, sg) if sigma else (phi, normal, ai, aj)


def apply_D(n, pairs, normal, gamma, arm_i=None, arm_j=None):
    pairs = _a(pairs, _i)
    F = np.empty((n, 3))
    T = np.empty((n, 3)) if arm_i is not None else None
    _call(orc, "orc_apply_D", n, len(pairs), pairs, _a(normal), _a(arm_i), _a(arm_j), _a(gamma), F, T)
    return F, T


def apply_M(r, mu, F, T=None, quat=None, kind=None):
    U = np.empty_like(_a(F))
    W = np.empty_like(U) if T is not None else None
    _call(orc, "orc_apply_M", len(r), _a(quat), _a(r), _a(kind, _i), float(mu), _a(F), _a(T), U, W)
    return U, W


def apply_DT(n, pairs, normal, U, W=None, arm_i=None, arm_j=None):
    pairs = _a(pairs, _i)
    g = np.empty(len(pairs))
    _call(orc, "orc_apply_DT", n, len(pairs), pairs, _a(normal), _a(arm_i), _a(arm_j), _a(U), _a(W), g)
    return g


def apply_A(n, pairs, normal, r, mu, dt, x, arm_i=None, arm_j=None, quat=None, kind=None):
    pairs = _a(pairs, _i)
    y = np.empty(len(pairs))
    _call(orc, "orc_apply_A", n, len(pairs), pairs, _a(normal), _a(arm_i), _a(arm_j), _a(quat), _a(r), _a(kind, _i),
          float(mu), float(dt), _a(x), y)
    return y


def resolve_collisions(n, pairs, normal, r, mu, dt, tol, max_it, phi0, gamma0=None, fixed_iters=-1, arm_i=None,
                       arm_j=None, quat=None, kind=None, U_ext=None, W_ext=None, sigma=None, length=None):
    """Returns dict(res, it, max_disp, gamma, phi0, F, T, U, W). Inputs are not modified.
    sigma (m, 2) + length: arms in axial form (what the device's step runs for spherocylinders) instead of arm_i/arm_j."""
    pairs = _a(pairs, _i)
    m = len(pairs)
    phi0 = _a(phi0).copy()
    gamma = np.zeros(m) if gamma0 is None else _a(gamma0).copy()
    six = arm_i is not None or sigma is not None
    F, U = np.zeros((n, 3)), np.zeros((n, 3))
    T = np.zeros((n, 3)) if six else None
    W = np.zeros((n, 3)) if six else None
    res = CollisionResult()
    if sigma is not None:
        fn = orc.orc_resolve_collisions_axial
        fn.restype = None
        fn(C.c_long(n), C.c_long(m), _p(pairs), _p(_a(normal)), _p(_a(sigma)), _p(_a(quat)), _p(_a(length)), _p(_a(r)),
           _p(_a(kind, _i)), C.c_double(mu), C.c_double(dt), C.c_double(tol), C.c_int(max_it), C.c_int(fixed_iters),
           _p(_a(U_ext)), _p(_a(W_ext)), _p(phi0), _p(gamma), _p(F), _p(T), _p(U), _p(W), C.byref(res))
        return dict(res=res.max_abs_projected_sep, it=res.ite_count, max_disp=res.max_displacement, gamma=gamma,
                    phi0=phi0, F=F, T=T, U=U, W=W)
    fn = orc.orc_resolve_collisions
    fn.restype = None
    fn(C.c_long(n), C.c_long(m), _p(pairs), _p(_a(normal)), _p(_a(arm_i)), _p(_a(arm_j)), _p(_a(quat)), _p(_a(r)),
       _p(_a(kind, _i)), C.c_double(mu), C.c_double(dt), C.c_double(tol), C.c_int(max_it), C.c_int(fixed_iters),
       _p(_a(U_ext)), _p(_a(W_ext)), _p(phi0), _p(gamma), _p(F), _p(T), _p(U), _p(W), C.byref(res))
    return dict(res=res.max_abs_projected_sep, it=res.ite_count, max_disp=res.max_displacement, gamma=gamma,
                phi0=phi0, F=F, T=T, U=U, W=W)


KERNEL_STOKES, KERNEL_RPY, KERNEL_RPYC, KERNEL_RPY_EXACT = 0, 1, 2, 3
MOBILITY_LOCAL_DRAG, MOBILITY_NEIGHBORS, MOBILITY_ALL = 0, 1, 2


def mobility_kernel(kind, mu, tpos, spos, srad, trad, sforce, tvel=None, lib=None):
    """tvel (+)= sum over all sources of the pair kernel `kind` (Periphery.hpp apply_{stokes,rpy,rpyc}_kernel; kind 3 =
    KokkosLCP.cpp RPYKernel). lib=oracle.ref runs the reference's own header (kinds 0-2)."""
    tpos, spos = _a(tpos), _a(spos)
    out = np.zeros_like(tpos) if tvel is None else _a(tvel).copy()
    if lib is None:
        _call(orc, "orc_mobility_kernel", int(kind), len(tpos), len(spos), float(mu), tpos, spos, _a(srad), _a(trad),
              _a(sforce), out)
    else:
        rc = _call(lib, "ref_mobility_kernel", int(kind), len(tpos), len(spos), float(mu), tpos, spos, _a(srad),
                   _a(trad), _a(sforce), out, restype=C.c_int)
        assert rc == 0
    return out


def mobility_kernel_neighbors(kind, mu, pos, rad, offsets, indices, force, vel=None):
    pos = _a(pos)
    out = np.zeros_like(pos) if vel is None else _a(vel).copy()
    _call(orc, "orc_mobility_kernel_neighbors", int(kind), len(pos), float(mu), pos, _a(rad), _a(offsets, _i),
          _a(indices, _i), _a(force), out)
    return out


def local_drag(kind, mu, rad, force, vel=None, lib=None):
    force = _a(force)
    out = np.zeros_like(force) if vel is None else _a(vel).copy()
    if lib is None:
        _call(orc, "orc_local_drag", int(kind), len(force), float(mu), _a(rad), force, out)
    else:
        _call(lib, "ref_local_drag", len(force), float(mu), _a(rad), force, out, restype=C.c_int)
    return out


def mobility_apply(mode, kind, mu, pos, rad, F, offsets=None, indices=None):
    pos = _a(pos)
    U = np.empty_like(pos)
    _call(orc, "orc_mobility_apply", int(mode), int(kind), len(pos), float(mu), pos, _a(rad), _a(offsets, _i),
          _a(indices, _i), _a(F), U)
    return U


def resolve_collisions_mobility(pairs, normal, mode, kind, mu, pos, rad, dt, tol, max_it, phi0, gamma0=None,
                                fixed_iters=-1, offsets=None, indices=None, F_ext=None):
    """BBPGD with the mobility behind an operator (NgpHP1.cpp:1487-1645). Returns dict(res, it, max_disp, gamma, phi0, F,
    U, U_ext)."""
    pairs, pos = _a(pairs, _i), _a(pos)
    n, m = len(pos), len(pairs)
    phi0 = _a(phi0).copy()
    gamma = np.zeros(m) if gamma0 is None else _a(gamma0).copy()
    F, U = np.zeros((n, 3)), np.zeros((n, 3))
    Ue = np.zeros((n, 3)) if F_ext is not None else None
    res = CollisionResult()
    fn = orc.orc_resolve_collisions_mobility
    fn.restype = None
    fn(C.c_long(n), C.c_long(m), _p(pairs), _p(_a(normal)), C.c_int(mode), C.c_int(kind), C.c_double(mu), _p(pos),
       _p(_a(rad)), _p(_a(offsets, _i)), _p(_a(indices, _i)), C.c_double(dt), C.c_double(tol), C.c_int(max_it),
       C.c_int(fixed_iters), _p(_a(F_ext)), _p(Ue), _p(phi0), _p(gamma), _p(F), _p(U), C.byref(res))
    return dict(res=res.max_abs_projected_sep, it=res.ite_count, max_disp=res.max_displacement, gamma=gamma, phi0=phi0,
                F=F, U=U, U_ext=Ue)


def euler_update(pos, U, dt, quat=None, kind=None, W=None):
    pos = _a(pos).copy()
    q = _a(quat).copy() if quat is not None else None
    _call(orc, "orc_euler_update", len(pos), pos, q, _a(kind, _i), _a(U), _a(W), float(dt))
    return pos, q


def abc_flow_force(pos):
    F = np.zeros_like(_a(pos))
    _call(orc, "orc_abc_flow_force", len(pos), _a(pos), F)
    return F


# ---- whole step (composition of the pieces above; the loop body of
# scrap/parameter_interface/NGPSpheresLCP.cpp:1135-1292) ---------------------------------------------------------
def collision_step(state, w, links, buffer, filt, mu, dt, tol, max_it, fixed, F_ext, cell=None):
    """One collision step composed from the oracle pieces; `links` carries (pairs, ref_center) between steps."""
    c, q = state
    kind, L, r = w.get("kind"), w.get("length"), w["radius"]
    if links is not None:  # rebuild trigger of GenNeighborLinks::generate (GenNeighborLinkers.hpp:603-615)
        if filt and kind is not None:
            if moved_too_much_shapes(c, q, L, kind, links[1], links[2], buffer, cell):
                links = None
        elif moved_too_much(c, links[1], buffer, cell):
            links = None
    rebuilt = links is None
    if links is None:
        S = search_radii(r, buffer, kind, L)
        pairs = search(c, S, cell=cell)
        if filt and kind is not None:
            pairs = filter_shape_distance(pairs, c, q, r, L, kind, 2 * buffer, cell)
        links = (pairs, c.copy(), None if q is None else q.copy())
    pairs = links[0]
    arms = kind is not None
    phi, nrm, _, _, sg = narrow_phase(pairs, c, r, q, L, kind, False, cell, sigma=True)
    Uext = Wext = None
    if F_ext is not None:
        Uext, Wext = apply_M(r, mu, F_ext, np.zeros_like(F_ext) if arms else None, q, kind)
    # spherocylinders: arms in axial form, as the device's step
    out = resolve_collisions(len(r), pairs, nrm, r, mu, dt, tol, max_it, phi, fixed_iters=fixed, quat=q, kind=kind,
                             U_ext=Uext, W_ext=Wext, sigma=sg if arms else None, length=L)
    U = out["U"] + (Uext if Uext is not None else 0.0)
    Wt = (out["W"] + (Wext if Wext is not None else 0.0)) if arms else None
    c2, q2 = euler_update(c, U, dt, q, kind, Wt)
    out["rebuilt"] = rebuilt
    return (c2, q2), links, out, U


# ---- math::convex (both libraries) ------------------------------------------------------------------------------------
def solve_cqpp_dense(A, q, x0, space_kind=1, lo=0.0, hi=0.0, max_iters=1000, tol=1e-8, resid_kind=0, lib=None,
                     prefix="orc_"):
    lib = lib or orc
    A, q = _a(A), _a(q)
    x = _a(x0).copy()
    iters, resid = C.c_uint(0), C.c_double(0.0)
    fn = getattr(lib, prefix + "solve_cqpp_dense")
    fn.restype = C.c_int
    conv = fn(C.c_long(len(q)), _p(A), _p(q), _p(x), C.c_int(space_kind), C.c_double(lo), C.c_double(hi),
              C.c_uint(max_iters), C.c_double(tol), C.c_int(resid_kind), C.byref(iters), C.byref(resid))
    return dict(x=x, iters=iters.value, residual=resid.value, converged=bool(conv))


def solve_contact_lcp_convex(n, pairs, normal, r, mu, dt, phi0, gamma0, max_iters=1000, tol=1e-8, resid_kind=0,
                             arm_i=None, arm_j=None, quat=None, kind=None):
    pairs = _a(pairs, _i)
    keep = [pairs, _a(normal), _a(arm_i), _a(arm_j), _a(quat), _a(r), _a(kind, _i)]
    ctx = ContactOpCtx(n, len(pairs), *[_p(k) for k in keep], mu, dt)
    gamma = _a(gamma0).copy()
    iters, resid = C.c_uint(0), C.c_double(0.0)
    fn = orc.orc_solve_contact_lcp_convex
    fn.restype = C.c_int
    conv = fn(C.byref(ctx), _p(_a(phi0)), _p(gamma), C.c_uint(max_iters), C.c_double(tol), C.c_int(resid_kind),
              C.byref(iters), C.byref(resid))
    return dict(x=gamma, iters=iters.value, residual=resid.value, converged=bool(conv))


# ---- driver LCP: summation order of the oracle, and the reference's own function bodies (oracle/ref_lcp.cpp) -----------
SUM_CHUNKS, SUM_SERIAL, SUM_DEVICE = 0, 1, 2


def set_sum_order(mode, grid=1, threads=256):
    """Order of the three Barzilai-Borwein sums inside oracle.resolve_collisions (see bb_sums in the oracle source):
    SUM_CHUNKS (default), SUM_SERIAL (= Kokkos Serial backend = oracle/_ref), SUM_DEVICE (= the rate pass's tree)."""
    _call(orc, "orc_set_sum_order", int(mode), int(grid), int(threads))


class sum_order:
    """with oracle.sum_order(oracle.SUM_SERIAL): ..."""

    def __init__(self, mode, grid=1, threads=256):
        self.args = (mode, grid, threads)

    def __enter__(self):
        set_sum_order(*self.args)

    def __exit__(self, *exc):
        set_sum_order(SUM_CHUNKS)


def diff_dots(x, xp, g, gp, dt):
    out = np.empty(3)
    _call(orc, "orc_diff_dots", len(x), _a(x), _a(xp), _a(g), _a(gp), float(dt), out)
    return out


def max_abs_projected_sep(gamma, phi0, gdot, dt):
    return _call(orc, "orc_max_abs_projected_sep", len(gamma), _a(gamma), _a(phi0), _a(gdot), float(dt),
                 restype=C.c_double)


# The reference's own kernels, scrap/lcp_spheres/KokkosLCP.cpp:449-876 (monodisperse spheres, dry mobility)
def ref_lcp_narrow(ids, radius, pos):
    ids = _a(ids, _i)
    sep, normal = np.empty(len(ids)), np.empty((len(ids), 3))
    _call(ref, "ref_lcp_signed_sep_and_normal", len(pos), len(ids), ids, float(radius), _a(pos), sep, normal)
    return sep, normal


def ref_lcp_D(n, ids, normal, gamma):
    F = np.empty((n, 3))
    _call(ref, "ref_lcp_sum_collision_force", n, len(ids), _a(ids, _i), _a(normal), _a(gamma), F)
    return F


def ref_lcp_M(radius, mu, F):
    U = np.zeros_like(_a(F))
    _call(ref, "ref_lcp_mobility", len(F), float(radius), float(mu), _a(F), U)
    return U


def ref_lcp_DT(n, ids, normal, U):
    g = np.empty(len(ids))
    _call(ref, "ref_lcp_rate_of_change_of_sep", n, len(ids), _a(ids, _i), _a(normal), _a(U), g)
    return g


def ref_lcp_residual(gamma, sep, sep_dot, dt):
    return _call(ref, "ref_lcp_max_abs_projected_sep", len(gamma), _a(gamma), _a(sep), _a(sep_dot), float(dt),
                 restype=C.c_double)


def ref_lcp_diff_dots(x, xp, g, gp, dt):
    out = np.empty(3)
    _call(ref, "ref_lcp_diff_dots", len(x), _a(x), _a(xp), _a(g), _a(gp), float(dt), out)
    return out


def ref_lcp_update_gammas(gamma_tmp, sep, sep_dot, dt, alpha):
    gamma = np.empty(len(gamma_tmp))
    _call(ref, "ref_lcp_update_con_gammas", len(gamma), gamma, _a(gamma_tmp), _a(sep), _a(sep_dot), float(dt),
          float(alpha))
    return gamma


def ref_lcp_resolve(ids, radius, mu, dt, tol, max_it, pos, sep, normal, gamma0=None):
    """resolve_collisions of KokkosLCP.cpp:675-876 (enable_hydro = false). A fixed iteration count K is (tol=0, max_it=K)."""
    ids = _a(ids, _i)
    n, m = len(pos), len(ids)
    gamma = np.zeros(m) if gamma0 is None else _a(gamma0).copy()
    F, U = np.zeros((n, 3)), np.zeros((n, 3))
    res = CollisionResult()
    fn = ref.ref_lcp_resolve_collisions
    fn.restype = None
    fn(C.c_long(n), C.c_long(m), _p(ids), C.c_double(radius), C.c_double(mu), C.c_double(dt), C.c_double(tol),
       C.c_int(max_it), _p(_a(pos)), _p(_a(sep)), _p(_a(normal)), _p(gamma), _p(F), _p(U), C.byref(res))
    return dict(res=res.max_abs_projected_sep, it=res.ite_count, max_disp=res.max_displacement, gamma=gamma, F=F, U=U)
