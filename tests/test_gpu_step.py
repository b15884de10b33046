"""GPU parity: math::convex solver, Euler update, and the whole collision step (device and host entry points) against an
oracle step composed in Python from the oracle's pieces (search -> filter -> narrow phase -> BBPGD -> Euler)."""
import os

import numpy as np
import pytest
import torch

import oracle
from mundy_b200 import workloads as wl

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def host(t):
    return t.detach().cpu().numpy()


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


# ---- math::convex ---------------------------------------------------------------------------------------------------
def test_convex_dense_reference_golden(cuda):
    """Planted-solution LCP solved by the reference's own convex.hpp (tests/golden/ref_outputs.npz)."""
    from mundy_b200 import ops
    REF = np.load(os.path.join(G, "ref_outputs.npz"))
    x = dev(np.full(50, 99.99))
    res = ops.solve_cqpp_dense(dev(REF["lcp_A"]), dev(REF["lcp_q"]), x, max_iters=1000, tol=1e-6, space_kind=1)
    meta = REF["lcp_meta_r0"]
    assert res.converged and res.num_iters == int(meta[0])
    assert np.allclose(host(x), REF["lcp_x_r0"], rtol=0, atol=1e-12)
    assert np.max(np.abs(host(x) - REF["lcp_xstar"])) <= 1e-5  # UnitTestConvex.cpp:529-606 acceptance


@pytest.mark.parametrize("space,lo,hi", [(0, 0, 0), (1, 0, 0), (2, 0, 0.1), (3, -0.01, 0.02)])
@pytest.mark.parametrize("rk", [0, 1])
def test_convex_dense_spaces(cuda, space, lo, hi, rk):
    from mundy_b200 import ops
    rng = np.random.default_rng(7)
    m = 200  # UnitTestConvex.cpp Kokkos-backend size
    A = rng.uniform(-1, 1, (m, m))
    A = A @ A.T + m * np.eye(m)
    q = rng.normal(size=m)
    want = oracle.solve_cqpp_dense(A, q, np.zeros(m), space, lo, hi, tol=1e-8, resid_kind=rk)
    x = dev(np.zeros(m))
    res = ops.solve_cqpp_dense(dev(A), dev(q), x, tol=1e-8, space_kind=space, lo=lo, hi=hi, residual_kind=rk)
    assert bool(res.converged) == want["converged"] and res.num_iters == want["iters"]
    assert np.allclose(host(x), want["x"], rtol=0, atol=1e-12)


def test_convex_contact_lcp(cuda):
    """The collision LCP posed through the math::convex interface: A = dt D^T M D, q = phi0, x >= 0."""
    from mundy_b200 import ops
    from test_gpu_collision import contact_system, gpu_op
    w = wl.polydisperse_spheres(3000, phi=0.2)
    pairs, phi, nrm, ai, aj = contact_system(w, 0.5)
    mu, dt = 1e-3, 1e-2
    want = oracle.solve_contact_lcp_convex(len(w["radius"]), pairs, nrm, w["radius"], mu, dt, phi, np.zeros(len(pairs)),
                                           max_iters=60, tol=1e-6, resid_kind=1)
    op = gpu_op(w, pairs, nrm, ai, aj, mu, dt)
    x = torch.zeros(len(pairs), dtype=torch.float64, device="cuda")
    res = ops.solve_lcp_contact(op, dev(phi), x, max_iters=60, tol=1e-6, residual_kind=1)
    assert res.num_iters == want["iters"] and bool(res.converged) == want["converged"]
    assert rel(host(x), want["x"]) <= 1e-9


# ---- Euler ----------------------------------------------------------------------------------------------------------
def test_euler_update(cuda):
    from mundy_b200 import ops
    rng = np.random.default_rng(3)
    n = 5000
    w = wl.mixed(n)
    U, W = rng.normal(size=(n, 3)), rng.normal(size=(n, 3))
    W[:10] = 0.0  # |omega| < 1e-15: quaternion untouched (Quaternion.hpp:1366)
    c, q = dev(w["center"]), dev(w["quat"])
    ops.euler_update(c, dev(U), 0.01, quat=q, kind=dev(w["kind"], torch.int32), W=dev(W))
    co, qo = oracle.euler_update(w["center"], U, 0.01, w["quat"], w["kind"], W)
    assert np.array_equal(host(c), co)          # x += dt U is bit-exact
    assert np.max(np.abs(host(q) - qo)) <= 1e-14  # sin/cos come from different libm implementations
    assert np.array_equal(host(q)[:10], w["quat"][:10])


def test_abc_flow_force(cuda):
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(5000)
    F0 = np.random.default_rng(0).normal(size=(5000, 3))
    F = ops.abc_flow_force(dev(w["center"]), dev(F0))
    Fo = F0.copy()
    Fo += oracle.abc_flow_force(w["center"])
    assert np.max(np.abs(host(F) - Fo)) <= 1e-14  # sin/cos: CUDA libdevice vs glibc, <= 2 ulp on O(1) values


# ---- whole step -----------------------------------------------------------------------------------------------------
oracle_step = oracle.collision_step


STEP_CASES = [("spheres", wl.polydisperse_spheres, dict(n=8000, phi=0.25), 0.5, False),
              ("rods_all_candidates", wl.rods, dict(n=3000, phi=0.1, length=4.0), 0.4, False),
              ("rods_shape_filter", wl.rods, dict(n=5000, phi=0.1), 0.4, True),
              ("mixed_shape_filter", wl.mixed, dict(n=4000, phi=0.2), 0.4, True)]


@pytest.mark.parametrize("name,maker,kw,buffer,filt", STEP_CASES, ids=[c[0] for c in STEP_CASES])
def test_collision_steps_match_oracle(cuda, name, maker, kw, buffer, filt):
    from mundy_b200 import ops
    w = maker(**kw)
    n = len(w["radius"])
    mu, dt, tol, fixed = 1e-3, 1e-4, 1e-5, 8
    rods = "kind" in w
    c, r = dev(w["center"]), dev(w["radius"])
    q = dev(w["quat"]) if rods else None
    b = ops.make_bodies(c, r, q, dev(w["length"]) if rods else None, dev(w["kind"], torch.int32) if rods else None)
    links = ops.NeighborLinks().set_search_buffer(buffer)
    if filt:
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    state, olinks = (w["center"].copy(), w["quat"].copy() if rods else None), None
    Fcol, Uout = torch.empty((n, 3), dtype=torch.float64, device="cuda"), torch.empty((n, 3), dtype=torch.float64, device="cuda")
    for step in range(3):
        F_ext = oracle.abc_flow_force(state[0])
        stats = links.step(b, mu, dt, tol, 10000, fixed_iters=fixed, F_ext=dev(F_ext), F_col=Fcol, U_out=Uout)
        state, olinks, out, U = oracle_step(state, w, olinks, buffer, filt, mu, dt, tol, 10000, fixed, F_ext)
        assert stats.rebuilt == int(out["rebuilt"]) and (step > 0 or stats.rebuilt == 1)
        assert stats.num_links == len(olinks[0]) and np.array_equal(host(links.pairs()), olinks[0])
        assert stats.lcp.ite_count == fixed and stats.kernel_launches > 0
        assert abs(stats.lcp.max_abs_projected_sep - out["res"]) <= 1e-9 * abs(out["res"])
        assert rel(host(Fcol), out["F"]) <= 1e-10, (step, rel(host(Fcol), out["F"]))
        assert rel(host(Uout), U) <= 1e-10
        assert rel(host(c), state[0]) <= 1e-12          # post-step positions
        if rods:
            assert np.max(np.abs(host(q) - state[1])) <= 1e-12
        # Continue the oracle from the GPU's post-step state: each step is then compared on identical inputs (the
        # last-bit differences of sin/cos in the quaternion update would otherwise be amplified by the next solve).
        state = (host(c).copy(), host(q).copy() if rods else None)


def test_multi_step_with_rebuilds(cuda):
    """Several steps with a displacement large enough to trigger neighbour rebuilds; same trigger steps as the oracle."""
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(5000, phi=0.2)
    mu, dt, tol, buffer = 1e-3, 3e-4, 1e-5, 0.5
    c = dev(w["center"])
    b = ops.make_bodies(c, dev(w["radius"]))
    links = ops.NeighborLinks().set_search_buffer(buffer)
    state, olinks, rebuilds = (w["center"].copy(), None), None, []
    for step in range(6):
        F_ext = oracle.abc_flow_force(state[0])
        stats = links.step(b, mu, dt, tol, 10000, fixed_iters=5, F_ext=dev(F_ext))
        state, olinks, out, U = oracle_step(state, w, olinks, buffer, False, mu, dt, tol, 10000, 5, F_ext)
        assert stats.rebuilt == int(out["rebuilt"]) and stats.num_links == len(olinks[0])
        assert rel(host(c), state[0]) <= 1e-12
        rebuilds.append(stats.rebuilt)
    assert rebuilds[0] == 1 and sum(rebuilds) >= 2, rebuilds


def test_collision_step_converges_and_host_entry_point(cuda):
    """Run to tolerance through the HOST entry point (H2D + step + D2H) and through the device entry point."""
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(6000, phi=0.2)
    n = len(w["radius"])
    mu, dt, tol = 1e-3, 1e-2, 1e-5
    c_h = torch.as_tensor(w["center"].copy()).pin_memory()
    r_h = torch.as_tensor(w["radius"].copy()).pin_memory()
    U_h = torch.empty((n, 3), dtype=torch.float64).pin_memory()
    links = ops.NeighborLinks().set_search_buffer(0.5)
    stats = links.step_host(c_h, r_h, None, None, None, mu, dt, tol, 10000, U_out=U_h)
    state, olinks, out, U = oracle_step((w["center"].copy(), None), w, None, 0.5, False, mu, dt, tol, 10000, -1, None)
    assert stats.lcp.converged and stats.lcp.ite_count == out["it"] and stats.num_links == len(olinks[0])
    assert rel(U_h.numpy(), U) <= 1e-8 and rel(c_h.numpy(), state[0]) <= 1e-10
    # post-step overlaps are resolved to the tolerance (the drivers' check_overlap, KokkosLCP.cpp:878-925)
    phi_after = oracle.narrow_phase(olinks[0], c_h.numpy(), w["radius"])[0]
    assert phi_after.min() > -10 * tol
    # device entry point gives the same answer
    c = dev(w["center"])
    b = ops.make_bodies(c, dev(w["radius"]))
    links2 = ops.NeighborLinks().set_search_buffer(0.5)
    stats2 = links2.step(b, mu, dt, tol, 10000)
    assert stats2.lcp.ite_count == stats.lcp.ite_count and np.array_equal(host(c), c_h.numpy())


def test_step_reports_non_convergence_like_the_reference_throw(cuda):
    from mundy_b200 import ops, _capi
    w = wl.monodisperse_spheres(3000, phi=0.4)
    b = ops.make_bodies(dev(w["center"]), dev(w["radius"]))
    links = ops.NeighborLinks().set_search_buffer(0.5)
    with pytest.raises(_capi.Mb200Error) as e:
        links.step(b, 1e-3, 5e-3, 1e-12, 3, throw_on_max_iters=True)
    assert e.value.status == -4 and "did not converge" in str(e.value)
