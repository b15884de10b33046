"""Parity at BASELINE.json's full sizes (10^6 bodies) through size-independent properties, plus oracle checks on the
parts the oracle finishes in seconds at that size (narrow phase, operator, stopping test: all O(C))."""
import numpy as np
import pytest
import torch

import oracle
from mundy_b200 import workloads as wl

pytestmark = pytest.mark.gpu
N = 1_000_000


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def host(t):
    return t.detach().cpu().numpy()


@pytest.fixture(scope="module")
def rods_step(cuda):
    """cfg3: 10^6 spherocylinders, aspect ratio 10; two relaxation steps to tolerance, then one measured step."""
    from mundy_b200 import ops
    w = wl.rods(N, phi=0.10)
    mu, dt, tol, buffer = 1e-3, 1e-4, 1e-3, 0.5
    c, q = dev(w["center"]), dev(w["quat"])
    b = ops.make_bodies(c, dev(w["radius"]), q, dev(w["length"]), dev(w["kind"], torch.int32))
    links = ops.NeighborLinks().set_search_buffer(buffer).set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    F = torch.zeros((N, 3), dtype=torch.float64, device="cuda")
    for _ in range(2):
        F.zero_()
        ops.abc_flow_force(c, F)
        links.step(b, mu, dt, tol, 10000, F_ext=F)
    F.zero_()
    ops.abc_flow_force(c, F)
    c0, q0 = host(c).copy(), host(q).copy()
    Fcol, U, Wv = torch.empty_like(F), torch.empty_like(F), torch.empty_like(F)
    stats = links.step(b, mu, dt, tol, 10000, F_ext=F, F_col=Fcol, U_out=U, W_out=Wv, force_rebuild=True)
    return dict(w=w, c0=c0, q0=q0, c1=host(c), q1=host(q), pairs=host(links.pairs()), stats=stats, F=host(F),
                Fcol=host(Fcol), U=host(U), W=host(Wv), p=(mu, dt, tol, buffer))


def test_cfg3_links_are_a_sorted_exact_candidate_set(rods_step):
    s, w = rods_step, rods_step["w"]
    pairs, (mu, dt, tol, buffer) = s["pairs"], s["p"]
    assert s["stats"].rebuilt == 1 and len(pairs) == s["stats"].num_links > 2 * N
    i, j = pairs[:, 0].astype(np.int64), pairs[:, 1].astype(np.int64)
    assert np.all(i < j) and i.min() >= 0 and j.max() < N
    key = i * N + j
    assert np.all(np.diff(key) > 0)                                   # sorted by (i, j), no duplicates
    # every link passes the shape filter evaluated by the oracle on the same bodies (phi <= 2 * buffer) ...
    phi, nrm, ai, aj = oracle.narrow_phase(pairs, s["c0"], w["radius"], s["q0"], w["length"], w["kind"], arms=True)
    assert phi.max() <= 2 * buffer
    # ... and the set is complete on a sample: the oracle's own search + filter around 20000 bodies finds the same links
    rng = np.random.default_rng(0)
    lo = rng.uniform(0, w["edge"] - 60.0, 3)
    inside = np.all((s["c0"] >= lo) & (s["c0"] <= lo + 60.0), axis=1)
    idx = np.nonzero(inside)[0]
    sub = oracle.search(s["c0"][idx], oracle.search_radii(w["radius"][idx], buffer, w["kind"][idx], w["length"][idx]))
    sub = oracle.filter_shape_distance(sub, s["c0"][idx], s["q0"][idx], w["radius"][idx], w["length"][idx],
                                       w["kind"][idx], 2 * buffer)
    want = np.stack([idx[sub[:, 0]], idx[sub[:, 1]]], axis=1)
    both = inside[i] & inside[j]
    assert np.array_equal(pairs[both], want)
    rods_step["narrow"] = (phi, nrm, ai, aj)


def test_cfg3_converged_step_satisfies_the_reference_stopping_test(rods_step):
    """The oracle re-evaluates the GPU's collision velocities: D^T U must leave every contact within tol (the
    reference's own residual), forces must balance (sum F_col = 0: D has one +n and one -n per contact), and the Euler
    update must be the reference's."""
    s, w = rods_step, rods_step["w"]
    mu, dt, tol, buffer = s["p"]
    st = s["stats"]
    assert st.lcp.converged and 0 < st.lcp.ite_count < 10000 and st.lcp.max_abs_projected_sep < tol
    phi, nrm, ai, aj = rods_step.get("narrow") or oracle.narrow_phase(s["pairs"], s["c0"], w["radius"], s["q0"],
                                                                      w["length"], w["kind"], arms=True)
    assert np.max(np.abs(s["Fcol"].sum(axis=0))) <= 1e-9 * np.abs(s["Fcol"]).sum()
    # U = U_ext + U_col ; U_ext = M F
    Uext, Wext = oracle.apply_M(w["radius"], mu, s["F"], np.zeros_like(s["F"]), s["q0"], w["kind"])
    # translational collision velocity is M F_col (isotropic in the prototype's mobility)
    Ucol, _ = oracle.apply_M(w["radius"], mu, s["Fcol"], np.zeros_like(s["F"]), s["q0"], w["kind"])
    assert np.max(np.abs(s["U"] - (Uext + Ucol))) <= 1e-10 * np.max(np.abs(s["U"]))
    # the reference's stopping test (KokkosLCP.cpp:500-521), evaluated by the oracle on the GPU's velocities:
    # phi0 + dt * D^T (U, W) >= -tol on every link (the constraint is the linearised one: normals and arms of the
    # start of the step; rods turn by a finite angle in a step, so the post-step geometric overlap is not bounded by tol)
    sep = phi + dt * oracle.apply_DT(N, s["pairs"], nrm, s["U"], s["W"], ai, aj)
    assert sep.min() > -tol, sep.min()
    assert abs(max(-sep.min(), 0.0) - 0.0) <= tol and st.lcp.max_abs_projected_sep >= max(-sep.min(), 0.0) - 1e-12
    assert np.max(np.abs(s["c1"] - (s["c0"] + dt * s["U"]))) <= 1e-12 * np.max(np.abs(s["c1"]))
    assert np.max(np.abs(np.linalg.norm(s["q1"], axis=1) - 1.0)) <= 1e-14


def test_cfg2_million_polydisperse_spheres(cuda):
    """cfg2 (free-space box): links exact on a sample, solver converges, forces balance, overlaps removed."""
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(N, phi=0.35)
    mu, dt, tol, buffer = 1e-3, 1e-4, 1e-3, 0.25
    c = dev(w["center"])
    b = ops.make_bodies(c, dev(w["radius"]))
    links = ops.NeighborLinks().set_search_buffer(buffer)
    Fcol = torch.empty((N, 3), dtype=torch.float64, device="cuda")
    c0 = w["center"].copy()
    st = links.step(b, mu, dt, tol, 10000, F_col=Fcol)
    pairs = host(links.pairs())
    i, j = pairs[:, 0].astype(np.int64), pairs[:, 1].astype(np.int64)
    assert st.lcp.converged and st.num_links == len(pairs) and np.all(i < j) and np.all(np.diff(i * N + j) > 0)
    d2 = np.sum((c0[j] - c0[i]) ** 2, axis=1)
    S = w["radius"] + buffer
    assert np.all(d2 <= (S[i] + S[j]) ** 2)
    rng = np.random.default_rng(1)
    lo = rng.uniform(0, w["edge"] - 30.0, 3)
    inside = np.all((c0 >= lo) & (c0 <= lo + 30.0), axis=1)
    idx = np.nonzero(inside)[0]
    sub = oracle.search(c0[idx], S[idx])
    assert np.array_equal(pairs[inside[i] & inside[j]], np.stack([idx[sub[:, 0]], idx[sub[:, 1]]], axis=1))
    F = host(Fcol)
    assert np.max(np.abs(F.sum(axis=0))) <= 1e-9 * np.abs(F).sum()
    phi1 = oracle.narrow_phase(pairs, host(c), w["radius"])[0]
    assert phi1.min() > -20 * tol


def test_cfg2_million_polydisperse_spheres_periodic(cuda):
    """cfg2 as BASELINE.json words it ("10^6 polydisperse spheres in a periodic box"): the whole link set equals the
    oracle's periodic cell-list search (OpenMP, seconds at this size), links across the boundary exist, the solver
    converges, forces balance, and after the step no nearest-image pair overlaps by more than the tolerance allows."""
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(N, phi=0.35)
    cell = [w["edge"]] * 3
    mu, dt, tol, buffer = 1e-3, 1e-4, 1e-3, 0.25
    c = dev(w["center"])
    b = ops.make_bodies(c, dev(w["radius"]))
    links = ops.NeighborLinks().set_search_buffer(buffer).set_periodic(cell)
    Fcol = torch.empty((N, 3), dtype=torch.float64, device="cuda")
    c0 = w["center"].copy()
    st = links.step(b, mu, dt, tol, 10000, F_col=Fcol)
    pairs = host(links.pairs())
    S = w["radius"] + buffer
    want = oracle.search(c0, S, cell=cell)
    assert st.lcp.converged and st.num_links == len(pairs)
    assert np.array_equal(pairs, want)
    free = oracle.search(c0, S)
    assert len(pairs) > len(free)  # the periodic images add links
    phi0, nrm, _, _ = oracle.narrow_phase(pairs, c0, w["radius"], cell=cell)
    across = np.abs(c0[pairs[:, 1]] - c0[pairs[:, 0]]).max(axis=1) > 0.5 * w["edge"]
    assert across.sum() == len(pairs) - len(free) and phi0[across].max() <= 2 * buffer
    F = host(Fcol)
    assert np.max(np.abs(F.sum(axis=0))) <= 1e-9 * np.abs(F).sum()
    phi1 = oracle.narrow_phase(pairs, host(c), w["radius"], cell=cell)[0]
    assert phi1.min() > -20 * tol


def test_cfg5_jammed_mixed_packing(cuda):
    """cfg5 at one GPU's share (5*10^5 bodies, half spheres half rods, volume fraction 0.55: dense, many contacts per
    body, many BBPGD iterations): links exact on a sample, the contact forces are pairwise-balanced, and the state the
    solver stops in passes the reference's own stopping test when the oracle evaluates it on the GPU's velocities."""
    from mundy_b200 import ops
    n = 500_000
    w = wl.mixed(n, phi=0.55)
    mu, dt, tol, buffer = 1e-3, 1e-4, 1e-3, 0.25
    c, q = dev(w["center"]), dev(w["quat"])
    b = ops.make_bodies(c, dev(w["radius"]), q, dev(w["length"]), dev(w["kind"], torch.int32))
    links = ops.NeighborLinks().set_search_buffer(buffer).set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    Fcol, U, Wv = (torch.empty((n, 3), dtype=torch.float64, device="cuda") for _ in range(3))
    c0, q0 = w["center"].copy(), w["quat"].copy()
    st = links.step(b, mu, dt, tol, 10000, F_col=Fcol, U_out=U, W_out=Wv)
    pairs = host(links.pairs())
    assert st.num_links == len(pairs) > 4 * n          # jammed: more than 8 contacts per body
    assert st.lcp.converged and st.lcp.ite_count > 100  # a hard LCP: hundreds of iterations
    rng = np.random.default_rng(2)
    lo = rng.uniform(0, w["edge"] - 25.0, 3)
    inside = np.all((c0 >= lo) & (c0 <= lo + 25.0), axis=1)
    idx = np.nonzero(inside)[0]
    sub = oracle.search(c0[idx], oracle.search_radii(w["radius"][idx], buffer, w["kind"][idx], w["length"][idx]))
    sub = oracle.filter_shape_distance(sub, c0[idx], q0[idx], w["radius"][idx], w["length"][idx], w["kind"][idx],
                                       2 * buffer)
    i, j = pairs[:, 0], pairs[:, 1]
    assert np.array_equal(pairs[inside[i] & inside[j]], np.stack([idx[sub[:, 0]], idx[sub[:, 1]]], axis=1))
    F = host(Fcol)
    assert np.max(np.abs(F.sum(axis=0))) <= 1e-9 * np.abs(F).sum()
    phi, nrm, ai, aj = oracle.narrow_phase(pairs, c0, w["radius"], q0, w["length"], w["kind"], arms=True)
    sep = phi + dt * oracle.apply_DT(n, pairs, nrm, host(U), host(Wv), ai, aj)
    assert sep.min() > -tol, sep.min()
