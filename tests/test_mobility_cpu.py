"""CPU tests (no GPU) of the non-local mobility row: the oracle's restatement of the reference's pair
kernels (scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp :686-1085, :1316-1354) is pinned
bit-for-bit against the reference's own header (stored outputs in tests/golden/mobility_ref.npz and, when oracle/_ref is
present, live); the Wilson-2013 three-sphere table of the reference's unit test (UnitTestPeriphery.cpp:671-775) is
reproduced; the KokkosLCP.cpp RPYKernel (:219-280) and the operator-form BBPGD (NgpHP1.cpp:1487-1645) are checked
through their properties."""
import os

import numpy as np
import pytest

import oracle

G = os.path.join(os.path.dirname(__file__), "golden")
MOB = np.load(os.path.join(G, "mobility_ref.npz"))
KINDS = [(oracle.KERNEL_STOKES, "stokes"), (oracle.KERNEL_RPY, "rpy"), (oracle.KERNEL_RPYC, "rpyc")]


@pytest.mark.parametrize("kind,name", KINDS)
def test_oracle_matches_reference_outputs(kind, name):
    got = oracle.mobility_kernel(kind, float(MOB["mu"]), MOB["tpos"], MOB["spos"], MOB["srad"], MOB["trad"],
                                 MOB["sforce"], MOB["vel0"])
    assert np.array_equal(got, MOB[name])


def test_local_drag_matches_reference_outputs():
    got = oracle.local_drag(oracle.KERNEL_RPY, float(MOB["mu"]), MOB["srad"][:80], MOB["sforce"][:80],
                            0.5 * MOB["sforce"][:80])
    assert np.array_equal(got, MOB["local_drag"][:80])  # rows 90, 91 have radius 0 (inf in both)


@pytest.mark.skipif(oracle.ref is None, reason="oracle/_ref/libmundy_ref.so not present")
@pytest.mark.parametrize("kind,name", KINDS)
def test_oracle_vs_live_reference_random(kind, name):
    rng = np.random.default_rng(77 + kind)
    nt, ns = 333, 1000
    tp, sp = rng.uniform(0, 25, (nt, 3)), rng.uniform(0, 25, (ns, 3))
    sp[:100] = tp[:100]
    sr, tr = rng.uniform(0.1, 2.0, ns), rng.uniform(0.1, 2.0, nt)
    f, v0 = rng.normal(size=(ns, 3)), rng.normal(size=(nt, 3))
    a = oracle.mobility_kernel(kind, 1.3, tp, sp, sr, tr, f, v0)
    b = oracle.mobility_kernel(kind, 1.3, tp, sp, sr, tr, f, v0, lib=oracle.ref)
    assert np.array_equal(a, b)


def test_three_spheres_wilson_table():
    """UnitTestPeriphery.cpp:671-775 (the reference prints the errors, it asserts nothing): stored reference outputs are
    reproduced exactly, and for well separated spheres RPY approaches the pseudo-analytical values."""
    s_list, radius = MOB["tri_s"], float(MOB["tri_radius"])
    U2 = [0.63461, 0.59718, 0.49545, 0.41694, 0.31859, 0.21586]
    U3 = [0.00498, 0.03517, 0.07393, 0.07824, 0.06925, 0.05078]
    for k, s in enumerate(s_list):
        pos = np.array([[0.0, 0.0, 0.0], [-s * radius / 2.0, -np.sqrt(3.0) * s * radius / 2.0, 0.0],
                        [s * radius / 2.0, -np.sqrt(3.0) * s * radius / 2.0, 0.0]])
        f = np.zeros((3, 3))
        f[0, 1] = -6.0 * np.pi * radius
        rad = np.full(3, radius)
        self_term = f / (6.0 * np.pi * radius * 1.0)
        for kind, key in ((oracle.KERNEL_RPY, "tri_rpy"), (oracle.KERNEL_RPYC, "tri_rpyc")):
            v = oracle.mobility_kernel(kind, 1.0, pos, pos, rad, rad, f) + self_term
            assert np.array_equal(v, MOB[key][k])
            assert v[0, 1] == -1.0  # U1: RPY has no reflection back onto the forced sphere
            if s >= 4.0:
                assert abs(-v[1, 1] - U2[k]) <= 5e-3 * U2[k]  # quake_inv_sqrt alone costs ~2e-3
                assert abs(v[2, 0] - U3[k]) <= 3e-2 * U3[k]


def _rpy_tensor_apply(mu, pos, a, f):
    """Textbook RPY for equal radii, non-overlapping: (1/(8 pi mu r)) [(1 + 2a^2/3r^2) I + (1 - 2a^2/r^2) rr] f."""
    out = np.zeros_like(f)
    for t in range(len(pos)):
        d = pos[t] - pos
        r = np.linalg.norm(d, axis=1)
        r[t] = np.inf
        rh = d / r[:, None]
        c1 = (1 + 2 * a * a / (3 * r * r)) / (8 * np.pi * mu * r)
        c2 = (1 - 2 * a * a / (r * r)) / (8 * np.pi * mu * r)
        out[t] = (c1[:, None] * f + c2[:, None] * rh * np.einsum("ij,ij->i", rh, f)[:, None]).sum(axis=0)
    return out


def test_exact_kernel_is_the_rpy_tensor_and_quake_is_close():
    rng = np.random.default_rng(3)
    n, a, mu = 200, 0.5, 0.9
    pos = rng.uniform(0, 20, (n, 3))
    rad = np.full(n, a)
    f = rng.normal(size=(n, 3))
    exact = oracle.mobility_kernel(oracle.KERNEL_RPY_EXACT, mu, pos, pos, rad, rad, f)
    want = _rpy_tensor_apply(mu, pos, a, f)
    assert np.max(np.abs(exact - want)) <= 1e-12 * np.max(np.abs(want))
    quake = oracle.mobility_kernel(oracle.KERNEL_RPY, mu, pos, pos, rad, rad, f)
    err = np.max(np.abs(quake - exact)) / np.max(np.abs(exact))
    assert 1e-6 < err < 1e-2  # one Newton iteration: ~1.7e-3 relative error in 1/r, by design of the reference
    corrected = oracle.mobility_kernel(oracle.KERNEL_RPYC, mu, pos, pos, rad, rad, f)
    d = np.linalg.norm(pos[:, None] - pos[None], axis=2) + np.eye(n) * 100
    if d.min() > 2 * a:  # no overlaps: RPYC is the plain tensor
        assert np.max(np.abs(corrected - want)) <= 1e-12 * np.max(np.abs(want))


def test_neighbor_form_equals_all_pairs_on_a_full_list():
    """apply_rpy_kernel_neighbors (KokkosLCP.cpp:305-333) over a CSR that lists every sphere in ascending order is the
    all-pairs kernel, bit for bit (same order of summation); a truncated list is the sum over that list."""
    rng = np.random.default_rng(5)
    n, mu = 150, 1.1
    pos, rad, f = rng.uniform(0, 9, (n, 3)), rng.uniform(0.2, 0.6, n), rng.normal(size=(n, 3))
    offsets = np.arange(n + 1, dtype=np.int32) * n
    indices = np.tile(np.arange(n, dtype=np.int32), n)
    v0 = rng.normal(size=(n, 3))
    for kind in (oracle.KERNEL_STOKES, oracle.KERNEL_RPY, oracle.KERNEL_RPYC, oracle.KERNEL_RPY_EXACT):
        a = oracle.mobility_kernel(kind, mu, pos, pos, rad, rad, f, v0)
        b = oracle.mobility_kernel_neighbors(kind, mu, pos, rad, offsets, indices, f, v0)
        assert np.array_equal(a, b)
    # one-directional list of the driver (row j lists i < j, KokkosLCP.cpp:404-412): only lower sources contribute
    lower_off = np.concatenate([[0], np.cumsum(np.arange(n))]).astype(np.int32)
    lower_idx = np.concatenate([np.arange(j) for j in range(n)]).astype(np.int32)
    got = oracle.mobility_kernel_neighbors(oracle.KERNEL_RPY_EXACT, mu, pos, rad, lower_off, lower_idx, f)
    fz = f.copy()
    for t in (0, 1, 17, n - 1):
        fm = np.where(np.arange(n)[:, None] < t, fz, 0.0)
        want = oracle.mobility_kernel(oracle.KERNEL_RPY_EXACT, mu, pos[t:t + 1], pos, rad, rad[t:t + 1], fm)
        assert np.allclose(got[t], want[0], rtol=1e-13, atol=1e-15)


def test_mobility_matrix_symmetric_positive_definite():
    """M = local drag + RPY (exact kernel, equal radii, no overlaps) is symmetric positive definite."""
    rng = np.random.default_rng(9)
    n, a, mu = 40, 0.5, 1.0
    g = np.stack(np.meshgrid(*[np.arange(4)] * 3, indexing="ij"), -1).reshape(-1, 3)[:n] * 1.3
    pos = g + rng.uniform(-0.1, 0.1, g.shape)
    rad = np.full(n, a)
    M = np.zeros((3 * n, 3 * n))
    for k in range(3 * n):
        e = np.zeros(3 * n)
        e[k] = 1.0
        M[:, k] = oracle.mobility_apply(oracle.MOBILITY_ALL, oracle.KERNEL_RPY_EXACT, mu, pos, rad,
                                        e.reshape(n, 3)).ravel()
    assert np.max(np.abs(M - M.T)) <= 1e-14 * np.abs(M).max()
    assert np.linalg.eigvalsh(0.5 * (M + M.T)).min() > 0


def _packing(n, phi, seed):
    from mundy_b200 import workloads as wl
    w = wl.polydisperse_spheres(n, phi=phi, seed=seed) if "seed" in wl.polydisperse_spheres.__code__.co_varnames \
        else wl.polydisperse_spheres(n, phi=phi)
    S = oracle.search_radii(w["radius"], 0.3)
    pairs = oracle.search(w["center"], S)
    phi0, nrm = oracle.narrow_phase(pairs, w["center"], w["radius"])[:2]
    return w, pairs, phi0, nrm


def test_operator_form_solve_with_local_drag_is_the_driver_solve():
    """NgpHP1.cpp:1487-1645 with SphereLocalDragMobilityOp == the driver BBPGD (KokkosLCP.cpp:675-876): same iterates up
    to the rounding of the drag coefficient ((1/(6 pi mu))/r versus (1/(6 pi mu)) * (1/r))."""
    w, pairs, phi0, nrm = _packing(1500, 0.3, 1)
    n, mu, dt = len(w["radius"]), 1e-3, 1e-4
    a = oracle.resolve_collisions(n, pairs, nrm, w["radius"], mu, dt, 1e-5, 10000, phi0, fixed_iters=25)
    b = oracle.resolve_collisions_mobility(pairs, nrm, oracle.MOBILITY_LOCAL_DRAG, oracle.KERNEL_RPY, mu, w["center"],
                                           w["radius"], dt, 1e-5, 10000, phi0, fixed_iters=25)
    assert a["it"] == b["it"] == 25
    assert np.allclose(a["gamma"], b["gamma"], rtol=1e-9, atol=1e-12 * np.abs(a["gamma"]).max())
    assert np.allclose(a["U"], b["U"], rtol=1e-9, atol=1e-12 * np.abs(a["U"]).max())


def test_operator_form_solve_converges_with_hydrodynamics():
    """With a symmetric positive definite M the LCP stays convex: BBPGD converges, gamma >= 0, no overlap beyond tol
    after the (linearised) step, complementarity on the separated contacts; the external force enters through
    phi0 += dt D^T M F_ext. (A kernel truncated to a neighbour list is in general NOT positive definite — the driver
    uses it only as an intermediate level, KokkosLCP.cpp:788-815 — so convergence is asserted for the full sum only.)"""
    mode = oracle.MOBILITY_ALL
    w, pairs, phi0, nrm = _packing(600, 0.25, 2)
    n, mu, dt, tol = len(w["radius"]), 1e-3, 1e-4, 1e-6
    rad = np.full(n, w["radius"].mean())
    # symmetric neighbour list from the pair list
    i, j = pairs[:, 0], pairs[:, 1]
    src = np.concatenate([i, j])
    dst = np.concatenate([j, i])
    order = np.lexsort((dst, src))
    offsets = np.concatenate([[0], np.cumsum(np.bincount(src, minlength=n))]).astype(np.int32)
    indices = dst[order].astype(np.int32)
    F_ext = np.random.default_rng(4).normal(size=(n, 3)) * 1e-3
    out = oracle.resolve_collisions_mobility(pairs, nrm, mode, oracle.KERNEL_RPYC, mu, w["center"], rad, dt, tol, 20000,
                                             phi0, offsets=offsets, indices=indices, F_ext=F_ext)
    assert out["res"] < tol and 0 < out["it"] < 20000
    assert out["gamma"].min() >= 0.0
    Ue = oracle.mobility_apply(mode, oracle.KERNEL_RPYC, mu, w["center"], rad, F_ext, offsets, indices)
    assert np.array_equal(out["U_ext"], Ue)
    want_phi = phi0 + dt * oracle.apply_DT(n, pairs, nrm, Ue)
    assert np.allclose(out["phi0"], want_phi, rtol=0, atol=1e-15)
    sep = out["phi0"] + dt * oracle.apply_DT(n, pairs, nrm, out["U"])
    assert sep.min() >= -tol
    assert np.all(np.abs(sep[out["gamma"] >= 1e-12]) < tol)
    assert np.array_equal(out["U"], oracle.mobility_apply(mode, oracle.KERNEL_RPYC, mu, w["center"], rad, out["F"],
                                                          offsets, indices))
