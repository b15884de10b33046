"""CPU tests (no GPU): the oracle's driver LCP (D, M, D^T, residual, BB sums, gamma update and the whole BBPGD loop) is
pinned to the reference's OWN function bodies, scrap/lcp_spheres/KokkosLCP.cpp:449-876, which oracle/Makefile cuts from
the reference tree at build time and compiles unmodified into oracle/_ref/libmundy_ref.so (oracle/ref_lcp.cpp).

Two forms: against the outputs of that code stored in tests/golden/lcp_ref.npz (made by tests/golden/make_golden_lcp.py;
runs anywhere) and, when the library is present (this container, and the GPU box through the prebuilt .so), live.
Everything is compared BIT FOR BIT: the reference on one thread adds each body's forces in ascending contact order, as
the oracle does, and oracle.SUM_SERIAL selects its ascending loop for the three Barzilai-Borwein sums.
Problem: the driver's own parameters (KokkosLCP.cpp:963-969): unit spheres, mu = 1e-3, dt = 5e-3, buffer 3 r, tol 1e-5.
"""
import hashlib
import os

import numpy as np
import pytest

import oracle

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "lcp_ref.npz"))
MU, DT, TOL, RADIUS, BUFFER = 1e-3, 5e-3, 1e-5, 1.0, 3.0
KS = (0, 1, 2, 5, 12, 150)
needs_ref = pytest.mark.skipif(oracle.ref is None or not hasattr(oracle.ref, "ref_lcp_resolve_collisions"),
                               reason="oracle/_ref/libmundy_ref.so not built (needs /root/reference)")


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def problem(n, seed=1234, phi=0.30):  # as in make_golden_lcp.py
    rng = np.random.default_rng(seed)
    edge = (n * 4.0 / 3.0 * np.pi * RADIUS ** 3 / phi) ** (1.0 / 3.0)
    pos = rng.uniform(0.0, edge, size=(n, 3))
    return pos, oracle.search(pos, np.full(n, RADIUS + BUFFER))


def mono_narrow(pos, ids):
    """KokkosLCP.cpp:466-489 in numpy, operation for operation: sep = |c_j - c_i| - 2.0 r (the monodisperse driver's form).
    The oracle and the device follow the polydisperse drivers' `d - r_i - r_j` (scrap/lcp_spheres/NGPSpheresLCP.cpp:
    384-406), which rounds twice: for centre distances below r / 2 the two can differ by one ulp, so the pin solves are
    fed the monodisperse form and the oracle's narrow phase is checked against it to one ulp (normals bit for bit)."""
    d3 = pos[ids[:, 1]] - pos[ids[:, 0]]
    d = np.sqrt(d3[:, 0] * d3[:, 0] + d3[:, 1] * d3[:, 1] + d3[:, 2] * d3[:, 2])
    return d - 2.0 * RADIUS, d3 * (1.0 / d)[:, None]


def oracle_solve(pos, ids, sep, normal, tol, max_it, fixed):
    with oracle.sum_order(oracle.SUM_SERIAL):
        return oracle.resolve_collisions(len(pos), ids, normal, np.full(len(pos), RADIUS), MU, DT, tol, max_it, sep,
                                         fixed_iters=fixed)


def test_operators_match_stored_reference_outputs():
    pos, ids, n = G["s_pos"], G["s_ids"], len(G["s_pos"])
    r = np.full(n, RADIUS)
    sep, normal, _, _ = oracle.narrow_phase(ids, pos, r)
    assert np.array_equal(sep, G["s_sep"]) and np.array_equal(normal, G["s_normal"])  # :449-491
    gam = G["s_gamma"]
    F, _ = oracle.apply_D(n, ids, normal, gam)
    assert np.array_equal(F, G["s_F"])  # :559-591
    U, _ = oracle.apply_M(r, MU, F)
    assert np.array_equal(U, G["s_U"])  # :593-614
    gd = oracle.apply_DT(n, ids, normal, U)
    assert np.array_equal(gd, G["s_gdot"])  # :616-647
    assert oracle.max_abs_projected_sep(gam, sep, gd, DT) == float(G["s_res"])  # :493-522
    with oracle.sum_order(oracle.SUM_SERIAL):
        assert np.array_equal(oracle.diff_dots(gam, 0.5 * gam, gd, 0.25 * gd, DT), G["s_dots"])  # :524-557
    # the other two orders agree to rounding
    for mode, grid in ((oracle.SUM_CHUNKS, 1), (oracle.SUM_DEVICE, 7)):
        with oracle.sum_order(mode, grid, 64):
            assert np.allclose(oracle.diff_dots(gam, 0.5 * gam, gd, 0.25 * gd, DT), G["s_dots"], rtol=1e-13, atol=0)
    # :649-665 gamma update = one fixed iteration's first statement; checked through K = 1 below and directly here
    assert np.array_equal(np.maximum(gam - 0.37 * (sep + DT * gd), 0.0), G["s_gamma_next"])


@pytest.mark.parametrize("K", KS)
def test_fixed_iterations_match_stored_reference_outputs(K):
    pos, ids = G["s_pos"], G["s_ids"]
    out = oracle_solve(pos, ids, G["s_sep"], G["s_normal"], 0.0, 10000, K)
    assert np.array_equal(out["gamma"], G[f"s_K{K}_gamma"])
    assert np.array_equal(out["F"], G[f"s_K{K}_F"]) and np.array_equal(out["U"], G[f"s_K{K}_U"])
    res, it, disp = G[f"s_K{K}_meta"]
    assert out["res"] == res and out["it"] == it and out["max_disp"] == disp


def test_converged_run_matches_stored_reference_outputs():
    out = oracle_solve(G["s_pos"], G["s_ids"], G["s_sep"], G["s_normal"], TOL, 10000, -1)
    res, it, disp = G["s_conv_meta"]
    assert out["it"] == it and out["res"] == res and out["max_disp"] == disp  # same stopping iteration
    assert np.array_equal(out["gamma"], G["s_conv_gamma"])


@pytest.mark.parametrize("K", KS)
def test_cfg1_fixed_iterations_match_stored_reference_digests(K):
    """10^4 monodisperse spheres (BASELINE.json configs[0]); the stored digests are of the reference's outputs."""
    pos, ids = problem(10000)
    n = len(pos)
    sep, normal = mono_narrow(pos, ids)
    assert len(ids) == int(G["c_count"][0])
    assert [digest(pos), digest(ids), digest(sep), digest(normal)] == list(G["c_digest_inputs"])
    osep, onormal, _, _ = oracle.narrow_phase(ids, pos, np.full(n, RADIUS))
    assert np.array_equal(onormal, normal) and np.max(np.abs(osep - sep)) <= 2.0 ** -52
    out = oracle_solve(pos, ids, sep, normal, 0.0, 10000, K)
    assert np.array_equal(out["gamma"][::97], G[f"c_K{K}_gamma_sample"])
    assert [digest(out["gamma"]), digest(out["F"] + 0.0), digest(out["U"] + 0.0)] == list(G[f"c_K{K}_digest"])
    res, it, disp = G[f"c_K{K}_meta"]
    assert out["res"] == res and out["it"] == it and out["max_disp"] == disp


def test_cfg1_converged_run_matches_stored_reference():
    pos, ids = problem(10000)
    sep, normal = mono_narrow(pos, ids)
    out = oracle_solve(pos, ids, sep, normal, TOL, 10000, -1)
    res, it, disp = G["c_conv_meta"]
    assert (out["it"], out["res"], out["max_disp"]) == (it, res, disp)
    assert np.array_equal(out["gamma"][::97], G["c_conv_gamma_sample"])


@needs_ref
@pytest.mark.parametrize("seed,n", [(11, 300), (12, 2000)])
def test_live_reference_operators_and_loop(seed, n):
    pos, ids = problem(n, seed=seed, phi=0.35)
    r = np.full(n, RADIUS)
    sep, normal = oracle.ref_lcp_narrow(ids, RADIUS, pos)
    msep, mnormal = mono_narrow(pos, ids)
    assert np.array_equal(sep, msep) and np.array_equal(normal, mnormal)
    osep, onormal, _, _ = oracle.narrow_phase(ids, pos, r)
    assert np.max(np.abs(sep - osep)) <= 2.0 ** -52 and np.array_equal(normal, onormal)
    rng = np.random.default_rng(seed)
    gam = np.where(rng.uniform(size=len(ids)) < 0.4, 0.0, np.abs(rng.normal(size=len(ids))))
    F = oracle.ref_lcp_D(n, ids, normal, gam)
    assert np.array_equal(F, oracle.apply_D(n, ids, normal, gam)[0])
    U = oracle.ref_lcp_M(RADIUS, MU, F)
    assert np.array_equal(U, oracle.apply_M(r, MU, F)[0])
    gd = oracle.ref_lcp_DT(n, ids, normal, U)
    assert np.array_equal(gd, oracle.apply_DT(n, ids, normal, U))
    assert oracle.ref_lcp_residual(gam, sep, gd, DT) == oracle.max_abs_projected_sep(gam, sep, gd, DT)
    for K in (0, 1, 2, 3, 4, 7, 40):
        ref = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, 0.0, K, pos, sep, normal)
        out = oracle_solve(pos, ids, sep, normal, 0.0, 10000, K)
        for key in ("gamma", "F", "U"):
            assert np.array_equal(out[key], ref[key]), (K, key)
        assert (out["res"], out["it"], out["max_disp"]) == (ref["res"], ref["it"], ref["max_disp"])
    ref = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, TOL, 10000, pos, sep, normal)
    out = oracle_solve(pos, ids, sep, normal, TOL, 10000, -1)
    assert out["it"] == ref["it"] and np.array_equal(out["gamma"], ref["gamma"])
    # warm start (gamma is in/out in the reference, :694)
    ref = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, 0.0, 3, pos, sep, normal, gamma0=gam)
    with oracle.sum_order(oracle.SUM_SERIAL):
        out = oracle.resolve_collisions(n, ids, normal, r, MU, DT, 0.0, 10000, sep, gamma0=gam, fixed_iters=3)
    assert np.array_equal(out["gamma"], ref["gamma"])
