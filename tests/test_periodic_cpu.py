"""CPU tests (no GPU) of the periodic extension: the oracle's restatement of
mundy/geom/src/mundy_geom/periodicity.hpp is pinned bit-for-bit against the reference's own classes (stored outputs in
tests/golden/periodic_ref.npz and, when oracle/_ref is present, live), the reference's property tests
(mundy/geom/tests/unit_tests/UnitTestPeriodicity.cpp:623-948) are reproduced, and the periodic search / narrow phase of
the oracle are checked against an explicit 27-image brute force."""
import itertools
import os

import numpy as np
import pytest

import oracle

G = os.path.join(os.path.dirname(__file__), "golden")
PER = np.load(os.path.join(G, "periodic_ref.npz"))
NAMES = ["cube100", "ortho", "x", "y", "xy", "yz"]
RELAXED = 1e-8  # math::get_relaxed_zero_tolerance<double>() (mundy/math/src/mundy_math/Tolerance.hpp:53-69)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_outputs(name):
    """Bit-for-bit against outputs of the reference's PeriodicScaledMetric / PeriodicMetricX,Y,XY,YZ."""
    cell, a, b, c = PER[f"cell_{name}"], PER["a"], PER["b"], PER["c"]
    assert np.array_equal(oracle.periodic(0, cell, a, b), PER[f"{name}_sep"])
    assert np.array_equal(oracle.periodic(1, cell, a), PER[f"{name}_wrap"])
    assert np.array_equal(oracle.periodic(2, cell, a, b), PER[f"{name}_unwrap"])
    s0, s1 = oracle.periodic(3, cell, a, b, c)
    assert np.array_equal(s0, PER[f"{name}_seg0"]) and np.array_equal(s1, PER[f"{name}_seg1"])


@pytest.mark.skipif(oracle.ref is None, reason="oracle/_ref/libmundy_ref.so not present")
@pytest.mark.parametrize("cell", [[100.0, 100.0, 100.0], [3.0, 7.5, 11.0], [64.0, 0, 0], [0, 5.0, 0], [9.0, 5.0, 0],
                                  [0, 5.0, 17.0]])
def test_oracle_vs_live_reference_random(cell):
    rng = np.random.default_rng(11)
    a, b = rng.uniform(-500, 500, (50000, 3)), rng.uniform(-500, 500, (50000, 3))
    c = a + rng.normal(scale=20.0, size=a.shape)
    for op in (0, 1, 2):
        assert np.array_equal(oracle.periodic(op, cell, a, b), oracle.periodic(op, cell, a, b, lib=oracle.ref))
    got, want = oracle.periodic(3, cell, a, b, c), oracle.periodic(3, cell, a, b, c, lib=oracle.ref)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_min_image_direct_vs_periodic():
    """UnitTestPeriodicity.cpp:623-660: min over the 27 images == |metric.sep| (tolerance 1e-8)."""
    rng = np.random.default_rng(1234)
    cell = np.array([100.0, 100.0, 100.0])
    p1, p2 = rng.uniform(0, 100, (100000, 3)), rng.uniform(0, 100, (100000, 3))
    best = np.full(len(p1), np.inf)
    for s in itertools.product((-1, 0, 1), repeat=3):
        best = np.minimum(best, np.linalg.norm(p2 + np.array(s) * cell - p1, axis=1))
    got = np.linalg.norm(oracle.periodic(0, cell, p1, p2), axis=1)
    assert np.max(np.abs(got - best)) <= RELAXED


def test_wrap_and_unwrap_properties():
    """UnitTestPeriodicity.cpp:701-791: wrapped points lie in the primary cell; unwrapped points lie within half a
    period of the reference point on periodic axes and differ from the input by whole periods."""
    rng = np.random.default_rng(5)
    for cell in ([100.0, 100.0, 100.0], [100.0, 80.0, 0.0]):
        cellv = np.array(cell)
        on = cellv > 0
        p = rng.uniform(-400, 400, (20000, 3))
        ref = rng.uniform(100, 200, (20000, 3))
        w = oracle.periodic(1, cell, p)
        assert np.all(w[:, on] >= 0.0) and np.all(w[:, on] < cellv[on]) and np.array_equal(w[:, ~on], p[:, ~on])
        k = (p[:, on] - w[:, on]) / cellv[on]
        assert np.max(np.abs(k - np.round(k))) <= RELAXED
        u = oracle.periodic(2, cell, p, ref)
        assert np.all(np.abs(u[:, on] - ref[:, on]) <= 0.5 * cellv[on] + RELAXED)
        k = (u[:, on] - p[:, on]) / cellv[on]
        assert np.max(np.abs(k - np.round(k))) <= RELAXED
        assert np.allclose(u[:, ~on], p[:, ~on], rtol=0, atol=1e-12)


def test_unwrap_spanning_segments():
    """UnitTestPeriodicity.cpp:856-907: a segment that spans the boundary, wrapped point by point (so that its
    Euclidean length explodes), is restored by unwrap_points_to_ref about its start point."""
    rng = np.random.default_rng(1234)
    cell = [100.0, 100.0, 100.0]
    start = rng.uniform(80, 100, (1000, 3))
    end = start + rng.uniform(0, 45, (1000, 3)) / np.sqrt(3)
    ws, we = oracle.periodic(1, cell, start), oracle.periodic(1, cell, end)
    s0, s1 = oracle.periodic(3, cell, ws, ws, we)
    assert np.max(np.abs((s1 - s0) - (end - start))) <= RELAXED
    assert np.max(np.abs(s0 - ws)) <= RELAXED


def _image_brute(pos, S, cell):
    """Pairs (i<j) whose nearest image satisfies |d|^2 <= (S_i + S_j)^2, by explicit enumeration of the 27 images."""
    cellv = np.where(np.array(cell) > 0, np.array(cell), 0.0)
    on = cellv > 0
    pos = pos.copy()
    pos[:, on] = np.mod(pos[:, on], cellv[on])  # into the primary cell, so that the nearest image is one period away at most
    n = len(S)
    out = []
    shifts = [np.array(s) * cellv for s in itertools.product((-1, 0, 1), repeat=3)]
    for i in range(n):
        d = pos[i + 1:] - pos[i]
        best = np.full(len(d), np.inf)
        for s in shifts:
            best = np.minimum(best, np.einsum("ij,ij->i", d + s, d + s))
        lim = (S[i] + S[i + 1:]) ** 2
        margin = 1e-9 * lim
        assert not np.any(np.abs(best - lim) < margin), "degenerate test input"
        for j in np.nonzero(best <= lim)[0]:
            out.append((i, i + 1 + j))
    return np.array(out, dtype=np.int32).reshape(-1, 2)


@pytest.mark.parametrize("cell,n", [([20.0, 20.0, 20.0], 600), ([9.0, 30.0, 14.0], 500), ([25.0, 0.0, 0.0], 500),
                                    ([0.0, 12.0, 18.0], 500), ([5.0, 5.0, 5.0], 60)])
def test_periodic_search_oracle(cell, n):
    """brute == cell list == explicit images; short axes (fewer than three cells) included. Positions are not wrapped:
    bodies lie in several images of the cell."""
    rng = np.random.default_rng(8)
    ext = np.where(np.array(cell) > 0, np.array(cell), 30.0)
    pos = rng.uniform(-1.0, 2.0, (n, 3)) * ext
    S = rng.uniform(0.6, 1.2, n)
    if min(ext) < 6:
        S *= 0.9  # keep 2 (S_i + S_j) below the shortest period so that the nearest image is unique
    brute = oracle.search(pos, S, brute=True, cell=cell)
    cells = oracle.search(pos, S, cell=cell)
    assert np.array_equal(brute, cells)
    assert np.array_equal(brute, _image_brute(pos, S, cell))
    assert len(brute) > len(oracle.search(pos, S, brute=True)) or n < 100  # periodicity adds links across the boundary


def test_periodic_narrow_phase_oracle():
    """Periodic narrow phase == free-space narrow phase with body j moved to its nearest image by hand."""
    from mundy_b200 import workloads as wl
    w = wl.mixed(800, phi=0.25)
    cell = [w["edge"]] * 3
    S = oracle.search_radii(w["radius"], 0.3, w["kind"], w["length"])
    pairs = oracle.search(w["center"], S, cell=cell)
    phi, nrm, ai, aj = oracle.narrow_phase(pairs, w["center"], w["radius"], w["quat"], w["length"], w["kind"], True, cell)
    for k in range(0, len(pairs), 7):
        i, j = pairs[k]
        pos2 = w["center"].copy()
        pos2[j] = w["center"][i] + oracle.periodic(0, cell, w["center"][i:i + 1], w["center"][j:j + 1])[0]
        p1, n1, a1, b1 = oracle.narrow_phase(pairs[k:k + 1], pos2, w["radius"], w["quat"], w["length"], w["kind"], True)
        assert p1[0] == phi[k] and np.array_equal(n1[0], nrm[k]) and np.array_equal(a1[0], ai[k]) and np.array_equal(b1[0], aj[k])
    # links across the boundary exist and are short
    far = np.linalg.norm(w["center"][pairs[:, 1]] - w["center"][pairs[:, 0]], axis=1) > 0.5 * w["edge"]
    assert far.any() and np.all(phi[far] < 2 * S.max())


def test_periodic_step_translation_invariance():
    """A periodic system has no boundary: shifting every body by a common vector (and re-wrapping) must leave pair
    set, separations and forces unchanged up to rounding. Free space fails this test by construction."""
    from mundy_b200 import workloads as wl
    w = wl.polydisperse_spheres(1500, phi=0.3)
    cell = [w["edge"]] * 3
    shift = np.array([0.37, -0.21, 0.55]) * w["edge"]
    outs = []
    for c0 in (w["center"], oracle.periodic(1, cell, w["center"] + shift)):
        state, links, out, U = oracle.collision_step((c0.copy(), None), w, None, 0.4, False, 1e-3, 1e-4, 1e-5, 10000,
                                                     12, None, cell=cell)
        outs.append((links[0], out))
    assert np.array_equal(outs[0][0], outs[1][0])
    assert np.allclose(outs[0][1]["phi0"], outs[1][1]["phi0"], rtol=0, atol=1e-10)
    assert np.allclose(outs[0][1]["F"], outs[1][1]["F"], rtol=1e-7, atol=1e-9 * np.abs(outs[0][1]["F"]).max())
