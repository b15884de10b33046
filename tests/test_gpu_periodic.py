"""GPU parity of the periodic extension (BASELINE configs[1] "periodic box"), all through the C-ABI:
metric leaves bit-exact against the reference's stored outputs and the oracle, periodic pair sets bit-exact, periodic
narrow phase bit-exact, whole periodic steps (spheres, rods with the shape filter, mixed) against the oracle step."""
import os

import numpy as np
import pytest
import torch

import oracle
from mundy_b200 import workloads as wl

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
NAMES = ["cube100", "ortho", "x", "y", "xy", "yz"]


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def host(t):
    return t.detach().cpu().numpy()


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


@pytest.mark.parametrize("name", NAMES)
def test_metric_leaves_reference_golden(cuda, name):
    """Bit-for-bit against outputs of the reference's own metric classes (tests/golden/periodic_ref.npz)."""
    from mundy_b200 import ops
    PER = np.load(os.path.join(G, "periodic_ref.npz"))
    cell, a, b, c = PER[f"cell_{name}"], dev(PER["a"]), dev(PER["b"]), dev(PER["c"])
    assert np.array_equal(host(ops.periodic_sep(cell, a, b)), PER[f"{name}_sep"])
    assert np.array_equal(host(ops.periodic_wrap(cell, a)), PER[f"{name}_wrap"])
    assert np.array_equal(host(ops.periodic_unwrap_points(cell, a, b)), PER[f"{name}_unwrap"])
    s0, s1 = ops.periodic_unwrap_segments(cell, a, c, b)
    assert np.array_equal(host(s0), PER[f"{name}_seg0"]) and np.array_equal(host(s1), PER[f"{name}_seg1"])


def test_metric_leaves_large_random(cuda):
    """Ragged size (not a multiple of the tile), points many periods away."""
    from mundy_b200 import ops
    rng = np.random.default_rng(4)
    n = 200003
    cell = [13.0, 0.0, 47.5]
    a, b = rng.uniform(-1e4, 1e4, (n, 3)), rng.uniform(-1e4, 1e4, (n, 3))
    c = a + rng.normal(scale=9.0, size=a.shape)
    assert np.array_equal(host(ops.periodic_sep(cell, dev(a), dev(b))), oracle.periodic(0, cell, a, b))
    assert np.array_equal(host(ops.periodic_wrap(cell, dev(a))), oracle.periodic(1, cell, a))
    assert np.array_equal(host(ops.periodic_unwrap_points(cell, dev(a), dev(b))), oracle.periodic(2, cell, a, b))
    s0, s1 = ops.periodic_unwrap_segments(cell, dev(a), dev(c), dev(b))
    w0, w1 = oracle.periodic(3, cell, a, b, c)
    assert np.array_equal(host(s0), w0) and np.array_equal(host(s1), w1)
    empty = ops.periodic_wrap(cell, dev(np.zeros((0, 3))))
    assert empty.shape == (0, 3)


@pytest.mark.parametrize("cell,n", [([40.0, 40.0, 40.0], 20000), ([9.0, 60.0, 31.0], 6000), ([55.0, 0.0, 0.0], 8000),
                                    ([0.0, 21.0, 33.0], 8000), ([5.0, 5.0, 5.0], 80), ([4.0, 70.0, 2.6], 900)])
def test_periodic_search_pairs(cuda, cell, n):
    """Pair set bit-exact against the oracle; bodies spread over several images; short axes with one or two cells."""
    from mundy_b200 import ops
    rng = np.random.default_rng(8)
    ext = np.where(np.array(cell) > 0, np.array(cell), 45.0)
    pos = rng.uniform(-1.0, 2.0, (n, 3)) * ext
    S = rng.uniform(0.3, 0.62, n)
    want = oracle.search(pos, S, cell=cell)
    got = host(ops.search_pairs(dev(pos), dev(S), cell_size=cell))
    assert np.array_equal(got, want)
    assert len(want) > len(oracle.search(pos, S))


def test_periodic_narrow_phase(cuda):
    from mundy_b200 import ops
    w = wl.mixed(6000, phi=0.25)
    cell = [w["edge"], w["edge"], 0.0]
    S = oracle.search_radii(w["radius"], 0.3, w["kind"], w["length"])
    pairs = oracle.search(w["center"], S, cell=cell)
    want = oracle.narrow_phase(pairs, w["center"], w["radius"], w["quat"], w["length"], w["kind"], True, cell)
    b = ops.make_bodies(dev(w["center"]), dev(w["radius"]), dev(w["quat"]), dev(w["length"]),
                        dev(w["kind"], torch.int32))
    got = ops.narrow_phase(b, dev(pairs, torch.int32), arms=True, cell_size=cell)
    for g, x in zip(got, want):
        assert np.array_equal(host(g), x)


@pytest.mark.parametrize("filt", [False, True])
def test_periodic_links_rods(cuda, filt):
    """GenNeighborLinks with a periodic box: candidates and the shape-distance filter; a box only ~2.6 search cells
    wide so that the short-axis rule of the cell walk is exercised by the whole pipeline."""
    from mundy_b200 import ops
    w = wl.rods(3000, phi=0.08)
    cell = [w["edge"], w["edge"], 0.35 * w["edge"]]
    c = w["center"].copy()
    c[:, 2] *= 0.35
    buffer = 0.4
    b = ops.make_bodies(dev(c), dev(w["radius"]), dev(w["quat"]), dev(w["length"]), dev(w["kind"], torch.int32))
    links = ops.NeighborLinks().set_search_buffer(buffer).set_periodic(cell)
    if filt:
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    assert links.generate(b)
    S = oracle.search_radii(w["radius"], buffer, w["kind"], w["length"])
    want = oracle.search(c, S, cell=cell)
    if filt:
        want = oracle.filter_shape_distance(want, c, w["quat"], w["radius"], w["length"], w["kind"], 2 * buffer, cell)
    assert np.array_equal(host(links.pairs()), want)
    assert not links.generate(b)  # nothing moved
    links.set_periodic(None)  # switching the metric invalidates the list
    assert links.generate(b)
    assert links.count() < len(want)


@pytest.mark.parametrize("name", ["spheres", "rods_filter", "mixed_xy"])
def test_periodic_collision_steps(cuda, name):
    """Several whole steps in a periodic box against the oracle step: rebuild decisions, link sets, forces, velocities,
    positions. Bodies start wrapped, leave the primary cell as they move and are NOT re-wrapped by the step."""
    from mundy_b200 import ops
    if name == "spheres":
        w, buffer, filt, fixed = wl.polydisperse_spheres(8000, phi=0.3), 0.3, False, 8
        cell = [w["edge"]] * 3
    elif name == "rods_filter":
        w, buffer, filt, fixed = wl.rods(5000, phi=0.1), 0.4, True, 8
        cell = [w["edge"]] * 3
    else:
        w, buffer, filt, fixed = wl.mixed(6000, phi=0.25), 0.3, False, 6
        cell = [w["edge"], w["edge"], 0.0]
    rods = "kind" in w
    n = len(w["radius"])
    mu, dt, tol = 1e-3, 2e-4, 1e-5
    # a drift on top of the ABC forcing pushes bodies through the boundary during the test
    drift = np.array([40.0, -25.0, 0.0])
    c, q = dev(w["center"]), dev(w["quat"]) if rods else None
    b = ops.make_bodies(c, dev(w["radius"]), q, dev(w["length"]) if rods else None,
                        dev(w["kind"], torch.int32) if rods else None)
    links = ops.NeighborLinks().set_search_buffer(buffer).set_periodic(cell)
    if filt:
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    state, olinks = (w["center"].copy(), w["quat"].copy() if rods else None), None
    Fcol = torch.empty((n, 3), dtype=torch.float64, device="cuda")
    Uout = torch.empty((n, 3), dtype=torch.float64, device="cuda")
    rebuilds = 0
    for step in range(4):
        F_ext = oracle.abc_flow_force(state[0]) + drift
        stats = links.step(b, mu, dt, tol, 10000, fixed_iters=fixed, F_ext=dev(F_ext), F_col=Fcol, U_out=Uout)
        state, olinks, out, U = oracle.collision_step(state, w, olinks, buffer, filt, mu, dt, tol, 10000, fixed, F_ext,
                                                      cell=cell)
        assert bool(stats.rebuilt) == out["rebuilt"], step
        rebuilds += int(out["rebuilt"])
        assert np.array_equal(host(links.pairs()), olinks[0]), step
        assert stats.lcp.ite_count == out["it"]
        assert rel(host(Fcol), out["F"]) <= 1e-10 and rel(host(Uout), U) <= 1e-10
        assert rel(host(c), state[0]) <= 1e-12
        if rods:
            assert np.max(np.abs(host(q) - state[1])) <= 1e-12
    assert rebuilds >= 2  # the drift forces at least one rebuild after the first build
    far = np.abs(state[0][olinks[0][:, 1]] - state[0][olinks[0][:, 0]]).max(axis=1) > 0.5 * w["edge"]
    assert far.any()  # links across the periodic boundary took part


def test_periodic_step_converged_overlap_removed(cuda):
    """Run to tolerance in a periodic box: afterwards no pair (nearest images included) overlaps by more than tol."""
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(20000, phi=0.3)
    cell = [w["edge"]] * 3
    mu, dt, tol, buffer = 1e-3, 1e-4, 1e-6, 0.3
    c = dev(w["center"])
    b = ops.make_bodies(c, dev(w["radius"]))
    links = ops.NeighborLinks().set_search_buffer(buffer).set_periodic(cell)
    stats = links.step(b, mu, dt, tol, 10000)
    assert stats.lcp.converged and stats.lcp.max_abs_projected_sep < tol
    pairs = host(links.pairs())
    phi = oracle.narrow_phase(pairs, host(c), w["radius"], cell=cell)[0]
    assert phi.min() >= -10 * tol
    across = np.abs(w["center"][pairs[:, 1]] - w["center"][pairs[:, 0]]).max(axis=1) > 0.5 * w["edge"]
    assert across.any() and phi[across].min() >= -10 * tol  # links through the boundary were resolved as well


def test_periodic_box_on_the_distributed_path(cuda):
    """A periodic box through the distributed step (refused in round 1). With one rank it must equal the single-GPU
    periodic step bit for bit; 2 / 4 / 8 ranks are covered by tests/test_gpu_dist.py (periodic spheres case)."""
    from mundy_b200 import dist as mbd
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(3000, phi=0.3)
    cell = [w["edge"]] * 3
    out = []
    for comm in (None, mbd.Comm()):
        c = dev(w["center"])
        b = ops.make_bodies(c, dev(w["radius"]))
        links = ops.NeighborLinks().set_search_buffer(0.3).set_periodic(cell)
        for _ in range(3):
            if comm is None:
                st = links.step(b, 1e-3, 1e-4, 1e-5, 10000, fixed_iters=6)
            else:
                st = mbd.step_dist(links, comm, b, 1e-3, 1e-4, 1e-5, 10000, fixed_iters=6)
        out.append((host(c), st.num_links, st.lcp.max_abs_projected_sep, host(links.pairs())))
    assert out[0][1] == out[1][1] and out[0][2] == out[1][2]
    assert np.array_equal(out[0][3], out[1][3]) and np.array_equal(out[0][0], out[1][0])
    free = oracle.search(w["center"], w["radius"] + 0.3)
    assert out[1][1] > len(free)  # the periodic images add links on the distributed path too
