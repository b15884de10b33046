"""GPU parity of the non-local mobility row, all through the C-ABI: pair kernels bit-exact against
the reference's stored outputs and the oracle (all-pairs and neighbour-list forms), the operator U = M F, and the
operator-form BBPGD against the oracle."""
import os

import numpy as np
import pytest
import torch

import oracle
from mundy_b200 import workloads as wl

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
KINDS = [(0, "stokes"), (1, "rpy"), (2, "rpyc")]


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def host(t):
    return t.detach().cpu().numpy()


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


@pytest.mark.parametrize("kind,name", KINDS)
def test_pair_kernels_reference_golden(cuda, kind, name):
    """Bit-for-bit against outputs of the reference's own Periphery.hpp kernels (tests/golden/mobility_ref.npz)."""
    from mundy_b200 import ops
    M = np.load(os.path.join(G, "mobility_ref.npz"))
    v = dev(M["vel0"])
    ops.mobility_kernel_apply(kind, float(M["mu"]), dev(M["tpos"]), dev(M["spos"]), dev(M["srad"]), dev(M["trad"]),
                              dev(M["sforce"]), v)
    assert np.array_equal(host(v), M[name])


def test_three_spheres_reference_golden(cuda):
    from mundy_b200 import ops
    M = np.load(os.path.join(G, "mobility_ref.npz"))
    radius = float(M["tri_radius"])
    for k, s in enumerate(M["tri_s"]):
        pos = np.array([[0.0, 0.0, 0.0], [-s * radius / 2.0, -np.sqrt(3.0) * s * radius / 2.0, 0.0],
                        [s * radius / 2.0, -np.sqrt(3.0) * s * radius / 2.0, 0.0]])
        f = np.zeros((3, 3))
        f[0, 1] = -6.0 * np.pi * radius
        rad = dev(np.full(3, radius))
        self_term = f / (6.0 * np.pi * radius * 1.0)
        for kind, key in ((1, "tri_rpy"), (2, "tri_rpyc")):
            v = torch.zeros((3, 3), dtype=torch.float64, device="cuda")
            ops.mobility_kernel_apply(kind, 1.0, dev(pos), dev(pos), rad, rad, dev(f), v)
            assert np.array_equal(host(v) + self_term, M[key][k])


@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_all_pairs_vs_oracle_ragged(cuda, kind):
    """Sizes that are not multiples of the target / source tiles; coincident points; accumulation into non-zero
    velocities; Stokes with NULL radii."""
    from mundy_b200 import ops
    rng = np.random.default_rng(40 + kind)
    nt, ns = 1301, 2777
    tp, sp = rng.uniform(0, 40, (nt, 3)), rng.uniform(0, 40, (ns, 3))
    sp[:300] = tp[:300]
    sr, tr = rng.uniform(0.1, 1.8, ns), rng.uniform(0.1, 1.8, nt)
    f, v0 = rng.normal(size=(ns, 3)), rng.normal(size=(nt, 3))
    want = oracle.mobility_kernel(kind, 0.37, tp, sp, sr, tr, f, v0)
    v = dev(v0)
    nul = kind == 0
    ops.mobility_kernel_apply(kind, 0.37, dev(tp), dev(sp), None if nul else dev(sr), None if nul else dev(tr), dev(f), v)
    assert np.array_equal(host(v), want)
    # empty source or target sets leave the velocities alone
    v2 = dev(v0)
    ops.mobility_kernel_apply(kind, 0.37, dev(tp), dev(sp[:0]), dev(sr[:0]), dev(tr), dev(f[:0]), v2)
    assert np.array_equal(host(v2), v0)


def _sym_csr(pairs, n):
    i, j = pairs[:, 0], pairs[:, 1]
    src, dst = np.concatenate([i, j]), np.concatenate([j, i])
    order = np.lexsort((dst, src))
    offsets = np.concatenate([[0], np.cumsum(np.bincount(src, minlength=n))]).astype(np.int32)
    return offsets, dst[order].astype(np.int32)


@pytest.mark.parametrize("kind", [0, 1, 2, 3])
def test_neighbor_form_vs_oracle(cuda, kind):
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(20000, phi=0.3)
    n = len(w["radius"])
    pairs = oracle.search(w["center"], oracle.search_radii(w["radius"], 1.0))
    offsets, indices = _sym_csr(pairs, n)
    assert offsets[-1] > 10 * n
    rng = np.random.default_rng(2)
    f, v0 = rng.normal(size=(n, 3)), rng.normal(size=(n, 3))
    want = oracle.mobility_kernel_neighbors(kind, 1e-3, w["center"], w["radius"], offsets, indices, f, v0)
    v = dev(v0)
    ops.mobility_kernel_apply_neighbors(kind, 1e-3, dev(w["center"]), dev(w["radius"]), dev(offsets, torch.int32),
                                        dev(indices, torch.int32), dev(f), v)
    assert np.array_equal(host(v), want)
    # the driver's one-directional list (row j lists i < j, KokkosLCP.cpp:404-412)
    lo_off = np.concatenate([[0], np.cumsum(np.bincount(pairs[:, 1], minlength=n))]).astype(np.int32)
    lo_idx = pairs[np.lexsort((pairs[:, 0], pairs[:, 1])), 0].astype(np.int32)
    want = oracle.mobility_kernel_neighbors(kind, 1e-3, w["center"], w["radius"], lo_off, lo_idx, f)
    v = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    ops.mobility_kernel_apply_neighbors(kind, 1e-3, dev(w["center"]), dev(w["radius"]), dev(lo_off, torch.int32),
                                        dev(lo_idx, torch.int32), dev(f), v)
    assert np.array_equal(host(v), want)


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("kernel", [1, 3])
def test_mobility_operator_vs_oracle(cuda, mode, kernel):
    """U = M F: local drag alone, + neighbour kernel, + all-pairs kernel; both forms of the drag coefficient."""
    from mundy_b200 import ops
    w = wl.polydisperse_spheres(3000, phi=0.3)
    n = len(w["radius"])
    pairs = oracle.search(w["center"], oracle.search_radii(w["radius"], 0.5))
    offsets, indices = _sym_csr(pairs, n)
    F = np.random.default_rng(6).normal(size=(n, 3))
    want = oracle.mobility_apply(mode, kernel, 2e-3, w["center"], w["radius"], F, offsets, indices)
    mob = ops.make_mobility_op(mode, kernel, 2e-3, dev(w["radius"]), dev(w["center"]), dev(offsets, torch.int32),
                               dev(indices, torch.int32))
    assert np.array_equal(host(ops.mobility_apply(mob, dev(F))), want)
    if mode == 0:  # the standalone local-drag entry point accumulates
        v = dev(0.5 * F)
        ops.apply_local_drag(kernel, 2e-3, dev(w["radius"]), dev(F), v)
        assert np.array_equal(host(v), oracle.local_drag(kernel, 2e-3, w["radius"], F, 0.5 * F))


def _contacts(n, phi, buffer=0.3):
    w = wl.polydisperse_spheres(n, phi=phi)
    pairs = oracle.search(w["center"], oracle.search_radii(w["radius"], buffer))
    phi0, nrm = oracle.narrow_phase(pairs, w["center"], w["radius"])[:2]
    return w, pairs, phi0, nrm


@pytest.mark.parametrize("mode,kernel", [(0, 1), (1, 3), (2, 2), (2, 3)])
def test_operator_form_solve_fixed_iterations(cuda, mode, kernel):
    """Same iterates as the oracle's restatement of NgpHP1.cpp:1487-1645 after a fixed number of iterations, with an
    external force (phi0 correction through the operator)."""
    from mundy_b200 import ops
    w, pairs, phi0, nrm = _contacts(4000, 0.3)
    n, mu, dt = len(w["radius"]), 1e-3, 1e-4
    offsets, indices = _sym_csr(pairs, n)
    F_ext = np.random.default_rng(8).normal(size=(n, 3)) * 1e-3
    iters = 6 if mode == 1 else 12  # the truncated operator is indefinite: it is compared before it blows up
    want = oracle.resolve_collisions_mobility(pairs, nrm, mode, kernel, mu, w["center"], w["radius"], dt, 1e-6, 10000, phi0,
                                              fixed_iters=iters, offsets=offsets, indices=indices, F_ext=F_ext)
    op = ops.make_contact_op(n, dev(pairs, torch.int32), dev(nrm), dev(w["radius"]), mu, dt)
    mob = ops.make_mobility_op(mode, kernel, mu, dev(w["radius"]), dev(w["center"]), dev(offsets, torch.int32),
                               dev(indices, torch.int32))
    p, g = dev(phi0), torch.zeros(len(pairs), dtype=torch.float64, device="cuda")
    res, F, U, Ue = ops.resolve_collisions_mobility(op, mob, p, g, 1e-6, 10000, fixed_iters=iters, F_ext=dev(F_ext))
    assert res.ite_count == want["it"] == iters
    assert np.array_equal(host(Ue), want["U_ext"]) and np.array_equal(host(p), want["phi0"])
    assert rel(host(g), want["gamma"]) <= 1e-9 and rel(host(F), want["F"]) <= 1e-9 and rel(host(U), want["U"]) <= 1e-9
    assert abs(res.max_abs_projected_sep - want["res"]) <= 1e-9 * abs(want["res"])
    assert abs(res.max_displacement - want["max_disp"]) <= 1e-9 * want["max_disp"]


def test_operator_form_solve_converges_all_pairs(cuda):
    """Run to tolerance with the full corrected-RPY mobility: iteration count equals the oracle's, the solution satisfies
    the LCP."""
    from mundy_b200 import ops
    w, pairs, phi0, nrm = _contacts(3000, 0.3)
    n, mu, dt, tol = len(w["radius"]), 1e-3, 1e-4, 1e-6
    want = oracle.resolve_collisions_mobility(pairs, nrm, 2, 2, mu, w["center"], w["radius"], dt, tol, 20000, phi0)
    op = ops.make_contact_op(n, dev(pairs, torch.int32), dev(nrm), dev(w["radius"]), mu, dt)
    mob = ops.make_mobility_op(2, 2, mu, dev(w["radius"]), dev(w["center"]))
    p, g = dev(phi0), torch.zeros(len(pairs), dtype=torch.float64, device="cuda")
    res, F, U, _ = ops.resolve_collisions_mobility(op, mob, p, g, tol, 20000)
    assert res.converged and res.max_abs_projected_sep < tol
    assert abs(res.ite_count - want["it"]) <= max(2, want["it"] // 50)
    gam = host(g)
    assert gam.min() >= 0.0
    sep = phi0 + dt * oracle.apply_DT(n, pairs, nrm, host(U))
    assert sep.min() >= -tol and np.all(np.abs(sep[gam >= 1e-12]) < tol)
    assert rel(host(U), oracle.mobility_apply(2, 2, mu, w["center"], w["radius"], host(F))) <= 1e-12
    assert rel(host(U), want["U"]) <= 1e-4  # both are tol-accurate solutions of the same strictly convex problem


def test_mobility_argument_checks(cuda):
    from mundy_b200 import ops
    from mundy_b200._capi import Mb200Error
    x = torch.zeros((4, 3), dtype=torch.float64, device="cuda")
    r = torch.ones(4, dtype=torch.float64, device="cuda")
    with pytest.raises(Mb200Error):
        ops.mobility_kernel_apply(7, 1.0, x, x, r, r, x, x.clone())
    with pytest.raises(Mb200Error):
        ops.mobility_kernel_apply(1, 1.0, x, x, None, None, x, x.clone())  # RPY needs radii
    with pytest.raises(Mb200Error):
        ops.mobility_kernel_apply(1, 0.0, x, x, r, r, x, x.clone())  # viscosity
    with pytest.raises(Mb200Error):
        ops.mobility_apply(ops.make_mobility_op(1, 1, 1.0, r, x), x)  # NEIGHBORS without a list
