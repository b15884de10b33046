"""GPU parity of the remaining primitives, all through the C-ABI.

Closed-form functions (three AABBs, Point-Sphere, Point-Line, Line-Line, Line-Sphere): BIT-EXACT against outputs of the
reference's own headers (tests/golden/primitives_ref.npz) and against the oracle on fresh inputs of ragged sizes.

Minimiser-based functions (Point-Ellipsoid, Ellipsoid-Ellipsoid, Circle3D-Circle3D): same operation order as the
reference, but sin / cos are CUDA's (<= 2 ulp from glibc's), and the L-BFGS path amplifies that. Perturbing the CPU
oracle's sin / cos by +-2 ulp moves the distance by <= 1.1e-9 and the closest points / normals (flat directions of the
minimum) by <= 1.1e-4 over 20000 random pairs, so the bounds below are: distance 1e-7 absolute (the reference's own
unit-test tolerance is 1e-4, UnitTestEllipsoidEllipsoid.cpp:53), closest points and normals 2e-3 absolute."""
import os
import sys

import numpy as np
import pytest
import torch

import oracle
import primitive_cases as pc

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_golden_primitives import make_inputs  # noqa: E402  (input generator only; needs no reference)

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "primitives_ref.npz"))
EXACT = ["aabb_point", "aabb_line_segment", "aabb_ellipsoid", "point_sphere", "point_line", "line_line", "line_sphere"]
MINIMISED = ["point_ellipsoid", "ellipsoid_ellipsoid", "circle_circle"]
DIST_TOL, POINT_TOL = 1e-7, 2e-3


def gpu(name, *ins):
    from mundy_b200 import ops
    fn = {"aabb_point": ops.compute_aabb_point, "aabb_line_segment": ops.compute_aabb_line_segment,
          "aabb_ellipsoid": ops.compute_aabb_ellipsoid, "point_sphere": ops.distance_point_sphere,
          "point_line": ops.distance_point_line, "line_line": ops.distance_line_line,
          "line_sphere": ops.distance_line_sphere, "point_ellipsoid": ops.distance_point_ellipsoid,
          "ellipsoid_ellipsoid": ops.distance_ellipsoid_ellipsoid, "circle_circle": ops.distance_circle3d_circle3d}[name]
    out = fn(*[torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float64, device="cuda") for x in ins])
    out = out if isinstance(out, tuple) else (out,)
    torch.cuda.synchronize()
    return tuple(o.cpu().numpy() for o in out)


def golden(name):
    ins, outs, k = [], [], 0
    while f"{name}.in{k}" in G:
        ins.append(G[f"{name}.in{k}"])
        k += 1
    k = 0
    while f"{name}.out{k}" in G:
        outs.append(G[f"{name}.out{k}"])
        k += 1
    return ins, outs


def _tuple(x):
    return x if isinstance(x, tuple) else (x,)


@pytest.mark.parametrize("name", EXACT)
def test_closed_form_reference_golden_bit_exact(cuda, name):
    ins, outs = golden(name)
    got = gpu(name, *ins)
    for k, w in enumerate(outs):
        assert np.array_equal(got[k], w), (name, k)


@pytest.mark.parametrize("n", [1, 255, 257, 5000])
@pytest.mark.parametrize("name", EXACT)
def test_closed_form_vs_oracle_ragged_sizes(cuda, name, n):
    ins = make_inputs(seed=1000 + n, n=n)[name]
    got, want = gpu(name, *ins), _tuple(oracle.primitive(name, *ins))
    for k, w in enumerate(want):
        assert np.array_equal(got[k], w), (name, n, k)


@pytest.mark.parametrize("name", EXACT + MINIMISED)
def test_empty_input(cuda, name):
    ins = [a[:0] for a in golden(name)[0]]
    for o in gpu(name, *ins):
        assert o.shape[0] == 0


@pytest.mark.parametrize("name", MINIMISED)
def test_minimised_reference_golden(cuda, name):
    ins, outs = golden(name)
    got = gpu(name, *ins)
    assert np.max(np.abs(got[0] - outs[0])) <= DIST_TOL, name
    for k in range(1, len(outs)):
        assert np.max(np.abs(got[k] - outs[k])) <= POINT_TOL, (name, k)


@pytest.mark.parametrize("name", MINIMISED)
def test_minimised_vs_oracle_random(cuda, name):
    n = 3001
    ins = make_inputs(seed=4242, n=n)[name]
    got, want = gpu(name, *ins), oracle.primitive(name, *ins)
    assert np.max(np.abs(got[0] - want[0])) <= DIST_TOL, name
    for k in range(1, len(want)):
        assert np.max(np.abs(got[k] - want[k])) <= POINT_TOL, (name, k)


@pytest.mark.parametrize("case", pc.aabb_cases() + [pc.ellipsoid_cases()], ids=lambda c: c[0])
def test_reference_known_answers(cuda, case):
    name, ins, want, tol = case
    got = gpu(name, *ins)
    for k, w in want.items():
        assert np.max(np.abs(got[k] - w)) <= tol, (name, got[k], w)


def test_ellipsoids_with_equal_radii_are_spheres(cuda):
    """UnitTestEllipsoidEllipsoid.cpp:54-117, tolerance 1e-4."""
    (pe_in, pe_want), (ee_in, ee_want) = pc.sphere_property_inputs()
    assert np.max(np.abs(gpu("point_ellipsoid", *pe_in)[0] - pe_want)) <= 1e-4
    assert np.max(np.abs(gpu("ellipsoid_ellipsoid", *ee_in)[0] - ee_want)) <= 1e-4


def test_shared_normal_property_at_scale(cuda):
    """Size-independent property at 10^5 pairs: the returned closest points lie on their ellipsoids, the returned normal
    is unit, and the signed distance equals (cp2 - cp1) . n."""
    n = 100_000
    ins = make_inputs(seed=7, n=n)["ellipsoid_ellipsoid"]
    dist, cp1, cp2, n1 = gpu("ellipsoid_ellipsoid", *ins)
    assert np.allclose(np.linalg.norm(n1, axis=1), 1.0, atol=1e-12)
    assert np.allclose(np.einsum("ij,ij->i", cp2 - cp1, n1), dist, atol=1e-12)
    for c, q, R, cp in ((ins[0], ins[1], ins[2], cp1), (ins[3], ins[4], ins[5], cp2)):
        qc = q * np.array([1.0, -1.0, -1.0, -1.0])
        body = oracle.quat_rotate(qc, cp - c)
        assert np.allclose(np.sum((body / R) ** 2, axis=1), 1.0, atol=1e-9)
