"""CPU tests (no GPU) of the remaining primitives. The oracle's restatement of compute_aabb(Point /
LineSegment / Ellipsoid), of the closed-form distances (Point-Sphere, Point-Line, Line-Line, Line-Sphere) and of the three
minimiser-based distances (Point-Ellipsoid, Ellipsoid-Ellipsoid, Circle3D-Circle3D, including the reference's L-BFGS
minimiser mundy/math/src/mundy_math/minimize.hpp) is pinned BIT-FOR-BIT against the reference's own headers: stored
outputs (tests/golden/primitives_ref.npz, generator make_golden_primitives.py) and, when oracle/_ref is present, live on
fresh random inputs. The reference's known-answer cases are reproduced to the reference's tolerances."""
import os

import numpy as np
import pytest

import oracle
import primitive_cases as pc

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "primitives_ref.npz"))
NAMES = list(oracle.PRIMITIVE_OUTPUTS)


def _ins(name):
    k, out = 0, []
    while f"{name}.in{k}" in G:
        out.append(G[f"{name}.in{k}"])
        k += 1
    return out


def _tuple(x):
    return x if isinstance(x, tuple) else (x,)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_matches_reference_outputs(name):
    got = _tuple(oracle.primitive(name, *_ins(name)))
    for k, g in enumerate(got):
        assert np.array_equal(g, G[f"{name}.out{k}"]), (name, k)


@pytest.mark.skipif(oracle.ref is None, reason="oracle/_ref/libmundy_ref.so not present")
@pytest.mark.parametrize("name", NAMES)
def test_oracle_vs_live_reference_random(name):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    from make_golden_primitives import make_inputs
    ins = make_inputs(seed=99, n=257)[name]
    a = _tuple(oracle.primitive(name, *ins))
    b = _tuple(oracle.primitive(name, *ins, lib=oracle.ref, prefix="ref_"))
    for k, (x, y) in enumerate(zip(a, b)):
        assert np.array_equal(x, y), (name, k)


@pytest.mark.parametrize("case", pc.aabb_cases() + [pc.ellipsoid_cases()], ids=lambda c: c[0])
def test_reference_known_answers(case):
    name, ins, want, tol = case
    got = _tuple(oracle.primitive(name, *ins))
    for k, w in want.items():
        assert np.max(np.abs(got[k] - w)) <= tol, (name, got[k], w)


def test_ellipsoids_with_equal_radii_are_spheres():
    """UnitTestEllipsoidEllipsoid.cpp:54-117, tolerance 1e-4."""
    (pe_in, pe_want), (ee_in, ee_want) = pc.sphere_property_inputs()
    assert np.max(np.abs(oracle.primitive("point_ellipsoid", *pe_in)[0] - pe_want)) <= 1e-4
    assert np.max(np.abs(oracle.primitive("ellipsoid_ellipsoid", *ee_in)[0] - ee_want)) <= 1e-4


def test_closed_form_properties():
    rng = np.random.default_rng(3)
    n = 500
    x, c, d = rng.normal(size=(n, 3)), rng.normal(size=(n, 3)), rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    dist, cp, arch = oracle.primitive("point_line", x, c, d)
    assert np.allclose(np.einsum("ij,ij->i", x - cp, d), 0.0, atol=1e-12)       # foot of the perpendicular
    assert np.allclose(dist, np.linalg.norm(np.cross(x - c, d), axis=1), atol=1e-12)
    c2, d2 = rng.normal(size=(n, 3)), rng.normal(size=(n, 3))
    dist, cp1, cp2, a1, a2, sep = oracle.primitive("line_line", c, d, c2, d2)
    nrm = np.cross(d, d2)
    assert np.allclose(dist, np.abs(np.einsum("ij,ij->i", c2 - c, nrm)) / np.linalg.norm(nrm, axis=1), atol=1e-10)
    assert np.allclose(sep, cp2 - cp1)
    di = rng.integers(1, 4, (n, 3)).astype(np.float64)   # small integers: D = a c - b^2 is exactly 0
    dist, cp1, cp2, a1, a2, sep = oracle.primitive("line_line", c, di, c2, 2.0 * di)  # parallel: the reference returns 0
    assert not dist.any() and not sep.any()
    r = rng.uniform(0.1, 1.0, n)
    ds, sepv = oracle.primitive("point_sphere", x, c, r)
    assert np.allclose(np.linalg.norm(sepv, axis=1), np.abs(ds))


def test_circle_distance_is_a_minimum_over_a_dense_scan():
    ins = [a[:40] for a in _ins("circle_circle")]
    dist, cp1, cp2, n1 = oracle.primitive("circle_circle", *ins)
    th = np.linspace(0.0, 2 * np.pi, 721)
    for i in range(40):
        def pts(c, q, r):
            local = np.stack([r * np.cos(th), r * np.sin(th), 0 * th], 1)
            return oracle.quat_rotate(np.repeat(q[None], len(th), 0), local) + c
        A, B = pts(ins[0][i], ins[1][i], ins[2][i]), pts(ins[3][i], ins[4][i], ins[5][i])
        scan = np.min(np.linalg.norm(A[:, None, :] - B[None, :, :], axis=2))
        assert dist[i] <= scan + 1e-6 and dist[i] >= scan - 2e-2   # the scan's grid is 0.5 degrees
        assert abs(np.linalg.norm(cp2[i] - cp1[i]) - dist[i]) < 1e-12


# ---- the product's own minimiser source, instantiated on the host -----------------------------------------------------
def _host_harness():
    """tests/cpp/host_minimize_check.cu: mundy_b200/csrc/minimize.cuh (the text the device kernels are compiled from,
    __host__ __device__) compiled for the HOST by nvcc with contraction off, behind the oracle's calling convention."""
    import ctypes
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    src = os.path.join(root, "tests", "cpp", "host_minimize_check.cu")
    hdr = os.path.join(root, "mundy_b200", "csrc", "minimize.cuh")
    out = os.path.join(root, "build", "libhost_minimize_check.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O2", "-fmad=false", "-Xcompiler", "-fPIC",
               "-Xcompiler", "-ffp-contract=off", "-shared", "-I", os.path.join(root, "include"), "-I",
               os.path.join(root, "mundy_b200", "csrc"), src, "-o", out, "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    return ctypes.CDLL(out)


@pytest.mark.parametrize("name", ["point_ellipsoid", "ellipsoid_ellipsoid", "circle_circle"])
def test_product_minimiser_source_on_the_host_matches_reference_bit_for_bit(name):
    """The device code path cannot be bit-exact (CUDA's sin / cos); the SOURCE is: compiled for the host, where sin / cos
    are glibc's like the reference's, it reproduces the reference's stored outputs exactly."""
    lib = _host_harness()
    got = _tuple(oracle.primitive(name, *_ins(name), lib=lib, prefix="hostchk_"))
    for k, g in enumerate(got):
        assert np.array_equal(g, G[f"{name}.out{k}"]), (name, k)
