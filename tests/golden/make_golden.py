"""Generates the committed golden fixtures in this directory. Run in the build container (needs /root/reference and
oracle/_ref/libmundy_ref.so = the reference's own headers compiled behind the Kokkos host shim):

    python tests/golden/make_golden.py

Two kinds of fixture:
  * reference_kats.npz  - known-answer vectors transcribed from the reference's own unit tests (file:line below)
  * ref_outputs.npz     - outputs of the REFERENCE code itself (libmundy_ref.so) on small seeded inputs, so that the
                          GPU box (which has no /root/reference) can still check the oracle and the kernels against
                          what the reference computes
  * segseg_cases.npz    - the hard-coded segment-segment cases with the reference's expected distances
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle  # noqa: E402

assert oracle.ref is not None, "oracle/_ref/libmundy_ref.so is required (make -C oracle ref)"
R = dict(lib=oracle.ref, prefix="ref_")
s = np.sqrt(0.5)

# ---- known-answer tests --------------------------------------------------------------------------------------------
# mundy/geom/tests/unit_tests/UnitTestComputeAABB.cpp:169-177 (spheres), :210-234 (spherocylinders),
# :239-261 (spherocylinder segments)
kats = dict(
    sphere_center=np.array([[0, 0, 0], [1, -2, 3]], float), sphere_radius=np.array([1, 4], float),
    sphere_aabb=np.array([[-1, -1, -1, 1, 1, 1], [-3, -6, -1, 5, 2, 7]], float),
    rod_center=np.array([[1, -2, 3]] * 4, float),
    rod_quat=np.array([[1, 0, 0, 0], [1, 0, 0, 0], [1, 0, 0, 0], [s, s, 0, 0]], float),
    rod_radius=np.array([4, 0, 2, 2], float), rod_length=np.array([0, 4, 4, 3], float),
    rod_aabb=np.array([[-3, -6, -1, 5, 2, 7], [1, -2, 1, 1, -2, 5], [-1, -4, -1, 3, 0, 7], [-1, -5.5, 1, 3, 1.5, 5]], float),
    seg_start=np.array([[1, -2, 3], [1, -2, 1], [1, -2, 1], [1, -3.5, 3]], float),
    seg_end=np.array([[1, -2, 3], [1, -2, 5], [1, -2, 5], [1, -0.5, 3]], float),
    seg_radius=np.array([4, 0, 2, 2], float),
    seg_aabb=np.array([[-3, -6, -1, 5, 2, 7], [1, -2, 1, 1, -2, 5], [-1, -4, -1, 3, 0, 7], [-1, -5.5, 1, 3, 1.5, 5]], float),
)
# scrap/parameter_interface/linkers/tests/unit_tests/UnitTestComputeSignedSeparationDistanceAndContactNormal.cpp:300-330
# sphere-sphere: centres (0,0,0) r 1.5 and (1,2,2) r 2 -> ssd -0.5, normal (1/3, 2/3, 2/3)
kats.update(ss_c1=np.array([[0.0, 0, 0]]), ss_r1=np.array([1.5]), ss_c2=np.array([[1.0, 2, 2]]), ss_r2=np.array([2.0]),
            ss_ssd=np.array([-0.5]), ss_normal=np.array([[1 / 3, 2 / 3, 2 / 3]]))
np.savez(os.path.join(HERE, "reference_kats.npz"), **kats)

# mundy/geom/tests/unit_tests/UnitTestSegmentSegment.cpp:417-472 (the two "pesky" hard-coded cases)
l0 = np.array([[0.2257294191072674, 0.30159862841764695, 0.12784820133135649], [9.64101615137754, 6, 3.18961417478521]])
l1 = np.array([[0.22572948671663273, 0.30159858045792487, 0.1278481814714105], [9.64101615137754, 6, 8.189614174785209]])
m0 = np.array([[0.5220039935659887, 0.88764831847472003, -0.2219484914838093], [10.39230484541326, 6, 0.6472696138825587]])
m1 = np.array([[0.50288066060587278, 0.66779290982621586, -0.5723507723323677], [10.39230484541326, 6, 5.647269613882559]])
np.savez(os.path.join(HERE, "segseg_cases.npz"), l0=l0, l1=l1, m0=m0, m1=m1,
         dist=np.array([0.74347757392471259, 0.7512886940357237]),
         s0=np.array([1.0]), t0=np.array([0.069641589451982497]),
         cp1_0=np.array([0.22572948671663273, 0.30159858045792487, 0.1278481814714105]),
         cp2_0=np.array([0.52067221426302679, 0.87233723836682309, -0.24635106326970288]))

# ---- reference outputs on seeded inputs ----------------------------------------------------------------------------
rng = np.random.default_rng(20261104)
n = 512
out = {}
c, r, L = rng.uniform(-20, 20, (n, 3)), rng.uniform(0.1, 2, n), rng.uniform(0, 10, n)
q = rng.normal(size=(n, 4))
out.update(c=c, r=r, L=L, q=q)
out["aabb_sphere"] = oracle.aabb_sphere(c, r, **R)
out["aabb_rod"] = oracle.aabb_spherocylinder(c, q, r, L, **R)
a, b, cc, d = (rng.uniform(-3, 3, (n, 3)) for _ in range(4))
k = n // 8
d[:k] = cc[:k] + 1.7 * (b[:k] - a[:k])      # parallel: collinear branch of LineSegmentLineSegment.hpp:215-268
cc[k:2 * k] = a[k:2 * k]                    # shared endpoint
b[2 * k:3 * k] = a[2 * k:3 * k]             # zero-length segment
out.update(sa=a, sb=b, sc=cc, sd=d)
for name, v in zip(["dist", "s", "t", "cp1", "cp2", "sep"], oracle.segseg(a, b, cc, d, **R)):
    out["segseg_" + name] = v
for name, v in zip(["dist", "t", "cp", "sep"], oracle.point_seg(a, b, cc, **R)):
    out["pseg_" + name] = v
for name, v in zip(["dist", "sep"], oracle.sphere_sphere(a, r, b, L + 0.1, **R)):
    out["ss_" + name] = v
for name, v in zip(["dist", "t", "cp", "sep"], oracle.seg_sphere(a, b, cc, r, **R)):
    out["segsph_" + name] = v
om = rng.normal(size=(n, 3))
qu = q / np.linalg.norm(q, axis=1, keepdims=True)
out.update(om=om, qu=qu)
out["quat_rot"] = oracle.quat_rotate(q, om, **R)
out["quat_step"] = oracle.rotate_quaternion(qu, om, 0.01, **R)
# math::convex: strictly diagonally dominant LCP with a planted complementary solution
# (mundy/math/tests/unit_tests/UnitTestConvex.cpp:145-229), size 50, both residual policies
m = 50
A = rng.uniform(-1, 1, (m, m))
A = 0.5 * (A + A.T)
A[np.arange(m), np.arange(m)] = np.abs(A).sum(axis=1) + 1.0
xs = np.where(rng.uniform(size=m) < 0.5, rng.uniform(0.5, 2.0, m), 0.0)
w = np.where(xs == 0.0, rng.uniform(0.5, 2.0, m), 0.0)
qv = w - A @ xs
out.update(lcp_A=A, lcp_q=qv, lcp_xstar=xs)
for rk in (0, 1):
    res = oracle.solve_cqpp_dense(A, qv, np.full(m, 99.99), space_kind=1, max_iters=1000, tol=1e-6, resid_kind=rk, **R)
    out[f"lcp_x_r{rk}"] = res["x"]
    out[f"lcp_meta_r{rk}"] = np.array([res["iters"], res["residual"], float(res["converged"])])
np.savez(os.path.join(HERE, "ref_outputs.npz"), **out)
print("golden fixtures written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
