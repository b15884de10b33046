"""Generates tests/golden/periodic_ref.npz: outputs of the REFERENCE's own periodic metric classes
(mundy/geom/src/mundy_geom/periodicity.hpp, through oracle/_ref/libmundy_ref.so :: ref_periodic) on seeded inputs.
Run in the build container (needs /root/reference to have built oracle/_ref): python tests/golden/make_golden_periodic.py

Inputs follow the reference's own property tests (mundy/geom/tests/unit_tests/UnitTestPeriodicity.cpp:623-948):
points in / around a 100^3 box, reference points in the disjoint box translated by one period, segments that span the
boundary; plus a non-cubic cell and the partially periodic metrics (X, Y, XY, YZ).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle  # noqa: E402

assert oracle.ref is not None, "oracle/_ref/libmundy_ref.so is needed (build with /root/reference present)"

rng = np.random.default_rng(20260101)
n = 1000
cells = {"cube100": [100.0, 100.0, 100.0], "ortho": [37.5, 81.0, 12.25], "x": [100.0, 0.0, 0.0], "y": [0.0, 81.0, 0.0],
         "xy": [100.0, 81.0, 0.0], "yz": [0.0, 81.0, 12.25]}
a = rng.uniform(-250.0, 350.0, (n, 3))          # anywhere within a few images
b = rng.uniform(100.0, 200.0, (n, 3))           # "disjoint box" reference points (UnitTestPeriodicity.cpp:745-747)
c = a + rng.uniform(-40.0, 40.0, (n, 3))        # second end point of a segment, may span a boundary
# exact half-period and boundary cases (round-half-away-from-zero, safe_unit_mod1 guard)
a[:8] = [[50, 50, 50], [-50, -50, -50], [150, 0, 0], [100, 100, 100], [0, 0, 0], [-1e-17, 1e-17, 0], [99.99999999999999, 0, 0],
         [250, -150, 50]]
b[:8] = [[0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 0], [100, 100, 100], [0, 0, 0], [0, 0, 0], [0, 0, 0]]
out = dict(a=a, b=b, c=c)
for name, cell in cells.items():
    out[f"cell_{name}"] = np.array(cell)
    out[f"{name}_sep"] = oracle.periodic(0, cell, a, b, lib=oracle.ref)
    out[f"{name}_wrap"] = oracle.periodic(1, cell, a, lib=oracle.ref)
    out[f"{name}_unwrap"] = oracle.periodic(2, cell, a, b, lib=oracle.ref)
    s0, s1 = oracle.periodic(3, cell, a, b, c, lib=oracle.ref)
    out[f"{name}_seg0"], out[f"{name}_seg1"] = s0, s1
    # Sphere / Spherocylinder overloads move the centre exactly like the Point overload
    assert np.array_equal(out[f"{name}_unwrap"], oracle.periodic(4, cell, a, b, lib=oracle.ref))
    assert np.array_equal(out[f"{name}_unwrap"], oracle.periodic(5, cell, a, b, lib=oracle.ref))
np.savez_compressed(os.path.join(HERE, "periodic_ref.npz"), **out)
print("wrote periodic_ref.npz", {k: v.shape for k, v in out.items() if k.endswith("_sep")})
