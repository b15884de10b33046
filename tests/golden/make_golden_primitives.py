"""Generates tests/golden/primitives_ref.npz: outputs of the REFERENCE's own headers for the remaining primitives
 on seeded inputs, through oracle/_ref/libmundy_ref.so :: ref_* of oracle/ref_primitives.cpp
(mundy/geom/src/mundy_geom/compute_aabb.hpp:47-104; distance/{PointSphere,PointLine,LineLine,LineSphere,PointEllipsoid,
EllipsoidEllipsoid,Circle3DCircle3D}.hpp; mundy/math/src/mundy_math/minimize.hpp).
Run in the build container (needs /root/reference): python tests/golden/make_golden_primitives.py

Inputs per function: n = 301 random items (not a multiple of any tile) with edge cases planted at the front: parallel and
identical lines (LineLine's D < 1e-15 branch), a point on the line, spherical and axis-aligned ellipsoids (the
alpha1 = alpha2 = 0 branches of the foot-point map), concentric ellipsoids, coplanar and concentric circles.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle  # noqa: E402


def make_inputs(seed=20260404, n=301):
    rng = np.random.default_rng(seed)

    def P(s=5.0):
        return rng.uniform(-s, s, (n, 3))

    def Q():
        q = rng.normal(size=(n, 4))
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        q[:8] = [1.0, 0.0, 0.0, 0.0]
        q[8:12] *= 1.7  # non-unit quaternions: the reference rotates with the general inverse
        return q

    def R3():
        r = rng.uniform(0.5, 2.0, (n, 3))
        r[:4] = 1.25  # spheres
        return r

    I = {}
    I["aabb_point"] = (P(),)
    I["aabb_line_segment"] = (P(), P())
    I["aabb_ellipsoid"] = (P(), Q(), R3())
    I["point_sphere"] = (P(), P(), rng.uniform(0.2, 2.0, n))
    x, lc, ld = P(), P(), P(1.0)
    x[:3] = lc[:3] + 0.7 * ld[:3]  # on the line
    I["point_line"] = (x, lc, ld)
    c1, d1, c2, d2 = P(), P(1.0), P(), P(1.0)
    d2[:4] = d1[:4]                # parallel
    d2[4:6] = -2.0 * d1[4:6]       # antiparallel, different length
    c2[6:8], d2[6:8] = c1[6:8], d1[6:8]  # identical
    I["line_line"] = (c1, d1, c2, d2)
    I["line_sphere"] = (P(), P(1.0), P(), rng.uniform(0.2, 2.0, n))
    x, c = P(8.0), P(2.0)
    x[:2] = c[:2] + [3.0, 0.0, 0.0]   # on the x axis of an identity-oriented sphere
    x[2:4] = c[2:4] + [0.0, 0.0, -3.0]
    I["point_ellipsoid"] = (x, c, Q(), R3())
    ca, cb = P(), P()
    cb[4:6] = ca[4:6]                 # concentric
    I["ellipsoid_ellipsoid"] = (ca, Q(), R3(), cb, Q(), R3())
    ca, cb, qa, qb = P(3.0), P(3.0), Q(), Q()
    cb[8:10] = ca[8:10]               # concentric
    I["circle_circle"] = (ca, qa, rng.uniform(0.5, 2.0, n), cb, qb, rng.uniform(0.5, 2.0, n))
    return I


if __name__ == "__main__":
    assert oracle.ref is not None, "oracle/_ref/libmundy_ref.so is needed (build with /root/reference present)"
    out = {}
    for name, ins in make_inputs().items():
        res = oracle.primitive(name, *ins, lib=oracle.ref, prefix="ref_")
        res = res if isinstance(res, tuple) else (res,)
        for k, a in enumerate(ins):
            out[f"{name}.in{k}"] = a
        for k, a in enumerate(res):
            out[f"{name}.out{k}"] = a
    path = os.path.join(HERE, "primitives_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
