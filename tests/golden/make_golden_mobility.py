"""Generates tests/golden/mobility_ref.npz: outputs of the REFERENCE's own mobility kernels
(scrap/parameter_interface/alens/src/mundy_alens/periphery/Periphery.hpp apply_stokes_kernel / apply_rpy_kernel /
apply_rpyc_kernel / apply_local_drag, through oracle/_ref/libmundy_ref.so :: ref_mobility_kernel, ref_local_drag) on
seeded inputs. Run in the build container (needs /root/reference): python tests/golden/make_golden_mobility.py

Inputs: polydisperse spheres, sources != targets; some sources coincide with targets (the r^2 < 1e-12 self guard), some
pairs overlap (a + b > r: the corrected branch of RPYC), a few are nested (r < |a - b|: local-drag branch) and two
sources have zero radius (RPYC falls through to the last branch). Also the three-sphere configuration of the reference's
unit test (UnitTestPeriphery.cpp:671-775, Wilson 2013 table) with the reference's outputs for RPY and RPYC.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle  # noqa: E402

assert oracle.ref is not None, "oracle/_ref/libmundy_ref.so is needed (build with /root/reference present)"

rng = np.random.default_rng(20260202)
nt, ns = 257, 391  # not multiples of any tile
mu = 0.77
tpos = rng.uniform(0.0, 12.0, (nt, 3))
spos = rng.uniform(0.0, 12.0, (ns, 3))
srad = rng.uniform(0.2, 1.1, ns)
trad = rng.uniform(0.2, 1.1, nt)
spos[:40] = tpos[:40]                                  # self pairs
spos[40:80] = tpos[40:80] + rng.normal(scale=0.3, size=(40, 3))   # overlapping
spos[80:90] = tpos[80:90] + rng.normal(scale=0.02, size=(10, 3))  # nested
srad[80:90] = 1.5
srad[90:92] = 0.0
sforce = rng.normal(size=(ns, 3))
vel0 = rng.normal(size=(nt, 3))                        # kernels accumulate into the target velocities
out = dict(mu=mu, tpos=tpos, spos=spos, srad=srad, trad=trad, sforce=sforce, vel0=vel0)
for kind, name in ((0, "stokes"), (1, "rpy"), (2, "rpyc")):
    out[name] = oracle.mobility_kernel(kind, mu, tpos, spos, srad, trad, sforce, vel0, lib=oracle.ref)
out["local_drag"] = oracle.local_drag(1, mu, srad, sforce, 0.5 * sforce, lib=oracle.ref)

# three spheres at the corners of an equilateral triangle (UnitTestPeriphery.cpp:671-775)
s_list = np.array([2.01, 2.10, 2.50, 3.00, 4.00, 6.00])
radius = 12.34
tri_rpy, tri_rpyc = [], []
for s in s_list:
    pos = np.array([[0.0, 0.0, 0.0], [-s * radius / 2.0, -np.sqrt(3.0) * s * radius / 2.0, 0.0],
                    [s * radius / 2.0, -np.sqrt(3.0) * s * radius / 2.0, 0.0]])
    f = np.zeros((3, 3))
    f[0, 1] = -6.0 * np.pi * radius
    rad = np.full(3, radius)
    self_term = f / (6.0 * np.pi * radius * 1.0)
    tri_rpy.append(oracle.mobility_kernel(1, 1.0, pos, pos, rad, rad, f, lib=oracle.ref) + self_term)
    tri_rpyc.append(oracle.mobility_kernel(2, 1.0, pos, pos, rad, rad, f, lib=oracle.ref) + self_term)
out.update(tri_s=s_list, tri_radius=radius, tri_rpy=np.array(tri_rpy), tri_rpyc=np.array(tri_rpyc))
np.savez_compressed(os.path.join(HERE, "mobility_ref.npz"), **out)
print("wrote mobility_ref.npz")
for s, a, b in zip(s_list, tri_rpy, tri_rpyc):
    print(s, "rpy U1,U2,U3 =", -a[0, 1], -a[1, 1], a[2, 0], "| rpyc", -b[0, 1], -b[1, 1], b[2, 0])
