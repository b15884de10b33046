"""Generates tests/golden/sfc_ref.npz: outputs of the REFERENCE's space-filling-curve code through
oracle/_ref/libmundy_ref.so :: ref_hilbert_3d, ref_zorder_less (oracle/ref_sfc.cpp):
math::hilbert_3d visiting orders for 2^3, 4^3, 8^3, 16^3 lattices (mundy/math/src/mundy_math/Hilbert.hpp:48-85) and the two
z-order comparators of mundy/math/src/mundy_math/zmort.hpp (zorder_knn::Less :195-220, math::zmorton_less :228-266) on
seeded integer lattice points and on signed floating-point points.
Run in the build container (needs /root/reference): python tests/golden/make_golden_sfc.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle  # noqa: E402

if __name__ == "__main__":
    assert oracle.ref is not None, "oracle/_ref/libmundy_ref.so is needed (build with /root/reference present)"
    out = {}
    for s in (2, 4, 8, 16):
        out[f"hilbert_{s}"] = oracle.ref_hilbert_3d(s).astype(np.int16)
    rng = np.random.default_rng(20260505)
    n = 4000
    a, b = rng.integers(0, 1024, (n, 3)), rng.integers(0, 1024, (n, 3))
    b[:200] = a[:200]                      # equal points
    b[200:600, 0] = a[200:600, 0]          # equal x
    a[600:1600], b[600:1600] = rng.integers(0, 4, (1000, 3)), rng.integers(0, 4, (1000, 3))  # many ties of the top bit
    fa, fb = rng.normal(size=(n, 3)) * 10.0, rng.normal(size=(n, 3)) * 10.0                    # signed floats
    out.update(int_p=a.astype(np.int16), int_q=b.astype(np.int16), float_p=fa, float_q=fb)
    for which, name in ((0, "knn_less"), (1, "zmorton_less")):
        out[f"int_{name}"] = oracle.ref_zorder_less(a, b, which)
        out[f"float_{name}"] = oracle.ref_zorder_less(fa, fb, which)
    path = os.path.join(HERE, "sfc_ref.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")
