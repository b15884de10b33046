"""Generates tests/golden/lcp_ref.npz: outputs of the reference's OWN driver-LCP code
(scrap/lcp_spheres/KokkosLCP.cpp:449-876, compiled into oracle/_ref/libmundy_ref.so by oracle/Makefile) on two seeded
monodisperse-sphere problems with the driver's parameters (KokkosLCP.cpp:963-969: mu = 1e-3, dt = 5e-3, r = 1,
search buffer 3 r, tol = 1e-5). Run in the build container (needs /root/reference for the _ref build):
    python tests/golden/make_golden_lcp.py
The small problem is stored whole (inputs and outputs); the cfg1-sized one (10^4 spheres) is regenerated from its seed by
the test and only its outputs' digests and a strided sample are stored (the file stays small).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402

MU, DT, TOL, RADIUS, BUFFER = 1e-3, 5e-3, 1e-5, 1.0, 3.0
KS = (0, 1, 2, 5, 12, 150)


def problem(n, seed=1234, phi=0.30):
    rng = np.random.default_rng(seed)
    edge = (n * 4.0 / 3.0 * np.pi * RADIUS ** 3 / phi) ** (1.0 / 3.0)
    pos = rng.uniform(0.0, edge, size=(n, 3))
    # every pair of spheres whose search spheres (r + buffer) touch: the GenNeighborLinks form of SURVEY 8a-B2
    ids = oracle.search(pos, np.full(n, RADIUS + BUFFER))
    return pos, ids


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    assert oracle.ref is not None, "oracle/_ref/libmundy_ref.so missing (needs /root/reference)"
    out = {}
    # small: stored whole
    pos, ids = problem(150)
    sep, normal = oracle.ref_lcp_narrow(ids, RADIUS, pos)
    out.update(s_pos=pos, s_ids=ids, s_sep=sep, s_normal=normal)
    rng = np.random.default_rng(7)
    gam = np.abs(rng.normal(size=len(ids)))
    F = oracle.ref_lcp_D(len(pos), ids, normal, gam)
    U = oracle.ref_lcp_M(RADIUS, MU, F)
    gd = oracle.ref_lcp_DT(len(pos), ids, normal, U)
    out.update(s_gamma=gam, s_F=F, s_U=U, s_gdot=gd, s_res=oracle.ref_lcp_residual(gam, sep, gd, DT),
               s_dots=oracle.ref_lcp_diff_dots(gam, 0.5 * gam, gd, 0.25 * gd, DT),
               s_gamma_next=oracle.ref_lcp_update_gammas(gam, sep, gd, DT, 0.37))
    for K in KS:
        r = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, 0.0, K, pos, sep, normal)
        out[f"s_K{K}_gamma"], out[f"s_K{K}_F"], out[f"s_K{K}_U"] = r["gamma"], r["F"], r["U"]
        out[f"s_K{K}_meta"] = np.array([r["res"], r["it"], r["max_disp"]])
    r = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, TOL, 10000, pos, sep, normal)
    out["s_conv_gamma"], out["s_conv_meta"] = r["gamma"], np.array([r["res"], r["it"], r["max_disp"]])
    # cfg1: 10^4 spheres, digests + strided samples
    pos, ids = problem(10000)
    sep, normal = oracle.ref_lcp_narrow(ids, RADIUS, pos)
    out["c_count"] = np.array([len(ids)])
    out["c_digest_inputs"] = np.array([digest(pos), digest(ids), digest(sep), digest(normal)])
    for K in KS:
        r = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, 0.0, K, pos, sep, normal)
        out[f"c_K{K}_digest"] = np.array([digest(r["gamma"]), digest(r["F"] + 0.0), digest(r["U"] + 0.0)])
        out[f"c_K{K}_gamma_sample"] = r["gamma"][::97]
        out[f"c_K{K}_meta"] = np.array([r["res"], r["it"], r["max_disp"]])
    r = oracle.ref_lcp_resolve(ids, RADIUS, MU, DT, TOL, 10000, pos, sep, normal)
    out["c_conv_gamma_sample"], out["c_conv_meta"] = r["gamma"][::97], np.array([r["res"], r["it"], r["max_disp"]])
    np.savez_compressed(os.path.join(HERE, "lcp_ref.npz"), **out)
    print("wrote lcp_ref.npz:", {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items() if "meta" in k})
    for k in out:
        if "meta" in k:
            print(k, out[k])


if __name__ == "__main__":
    main()
