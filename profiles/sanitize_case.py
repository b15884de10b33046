#!/usr/bin/env python
"""Small collision steps for compute-sanitizer: the <= 2*10^4-body twins of cfg1 (monodisperse spheres),
cfg2 (polydisperse spheres, periodic box), cfg3 (spherocylinders, shape filter) and cfg5 (mixed, dense), three steps each
through mb200_collision_step, plus one solve through each caller-layout entry point.
usage (one tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|initcheck python profiles/sanitize_case.py [n]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mundy_b200 import ops, workloads as wl  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000


def dev(x, dt=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dt, device="cuda")


cases = [("cfg1 spheres", wl.monodisperse_spheres(n), 3.0, False, False),
         ("cfg2 periodic spheres", wl.polydisperse_spheres(n, phi=0.3), 0.5, False, True),
         ("cfg3 rods", wl.rods(n, phi=0.1), 0.5, True, False),
         ("cfg5 mixed", wl.mixed(n, phi=0.45), 0.4, True, False)]
for name, w, buffer, filt, periodic in cases:
    rods = "kind" in w
    c, r = dev(w["center"]), dev(w["radius"])
    q = dev(w["quat"]) if rods else None
    b = ops.make_bodies(c, r, q, dev(w["length"]) if rods else None, dev(w["kind"], torch.int32) if rods else None)
    links = ops.NeighborLinks().set_search_buffer(buffer)
    if filt:
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    if periodic:
        links.set_periodic([w["edge"]] * 3)
    F = torch.zeros((len(w["radius"]), 3), dtype=torch.float64, device="cuda")
    for s in range(3):
        F.zero_()
        ops.abc_flow_force(c, F)
        st = links.step(b, 1e-3, 1e-4, 1e-3, 200, F_ext=F)
    torch.cuda.synchronize()
    print(f"{name}: n={len(w['radius'])} links={st.num_links} iterations={st.lcp.ite_count} residual={st.lcp.max_abs_projected_sep:.3e}")
# caller-layout solver entry points (general arm vectors, axial arms) on the last case's geometry
pairs = links.pairs()
phi, nrm, ai, aj = ops.narrow_phase(b, pairs, arms=True)
op = ops.make_contact_op(len(w["radius"]), pairs, nrm, r, 1e-3, 1e-4, ai, aj, q, dev(w["kind"], torch.int32))
g = torch.zeros(len(pairs), dtype=torch.float64, device="cuda")
res, *_ = ops.resolve_collisions(op, phi.clone(), g, 1e-3, 50)
torch.cuda.synchronize()
print("resolve_collisions (general arms): iterations", res.ite_count)
