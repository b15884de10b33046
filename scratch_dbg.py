import sys, os
sys.path.insert(0, "tests")
import numpy as np, torch
import oracle
from mundy_b200 import ops, workloads as wl
def dev(x, dtype=torch.float64): return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")
def host(t): return t.detach().cpu().numpy()
for name, maker, kw, buffer, filt in [("spheres", wl.polydisperse_spheres, dict(n=8000, phi=0.25), 0.5, False),
                                      ("rods_shape_filter", wl.rods, dict(n=5000, phi=0.1), 0.4, True)]:
    w = maker(**kw); n = len(w["radius"]); mu, dt, tol, fixed = 1e-3, 1e-4, 1e-5, 8
    rods = "kind" in w
    c, r = dev(w["center"]), dev(w["radius"]); q = dev(w["quat"]) if rods else None
    b = ops.make_bodies(c, r, q, dev(w["length"]) if rods else None, dev(w["kind"], torch.int32) if rods else None)
    links = ops.NeighborLinks().set_search_buffer(buffer)
    if filt: links.set_search_filter(1)
    state, olinks = (w["center"].copy(), w["quat"].copy() if rods else None), None
    for step in range(3):
        F_ext = oracle.abc_flow_force(state[0])
        c_before = host(c).copy()
        stats = links.step(b, mu, dt, tol, 10000, fixed_iters=fixed, F_ext=dev(F_ext))
        prev = state[0]
        state, olinks, out, U = oracle.collision_step(state, w, olinks, buffer, filt, mu, dt, tol, 10000, fixed, F_ext)
        gp = host(links.pairs())
        print(name, step, "gpu links", stats.num_links, "rebuilt", stats.rebuilt, "| oracle", len(olinks[0]), out["rebuilt"],
              "| max step disp", np.max(np.linalg.norm(state[0]-prev,axis=1)), "pos err", np.max(np.abs(host(c)-state[0])),
              "pairs equal", gp.shape == olinks[0].shape and np.array_equal(gp, olinks[0]))
