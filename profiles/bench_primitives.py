"""Micro-benchmark of the remaining-primitive kernels on one GPU: CUDA events on the launching
stream after warm-up, inputs larger than L2 for the closed-form kernels. Closed-form kernels report algorithmic GB/s
against MEASURED_PEAKS.json hbm_gbs; the minimiser-based ones (fp64-compute-bound) report pairs/s with the CPU oracle
(all host threads, bounded sample) beside them. One JSON line per kernel. Run on the GPU box:
python profiles/bench_primitives.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from mundy_b200 import ops  # noqa: E402
import oracle  # noqa: E402
from make_golden_primitives import make_inputs  # noqa: E402

FN = {"aabb_point": ops.compute_aabb_point, "aabb_line_segment": ops.compute_aabb_line_segment,
      "aabb_ellipsoid": ops.compute_aabb_ellipsoid, "point_sphere": ops.distance_point_sphere,
      "point_line": ops.distance_point_line, "line_line": ops.distance_line_line,
      "line_sphere": ops.distance_line_sphere, "point_ellipsoid": ops.distance_point_ellipsoid,
      "ellipsoid_ellipsoid": ops.distance_ellipsoid_ellipsoid, "circle_circle": ops.distance_circle3d_circle3d}
MINIMISED = ("point_ellipsoid", "ellipsoid_ellipsoid", "circle_circle")


def gpu_ms(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    peak = 6549.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for name, fn in FN.items():
        mini = name in MINIMISED
        n = 262_144 if mini else 8_000_000
        ins = make_inputs(seed=11, n=n)[name]
        d = [torch.as_tensor(np.ascontiguousarray(x), dtype=torch.float64, device="cuda") for x in ins]
        # the wrapper allocates its outputs per call from torch's caching allocator (no cudaMalloc after warm-up)
        ms = gpu_ms(lambda: fn(*d), 3 if mini else 10)
        in_bytes = sum(x.shape[1] if x.ndim == 2 else 1 for x in ins) * 8
        out_bytes = sum(oracle.PRIMITIVE_OUTPUTS[name]) * 8
        line = {"kernel": name, "n": n, "ms": round(ms, 4), "items_per_s": n / (ms * 1e-3),
                "alg_bytes_per_item": in_bytes + out_bytes}
        if mini:
            ncpu = 20_000
            t0 = time.perf_counter()
            oracle.primitive(name, *[x[:ncpu] for x in ins])
            cpu_s = time.perf_counter() - t0
            line.update({"bound": "fp64 compute", "cpu_items_per_s": ncpu / cpu_s, "cpu_threads": os.cpu_count(),
                         "cpu_sample": f"{ncpu} items, oracle port, OpenMP"})
        else:
            gbs = n * (in_bytes + out_bytes) / (ms * 1e-3) / 1e9
            line.update({"bound": "hbm", "gbs": round(gbs, 1), "peak_gbs": peak, "frac": round(gbs / peak, 4)})
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
