"""Micro-benchmark of the non-local mobility kernels: pair interactions per second of the all-pairs and
the neighbour-list forms on one GPU (CUDA events on the launching stream, after warm-up), with the CPU oracle timed on a
bounded sample beside it. Prints one JSON line per case. Run on the GPU box: python profiles/bench_mobility.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from mundy_b200 import ops  # noqa: E402
import oracle  # noqa: E402

NAMES = {0: "stokes", 1: "rpy", 2: "rpyc", 3: "rpy_exact"}


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def gpu_ms(fn, reps):
    for _ in range(2):
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
    rng = np.random.default_rng(0)
    n = 65536
    pos = rng.uniform(0, 80, (n, 3))
    rad = rng.uniform(0.3, 0.7, n)
    f = rng.normal(size=(n, 3))
    dp, dr, df = dev(pos), dev(rad), dev(f)
    v = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    ncpu = 4096
    for kind in (0, 1, 2, 3):
        ms = gpu_ms(lambda: ops.mobility_kernel_apply(kind, 1.0, dp, dp, dr, dr, df, v), 5)
        t0 = time.perf_counter()
        oracle.mobility_kernel(kind, 1.0, pos[:ncpu], pos, rad, rad[:ncpu], f)
        cpu_s = time.perf_counter() - t0
        print(json.dumps({"case": "all_pairs", "kernel": NAMES[kind], "n": n, "ms": round(ms, 3),
                          "gpu_pairs_per_s": n * n / (ms * 1e-3), "cpu_pairs_per_s": ncpu * n / cpu_s,
                          "cpu_threads": os.cpu_count(), "cpu_sample": f"{ncpu} targets x {n} sources"}))
    # neighbour-list form: 10^6 spheres, ~27 neighbours each (uniform random lists: worst-case locality)
    n, deg = 1_000_000, 27
    pos = rng.uniform(0, 200, (n, 3))
    rad = rng.uniform(0.3, 0.7, n)
    f = rng.normal(size=(n, 3))
    offsets = (np.arange(n + 1, dtype=np.int64) * deg).astype(np.int32)
    local = (np.repeat(np.arange(n), deg) + rng.integers(-2000, 2000, n * deg)) % n  # Morton-like locality
    indices = local.astype(np.int32)
    dp, dr, df = dev(pos), dev(rad), dev(f)
    do, di = dev(offsets, torch.int32), dev(indices, torch.int32)
    v = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    for kind in (1, 3):
        ms = gpu_ms(lambda: ops.mobility_kernel_apply_neighbors(kind, 1.0, dp, dr, do, di, df, v), 10)
        t0 = time.perf_counter()
        oracle.mobility_kernel_neighbors(kind, 1.0, pos, rad, offsets, indices, f)
        cpu_s = time.perf_counter() - t0
        print(json.dumps({"case": "neighbors", "kernel": NAMES[kind], "n": n, "pairs": n * deg, "ms": round(ms, 3),
                          "gpu_pairs_per_s": n * deg / (ms * 1e-3), "cpu_pairs_per_s": n * deg / cpu_s,
                          "cpu_threads": os.cpu_count(),
                          "gather_GBs": n * deg * 64 / (ms * 1e-3) / 1e9}))


if __name__ == "__main__":
    main()
