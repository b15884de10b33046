#!/usr/bin/env python
"""bench.py — particle-timesteps/s of the collision hot path (BASELINE.json metric) on B200.

A "step" is one collision timestep of the reference's multi-step driver loop
(scrap/parameter_interface/NGPSpheresLCP.cpp:1135-1292): external ABC-flow forcing -> rebuild trigger / neighbour-link
rebuild -> narrow phase -> BBPGD LCP solve to tolerance -> Euler update, through the C-ABI of libmundy_b200.so.

Workload at one GPU = BASELINE.json configs[2]: 10^6 spherocylinders of aspect ratio 10 (r = 0.5, L = 9) uniformly random
in a cube at volume fraction 0.10, driver parameters mu = 1e-3, dt = 1e-4, search buffer = r_max, tol = 1e-3.

  value  particle-timesteps/s with the bodies resident in HBM (device entry point mb200_collision_step)
  e2e    the same through mb200_collision_step_host: pinned HOST arrays in, H2D + step + D2H inside the timed region
  roofline   dominant kernel (the BBPGD rate pass): algorithmic bytes / CUDA-event time of its launches in the timed
             region, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline   the CPU oracle step (OpenMP restatement of the reference loop, "port") on a bounded sample

`--impl reference` times that CPU implementation alone (rank 0 only under torchrun).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "particle-timesteps/s, 1M-spherocylinder LCP collision step"
UNIT = "particle-timesteps/s"
# driver parameters (scrap/parameter_interface/NGPSpheresLCP.cpp:1034-1046: dt = 1e-4, buffer = r_max, tol = 1e-3)
MU, DT, TOL, MAX_ITERS = 1e-3, 1e-4, 1e-3, 10000
RADIUS, LENGTH, PHI = 0.5, 9.0, 0.10
BUFFER = RADIUS


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--bodies", type=int, default=1_000_000, help="spherocylinders per GPU")
    ap.add_argument("--filter", default="shape", choices=["shape", "none"],
                    help="search filter: shape-distance (default) or every bounding-sphere candidate")
    ap.add_argument("--prep", type=int, default=2, help="untimed steps that relax the random initial overlaps")
    ap.add_argument("--cpu-bodies", type=int, default=100_000, help="bodies of the CPU sample")
    ap.add_argument("--cpu-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-stage-timers", action="store_true")
    ap.add_argument("--stage-sample", type=int, default=8,
                    help="bracket every k-th launch of the per-iteration LCP kernels with CUDA events (1 = every launch)")
    return ap.parse_args()


# ---- clocks ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    # Started BEFORE the warm-up steps (NVML start-up stalls the driver for tens of ms and would otherwise land inside the
    # timed region); only samples stamped after mark() are used.
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.t0 = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                                       "50", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def wait_ready(self, timeout=8.0):
        """Blocks until nvidia-smi has written its first sample: its NVML start-up (process start + driver attach,
        200-400 ms, during which CUDA calls of this process can stall for tens of ms) is then over. Observed without
        this wait: a +24..58 ms hiccup in the 4th timed step of some runs."""
        if self.p is None:
            return
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < timeout and self.p.poll() is None:
            if os.path.getsize(self.f.name) > 0:
                return
            time.sleep(0.02)

    def mark(self):
        import datetime
        self.t0 = datetime.datetime.now()

    def stop(self):
        import datetime
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 8:
                continue
            try:
                if self.t0 is not None and datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f") < self.t0:
                    continue
                sm.append(float(c[1])), smax.append(float(c[2])), power.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(names, c[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "reasons": sorted(reasons),
                   "power_w_max": max(power), "samples": len(sm)}
        return out


# ---- CPU arm (the oracle; test infrastructure used here only as the timed baseline) -----------------------------------
def cpu_steps(n, steps, warmup, filt, seed=3):
    """Times `steps` oracle collision steps on n bodies after `warmup` untimed ones. Returns (seconds, info)."""
    import numpy as np  # noqa: F401
    import oracle
    from mundy_b200 import workloads as wl

    w = wl.rods(n, phi=PHI, radius=RADIUS, length=LENGTH, seed=seed)
    state, links = (w["center"].copy(), w["quat"].copy()), None
    iters = []

    def one():
        nonlocal state, links
        F = oracle.abc_flow_force(state[0])
        state, links, out, _ = oracle.collision_step(state, w, links, BUFFER, filt, MU, DT, TOL, MAX_ITERS, -1, F)
        return out

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(steps):
        iters.append(one()["it"])
    dt = time.perf_counter() - t0
    return dt, {"links": int(len(links[0])), "iters_per_step": iters}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.cpu_bodies
    filt = args.filter == "shape"
    secs, info = cpu_steps(n, args.steps, max(args.warmup, 1), filt)
    value = n * args.steps / secs
    cores = int(os.environ.get("OMP_NUM_THREADS", host_threads()))
    sample = (f"{n} spherocylinders (same density/parameters as the 10^6 workload), {args.steps} steps after "
              f"{max(args.warmup, 1)} warm-up steps, OpenMP oracle step")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, n),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "lcp_iters_per_step": info["iters_per_step"], "links": info["links"],
    }
    emit(line)


def workload_config(args, n):
    return {"workload": f"{n} spherocylinders/GPU, aspect ratio 10 (r=0.5, L=9), volume fraction {PHI}, uniform random "
                        "(BASELINE.json configs[2])",
            "bodies_per_gpu": n, "mu": MU, "dt": DT, "tol": TOL, "max_iters": MAX_ITERS, "search_buffer": BUFFER,
            "search_filter": "shape_distance<=2*buffer" if args.filter == "shape" else "none (all bounding-sphere candidates)",
            "forcing": "attractive ABC flow", "l2": "inputs larger than L2 (contact arrays >= 400 MB per GPU)",
            "parallelism": "1 GPU" if args.gpus == 1 else
            f"{args.gpus} ranks, one box of {args.gpus}x{n} bodies, Hilbert-curve blocks, ghost velocity halo (NCCL send/recv) "
            "+ all-gathered solver sums per iteration, state all-gather per step"}


# ---- B200 arm ---------------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py --impl b200 needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from mundy_b200 import ops, workloads as wl

    n = args.bodies
    filt = args.filter == "shape"
    dev = torch.device("cuda", local)
    comm = None
    if world > 1:
        # weak scaling: world * n bodies in one box at the same density, ordered along the Hilbert curve; every rank
        # builds the same replica (same seed) and owns one contiguous block of n bodies (mundy_b200/dist.py)
        from mundy_b200 import dist as mbd
        w = wl.rods(n * world, phi=PHI, radius=RADIUS, length=LENGTH, seed=3)
        w = mbd.reorder(w, mbd.hilbert_order(w["center"]))
        comm = mbd.Comm(dev)
    else:
        w = wl.rods(n, phi=PHI, radius=RADIUS, length=LENGTH, seed=3)
    ntot = len(w["radius"])

    def to_dev(x, dtype=torch.float64):
        return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device=dev)

    c, q = to_dev(w["center"]), to_dev(w["quat"])
    r, L, kind = to_dev(w["radius"]), to_dev(w["length"]), to_dev(w["kind"], torch.int32)
    bodies = ops.make_bodies(c, r, q, L, kind)
    F = torch.zeros((ntot, 3), dtype=torch.float64, device=dev)
    links = ops.NeighborLinks().set_search_buffer(BUFFER)
    if filt:
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)

    log = []

    def step():
        F.zero_()
        ops.abc_flow_force(c, F)
        if comm is None:
            st = links.step(bodies, MU, DT, TOL, MAX_ITERS, F_ext=F)
        else:
            st = mbd.step_dist(links, comm, bodies, MU, DT, TOL, MAX_ITERS, F_ext=F)
        log.append((st.lcp.ite_count, st.num_links, st.rebuilt, st.lcp.converged))
        if os.environ.get("MB200_BENCH_VERBOSE"):
            print(f"[rank {rank}] step {len(log)}: iters {st.lcp.ite_count} links {st.num_links} rebuilt {st.rebuilt} "
                  f"res {st.lcp.max_abs_projected_sep:.3e} t {time.perf_counter():.1f}", file=sys.stderr, flush=True)
        return st

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.prep):
        step()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.wait_ready()
    barrier()
    for _ in range(args.warmup):
        step()
    log.clear()

    # ---- timed region: device-resident -----------------------------------------------------------------------------
    if not args.no_stage_timers:
        ops.profile_enable(max(1, args.stage_sample))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = ops.kernel_launch_count()
    barrier()
    if sampler:
        sampler.mark()
    ev0.record()
    marks = [ev0]
    for _ in range(args.steps):
        step()
        marks.append(torch.cuda.Event(enable_timing=True))
        marks[-1].record()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    ms_steps = [round(marks[k].elapsed_time(marks[k + 1]), 3) for k in range(args.steps)]
    launches = ops.kernel_launch_count() - launches0
    clocks = sampler.stop() if sampler else None
    stages = ops.profile_read() if not args.no_stage_timers else {}
    stage_calls = ops.profile_calls() if not args.no_stage_timers else {}
    ops.profile_enable(False)
    tlog = list(log)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = ntot * args.steps / (ms * 1e-3)

    # ---- e2e: host arrays through mb200_collision_step_host ---------------------------------------------------------
    e2e = None
    if not args.no_e2e and comm is not None:
        # every rank keeps its OWNED block in pinned host memory: each step uploads it (centres, quaternions, forcing),
        # all-gathers the blocks into the device replicas, runs the distributed step and reads the owned block back
        o0 = rank * n
        hc, hq, hF = (x[o0:o0 + n].cpu().pin_memory() for x in (c, q, F))
        dc, dq, dF = (torch.empty_like(x[o0:o0 + n]) for x in (c, q, F))
        e2e_steps = max(3, min(args.steps, 10))

        def host_step():
            for d_, h_, g_ in ((dc, hc, c), (dq, hq, q), (dF, hF, F)):
                d_.copy_(h_, non_blocking=True)
                dist.all_gather_into_tensor(g_, d_)
            mbd.step_dist(links, comm, bodies, MU, DT, TOL, MAX_ITERS, F_ext=F)
            hc.copy_(c[o0:o0 + n], non_blocking=True)
            hq.copy_(q[o0:o0 + n], non_blocking=True)
            torch.cuda.synchronize()

        host_step()
        barrier()
        ev0.record()
        for _ in range(e2e_steps):
            host_step()
        ev1.record()
        barrier()
        t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e = float(t.item())
        e2e = {"value": ntot * e2e_steps / (ms_e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": (3 + 4 + 3) * 8 * ntot,
               "d2h_bytes_per_step": (3 + 4) * 8 * ntot, "steps": e2e_steps, "ms_per_step": ms_e / e2e_steps,
               "api": "per rank: pinned owned block -> H2D -> all-gather into the replicas -> mb200_collision_step_dist "
                      "-> D2H of the owned block (bytes summed over ranks)"}
    elif not args.no_e2e:
        hc, hq = c.cpu().pin_memory(), q.cpu().pin_memory()
        hr, hL, hk = r.cpu().pin_memory(), L.cpu().pin_memory(), kind.cpu().pin_memory()
        hF = F.cpu().pin_memory()  # forcing sampled at the current positions, kept for the e2e steps
        hlinks = ops.NeighborLinks().set_search_buffer(BUFFER)
        if filt:
            hlinks.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
        e2e_steps = max(3, min(args.steps, 10))
        hlinks.step_host(hc, hr, hq, hL, hk, MU, DT, TOL, MAX_ITERS, F_ext=hF)  # warm-up (first build, allocations)
        barrier()
        t0 = time.perf_counter()
        ev0.record()
        for _ in range(e2e_steps):
            hlinks.step_host(hc, hr, hq, hL, hk, MU, DT, TOL, MAX_ITERS, F_ext=hF)
        ev1.record()
        barrier()
        ms_e = max(ev0.elapsed_time(ev1), 1e3 * (time.perf_counter() - t0))  # device events vs host wall clock
        t = torch.tensor([ms_e], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e = float(t.item())
        h2d = (3 + 4 + 1 + 1 + 3) * 8 * n + 4 * n
        d2h = (3 + 4) * 8 * n
        e2e = {"value": n * world * e2e_steps / (ms_e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": e2e_steps, "ms_per_step": ms_e / e2e_steps,
               "api": "mb200_collision_step_host (pinned host arrays in/out)"}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ---------------------------------------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "6650 GB/s (of fallback)"
    C = statistics.mean(x[1] for x in tlog)
    roofline = None
    stage_out = {}
    if stages:
        # algorithmic bytes per launch (DESIGN.md "Kernels"): C contacts, n bodies, 6-DOF
        nloc = mbd.local_counts(links)[0] if comm is not None else n  # bodies this rank works on (owned + ghosts)
        alg = {"lcp_rate": 128.0 * C + 64.0 * nloc, "lcp_body": 120.0 * C + 104.0 * nloc}
        for name, (tot_ms, cnt) in stages.items():
            stage_out[name] = {"ms_total": round(tot_ms, 4), "regions": cnt, "calls": stage_calls.get(name, cnt)}
            if name in alg and cnt > 0:
                gbs = alg[name] / (tot_ms / cnt * 1e-3) / 1e9
                stage_out[name].update({"alg_bytes": alg[name], "gbs": round(gbs, 1), "frac": round(gbs / peak, 4)})
        dom = max(("lcp_rate", "lcp_body"), key=lambda k: stages[k][0])
        tot_ms, cnt = stages[dom]
        if cnt > 0:
            achieved = alg[dom] / (tot_ms / cnt * 1e-3) / 1e9
            traffic = None  # DRAM bytes per launch from the committed ncu capture (same workload), if there is one
            try:
                if n == 1_000_000 and args.filter == "shape":
                    traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(dom)
            except (OSError, ValueError):
                pass
            roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                        "launches": stage_calls.get(dom, cnt), "launches_timed": cnt, "avg_us": 1e3 * tot_ms / cnt,
                        "share_of_step": (tot_ms / cnt) * stage_calls.get(dom, cnt) / ms,
                        "timing": f"CUDA events around every {max(1, args.stage_sample)}-th launch inside the timed region"}
            if stages["lcp_rate"][1] > 0 and stages["lcp_body"][1] > 0:
                # one operator application = one body pass + one contact pass
                it_ms = sum(stages[k][0] / stages[k][1] for k in ("lcp_rate", "lcp_body"))
                it_bytes = 120.0 * C + 240.0 * nloc  # fused-ideal bytes of one rod BBPGD iteration
                roofline["lcp_iteration"] = {"alg_bytes": it_bytes, "avg_us": 1e3 * it_ms,
                                             "gbs": it_bytes / (it_ms * 1e-3) / 1e9,
                                             "frac": it_bytes / (it_ms * 1e-3) / 1e9 / peak}

    cpu = None
    if not args.no_cpu_baseline:
        secs, info = cpu_steps(args.cpu_bodies, args.cpu_steps, 1, filt)
        cpu = {"value": args.cpu_bodies * args.cpu_steps / secs, "unit": UNIT,
               "cores": int(os.environ.get("OMP_NUM_THREADS", host_threads())), "kind": "port",
               "sample": f"{args.cpu_bodies} spherocylinders at the same density/parameters, {args.cpu_steps} steps after 1 "
                         f"warm-up step; OpenMP oracle step (LCP iterations/step {info['iters_per_step']})"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(args, n),
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "lcp_iters_per_step": [x[0] for x in tlog], "ms_steps": ms_steps, "links_per_step": int(C), "rebuilds": sum(x[2] for x in tlog),
        "all_converged": all(x[3] for x in tlog), "stages": stage_out, "prep_steps": args.prep,
    }
    if comm is not None:
        nl, ng = mbd.local_counts(links)
        line["dist"] = {"rank0_local_bodies": nl, "rank0_ghost_rows": ng, "owned_per_rank": n,
                        "exchange": comm.exchange_mode,
                        "note": "links_per_step, stages and roofline are rank 0's"}
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else that writes to fd 1 (NCCL prints its
    version banner there from C) was redirected to stderr in main()."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    args = parse_args()
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
