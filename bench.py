#!/usr/bin/env python
"""bench.py — particle-timesteps/s of the collision hot path (BASELINE.json metric) on B200.

A "step" is one collision timestep of the reference's multi-step driver loop
(scrap/parameter_interface/NGPSpheresLCP.cpp:1135-1292): external ABC-flow forcing -> rebuild trigger / neighbour-link
rebuild -> narrow phase -> BBPGD LCP solve to tolerance -> Euler update, through the C-ABI of libmundy_b200.so.

Workload at one GPU = BASELINE.json configs[2]: 10^6 spherocylinders of aspect ratio 10 (r = 0.5, L = 9) uniformly random
in a cube at volume fraction 0.10, driver parameters mu = 1e-3, dt = 1e-4, search buffer = r_max, tol = 1e-3.

  value  particle-timesteps/s with the bodies resident in HBM (device entry point mb200_collision_step)
  e2e    the same through mb200_collision_step_host: pinned HOST arrays in, H2D + step + D2H inside the timed region
  roofline   the dominant kernel group = one BBPGD iteration (body pass + rate pass): SURVEY.md 8(d)'s algorithmic bytes
             (120 C + 240 N for rods) / CUDA-event time of the two launches in the timed region, against
             MEASURED_PEAKS.json hbm_gbs; `stages` carries the 8(d) bytes and fraction of every other stage
  cpu_baseline   the CPU oracle step (OpenMP restatement of the reference loop, "port", pinned bit for bit to the
             reference's own code compiled in oracle/_ref) on a bounded sample, all host threads

`--impl reference` times that CPU implementation alone on the SAME workload (rank 0 only under torchrun; bodies capped
by --cpu-cap for multi-GPU configurations so that the run ends within minutes).
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
    ap.add_argument("--workload", default="rods", choices=["rods", "mixed"],
                    help="rods: BASELINE.json configs[2]/[3] (the metric's workload); mixed: configs[4], a dense packing "
                         "of half spheres half spherocylinders at volume fraction 0.55, search buffer 0.25, no forcing")
    ap.add_argument("--filter", default="shape", choices=["shape", "none"],
                    help="search filter: shape-distance (default) or every bounding-sphere candidate")
    ap.add_argument("--prep", type=int, default=2, help="untimed steps that relax the random initial overlaps")
    ap.add_argument("--cpu-bodies", type=int, default=100_000, help="bodies of the in-line cpu_baseline sample")
    ap.add_argument("--cpu-cap", type=int, default=1_000_000,
                    help="--impl reference: at most this many bodies (the one-GPU workload); larger multi-GPU "
                         "configurations are sampled at this size, same density and parameters")
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
    use_all_host_threads()  # before the oracle (and with it libgomp) is loaded
    import oracle
    wl = load_workloads()
    use_all_host_threads()

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


# taken at import: with OMP_PROC_BIND set, libgomp pins the main thread to one place when it loads, after which the
# affinity mask of this process no longer says how many cores it may use
_HOST_THREADS = host_threads()


def use_all_host_threads():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU arm is meant to use every host core the
    process may run on. Sets the environment (read by libgomp when the oracle loads it) and, once it is loaded, the
    runtime itself."""
    import ctypes
    n = _HOST_THREADS
    os.environ["OMP_NUM_THREADS"] = str(n)
    os.environ.setdefault("OMP_PROC_BIND", "spread")  # the reference's own job template (mundy_brownian_*.run.in)
    os.environ.setdefault("OMP_PLACES", "threads")
    try:
        ctypes.CDLL("libgomp.so.1", mode=ctypes.RTLD_GLOBAL).omp_set_num_threads(n)
    except (OSError, AttributeError):
        pass
    return n


def omp_threads_in_use():
    import ctypes
    try:
        return int(ctypes.CDLL("libgomp.so.1").omp_get_max_threads())
    except (OSError, AttributeError):
        return 1


def load_workloads():
    """mundy_b200/workloads.py by path: importing the package would map libmundy_b200.so into the CPU arm."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("mb200_workloads", os.path.join(ROOT, "mundy_b200", "workloads.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    n_cfg = args.bodies * max(world, 1)          # the B200 arm's workload: `bodies` per GPU in one box
    n = min(n_cfg, args.cpu_cap)
    filt = args.filter == "shape"
    secs, info = cpu_steps(n, args.steps, max(args.warmup, 1), filt)
    value = n * args.steps / secs
    cores = omp_threads_in_use()
    sample = (f"{n} spherocylinders, {args.steps} steps after {max(args.warmup, 1)} warm-up steps, OpenMP oracle step on "
              f"{cores} threads" + ("" if n == n_cfg else f" (sample of the {n_cfg}-body configuration: same density and "
                                                          "parameters, bodies capped by --cpu-cap)"))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, args.bodies, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "lcp_iters_per_step": info["iters_per_step"], "links": info["links"],
        "same_config": n == n_cfg, "bodies": n,
    }
    emit(line)


def make_workload(wl, args, n):
    if args.workload == "mixed":
        return wl.mixed(n, phi=0.55)
    return wl.rods(n, phi=PHI, radius=RADIUS, length=LENGTH, seed=3)


def workload_config(args, n, world, exchange=None):
    if world == 1:
        par = "1 GPU"
    else:
        how = {"peer": "velocity halo stored into the peers' arenas by the body pass and solver sums reduced inside the "
                       "rate pass over peer memory (NVLink)",
               "nccl": "velocity halo by NCCL send/recv and solver sums by NCCL all-gather, per iteration"}.get(exchange)
        par = (f"{world} ranks, one box of {world}x{n} bodies, Hilbert-curve blocks, each rank holds its block + ghosts "
               "fetched from the neighbouring ranks at every rebuild" + (f"; {how}" if how else ""))
    what = (f"{n} spherocylinders/GPU, aspect ratio 10 (r=0.5, L=9), volume fraction {PHI}, uniform random "
            "(BASELINE.json configs[2])") if args.workload == "rods" else (
        f"{n} bodies/GPU, half spheres (r=1) half spherocylinders (r=0.5, L=4), volume fraction 0.55, uniform random: dense "
        "packing relaxed by the solver (BASELINE.json configs[4])")
    return {"workload": what,
            "bodies_per_gpu": n, "mu": MU, "dt": DT, "tol": TOL, "max_iters": MAX_ITERS,
            "search_buffer": BUFFER if args.workload == "rods" else 0.25,
            "search_filter": "shape_distance<=2*buffer" if args.filter == "shape" else "none (all bounding-sphere candidates)",
            "forcing": "attractive ABC flow" if args.workload == "rods" else "none", "l2": "inputs larger than L2 (contact arrays >= 250 MB per GPU)",
            "parallelism": par}


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
    if world > 1 or os.environ.get("MB200_BENCH_FORCE_DIST"):  # (the variable: the distributed code path on one GPU, a diagnostic)
        # weak scaling: world * n bodies in one box at the same density, ordered along the Hilbert curve; rank r owns the
        # r-th block of n bodies and holds ONLY that block (mundy_b200/dist.py): the library fetches the ghosts
        from mundy_b200 import dist as mbd
        w = make_workload(wl, args, n * world)
        w = mbd.reorder(w, mbd.hilbert_order(w["center"]))
        lo, hi = mbd.owned_range(n * world, world, rank)
        w = {k: (v[lo:hi] if isinstance(v, np.ndarray) and v.shape[:1] == (n * world,) else v) for k, v in w.items()}
        comm = mbd.Comm(dev)
    else:
        w = make_workload(wl, args, n)
    ntot = n * world
    buffer = BUFFER if args.workload == "rods" else 0.25
    forcing = args.workload == "rods"

    def to_dev(x, dtype=torch.float64):
        return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device=dev)

    c, q = to_dev(w["center"]), to_dev(w["quat"])
    r, L, kind = to_dev(w["radius"]), to_dev(w["length"]), to_dev(w["kind"], torch.int32)
    bodies = ops.make_bodies(c, r, q, L, kind)
    F = torch.zeros((n, 3), dtype=torch.float64, device=dev)
    links = ops.NeighborLinks().set_search_buffer(buffer)
    if filt:
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)

    log = []

    def step():
        F.zero_()
        if forcing:
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
        # runs the distributed step and reads the advanced block back
        hc, hq, hF = (x.cpu().pin_memory() for x in (c, q, F))
        e2e_steps = max(3, min(args.steps, 10))

        def host_step():
            for d_, h_ in ((c, hc), (q, hq), (F, hF)):
                d_.copy_(h_, non_blocking=True)
            mbd.step_dist(links, comm, bodies, MU, DT, TOL, MAX_ITERS, F_ext=F)
            hc.copy_(c, non_blocking=True)
            hq.copy_(q, non_blocking=True)
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
               "api": "per rank: pinned owned block -> H2D -> mb200_collision_step_dist -> D2H of the owned block "
                      "(bytes summed over ranks)"}
    elif not args.no_e2e:
        hc, hq = c.cpu().pin_memory(), q.cpu().pin_memory()
        hr, hL, hk = r.cpu().pin_memory(), L.cpu().pin_memory(), kind.cpu().pin_memory()
        hF = F.cpu().pin_memory()  # forcing sampled at the current positions, kept for the e2e steps
        hlinks = ops.NeighborLinks().set_search_buffer(buffer)
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
        # Algorithmic bytes per launch = SURVEY.md 8(d)'s per-unit figures (C contacts, N bodies, rods):
        #   rebuild 32 N + 64 N + 8 C | trigger 48 N | narrow phase 88 C + 72 N | Euler 72 N + 88 N
        #   one BBPGD iteration 120 C + 240 N, which the two kernels of an iteration share: the per-kernel entries carry
        #   the bytes this implementation's layout needs, the iteration carries 8(d)'s.
        nloc = mbd.local_counts(links)[0] if comm is not None else n  # bodies this rank works on (owned + ghosts)
        alg = {"broadphase": 96.0 * nloc + 8.0 * C, "trigger": 48.0 * nloc, "narrow": 88.0 * C + 72.0 * nloc,
               "euler": 160.0 * nloc,
               "lcp_rate": 88.0 * C + 64.0 * nloc,   # state 16 r + 16 w, {n, phi0} 32, sigma 16, ids 8; velocity records once
               "lcp_body": 32.0 * C + 176.0 * nloc}  # 8(d) iteration bytes minus the rate pass's
        # The solver enqueues iterations ahead of the convergence test; launches queued behind the converged one return
        # at once. They are in `calls` and in the sampled regions, so a plain average over the regions would be
        # diluted by them: the sampled time is attributed to the launches that did work (iterations + the initial
        # pass [+ the final body pass] of every solve), which charges the no-op launches' few microseconds to them.
        iters_total = sum(x[0] for x in tlog)
        executed = {"lcp_body": iters_total + 2 * len(tlog), "lcp_rate": iters_total + len(tlog)}
        per_launch_ms = {}
        for name, (tot_ms, cnt) in stages.items():
            calls_n = stage_calls.get(name, cnt)
            stage_out[name] = {"ms_total": round(tot_ms, 4), "regions": cnt, "calls": calls_n}
            if name in alg and cnt > 0:
                avg_ms = tot_ms / cnt
                if name in executed and 0 < executed[name] <= calls_n:
                    stage_out[name].update({"executed": executed[name], "noop_launches": calls_n - executed[name],
                                            "avg_us_all_launches": round(1e3 * avg_ms, 2)})
                    avg_ms = avg_ms * calls_n / executed[name]
                per_launch_ms[name] = avg_ms
                gbs = alg[name] / (avg_ms * 1e-3) / 1e9
                stage_out[name].update({"avg_us": round(1e3 * avg_ms, 2), "alg_bytes": alg[name],
                                        "gbs": round(gbs, 1), "frac": round(gbs / peak, 4)})
        if stages["lcp_rate"][1] > 0 and stages["lcp_body"][1] > 0:
            # one operator application = one body pass + one rate pass: the dominant kernel group of the step
            it_ms = per_launch_ms["lcp_rate"] + per_launch_ms["lcp_body"]
            it_bytes = 120.0 * C + 240.0 * nloc
            achieved = it_bytes / (it_ms * 1e-3) / 1e9
            calls = executed["lcp_rate"]
            traffic = None  # DRAM bytes per iteration (body + rate) from the committed ncu capture of this workload
            try:
                if n == 1_000_000 and args.filter == "shape":
                    tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
                    traffic = tj.get("lcp_body", 0) + tj.get("lcp_rate", 0) or None
            except (OSError, ValueError, TypeError):
                pass
            roofline = {"bound": "hbm", "kernel": "lcp_iteration (body_pass_kernel + rate_pass_kernel)",
                        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                        "alg_bytes": it_bytes, "alg_bytes_model": "SURVEY.md 8(d): 120 C + 240 N per rod iteration",
                        "peak_source": peak_src, "launches": 2 * calls,
                        "launches_timed": stages["lcp_rate"][1] + stages["lcp_body"][1], "avg_us": 1e3 * it_ms,
                        "share_of_step": it_ms * calls / ms,
                        "body_pass_us": 1e3 * per_launch_ms["lcp_body"],
                        "rate_pass_us": 1e3 * per_launch_ms["lcp_rate"],
                        "timing": f"CUDA events around every {max(1, args.stage_sample)}-th launch inside the timed region; "
                                  "sampled time attributed to the launches that did work (no-op launches queued behind "
                                  "the converged iteration excluded from the count)"}

            if roofline["frac"] > 1.0:
                roofline["note"] = ("SURVEY 8(d)'s bytes charge every contact to every pass; the body pass visits only the "
                                    "contacts whose activity bit is set, so with few active contacts the kernels move "
                                    "fewer bytes than the model and the fraction exceeds 1: see stages.lcp_rate (which "
                                    "streams every contact) for a bandwidth figure")

    cpu = None
    if not args.no_cpu_baseline:
        secs, info = cpu_steps(args.cpu_bodies, args.cpu_steps, 1, filt)
        cpu = {"value": args.cpu_bodies * args.cpu_steps / secs, "unit": UNIT, "cores": omp_threads_in_use(), "kind": "port",
               "sample": f"{args.cpu_bodies} spherocylinders at the same density/parameters, {args.cpu_steps} steps after 1 "
                         f"warm-up step; OpenMP oracle step (LCP iterations/step {info['iters_per_step']})"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, n, world, comm.exchange_mode if comm is not None else None),
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "lcp_iters_per_step": [x[0] for x in tlog], "ms_steps": ms_steps, "links_per_step": int(C), "rebuilds": sum(x[2] for x in tlog),
        "all_converged": all(x[3] for x in tlog), "stages": stage_out, "prep_steps": args.prep,
        "lcp_iters_mean": statistics.mean(x[0] for x in tlog),
        "us_per_lcp_iteration": (roofline or {}).get("avg_us"),
    }
    if args.filter != "shape":
        line["metric"] = METRIC + " (search_filter: none, every bounding-sphere candidate is a constraint)"
    if args.workload != "rods":
        line["metric"] = "particle-timesteps/s, dense mixed sphere/spherocylinder packing (BASELINE.json configs[4])"
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
