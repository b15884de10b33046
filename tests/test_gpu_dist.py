"""GPU tests of the multi-GPU step (every rank passes only the bodies it owns). Parity target = the single-GPU step on the
same bodies (the reference's own MPI run never updates ghost fields, F8). world = 1 runs on any GPU box;
world = 2 / 4 / 8 need that many GPUs (gpurun --gpus N; the logs of those runs are kept under profiles/)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from mundy_b200 import workloads as wl

pytestmark = pytest.mark.gpu
MU, DT, TOL, BUF = 1e-3, 1e-4, 1e-5, 0.4


def _dev(x, device, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device=device)


def _workload(kind, n):
    from mundy_b200 import dist as mbd
    w = wl.rods(n, phi=0.1) if kind == "rods" else wl.polydisperse_spheres(n, phi=0.25)
    return mbd.reorder(w, mbd.hilbert_order(w["center"]))


def _bodies(w, device):
    from mundy_b200 import ops
    rods = "kind" in w
    c, r = _dev(w["center"], device), _dev(w["radius"], device)
    q = _dev(w["quat"], device) if rods else None
    b = ops.make_bodies(c, r, q, _dev(w["length"], device) if rods else None,
                        _dev(w["kind"], device, torch.int32) if rods else None)
    return b, c, q


def _run(kind, n, steps, fixed, device, comm=None, periodic=False):
    """Returns per-step (centres, quats, links, iterations, residual, rebuilt) of the bodies this rank owns (all of them
    when comm is None: the single-GPU step)."""
    from mundy_b200 import dist as mbd
    from mundy_b200 import ops
    w = _workload(kind, n)
    cell = [w["edge"]] * 3 if periodic else None
    lo, hi = (0, n) if comm is None else mbd.owned_range(n, comm.world, comm.rank)
    w = {k: (v[lo:hi] if isinstance(v, np.ndarray) and v.shape[:1] == (n,) else v) for k, v in w.items()}
    b, c, q = _bodies(w, device)
    links = ops.NeighborLinks().set_search_buffer(BUF)
    if cell is not None:
        links.set_periodic(cell)
    if kind == "rods":
        links.set_search_filter(ops.NeighborLinks.FILTER_SHAPE_DISTANCE)
    F = torch.zeros((hi - lo, 3), dtype=torch.float64, device=device)
    out = []
    for _ in range(steps):
        F.zero_()
        ops.abc_flow_force(c, F)
        if comm is None:
            st = links.step(b, MU, DT, TOL, 10000, fixed_iters=fixed, F_ext=F)
        else:
            st = mbd.step_dist(links, comm, b, MU, DT, TOL, 10000, fixed_iters=fixed, F_ext=F)
        out.append((c.cpu().numpy().copy(), None if q is None else q.cpu().numpy().copy(), st.num_links,
                    st.lcp.ite_count, st.lcp.max_abs_projected_sep, st.rebuilt))
    return out


@pytest.mark.parametrize("kind", ["rods", "spheres"])
def test_world1_distributed_step_equals_single_gpu_step(cuda, kind):
    """With one rank the distributed code path (ghost selection, local set, counted flags, reduction hooks) must reproduce
    the single-GPU step bit for bit."""
    from mundy_b200 import dist as mbd
    comm = mbd.Comm()
    for fixed in (6, -1):
        a = _run(kind, 4000, 3, fixed, "cuda:0")
        b = _run(kind, 4000, 3, fixed, "cuda:0", comm)
        for x, y in zip(a, b):
            assert x[2:] == y[2:]
            assert np.array_equal(x[0], y[0]) and (x[1] is None or np.array_equal(x[1], y[1]))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def _world_worker(rank, world, port, out, exchange):
    try:
        import torch.distributed as dist
        from mundy_b200 import dist as mbd
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        os.environ["MB200_DIST_EXCHANGE"] = exchange
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        dev = f"cuda:{rank}"
        comm = mbd.Comm()
        msgs = []
        cases = [("rods", 6000 // world * world, False), ("spheres", 8000 // world * world, False),
                 ("spheres", 8000 // world * world, True)]
        for kind, n, periodic in cases:
            lo, hi = mbd.owned_range(n, world, rank)
            # fixed iteration count: iterates must agree with the single-GPU run to 1e-10 (positions 1e-12)
            ref = _run(kind, n, 3, 8, dev, periodic=periodic)
            got = _run(kind, n, 3, 8, dev, comm, periodic=periodic)
            for s, (x, y) in enumerate(zip(ref, got)):
                assert _rel(y[0], x[0][lo:hi]) <= 1e-12, (kind, s, _rel(y[0], x[0][lo:hi]))
                assert x[1] is None or np.max(np.abs(y[1] - x[1][lo:hi])) <= 1e-12
                assert abs(y[4] - x[4]) <= 1e-9 * abs(x[4]) and y[3] == x[3] and y[5] == x[5]
                assert y[2] <= x[2]  # a rank holds a subset of the links
                # every link is counted by exactly one rank: the ranks' link counts add up to at least the global count
                total = torch.tensor([float(y[2])], device=dev)
                dist.all_reduce(total)
                assert total.item() >= x[2]
            # to tolerance: residual below tol on both; BBPGD is non-monotone, so the stopping iteration moves with the
            # rounding order of the global sums (observed 611 vs 514) and the two converged states differ by O(tol)
            ref = _run(kind, n, 2, -1, dev, periodic=periodic)
            got = _run(kind, n, 2, -1, dev, comm, periodic=periodic)
            for x, y in zip(ref, got):
                assert y[4] < TOL and x[4] < TOL and abs(y[3] - x[3]) <= max(2, x[3] // 3), (kind, x[3], y[3])
                assert np.max(np.abs(y[0] - x[0][lo:hi])) <= 20 * TOL
            msgs.append((kind, periodic, [g[2] for g in got], [g[3] for g in got], comm.exchange_mode))
        # "peer": the kernels exchanged through CUDA-IPC mapped memory (no silent fall back to NCCL on this box)
        assert comm.exchange_mode == exchange, (comm.exchange_mode, exchange)
        out.put((rank, "ok", msgs))
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        import traceback
        out.put((rank, f"{type(e).__name__}: {e}\n{traceback.format_exc()}", None))


@pytest.mark.parametrize("exchange", ["peer", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_step_matches_single_gpu(cuda, world, exchange):
    """Both forms of the per-iteration exchange: peer memory (remote stores + system-scope flags from inside the kernels)
    and NCCL send/recv + all-gather."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs (run with gpurun --gpus {world})")
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_world_worker, args=(r, world, port, out, exchange)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
    print(f"world {world} {exchange}:", sorted(res)[0][2])
