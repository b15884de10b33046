"""CPU tests of the multi-GPU host logic: Hilbert ordering, the decomposition plan (numpy statement of what
csrc/dist_step.cu builds) checked for cross-rank consistency by two gloo ranks, and the NCCL-free parts of the ABI."""
import itertools
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from mundy_b200 import dist as mbd
from mundy_b200 import workloads as wl


@pytest.mark.parametrize("bits", [1, 2, 3, 4])
def test_hilbert_index_is_a_hilbert_curve(bits):
    n = 1 << bits
    cells = np.array(list(itertools.product(range(n), repeat=3)))
    h = mbd.hilbert_index(cells, bits)
    assert sorted(h.tolist()) == list(range(n ** 3))              # bijection onto [0, 8^bits)
    path = cells[np.argsort(h)]
    assert np.all(np.abs(np.diff(path, axis=0)).sum(axis=1) == 1)  # consecutive cells share a face


def test_hilbert_order_reproduces_the_reference_2cube_golden():
    # mundy/math/tests/unit_tests/UnitTestHilbert.cpp:47-66
    golden = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 1, 1), (1, 1, 1), (1, 0, 1), (0, 0, 1)]
    cells = np.array(list(itertools.product(range(2), repeat=3)), dtype=float)
    order = mbd.hilbert_order(cells, bits=1)
    assert [tuple(int(v) for v in cells[k]) for k in order] == golden


def test_hilbert_blocks_are_compact():
    w = wl.rods(16000, phi=0.1)
    w = mbd.reorder(w, mbd.hilbert_order(w["center"]))
    world, n = 8, len(w["radius"])
    total = np.prod(w["center"].max(0) - w["center"].min(0))
    vols = []
    for r in range(world):
        a, b = mbd.owned_range(n, world, r)
        c = w["center"][a:b]
        vols.append(np.prod(c.max(0) - c.min(0)))
    assert sum(vols) < 1.6 * total  # 8 Hilbert blocks of a uniform cube ~ 8 octants (random order would give 8x)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _plan_worker(rank, world, port, out):
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        w = wl.rods(3000, phi=0.1, length=4.0)
        w = mbd.reorder(w, mbd.hilbert_order(w["center"]))
        n = len(w["radius"])
        S = oracle.search_radii(w["radius"], 0.4, w["kind"], w["length"])
        p = mbd.plan(w["center"], S, world, rank, oracle.search)
        plans = [None] * world
        dist.all_gather_object(plans, {k: p[k] for k in ("links", "counted", "send", "recv")})
        # (1) what rank a sends to rank b is exactly what b expects from a, in the same order
        for a in range(world):
            for b in range(world):
                if a != b:
                    assert np.array_equal(plans[a]["send"][b], plans[b]["recv"][a]), (a, b)
        # (2) every link of the global problem is counted by exactly one rank
        glob = oracle.search(w["center"], S)
        counted = np.concatenate([pl["links"][pl["counted"]] for pl in plans])
        counted = counted[np.lexsort((counted[:, 1], counted[:, 0]))]
        assert np.array_equal(counted, glob)
        # (3) a rank holds every link incident to its owned bodies
        a0, a1 = mbd.owned_range(n, world, rank)
        mine = glob[((glob[:, 0] >= a0) & (glob[:, 0] < a1)) | ((glob[:, 1] >= a0) & (glob[:, 1] < a1))]
        have = p["links"][np.lexsort((p["links"][:, 1], p["links"][:, 0]))]
        assert np.array_equal(have, mine)
        # (4) the scalar reductions of the solver: max / sum over ranks of per-rank partials of counted links
        x = torch.tensor([float(p["counted"].sum())])
        dist.all_reduce(x)
        assert int(x.item()) == len(glob)
        out.put((rank, "ok", int(p["counted"].sum()), len(p["gidx"])))
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        out.put((rank, f"{type(e).__name__}: {e}", 0, 0))


@pytest.mark.parametrize("world", [2, 4])
def test_decomposition_plan_is_consistent_across_gloo_ranks(world):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_plan_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
    assert all(r[3] < 3000 for r in res) or world == 2  # with 4 ranks nobody needs the whole system


def test_comm_abi_without_device():
    from mundy_b200 import _capi
    assert _capi.lib.mb200_comm_world(None) == 1 and _capi.lib.mb200_comm_rank(None) == 0
    if not torch.cuda.is_available():
        h = _capi.vp()
        import ctypes as C
        buf = C.create_string_buffer(128)
        st = _capi.lib.mb200_comm_create(C.byref(h), 1, 0, buf)
        assert st == -2  # MB200_ERR_CUDA: no device, no CPU fallback


# ---- re-partitioning: the numpy statement of csrc/repartition.cu's decisions --------------------------------------------
def test_repartition_keys_order_bodies_like_hilbert_order():
    rng = np.random.default_rng(1)
    c = rng.uniform(-3.0, 40.0, (20000, 3))
    lo, span = mbd.lattice_of([c[:7000], c[7000:]])
    keys = mbd.hilbert_keys(c, lo, span)
    assert keys.max() < (1 << 30)
    assert np.array_equal(np.argsort(keys, kind="stable"), mbd.hilbert_order(c))


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_repartition_plan_properties(world):
    rng = np.random.default_rng(world)
    n = 30000
    c = rng.uniform(0.0, 60.0, (n, 3))
    owner = rng.choice(world, size=n, p=np.arange(1, world + 1) / np.arange(1, world + 1).sum())
    per = [np.nonzero(owner == r)[0] for r in range(world)]
    lo, span = mbd.lattice_of([c[ix] for ix in per])
    keys = [mbd.hilbert_keys(c[ix], lo, span) for ix in per]
    split, dest = mbd.repartition_plan(keys)
    assert split[0] == 0 and split[-1] == 1 << 18 and all(a <= b for a, b in zip(split, split[1:]))
    gid = [np.array([per[s][i] for s, i in d], dtype=np.int64) for d in dest]
    assert sorted(np.concatenate(gid).tolist()) == list(range(n))          # every body has exactly one new owner
    allk = np.concatenate([mbd.hilbert_keys(c[g], lo, span).astype(np.int64) for g in gid])
    assert np.all(np.diff(allk) >= 0)                                        # rank order = curve order, sorted inside
    biggest = np.bincount(np.concatenate(keys).astype(np.int64) >> mbd.BUCKET_SHIFT).max()
    sizes = [len(g) for g in gid]
    assert max(sizes) - min(sizes) <= 2 * biggest                            # equal up to one bucket's population
    # bodies of one bucket stay together
    for r, g in enumerate(gid):
        b = mbd.hilbert_keys(c[g], lo, span).astype(np.int64) >> mbd.BUCKET_SHIFT
        assert len(g) == 0 or (b.min() >= split[r] and b.max() < split[r + 1])


def _splitter_worker(rank, world, port, out):
    """What every rank does with the all-reduced histogram must give the splitters of the global statement."""
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        rng = np.random.default_rng(11)
        n = 20000
        c = rng.uniform(0.0, 30.0, (n, 3))
        owner = rng.integers(0, world, n)
        per = [np.nonzero(owner == r)[0] for r in range(world)]
        # bounding boxes all-gathered (6 doubles per rank), as the device code does
        mine = c[per[rank]]
        box = torch.tensor(np.concatenate([mine.min(axis=0), mine.max(axis=0)]))
        boxes = [torch.zeros(6, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(boxes, box)
        lo = np.min([b[:3].numpy() for b in boxes], axis=0)
        span = float(np.max(np.max([b[3:].numpy() for b in boxes], axis=0) - lo))
        lo_g, span_g = mbd.lattice_of([c[ix] for ix in per])
        assert np.array_equal(lo, lo_g) and span == span_g
        keys = mbd.hilbert_keys(mine, lo, span).astype(np.int64)
        hist = torch.tensor(np.bincount(keys >> mbd.BUCKET_SHIFT, minlength=1 << 18))
        dist.all_reduce(hist)
        below = np.concatenate([[0], np.cumsum(hist.numpy())])
        split = [0] + [int(np.searchsorted(below, (n * r) // world, side="left")) for r in range(1, world)] + [1 << 18]
        want_split, dest = mbd.repartition_plan([mbd.hilbert_keys(c[ix], lo, span) for ix in per])
        assert split == want_split
        # send counts of this rank = what the destination ranks expect from it
        bucket = keys >> mbd.BUCKET_SHIFT
        send = [int(np.sum((bucket >= split[r]) & (bucket < split[r + 1]))) for r in range(world)]
        assert send == [int(np.sum(d[:, 0] == rank)) for d in dest]
        out.put((rank, "ok"))
        dist.destroy_process_group()
    except Exception as e:  # noqa: BLE001
        out.put((rank, f"{type(e).__name__}: {e}"))


def test_repartition_splitters_agree_across_gloo_ranks():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_splitter_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = [out.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res
