"""Multi-GPU host side: Hilbert-curve ordering of the bodies, the NCCL communicator, the distributed step wrapper, and a
numpy statement of the decomposition plan (what the device code in csrc/dist_step.cu builds), used by the CPU tests.

Decomposition (include/mundy_b200.h "Multi-GPU"): bodies are sorted along a 3-D Hilbert curve of their centres
(`hilbert_order`); rank r owns the contiguous block [r n/W, (r+1) n/W) of that order and passes ONLY that block to the
library, which finds, fetches and refreshes the ghost bodies it needs from the other ranks. The reference decomposes with stk::balance RCB and never names a curve index (mundy/math/Hilbert.hpp
is a curve *generator*); its Hilbert generator and this index function describe the same family of curves: consecutive
indices are face-adjacent lattice cells (tests/test_dist_cpu.py).
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import check, lib


# ---- Hilbert ordering (Skilling, "Programming the Hilbert curve", AIP Conf. Proc. 707, 2004: transpose form) ---------
def hilbert_index(cells, bits):
    """cells: (n, 3) integer lattice coordinates in [0, 2^bits). Returns the uint64 Hilbert index of every cell."""
    X = np.array(cells, dtype=np.uint64).T.copy()  # (3, n)
    n = 3
    M = np.uint64(1) << np.uint64(bits - 1)
    Q = M
    while Q > np.uint64(1):  # inverse undo excess work
        P = Q - np.uint64(1)
        for i in range(n):
            hit = (X[i] & Q) != 0
            X[0] = np.where(hit, X[0] ^ P, X[0])          # invert low bits of X[0]
            t = (X[0] ^ X[i]) & P                          # exchange low bits of X[0] and X[i]
            X[0] = np.where(hit, X[0], X[0] ^ t)
            X[i] = np.where(hit, X[i], X[i] ^ t)
        Q >>= np.uint64(1)
    for i in range(1, n):  # Gray encode
        X[i] ^= X[i - 1]
    t = np.zeros_like(X[0])
    Q = M
    while Q > np.uint64(1):
        t = np.where((X[n - 1] & Q) != 0, t ^ (Q - np.uint64(1)), t)
        Q >>= np.uint64(1)
    for i in range(n):
        X[i] ^= t
    # interleave the transposed bits: bit b of X[i] -> bit (b * 3 + (2 - i)) of the index
    h = np.zeros(X.shape[1], dtype=np.uint64)
    for b in range(bits - 1, -1, -1):
        for i in range(n):
            h = (h << np.uint64(1)) | ((X[i] >> np.uint64(b)) & np.uint64(1))
    return h


# The eight sub-cubes of the reference's recursive generator (mundy/math/src/mundy_math/Hilbert.hpp:48-85), in visiting
# order: (coefficients of (dr1, dr2, dr3) locating the sub-cube, its frame as signed picks (sign, index) of the parent's
# directors). E.g. the first recursive call is hilbert_3d(snew, i, ..., current_position_new, dr2, dr3, dr1).
_HILBERT_CHILDREN = (
    ((0, 0, 0), ((+1, 1), (+1, 2), (+1, 0))),
    ((1, 0, 0), ((+1, 2), (+1, 0), (+1, 1))),
    ((1, 1, 0), ((+1, 2), (+1, 0), (+1, 1))),
    ((0, 1, 0), ((-1, 0), (-1, 1), (+1, 2))),
    ((0, 1, 1), ((-1, 0), (-1, 1), (+1, 2))),
    ((1, 1, 1), ((-1, 2), (+1, 0), (-1, 1))),
    ((1, 0, 1), ((-1, 2), (+1, 0), (-1, 1))),
    ((0, 0, 1), ((+1, 1), (-1, 2), (-1, 0))),
)


def _hilbert_tables():
    """The recursion as an automaton: state = the frame (dr1, dr2, dr3), one of the signed axis permutations reachable from
    (x, y, z); input = the octant of the cell inside the current cube (one bit per axis); output = which of the eight
    recursive calls covers that octant, and the frame it is called with. Built by running the recursion's own arithmetic
    on unit cubes (snew = 1)."""
    eye = np.eye(3, dtype=np.int64)
    states, index, todo = [], {}, [tuple(map(tuple, eye))]
    rows_k, rows_next = [], []
    while todo:
        f = todo.pop()
        if f in index:
            continue
        index[f] = len(states)
        states.append(f)
        D = np.array(f, dtype=np.int64)
        start = -np.minimum(D, 0).sum(axis=0)  # entry corner relative to the minimum corner, in units of snew
        ks, nxt = [None] * 8, [None] * 8
        for k, (combo, frame) in enumerate(_HILBERT_CHILDREN):
            off = start + combo[0] * D[0] + combo[1] * D[1] + combo[2] * D[2]  # child's minimum corner / snew
            assert off.min() >= 0 and off.max() <= 1
            octant = int(off[0]) | (int(off[1]) << 1) | (int(off[2]) << 2)
            child = tuple(tuple(int(v) for v in sg * D[j]) for sg, j in frame)
            assert ks[octant] is None
            ks[octant], nxt[octant] = k, child
            todo.append(child)
        rows_k.append(ks)
        rows_next.append(nxt)
    T_k = np.array(rows_k, dtype=np.uint64)
    T_next = np.array([[index[c] for c in row] for row in rows_next], dtype=np.int64)
    return T_k, T_next


_HILBERT_TABLES = None


def hilbert_index_reference(cells, bits):
    """Position of every lattice cell (n, 3) in [0, 2^bits)^3 along the REFERENCE's Hilbert curve: the visiting order of
    math::hilbert_3d(s = 2^bits, start (0,0,0), frame (x, y, z)) (Hilbert.hpp:48-85; goldens UnitTestHilbert.cpp:48-295),
    found by descending that recursion for all cells at once (one table look-up per level, see _hilbert_tables) instead
    of enumerating the 8^bits lattice points. Invariant of the recursion used: `current_position` is always the minimum
    corner of the sub-cube; the curve enters at that corner shifted by snew along every axis a director points down
    (Hilbert.hpp:63-66)."""
    global _HILBERT_TABLES
    if _HILBERT_TABLES is None:
        _HILBERT_TABLES = _hilbert_tables()
    T_k, T_next = _HILBERT_TABLES
    p = np.asarray(cells, dtype=np.int64)
    state = np.zeros(len(p), dtype=np.int64)
    idx = np.zeros(len(p), dtype=np.uint64)
    for level in range(bits - 1, -1, -1):
        octant = ((p[:, 0] >> level) & 1) | (((p[:, 1] >> level) & 1) << 1) | (((p[:, 2] >> level) & 1) << 2)
        idx = (idx << np.uint64(3)) | T_k[state, octant]
        state = T_next[state, octant]
    return idx


def morton_key(cells):
    """The device's cell key (csrc/broadphase.cu:42-44 morton3): bit b of x, y, z -> bits 3b, 3b + 1, 3b + 2, i.e. z is the
    most significant axis. Sorting by it is the order of zorder_knn::Less<Point, 3> (mundy/math/src/mundy_math/zmort.hpp:
    195-220) on non-negative integer coordinates (tests/test_sfc_cpu.py). 10 bits per axis."""
    c = np.asarray(cells, dtype=np.uint64)
    key = np.zeros(len(c), dtype=np.uint64)
    for b in range(10):
        for d in range(3):
            key |= ((c[:, d] >> np.uint64(b)) & np.uint64(1)) << np.uint64(3 * b + d)
    return key


def hilbert_order(center, bits=10):
    """Permutation that sorts the bodies along the Hilbert curve of their centres (2^bits cells per axis)."""
    center = np.asarray(center, dtype=np.float64)
    lo, hi = center.min(axis=0), center.max(axis=0)
    span = float(np.max(hi - lo))
    span = span if span > 0 else 1.0
    cells = np.minimum(((center - lo) / span * (1 << bits)).astype(np.int64), (1 << bits) - 1)
    # the reference's own curve (Hilbert.hpp:48-85): visiting order pinned against hilbert_3d for 2^3 .. 16^3 lattices
    return np.argsort(hilbert_index_reference(cells, bits), kind="stable")


def reorder(workload, order):
    """Apply a body permutation to every per-body array of a workload dict."""
    n = len(workload["radius"])
    return {k: (v[order] if isinstance(v, np.ndarray) and v.shape[:1] == (n,) else v) for k, v in workload.items()}


def owned_range(n, world, rank):
    assert n % world == 0, "the number of bodies must be a multiple of the world size"
    m = n // world
    return rank * m, (rank + 1) * m


# ---- the decomposition plan, in numpy (specification of csrc/dist_step.cu; O(n_local^2)-free but host-speed) ---------
def plan(center, S, world, rank, find_pairs):
    """center (n,3), S (n,) search radii of ALL bodies. find_pairs(center_local, S_local) -> sorted (m,2) int pairs i<j.
    Returns dict(gidx, owned, links (global ids, with >= 1 owned end), counted, send {peer: gids}, recv {peer: gids})."""
    n = len(S)
    o0, o1 = owned_range(n, world, rank)
    own = np.zeros(n, dtype=bool)
    own[o0:o1] = True
    lo, hi = center[o0:o1].min(axis=0), center[o0:o1].max(axis=0)
    reach = ((S + S[o0:o1].max()) * (1.0 + 1e-9))[:, None]
    near = np.all((center >= lo - reach) & (center <= hi + reach), axis=1)
    gidx = np.nonzero(own | near)[0]                      # ascending global index: local order = global order
    pairs = find_pairs(center[gidx], S[gidx])
    gi, gj = gidx[pairs[:, 0]], gidx[pairs[:, 1]]
    keep = own[gi] | own[gj]                              # ghost-ghost links belong to other ranks
    gi, gj = gi[keep], gj[keep]
    m = n // world
    oi, oj = gi // m, gj // m
    counted = oi == rank                                  # gi < gj: the owner of the lower index counts the link
    cut = oi != oj
    send, recv = {}, {}
    for p in range(world):
        if p == rank:
            continue
        s = np.concatenate([gi[cut & (oi == rank) & (oj == p)], gj[cut & (oj == rank) & (oi == p)]])
        r = np.concatenate([gj[cut & (oi == rank) & (oj == p)], gi[cut & (oj == rank) & (oi == p)]])
        send[p], recv[p] = np.unique(s), np.unique(r)
    return dict(gidx=gidx, owned=own[gidx], links=np.stack([gi, gj], axis=1), counted=counted, send=send, recv=recv)


# ---- communicator and distributed step --------------------------------------------------------------------------------
class Comm:
    """NCCL communicator of libmundy_b200 (its own ncclComm; the unique id travels over torch.distributed)."""

    def __init__(self, device=None):
        import torch
        import torch.distributed as dist

        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        buf = C.create_string_buffer(128)
        if self.rank == 0:
            check(lib.mb200_comm_get_unique_id(buf), "comm_get_unique_id")
        if self.world > 1:
            dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
            t = torch.tensor(list(buf.raw), dtype=torch.uint8, device=dev if dist.get_backend() == "nccl" else "cpu")
            dist.broadcast(t, 0)
            buf = C.create_string_buffer(bytes(t.cpu().tolist()), 128)
        self._h = C.c_void_p()
        check(lib.mb200_comm_create(C.byref(self._h), self.world, self.rank, buf), "comm_create")

    @property
    def exchange_mode(self):
        """'peer' (kernels store into the peers' memory over NVLink) or 'nccl'; decided at the first distributed step."""
        return "peer" if lib.mb200_comm_exchange_mode(self._h) else "nccl"

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            lib.mb200_comm_destroy(self._h)
            self._h = None


def step_dist(links, comm, bodies, viscosity, dt, tol, max_iters, fixed_iters=-1, force_rebuild=False, F_ext=None,
              T_ext=None, F_col=None, U_out=None, throw_on_max_iters=False):
    """One distributed collision step. `bodies` = ops.make_bodies of the bodies THIS RANK OWNS (any number, also zero);
    F_ext / T_ext / F_col / U_out are per owned body. Collective over the communicator."""
    from .ops import _f64, _stream
    params = capi.StepParams(viscosity, dt, tol, max_iters, fixed_iters, int(throw_on_max_iters), int(force_rebuild))
    stats = capi.StepStats()
    check(lib.mb200_collision_step_dist(links._h, comm._h, _stream(), C.byref(bodies), C.byref(params), _f64(F_ext),
                                        _f64(T_ext), _f64(F_col), _f64(U_out), C.byref(stats)), "collision_step_dist")
    return stats


def local_counts(links):
    a, b = C.c_int64(0), C.c_int64(0)
    check(lib.mb200_dist_local_counts(links._h, C.byref(a), C.byref(b)), "dist_local_counts")
    return a.value, b.value


# ---- re-partitioning (csrc/repartition.cu) ------------------------------------------------------------------------------
HILBERT_BITS = 10       # lattice cells per axis of the repartitioning keys
BUCKET_SHIFT = 12       # keys >> 12: the 2^18 histogram buckets the splitters are chosen among


def lattice_of(centers):
    """(lo, span) of the repartitioning lattice: the bounding box of ALL ranks' centres (list of (n_r, 3) arrays)."""
    pts = [np.asarray(c, dtype=np.float64) for c in centers if len(c)]
    if not pts:
        return np.zeros(3), 1.0
    lo = np.min([p.min(axis=0) for p in pts], axis=0)
    hi = np.max([p.max(axis=0) for p in pts], axis=0)
    span = float(np.max(hi - lo))
    return lo, (span if span > 0 else 1.0)


def hilbert_keys(center, lo, span, bits=HILBERT_BITS):
    """numpy statement of mb200_hilbert_keys: the reference curve's index of every centre's lattice cell."""
    center = np.asarray(center, dtype=np.float64).reshape(-1, 3)
    cells = np.clip(((center - lo) / span * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    return hilbert_index_reference(cells, bits).astype(np.uint32)


def repartition_plan(keys_per_rank):
    """numpy statement of mb200_dist_repartition's decisions. keys_per_rank: list (one entry per rank) of uint32 key arrays
    in the ranks' owned order. Returns (split, dest): split[r] = first bucket of rank r (split[W] = 2^18), and per rank
    the list `dest[r]` of (source rank, source index) pairs it owns afterwards, in its final order (ascending key; ties by
    source rank, then source order)."""
    W = len(keys_per_rank)
    nb = 1 << (3 * HILBERT_BITS - BUCKET_SHIFT)
    hist = np.zeros(nb, dtype=np.int64)
    for k in keys_per_rank:
        hist += np.bincount(np.asarray(k, dtype=np.int64) >> BUCKET_SHIFT, minlength=nb)
    total = int(hist.sum())
    below = np.concatenate([[0], np.cumsum(hist)])          # below[b] = bodies in buckets < b
    split = [0] * (W + 1)
    for r in range(1, W):
        split[r] = int(np.searchsorted(below, (total * r) // W, side="left"))  # first b with below[b] >= r N / W
    split[W] = nb
    dest = []
    for r in range(W):
        rows = []
        for src, k in enumerate(keys_per_rank):
            k = np.asarray(k, dtype=np.int64)
            sel = np.nonzero(((k >> BUCKET_SHIFT) >= split[r]) & ((k >> BUCKET_SHIFT) < split[r + 1]))[0]
            sel = sel[np.argsort(k[sel], kind="stable")]
            rows.append(np.stack([np.full(len(sel), src, dtype=np.int64), sel.astype(np.int64), k[sel]], axis=1))
        allr = np.concatenate(rows) if rows else np.zeros((0, 3), dtype=np.int64)
        allr = allr[np.argsort(allr[:, 2], kind="stable")]
        dest.append(allr[:, :2])
    return split, dest


def device_hilbert_keys(center, lo, span):
    import torch
    from .ops import _f64, _stream
    n = center.shape[0]
    keys = torch.empty(n, dtype=torch.int32, device=center.device)
    lo_h = (C.c_double * 3)(*[float(x) for x in lo])
    check(lib.mb200_hilbert_keys(_stream(), n, _f64(center), lo_h, float(span), C.c_void_p(keys.data_ptr())),
          "hilbert_keys")
    return keys


def repartition(comm, center, radius, quat=None, length=None, kind=None, extra=None, capacity=None, links=None):
    """Collective. Tensors of the bodies THIS RANK owns -> dict of the tensors it owns afterwards (views of
    capacity-sized buffers cut to the new count), plus 'origin' ((source rank << 32) | source index) and 'keys'."""
    import torch
    from .ops import _f64, _i32, _stream, make_bodies
    n = radius.numel()
    cap = int(capacity) if capacity is not None else max(2 * n, 1024)
    dev = radius.device
    out = {"center": torch.empty((cap, 3), dtype=torch.float64, device=dev),
           "radius": torch.empty(cap, dtype=torch.float64, device=dev),
           "quat": torch.empty((cap, 4), dtype=torch.float64, device=dev) if quat is not None else None,
           "length": torch.empty(cap, dtype=torch.float64, device=dev) if length is not None else None,
           "kind": torch.empty(cap, dtype=torch.int32, device=dev) if kind is not None else None}
    edim = 0 if extra is None else int(extra.shape[1])
    extra_out = torch.empty((cap, edim), dtype=torch.float64, device=dev) if edim else None
    origin = torch.empty(cap, dtype=torch.int64, device=dev)
    keys = torch.empty(cap, dtype=torch.int32, device=dev)
    b_in = make_bodies(center, radius, quat, length, kind)
    b_out = make_bodies(out["center"], out["radius"], out["quat"], out["length"], out["kind"])
    check(lib.mb200_dist_repartition(comm._h, _stream(), links._h if links is not None else None, C.byref(b_in),
                                     _f64(extra), edim, cap, C.byref(b_out), _f64(extra_out),
                                     C.c_void_p(origin.data_ptr()), C.c_void_p(keys.data_ptr())), "dist_repartition")
    m = int(b_out.n)
    res = {k: (v[:m] if v is not None else None) for k, v in out.items()}
    res["extra"] = extra_out[:m] if extra_out is not None else None
    res["origin"], res["keys"] = origin[:m], keys[:m]
    return res
