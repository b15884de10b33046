"""mb200_narrow_phase_no_alloc (static device scratch, no allocation) against mb200_narrow_phase, through the C-ABI.

Collected last on purpose. Bar: bit-identical (the narrow phase is a per-pair fp64
stage; both entry points run the same kernels). Two calls in a row: the second on a smaller, different body set, so
that the shared array still holds the first call's records beyond the second call's range.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def make_case(seed, n, npairs, rods):
    rng = np.random.default_rng(seed)
    c = rng.uniform(0.0, 20.0, (n, 3))
    r = rng.uniform(0.2, 0.6, n)
    i = rng.integers(0, n, npairs)
    j = (i + 1 + rng.integers(0, n - 1, npairs)) % n  # j != i
    pairs = np.stack([np.minimum(i, j), np.maximum(i, j)], axis=1).astype(np.int32)
    case = dict(center=dev(c), radius=dev(r), pairs=dev(pairs, torch.int32))
    if rods:
        q = rng.normal(size=(n, 4))
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        kind = (rng.uniform(size=n) < 0.7).astype(np.int32)  # spherocylinders among spheres
        case.update(quat=dev(q), length=dev(rng.uniform(0.0, 3.0, n)), kind=dev(kind, torch.int32))
    return case


def run_both(case, arms):
    from mundy_b200 import ops
    b = ops.make_bodies(case["center"], case["radius"], case.get("quat"), case.get("length"), case.get("kind"))
    npairs = case["pairs"].shape[0]
    want = ops.narrow_phase(b, case["pairs"], arms=arms)
    # outputs are the caller's: filled with NaN so that an unwritten element cannot pass
    phi = torch.full((npairs,), float("nan"), dtype=torch.float64, device="cuda")
    normal = torch.full((npairs, 3), float("nan"), dtype=torch.float64, device="cuda")
    ai = torch.full((npairs, 3), float("nan"), dtype=torch.float64, device="cuda") if arms else None
    aj = torch.full((npairs, 3), float("nan"), dtype=torch.float64, device="cuda") if arms else None
    got = ops.narrow_phase_no_alloc(b, case["pairs"], phi, normal, ai, aj)
    torch.cuda.synchronize()
    for name, g, w in zip(("phi", "normal", "arm_i", "arm_j"), got, want):
        if w is None:
            assert g is None
            continue
        assert torch.equal(g, w), f"{name}: max |diff| = {(g - w).abs().max().item()}"


@pytest.mark.parametrize("rods", [True, False])
def test_no_alloc_matches_narrow_phase_on_two_calls_in_a_row(cuda, rods):
    run_both(make_case(11, 20_000, 150_001, rods), arms=rods)
    run_both(make_case(12, 3_000, 40_000, rods), arms=rods)   # stale records of the first call beyond [0, 3000)
    run_both(make_case(11, 20_000, 150_001, rods), arms=False)


def test_no_alloc_makes_no_allocation(cuda):
    """Device memory in use is the same before and after (the stream-ordered pool of the default device included)."""
    from mundy_b200 import ops
    case = make_case(13, 5_000, 30_000, True)
    b = ops.make_bodies(case["center"], case["radius"], case["quat"], case["length"], case["kind"])
    npairs = case["pairs"].shape[0]
    phi = torch.empty(npairs, dtype=torch.float64, device="cuda")
    normal = torch.empty((npairs, 3), dtype=torch.float64, device="cuda")
    ops.narrow_phase_no_alloc(b, case["pairs"], phi, normal)  # set-up call: symbol address looked up here
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        ops.narrow_phase_no_alloc(b, case["pairs"], phi, normal)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0


def test_no_alloc_rejects_more_than_the_static_maximum(cuda):
    from mundy_b200 import ops
    from mundy_b200._capi import Mb200Error
    n = ops.NO_ALLOC_MAX_BODIES + 1
    case = make_case(14, n, 1_000, False)
    b = ops.make_bodies(case["center"], case["radius"])
    phi = torch.empty(1_000, dtype=torch.float64, device="cuda")
    normal = torch.empty((1_000, 3), dtype=torch.float64, device="cuda")
    with pytest.raises(Mb200Error, match="at most 65536 bodies"):
        ops.narrow_phase_no_alloc(b, case["pairs"], phi, normal)
    # exactly the maximum is accepted
    case = make_case(15, ops.NO_ALLOC_MAX_BODIES, 50_000, True)
    run_both(case, arms=True)
