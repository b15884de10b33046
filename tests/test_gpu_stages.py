"""GPU parity tests, stage by stage, through the C-ABI (mundy_b200.ops -> libmundy_b200.so) against the CPU oracle.

Bar: bit-exact for index work (pair sets) and for every per-element fp64 stage whose operation order is pinned
(AABB, distance, narrow phase, D / M / D^T); 1e-10 relative where a reduction order or libm (sin/cos) differs.
"""
import numpy as np
import pytest
import torch

import oracle
from mundy_b200 import workloads as wl

pytestmark = pytest.mark.gpu

SIZES = [0, 1, 2, 255, 256, 257, 10_000]


def dev(x, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype, device="cuda")


def host(t):
    return t.detach().cpu().numpy()


def rand_quat(rng, n, unit=True):
    q = rng.normal(size=(n, 4))
    if unit:
        q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q


# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", SIZES)
def test_aabb_sphere_bit_exact(cuda, n):
    from mundy_b200 import ops
    rng = np.random.default_rng(n)
    c, r = rng.uniform(-50, 50, (n, 3)), rng.uniform(0.1, 2, n)
    got = host(ops.compute_aabb_sphere(dev(c), dev(r)))
    assert np.array_equal(got, oracle.aabb_sphere(c, r).reshape(n, 6))


@pytest.mark.parametrize("n", SIZES)
def test_aabb_spherocylinder_bit_exact(cuda, n):
    from mundy_b200 import ops
    rng = np.random.default_rng(100 + n)
    c, r, L = rng.uniform(-50, 50, (n, 3)), rng.uniform(0.1, 2, n), rng.uniform(0, 10, n)
    q = rand_quat(rng, n, unit=False)  # the reference does not assume unit norm
    got = host(ops.compute_aabb_spherocylinder(dev(c), dev(q), dev(r), dev(L)))
    assert np.array_equal(got, oracle.aabb_spherocylinder(c, q, r, L).reshape(n, 6))


@pytest.mark.parametrize("n", SIZES)
def test_aabb_segment_bit_exact(cuda, n):
    from mundy_b200 import ops
    rng = np.random.default_rng(200 + n)
    a, b, r = rng.uniform(-5, 5, (n, 3)), rng.uniform(-5, 5, (n, 3)), rng.uniform(0.1, 2, n)
    got = host(ops.compute_aabb_segment(dev(a), dev(b), dev(r)))
    assert np.array_equal(got, oracle.aabb_segment(a, b, r).reshape(n, 6))


def test_aabb_reference_goldens(cuda):
    """mundy/geom/tests/unit_tests/UnitTestComputeAABB.cpp:180-240 (sphere and spherocylinder cases)."""
    from mundy_b200 import ops
    got = host(ops.compute_aabb_sphere(dev([[1.0, -2.0, 3.0]]), dev([2.0])))
    assert np.allclose(got, [[-1, -4, 1, 3, 0, 5]], atol=1e-12)
    s = np.sqrt(0.5)
    q = np.array([[s, s, 0.0, 0.0]])  # 90 degrees about x: z_hat -> -y_hat
    got = host(ops.compute_aabb_spherocylinder(dev([[1.0, -2.0, 3.0]]), dev(q), dev([2.0]), dev([3.0])))
    assert np.allclose(got, [[-1, -5.5, 1, 3, 1.5, 5]], atol=1e-12)


# ---------------------------------------------------------------------------------------------------------------
def _segments(rng, n):
    l0, l1 = rng.uniform(-3, 3, (n, 3)), rng.uniform(-3, 3, (n, 3))
    m0, m1 = rng.uniform(-3, 3, (n, 3)), rng.uniform(-3, 3, (n, 3))
    k = n // 8
    if k:
        m1[:k] = m0[:k] + 1.7 * (l1[:k] - l0[:k])           # parallel -> collinear branch
        m0[k:2 * k] = l0[k:2 * k]                          # shared endpoint
        l1[2 * k:3 * k] = l0[2 * k:3 * k]                  # degenerate (zero-length) segment
        m1[3 * k:4 * k] = m0[3 * k:4 * k] = l0[3 * k:4 * k]  # both degenerate and coincident-ish
        m1[3 * k:4 * k] += 1e-3
    return l0, l1, m0, m1


@pytest.mark.parametrize("n", SIZES)
def test_distance_segment_segment_bit_exact(cuda, n):
    from mundy_b200 import ops
    l0, l1, m0, m1 = _segments(np.random.default_rng(300 + n), n)
    got = ops.distance_segment_segment(dev(l0), dev(l1), dev(m0), dev(m1))
    want = oracle.segseg(l0, l1, m0, m1)
    for g, w in zip(got, want):
        assert np.array_equal(host(g), w.reshape(host(g).shape))


@pytest.mark.parametrize("n", [0, 1, 1000])
def test_distance_other_pairs_bit_exact(cuda, n):
    from mundy_b200 import ops
    rng = np.random.default_rng(400 + n)
    a, b, c = rng.uniform(-3, 3, (n, 3)), rng.uniform(-3, 3, (n, 3)), rng.uniform(-3, 3, (n, 3))
    r1, r2 = rng.uniform(0.1, 1, n), rng.uniform(0.1, 1, n)
    for g, w in zip(ops.distance_sphere_sphere(dev(a), dev(r1), dev(b), dev(r2)), oracle.sphere_sphere(a, r1, b, r2)):
        assert np.array_equal(host(g), w.reshape(host(g).shape))
    for g, w in zip(ops.distance_point_segment(dev(a), dev(b), dev(c)), oracle.point_seg(a, b, c)):
        assert np.array_equal(host(g), w.reshape(host(g).shape))
    for g, w in zip(ops.distance_segment_sphere(dev(a), dev(b), dev(c), dev(r1)), oracle.seg_sphere(a, b, c, r1)):
        assert np.array_equal(host(g), w.reshape(host(g).shape))


def test_segment_segment_reference_golden(cuda):
    """The 'pesky' collinear case of mundy/geom/tests/unit_tests/UnitTestSegmentSegment.cpp:459-468."""
    from mundy_b200 import ops
    golden = np.load(__import__("os").path.join(__import__("os").path.dirname(__file__), "golden", "segseg_cases.npz"))
    got = ops.distance_segment_segment(dev(golden["l0"]), dev(golden["l1"]), dev(golden["m0"]), dev(golden["m1"]))
    assert np.allclose(host(got[0]), golden["dist"], rtol=0, atol=1e-12)


# ---------------------------------------------------------------------------------------------------------------
def _bodies_dev(w):
    from mundy_b200 import ops
    t = dict(center=dev(w["center"]), radius=dev(w["radius"]))
    if "kind" in w:
        t.update(quat=dev(w["quat"]), length=dev(w["length"]), kind=dev(w["kind"], torch.int32))
    return ops.make_bodies(t["center"], t["radius"], t.get("quat"), t.get("length"), t.get("kind")), t


@pytest.mark.parametrize("maker,kw", [(wl.monodisperse_spheres, dict(n=3000)), (wl.polydisperse_spheres, dict(n=3000)),
                                      (wl.rods, dict(n=2000, phi=0.2)), (wl.mixed, dict(n=2000))])
def test_narrow_phase_bit_exact(cuda, maker, kw):
    from mundy_b200 import ops
    w = maker(**kw)
    S = oracle.search_radii(w["radius"], 0.3, w.get("kind"), w.get("length"))
    pairs = oracle.search(w["center"], S)
    assert len(pairs) > 100
    arms = "kind" in w
    want = oracle.narrow_phase(pairs, w["center"], w["radius"], w.get("quat"), w.get("length"), w.get("kind"), arms)
    bodies, _ = _bodies_dev(w)
    got = ops.narrow_phase(bodies, dev(pairs, torch.int32), arms=arms)
    for g, x in zip(got, want):
        if x is not None:
            assert np.array_equal(host(g), x.reshape(host(g).shape))


@pytest.mark.parametrize("n,m,empty", [(1, 0, 0.0), (50, 400, 0.0), (20000, 130001, 0.1), (7, 5000, 0.5)])
def test_links_coo_to_crs(cuda, n, m, empty):
    """entity -> connected links from link -> linked entities; a stable argsort in numpy is the
    statement of the result (ascending (link, slot) per entity), the counts are the reference's atomic counts
    (NgpCOOToCRSSynchronizer.hpp:405-468), empty slots are skipped (:449)."""
    from mundy_b200 import ops
    rng = np.random.default_rng(n + m)
    linked = rng.integers(0, n, size=(m, 2)).astype(np.int32)
    if empty > 0:
        linked[rng.uniform(size=(m, 2)) < empty] = -1
    off, links, slots = ops.links_coo_to_crs(n, dev(linked, torch.int32))
    flat = linked.reshape(-1)
    valid = np.nonzero(flat >= 0)[0]
    order = valid[np.argsort(flat[valid], kind="stable")]
    assert np.array_equal(host(off), np.concatenate([[0], np.cumsum(np.bincount(flat[valid], minlength=n))]))
    assert np.array_equal(host(links), order // 2) and np.array_equal(host(slots), order % 2)
    # every entry points back at its entity
    o = host(off)
    ent = np.repeat(np.arange(n), np.diff(o))
    assert np.array_equal(linked[host(links), host(slots)], ent)


def test_links_coo_to_crs_reference_unit_test_case(cuda):
    """The reference's own check of the synchroniser, restated for pair links
    (mundy/mesh/tests/unit_tests/UnitTestNgpLinkData.cpp): links are declared on freshly made entities
    (initialize_links :115-129, declare_and_validate_relations :131-160), the expected CRS connectivity is the
    inversion of that initialisation data and is compared per entity AS A SET (validate_crs_connectivity :276-327);
    then the relations of every link are destroyed and re-declared in swapped slots on the device
    (modify_ngp_link_data :241-273), which must leave the sets unchanged; a destroyed relation (empty slot,
    NgpCOOToCRSSynchronizer.hpp:449) drops out of its entity's row and of the atomic counts (:405-468)."""
    from mundy_b200 import ops
    n_links = 37
    # link k on the new entities (2k, 2k + 1), as initialize_links makes them; some entities get further links so that
    # rows hold more than one link (the reference's sets can)
    init = [[2 * k, 2 * k + 1] for k in range(n_links)]
    extra = [[0, 5], [5, 9], [9, 0], [73, 1]]
    link_and_linked = init + extra
    n_ent = 2 * n_links

    def expected(lk):  # validate_crs_connectivity :283-290
        conn = {}
        for link, ents in enumerate(lk):
            for e in ents:
                if e >= 0:
                    conn.setdefault(e, set()).add(link)
        return conn

    def check(lk):
        off, links, slots = (host(t) for t in ops.links_coo_to_crs(n_ent, dev(np.array(lk), torch.int32)))
        exp = expected(lk)
        for e in range(n_ent):
            row = links[off[e]:off[e + 1]]
            assert set(row.tolist()) == exp.get(e, set())
            assert len(row) == sum(1 for ents in lk for x in ents if x == e)  # the count of gather_part_1_count
            for l, sl in zip(row, slots[off[e]:off[e + 1]]):
                assert lk[l][sl] == e
        return off

    check(link_and_linked)
    swapped = [[b, a] for a, b in link_and_linked]  # modify_ngp_link_data: destroy both, declare in the other order
    check(swapped)
    destroyed = [list(x) for x in swapped]
    destroyed[3][0] = -1   # destroy_relation without a new declaration: an empty slot
    destroyed[38][1] = -1
    off = check(destroyed)
    assert off[-1] == 2 * len(link_and_linked) - 2


def test_links_coo_to_crs_rejects_out_of_range(cuda):
    from mundy_b200 import ops
    from mundy_b200._capi import Mb200Error
    with pytest.raises(Mb200Error):
        ops.links_coo_to_crs(4, dev(np.array([[0, 4]]), torch.int32))


@pytest.mark.parametrize("n", [0, 1, 257, 50_000])
def test_hilbert_keys_match_the_reference_curve(cuda, n):
    """mb200_hilbert_keys against dist.hilbert_index_reference (pinned to math::hilbert_3d's visiting order and the
    goldens of UnitTestHilbert.cpp in tests/test_dist_cpu.py): integer work, exact."""
    from mundy_b200 import dist as mbd
    rng = np.random.default_rng(n + 3)
    c = rng.uniform(-7.0, 31.0, (n, 3))
    if n > 1:
        c[0], c[1] = c.min(axis=0), c.max(axis=0)  # both corners of the lattice
    lo, span = mbd.lattice_of([c])
    got = host(mbd.device_hilbert_keys(dev(c), lo, span)).astype(np.int64)
    assert np.array_equal(got, mbd.hilbert_keys(c, lo, span).astype(np.int64))
