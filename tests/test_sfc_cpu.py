"""CPU tests of the space-filling-curve specifications against the reference's own code (stored
outputs tests/golden/sfc_ref.npz from mundy/math/src/mundy_math/{Hilbert,zmort}.hpp compiled in oracle/_ref; live when that
library is present):
* mundy_b200.dist.hilbert_index_reference is the visiting order of the reference's math::hilbert_3d (2^3 .. 16^3 lattices;
  the 2^3 and 4^3 orders are also the literal goldens of UnitTestHilbert.cpp:48-100), and it is the curve the multi-GPU
  decomposition sorts bodies along;
* the device's cell key (morton3, csrc/broadphase.cu:42-44, restated as mundy_b200.dist.morton_key) sorts lattice cells in
  the order of zorder_knn::Less<Point, 3> (zmort.hpp:195-220). mundy::math::zmorton_less (zmort.hpp:228-266) breaks ties of
  the most significant differing bit towards the LOWEST axis instead of the highest, so it is a different (also valid)
  z-order; the test records the difference."""
import os

import numpy as np
import pytest

import oracle
from mundy_b200 import dist as mbd

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "sfc_ref.npz"))


@pytest.mark.parametrize("bits", [1, 2, 3, 4])
def test_hilbert_index_is_the_reference_visiting_order(bits):
    pts = G[f"hilbert_{1 << bits}"].astype(np.int64)
    assert np.array_equal(mbd.hilbert_index_reference(pts, bits), np.arange(len(pts), dtype=np.uint64))


def test_reference_unit_test_goldens():
    # UnitTestHilbert.cpp:59-62 (2^3) and :79-83 (first eight points of the 4^3 lattice)
    want2 = [(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 1, 1), (1, 1, 1), (1, 0, 1), (0, 0, 1)]
    want4 = [(0, 0, 0), (0, 1, 0), (0, 1, 1), (0, 0, 1), (1, 0, 1), (1, 1, 1), (1, 1, 0), (1, 0, 0)]
    assert [tuple(p) for p in G["hilbert_2"]] == want2
    assert [tuple(p) for p in G["hilbert_4"][:8]] == want4
    order = np.argsort(mbd.hilbert_index_reference(np.array(want2[::-1]), 1))
    assert [want2[::-1][k] for k in order] == want2


@pytest.mark.skipif(oracle.ref is None, reason="oracle/_ref/libmundy_ref.so not present")
def test_hilbert_live_reference_32_cubed():
    pts = oracle.ref_hilbert_3d(32)
    assert np.array_equal(mbd.hilbert_index_reference(pts, 5), np.arange(32 ** 3, dtype=np.uint64))
    assert np.all(np.abs(np.diff(pts, axis=0)).sum(axis=1) == 1)  # a Hilbert curve: consecutive cells share a face


def test_hilbert_index_at_decomposition_resolution_is_a_continuous_bijection():
    bits = 10
    rng = np.random.default_rng(4)
    cells = rng.integers(0, 1 << bits, (20000, 3))
    h = mbd.hilbert_index_reference(cells, bits)
    assert h.max() < np.uint64(1) << np.uint64(3 * bits)
    uniq = np.unique(cells, axis=0)
    assert len(np.unique(mbd.hilbert_index_reference(uniq, bits))) == len(uniq)
    # refinement: the index at 2^b cells per axis is the top 3b bits of the index at 2^(b+1)... only when the frames agree,
    # which the recursion guarantees: a parent cube is finished before the curve leaves it
    coarse = mbd.hilbert_index_reference(cells >> 1, bits - 1)
    assert np.array_equal(h >> np.uint64(3), coarse)


def test_device_cell_key_is_the_knn_z_order():
    p, q = G["int_p"].astype(np.int64), G["int_q"].astype(np.int64)
    mine = mbd.morton_key(p) < mbd.morton_key(q)
    assert np.array_equal(mine, G["int_knn_less"])
    differs = mine != G["int_zmorton_less"]
    assert differs.any()  # math::zmorton_less: ties of the top differing bit go to the lowest axis
    # ... and only there: whenever one axis alone carries the most significant differing bit, all three agree
    top = np.array([[int(v).bit_length() for v in row] for row in (p ^ q)])
    unique_top = (top == top.max(axis=1, keepdims=True)).sum(axis=1) == 1
    assert not differs[unique_top].any()


@pytest.mark.skipif(oracle.ref is None, reason="oracle/_ref/libmundy_ref.so not present")
def test_zorder_live_reference():
    rng = np.random.default_rng(8)
    a, b = rng.integers(0, 1024, (5000, 3)), rng.integers(0, 1024, (5000, 3))
    assert np.array_equal(mbd.morton_key(a) < mbd.morton_key(b), oracle.ref_zorder_less(a, b, 0))
    # sorting by the key is a total order consistent with the comparator: adjacent sorted cells are never out of order
    cells = np.unique(a, axis=0)
    s = cells[np.argsort(mbd.morton_key(cells))]
    assert not oracle.ref_zorder_less(s[1:], s[:-1], 0).any()
