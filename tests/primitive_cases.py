"""Known-answer cases of the reference's unit tests for the remaining primitives, shared by the CPU
(oracle) and GPU (C-ABI) tests. Each entry: (function name, input arrays, {output index: expected}, tolerance).

AABB cases: mundy/geom/tests/unit_tests/UnitTestComputeAABB.cpp:137-205 (Point, LineSegment, Ellipsoid), tolerance
get_relaxed_zero_tolerance = 1e-8. Ellipsoid distances: mundy/geom/tests/unit_tests/UnitTestEllipsoidEllipsoid.cpp
(analytical ellipsoid cases 1-4 and the sphere property tests), tolerance TEST_DOUBLE_EPSILON = 1e-4 (:53)."""
import numpy as np

S = 1.0 / np.sqrt(2.0)
X90 = [S, S, 0.0, 0.0]       # get_quaternion_x_90, UnitTestComputeAABB.cpp:58-63
ID = [1.0, 0.0, 0.0, 0.0]
X_TO_Y = [S, 0.0, 0.0, S]    # quat_from_parallel_transport(x_hat, y_hat), Quaternion.hpp:1489-1506


def _a(x):
    return np.atleast_2d(np.asarray(x, dtype=np.float64))


def aabb_cases():
    return [
        ("aabb_point", (_a([[0, 0, 0], [1, -2, 3]]),), {0: _a([[0, 0, 0, 0, 0, 0], [1, -2, 3, 1, -2, 3]])}, 1e-8),
        ("aabb_line_segment", (_a([[1, -2, 3], [1, -2, 3], [4, 5, 6]]), _a([[1, -2, 3], [4, 5, 6], [1, -2, 3]])),
         {0: _a([[1, -2, 3, 1, -2, 3], [1, -2, 3, 4, 5, 6], [1, -2, 3, 4, 5, 6]])}, 1e-8),
        ("aabb_ellipsoid", (_a([[1, -2, 3], [1, -2, 3], [0, 0, 0], [1, -2, 3]]), _a([ID, ID, X90, X90]),
                            _a([[4, 4, 4], [4, 5, 6], [4, 5, 6], [4, 5, 6]])),
         {0: _a([[-3, -6, -1, 5, 2, 7], [-3, -7, -3, 5, 3, 9], [-4, -6, -5, 4, 6, 5], [-3, -8, -2, 5, 4, 8]])}, 1e-8),
    ]


def ellipsoid_cases():
    """UnitTestEllipsoidEllipsoid.cpp:119-235: (1) identical concentric -> -2 r2; (2) concentric, second twice as large
    -> -3 r2; (3) tip to tip along x at separation d; (4) first turned onto y above the second's flank at separation d."""
    r = [3.0, 1.0, 2.0]
    c1, q1, R1, c2, q2, R2, want = [], [], [], [], [], [], []

    def add(ca, qa, ra, cb, qb, rb, w):
        c1.append(ca), q1.append(qa), R1.append(ra), c2.append(cb), q2.append(qb), R2.append(rb), want.append(w)

    add([0, 0, 0], ID, r, [0, 0, 0], ID, r, -2.0)
    add([0, 0, 0], ID, r, [0, 0, 0], ID, [6.0, 2.0, 4.0], -3.0)
    for d in (0.2, -0.2, 0.0):
        add([-3.0 - 0.5 * d, 0, 0], ID, r, [3.0 + 0.5 * d, 0, 0], ID, r, d)
    for d in (0.2, -0.2, 0.0):
        add([0, 3.0 + 1.0 + d, 0], X_TO_Y, r, [0, 0, 0], ID, r, d)
    return ("ellipsoid_ellipsoid", (_a(c1), _a(q1), _a(R1), _a(c2), _a(q2), _a(R2)), {0: np.asarray(want)}, 1e-4)


def sphere_property_inputs(n=400, seed=5):
    """The reference's sphere property tests (:54-117): ellipsoids with three equal radii and random orientation behave as
    spheres. Returns (point-ellipsoid inputs, expected), (ellipsoid-ellipsoid inputs, expected)."""
    rng = np.random.default_rng(seed)

    def Q():
        q = rng.normal(size=(n, 4))
        return q / np.linalg.norm(q, axis=1, keepdims=True)

    c0, c1, x = rng.uniform(-10, 10, (n, 3)), rng.uniform(-10, 10, (n, 3)), rng.uniform(-10, 10, (n, 3))
    r0, r1 = rng.uniform(0.1, 10.0, n), rng.uniform(0.1, 10.0, n)
    R0, R1 = np.repeat(r0[:, None], 3, 1), np.repeat(r1[:, None], 3, 1)
    pe = ((x, c0, Q(), R0), np.linalg.norm(x - c0, axis=1) - r0)
    ee = ((c0, Q(), R0, c1, Q(), R1), np.linalg.norm(c1 - c0, axis=1) - r0 - r1)
    return pe, ee
