"""Synthetic inputs for the BASELINE.json configurations. numpy only, fixed seeds.

The reference draws its inputs with OpenRAND Philox streams (scrap/lcp_spheres/KokkosLCP.cpp:933-950); OpenRAND is
not available, and parity does not depend on the generator because the same arrays feed the oracle and the GPU.
Orientations follow mundy/geom/src/mundy_geom/randomize.hpp:67-84 (uniform on S^2: z = 2u-1, theta = 2 pi u',
quaternion = shortest rotation taking z_hat to the axis).
"""
import math

import numpy as np

SPHERE, ROD = 0, 1


def _quat_from_z_to(axis):
    """Unit quaternion (w,x,y,z) rotating z_hat onto `axis` (n x 3, unit)."""
    w = 1.0 + axis[:, 2]
    q = np.stack([w, -axis[:, 1], axis[:, 0], np.zeros_like(w)], axis=1)  # (1 + z.a, z x a)
    flip = w < 1e-12
    q[flip] = np.array([0.0, 1.0, 0.0, 0.0])
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q


def box_edge(n, volume_per_body, phi):
    return (n * volume_per_body / phi) ** (1.0 / 3.0)


def monodisperse_spheres(n, phi=0.30, radius=1.0, seed=1234):
    """cfg1: n unit spheres uniform in a cube at volume fraction phi (KokkosLCP.cpp:963-992)."""
    rng = np.random.default_rng(seed)
    edge = box_edge(n, 4.0 / 3.0 * math.pi * radius ** 3, phi)
    center = rng.uniform(0.0, edge, size=(n, 3))
    return dict(center=center, radius=np.full(n, radius), edge=edge)


def polydisperse_spheres(n, phi=0.35, rmin=0.5, rmax=1.0, seed=2):
    """cfg2: radii ~ U(rmin, rmax)."""
    rng = np.random.default_rng(seed)
    radius = rng.uniform(rmin, rmax, size=n)
    vol = 4.0 / 3.0 * math.pi * float(np.mean(radius ** 3))
    edge = box_edge(n, vol, phi)
    center = rng.uniform(0.0, edge, size=(n, 3))
    return dict(center=center, radius=radius, edge=edge)


def rods(n, phi=0.10, radius=0.5, length=9.0, seed=3):
    """cfg3: spherocylinders, aspect ratio (L + 2r) / 2r = 10 by default."""
    rng = np.random.default_rng(seed)
    vol = math.pi * radius ** 2 * length + 4.0 / 3.0 * math.pi * radius ** 3
    edge = box_edge(n, vol, phi)
    center = rng.uniform(0.0, edge, size=(n, 3))
    z = 2.0 * rng.uniform(size=n) - 1.0
    th = 2.0 * math.pi * rng.uniform(size=n)
    s = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    axis = np.stack([s * np.cos(th), s * np.sin(th), z], axis=1)
    return dict(center=center, radius=np.full(n, radius), length=np.full(n, length), quat=_quat_from_z_to(axis),
                kind=np.full(n, ROD, dtype=np.int32), edge=edge)


def mixed(n, phi=0.30, seed=5, sphere_radius=1.0, rod_radius=0.5, rod_length=4.0):
    """cfg5-like: half spheres, half rods, interleaved."""
    rng = np.random.default_rng(seed)
    kind = (np.arange(n) % 2).astype(np.int32)
    radius = np.where(kind == ROD, rod_radius, sphere_radius)
    length = np.where(kind == ROD, rod_length, 0.0)
    vol_s = 4.0 / 3.0 * math.pi * sphere_radius ** 3
    vol_r = math.pi * rod_radius ** 2 * rod_length + 4.0 / 3.0 * math.pi * rod_radius ** 3
    edge = box_edge(n, 0.5 * (vol_s + vol_r), phi)
    center = rng.uniform(0.0, edge, size=(n, 3))
    z = 2.0 * rng.uniform(size=n) - 1.0
    th = 2.0 * math.pi * rng.uniform(size=n)
    s = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    axis = np.stack([s * np.cos(th), s * np.sin(th), z], axis=1)
    return dict(center=center, radius=radius, length=length, quat=_quat_from_z_to(axis), kind=kind, edge=edge)


def abc_flow_force(center):
    """'Attractive ABC flow' forcing of scrap/parameter_interface/NGPSpheresLCP.cpp:971-998 (input generation)."""
    a, b, c, k = 1.0, math.sqrt(2.0), math.sqrt(3.0), 2.0
    x, y, z = center[:, 0], center[:, 1], center[:, 2]
    r = np.sqrt(x * x + y * y + z * z)
    r = np.where(r > 0, r, 1.0)
    return np.stack([a * np.sin(z) + c * np.cos(y) - k * x / r,
                     b * np.sin(x) + a * np.cos(z) - k * y / r,
                     c * np.sin(y) + b * np.cos(x) - k * z / r], axis=1)
