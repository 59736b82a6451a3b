"""Shared helpers for the parity tests: seeded synthetic clouds shaped like the reference's object clouds
(surface samples ~6-7 mm apart, coloured) and thin numpy<->device adapters."""
from __future__ import annotations

import numpy as np


def surface_cloud(n, seed, scale=0.6, noise=0.002):
    """Points on an ellipsoid + box-face mixture with smooth colour field (+ noise): a stand-in for a segmented
    object cloud (SURVEY.md 8(d) synthetic inputs)."""
    rng = np.random.default_rng(seed)
    n1 = n // 2
    u = rng.normal(size=(n1, 3))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    ell = u * np.array([0.5, 0.3, 0.2]) * scale
    n2 = n - n1
    face = rng.integers(0, 3, n2)
    box = rng.uniform(-0.5, 0.5, (n2, 3)) * scale
    box[np.arange(n2), face] = 0.5 * scale * np.sign(rng.normal(size=n2))
    box = box * np.array([0.8, 0.5, 0.4]) + np.array([0.9 * scale, 0.1, 0.0])
    p = np.concatenate([ell, box]) + rng.normal(0, noise, (n, 3))
    c = 0.5 + 0.5 * np.stack([np.sin(7 * p[:, 0] + 1), np.cos(5 * p[:, 1]), np.sin(9 * p[:, 2] + 2)], axis=1)
    c = np.clip(c + rng.normal(0, 0.02, c.shape), 0, 1)
    perm = rng.permutation(n)
    return np.ascontiguousarray(p[perm]), np.ascontiguousarray(c[perm])


def random_rigid(seed, max_angle=0.5, max_t=0.2):
    rng = np.random.default_rng(seed)
    ax = rng.normal(size=3)
    ax /= np.linalg.norm(ax)
    ang = rng.uniform(-max_angle, max_angle)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    R = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = rng.uniform(-max_t, max_t, 3)
    return T


def rot_angle_deg(Ra, Rb):
    c = (np.trace(Ra.T @ Rb) - 1) / 2
    return float(np.degrees(np.arccos(np.clip(c, -1, 1))))


def dev(a, dtype=None):
    import torch
    from instance_based_loc_b200 import kernels as K
    if dtype is None:
        dtype = torch.float64
    return K.to_dev(a, dtype)
