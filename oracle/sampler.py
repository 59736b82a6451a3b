"""oracle/sampler.py -- TEST INFRASTRUCTURE ONLY.

numpy restatement of the counter-based hypothesis sampler ``ibl_sample`` (csrc/rigid.cuh): the
RANSAC draw j (0..2) of hypothesis h over C correspondences.  Open3D 0.17 draws
``corres[rand_gen()]`` from a thread-local mt19937 (Registration.cpp, A.7) -- nondeterministic
under OpenMP -- so the build feeds BOTH paths the same host-seeded stream; this module is the
oracle-side copy of that stream ("host-seeded identical hypothesis samples").
"""
from __future__ import annotations

import numpy as np

_M = np.uint64(0xFFFFFFFFFFFFFFFF)


def samples(seed: int, h0: int, H: int, C: int) -> np.ndarray:
    """-> int32 [H, 3] correspondence indices for hypotheses h0 .. h0+H-1."""
    with np.errstate(over="ignore"):
        h = np.arange(h0, h0 + H, dtype=np.uint64)[:, None]
        j = np.arange(3, dtype=np.uint64)[None, :]
        z = np.uint64(seed) + (np.uint64(3) * h + j + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        hi = z >> np.uint64(32)
        return ((hi * np.uint64(C)) >> np.uint64(32)).astype(np.int32)
