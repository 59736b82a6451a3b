"""Host-side mirror of the reference's ``utils/quaternion_ops.py`` (the evaluation step
right after ``ObjectMemory.localise``).  Quaternions are ``[w, x, y, z]`` as in the reference (the pose vectors are
``[x, y, z, qw, qx, qy, qz]``).

Every function takes a single quaternion ``[4]`` (the reference's call shape: quaternion_ops.py:5-24) or a stack
``[N, 4]`` (a whole 1000-frame run at once); the arithmetic is element-wise f64 in the reference's operation order, so
the scalar call gives the reference's bits for the product and the conjugate.  This is host bookkeeping on N x 7 pose
vectors (28 kB for 1000 frames): it stays numpy, no device work.
"""
from __future__ import annotations

import numpy as np


def _wxyz(q):
    q = np.asarray(q, dtype=np.float64)
    if q.shape[-1] != 4:
        raise ValueError(f"quaternion must have 4 components [w, x, y, z], got shape {q.shape}")
    return q[..., 0], q[..., 1], q[..., 2], q[..., 3]


class QuaternionOps:
    @staticmethod
    def quaternion_multiply(q1, q2):
        """Hamilton product q1 * q2 (quaternion_ops.py:5-12), terms accumulated left to right as there."""
        a, b, c, d = _wxyz(q1)
        e, f, g, h = _wxyz(q2)
        return np.stack([a * e - b * f - c * g - d * h,
                         a * f + b * e + c * h - d * g,
                         a * g - b * h + c * e + d * f,
                         a * h + b * g - c * f + d * e], axis=-1)

    @staticmethod
    def quaternion_conjugate(q):
        """[w, -x, -y, -z] (quaternion_ops.py:14-17)."""
        return np.asarray(q, dtype=np.float64) * np.array([1.0, -1.0, -1.0, -1.0])

    @staticmethod
    def quaternion_error(q1, q2):
        """Angle between two orientations, sign-ambiguity removed (quaternion_ops.py:20-24).

        The reference evaluates ``|atan2(|v|, w)|`` for conj(q1)*q2 and for conj(q1)*(-q2) and keeps the smaller.  The
        second product is the exact negation of the first (IEEE rounding is sign-symmetric), |v| >= 0, so the smaller
        of atan2(|v|, w) and atan2(|v|, -w) is atan2(|v|, |w|): one product, same value.  Returns a float for a single
        pair, an array [N] for stacks.
        """
        d = QuaternionOps.quaternion_multiply(QuaternionOps.quaternion_conjugate(q1), q2)
        v = d[..., 1:]
        n = np.sqrt((v[..., 0] * v[..., 0] + v[..., 1] * v[..., 1]) + v[..., 2] * v[..., 2])
        return np.arctan2(n, np.abs(d[..., 0]))
