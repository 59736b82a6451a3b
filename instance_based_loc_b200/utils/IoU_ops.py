"""Mirror of the reference's utils/IoU_ops.py axis-aligned overlap measures (``calculate_3d_IoU`` :9-52,
``calculate_strict_overlap`` :54-97), used while the memory is built.

The bounds come from one min/max reduction kernel per cloud (``ibl_bounds3``; min and max are exact in any order, so
they equal numpy's); the dozen scalar operations after that follow the reference line by line in f64 on the host.

``calculate_obj_aligned_3d_IoU`` (:99-144) is NOT built: it chains Open3D's ``OrientedBoundingBox.create_from_points``
(Qhull convex hull + PCA), Objectron's box clipping and another Qhull volume -- three third-party algorithms whose
sources are not under the reference tree.  It raises instead of approximating.
"""
from __future__ import annotations

import numpy as np

from .. import kernels as K
from ..o3d.pointcloud import PointCloud


def _bounds(pcd):
    """pcd.min(axis=-1), pcd.max(axis=-1) of the reference's 3xN view; an empty cloud makes numpy raise there, which the
    reference turns into ``return 0``."""
    if isinstance(pcd, PointCloud):
        if pcd.n() == 0:
            return None
        return K.bounds3(pcd._p)
    a = np.asarray(pcd.points, dtype=np.float64)
    if a.shape[0] == 0:
        return None
    return K.bounds3(K.to_dev(a))


def _overlap(pcd1, pcd2):
    b1, b2 = _bounds(pcd1), _bounds(pcd2)
    if b1 is None or b2 is None:
        return None
    (bb1_min, bb1_max), (bb2_min, bb2_max) = b1, b2
    overlap_min_corner = np.stack([bb1_min, bb2_min], axis=0).max(axis=0)
    overlap_max_corner = np.stack([bb1_max, bb2_max], axis=0).min(axis=0)
    if (overlap_min_corner > overlap_max_corner).any():
        return None
    v = overlap_max_corner - overlap_min_corner
    overlap_volume = v[0] * v[1] * v[2]
    bb1 = bb1_max - bb1_min
    bb2 = bb2_max - bb2_min
    v1 = bb1[0] * bb1[1] * bb1[2]
    v2 = bb2[0] * bb2[1] * bb2[2]
    return overlap_volume, v1, v2


def calculate_3d_IoU(pcd1, pcd2):
    """:9-52 -- axis-aligned-box IoU of two clouds; 0 when the boxes are disjoint or a cloud is empty."""
    r = _overlap(pcd1, pcd2)
    if r is None:
        return 0
    overlap_volume, v1, v2 = r
    return overlap_volume / (v1 + v2 - overlap_volume)


def calculate_strict_overlap(pcd1, pcd2):
    """:54-97 -- overlap volume over the smaller box's volume."""
    r = _overlap(pcd1, pcd2)
    if r is None:
        return 0
    overlap_volume, v1, v2 = r
    return overlap_volume / min(v1, v2)


def calculate_obj_aligned_3d_IoU(pcd1, pcd2):
    raise NotImplementedError(
        "calculate_obj_aligned_3d_IoU (utils/IoU_ops.py:99-144) needs Open3D's Qhull-based oriented bounding box and "
        "Objectron's box IoU; neither source is in the reference tree and neither is restated here (DESIGN.md section 7)")
