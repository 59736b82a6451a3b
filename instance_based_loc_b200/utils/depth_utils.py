"""Drop-in for the reference's utils/depth_utils.py (same names, argument meaning and return types) on
libibl.so kernels.  Reference lines cited per function are relative to /root/reference/utils/depth_utils.py.

Differences that do not change results: the per-mask loop of the reference (n full-frame numpy passes and n
D2H mask copies, :198-204) is ONE kernel pass over the frame for all masks; clouds stay in HBM.
"""
from __future__ import annotations

import numpy as np
import torch
from scipy.spatial.transform import Rotation

from .. import kernels as K
from ..o3d.pointcloud import PointCloud, remove_radius_outlier_many

DEFAULT_OUTLIER_REMOVAL_CONFIG = \
    {
        "radius_nb_points": 12,
        "radius": 0.05,
    }


def _depth_to_dev(depth_image):
    """float32 depth stays float32 (numpy keeps f32*f32/pyfloat in f32, :65-67); anything else goes through
    float64 like numpy's promotion of e.g. uint16/5000. (already float64 when it reaches this module)."""
    if isinstance(depth_image, torch.Tensor):
        d = depth_image
        if d.dtype not in (torch.float32, torch.float64):
            d = d.to(torch.float64)
        return d.to(K._dev()).contiguous()
    d = np.asarray(depth_image)
    if d.dtype != np.float32:
        d = d.astype(np.float64)
    return torch.from_numpy(np.ascontiguousarray(d)).to(K._dev())


def _rgb_to_dev(rgb_image):
    if rgb_image is None:
        return None
    if isinstance(rgb_image, torch.Tensor):
        r = rgb_image.to(K._dev())
    else:
        r = torch.from_numpy(np.ascontiguousarray(rgb_image)).to(K._dev())
    if r.dtype != torch.uint8:
        raise TypeError("rgb_image must be uint8 [H,W,3] (the reference divides the uint8 image by 255 in float32)")
    if r.shape[-1] != 3:
        r = r[..., :3]
    return r.contiguous()


def _masks_to_dev(masks, H, W):
    if isinstance(masks, torch.Tensor):
        m = masks.to(K._dev())
    else:
        m = torch.from_numpy(np.ascontiguousarray(masks)).to(K._dev())
    n = int(m.shape[0])
    m = m.reshape(n, H, W)
    if m.dtype != torch.bool and m.dtype != torch.uint8:
        raise TypeError("masks must be bool/uint8 (SAM masks); arbitrary float masks are not supported")
    return m


def _clouds_from_masks(depth_image, rgb_image, masks, fx, fy, outlier_removal_config):
    d = _depth_to_dev(depth_image)
    H, W = d.shape
    if rgb_image is not None:
        assert tuple(np.shape(rgb_image)[:2]) == (H, W), "Depth and RGB image dimensions do not match"
    rgb = _rgb_to_dev(rgb_image)
    m = _masks_to_dev(masks, H, W)
    pts, cols, offs = K.backproject_masks(d, rgb, m, fx, fy)
    # point spacing of every mask's cloud ~ mean depth / focal length: a performance hint for the search grids built on
    # these clouds later (one small read-back per frame instead of a probe grid per cloud)
    spacing = [None] * (len(offs) - 1)
    if pts.shape[0]:
        cz = torch.cumsum(pts[:, 2], dim=0)
        ends = torch.as_tensor([max(o - 1, 0) for o in offs], device=pts.device)
        zs = cz[ends].cpu().numpy()
        for i in range(len(offs) - 1):
            cnt = offs[i + 1] - offs[i]
            if cnt > 0:
                zsum = zs[i + 1] - (zs[i] if offs[i] > 0 else 0.0)
                spacing[i] = float(max(zsum / cnt, 1e-3) / max(abs(float(fx)), 1e-9))
    out = [PointCloud.from_tensors(pts[offs[i]:offs[i + 1]], cols[offs[i]:offs[i + 1]] if cols is not None else None,
                                   spacing=spacing[i]) for i in range(len(offs) - 1)]
    if outlier_removal_config is not None:                     # :41 / :87, every mask's cloud on its own -- one batched pass
        out = remove_radius_outlier_many(out, outlier_removal_config["radius_nb_points"], outlier_removal_config["radius"])
    return out


def get_pointcloud_from_depth(depth_image, focal_lenth_x, focal_lenth_y,
                              outlier_removal_config=DEFAULT_OUTLIER_REMOVAL_CONFIG):
    """:12-44."""
    d = _depth_to_dev(depth_image)
    ones = torch.ones((1,) + tuple(d.shape), dtype=torch.uint8, device=d.device)
    return _clouds_from_masks(d, None, ones, focal_lenth_x, focal_lenth_y, outlier_removal_config)[0]


def get_coloured_pointcloud_from_depth(depth_image, rgb_image, focal_lenth_x, focal_lenth_y,
                                       outlier_removal_config=DEFAULT_OUTLIER_REMOVAL_CONFIG):
    """:46-90."""
    d = _depth_to_dev(depth_image)
    ones = torch.ones((1,) + tuple(d.shape), dtype=torch.uint8, device=d.device)
    return _clouds_from_masks(d, rgb_image, ones, focal_lenth_x, focal_lenth_y, outlier_removal_config)[0]


def get_mask_pointclouds_from_depth(depth_image, masks, focal_length_x, focal_length_y,
                                    outlier_removal_config=DEFAULT_OUTLIER_REMOVAL_CONFIG):
    """:146-174."""
    return _clouds_from_masks(depth_image, None, masks, focal_length_x, focal_length_y, outlier_removal_config)


def get_mask_coloured_pointclouds_from_depth(depth_image, rgb_image, masks, focal_length_x, focal_length_y,
                                             outlier_removal_config=DEFAULT_OUTLIER_REMOVAL_CONFIG):
    """:176-206."""
    return _clouds_from_masks(depth_image, rgb_image, masks, focal_length_x, focal_length_y, outlier_removal_config)


def _pose_matrix(pose, kinect=False):
    t = pose[:3]
    q = pose[3:]
    q /= np.linalg.norm(q)          # in place, like the reference (:102, :127)
    R = Rotation.from_quat(q).as_matrix()
    T = np.eye(4)
    if kinect:
        R2 = Rotation.from_euler('xyz', [0, np.pi, 0]).as_matrix()
        T[:3, :3] = R @ R2
        T[:3, 3] = -np.asarray(t, np.float64)
    else:
        T[:3, :3] = R
        T[:3, 3] = t
    return T


def transform_pointcloud(pointcloud, pose):
    """:92-116  (R p + t).  numpy evaluates R @ P.T with BLAS (unspecified summation order); the kernel uses
    ((R0 x + R1 y) + R2 z) + t -- equal to within 1 ulp."""
    T = _pose_matrix(pose)
    out = PointCloud.from_tensors(K.transform_points(T, pointcloud._p) if pointcloud.n() else pointcloud._p,
                                  pointcloud._c)
    return out


def transform_pointcloud_kinect(pointcloud, pose):
    """:118-144."""
    T = _pose_matrix(pose, kinect=True)
    return PointCloud.from_tensors(K.transform_points(T, pointcloud._p) if pointcloud.n() else pointcloud._p,
                                   pointcloud._c)


def voxel_down_sample_with_colors(pcd, voxel_size):
    """:211-265: key = floor(p / voxel) (int64 per axis), voxels in first-occurrence order, sequential means."""
    if pcd.is_empty():
        return PointCloud()
    p, c, n, _, _ = K.voxel_downsample(pcd._p, pcd._c if pcd.has_colors() else None,
                                       pcd._n if pcd.has_normals() else None, voxel_size)
    return PointCloud.from_tensors(p.contiguous(), None if c is None else c.contiguous(),
                                   None if n is None else n.contiguous())


def combine_point_clouds(pcds):
    """:268-272."""
    combined_pcd = PointCloud()
    for pcd in pcds:
        combined_pcd += pcd
    return combined_pcd


def compute_center(point_cloud):
    """:274-277."""
    return point_cloud.get_center()


def decompose_pose_matrix(pose_matrix):
    """:279-288."""
    translation = pose_matrix[:3, 3]
    rotation_matrix = pose_matrix[:3, :3]
    rotation = Rotation.from_matrix(rotation_matrix)
    quaternion = rotation.as_quat()
    result = np.concatenate((translation, quaternion))
    return result
