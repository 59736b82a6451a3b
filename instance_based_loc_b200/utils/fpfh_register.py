"""Drop-in for the reference's utils/fpfh_register.py (:86-150): same function names, arguments, return values
and error behaviour (any failure in the FPFH/RANSAC/coloured-ICP branch falls back to identity-initialised
point-to-point ICP on the raw clouds), with every Open3D call replaced by a libibl.so kernel through
instance_based_loc_b200.o3d."""
from __future__ import annotations

import copy

import numpy as np

from .. import o3d
from .._lib import IblError


def get_SVD_transform(p, q):
    """:67-82 (helper, unused by localise): tiny 3x3 host algebra, kept verbatim in behaviour."""
    p, q = np.asarray(p, np.float64), np.asarray(q, np.float64)
    u_p = np.mean(p, axis=0)
    u_q = np.mean(q, axis=0)
    W = (q - u_q).T @ (p - u_p)
    u, s, vh = np.linalg.svd(W, full_matrices=True)
    M = np.diag([1, 1, np.linalg.det(u) * np.linalg.det(vh)])
    R_recovered = u @ M @ vh
    T = np.eye(4)
    T[:3, :3] = R_recovered
    T[:3, 3] = u_q - R_recovered @ u_p
    return T


def get_transformation(source, target):
    """:24-65 (helper, unused by localise)."""
    source, target = np.array(source), np.array(target)
    cs, ct = np.mean(source, axis=0), np.mean(target, axis=0)
    H = np.dot((source - cs).T, target - ct)
    U, _, Vt = np.linalg.svd(H)
    R_matrix = np.dot(Vt.T, U.T)
    if np.linalg.det(R_matrix) < 0:
        Vt[2, :] *= -1
        R_matrix = np.dot(Vt.T, U.T)
    return R_matrix, ct - np.dot(cs, R_matrix)


def downsample_and_compute_fpfh(pcd, voxel_size):
    """:86-98.  NB the reference does NOT downsample (deep copy only); voxel_size only scales the radii."""
    pcd_down = copy.deepcopy(pcd)
    # the FPFH search (5 vs / 100) contains the normal search (2 vs / 30): announce it, so ONE search serves both
    pcd_down._knn_hint = (voxel_size * 5, 100)
    radius_normal = voxel_size * 2
    pcd_down.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=radius_normal, max_nn=30))
    radius_feature = voxel_size * 5
    pcd_fpfh = o3d.pipelines.registration.compute_fpfh_feature(
        pcd_down, o3d.geometry.KDTreeSearchParamHybrid(radius=radius_feature, max_nn=100))
    return pcd_down, pcd_fpfh


def register_point_clouds(source, target, voxel_size, global_dist_factor=1.5, local_dist_factor=0.4,
                          target_features=None):
    """:100-143.  ``target_features`` (not in the reference) lets a caller that registers many sources against the
    same memory cloud pass a cached ``downsample_and_compute_fpfh(target, voxel_size)`` result."""
    reg = o3d.pipelines.registration
    try:    # catch cases where normals cant be computed  (reference: bare except, :101/:137)
        source_down, source_fpfh = downsample_and_compute_fpfh(source, voxel_size)
        if target_features is not None:
            target_down, target_fpfh = target_features
        else:
            target_down, target_fpfh = downsample_and_compute_fpfh(target, voxel_size)
        distance_threshold = voxel_size * global_dist_factor
        result_ransac = reg.registration_ransac_based_on_feature_matching(
            source_down, target_down, source_fpfh, target_fpfh, True,
            distance_threshold,
            reg.TransformationEstimationPointToPoint(False),
            3, [
                reg.CorrespondenceCheckerBasedOnEdgeLength(0.9),
                reg.CorrespondenceCheckerBasedOnDistance(distance_threshold)
            ], reg.RANSACConvergenceCriteria(4000000, 0.99))
        result_icp = reg.registration_colored_icp(
            source_down, target_down, voxel_size * local_dist_factor, result_ransac.transformation,
            reg.TransformationEstimationForColoredICP()
        )
    except (KeyboardInterrupt, SystemExit, IblError):
        raise           # a CUDA / ABI failure must not turn into a plausible fallback pose
    except Exception:   # Open3D's own failures (no normals, no colours, empty clouds): the reference's bare except
        result_icp = reg.registration_icp(
            source, target, voxel_size * local_dist_factor, np.eye(4),
            reg.TransformationEstimationPointToPoint()
        )
    return result_icp.transformation, result_icp.inlier_rmse, result_icp.fitness


def evaluate_transform(source, target, trans_init, threshold=0.02):
    """:145-150."""
    res = o3d.pipelines.registration.evaluate_registration(source, target, threshold, trans_init)
    return res.inlier_rmse, res.fitness
