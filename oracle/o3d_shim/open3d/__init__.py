"""oracle/o3d_shim/open3d -- TEST INFRASTRUCTURE ONLY (PARITY UNPINNED).

A numpy/scipy stand-in for the slice of ``open3d==0.17.0`` (environment.yml:229) that the
reference's hot path touches, so that the reference's OWN modules (utils/depth_utils.py,
utils/fpfh_register.py, object_memory.localise) can be imported unmodified in this container for
oracle validation and golden-vector generation.  Semantics follow SURVEY.md Appendix A via
oracle/o3d_core.py.  Put ``oracle/o3d_shim`` on sys.path to make ``import open3d`` resolve here.
"""
from __future__ import annotations

import copy as _copy
import types as _types

import numpy as _np

from oracle import o3d_core as _oc

__version__ = "0.17.0-oracle-shim"


class _Vec(_np.ndarray):
    """ndarray subclass standing in for Vector3dVector / Vector2iVector."""

    def __new__(cls, arr=None, dtype=_np.float64, cols=3):
        if arr is None:
            a = _np.zeros((0, cols), dtype)
        else:
            a = _np.array(arr, dtype=dtype).reshape(-1, cols)
        return a.view(cls)


def _Vector3dVector(arr=None):
    return _Vec(arr, _np.float64, 3)


def _Vector2iVector(arr=None):
    return _Vec(arr, _np.int32, 2)


class KDTreeSearchParamHybrid:
    def __init__(self, radius, max_nn):
        self.radius = float(radius)
        self.max_nn = int(max_nn)


class KDTreeSearchParamKNN:
    def __init__(self, knn=30):
        self.radius = _np.inf
        self.max_nn = int(knn)


class PointCloud:
    def __init__(self, points=None):
        self._p = _Vector3dVector(points)
        self._c = _Vector3dVector()
        self._n = _Vector3dVector()

    points = property(lambda s: s._p, lambda s, v: setattr(s, "_p", _Vector3dVector(v)))
    colors = property(lambda s: s._c, lambda s, v: setattr(s, "_c", _Vector3dVector(v)))
    normals = property(lambda s: s._n, lambda s, v: setattr(s, "_n", _Vector3dVector(v)))

    def has_points(self):
        return len(self._p) > 0

    def has_colors(self):
        return len(self._p) > 0 and len(self._c) == len(self._p)

    def has_normals(self):
        return len(self._p) > 0 and len(self._n) == len(self._p)

    def is_empty(self):
        return len(self._p) == 0

    def __len__(self):
        return len(self._p)

    def __add__(self, o):  # A.10
        r = PointCloud()
        r += self
        r += o
        return r

    def __iadd__(self, o):
        if o.is_empty():
            return self
        old = len(self._p)
        if (self.is_empty() or self.has_normals()) and o.has_normals():
            self._n = _Vector3dVector(_np.concatenate([_np.asarray(self._n).reshape(-1, 3)[:old], o._n]))
        else:
            self._n = _Vector3dVector()
        if (self.is_empty() or self.has_colors()) and o.has_colors():
            self._c = _Vector3dVector(_np.concatenate([_np.asarray(self._c).reshape(-1, 3)[:old], o._c]))
        else:
            self._c = _Vector3dVector()
        self._p = _Vector3dVector(_np.concatenate([self._p, o._p]))
        return self

    def translate(self, v, relative=True):
        v = _np.asarray(v, _np.float64).reshape(3)
        if not relative:
            v = v - _np.asarray(self.get_center())
        self._p = _Vector3dVector(_np.asarray(self._p) + v)
        return self

    def transform(self, T):
        T = _np.asarray(T, _np.float64)
        self._p = _Vector3dVector(_oc.transform_points(T, _np.asarray(self._p)))
        if self.has_normals():
            n = _np.asarray(self._n)
            R = T[:3, :3]
            self._n = _Vector3dVector(_np.stack([(R[a, 0] * n[:, 0] + R[a, 1] * n[:, 1]) + R[a, 2] * n[:, 2]
                                                  for a in range(3)], axis=1))
        return self

    def get_center(self):
        return _np.mean(_np.asarray(self._p), axis=0) if len(self._p) else _np.zeros(3)

    def select_by_index(self, idx, invert=False):
        idx = _np.asarray(idx, _np.int64)
        if invert:
            m = _np.ones(len(self._p), bool)
            m[idx] = False
            idx = _np.nonzero(m)[0]
        r = PointCloud(_np.asarray(self._p)[idx])
        if self.has_colors():
            r.colors = _np.asarray(self._c)[idx]
        if self.has_normals():
            r.normals = _np.asarray(self._n)[idx]
        return r

    def remove_radius_outlier(self, nb_points, radius, print_progress=False):  # A.2
        if nb_points < 1 or radius <= 0:
            raise RuntimeError("Illegal input parameters, number of points and radius must be positive")
        keep = _np.nonzero(_oc.remove_radius_outlier_mask(_np.asarray(self._p), nb_points, radius))[0]
        return self.select_by_index(keep), list(keep)

    def cluster_dbscan(self, eps, min_points, print_progress=False):  # PointCloudCluster.cpp (memory build)
        return [int(v) for v in _oc.cluster_dbscan(_np.asarray(self._p), eps, min_points)]

    def estimate_normals(self, search_param=None, fast_normal_computation=True):  # A.3
        sp = search_param or KDTreeSearchParamKNN(30)
        self._n = _Vector3dVector(_oc.estimate_normals(_np.asarray(self._p), sp.radius, sp.max_nn))

    def paint_uniform_color(self, c):
        self._c = _Vector3dVector(_np.tile(_np.asarray(c, _np.float64).reshape(1, 3), (len(self._p), 1)))
        return self

    def __deepcopy__(self, memo):
        r = PointCloud(_np.array(self._p))
        r._c = _Vector3dVector(_np.array(self._c))
        r._n = _Vector3dVector(_np.array(self._n))
        return r


# ------------------------------------------------------------------------------------------------
class Feature:
    def __init__(self):
        self.data = _np.zeros((33, 0))

    def dimension(self):
        return self.data.shape[0]

    def num(self):
        return self.data.shape[1]


def compute_fpfh_feature(pcd, search_param):
    if not pcd.has_normals():
        raise RuntimeError("Failed because input point cloud has no normal.")
    f = Feature()
    f.data = _oc.compute_fpfh(_np.asarray(pcd.points), _np.asarray(pcd.normals), search_param.radius,
                              search_param.max_nn)
    return f


class RegistrationResult:
    def __init__(self, r=None):
        r = r or _oc.RegResult()
        self.transformation = r.transformation
        self.fitness = r.fitness
        self.inlier_rmse = r.inlier_rmse
        self.correspondence_set = r.correspondence_set


class TransformationEstimationPointToPoint:
    def __init__(self, with_scaling=False):
        assert not with_scaling
        self.with_scaling = with_scaling


class TransformationEstimationForColoredICP:
    def __init__(self, lambda_geometric=0.968):
        self.lambda_geometric = lambda_geometric


class CorrespondenceCheckerBasedOnEdgeLength:
    def __init__(self, similarity_threshold=0.9):
        self.similarity_threshold = similarity_threshold


class CorrespondenceCheckerBasedOnDistance:
    def __init__(self, distance_threshold):
        self.distance_threshold = distance_threshold


class RANSACConvergenceCriteria:
    def __init__(self, max_iteration=100000, confidence=0.999):
        self.max_iteration = int(max_iteration)
        self.confidence = confidence


class ICPConvergenceCriteria:
    def __init__(self, relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=30):
        self.relative_fitness = relative_fitness
        self.relative_rmse = relative_rmse
        self.max_iteration = max_iteration


# host-seeded sample source (A.7): tests set this to a callable (C, H) -> int32 [H,3] array, or -> a callable
# (h0, h1) -> int32 [h1-h0,3] block of the stream (oracle/sampler.py)
ransac_sampler = lambda C, H: _np.random.default_rng(0).integers(0, C, (H, 3), dtype=_np.int32)  # noqa: E731
last_ransac_info = {}


def registration_ransac_based_on_feature_matching(source, target, source_feature, target_feature, mutual_filter,
                                                  max_correspondence_distance, estimation_method=None, ransac_n=3,
                                                  checkers=(), criteria=None):
    criteria = criteria or RANSACConvergenceCriteria()
    if ransac_n < 3 or max_correspondence_distance <= 0:
        return RegistrationResult()
    assert ransac_n == 3
    cor = _oc.match_features(source_feature.data, target_feature.data, mutual_filter, ransac_n)
    edge = 0.0
    use_checkers = False
    for c in checkers:
        if isinstance(c, CorrespondenceCheckerBasedOnEdgeLength):
            edge = c.similarity_threshold
            use_checkers = True
        if isinstance(c, CorrespondenceCheckerBasedOnDistance):
            assert c.distance_threshold == max_correspondence_distance
            use_checkers = True
    samples = ransac_sampler(len(cor), criteria.max_iteration)
    res, info = _oc.ransac(_np.asarray(source.points), _np.asarray(target.points), cor, samples,
                           max_correspondence_distance, edge_thr=edge, confidence=criteria.confidence,
                           checkers=use_checkers, max_iteration=criteria.max_iteration)
    last_ransac_info.clear()
    last_ransac_info.update(info, n_corres=len(cor))
    return RegistrationResult(res)


def registration_icp(source, target, max_correspondence_distance, init=None, estimation_method=None, criteria=None):
    init = _np.eye(4) if init is None else _np.asarray(init, _np.float64)
    criteria = criteria or ICPConvergenceCriteria()
    est = estimation_method or TransformationEstimationPointToPoint()
    S, Tg = _np.asarray(source.points), _np.asarray(target.points)
    if isinstance(est, TransformationEstimationForColoredICP):
        return registration_colored_icp(source, target, max_correspondence_distance, init, est, criteria)
    fn = lambda pcd, cor: _oc.p2p_update(pcd, Tg, cor)  # noqa: E731
    return RegistrationResult(_oc.icp(S, Tg, max_correspondence_distance, init, fn, criteria.max_iteration,
                                      criteria.relative_fitness, criteria.relative_rmse))


def registration_colored_icp(source, target, max_correspondence_distance, init=None, estimation_method=None,
                             criteria=None):
    init = _np.eye(4) if init is None else _np.asarray(init, _np.float64)
    criteria = criteria or ICPConvergenceCriteria()
    est = estimation_method or TransformationEstimationForColoredICP()
    if not target.has_normals():
        raise RuntimeError("ColoredICP requires target pointcloud to have normals.")
    if not source.has_colors() or not target.has_colors():
        raise RuntimeError("ColoredICP requires colored pointclouds.")
    S, Sc = _np.asarray(source.points), _np.asarray(source.colors)
    Tg, Tn, Tc = _np.asarray(target.points), _np.asarray(target.normals), _np.asarray(target.colors)
    grad = _oc.color_gradient(Tg, Tn, Tc, max_correspondence_distance)
    fn = lambda pcd, cor: _oc.colored_update(pcd, Sc, Tg, Tn, Tc, grad, cor, est.lambda_geometric)[0]  # noqa: E731
    return RegistrationResult(_oc.icp(S, Tg, max_correspondence_distance, init, fn, criteria.max_iteration,
                                      criteria.relative_fitness, criteria.relative_rmse))


def evaluate_registration(source, target, max_correspondence_distance, transformation=None):
    return RegistrationResult(_oc.evaluate_registration(_np.asarray(source.points), _np.asarray(target.points),
                                                        max_correspondence_distance, transformation))


def _write_point_cloud(path, pcd, **kw):
    from oracle import ply_io
    ply_io.write_ply(path, _np.asarray(pcd.points), _np.asarray(pcd.colors) if pcd.has_colors() else None)
    return True


def _read_point_cloud(path, **kw):
    from oracle import ply_io
    p, c = ply_io.read_ply(path)
    r = PointCloud(p)
    if c is not None:
        r.colors = c
    return r


class OrientedBoundingBox:
    """open3d.geometry.OrientedBoundingBox as utils/IoU_ops.py:117-135 uses it (create_from_points, get_center,
    get_box_points) -- upstream cpp/open3d/geometry/BoundingVolume.cpp (0.17): Qhull convex hull of the points (scipy's
    ConvexHull is the same Qhull with the same 'Qt' option; robust=True adds 'QJ'), mean and covariance of the HULL
    VERTICES from raw cumulants, eigenvectors sorted by descending eigenvalue, third axis = first x second, extents from
    the axis-aligned box of the rotated hull vertices.  PARITY UNPINNED like every Open3D row: the eigenvector signs and the
    hull vertex order (rounding of the sums) are Eigen's / Qhull's upstream; the corner SET does not depend on the signs."""

    def __init__(self):
        self.center = _np.zeros(3)
        self.R = _np.eye(3)
        self.extent = _np.zeros(3)

    @staticmethod
    def create_from_points(points, robust=False):
        from scipy.spatial import ConvexHull
        pts = _np.asarray(points, _np.float64).reshape(-1, 3)
        if len(pts) < 4:
            raise RuntimeError("Qhull needs at least 4 points")
        hull = ConvexHull(pts, qhull_options="QJ" if robust else None)
        hp = pts[hull.vertices]
        n = float(len(hp))
        c = _np.zeros(9)
        for q in hp:                                               # raw cumulants, accumulated point by point
            c[0] += q[0]; c[1] += q[1]; c[2] += q[2]
            c[3] += q[0] * q[0]; c[4] += q[0] * q[1]; c[5] += q[0] * q[2]
            c[6] += q[1] * q[1]; c[7] += q[1] * q[2]; c[8] += q[2] * q[2]
        c /= n
        mean = c[:3].copy()
        cov = _np.empty((3, 3))
        cov[0, 0] = c[3] - c[0] * c[0]; cov[1, 1] = c[6] - c[1] * c[1]; cov[2, 2] = c[8] - c[2] * c[2]
        cov[0, 1] = cov[1, 0] = c[4] - c[0] * c[1]
        cov[0, 2] = cov[2, 0] = c[5] - c[0] * c[2]
        cov[1, 2] = cov[2, 1] = c[7] - c[1] * c[2]
        evals, R = _np.linalg.eigh(cov)                            # ascending, like Eigen's SelfAdjointEigenSolver
        evals, R = evals.copy(), R.copy()
        for a, b in ((1, 0), (2, 0), (2, 1)):                      # the three conditional swaps of the upstream code
            if evals[a] > evals[b]:
                evals[[a, b]] = evals[[b, a]]
                R[:, [a, b]] = R[:, [b, a]]
        R[:, 0] /= _np.linalg.norm(R[:, 0])
        R[:, 1] /= _np.linalg.norm(R[:, 1])
        R[:, 2] = _np.cross(R[:, 0], R[:, 1])
        local = (hp - mean) @ R                                    # R^T (p - mean) per point
        lo, hi = local.min(axis=0), local.max(axis=0)
        box = OrientedBoundingBox()
        box.center = R @ ((lo + hi) * 0.5) + mean
        box.R = R
        box.extent = hi - lo
        return box

    def get_center(self):
        return self.center

    def volume(self):
        return float(self.extent[0] * self.extent[1] * self.extent[2])

    def get_box_points(self):
        x = self.R @ _np.array([self.extent[0] / 2, 0, 0])
        y = self.R @ _np.array([0, self.extent[1] / 2, 0])
        z = self.R @ _np.array([0, 0, self.extent[2] / 2])
        c = self.center
        return _Vector3dVector(_np.stack([c - x - y - z, c + x - y - z, c - x + y - z, c - x - y + z,
                                          c + x + y + z, c - x + y + z, c + x - y + z, c + x + y - z]))


geometry = _types.SimpleNamespace(PointCloud=PointCloud, OrientedBoundingBox=OrientedBoundingBox, KDTreeSearchParamHybrid=KDTreeSearchParamHybrid,
                                  KDTreeSearchParamKNN=KDTreeSearchParamKNN)
utility = _types.SimpleNamespace(Vector3dVector=_Vector3dVector, Vector2iVector=_Vector2iVector)
registration = _types.SimpleNamespace(
    Feature=Feature, compute_fpfh_feature=compute_fpfh_feature, RegistrationResult=RegistrationResult,
    TransformationEstimationPointToPoint=TransformationEstimationPointToPoint,
    TransformationEstimationForColoredICP=TransformationEstimationForColoredICP,
    CorrespondenceCheckerBasedOnEdgeLength=CorrespondenceCheckerBasedOnEdgeLength,
    CorrespondenceCheckerBasedOnDistance=CorrespondenceCheckerBasedOnDistance,
    RANSACConvergenceCriteria=RANSACConvergenceCriteria, ICPConvergenceCriteria=ICPConvergenceCriteria,
    registration_ransac_based_on_feature_matching=registration_ransac_based_on_feature_matching,
    registration_icp=registration_icp, registration_colored_icp=registration_colored_icp,
    evaluate_registration=evaluate_registration)
pipelines = _types.SimpleNamespace(registration=registration)
io = _types.SimpleNamespace(write_point_cloud=_write_point_cloud, read_point_cloud=_read_point_cloud)
