"""open3d.pipelines.registration surface used by utils/fpfh_register.py:86-150, on libibl.so kernels
(SURVEY.md Appendix A.4-A.9).

RANSAC sampling: Open3D 0.17 draws correspondence triplets from a thread-local mt19937 under OpenMP
(nondeterministic).  Here the draw stream is explicit: either the counter-based sampler
``ibl_sample(ransac_seed, h, j, C)`` (csrc/rigid.cuh) or a host-provided int32 [H,3] array returned by
``ransac_sampler(C, H)`` -- the same stream is fed to the oracle in the parity tests.
"""
from __future__ import annotations

import os
import threading

import numpy as np
import torch

from .. import kernels as K
from .pointcloud import DeviceVec, PointCloud

# ---- module-level knobs (not part of open3d) ----------------------------------------------------
ransac_seed = 0                  # seed of the counter-based sample stream
ransac_sampler = None            # optional callable (C, H) -> int32 [H,3] host array (overrides the seed)
ransac_block = 1 << 14           # hypotheses checked per device block before the ordered host scan
ransac_batch_block = int(os.environ.get("IBL_RANSAC_BATCH_BLOCK", 1 << 14))     # first block per problem of the batched RANSAC (grows x4 per round up to 2^20; 2^14 measured best on configs[1])
ransac_max_iteration = None      # benchmark knob: overrides criteria.max_iteration (configs[2]/[4]: a FIXED hypothesis count)
ransac_early_exit = True         # benchmark knob: False disables the adaptive stop (every hypothesis of the stream is drawn)
last_ransac_info = {}            # most recent RANSAC of ANY thread (kept for tools); per-thread copy: last_info()
_tls = threading.local()


def last_info():
    """What the calling thread's most recent registration did: RANSAC {drawn, validated, n_corres, ...} and ICP
    {iterations, ...}.  Thread-local, because localise_batch registers frames on several host threads."""
    d = getattr(_tls, "info", None)
    if d is None:
        d = _tls.info = {}
    return d


class Feature:
    """[33, N] feature like open3d's; the device copy is point-major [N, 33]."""

    def __init__(self, dev=None):
        self._dev = dev
        self._host = None
        self._pack = None

    def pack(self):
        """The matcher's cached operands of this feature set (ibl_feature_pack_create), built on first use.  Worth it for
        a set that is matched many times -- the memory side; match_features uses it whenever it exists."""
        if self._pack is None and self._dev is not None and self._dev.shape[0] > 0:
            self._pack = K.FeaturePack(self._dev)
        return self._pack

    @property
    def data(self):
        if self._host is None:
            self._host = (np.zeros((33, 0)) if self._dev is None else self._dev.cpu().numpy().T.copy())
        return self._host

    @data.setter
    def data(self, v):
        self._host = np.asarray(v, np.float64)
        self._dev = K.to_dev(np.ascontiguousarray(self._host.T))
        self._pack = None

    def dimension(self):
        return 33 if self._dev is None else int(self._dev.shape[1])

    def num(self):
        return 0 if self._dev is None else int(self._dev.shape[0])


def compute_fpfh_feature(pcd, search_param):
    """A.4 ComputeFPFHFeature (fpfh_register.py:94-97)."""
    if not pcd.has_normals():
        raise RuntimeError("Failed because input point cloud has no normal.")
    idx, d2, cnt = pcd.knn(search_param.radius, search_param.max_nn, want_d2=True)
    sp = K.spfh(pcd._p, pcd._n, idx, cnt)
    return Feature(K.fpfh(idx, d2, cnt, sp))


class RegistrationResult:
    def __init__(self, T=None, fitness=0.0, inlier_rmse=0.0, corres=None, info=None):
        self.transformation = np.eye(4) if T is None else np.array(T, dtype=np.float64).reshape(4, 4)
        self.fitness = float(fitness)
        self.inlier_rmse = float(inlier_rmse)
        self._corres = corres
        self.info = info or {}

    @property
    def correspondence_set(self):
        if self._corres is None:
            return np.zeros((0, 2), np.int32)
        return self._corres() if callable(self._corres) else self._corres

    def __repr__(self):
        return (f"RegistrationResult with fitness={self.fitness:e}, inlier_rmse={self.inlier_rmse:e}, "
                f"and correspondence_set size of {self.info.get('n_inliers', 0)}")


class TransformationEstimationPointToPoint:
    def __init__(self, with_scaling=False):
        if with_scaling:
            raise NotImplementedError("with_scaling=True is not used by the reference path")
        self.with_scaling = False


class TransformationEstimationForColoredICP:
    def __init__(self, lambda_geometric=0.968):
        self.lambda_geometric = float(lambda_geometric)


class CorrespondenceCheckerBasedOnEdgeLength:
    def __init__(self, similarity_threshold=0.9):
        self.similarity_threshold = float(similarity_threshold)


class CorrespondenceCheckerBasedOnDistance:
    def __init__(self, distance_threshold):
        self.distance_threshold = float(distance_threshold)


class RANSACConvergenceCriteria:
    def __init__(self, max_iteration=100000, confidence=0.999):
        self.max_iteration = int(max_iteration)
        self.confidence = float(confidence)


class ICPConvergenceCriteria:
    def __init__(self, relative_fitness=1e-6, relative_rmse=1e-6, max_iteration=30):
        self.relative_fitness = float(relative_fitness)
        self.relative_rmse = float(relative_rmse)
        self.max_iteration = int(max_iteration)


# voxel side of the NN candidate field = max_d / nn_voxel_div (clamped below): finer voxels give shorter candidate lists
# per query at the price of a larger (cached, memory-side) table.  Measured on B200 (configs[1]): /8 instead of /4 cuts
# ransac_score time by a third.  IBL_NN_VOXEL_DIV / IBL_NN_VOXEL_MIN override for experiments.
nn_voxel_div = float(os.environ.get("IBL_NN_VOXEL_DIV", "8"))
nn_voxel_min = float(os.environ.get("IBL_NN_VOXEL_MIN", "0.005"))


def nn_voxel(max_d):
    """Voxel side of the exact NN candidate field for 1-NN-within-max_d queries (ibl_grid_build_nnfield)."""
    return max(max_d / nn_voxel_div, nn_voxel_min)


def _nn_grid(target, max_d):
    """Target-side search structure for GetRegistrationResultAndCorrespondences-style queries.  Cached on the
    cloud, so a memory-side target pays for it once, not once per frame."""
    return target.grid(max(max_d, 1e-4), nnfield=max_d, nnfield_voxel=nn_voxel(max_d))


# reverse matching restricted to the rows the mutual filter reads (ibl_feature_match_packed_mutual); IBL_MUTUAL_SUBSET=0
# runs the full reverse direction again (measurement switch)
mutual_reverse_subset = os.environ.get("IBL_MUTUAL_SUBSET", "1") != "0"


def match_features(source_feature, target_feature, mutual_filter=True, ransac_n=3):
    """A.5 -> device int32 [C,2]."""
    fa, fb = source_feature._dev, target_feature._dev
    if mutual_filter and target_feature._pack is not None and fa.shape[0] > 0:
        # the target's packed operands are cached: both directions from one read of the source rows
        # (the reverse direction only for the target rows some source row chose: the mutual filter reads no others)
        ij, ji, flag = (K.feature_match_packed_mutual if mutual_reverse_subset else K.feature_match_packed)(fa, target_feature._pack)
        cor = K.mutual_corres(ij, ji, ransac_n)              # synchronises
        if int(flag.item()) == 0:
            return cor
        # source values outside the pack's fp16 range (never the case for FPFH): the general path below
    ij = K.feature_nn(fa, fb)
    if not mutual_filter:
        ar = torch.arange(fa.shape[0], dtype=torch.int32, device=fa.device)
        return torch.stack([ar, ij], dim=1).contiguous()
    ji = K.feature_nn(fb, fa)
    return K.mutual_corres(ij, ji, ransac_n)


def registration_ransac_based_on_feature_matching(source, target, source_feature, target_feature, mutual_filter,
                                                  max_correspondence_distance, estimation_method=None, ransac_n=3,
                                                  checkers=(), criteria=None):
    """A.5-A.7 (fpfh_register.py:110-119)."""
    criteria = criteria or RANSACConvergenceCriteria()
    if ransac_n < 3 or max_correspondence_distance <= 0:
        return RegistrationResult()
    if ransac_n != 3:
        raise NotImplementedError("ransac_n != 3 is not used by the reference path")
    cor = match_features(source_feature, target_feature, mutual_filter, ransac_n)
    return registration_ransac_based_on_correspondence(source, target, cor, max_correspondence_distance,
                                                       estimation_method, ransac_n, checkers, criteria)


def registration_ransac_based_on_correspondence(source, target, corres, max_correspondence_distance,
                                                estimation_method=None, ransac_n=3, checkers=(), criteria=None):
    criteria = criteria or RANSACConvergenceCriteria()
    if not isinstance(corres, torch.Tensor):
        corres = K.to_dev(np.asarray(corres, np.int32), torch.int32)
    edge, use_checkers = 0.0, False
    for c in checkers:
        if isinstance(c, CorrespondenceCheckerBasedOnEdgeLength):
            edge, use_checkers = c.similarity_threshold, True
        elif isinstance(c, CorrespondenceCheckerBasedOnDistance):
            if c.distance_threshold != max_correspondence_distance:
                raise NotImplementedError("distance checker threshold must equal max_correspondence_distance")
            use_checkers = True
    C = int(corres.shape[0])
    if C < ransac_n:
        last_info()["ransac"] = {"drawn": 0, "validated": 0, "best_hyp": -1, "n_corres": C, "n_inliers": 0, "err2": 0.0}
        return RegistrationResult()
    samples = None
    if ransac_sampler is not None:
        samples = K.to_dev(np.ascontiguousarray(ransac_sampler(C, criteria.max_iteration), np.int32), torch.int32)
    max_it = criteria.max_iteration if ransac_max_iteration is None else int(ransac_max_iteration)
    r = K.ransac_run(_nn_grid(target, max_correspondence_distance), source._p, target._p, corres,
                     max_it, criteria.confidence, max_correspondence_distance, edge_thr=edge,
                     samples=samples, seed=ransac_seed, use_checkers=use_checkers, early_exit=bool(ransac_early_exit),
                     block=ransac_block)
    rinfo = {k: r[k] for k in ("drawn", "validated", "best_hyp", "n_corres", "n_inliers", "err2")}
    last_ransac_info.clear()
    last_ransac_info.update(rinfo)
    last_info()["ransac"] = rinfo
    return RegistrationResult(r["T"], r["fitness"], r["inlier_rmse"], info=r)


def _corres_lazy(grid, src, T, max_d):
    def get():
        _, _, nn, _ = K.nn_score(grid, src, max_d, T=T, want_idx=True)
        nn = nn.cpu().numpy()
        i = np.nonzero(nn >= 0)[0].astype(np.int32)
        return np.stack([i, nn[i]], axis=1)
    return get


def evaluate_registration(source, target, max_correspondence_distance, transformation=None):
    """A.6 (fpfh_register.py:145-150)."""
    T = np.eye(4) if transformation is None else np.asarray(transformation, np.float64)
    if source.is_empty() or target.is_empty() or max_correspondence_distance <= 0:
        return RegistrationResult(T)
    g = _nn_grid(target, max_correspondence_distance)
    Tq = None if np.array_equal(T, np.eye(4)) else T
    n_in, err2, _, _ = K.nn_score(g, source._p, max_correspondence_distance, T=Tq)
    fit = n_in / source.n()
    rmse = float(np.sqrt(err2 / n_in)) if n_in else 0.0
    return RegistrationResult(T, fit, rmse, _corres_lazy(g, source._p, Tq, max_correspondence_distance),
                              {"n_inliers": n_in, "err2": err2})


def registration_icp(source, target, max_correspondence_distance, init=None, estimation_method=None, criteria=None):
    """A.9 (fpfh_register.py:138-141)."""
    est = estimation_method or TransformationEstimationPointToPoint()
    if isinstance(est, TransformationEstimationForColoredICP):
        return registration_colored_icp(source, target, max_correspondence_distance, init, est, criteria)
    criteria = criteria or ICPConvergenceCriteria()
    init = np.eye(4) if init is None else np.asarray(init, np.float64)
    if max_correspondence_distance <= 0:
        raise RuntimeError("Invalid max_correspondence_distance.")
    if source.is_empty() or target.is_empty():
        return RegistrationResult(init)
    g = _nn_grid(target, max_correspondence_distance)
    r = K.icp(g, source._p, target._p, max_correspondence_distance, init, mode=0,
              max_iteration=criteria.max_iteration, rel_fitness=criteria.relative_fitness,
              rel_rmse=criteria.relative_rmse)
    last_info()["icp"] = {"mode": "p2p", "iterations": r["iterations"], "n_inliers": r["n_inliers"]}
    return RegistrationResult(r["T"], r["fitness"], r["inlier_rmse"], info=r)


def color_gradient(target, max_correspondence_distance):
    """A.8 InitializePointCloudForColoredICP: Hybrid(2*max_d, 30) neighbourhoods."""
    idx, _, cnt = target.knn(max_correspondence_distance * 2.0, 30, want_d2=False)
    return K.color_gradient(target._p, target._n, target._c, idx, cnt)


def ensure_color_gradient(target, max_correspondence_distance):
    """colour gradients of a target, cached on the cloud (they depend only on the target and max_d)."""
    grad = getattr(target, "_color_grad", None)
    if grad is None or grad[0] != max_correspondence_distance or grad[1].shape[0] != target.n():
        grad = (max_correspondence_distance, color_gradient(target, max_correspondence_distance))
        target._color_grad = grad
    return grad


def registration_colored_icp(source, target, max_correspondence_distance, init=None, estimation_method=None,
                             criteria=None):
    """A.8 (fpfh_register.py:132-135).  Raises like open3d when normals / colours are missing, which is what
    sends the reference into its point-to-point fallback."""
    criteria = criteria or ICPConvergenceCriteria()
    est = estimation_method or TransformationEstimationForColoredICP()
    init = np.eye(4) if init is None else np.asarray(init, np.float64)
    if not target.has_normals():
        raise RuntimeError("ColoredICP requires target pointcloud to have normals.")
    if not source.has_colors() or not target.has_colors():
        raise RuntimeError("ColoredICP requires colored pointclouds.")
    if max_correspondence_distance <= 0:
        raise RuntimeError("Invalid max_correspondence_distance.")
    grad = ensure_color_gradient(target, max_correspondence_distance)
    g = _nn_grid(target, max_correspondence_distance)
    r = K.icp(g, source._p, target._p, max_correspondence_distance, init, mode=1, scol=source._c, tnrm=target._n,
              tcol=target._c, tgrad=grad[1], max_iteration=criteria.max_iteration,
              rel_fitness=criteria.relative_fitness, rel_rmse=criteria.relative_rmse,
              lambda_geometric=est.lambda_geometric)
    last_info()["icp"] = {"mode": "colored", "iterations": r["iterations"], "n_inliers": r["n_inliers"]}
    return RegistrationResult(r["T"], r["fitness"], r["inlier_rmse"], info=r)
