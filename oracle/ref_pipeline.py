"""oracle/ref_pipeline.py -- TEST INFRASTRUCTURE ONLY (and the timed CPU baseline of bench.py).

Array-level restatement of the reference's per-frame flow: ``_get_object_info`` cloud extraction
(object_memory/object_memory.py:147-154 -> utils/depth_utils.py:176-206), ``register_point_clouds`` /
``evaluate_transform`` (utils/fpfh_register.py:100-150) and the ``localise`` orchestration
(object_memory/object_memory.py:898-1169) on top of ref_port.py (the reference's numpy code, pinned) and
o3d_core.py (Open3D 0.17 semantics, PARITY UNPINNED).  tests/test_oracle_vs_reference.py checks this file
against the UNMODIFIED reference ``ObjectMemory.localise`` running on the oracle's open3d shim.

The RANSAC draw stream is the counter-based sampler (oracle/sampler.py) with an explicit seed -- the product is fed
the same stream.
"""
from __future__ import annotations

import numpy as np
from scipy.spatial.transform import Rotation

from . import o3d_core as oc
from . import ref_port, sampler

MAX_ITERATION = 4000000     # fpfh_register.py:119
CONFIDENCE = 0.99


def detect_clouds(depth, rgb, masks, fx, fy, outlier_cfg):
    """-> list of (pts, cols) after the first remove_radius_outlier (depth_utils.py:85-88)."""
    out = []
    for p, c in ref_port.backproject_masks(depth, rgb, masks, fx, fy):
        if outlier_cfg is not None and len(p):
            keep = oc.remove_radius_outlier_mask(p, outlier_cfg["radius_nb_points"], outlier_cfg["radius"])
            p, c = p[keep], c[keep]
        out.append((p, c))
    return out


def fpfh_features(pts, voxel_size):
    """downsample_and_compute_fpfh (fpfh_register.py:86-98): no downsampling; -> (normals, fpfh [33,N])."""
    nrm = oc.estimate_normals(pts, voxel_size * 2, 30)
    return nrm, oc.compute_fpfh(pts, nrm, voxel_size * 5, 100)


def register_point_clouds(src_p, src_c, tgt_p, tgt_c, voxel_size, global_dist_factor=1.5, local_dist_factor=0.4,
                          seed=0, tgt_feats=None, info=None, max_iteration=MAX_ITERATION, early_exit=True):
    """fpfh_register.py:100-143 -> (T, inlier_rmse, fitness).  ``max_iteration`` / ``early_exit`` only differ from the
    reference's RANSACConvergenceCriteria(4000000, 0.99) in the fixed-hypothesis benchmark configurations."""
    info = info if info is not None else {}
    try:
        if len(src_p) == 0 or len(tgt_p) == 0:
            raise RuntimeError("empty cloud")
        sn, sf = fpfh_features(src_p, voxel_size)
        tn, tf = tgt_feats if tgt_feats is not None else fpfh_features(tgt_p, voxel_size)
        thr = voxel_size * global_dist_factor
        cor = oc.match_features(sf, tf, True, 3)
        C = len(cor)
        res, rinfo = oc.ransac(src_p, tgt_p, cor, lambda a, b: sampler.samples(seed, a, b - a, C), thr, 0.9, CONFIDENCE,
                               early_exit=early_exit, max_iteration=max_iteration)
        info.update(rinfo, n_corres=C, ransac_T=res.transformation, ransac_inliers=res.n_inliers, corres=cor)
        if src_c is None or tgt_c is None:
            raise RuntimeError("ColoredICP requires colored pointclouds.")
        r = oc.icp_colored(src_p, src_c, tgt_p, tn, tgt_c, voxel_size * local_dist_factor, res.transformation)
    except Exception as e:      # the reference's bare except (fpfh_register.py:137)
        info["fallback"] = repr(e)
        r = oc.icp_p2p(src_p, tgt_p, voxel_size * local_dist_factor, np.eye(4))
    info.update(icp_iterations=getattr(r, "iterations", 0), icp_inliers=r.n_inliers)
    return r.transformation, r.inlier_rmse, r.fitness


def evaluate_transform(src_p, tgt_p, T, threshold=0.02):
    r = oc.evaluate_registration(src_p, tgt_p, threshold, T)
    return r.inlier_rmse, r.fitness


def localise_core(detected, detected_embs, memory, outlier_cfg=None, fpfh_global_dist_factor=2,
                  fpfh_local_dist_factor=0.4, fpfh_voxel_size=0.05, max_detected_object_num=7, seed=0,
                  bug_compatible=True, cache=None):
    """object_memory.py:898-1169.  detected: list of (pts, cols); memory: list of dicts {points, colors, embeddings}.
    -> (pose[7], [assn, None], info)."""
    outlier_cfg = outlier_cfg or {"radius_nb_points": 8, "radius": 0.05}
    info = {"assignments": [], "registrations": []}
    if detected is None or len(detected) == 0:
        return np.array([0., 0., 0., 0., 0., 0., 1.]), [[], []], info
    embs = np.array(detected_embs, dtype=np.float32)
    if len(detected) > max_detected_object_num:
        order = sorted(range(len(detected)), key=lambda i: len(detected[i][0]), reverse=True)[:max_detected_object_num]
        detected = [detected[i] for i in order]
        embs = embs[order]
    if len(embs) > len(memory):
        embs = embs[:len(memory)]
    S = ref_port.closest_similarities(embs, [m["embeddings"] for m in memory])
    info["closest_similarities"] = S
    assert S.shape[0] >= min(len(detected), 3) or S.shape[0] == 1
    assns = ref_port.simvolume_assignments(S, 4)
    info["assignments"] = [list(map(list, a)) for a in assns]
    det2 = []
    for p, c in detected:                                                     # :992-998
        if len(p):
            keep = oc.remove_radius_outlier_mask(p, outlier_cfg["radius_nb_points"], outlier_cfg["radius"])
            p, c = p[keep], c[keep]
        det2.append((p, c))
    all_mem = np.concatenate([m["points"] for m in memory])
    all_det = np.concatenate([p for p, _ in det2])
    cache = cache if cache is not None else {}
    data = []
    detected_mean = memory_mean = None
    assn = None
    for assn in assns:
        dp = np.concatenate([det2[d][0] for d, _ in assn])
        dc = np.concatenate([det2[d][1] for d, _ in assn])
        key = tuple(m for _, m in assn)
        if key not in cache:
            mp = np.concatenate([memory[m]["points"] for m in key])
            mc = np.concatenate([memory[m]["colors"] for m in key])
            mm = np.mean(mp, axis=0)
            mp = mp + (-mm)
            try:
                feats = fpfh_features(mp, fpfh_voxel_size)
            except Exception:
                feats = None
            cache[key] = (mp, mc, mm, feats)
        mp, mc, memory_mean, feats = cache[key]
        detected_mean = np.mean(dp, axis=0)
        dp = dp + (-detected_mean)
        rinfo = {}
        T, rmse, fit = register_point_clouds(dp, dc, mp, mc, fpfh_voxel_size, fpfh_global_dist_factor,
                                             fpfh_local_dist_factor, seed=seed, tgt_feats=feats, info=rinfo)
        G = np.array(T)
        G[:3, 3] = T[:3, 3] + memory_mean - T[:3, :3] @ detected_mean
        frmse, ffit = evaluate_transform(all_det, all_mem, G)
        data.append([assn, T, rmse, fit, frmse, ffit, detected_mean, memory_mean])
        info["registrations"].append({"assn": assn, "transform": T, "rmse": rmse, "fitness": fit, "full_rmse": frmse,
                                      "full_fitness": ffit, "detail": rinfo})
    best = sorted(data, key=lambda x: x[5], reverse=True)
    bt, ba = best[0][1], best[0][0]
    if not bug_compatible:
        detected_mean, memory_mean, assn = best[0][6], best[0][7], ba
    R = bt[:3, :3]
    pose = np.concatenate((bt[:3, 3] + memory_mean - R @ detected_mean, Rotation.from_matrix(R).as_quat()))
    info["best_assn"], info["best_transform"] = ba, bt
    return pose, [assn, None], info


def localise_frame(frame, memory, fx, fy, seed=0, **kw):
    """frame: dict {rgb, depth, masks [n,1,H,W], embs}."""
    cfg = kw.get("outlier_cfg") or {"radius_nb_points": 8, "radius": 0.05}
    det = detect_clouds(frame["depth"], frame["rgb"], frame["masks"], fx, fy, cfg)
    return localise_core(det, frame["embs"], memory, seed=seed, **kw)
