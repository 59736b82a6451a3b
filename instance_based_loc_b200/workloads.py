"""The five benchmark configurations of BASELINE.json / SURVEY.md section 8(d), as data: scene, frame and memory
parameters, the localise() arguments and the RANSAC mode.  bench.py, the parity tests and the tools all build their
inputs from here, so "configs[1]" means the same arrays everywhere.  Nothing in this module touches oracle/.

  0  single TUM-shaped 640x480 frame (fx = fy = 525) against a 20-object memory          -- single-frame latency
  1  8room-style batch: 100 distinct 600x600 (f = 300) frames against a 40-object memory -- the headline (driver default)
  2  real-scale: 1280x960 frames whose instance clouds hold ~50 k points, 1e5 hypotheses FIXED (no adaptive stop)
  3  HM3D-scale: 500-object memory (~5e6 points), frames sharded over the ranks
  4  RANSAC / ICP microbenchmark sweep: 1e4-1e6 hypotheses x 10 k-200 k points (hypothesis-block sharding over ranks)
"""
from __future__ import annotations

import numpy as np

SUCCESS_TRANS_M = 0.6      # the reference's success criterion (8room_localisation_trial.py:213)
SUCCESS_ROT_RAD = 0.3

_COMMON = dict(D=768, fpfh_voxel_size=0.05, fpfh_global_dist_factor=1.5, fpfh_local_dist_factor=1.5,
               ransac_max_iteration=None, ransac_early_exit=True, size=(0.3, 0.9), max_objects=12, min_pixels=300)

CONFIGS = {
    0: dict(_COMMON, name="single TUM-shaped 640x480 f=525 frame vs 20-object memory (configs[0])", m=20, H=480, W=640,
            f=525.0, scene_seed=1001, ransac_seed=1001, mem_spacing=0.012, mem_voxel=0.005, pool_frames=1,
            frames_per_step=1, workers=1),
    1: dict(_COMMON, name="8room-style: 600x600 f=300 query frames vs 40-object memory (configs[1])", m=40, H=600, W=600,
            f=300.0, scene_seed=1002, ransac_seed=1002, mem_spacing=0.01, mem_voxel=0.01, pool_frames=100,
            frames_per_step=16, workers=4),
    2: dict(_COMMON, name="real-scale: 1280x960 f=1050 frames, ~50k-point instance clouds, 1e5 RANSAC hypotheses fixed "
                          "(configs[2])", m=40, H=960, W=1280, f=1050.0, scene_seed=1003, ransac_seed=1003,
            mem_spacing=0.006, mem_voxel=0.005, pool_frames=10, frames_per_step=2, workers=2, max_objects=7,
            min_pixels=2000, ransac_max_iteration=100000, ransac_early_exit=False),
    3: dict(_COMMON, name="HM3D-scale: 600x600 f=300 query frames vs 500-object memory (configs[3])", m=500, H=600, W=600,
            f=300.0, scene_seed=1004, ransac_seed=1004, mem_spacing=0.015, mem_voxel=0.01, pool_frames=64,
            frames_per_step=16, workers=4),
    4: dict(name="RANSAC/ICP microbench sweep: 1e4-1e6 hypotheses x 10k-200k points per cloud (configs[4])",
            points=(10000, 50000, 100000, 200000), hypotheses=(10000, 100000, 1000000), outliers=0.3, seed=1005,
            fpfh_voxel_size=0.05, max_d=0.075),
}


def localise_kwargs(cfg):
    return dict(fpfh_global_dist_factor=cfg["fpfh_global_dist_factor"], fpfh_local_dist_factor=cfg["fpfh_local_dist_factor"],
                fpfh_voxel_size=cfg["fpfh_voxel_size"], perform_semantic_icp=False)


def apply_ransac_mode(cfg):
    """Set the registration module's benchmark knobs for this configuration."""
    from .o3d import registration as reg
    reg.ransac_seed = cfg.get("ransac_seed", 0)
    reg.ransac_max_iteration = cfg.get("ransac_max_iteration")
    reg.ransac_early_exit = cfg.get("ransac_early_exit", True)


def make_scene(cfg):
    from . import synth
    return synth.make_scene(m=cfg["m"], seed=cfg["scene_seed"], D=cfg["D"], size=cfg["size"])


def make_frames(cfg, scene, n_frames, rank=0, device="cpu"):
    """A pool of distinct frames (host arrays) for this rank; on a CUDA device the first hits come from the
    ibl_raycast_scene kernel."""
    from . import synth
    return synth.make_frames(scene, n_frames, H=cfg["H"], W=cfg["W"], f=cfg["f"], device=device, seed=100 + rank,
                             max_objects=cfg["max_objects"], min_pixels=cfg["min_pixels"])


def pose_success(pose, T_wc):
    """(translation error m, rotation error rad, success by the reference's criterion) of a 7-vector pose against the
    true camera-to-world transform."""
    from scipy.spatial.transform import Rotation
    R = Rotation.from_quat(pose[3:]).as_matrix()
    dt = float(np.linalg.norm(np.asarray(pose[:3]) - T_wc[:3, 3]))
    c = (np.trace(R.T @ T_wc[:3, :3]) - 1.0) / 2.0
    dr = float(np.arccos(np.clip(c, -1.0, 1.0)))
    return dt, dr, bool(dt < SUCCESS_TRANS_M and dr < SUCCESS_ROT_RAD)


# ------------------------------------------------------------------------------------------------ configs[4]
def microbench_pair(n_points, outliers=0.3, seed=1005, noise=0.002):
    """Random surface cloud of ~n_points (a union of boxes / ellipsoids, sampled like a memory object), a rigidly moved
    noisy copy as the target, and ``outliers`` of the SOURCE replaced by uniform clutter -> (src_p, src_c, tgt_p, tgt_c,
    T_gt), host float64."""
    from . import synth
    rng = np.random.default_rng(seed + n_points)
    scene = synth.make_scene(m=3, seed=seed, D=8, size=(0.6, 1.2))
    area = sum(synth._prim_area(scene.kind[q], scene.half[q], False) for q in range(scene.n_prim))
    spacing = float(np.sqrt(area / (n_points * 1.15)))
    P, C = zip(*[synth.memory_cloud(scene, j, spacing) for j in range(scene.m)])
    p, c = np.concatenate(P), np.concatenate(C)
    sel = rng.permutation(len(p))[:n_points]
    sel.sort()
    p, c = p[sel], c[sel]
    p = p - p.mean(0)
    ang = rng.uniform(-np.pi, np.pi, 3)
    cx, sx, cy, sy, cz, sz = np.cos(ang[0]), np.sin(ang[0]), np.cos(ang[1]), np.sin(ang[1]), np.cos(ang[2]), np.sin(ang[2])
    R = np.array([[cz * cy, cz * sy * sx - sz * cx, cz * sy * cx + sz * sx],
                  [sz * cy, sz * sy * sx + cz * cx, sz * sy * cx - cz * sx],
                  [-sy, cy * sx, cy * cx]])
    T = np.eye(4)
    T[:3, :3], T[:3, 3] = R, rng.uniform(-0.5, 0.5, 3)
    tgt = p @ R.T + T[:3, 3] + rng.normal(0, noise, p.shape)
    src = p.copy()
    n_out = int(outliers * len(src))
    bad = rng.permutation(len(src))[:n_out]
    lo, hi = src.min(0) - 0.3, src.max(0) + 0.3
    src[bad] = rng.uniform(lo, hi, (n_out, 3))
    sc = c.copy()
    sc[bad] = rng.uniform(0, 1, (n_out, 3))
    return np.ascontiguousarray(src), np.ascontiguousarray(sc), np.ascontiguousarray(tgt), np.ascontiguousarray(c), T
