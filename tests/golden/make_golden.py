"""Generates tests/golden/ref_golden.npz by running the UNMODIFIED reference modules
(/root/reference/utils/{depth_utils,similarity_volume,fpfh_register}.py) in the build container.

The reference tree does not exist on the GPU box, so the outputs are committed as small fixtures;
this script is the provenance.  Run:  python tests/golden/make_golden.py

What is pinned here is the reference's OWN numpy code (a1-a3, a7, get_SVD_transform).  The Open3D
0.17 calls (a5, a8-a11) cannot be run here (open3d is not installable) -- see oracle/__init__.py.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ply_io, ref_loader  # noqa: E402


def main():
    import torch
    du = ref_loader.load("utils/depth_utils.py")
    sv = ref_loader.load("utils/similarity_volume.py")
    fr = ref_loader.load("utils/fpfh_register.py")
    o3d = ref_loader.open3d()
    out = {}
    rng = np.random.default_rng(20240)

    # ---- a1/a2: back-projection, float32 metres (8room/synthetic convention) and f64 (TUM /5000.)
    H, W = 48, 64
    depth = rng.uniform(0.4, 4.0, (H, W)).astype(np.float32)
    depth[rng.random((H, W)) < 0.25] = 0.0
    rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    masks = np.zeros((3, 1, H, W), bool)
    masks[0, 0, 5:30, 4:40] = True
    masks[1, 0, 20:48, 30:64] = True
    masks[2, 0] = rng.random((H, W)) < 0.5
    out["bp_depth"], out["bp_rgb"], out["bp_masks"] = depth, rgb, masks
    out["bp_fx"], out["bp_fy"] = np.float64(52.5), np.float64(48.0)
    clouds = du.get_mask_coloured_pointclouds_from_depth(depth, rgb, torch.from_numpy(masks), 52.5, 48.0, None)
    out["bp_offsets"] = np.cumsum([0] + [len(c.points) for c in clouds]).astype(np.int64)
    out["bp_points"] = np.concatenate([np.asarray(c.points) for c in clouds])
    out["bp_colors"] = np.concatenate([np.asarray(c.colors) for c in clouds])
    d16 = (depth * 5000).astype(np.uint16)
    depth64 = d16 / 5000.0
    c64 = du.get_coloured_pointcloud_from_depth(depth64, rgb, 52.5, 48.0, None)
    out["bp64_depth"] = depth64
    out["bp64_points"], out["bp64_colors"] = np.asarray(c64.points), np.asarray(c64.colors)

    # ---- a3: voxel_down_sample_with_colors
    pts = rng.normal(0, 0.03, (2500, 3)) + rng.integers(-2, 3, (2500, 1)) * 0.02
    cols = rng.random((2500, 3))
    pc = o3d.geometry.PointCloud()
    pc.points = o3d.utility.Vector3dVector(pts)
    pc.colors = o3d.utility.Vector3dVector(cols)
    ds = du.voxel_down_sample_with_colors(pc, 0.01)
    out["vox_pts"], out["vox_cols"], out["vox_size"] = pts, cols, np.float64(0.01)
    out["vox_out_pts"], out["vox_out_cols"] = np.asarray(ds.points), np.asarray(ds.colors)
    out["vox_keys"] = np.floor(pts / 0.01).astype(np.int64)

    # ---- a7: SimVolume assignments (reference prints a tqdm bar: silence it)
    cases = [(1, 5), (2, 4), (3, 3), (3, 7), (4, 6), (5, 9), (7, 12), (3, 2), (2, 1)]
    for ci, (nd, m) in enumerate(cases):
        S = rng.uniform(-0.3, 0.98, (nd, m)).astype(np.float32)
        if ci == 4:
            S[1] = S[0]          # exact ties between detections
            S[:, 3] = S[:, 2]    # and between memory objects
        with contextlib.redirect_stderr(io.StringIO()):
            v = sv.SimVolume(S.copy())
            v.fast_construct_volume(min(nd, 3))
            subs = [np.array(s, dtype=np.float16) for s in v.subvolumes]
            a = v.get_top_indices_from_subvolumes(num_per_length=4)
        out[f"sv{ci}_S"] = S
        out[f"sv{ci}_sub0"] = subs[0]
        flat = [[int(x) for pair in assn for x in pair] for assn in a]
        out[f"sv{ci}_assn_len"] = np.array([len(f) // 2 for f in flat], np.int32)
        out[f"sv{ci}_assn_flat"] = np.array([x for f in flat for x in f], np.int32)
    out["sv_ncases"] = np.int32(len(cases))
    # the reference's own __main__ smoke (similarity_volume.py:418-452): cs[i,j] = i + j, 10 x 4
    cs = np.add.outer(np.arange(10), np.arange(4)).astype(np.float64)
    with contextlib.redirect_stderr(io.StringIO()):
        v = sv.SimVolume(cs.copy())
        v.fast_construct_volume(3)
        a = v.get_top_indices_from_subvolumes(3)
    out["svmain_assn_len"] = np.array([len(x) for x in a], np.int32)
    out["svmain_assn_flat"] = np.array([int(y) for x in a for pair in x for y in pair], np.int32)

    # ---- get_SVD_transform known answers (test.py:1-60)
    p = [np.eye(3), np.eye(3), np.array([[1., 2, 3], [4, 5, 6], [7, 8, 9]])]
    q = [np.array([[0., 1, 0], [-1, 0, 0], [0, 0, 1]]), np.array([[1., 1, 0], [0, 0, 0], [-1, -1, 1]]), p[2] + 0.5]
    for k in range(3):
        out[f"svd{k}_p"], out[f"svd{k}_q"] = p[k], q[k]
        with contextlib.redirect_stdout(io.StringIO()):
            out[f"svd{k}_T"] = fr.get_SVD_transform(p[k], q[k])

    # ---- a real Open3D-written memory cloud (inputs only; out/360_trial_with_floor/objects/0), subsampled
    ply = os.path.join(ref_loader.REF, "out/360_trial_with_floor/objects/0/pointcloud.ply")
    P, C = ply_io.read_ply(ply)
    sel = np.sort(rng.choice(len(P), 4000, replace=False))
    out["fix_points"], out["fix_colors"] = P[sel], C[sel]
    out["fix_npoints_full"] = np.int64(len(P))

    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz"), **out)
    print("wrote ref_golden.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
