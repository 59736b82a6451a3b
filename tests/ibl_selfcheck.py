"""The checker behind __graft_entry__.smoke() and tests/test_pipeline_gpu.py: runs the product path
(ObjectMemory.localise_arrays on cuda) and the CPU oracle (oracle/ref_pipeline.py) on the same synthetic frame and
compares them.  Lives under tests/ because only tests/, smoke() and bench.py's CPU legs may touch oracle/."""
from __future__ import annotations

import numpy as np
import torch

from instance_based_loc_b200.selfcheck import build_memory, pose_matrix, rot_deg


def compare_frame(om, host_memory, frame, f, seed=0, **kw):
    """-> dict of parity facts for one frame."""
    from oracle import ref_pipeline
    from instance_based_loc_b200.o3d import registration as reg
    reg.ransac_seed = seed
    pose, assn = om.localise_arrays(frame["rgb"], frame["depth"], frame["masks"], frame["embs"],
                                    perform_semantic_icp=False, **kw)
    gi = om.last_localise_info
    okw = {k: v for k, v in kw.items() if k.startswith("fpfh_")}
    rpose, rassn, ri = ref_pipeline.localise_frame(frame, host_memory, f, f, seed=seed, **okw)
    rep = {"assignments_equal": gi["assignments"] == ri["assignments"], "n_assn": len(ri["assignments"]),
           "assn_equal": assn[0] == rassn[0]}
    S, rS = gi["closest_similarities"].cpu().numpy(), ri["closest_similarities"]
    rep["sim_maxdiff"] = float(np.abs(S - rS).max())
    rot, tr, fit = 0.0, 0.0, True
    for a, b in zip(gi["registrations"], ri["registrations"]):
        rot = max(rot, rot_deg(a["transform"][:3, :3], b["transform"][:3, :3]))
        tr = max(tr, float(np.linalg.norm(a["transform"][:3, 3] - b["transform"][:3, 3])))
        fit = fit and abs(a["full_fitness"] - b["full_fitness"]) < 1e-12 and abs(a["fitness"] - b["fitness"]) < 1e-12
    rep.update(reg_rot_deg=rot, reg_trans_m=tr, fitness_equal=bool(fit))
    Tp, Tr = pose_matrix(pose), pose_matrix(rpose)
    rep["pose_rot_deg"] = rot_deg(Tp[:3, :3], Tr[:3, :3])
    rep["pose_trans_m"] = float(np.linalg.norm(Tp[:3, 3] - Tr[:3, 3]))
    Tg = frame["T_wc"]
    rep["gt_rot_deg"] = rot_deg(Tp[:3, :3], Tg[:3, :3])
    rep["gt_trans_m"] = float(np.linalg.norm(Tp[:3, 3] - Tg[:3, 3]))
    rep["ok"] = bool(rep["assignments_equal"] and rep["assn_equal"] and rep["fitness_equal"] and
                     rep["pose_rot_deg"] < 0.1 and rep["pose_trans_m"] < 1e-3)
    return rep


def run_smoke(device="cuda:0", n_points=1500, seed=7):
    from instance_based_loc_b200 import synth
    torch.cuda.set_device(torch.device(device))
    scene = synth.make_scene(m=6, seed=seed, D=64, size=(0.35, 0.7))
    om, host = build_memory(scene, spacing=0.02, voxel=0.02, f=120.0)
    frame = synth.make_frames(scene, 1, H=180, W=240, f=120.0, device="cpu", seed=seed, min_pixels=150)[0]
    return compare_frame(om, host, frame, 120.0, seed=seed, fpfh_global_dist_factor=1.5, fpfh_local_dist_factor=1.5,
                         fpfh_voxel_size=0.05)
