"""GPU probe: hybrid k-NN time on a DETECTED cloud of the bench workload (union of the 3 largest detections of frame 0,
as register_point_clouds sees it) for several grid cell factors (kernels.knn_cell_factor)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from instance_based_loc_b200 import kernels as K
from instance_based_loc_b200.utils.depth_utils import get_mask_coloured_pointclouds_from_depth

w = bench.WORKLOAD
scene, frames = bench.make_inputs(2, 0, "cuda")
for fi, fr in enumerate(frames):
    clouds = get_mask_coloured_pointclouds_from_depth(fr["depth"], fr["rgb"], fr["masks"], w["f"], w["f"],
                                                      {"radius_nb_points": 8, "radius": 0.05})
    clouds = sorted(clouds, key=lambda c: c.n(), reverse=True)[:3]
    P = torch.cat([c._p for c in clouds]).contiguous()
    P = P - P.mean(0, keepdim=True)
    print("frame", fi, "points", P.shape[0], flush=True)
    sp = K.surface_spacing(P, 0.05)
    ref = {}
    for fac in (0.6, 0.8, 1.0, 1.25, 1.5, 2.0, 2.5):
        K.knn_cell_factor = fac
        for (r, k) in ((0.25, 100), (0.1, 30)):
            cell = K.knn_cell_from_spacing(sp, k, r)
            g = K.Grid(P, cell)
            out = K.knn_hybrid(g, r, k)
            for _ in range(2):
                K.knn_hybrid(g, r, k)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                K.knn_hybrid(g, r, k)
            e1.record(); torch.cuda.synchronize()
            idx = out[0].cpu().numpy()
            same = bool(np.array_equal(ref[(r, k)], idx)) if (r, k) in ref else True
            ref.setdefault((r, k), idx)
            print(f"factor {fac} r={r} k={k} spacing={sp:.4f} cell={cell:.4f}: {e0.elapsed_time(e1) / 5:.3f} ms  same_as_first={same}", flush=True)
