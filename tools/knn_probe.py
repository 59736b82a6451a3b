"""GPU probe: hybrid k-NN time on a bench-scene cloud for several grid cell factors (IBL_KNN_CELL_FACTOR)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instance_based_loc_b200 import kernels as K, synth, o3d

scene = synth.make_scene(m=40, seed=1002, D=768)
pts = np.concatenate([synth.memory_cloud(scene, j, 0.01)[0] for j in (0, 1, 2)])
P = torch.from_numpy(pts - pts.mean(0)).cuda()
print("points", P.shape[0], flush=True)
sp = K.surface_spacing(P, 0.05)
ref = {}
for fac in (0.6, 0.5, 0.42, 0.34, 0.28):
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
        same = True
        if (r, k) in ref:
            same = bool(np.array_equal(ref[(r, k)], idx))
        else:
            ref[(r, k)] = idx
        print(f"factor {fac} r={r} k={k} cell={cell:.4f}: {e0.elapsed_time(e1) / 5:.3f} ms  same_as_first={same}", flush=True)
