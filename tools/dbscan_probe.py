"""Times ibl_dbscan (grid build separately) on synthetic surface clouds with CUDA events on the launching stream, after
a warm-up call: python tools/dbscan_probe.py [n_points] -> one JSON line per (eps, min_points) setting."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ibl_testutil import surface_cloud  # noqa: E402
from instance_based_loc_b200 import kernels as K  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    offs = ([0, 0, 0], [1.3, 0, 0], [4, 0, 0], [4, 3, 0], [-4, 1, 2])
    pts = np.concatenate([surface_cloud(n // 5, 100 + k, scale=0.6)[0] + np.asarray(o, float) for k, o in enumerate(offs)])
    pts = K.to_dev(pts[np.random.default_rng(0).permutation(len(pts))])
    for eps, mp in ((0.2, 300), (0.4, 150), (0.05, 12)):
        grid = K.Grid(pts, eps * 1.001)
        cnt = K.radius_count(grid, eps)
        K.dbscan(grid, eps, mp)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        torch.cuda.synchronize()
        ev[0].record()
        reps = 3
        for _ in range(reps):
            lab, nc = K.dbscan(grid, eps, mp)
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / reps
        pairs = int(cnt.to(torch.int64).sum().item())
        print(json.dumps({"probe": "ibl_dbscan", "points": int(pts.shape[0]), "eps": eps, "min_points": mp, "clusters": nc,
                          "noise": int((lab < 0).sum().item()), "neighbour_pairs": pairs, "ms": round(ms, 3),
                          "pair_tests_per_s": round(2.0 * pairs / (ms * 1e-3), 1)}))


if __name__ == "__main__":
    main()
