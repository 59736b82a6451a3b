"""Times ibl_raycast_scene (600x600 frame, 40-object scene = configs[1]'s frame shape) with CUDA events on the launching
stream after a warm-up: python tools/raycast_probe.py -> one JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from instance_based_loc_b200 import kernels as K, synth  # noqa: E402


def main():
    sc = synth.make_scene(m=40, seed=1002, D=16)
    T = synth.camera_pose(sc, np.random.default_rng(1))
    H = W = 600
    args = (T, H, W, 300.0, K.to_dev(sc.kind, torch.int32), K.to_dev(sc.half), K.to_dev(sc.R), K.to_dev(sc.t), sc.room, sc.height)
    K.raycast_scene(*args)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    reps = 20
    ev[0].record()
    for _ in range(reps):
        K.raycast_scene(*args)
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1]) / reps
    print(json.dumps({"probe": "ibl_raycast_scene", "H": H, "W": W, "objects": 40, "ms": round(ms, 4),
                      "ray_object_tests_per_s": round(H * W * 46 / (ms * 1e-3), 1),
                      "output_GBps": round(44.0 * H * W / (ms * 1e-3) / 1e9, 2)}))


if __name__ == "__main__":
    main()
