#!/usr/bin/env python
"""Where does a localise_batch step spend its wall time?  Runs configs[1] steps for several (workers, pipeline_chunks)
settings and prints wall ms per step together with the main thread's wait for stage 1 and its time in stage 2
(ObjectMemory.last_batch_timing).  Usage: python tools/step_profile.py [--steps 4] [--settings 4x1,4x2,6x2,8x2,8x4]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from instance_based_loc_b200 import selfcheck, workloads as WL  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--settings", default="4x1,4x2,6x2,8x2,8x4")
    args = ap.parse_args()
    cfg = WL.CONFIGS[1]
    WL.apply_ransac_mode(cfg)
    kw = WL.localise_kwargs(cfg)
    scene = WL.make_scene(cfg)
    frames = WL.make_frames(cfg, scene, 64, rank=0, device="cuda")
    om, _ = selfcheck.build_memory(scene, cfg["mem_spacing"], cfg["mem_voxel"], f=cfg["f"])
    dev = torch.device("cuda")
    res = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in ("rgb", "depth", "masks", "embs")} for fr in frames]
    B = args.frames
    for i in range(0, 64, B):                                   # warm every assignment tuple of the pool
        om.localise_batch(res[i:i + B], workers=4, **kw)
    for s in args.settings.split(","):
        w, c = [int(x) for x in s.split("x")]
        om.localise_batch(res[:B], workers=w, pipeline_chunks=c, **kw)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tim = {"stage1_wait_s": 0.0, "stage2_s": 0.0}
        for k in range(args.steps):
            i = (k * B) % 64
            om.localise_batch(res[i:i + B], workers=w, pipeline_chunks=c, **kw)
            for key in tim:
                tim[key] += om.last_batch_timing[key]
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / args.steps
        print(json.dumps({"workers": w, "chunks": c, "ms_per_step": round(wall * 1e3, 1), "frames_per_s": round(B / wall, 1),
                          "stage1_wait_ms": round(tim["stage1_wait_s"] / args.steps * 1e3, 1),
                          "stage2_ms": round(tim["stage2_s"] / args.steps * 1e3, 1)}), flush=True)


if __name__ == "__main__":
    main()
