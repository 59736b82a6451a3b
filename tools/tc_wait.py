#!/usr/bin/env python
"""Who paces the tcgen05 matcher kernel on the bench workload?  Runs configs[1] frames through localise_batch (one frame in
flight) with the kernel's wait-clock counters on (ibl_feature_nn_wait_stats) and prints, per streamed target tile, the MMA
thread's clocks and what it waited for."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from instance_based_loc_b200 import kernels as K, selfcheck, workloads as WL  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    cfg = WL.CONFIGS[1]
    WL.apply_ransac_mode(cfg)
    kw = WL.localise_kwargs(cfg)
    scene = WL.make_scene(cfg)
    frames = WL.make_frames(cfg, scene, n, rank=0, device="cuda")
    om, _ = selfcheck.build_memory(scene, cfg["mem_spacing"], cfg["mem_voxel"], f=cfg["f"])
    dev = torch.device("cuda")
    res = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in ("rgb", "depth", "masks", "embs")} for fr in frames]
    om.localise_batch(res, workers=1, **kw)                      # warm: memory-side packs cached
    K.feature_nn_wait_stats(start=True)
    K.feature_nn_stats(True)
    om.localise_batch(res, workers=1, **kw)
    w = K.feature_nn_wait_stats()
    cand, evals = K.feature_nn_stats(True)
    nreg = sum(len(i["registrations"]) for i in om.last_batch_infos) if hasattr(om, "last_batch_infos") else 0
    t = max(w["tiles"], 1)
    out = dict(w)
    out.update(frames=n, tiles_per_cta=w["tiles"] / max(w["ctas"], 1),
               mma_clk_per_tile=w["mma_loop_clk"] / t, mma_wait_accumulator_per_tile=w["mma_wait_accumulator_clk"] / t,
               mma_wait_tile_per_tile=w["mma_wait_tile_clk"] / t, epilogue_clk_per_tile=w["epilogue_loop_clk"] / t,
               epilogue_wait_per_tile=w["epilogue_wait_accumulator_clk"] / t,
               producer_wait_per_tile=w["producer_wait_stage_clk"] / t,
               mma_floor_clk_per_tile=2 * 7 * 64, band_candidates=cand, exact_evaluations=evals,
               query_rows_upper_bound=w["ctas"] * 256, candidates_per_row_lower_bound=cand / max(w["ctas"] * 256, 1),
               candidates_per_row_tile=cand / max(w["tiles"] * 256, 1))
    print(json.dumps(out))


main()
