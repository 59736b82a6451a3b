"""Diagnostic (GPU box): wall time and per-kernel device time of resident vs host-input steps of bench.py."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from instance_based_loc_b200 import kernels as K, selfcheck
from instance_based_loc_b200.o3d import registration as reg
w = bench.WORKLOAD
B = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
scene, frames = bench.make_inputs(B, 0, "cuda")
om, host_mem = selfcheck.build_memory(scene, w["mem_spacing"], w["mem_voxel"], f=w["f"])
reg.ransac_seed = w["ransac_seed"]
keys = ("rgb", "depth", "masks", "embs")
resident = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in keys} for fr in frames]
pinned = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).pin_memory() for k in keys} for fr in frames]
K.prof_enable(True)
for rep in range(3):
    for mode in ("resident", "host"):
        K.prof_read()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        per = []
        for i in range(B):
            t1 = time.perf_counter()
            fr = resident[i] if mode == "resident" else {k: v.to(dev, non_blocking=True) for k, v in pinned[i].items()}
            om.localise_arrays(fr["rgb"], fr["depth"], fr["masks"], fr["embs"], **bench.LOCALISE_KW)
            per.append(round((time.perf_counter() - t1) * 1e3, 1))
        torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3
        p = K.prof_read()
        top = sorted(((v["ms"], k, v["launches"]) for k, v in p.items()), reverse=True)[:6]
        print(rep, mode, f"{ms:.1f} ms", per, [(k, round(m, 1), n) for m, k, n in top], flush=True)
for workers in (1, 2, 4, 8):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    res = om.localise_batch(resident * (8 // B), workers=workers, **bench.LOCALISE_KW)
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3
    print("batch of", len(res), "workers", workers, f"{ms:.1f} ms  -> {len(res) / ms * 1e3:.2f} frames/s", flush=True)
ref = [om.localise_arrays(fr["rgb"], fr["depth"], fr["masks"], fr["embs"], **bench.LOCALISE_KW)[0] for fr in resident]
print("batch == sequential:", all(np.array_equal(a, r[0]) for a, r in zip(ref, res)))
info = om.last_localise_info
print([ (r["assn"], round(r["fitness"],3), round(r["full_fitness"],3)) for r in info["registrations"]])
