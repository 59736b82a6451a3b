"""Diagnostic (GPU box): per-batch wall times of localise_batch at 4 frames in flight, no nvidia-smi running, with GC pauses
timed -- looks for sporadic stalls."""
import gc, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from instance_based_loc_b200 import kernels as K, selfcheck
from instance_based_loc_b200.o3d import registration as reg
w = bench.WORKLOAD
dev = torch.device("cuda", 0)
B = 8
scene, frames = bench.make_inputs(B, 0, "cuda")
om, host_mem = selfcheck.build_memory(scene, w["mem_spacing"], w["mem_voxel"], f=w["f"])
reg.ransac_seed = w["ransac_seed"]
keys = ("rgb", "depth", "masks", "embs")
resident = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in keys} for fr in frames]
gc_t = {"t0": 0.0, "tot": 0.0, "n": 0, "max": 0.0}
def cb(phase, info):
    if phase == "start":
        gc_t["t0"] = time.perf_counter()
    else:
        d = time.perf_counter() - gc_t["t0"]
        gc_t["tot"] += d; gc_t["n"] += 1; gc_t["max"] = max(gc_t["max"], d)
gc.callbacks.append(cb)
print("pool MB after build", {k: v >> 20 for k, v in K.scratch_stats().items()}, flush=True)
for mode in ("gc off", "gc off again"):
    gc.collect(); gc.disable()
    print("pool MB", mode, {k: v >> 20 for k, v in K.scratch_stats().items()}, flush=True)
    om.localise_batch(resident, workers=4, **bench.LOCALISE_KW)
    torch.cuda.synchronize()
    times = []
    gc_t.update(tot=0.0, n=0, max=0.0)
    for rep in range(60):
        r0, c0, t0 = K.scratch_stats()["reserved"], time.process_time(), time.perf_counter()
        om.localise_batch(resident, workers=4, **bench.LOCALISE_KW)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        times.append(dt)
        if len(times) > 5 and dt > 1.25 * np.median(times):
            print(f"   SLOW batch {rep}: {dt:.0f} ms, process cpu {1e3 * (time.process_time() - c0):.0f} ms, pool reserved "
                  f"{r0 >> 20} -> {K.scratch_stats()['reserved'] >> 20} MB, torch reserved {torch.cuda.memory_reserved() >> 20} MB, "
                  f"loadavg {os.getloadavg()}", flush=True)
    t = np.array(times)
    print("   slow batch indices", [int(i) for i in np.nonzero(t > 1.3 * np.median(t))[0]], "pool MB",
          {k: v >> 20 for k, v in K.scratch_stats().items()}, flush=True)
    print(mode, "batch ms: median %.1f  min %.1f  max %.1f  p90 %.1f" % (np.median(t), t.min(), t.max(), np.percentile(t, 90)),
          " slow(>1.3x median):", [round(x) for x in t[t > 1.3 * np.median(t)]],
          " gc: n=%d total %.1f ms max %.1f ms" % (gc_t["n"], gc_t["tot"] * 1e3, gc_t["max"] * 1e3), flush=True)
