"""Diagnostic (GPU box): where does the single-frame wall time go on the host?  cProfile of localise_arrays after
warm-up, process CPU time vs wall, and a workers sweep with nvidia-smi utilisation."""
import cProfile, io, os, pstats, subprocess, sys, threading, time
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
for i in range(B):
    om.localise_arrays(*(resident[i][k] for k in keys), **bench.LOCALISE_KW)
torch.cuda.synchronize()
for rep in range(2):
    c0, t0 = time.process_time(), time.perf_counter()
    for i in range(B):
        om.localise_arrays(*(resident[i][k] for k in keys), **bench.LOCALISE_KW)
    torch.cuda.synchronize()
    print(f"sequential: wall {(time.perf_counter()-t0)*1e3/B:.1f} ms/frame, process cpu {(time.process_time()-c0)*1e3/B:.1f} ms/frame", flush=True)
pr = cProfile.Profile()
pr.enable()
for i in range(B):
    om.localise_arrays(*(resident[i][k] for k in keys), **bench.LOCALISE_KW)
torch.cuda.synchronize()
pr.disable()
for sort in ("cumulative", "tottime"):
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats(sort).print_stats(45)
    print(s.getvalue()[:9000], flush=True)


class Util:
    def __init__(self):
        self.rows = []
        self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=utilization.gpu", "--format=csv,noheader,nounits", "-lms", "100", "-i", "0"],
                                  stdout=subprocess.PIPE, text=True)
        threading.Thread(target=lambda: [self.rows.append(l.strip()) for l in self.p.stdout], daemon=True).start()
    def stop(self):
        self.p.terminate()
        v = [float(x) for x in self.rows if x.replace(".", "").isdigit()]
        return (np.mean(v) if v else -1, len(v))


for workers in (1, 2, 3, 4, 6, 8):
    om.localise_batch(resident, workers=workers, **bench.LOCALISE_KW)
    torch.cuda.synchronize(); u = Util(); c0, t0 = time.process_time(), time.perf_counter()
    n = 0
    for _ in range(3):
        n += len(om.localise_batch(resident, workers=workers, **bench.LOCALISE_KW))
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 1e3; cpu = (time.process_time() - c0) * 1e3
    print("workers", workers, f"{n / ms * 1e3:.2f} frames/s  wall {ms/n:.1f} ms/frame  cpu {cpu/n:.1f} ms/frame  gpu util {u.stop()}", flush=True)
