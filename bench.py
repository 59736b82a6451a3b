#!/usr/bin/env python
"""bench.py -- localised frames/s of the per-frame object-localisation hot path (BASELINE.json metric) on
synthetic 8room-style frames (configs[1]: query frames x 40-object memory, DINO-like 768-d embeddings).

  python bench.py [--gpus N] [--steps K] [--warmup W]            product arm (libibl.so kernels through the C ABI)
  python bench.py --impl reference ...                           the reference's CPU path (oracle restatement, host cores)
  torchrun ... bench.py --gpus N ...                             one rank per GPU; frames are sharded by rank (weak scaling)

A step = one pass of ObjectMemory.localise_arrays over a batch of --frames-per-step synthetic frames.
value  : frames/s, inputs resident in HBM when the timed region starts.
e2e    : frames/s through the same public call with HOST (pinned) rgb/depth/masks/embeddings, H2D copies and the D2H
         of the pose inside the timed region.
One JSON line on stdout (rank 0)."""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = dict(name="8room-style: 600x600 f=300 query frames vs 40-object memory (configs[1])", m=40, H=600, W=600,
                f=300.0, D=768, scene_seed=1002, mem_spacing=0.01, mem_voxel=0.01, fpfh_voxel_size=0.05,
                fpfh_global_dist_factor=1.5, fpfh_local_dist_factor=1.5, ransac_seed=1002)
LOCALISE_KW = dict(fpfh_global_dist_factor=WORKLOAD["fpfh_global_dist_factor"],
                   fpfh_local_dist_factor=WORKLOAD["fpfh_local_dist_factor"],
                   fpfh_voxel_size=WORKLOAD["fpfh_voxel_size"], perform_semantic_icp=False)


def peaks():
    """(HBM GB/s, dense bf16 TFLOP/s, source).  Burst figures: the roofline kernel is timed alone with its own events."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 2250.0, "fallback (B200_PROFILING.md: ~6.65 TB/s copy, 2.25 PFLOP/s nominal dense bf16)"


def ncu_traffic(tag):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of a kernel group, from the round's `ncu --set full`
    capture (profiles/ncu_traffic.json: {tag: {"bytes_per_launch": .., "source": .., "launch": ..}}); None if absent."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None, None
    e = json.load(open(p)).get(tag)
    return (e["bytes_per_launch"], f'{e["source"]} ({e["launch"]})') if e else (None, None)


# kernels whose IBL_PROF work unit is flops (csrc/match.cu); every other tag counts algorithmic bytes
TENSOR_KERNELS = {"feature_nn"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, gpu_index):
        self.rows, self.p = [], None
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "200",
                                       "-i", str(gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 6 and r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_inputs(n_frames, rank, device):
    """Scene + memory arrays + a pool of distinct frames for this rank (host arrays; deterministic per rank)."""
    from instance_based_loc_b200 import synth
    w = WORKLOAD
    scene = synth.make_scene(m=w["m"], seed=w["scene_seed"], D=w["D"])
    frames = synth.make_frames(scene, n_frames, H=w["H"], W=w["W"], f=w["f"], device=device, seed=100 + rank)
    return scene, frames


def cpu_reference_sample(scene_memory_host, frames, budget_s, seed):
    """Times the oracle (CPU restatement of the reference path: numpy + scipy cKDTree + C, all host threads the
    libraries use) on frame 0 until ~budget_s of work is done.  Frames/s is extrapolated from the fraction of the
    frame's assignments that were registered inside the budget (memory-side FPFH included, as the reference
    recomputes it per assignment)."""
    from oracle import o3d_core, ref_pipeline
    o3d_core.lib()
    fr = frames[0]
    w = WORKLOAD
    cfg = {"radius_nb_points": 8, "radius": 0.05}
    t0 = time.perf_counter()
    det = ref_pipeline.detect_clouds(fr["depth"], fr["rgb"], fr["masks"], w["f"], w["f"], cfg)
    t_front = time.perf_counter() - t0
    # orchestration up to the assignment list (cheap) -- then time registrations one by one
    embs = np.array(fr["embs"], np.float32)
    if len(det) > 7:
        order = sorted(range(len(det)), key=lambda i: len(det[i][0]), reverse=True)[:7]
        det, embs = [det[i] for i in order], embs[order]
    from oracle import ref_port
    t1 = time.perf_counter()
    S = ref_port.closest_similarities(embs, [m["embeddings"] for m in scene_memory_host])
    assns = ref_port.simvolume_assignments(S, 4)
    det2 = []
    for p, c in det:
        keep = o3d_core.remove_radius_outlier_mask(p, 8, 0.05)
        det2.append((p[keep], c[keep]))
    all_mem = np.concatenate([m["points"] for m in scene_memory_host])
    all_det = np.concatenate([p for p, _ in det2])
    t_mid = time.perf_counter() - t1
    done, t_reg = 0, 0.0
    for assn in assns:
        t2 = time.perf_counter()
        dp = np.concatenate([det2[d][0] for d, _ in assn]); dc = np.concatenate([det2[d][1] for d, _ in assn])
        mp = np.concatenate([scene_memory_host[m]["points"] for _, m in assn])
        mc = np.concatenate([scene_memory_host[m]["colors"] for _, m in assn])
        mm, dm = mp.mean(0), dp.mean(0)
        T, _, _ = ref_pipeline.register_point_clouds(dp - dm, dc, mp - mm, mc, w["fpfh_voxel_size"],
                                                     w["fpfh_global_dist_factor"], w["fpfh_local_dist_factor"], seed=seed)
        G = np.array(T); G[:3, 3] = T[:3, 3] + mm - T[:3, :3] @ dm
        ref_pipeline.evaluate_transform(all_det, all_mem, G)
        t_reg += time.perf_counter() - t2
        done += 1
        if t_front + t_mid + t_reg > budget_s:
            break
    per_frame = t_front + t_mid + t_reg * (len(assns) / max(done, 1))
    return {"value": 1.0 / per_frame, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"frame 0 of the workload: front end + {done} of {len(assns)} assignment registrations timed "
                      f"({t_front + t_mid + t_reg:.1f} s), extrapolated to the whole frame",
            "seconds_per_frame": per_frame}


def run_product(args):
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from instance_based_loc_b200 import kernels as K
    from instance_based_loc_b200 import selfcheck
    from instance_based_loc_b200.o3d import registration as reg
    w = WORKLOAD
    B = args.frames_per_step
    host_wait = args.host_wait
    K.set_host_wait(host_wait)
    scene, frames = make_inputs(B, rank, "cuda")
    om, host_mem = selfcheck.build_memory(scene, w["mem_spacing"], w["mem_voxel"], f=w["f"])
    reg.ransac_seed = w["ransac_seed"]
    resident = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in ("rgb", "depth", "masks", "embs")}
                for fr in frames]
    pinned = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).pin_memory() for k in ("rgb", "depth", "masks", "embs")}
              for fr in frames]
    h2d = sum(t.numel() * t.element_size() for fr in pinned for t in fr.values())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    gathered = [torch.zeros((B, 7), dtype=torch.float64, device=dev) for _ in range(world)]

    def step(host_inputs, workers=None):
        flush.zero_()                                   # L2 flush between iterations (126 MB L2 < 256 MiB)
        if host_inputs:                                 # H2D of this step's inputs from pinned memory
            batch = [{k: v.to(dev, non_blocking=True) for k, v in pinned[i].items()} for i in range(B)]
        else:
            batch = resident
        res = om.localise_batch(batch, workers=args.workers if workers is None else workers, **LOCALISE_KW)
        out = torch.from_numpy(np.stack([r[0] for r in res])).to(dev)
        if world > 1:                                   # the only traffic that crosses NVLink: 56 B per frame
            dist.all_gather(gathered, out)
        return out.cpu()                                # D2H of the step's result

    counted = {"launches": 0}

    def timed(host_inputs, steps, workers=None, profiler=False):
        step(host_inputs, workers)                      # untimed: every timed pass starts from the steady state of ITS mode
        if profiler:                                    # ncu --profile-from-start off: capture the timed region only
            torch.cuda.synchronize()
            torch.cuda.cudart().cudaProfilerStart()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = K.launch_count()
        e0.record()
        for _ in range(steps):
            step(host_inputs, workers)
        e1.record()
        torch.cuda.synchronize()
        counted["launches"] = K.launch_count() - n0      # libibl.so kernels launched inside this timed region
        if profiler:
            torch.cuda.cudart().cudaProfilerStop()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # nvidia-smi attaches to the driver when it starts, which can stall this process's CUDA calls for a few hundred ms
    # (tools/stall_probe.py vs tools/host_profile.py): start the sampler BEFORE the warm-up so that its start-up is over
    # when the timed region begins; it keeps sampling every 200 ms through all timed passes.
    clocks = ClockSampler(local)
    for _ in range(args.warmup):
        step(False)
    step(True)
    # a serving process freezes the heap after warm-up: a gen-2 collection is a 30-40 ms pause with 4 frame threads waiting
    gc.collect()
    gc.freeze()
    gc.disable()
    ms = timed(False, args.steps, profiler=args.profiler_range)   # `value`: no per-kernel events inside this region
    launches = counted["launches"]
    # per-kernel breakdown and the roofline kernel's launch time: the same steps again with CUDA events around every
    # kernel group (csrc/prof.cu), ONE frame in flight -- with several frames in flight an event pair on one stream also
    # spans whatever the other streams' kernels were doing, so a launch's duration would be over-stated.  The events cost
    # time and nothing overlaps, so this pass is not the one `value` is taken from.
    K.prof_enable(True)
    K.prof_read()
    ms_prof = timed(False, args.steps, workers=1)
    prof = K.prof_read()
    K.prof_enable(False)
    ms_e2e = timed(True, args.steps)
    clk = clocks.stop()
    total_frames = B * args.steps * world
    value = total_frames / (ms / 1e3)
    e2e = total_frames / (ms_e2e / 1e3)

    hbm_peak, tensor_peak, peak_src = peaks()
    gemm = prof.pop("feature_gemm", None)                 # nested inside feature_nn: not a group of its own
    top = max(prof.items(), key=lambda kv: kv[1]["ms"]) if prof else ("none", {"ms": 0.0, "launches": 0, "work": 0.0})
    tname, t = top
    per_launch_ms = t["ms"] / max(t["launches"], 1)
    work_per_s = (t["work"] / max(t["launches"], 1)) / (per_launch_ms / 1e3) if per_launch_ms > 0 else 0.0
    if tname in TENSOR_KERNELS:
        bound, unit, achieved, peak = "tensor", "TFLOP/s", work_per_s / 1e12, tensor_peak
        note = ("achieved = ALGORITHMIC flops 2*33*N_s*N_t of the FPFH distance matrix per launch / launch time (pack, the "
                "one-pass tcgen05 kind::f16 score GEMM of depth 112 (3-way fp16 split), the fp32-difference filter and the exact f64 "
                "re-evaluation are all inside the launch); peak is the measured dense bf16 figure")
    else:
        bound, unit, achieved, peak = "hbm", "GB/s", work_per_s / 1e9, hbm_peak
        note = "achieved = algorithmic bytes per launch (DESIGN.md section 4) / launch time"
    traffic, traffic_src = ncu_traffic(tname)
    roofline = {"kernel": tname, "bound": bound, "achieved": achieved, "peak": peak, "unit": unit,
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "avg_launch_ms": per_launch_ms, "launches": t["launches"],
                "share_of_step": t["ms"] / max(sum(v["ms"] for v in prof.values()), 1e-9), "note": note,
                "timed_in": "a second pass of the same K steps in this run, one frame in flight, with CUDA events around "
                            "every kernel group (`value` comes from the first, un-instrumented pass); share_of_step = this "
                            "kernel's event time / the sum over all kernel groups",
                "ms_per_step_instrumented": ms_prof / args.steps}
    if gemm and gemm["ms"] > 0:
        # the tcgen05 kernel alone: flops it issues (K = 112 over padded 128 x 128 tiles) against the tensor peak, and its time
        # against the floor set by reading every fp32 score out of TMEM once (64 B/clk/SM: 1024 clk per 128 x 128 tile)
        tiles = gemm["work"] / (2.0 * 112 * 128 * 128)
        floor_ms = tiles * 1024 / 148 / (float(clk["sm_mhz"] or 1965.0) * 1e3)
        roofline["gemm_kernel"] = {"kernel": "feat_tc_kernel<1> (tcgen05 kind::f16, one pass)", "launches": gemm["launches"],
                                   "avg_launch_ms": gemm["ms"] / gemm["launches"],
                                   "issued_tflops": gemm["work"] / gemm["ms"] / 1e9,
                                   "issued_frac_of_tensor_peak": gemm["work"] / gemm["ms"] / 1e9 / tensor_peak,
                                   "tmem_read_floor_ms_total": floor_ms, "frac_of_tmem_read_floor": floor_ms / gemm["ms"],
                                   "note": "MMA floor is 448 clk per tile, TMEM read-back of the accumulator 1024 clk per tile "
                                           "(DESIGN.md 4.1): the kernel is TMEM-read-bound, so ~0.44 of the tensor peak is its ceiling"}
    kernels = {k: {"ms": round(v["ms"], 3), "launches": v["launches"],
                   "gwork_per_s": round(v["work"] / max(v["ms"], 1e-9) / 1e6, 2)} for k, v in
               sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}
    line = {"metric": "localised frames/s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["name"], "frames_per_step_per_gpu": B, "frames_in_flight": args.workers,
                       "host_wait": host_wait, "host_cores": os.cpu_count(), "memory_objects": w["m"],
                       "memory_points": int(sum(len(m["points"]) for m in host_mem)), "embedding_dim": w["D"],
                       "ransac": "max_iteration 4e6, confidence 0.99, counter-based sample stream",
                       "l2_flush": "256 MiB memset before every step", "python_gc": "frozen and disabled after warm-up", "parallelism": f"frames sharded over {world} rank(s)"},
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": B * 56,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clk, "roofline": roofline, "kernels": kernels}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_reference_sample(host_mem, frames, args.cpu_budget, w["ransac_seed"])
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle restatement; Open3D itself is not installable here) on the
    same workload.  Each step is a bounded sample of one frame; rank 0 only."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    from oracle import ref_port
    from instance_based_loc_b200 import synth
    w = WORKLOAD
    scene, frames = make_inputs(1, 0, "cpu")
    host_mem = []
    for j in range(scene.m):
        p, c = synth.memory_cloud(scene, j, w["mem_spacing"])
        p2, c2, _, _ = ref_port.voxel_down_sample(p, c, w["mem_voxel"])
        host_mem.append({"points": p2, "colors": c2, "embeddings": scene.memory_embs[j]})
    budget = max(5.0, args.cpu_budget / max(args.steps, 1))
    for _ in range(min(args.warmup, 1)):
        cpu_reference_sample(host_mem, frames, 2.0, w["ransac_seed"])
    vals = [cpu_reference_sample(host_mem, frames, budget, w["ransac_seed"]) for _ in range(args.steps)]
    spf = float(np.mean([v["seconds_per_frame"] for v in vals]))
    cb = dict(vals[-1], value=1.0 / spf)
    line = {"impl": "reference", "metric": "localised frames/s", "value": 1.0 / spf, "unit": "frames/s",
            "n_gpus": int(os.environ.get("WORLD_SIZE", 1)), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": spf * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": w["name"], "memory_objects": w["m"], "embedding_dim": w["D"]},
            "cpu_baseline": cb, "e2e": {"value": 1.0 / spf, "unit": "frames/s", "h2d_bytes_per_step": 0,
                                        "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--frames-per-step", type=int, default=16)
    ap.add_argument("--workers", type=int, default=4, help="frames in flight per GPU (host threads / CUDA streams)")
    ap.add_argument("--host-wait", default="auto", choices=["auto", "spin", "yield", "block", "hybrid"],
                    help="how host threads wait in stream syncs (ibl_set_host_wait).  Measured at 8 GPUs / 32 cores: auto "
                         "(spin) 227-246 frames/s, yield 205, block 70 -- the path syncs hundreds of times per frame")
    ap.add_argument("--cpu-budget", type=float, default=25.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profiler-range", action="store_true",
                    help="bracket the device-resident timed region with cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
