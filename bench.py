#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric ("localised frames/s ...; RANSAC hyps/s") for the per-frame object-localisation
hot path, on the synthetic configurations of SURVEY.md section 8(d) (instance_based_loc_b200/workloads.py).

  python bench.py [--gpus N] [--steps K] [--warmup W]            product arm, configs[1] (the driver's call)
  python bench.py --config {0,1,2,3,4} ...                       the other BASELINE configurations
  python bench.py --impl reference ...                           the reference's CPU path (oracle restatement, host cores)
  torchrun ... bench.py --gpus N ...                             one rank per GPU; frames sharded by rank (weak scaling);
                                                                 --config 4 shards every registration by hypothesis block

Frame configurations (0-3).  A step = ObjectMemory.localise_batch over --frames-per-step frames taken, in order, from a
pool of DISTINCT frames per rank (100 for configs[1]); successive steps continue through the pool and wrap around.
  value  : frames/s over the K timed steps, inputs resident in HBM.
  e2e    : frames/s through the same public call with HOST (pinned) rgb / depth / masks / embeddings: H2D copies of the
           step's frames and the D2H of its poses inside the timed region.  The same frames as the `value` pass.
  cold   : one pass over the whole pool right after the per-assignment memory-side caches were emptied (every assignment
           tuple pays for its target FPFH / search structures once), with the cache hit rate of that pass.
  ransac : hypotheses drawn / validated per second over the timed region (RegResultT.drawn / .validated), means per
           registration; gt: share of frames within the reference's success criterion (0.6 m, 0.3 rad) of the true pose.
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
# torch's own caching allocator grows by cudaMalloc whenever a frame needs a block size it has not cached for that stream;
# with several frame streams that happened every few steps and stalled whole steps (measured: 150-220 ms steps with
# occasional 400-600 ms ones).  The stream-ordered allocator shares the driver pool libibl.so already uses (release
# threshold = max), so after warm-up nothing on the path calls cudaMalloc.  Must be set before torch touches CUDA.
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "backend:cudaMallocAsync")

from instance_based_loc_b200 import workloads as WL  # noqa: E402


def peaks():
    """(HBM GB/s, dense bf16 TFLOP/s, source).  Burst figures: the roofline kernel is timed alone with its own events."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d["bf16_tflops"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 2250.0, "fallback (B200_PROFILING.md: ~6.65 TB/s copy, 2.25 PFLOP/s nominal dense bf16)"


def _profile_json(name):
    p = os.path.join(ROOT, "profiles", name)
    return json.load(open(p)) if os.path.exists(p) else {}


def ncu_traffic(tag):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of a kernel group, from the round's `ncu --set full`
    capture (profiles/ncu_traffic.json: {tag: {"bytes_per_launch": .., "source": .., "launch": ..}}); None if absent."""
    e = _profile_json("ncu_traffic.json").get(tag)
    return (e["bytes_per_launch"], f'{e["source"]} ({e["launch"]})') if e else (None, None)


# kernels whose IBL_PROF work unit is flops (csrc/match_tc.cu); every other tag counts algorithmic bytes
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


def config_keys(cfg, extra=None):
    """The `config` object both arms print (same keys, so the driver's same_config check can hold)."""
    c = {"workload": cfg["name"], "memory_objects": cfg["m"], "frame": f'{cfg["W"]}x{cfg["H"]} f={cfg["f"]}',
         "embedding_dim": cfg["D"], "pool_frames_per_rank": cfg["pool_frames"],
         "fpfh_voxel_size": cfg["fpfh_voxel_size"], "fpfh_global_dist_factor": cfg["fpfh_global_dist_factor"],
         "fpfh_local_dist_factor": cfg["fpfh_local_dist_factor"],
         "ransac": ("max_iteration 4e6, confidence 0.99 (adaptive stop)" if cfg["ransac_max_iteration"] is None else
                    f'max_iteration {cfg["ransac_max_iteration"]} fixed, no adaptive stop') + ", counter-based sample stream"}
    c.update(extra or {})
    return c


# ======================================================================================================================
#  CPU arm: the oracle's restatement of the reference path, timed on the host cores
# ======================================================================================================================
class CpuFrame:
    """Frame 0 of the workload cut into SEGMENTS that can be timed one by one: segment 0 is the front end
    (back-projection, both outlier filters, closest similarities, SimVolume), segment k >= 1 is the k-th assignment's
    registration (source AND memory-side FPFH -- the reference recomputes both per assignment -- matching, RANSAC,
    coloured ICP) plus its evaluate_transform.  frames/s = 1 / sum of the segments' mean times."""

    def __init__(self, cfg, host_mem, frame):
        from oracle import o3d_core
        o3d_core.lib()
        self.cfg, self.mem, self.fr = cfg, host_mem, frame
        self.times = {}
        self.state = None

    def front(self):
        from oracle import o3d_core, ref_pipeline, ref_port
        cfg, fr = self.cfg, self.fr
        t0 = time.perf_counter()
        det = ref_pipeline.detect_clouds(fr["depth"], fr["rgb"], fr["masks"], cfg["f"], cfg["f"],
                                         {"radius_nb_points": 8, "radius": 0.05})
        embs = np.array(fr["embs"], np.float32)
        if len(det) > 7:
            order = sorted(range(len(det)), key=lambda i: len(det[i][0]), reverse=True)[:7]
            det, embs = [det[i] for i in order], embs[order]
        S = ref_port.closest_similarities(embs, [m["embeddings"] for m in self.mem])
        assns = ref_port.simvolume_assignments(S, 4)
        det2 = []
        for p, c in det:
            keep = o3d_core.remove_radius_outlier_mask(p, 8, 0.05)
            det2.append((p[keep], c[keep]))
        all_mem = np.concatenate([m["points"] for m in self.mem])
        all_det = np.concatenate([p for p, _ in det2])
        dt = time.perf_counter() - t0
        self.state = (assns, det2, all_mem, all_det)
        self.times.setdefault(0, []).append(dt)
        return dt

    def n_segments(self):
        if self.state is None:
            self.front()
        return 1 + len(self.state[0])

    def registration(self, k):
        from oracle import ref_pipeline
        cfg = self.cfg
        assns, det2, all_mem, all_det = self.state
        assn = assns[k - 1]
        t0 = time.perf_counter()
        dp = np.concatenate([det2[d][0] for d, _ in assn]); dc = np.concatenate([det2[d][1] for d, _ in assn])
        mp = np.concatenate([self.mem[m]["points"] for _, m in assn])
        mc = np.concatenate([self.mem[m]["colors"] for _, m in assn])
        mm, dm = mp.mean(0), dp.mean(0)
        kw = {}
        if cfg["ransac_max_iteration"] is not None:
            kw = dict(max_iteration=cfg["ransac_max_iteration"], early_exit=cfg["ransac_early_exit"])
        T, _, _ = ref_pipeline.register_point_clouds(dp - dm, dc, mp - mm, mc, cfg["fpfh_voxel_size"],
                                                     cfg["fpfh_global_dist_factor"], cfg["fpfh_local_dist_factor"],
                                                     seed=cfg["ransac_seed"], **kw)
        G = np.array(T); G[:3, 3] = T[:3, 3] + mm - T[:3, :3] @ dm
        ref_pipeline.evaluate_transform(all_det, all_mem, G)
        dt = time.perf_counter() - t0
        self.times.setdefault(k, []).append(dt)
        return dt

    def run_segment(self, k):
        return self.front() if k == 0 else self.registration(k)

    def summary(self):
        """seconds per frame from the segment means; registrations never timed are estimated at the mean seconds per
        source point of the timed ones (and reported)."""
        assns, det2 = self.state[0], self.state[1]
        n_src = [sum(len(det2[d][0]) for d, _ in a) for a in assns]
        mean = {k: float(np.mean(v)) for k, v in self.times.items()}
        timed = [k for k in range(1, len(assns) + 1) if k in mean]
        rate = (sum(mean[k] for k in timed) / max(sum(n_src[k - 1] for k in timed), 1)) if timed else 0.0
        total, missing = mean.get(0, 0.0), []
        for k in range(1, len(assns) + 1):
            if k in mean:
                total += mean[k]
            else:
                total += rate * n_src[k - 1]
                missing.append(k)
        return total, missing, {str(k): round(v, 3) for k, v in sorted(mean.items())}


def cpu_host_memory(cfg, scene):
    from instance_based_loc_b200 import synth
    from oracle import ref_port
    host_mem = []
    for j in range(scene.m):
        p, c = synth.memory_cloud(scene, j, cfg["mem_spacing"])
        p2, c2, _, _ = ref_port.voxel_down_sample(p, c, cfg["mem_voxel"])
        host_mem.append({"points": p2, "colors": c2, "embeddings": scene.memory_embs[j]})
    return host_mem


def segment_order(n):
    """front end first, then registrations alternating from both ends of the assignment list (assignments come out
    ordered by length, so a bounded sample that stops early still holds short AND long ones)."""
    regs, lo, hi = [], 1, n - 1
    while lo <= hi:
        regs.append(hi)
        if lo != hi:
            regs.append(lo)
        lo, hi = lo + 1, hi - 1
    return [0] + regs


def cpu_bounded_sample(cfg, host_mem, frame, budget_s):
    """cpu_baseline of the product arm: segments of frame 0 timed in segment_order until ~budget_s is spent."""
    cf = CpuFrame(cfg, host_mem, frame)
    spent = cf.front()
    order = segment_order(cf.n_segments())[1:]
    done = 0
    for k in order:
        if spent > budget_s and done >= 1:
            break
        spent += cf.registration(k)
        done += 1
    spf, missing, seg = cf.summary()
    return {"value": 1.0 / spf, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"frame 0 of the workload: front end + {done} of {cf.n_segments() - 1} assignment registrations timed "
                      f"({spent:.1f} s of CPU work)" + (f"; registrations {missing} estimated from the timed ones' "
                                                        f"seconds per source point" if missing else ""),
            "seconds_per_frame": spf, "segment_seconds": seg}


def run_reference(args, cfg):
    """--impl reference: the reference's CPU path on the same workload (the oracle restatement: Open3D itself is not
    installable here).  One STEP = one segment of frame 0 (CpuFrame), taken round-robin, timed for real -- ms_per_step is
    the measured mean wall time of the K timed steps; value = 1 / (sum over segments of their mean time, warm-up steps
    included in the means).  Rank 0 only."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    if args.config == 4:
        return run_micro_reference(args, cfg)
    scene = WL.make_scene(cfg)
    frames = WL.make_frames(cfg, scene, 1, rank=0, device="cpu")
    host_mem = cpu_host_memory(cfg, scene)
    cf = CpuFrame(cfg, host_mem, frames[0])
    n_seg = None
    order, pos, walls = [0], 0, []
    for i in range(args.warmup + args.steps):
        k = order[pos % len(order)]
        pos += 1
        dt = cf.run_segment(k)
        if n_seg is None:
            n_seg = cf.n_segments()
            order = segment_order(n_seg)
        if i >= args.warmup:
            walls.append(dt)
    spf, missing, seg = cf.summary()
    cb = {"value": 1.0 / spf, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
          "sample": f"frame 0 of the workload cut into {n_seg} segments (front end + one per assignment registration); "
                    f"{args.warmup + args.steps} steps of one segment each, round-robin"
                    + (f"; segments {missing} never timed, estimated per source point" if missing else ""),
          "seconds_per_frame": spf, "segment_seconds": seg}
    line = {"impl": "reference", "metric": "localised frames/s", "value": 1.0 / spf, "unit": "frames/s",
            "n_gpus": int(os.environ.get("WORLD_SIZE", 1)), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": float(np.mean(walls)) * 1e3 if walls else 0.0, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_keys(cfg, {"step": "one segment of one frame (see cpu_baseline.sample)"}),
            "cpu_baseline": cb, "e2e": {"value": 1.0 / spf, "unit": "frames/s", "h2d_bytes_per_step": 0,
                                        "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ======================================================================================================================
#  product arm, frame configurations (0-3)
# ======================================================================================================================
def run_frames(args, cfg):
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
    B = args.frames_per_step or cfg["frames_per_step"]
    workers = args.workers or cfg["workers"]
    K.set_host_wait(args.host_wait)
    WL.apply_ransac_mode(cfg)
    kw = WL.localise_kwargs(cfg)
    scene = WL.make_scene(cfg)
    n_pool = max(args.pool or cfg["pool_frames"], 1)
    t_setup = time.perf_counter()
    frames = WL.make_frames(cfg, scene, n_pool, rank=rank, device="cuda")
    om, host_mem = selfcheck.build_memory(scene, cfg["mem_spacing"], cfg["mem_voxel"], f=cfg["f"])
    keys = ("rgb", "depth", "masks", "embs")
    resident = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).to(dev) for k in keys} for fr in frames]
    pinned = [{k: torch.from_numpy(np.ascontiguousarray(fr[k])).pin_memory() for k in keys} for fr in frames]
    frame_bytes = [sum(t.numel() * t.element_size() for t in fr.values()) for fr in pinned]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    # The poses of a step are all-gathered ASYNCHRONOUSLY (NCCL's own stream) and only waited for at the end of the timed
    # region: frames are independent, so nothing on the path needs the other ranks' results, and a per-step wait made every
    # step last as long as the slowest rank's 16 frames (measured at N = 8: steps of ~175 ms against ~150 ms at N = 1, the
    # mean plus ~1.4 sigma of the per-step spread -- that, not host contention, was the scaling loss; `taskset -c 0-3` at
    # N = 1 costs nothing).
    pending = []
    t_setup = time.perf_counter() - t_setup
    cursor = {"i": 0}
    log = {"on": False, "rows": []}
    trace = bool(os.environ.get("IBL_BENCH_TRACE"))

    def step(host_inputs, nworkers=None):
        flush.zero_()                                   # L2 flush between iterations (126 MB L2 < 256 MiB)
        idx = [(cursor["i"] + k) % n_pool for k in range(B)]
        cursor["i"] = (cursor["i"] + B) % n_pool
        if host_inputs:                                 # H2D of this step's frames from pinned memory
            batch = [{k: v.to(dev, non_blocking=True) for k, v in pinned[i].items()} for i in idx]
        else:
            batch = [resident[i] for i in idx]
        t_call = time.perf_counter()
        res = om.localise_batch(batch, workers=workers if nworkers is None else nworkers, **kw)
        if trace:
            ms_ = torch.cuda.memory_stats()
            sc_ = K.scratch_stats()
            print(f"[trace] torch segments={ms_.get('num_device_alloc', 0)} frees={ms_.get('num_device_free', 0)} "
                  f"reserved={ms_.get('reserved_bytes.all.current', 0) >> 20} MiB; pool reserved={sc_['reserved'] >> 20} MiB "
                  f"used_high={sc_['used_high'] >> 20} MiB", file=sys.stderr, flush=True)
            print(f"[trace] step host_inputs={bool(host_inputs)} idx0={idx[0]} call_ms={(time.perf_counter() - t_call) * 1e3:.1f} "
                  f"timing={getattr(om, 'last_batch_timing', None)}", file=sys.stderr, flush=True)
            for i, r in zip(idx, res):
                s1 = r[2].get("stage1_ms", (0, 0))
                if s1[0] + s1[1] > 60:
                    print(f"[trace]   slow frame {i}: plan {s1[0]:.0f} ms, features+matching {s1[1]:.0f} ms, regs "
                          f"{[(g['n_source'], g['n_target'], g['ransac'].get('n_corres')) for g in r[2]['registrations']]}",
                          file=sys.stderr, flush=True)
        out = torch.from_numpy(np.stack([r[0] for r in res])).to(dev)
        if world > 1:                                   # the only traffic that crosses NVLink: 56 B per frame
            bufs = [torch.empty((B, 7), dtype=torch.float64, device=dev) for _ in range(world)]
            pending.append((dist.all_gather(bufs, out, async_op=True), bufs, out))
        host = out.cpu()                                # D2H of the step's result
        if log["on"]:
            log["rows"].extend(zip(idx, res))
        return host, idx

    counted = {"launches": 0}

    def drain():
        for w_, _, _ in pending:
            w_.wait()
        del pending[:]

    def timed(host_inputs, steps, nworkers=None, profiler=False, start=0, keep=False, before=None):
        cursor["i"] = (start - B) % n_pool
        step(host_inputs, nworkers)                     # untimed: every timed pass starts from the steady state of ITS mode
        if before is not None:
            before()
        if profiler:                                    # ncu --profile-from-start off: capture the timed region only
            torch.cuda.synchronize()
            torch.cuda.cudart().cudaProfilerStart()
        drain()                                         # gathers of the untimed steps before this point
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = K.launch_count()
        log["on"], log["rows"] = keep, []
        st0 = dict(om.cache_stats)
        e0.record()
        walls = []
        for _ in range(steps):
            t_s = time.perf_counter()
            step(host_inputs, nworkers)                 # ends with the D2H read of the step's poses: the host time is the step's
            walls.append(round((time.perf_counter() - t_s) * 1e3, 1))
        drain()                                         # every rank's poses of every step have arrived: inside the timed region
        e1.record()
        counted["step_wall_ms"] = walls
        torch.cuda.synchronize()
        counted["launches"] = K.launch_count() - n0      # libibl.so kernels launched inside this timed region
        log["on"] = False
        if profiler:
            torch.cuda.cudart().cudaProfilerStop()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        look = om.cache_stats["lookups"] - st0["lookups"]
        miss = om.cache_stats["misses"] - st0["misses"]
        return float(ms.item()), (1.0 - miss / look) if look else None

    # nvidia-smi attaches to the driver when it starts, which can stall this process's CUDA calls for a few hundred ms:
    # start the sampler BEFORE the warm-up so that its start-up is over when the timed region begins.
    clocks = ClockSampler(local)
    step(False)                                          # CUDA / library first-call costs stay out of the cold pass
    # ---- cold pass: every frame of the pool once, per-assignment memory-side caches emptied first
    cold_steps = max(1, n_pool // B)
    ms_cold, hit_cold = timed(False, cold_steps, start=0, before=om.clear_assignment_cache)
    for _ in range(args.warmup):
        step(False)
    step(True)
    # a serving process freezes the heap after warm-up: a gen-2 collection is a 30-40 ms pause with frame threads waiting
    gc.collect()
    gc.freeze()
    gc.disable()
    start = cursor["i"]
    ms, hit_warm = timed(False, args.steps, profiler=args.profiler_range, start=start, keep=True)
    launches = counted["launches"]
    walls_value = counted["step_wall_ms"]
    rows = log["rows"]
    # per-kernel breakdown and the roofline kernel's launch time: the same steps again with CUDA events around every
    # kernel group (csrc/prof.cu), ONE frame in flight -- with several frames in flight an event pair on one stream also
    # spans whatever the other streams' kernels were doing.  Not the pass `value` is taken from.
    K.prof_enable(True)
    K.prof_read()
    ms_prof, _ = timed(False, args.steps, nworkers=1, start=start)
    prof = K.prof_read()
    K.prof_enable(False)
    ms_e2e, _ = timed(True, args.steps, start=start)
    walls_e2e = counted["step_wall_ms"]
    clk = clocks.stop()
    total_frames = B * args.steps * world
    value = total_frames / (ms / 1e3)
    e2e = total_frames / (ms_e2e / 1e3)
    idx_timed = [(start + k) % n_pool for k in range(B * args.steps)]
    h2d = float(np.mean([sum(frame_bytes[i] for i in idx_timed[s * B:(s + 1) * B]) for s in range(args.steps)]))

    # ---- what the timed region computed: RANSAC volume, success against ground truth
    drawn = validated = corres = nreg = 0
    succ = succ_best = 0
    for i, r in rows:
        info = r[2]
        for g in info.get("registrations", []):
            ra = g.get("ransac") or {}
            drawn += ra.get("drawn", 0); validated += ra.get("validated", 0); corres += ra.get("n_corres", 0)
            nreg += 1
        succ += WL.pose_success(r[0], frames[i]["T_wc"])[2]
        if "pose_best_means" in info:
            succ_best += WL.pose_success(info["pose_best_means"], frames[i]["T_wc"])[2]
    sec = ms / 1e3
    stats = torch.tensor([drawn, validated, corres, nreg, succ, succ_best, len(rows)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(stats)
    drawn, validated, corres, nreg, succ, succ_best, nrows = [float(x) for x in stats.tolist()]
    ransac = {"hyps_drawn_per_s": drawn / sec, "hyps_validated_per_s": validated / sec,
              "drawn_per_registration": drawn / max(nreg, 1), "validated_per_registration": validated / max(nreg, 1),
              "correspondences_per_registration": corres / max(nreg, 1), "registrations_per_frame": nreg / max(nrows, 1)}
    gt = {"success_rate": succ / max(nrows, 1), "success_rate_best_assignment_means": succ_best / max(nrows, 1),
          "criterion": f"trans < {WL.SUCCESS_TRANS_M} m and rot < {WL.SUCCESS_ROT_RAD} rad (8room_localisation_trial.py:213)",
          "note": "success_rate is the pose localise() returns, which composes the translation with the LAST assignment's "
                  "means (object_memory.py:1127, reproduced under bug_compatible); the second figure composes it with the "
                  "best assignment's own means", "frames": int(nrows)}

    hbm_peak, tensor_peak, peak_src = peaks()
    gemm = prof.pop("feature_gemm", None)                 # nested inside feature_nn: not a group of its own
    top = max(prof.items(), key=lambda kv: kv[1]["ms"]) if prof else ("none", {"ms": 0.0, "launches": 0, "work": 0.0})
    tname, t = top
    per_launch_ms = t["ms"] / max(t["launches"], 1)
    work_per_s = (t["work"] / max(t["launches"], 1)) / (per_launch_ms / 1e3) if per_launch_ms > 0 else 0.0
    if tname in TENSOR_KERNELS:
        bound, unit, achieved, peak = "tensor", "TFLOP/s", work_per_s / 1e12, tensor_peak
        note = ("achieved = ALGORITHMIC flops 2*33*N_s*N_t of the FORWARD search per launch / launch time; the reverse search "
                "runs only for the memory rows the mutual filter reads, its row count lives on the device and its flops are "
                "not counted although its time is (a lower bound) (operand "
                "packing, the one-pass tcgen05 kind::f16 score GEMM of depth 112, the fp32-difference filter and the exact "
                "f64 re-evaluation are all inside the launch); peak is the measured dense bf16 figure")
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
        # the kernel's own floor: its 128x128x16 MMAs back to back at the measured dense rate (tools/probes/mma_probe.cu: 64 clk
        # per MMA on zero operands; the measured bf16 peak corresponds to ~90 clk per MMA on real data, which is what the
        # kernel's MMA thread sees inside its tile loop, profiles/r02_tc_wait_after.json)
        floor_ms = gemm["work"] / (tensor_peak * 1e12) * 1e3
        roofline["gemm_kernel"] = {"kernel": "feat_tc_kernel<1> (tcgen05 kind::f16, one pass)", "launches": gemm["launches"],
                                   "avg_launch_ms": gemm["ms"] / gemm["launches"],
                                   "issued_tflops": gemm["work"] / gemm["ms"] / 1e9,
                                   "issued_frac_of_tensor_peak": gemm["work"] / gemm["ms"] / 1e9 / tensor_peak,
                                   "mma_floor_ms_total_at_measured_peak": floor_ms,
                                   "note": "issued flops = 2 * 112 * padded rows * padded columns (the depth-112 split GEMM); the "
                                           "rest of the launch time is per-CTA prologue, wave quantisation and the epilogue's "
                                           "candidate path (DESIGN.md 4.1)"}
    bounds = _profile_json("r02_kernel_bounds.json")      # per kernel group: what it is bound by, from ncu (tools/ncu_summary.py)
    kernels = {}
    for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
        kernels[k] = {"ms": round(v["ms"], 3), "launches": v["launches"],
                      "gwork_per_s": round(v["work"] / max(v["ms"], 1e-9) / 1e6, 2)}
        if k in bounds:
            kernels[k].update(bounds[k])
    line = {"metric": "localised frames/s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_keys(cfg, {"frames_per_step_per_gpu": B, "frames_in_flight": workers, "host_wait": args.host_wait,
                                        "host_cores": os.cpu_count(),
                                        "memory_points": int(sum(len(m["points"]) for m in host_mem)),
                                        "frames": f"pool of {n_pool} distinct frames per rank, taken in order and wrapped",
                                        "l2_flush": "256 MiB memset before every step",
                                        "python_gc": "frozen and disabled after warm-up",
                                        "parallelism": f"frames sharded over {world} rank(s)"}),
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": B * 56,
                    "ms_per_step": ms_e2e / args.steps, "step_wall_ms": walls_e2e},
            "step_wall_ms": walls_value,                 # host wall time of every timed step (rank 0): outliers are visible
            "gpu_launches": launches, "launches_per_frame": launches / max(B * args.steps, 1), "clocks": clk,
            "ransac": ransac, "gt": gt, "memory_cache_hit_rate": hit_warm, "memory_cache": om.cache_info(),
            "cold": {"value": B * cold_steps * world / (ms_cold / 1e3), "unit": "frames/s", "frames_per_rank": B * cold_steps,
                     "memory_cache_hit_rate": hit_cold,
                     "note": "first pass over the pool with the per-assignment memory-side caches (target normals / FPFH / "
                             "NN fields / colour gradients) emptied; per-object and all-memory structures stay built"},
            "setup_s": t_setup, "roofline": roofline, "kernels": kernels}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_bounded_sample(cfg, host_mem, frames[0], args.cpu_budget)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ======================================================================================================================
#  configs[4]: RANSAC / ICP microbenchmark sweep
# ======================================================================================================================
def run_micro(args, cfg):
    """For every cloud size N: FPFH of both clouds + mutual matching once, then for every hypothesis count H two RANSAC
    runs without adaptive stop -- checkers ON (H drawn, H_v validated) and checkers OFF (H_v = H: every hypothesis is
    scored over the whole source) -- and one coloured ICP from the winner.  With W ranks every run is sharded by
    hypothesis block (dist.hypothesis_block) and the winners merged by dist.best_of_ranks; times are max over ranks."""
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from instance_based_loc_b200 import dist as D
    from instance_based_loc_b200 import kernels as K
    from instance_based_loc_b200.o3d import registration as reg
    from instance_based_loc_b200.o3d.pointcloud import PointCloud
    from instance_based_loc_b200.utils.fpfh_register import downsample_and_compute_fpfh
    K.set_host_wait(args.host_wait)
    clocks = ClockSampler(local)
    vs, max_d = cfg["fpfh_voxel_size"], cfg["max_d"]
    points = cfg["points"] if not args.micro_points else tuple(int(x) for x in args.micro_points.split(","))
    hyps = cfg["hypotheses"] if not args.micro_hyps else tuple(int(x) for x in args.micro_hyps.split(","))

    def dev_time(fn, reps):
        """max-over-ranks CUDA-event time of reps calls, after one untimed call."""
        out = fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), out

    n0 = K.launch_count()
    cells, total_drawn, total_ms = [], 0.0, 0.0
    for n in points:
        sp, sc, tp, tc, T_gt = WL.microbench_pair(n, cfg["outliers"], cfg["seed"])
        src, tgt = PointCloud(sp), PointCloud(tp)
        src.colors, tgt.colors = sc, tc
        ms_feat, (sf, tf) = dev_time(lambda: (downsample_and_compute_fpfh(src, vs), downsample_and_compute_fpfh(tgt, vs)), 1)
        (sd, sfe), (td, tfe) = sf, tf
        ms_match, cor = dev_time(lambda: reg.match_features(sfe, tfe, True, 3), max(1, args.steps))
        grid = reg._nn_grid(td, max_d)
        reg.ensure_color_gradient(td, max_d)
        for H in hyps:
            h0, h1 = D.hypothesis_block(H, rank, world)
            row = {"points": n, "hypotheses": H, "correspondences": int(cor.shape[0])}
            for name, chk in (("checkers_on", True), ("checkers_off", False)):
                if not chk and H * n > args.micro_max_evals:
                    row[name] = {"skipped": f"H*N = {H * n:.2e} point-hypothesis evaluations > --micro-max-evals"}
                    continue

                def run():
                    r = K.ransac_run(grid, sd._p, td._p, cor, h1, 0.99, max_d, edge_thr=0.9, seed=cfg["seed"],
                                     use_checkers=chk, early_exit=False, block=reg.ransac_block, h0=h0)
                    best = D.best_of_ranks(r["fitness"], r["inlier_rmse"], r["best_hyp"], r["T"], device=dev)
                    return r, best
                ms, (r, best) = dev_time(run, max(1, args.steps))
                val = torch.tensor([float(r["validated"])], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(val)
                v = float(val.item())
                R = best[3][:3, :3]
                rot_err = float(np.degrees(np.arccos(np.clip((np.trace(R.T @ T_gt[:3, :3]) - 1) / 2, -1, 1))))
                row[name] = {"ms": ms, "hyps_drawn_per_s": H / (ms / 1e3), "validated": v,
                             "hyps_validated_per_s": v / (ms / 1e3), "point_hyp_evals_per_s": v * n / (ms / 1e3),
                             "fitness": best[0], "rot_err_deg_vs_gt": rot_err}
                if chk:
                    total_drawn += H
                    total_ms += ms
                    T0 = best[3]
            ms_icp, ri = dev_time(lambda: reg.registration_colored_icp(sd, td, max_d, T0), max(1, args.steps))
            row["colored_icp"] = {"ms": ms_icp, "iterations": ri.info["iterations"], "fitness": ri.fitness,
                                  "ms_per_iteration": ms_icp / max(ri.info["iterations"] + 1, 1)}
            row["fpfh_both_clouds_ms"], row["mutual_match_ms"] = ms_feat, ms_match
            cells.append(row)
    clk = clocks.stop()
    value = total_drawn / (total_ms / 1e3) if total_ms else 0.0
    line = {"metric": "RANSAC hyps/s", "value": value, "unit": "hypotheses drawn/s (checkers on, sweep aggregate)",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": cfg["name"], "points": list(points), "hypotheses": list(hyps), "outliers": cfg["outliers"],
                       "max_d": max_d, "parallelism": f"hypothesis blocks over {world} rank(s), winners merged by one "
                                                      f"19-double all_gather per run"},
            "gpu_launches": K.launch_count() - n0, "clocks": clk, "sweep": cells}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_micro_reference(args, cfg):
    """CPU arm of configs[4]: the oracle's RANSAC on the smallest cell (bounded)."""
    from oracle import o3d_core as oc, ref_pipeline, sampler
    n, H = cfg["points"][0], cfg["hypotheses"][0]
    sp, sc, tp, tc, _ = WL.microbench_pair(n, cfg["outliers"], cfg["seed"])
    sn, sf = ref_pipeline.fpfh_features(sp, cfg["fpfh_voxel_size"])
    tn, tf = ref_pipeline.fpfh_features(tp, cfg["fpfh_voxel_size"])
    cor = oc.match_features(sf, tf, True, 3)
    C = len(cor)
    walls = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        oc.ransac(sp, tp, cor, lambda a, b: sampler.samples(cfg["seed"], a, b - a, C), cfg["max_d"], 0.9, 0.99,
                  early_exit=False, max_iteration=H)
        if i >= args.warmup:
            walls.append(time.perf_counter() - t0)
    v = H / float(np.mean(walls))
    line = {"impl": "reference", "metric": "RANSAC hyps/s", "value": v, "unit": "hypotheses drawn/s (checkers on)", "n_gpus": 1,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": float(np.mean(walls)) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": cfg["name"], "points": [n], "hypotheses": [H]},
            "cpu_baseline": {"value": v, "unit": "hypotheses drawn/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": f"the sweep's smallest cell only: {n} points, {H} hypotheses, checkers on"},
            "e2e": {"value": v, "unit": "hypotheses drawn/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=sorted(WL.CONFIGS),
                    help="BASELINE.json configs index (default 1: the configuration the metric is quoted on)")
    ap.add_argument("--frames-per-step", type=int, default=0, help="0 = the configuration's own")
    ap.add_argument("--pool", type=int, default=0, help="distinct frames per rank (0 = the configuration's own)")
    ap.add_argument("--workers", type=int, default=0, help="frames in flight per GPU (0 = the configuration's own)")
    ap.add_argument("--host-wait", default="auto", choices=["auto", "spin", "yield", "block", "hybrid"],
                    help="how host threads wait in stream syncs (ibl_set_host_wait)")
    ap.add_argument("--cpu-budget", type=float, default=25.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profiler-range", action="store_true",
                    help="bracket the device-resident timed region with cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    ap.add_argument("--micro-points", default="", help="configs[4]: comma-separated cloud sizes (default: the sweep's)")
    ap.add_argument("--micro-hyps", default="", help="configs[4]: comma-separated hypothesis counts")
    ap.add_argument("--micro-max-evals", type=float, default=2.5e10,
                    help="configs[4]: skip checkers-off cells above this many point-hypothesis evaluations")
    args = ap.parse_args()
    cfg = WL.CONFIGS[args.config]
    if args.impl == "reference":
        run_reference(args, cfg)
    elif args.config == 4:
        run_micro(args, cfg)
    else:
        run_frames(args, cfg)


if __name__ == "__main__":
    main()
