"""Summaries of ncu output for profiles/ (read in the build container; ncu -i needs no GPU).

  python tools/ncu_summary.py launches gpurun_out/launches.csv "header text"      per-kernel time table of a launch list
  python tools/ncu_summary.py full gpurun_out/full_X.ncu-rep [...]                key metrics of --set full captures
  python tools/ncu_summary.py traffic TAG=gpurun_out/full_X.ncu-rep [...]         profiles/ncu_traffic.json: DRAM bytes per
                                                                                  launch of a kernel group (bench.py roofline.traffic)
"""
from __future__ import annotations

import collections
import csv
import io
import re
import subprocess
import sys

FULL_METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tmem.sum",
    "sm__inst_executed_pipe_tensor.sum",
]

# extra metrics to pass to ncu beside --set full (comma-joined): the pipes the path is actually bound by
EXTRA_METRICS = ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active,"
                 "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,"
                 "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tmem.sum,"
                 "sm__inst_executed_pipe_tensor.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,"
                 "lts__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,"
                 "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")

# kernel function -> bench.py kernel group (csrc IBL_PROF tags)
GROUPS = [("feat_tc_kernel", "feature_nn"), ("feat_tc_", "feature_nn"), ("knn_kernel", "knn"), ("fpfh_kernel", "fpfh"),
          ("spfh_kernel", "spfh"), ("normals_kernel", "normals"), ("ransac_count", "ransac_score"), ("ransac_score", "ransac_score"),
          ("ransac_finalize", "ransac_score"), ("ransac_check", "ransac_check"), ("colored_step", "icp_reduce"),
          ("reduceK", "icp_reduce"), ("nn_score", "nn_score"), ("score_final", "nn_score"), ("radius_count", "radius_count"),
          ("cellid", "grid_build"), ("gather_sorted", "grid_build"), ("bbox", "grid_build"), ("DeviceRadixSort", "grid_build"),
          ("DeviceScan", "grid_build"), ("nf_", "grid_build"), ("seqsum3", "mean3"), ("simvol", "simvolume"),
          ("cosine", "cosine_segmax"), ("backproject", "backproject"), ("voxel", "voxel"), ("color_gradient", "color_gradient")]


def bounds(path, out_json="profiles/r02_kernel_bounds.json"):
    """Per kernel of a --set full capture: which unit it is bound by.  The candidates are the percent-of-peak counters of
    DRAM, L2, L1/TEX, the fp64 pipe, the tensor pipe and instruction issue; the largest one names the bound, and anything
    whose largest counter is below 35 % is latency-bound (dependent loads / small grids).  Writes the table bench.py merges
    into its `kernels` object (per kernel group: the group's heaviest kernel)."""
    import json
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    col = {m: hdr.index(m) for m in hdr}
    kn = hdr.index("Kernel Name")

    def val(r, m):
        try:
            return float(r[col[m]].replace(",", ""))
        except Exception:
            return 0.0
    cand = {"hbm": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l2": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "l1": "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "fp64": "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
            "tensor": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active"}
    per_kernel, groups = [], {}
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[kn])[:70]
        pct = {k: val(r, m) for k, m in cand.items() if m in col}
        top = max(pct, key=pct.get) if pct else "latency"
        bound = top if pct.get(top, 0) >= 35.0 else "latency"
        us = val(r, "gpu__time_duration.sum") * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(units[col["gpu__time_duration.sum"]], 1.0)
        rec = {"kernel": name, "us": round(us, 1), "grid": int(val(r, "launch__grid_size")), "regs": int(val(r, "launch__registers_per_thread")),
               "bound": bound, "pct_of_peak": {k: round(v, 1) for k, v in pct.items()},
               "warps_active_pct": round(val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
               "dram_bytes": val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")}
        per_kernel.append(rec)
        for pat, grp in GROUPS:
            if pat in name:
                if grp not in groups or groups[grp]["_us"] < us:
                    groups[grp] = {"bound": bound, "frac_of_that_peak": round(pct.get(top, 0) / 100.0, 3), "bound_kernel": name,
                                   "bound_counters_pct": rec["pct_of_peak"], "_us": us}
                break
    for g in groups.values():
        g.pop("_us")
    print(f"{'kernel':60s} {'us':>9s} {'grid':>7s} {'bound':>8s}  " + " ".join(f"{k:>6s}" for k in cand))
    for rec in sorted(per_kernel, key=lambda x: -x["us"]):
        print(f"{rec['kernel'][:60]:60s} {rec['us']:9.1f} {rec['grid']:7d} {rec['bound']:>8s}  " +
              " ".join(f"{rec['pct_of_peak'].get(k, 0):6.1f}" for k in cand))
    json.dump(groups, open(out_json, "w"), indent=1, sort_keys=True)
    json.dump(per_kernel, open(out_json.replace(".json", "_per_kernel.json"), "w"), indent=1)



def launches(path, header=""):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0.0, 0])
    tot, n = 0.0, 0
    for row in r:
        if len(row) <= vi:
            continue
        try:
            v = float(row[vi].replace(",", ""))
        except ValueError:
            continue
        u = row[ui]
        us = v / 1e3 if u in ("ns", "nsecond") else (v if u in ("us", "usecond") else v * 1e3)
        name = re.sub(r"\(.*", "", row[ki])[:80]
        agg[name][0] += us
        agg[name][1] += 1
        tot += us
        n += 1
    if header:
        print(header)
    print(f"# total {tot / 1e3:.2f} ms of kernel time in {n} launches (cold-cache, serialised per-launch times: compare SHARES)")
    for k, (t, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:32]:
        print(f"{t / 1e3:10.3f} ms {100 * t / tot:6.2f}%  n={c:5d} avg={t / c:9.1f} us  {k}")


def full(paths):
    for p in paths:
        out = subprocess.run(["ncu", "-i", p, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        if len(rows) < 3:
            print(f"== {p}: no data")
            continue
        hdr, units = rows[0], rows[1]
        kn = hdr.index("Kernel Name")
        print(f"== {p}")
        for r in rows[2:]:
            print("  kernel:", re.sub(r"\(.*", "", r[kn])[:90])
            for m in FULL_METRICS:
                if m in hdr:
                    i = hdr.index(m)
                    print(f"    {m:78s} {r[i]:>16s} {units[i]}")


def _bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def traffic(specs):
    """TAG=file.ncu-rep: the capture holds the kernels of ONE launch of the kernel group TAG (e.g. the 8 kernels of one
    ibl_feature_nn call, or one ransac_score_kernel launch); traffic = sum of dram read + write bytes over them."""
    import json
    import os
    out = {}
    for spec in specs:
        tag, path = spec.split("=", 1)
        nlaunch = 1                                       # TAG=path:N -- the capture holds N launches of the group
        if ":" in path and path.rsplit(":", 1)[1].isdigit():
            path, nl = path.rsplit(":", 1)
            nlaunch = int(nl)
        txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        ir, iw, kn, it = (hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("Kernel Name"),
                          hdr.index("gpu__time_duration.sum"))
        tot, names, dur = 0.0, [], 0.0
        for r in rows[2:]:
            tot += _bytes(r[ir], units[ir]) + _bytes(r[iw], units[iw])
            names.append(re.sub(r"\(.*", "", r[kn])[:40])
            dur += float(r[it].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(units[it], 1.0)
        out[tag] = {"bytes_per_launch": tot / nlaunch, "source": f"ncu --set full, {os.path.basename(path)}",
                    "launch": f"{nlaunch} launch(es) of the group captured, {len(names)} kernel(s): {', '.join(names)}; "
                              f"{dur:.1f} us under ncu in total"}
    os.makedirs("profiles", exist_ok=True)
    json.dump(out, open("profiles/ncu_traffic.json", "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "bounds":
        bounds(*sys.argv[2:4])
    elif sys.argv[1] == "extra-metrics":
        print(EXTRA_METRICS)
    elif sys.argv[1] == "traffic":
        traffic(sys.argv[2:])
    elif sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
    else:
        full(sys.argv[2:])
