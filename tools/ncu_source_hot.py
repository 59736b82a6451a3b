#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump (SASS rows with executed-instruction counts and stall samples) per kernel:
totals, and the SASS address ranges mapped to CUDA source lines through `nvdisasm --print-line-info` of the object file.

  python tools/ncu_source_hot.py gpurun_out/src_top_source.csv.gz [kernel-substring] [--obj build/x.o --func mangled-substring]
"""
import csv, gzip, io, re, subprocess, sys, collections


def kernels(path):
    txt = gzip.open(path, "rt").read() if path.endswith(".gz") else open(path).read()
    blocks = re.split(r'(?m)^"Kernel Name",', txt)
    for b in blocks[1:]:
        name, _, rest = b.partition("\n")
        rows = list(csv.reader(io.StringIO(rest)))
        hdr = rows[0]
        yield name.strip('",'), hdr, [r for r in rows[1:] if len(r) == len(hdr)]


def line_map(obj, func):
    """offset -> (file, line) for one function from nvdisasm -g"""
    if not obj.endswith(".cubin"):          # extract the sm_100a cubin of an object file next to it
        import os, tempfile
        d = tempfile.mkdtemp()
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
        obj = os.path.join(d, [f for f in os.listdir(d) if f.endswith(".cubin")][0])
    out = subprocess.run(["nvdisasm", "-g", "-c", obj], capture_output=True, text=True).stdout
    m, cur, on = {}, None, False
    for ln in out.splitlines():
        if ".section" in ln and ".text." in ln:
            on = func in ln
        if not on:
            continue
        mm = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if mm:
            cur = (mm.group(1).split("/")[-1], int(mm.group(2)))
        mm = re.search(r'/\*([0-9a-f]{4,})\*/\s+\S', ln)
        if mm and cur:
            m[int(mm.group(1), 16)] = cur
    return m


def main():
    path = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else ""
    obj = func = None
    if "--obj" in sys.argv:
        obj = sys.argv[sys.argv.index("--obj") + 1]
        func = sys.argv[sys.argv.index("--func") + 1]
    for name, hdr, rows in kernels(path):
        if want not in name:
            continue
        ia, isamp, iinst, isrc = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
        base = int(rows[0][ia], 16)
        tot_i = sum(int(r[iinst]) for r in rows)
        tot_s = sum(int(r[isamp]) for r in rows)
        print(f"== {name[:90]}\n   {len(rows)} SASS rows, {tot_i:,} warp instructions executed, {tot_s:,} samples")
        if obj:
            lm = line_map(obj, func)
            by = collections.defaultdict(lambda: [0, 0])
            for r in rows:
                k = lm.get(int(r[ia], 16) - base, ("?", 0))
                by[k][0] += int(r[iinst]); by[k][1] += int(r[isamp])
            for k, (i, s) in sorted(by.items(), key=lambda kv: -kv[1][1])[:40]:
                print(f"   {k[0]}:{k[1]:<5d} inst {100.0 * i / tot_i:5.1f}%  samples {100.0 * s / max(tot_s, 1):5.1f}%")
        else:
            top = sorted(rows, key=lambda r: -int(r[isamp]))[:25]
            for r in top:
                print(f"   +{int(r[ia], 16) - base:05x} samples {100.0 * int(r[isamp]) / max(tot_s, 1):5.1f}% inst {int(r[iinst]):>10,}  {r[isrc].strip()[:70]}")


main()
