"""oracle/ply_io.py -- TEST INFRASTRUCTURE ONLY.

Minimal PLY reader/writer for the layout Open3D 0.17 writes for the reference's memory clouds
(object_memory/object_info.py:109-118, out/360_trial_with_floor/objects/*/pointcloud.ply):
``binary_little_endian``, ``double x,y,z`` and optional ``uchar red,green,blue``; ascii is read too.
"""
from __future__ import annotations

import numpy as np

_T = {"double": "<f8", "float": "<f4", "float64": "<f8", "float32": "<f4", "uchar": "u1", "uint8": "u1",
      "int": "<i4", "int32": "<i4", "uint": "<u4", "short": "<i2", "ushort": "<u2", "char": "i1"}


def read_ply(path):
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, n, props, in_vertex = None, 0, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("unterminated PLY header")
            tok = line.decode("ascii", "replace").split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    n = int(tok[2])
            elif tok[0] == "property" and in_vertex:
                props.append((tok[2], _T[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt == "binary_little_endian":
            data = np.frombuffer(f.read(n * np.dtype(props).itemsize), dtype=np.dtype(props), count=n)
        elif fmt == "ascii":
            raw = np.loadtxt(f, max_rows=n, ndmin=2)
            data = np.zeros(n, dtype=np.dtype(props))
            for k, (name, _) in enumerate(props):
                data[name] = raw[:, k]
        else:
            raise ValueError(f"unsupported PLY format {fmt}")
    pts = np.stack([data["x"], data["y"], data["z"]], axis=1).astype(np.float64)
    cols = None
    if "red" in data.dtype.names:
        cols = np.stack([data["red"], data["green"], data["blue"]], axis=1).astype(np.float64) / 255.0
    return pts, cols


def write_ply(path, pts, cols=None):
    pts = np.asarray(pts, np.float64).reshape(-1, 3)
    props = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")]
    if cols is not None:
        props += [("red", "u1"), ("green", "u1"), ("blue", "u1")]
    data = np.zeros(len(pts), dtype=np.dtype(props))
    data["x"], data["y"], data["z"] = pts[:, 0], pts[:, 1], pts[:, 2]
    if cols is not None:
        c = np.round(np.clip(np.asarray(cols, np.float64).reshape(-1, 3), 0.0, 1.0) * 255.0).astype(np.uint8)
        data["red"], data["green"], data["blue"] = c[:, 0], c[:, 1], c[:, 2]
    hdr = ["ply", "format binary_little_endian 1.0", "comment Created by instance_based_loc_b200",
           f"element vertex {len(pts)}", "property double x", "property double y", "property double z"]
    if cols is not None:
        hdr += ["property uchar red", "property uchar green", "property uchar blue"]
    hdr.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode("ascii"))
        f.write(data.tobytes())
