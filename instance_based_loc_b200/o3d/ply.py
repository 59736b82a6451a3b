"""open3d.io.{read,write}_point_cloud for the PLY layout the reference's memory uses
(object_memory/object_info.py:109-118: binary little-endian, double x,y,z + uchar r,g,b).  Host I/O only."""
from __future__ import annotations

import numpy as np

from .pointcloud import PointCloud

_T = {"double": "<f8", "float": "<f4", "float64": "<f8", "float32": "<f4", "uchar": "u1", "uint8": "u1",
      "int": "<i4", "int32": "<i4", "uint": "<u4", "short": "<i2", "ushort": "<u2", "char": "i1"}


def read_ply_arrays(path):
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError(f"{path}: not a PLY file")
        fmt, n, props, in_vertex = None, 0, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError(f"{path}: unterminated PLY header")
            tok = line.decode("ascii", "replace").split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                n = int(tok[2]) if in_vertex else n
            elif tok[0] == "property" and in_vertex:
                props.append((tok[2], _T[tok[1]]))
            elif tok[0] == "end_header":
                break
        dt = np.dtype(props)
        if fmt == "binary_little_endian":
            data = np.frombuffer(f.read(n * dt.itemsize), dtype=dt, count=n)
        elif fmt == "ascii":
            raw = np.loadtxt(f, max_rows=n, ndmin=2)
            data = np.zeros(n, dtype=dt)
            for k, (name, _) in enumerate(props):
                data[name] = raw[:, k]
        else:
            raise ValueError(f"{path}: unsupported PLY format {fmt}")
    names = data.dtype.names
    pts = np.stack([data["x"], data["y"], data["z"]], axis=1).astype(np.float64)
    cols = nrm = None
    if "red" in names:
        cols = np.stack([data["red"], data["green"], data["blue"]], axis=1).astype(np.float64) / 255.0
    if "nx" in names:
        nrm = np.stack([data["nx"], data["ny"], data["nz"]], axis=1).astype(np.float64)
    return pts, cols, nrm


def read_point_cloud(path, **kw):
    pts, cols, nrm = read_ply_arrays(path)
    r = PointCloud(pts)
    if cols is not None:
        r.colors = cols
    if nrm is not None:
        r.normals = nrm
    return r


def write_point_cloud(path, pcd, write_ascii=False, compressed=False, print_progress=False):
    pts = np.asarray(pcd.points)
    has_c, has_n = pcd.has_colors(), pcd.has_normals()
    props = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")]
    if has_n:
        props += [("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8")]
    if has_c:
        props += [("red", "u1"), ("green", "u1"), ("blue", "u1")]
    data = np.zeros(len(pts), dtype=np.dtype(props))
    data["x"], data["y"], data["z"] = pts[:, 0], pts[:, 1], pts[:, 2]
    if has_n:
        nn = np.asarray(pcd.normals)
        data["nx"], data["ny"], data["nz"] = nn[:, 0], nn[:, 1], nn[:, 2]
    if has_c:
        c = np.round(np.clip(np.asarray(pcd.colors), 0.0, 1.0) * 255.0).astype(np.uint8)
        data["red"], data["green"], data["blue"] = c[:, 0], c[:, 1], c[:, 2]
    tname = {"<f8": "double", "u1": "uchar"}
    hdr = ["ply", "format binary_little_endian 1.0", "comment Created by instance_based_loc_b200",
           f"element vertex {len(pts)}"] + [f"property {tname[t]} {nm}" for nm, t in props] + ["end_header"]
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode("ascii"))
        f.write(data.tobytes())
    return True
