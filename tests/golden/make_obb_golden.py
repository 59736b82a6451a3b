"""Generates tests/golden/obb_iou_golden.npz: outputs of the UNMODIFIED reference ``calculate_obj_aligned_3d_IoU``
(utils/IoU_ops.py:99-144) on seeded cloud pairs, run in the build container with

 * the reference's own vendored Objectron code (/root/reference/Objectron/objectron/dataset/{box,iou}.py) for the
   box-box intersection (Sutherland-Hodgman clipping + Qhull volume), and
 * the oracle shim's ``OrientedBoundingBox.create_from_points`` (oracle/o3d_shim: Open3D 0.17's published algorithm on
   scipy's Qhull) for the oriented boxes -- Open3D itself is not installable here, so that half is "parity unpinned".

The product does not compute this measure yet (a GPU convex hull is the missing piece); the fixture
is the target the next step is built against.  Run: python tests/golden/make_obb_golden.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "obb_iou_golden.npz")


def box_cloud(rng, centre, ext, yaw, n):
    p = rng.uniform(-0.5, 0.5, (n, 3)) * np.asarray(ext, float)
    R = np.array([[np.cos(yaw), -np.sin(yaw), 0], [np.sin(yaw), np.cos(yaw), 0], [0, 0, 1.0]])
    return p @ R.T + np.asarray(centre, float)


def pairs(rng):
    a = box_cloud(rng, (0, 0, 0), (1, 2, 3), 0.0, 400)
    return [
        (a, box_cloud(rng, (0.5, 0, 0), (1, 2, 3), 0.0, 400)),                # shifted copy
        (a, box_cloud(rng, (0, 0, 0), (1, 2, 3), 0.6, 400)),                  # rotated about z
        (a, a + 10.0),                                                       # disjoint -> 0
        (a, a.copy()),                                                       # identical -> 1 (to float32 rounding)
        (a, box_cloud(rng, (0.1, 0.2, 0.3), (0.3, 0.4, 0.5), 1.1, 300)),      # nested small box
        (rng.normal(size=(350, 3)) * (0.5, 0.2, 0.1), rng.normal(size=(350, 3)) * (0.2, 0.5, 0.1) + (0.2, 0, 0)),
    ]


def main():
    sys.path.insert(0, os.path.join(ref_loader.REF, "Objectron"))            # IoU_ops.py appends ./Objectron (cwd-relative)
    iou_mod = ref_loader.load("utils/IoU_ops.py")
    o3d = ref_loader.open3d()
    rng = np.random.default_rng(77)
    out = {}
    ps = pairs(rng)
    out["n"] = np.array(len(ps))
    for k, (p, q) in enumerate(ps):
        out[f"a{k}"], out[f"b{k}"] = p, q
        out[f"iou{k}"] = np.array(iou_mod.calculate_obj_aligned_3d_IoU(p, q), np.float64)
        for tag, pts in (("a", p), ("b", q)):
            bb = o3d.geometry.OrientedBoundingBox.create_from_points(o3d.utility.Vector3dVector(pts))
            out[f"{tag}{k}_center"], out[f"{tag}{k}_extent"] = np.asarray(bb.get_center()), np.asarray(bb.extent)
            out[f"{tag}{k}_corners"] = np.asarray(bb.get_box_points())
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes; IoU:", [round(float(out[f"iou{k}"]), 4) for k in range(len(ps))])


if __name__ == "__main__":
    main()
