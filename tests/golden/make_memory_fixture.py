"""Generates tests/golden/ref_memory.pkl: a small memory bank written by the UNMODIFIED reference
``ObjectMemory.save_to_pkl`` (object_memory/object_memory.py:779-826, imported from /root/reference on the oracle's
open3d shim) -- the on-disk format on the input side of the path.  The pickle names the
reference's own class path ``object_memory.object_info.ObjectInfo`` and carries its quirks (``id`` is the BUILTIN
function, :793; the floors tuple is built from the last object, :804-824).  Also writes the arrays that went in
(ref_memory_expect.npz).  Run: python tests/golden/make_memory_fixture.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402


def main():
    om_mod, info_mod = ref_loader.load_object_memory()
    o3d = ref_loader.open3d()
    rng = np.random.default_rng(77)
    mem = om_mod.ObjectMemory.__new__(om_mod.ObjectMemory)          # no detector set-up
    mem.memory = []
    mem.log_enabled = False
    expect = {}
    for j, (n, k) in enumerate([(300, 1), (120, 3), (57, 2)]):
        p, c = rng.normal(j, 0.2, (n, 3)), rng.random((n, 3))
        embs = [rng.normal(size=24).astype(np.float32) for _ in range(k)]
        pc = o3d.geometry.PointCloud()
        pc.points = o3d.utility.Vector3dVector(p)
        pc.colors = o3d.utility.Vector3dVector(c)
        oi = info_mod.ObjectInfo(j, f"obj{j}", embs[0], pc, 10)
        for e in embs[1:]:
            oi._add_embedding(e)
        oi._add_name(f"alias{j}")
        oi._compute_means()
        mem.memory.append(oi)
        expect[f"p{j}"], expect[f"c{j}"], expect[f"e{j}"] = p, c, np.stack(embs)
        expect[f"mean{j}"], expect[f"centroid{j}"] = oi.mean_emb, oi.centroid
    here = os.path.dirname(os.path.abspath(__file__))
    # pickle records classes by module path: the loader un-registers the reference's modules after import, so put them
    # back for the duration of the dump
    import types
    saved = {k: sys.modules.get(k) for k in ("object_memory", "object_memory.object_info")}
    pkg = types.ModuleType("object_memory")
    pkg.__path__ = []
    sys.modules.update({"object_memory": pkg, "object_memory.object_info": info_mod})
    try:
        mem.save_to_pkl(os.path.join(here, "ref_memory.pkl"))
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    np.savez_compressed(os.path.join(here, "ref_memory_expect.npz"), **expect)
    print("wrote ref_memory.pkl", os.path.getsize(os.path.join(here, "ref_memory.pkl")), "bytes")


if __name__ == "__main__":
    main()
