"""SURVEY.md 8(f) row N1 -- the memory-bank on-disk format on the input side of the path: a pkl written by the UNMODIFIED
reference ObjectMemory.save_to_pkl (tests/golden/make_memory_fixture.py; the pickle names the reference's own class path
object_memory.object_info.ObjectInfo) is read WITHOUT the reference tree, records mapped onto this package's ObjectInfo."""
import os
import sys

import numpy as np

from instance_based_loc_b200.object_memory.object_info import ObjectInfo
from instance_based_loc_b200.object_memory.object_memory import ObjectMemory

G = os.path.join(os.path.dirname(__file__), "golden")


def test_reference_written_pkl_reads_without_reference_tree(monkeypatch):
    for k in ("object_memory", "object_memory.object_info"):          # nothing of the reference is importable here
        monkeypatch.delitem(sys.modules, k, raising=False)
    exp = np.load(os.path.join(G, "ref_memory_expect.npz"))
    mem, floors = ObjectMemory.read_pkl(os.path.join(G, "ref_memory.pkl"))
    assert len(mem) == 3
    for j, (info, pts, cols) in enumerate(mem):
        assert type(info) is ObjectInfo and info.pointcloud is None
        assert "pcd" not in info.__dict__ and "pcd_colors" not in info.__dict__       # the reference's eager copies
        assert info.names == [f"obj{j}", f"alias{j}"]
        assert np.array_equal(np.stack(info.embeddings), exp[f"e{j}"])
        assert np.array_equal(info.mean_emb, exp[f"mean{j}"]) and np.array_equal(info.centroid, exp[f"centroid{j}"])
        assert pts.dtype == np.float64 and np.array_equal(pts, exp[f"p{j}"]) and np.array_equal(cols, exp[f"c{j}"])
        assert info.max_embeddings_num == 10
        assert info.id is id                                          # reference quirk (:793): the builtin, not an int
    # reference quirk (:804-824): the floors record is rebuilt from the LAST object
    finfo, fp, fc = floors
    assert type(finfo) is ObjectInfo and finfo.names == mem[-1][0].names and np.array_equal(fp, exp["p2"]) and np.array_equal(fc, exp["c2"])
