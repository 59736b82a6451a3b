"""Generates tests/golden/build_golden.npz (memory-build geometry) by running the UNMODIFIED
reference code in the build container on top of the oracle's open3d shim:

 * ``utils/IoU_ops.py`` ``calculate_3d_IoU`` / ``calculate_strict_overlap`` on seeded cloud pairs (overlapping, nested,
   touching, disjoint, one empty) -- pure numpy in the reference, so these values ARE the reference's;
 * ``ObjectMemory.remove_points_below_floor``, ``recluster_objects_with_dbscan``, ``recluster_via_agglomerative_clustering``,
   ``recluster_via_combined``, ``_recluster_IoU`` / ``recluster_via_clustering_and_IoU`` (pair measure swapped for the
   reference's axis-aligned ``calculate_3d_IoU``) and ``process_image`` (with and without the noise option, numpy's
   global generator seeded) from object_memory/object_memory.py.  ``cluster_dbscan`` under
   them is the oracle's restatement of Open3D's loop (oracle/o3d_core.c: orc_dbscan), so the labels themselves are
   "parity unpinned" like the other Open3D rows; the reference's glue around them is what this pins.

The detector is out of scope: ``_get_object_info`` is replaced on the instance by a function that returns prepared
(phrases, embeddings, clouds).  The reference tree does not exist on the GPU box; this script is the provenance.
Run: python tests/golden/make_build_golden.py
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import o3d_core, ref_loader  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "build_golden.npz")


def blob(rng, centre, n, sigma):
    return np.asarray(centre, np.float64) + rng.normal(0, sigma, (n, 3))


def recluster_scene(rng):
    """7 objects: two touching pairs / chains that DBSCAN joins, one isolated object, one too sparse to hold a core."""
    return [
        ("mug", blob(rng, (0.0, 0.5, 0.0), 420, 0.04)),
        ("cup", blob(rng, (0.09, 0.52, 0.0), 380, 0.04)),          # touches "mug"
        ("sofa", blob(rng, (2.0, 0.4, 1.0), 500, 0.06)),
        ("shelf", blob(rng, (-1.5, 0.9, -1.0), 450, 0.05)),
        ("book", blob(rng, (-1.38, 0.95, -1.0), 300, 0.05)),       # touches "shelf"
        ("vase", blob(rng, (-1.26, 0.9, -1.02), 310, 0.05)),       # touches "book": a chain of three
        ("dust", rng.uniform(-3, 3, (60, 3)) + np.array([0, 6.0, 0])),   # sparse: all noise, dropped by the reference
    ]


def main():
    rng = np.random.default_rng(20261)
    out = {}
    o3d = ref_loader.open3d()

    def mk(p, c=None):
        pc = o3d.geometry.PointCloud()
        pc.points = o3d.utility.Vector3dVector(np.asarray(p, np.float64).reshape(-1, 3))
        if c is not None:
            pc.colors = o3d.utility.Vector3dVector(c)
        return pc

    # ---------------------------------------------------------------- IoU_ops (pure numpy in the reference)
    ref_loader._stub_force("objectron")
    ref_loader._stub_force("objectron.dataset", iou=None, box=None)
    iou_mod = ref_loader.load("utils/IoU_ops.py")
    a = rng.uniform(-1, 1, (300, 3))
    pairs = [
        (a, rng.uniform(-0.5, 1.5, (250, 3))),                     # partial overlap
        (a, 0.3 * rng.uniform(-1, 1, (100, 3))),                   # nested
        (a, a + np.array([5.0, 0, 0])),                            # disjoint
        (a, rng.uniform(-1, 1, (50, 3)) * np.array([1, 1, 0.01]) + np.array([0, 0, 0.5])),   # thin slab
        (a, np.zeros((0, 3))),                                     # one empty cloud -> 0
        (a[:1], a[:1]),                                            # single point: zero volumes -> nan (0/0)
    ]
    out["iou_n"] = np.array(len(pairs))
    with np.errstate(all="ignore"):
        for k, (p, q) in enumerate(pairs):
            out[f"iou_a{k}"], out[f"iou_b{k}"] = p, q
            out[f"iou_val{k}"] = np.array([iou_mod.calculate_3d_IoU(mk(p), mk(q)),
                                           iou_mod.calculate_strict_overlap(mk(p), mk(q))], np.float64)

    # ---------------------------------------------------------------- ObjectMemory memory-build methods
    om_mod, info_mod = ref_loader.load_object_memory()
    phrases_mod = ref_loader.load("object_memory/object_finder_phrases.py")
    om_mod.check_if_floor = phrases_mod.check_if_floor               # the loader stubs it out for the localise tests

    def new_memory():
        mem = om_mod.ObjectMemory.__new__(om_mod.ObjectMemory)     # no detector set-up
        mem.memory, mem.floors = [], None
        mem.log_enabled = False
        mem.dataset_floor_thickness = 0.1
        mem.object_info_max_embeddings_num = 10 ** 6
        return mem

    def fill(mem, scene, embs):
        for j, (name, p) in enumerate(scene):
            mem.memory.append(info_mod.ObjectInfo(j, name, embs[j], mk(p, np.clip(p * 0.1 + 0.5, 0, 1)), 10 ** 6))

    scene = recluster_scene(rng)
    embs = rng.normal(size=(len(scene), 16)).astype(np.float32)
    out["scene_n"] = np.array(len(scene))
    out["scene_embs"] = embs
    for j, (name, p) in enumerate(scene):
        out[f"scene_name{j}"], out[f"scene_p{j}"] = np.array(name), p

    eps, min_pts = 0.08, 10
    union = np.concatenate([p for _, p in scene])
    out["dbscan_eps_minpts"] = np.array([eps, min_pts])
    out["dbscan_labels"] = o3d_core.cluster_dbscan(union, eps, min_pts)

    def dump(prefix, mem, colors=True):
        out[f"{prefix}_n"] = np.array(len(mem.memory))
        for j, o in enumerate(mem.memory):
            out[f"{prefix}_names{j}"] = np.array(o.names)
            out[f"{prefix}_id{j}"] = np.array(o.id)
            out[f"{prefix}_p{j}"] = np.asarray(o.pointcloud.points)
            if colors:
                out[f"{prefix}_c{j}"] = np.asarray(o.pointcloud.colors)
            out[f"{prefix}_e{j}"] = np.stack([np.asarray(e) for e in o.embeddings])

    with contextlib.redirect_stdout(io.StringIO()):
        mem = new_memory()
        fill(mem, scene, embs)
        mem.recluster_objects_with_dbscan(eps=eps, min_points_per_cluster=min_pts)
        dump("recl", mem)

        mem = new_memory()
        fill(mem, scene[:6], embs)
        mem.remove_points_below_floor()
        dump("floor", mem)

    # embedding-driven reclustering: three embedding groups {mug, cup, dust}, {sofa}, {shelf, book, vase}
    plt = sys.modules["matplotlib.pyplot"]
    plt.imshow = lambda *a, **k: None
    plt.colorbar = lambda *a, **k: None
    base = rng.normal(size=(3, 16))
    grp = [0, 0, 1, 2, 2, 2, 0]
    embs2 = np.stack([base[k] + 0.05 * rng.normal(size=16) for k in grp]).astype(np.float32)
    out["scene_embs2"] = embs2
    aabb_iou = lambda a, b: iou_mod.calculate_3d_IoU(mk(a), mk(b))        # noqa: E731  (the reference's own function)
    with contextlib.redirect_stdout(io.StringIO()):
        mem = new_memory()
        fill(mem, scene, embs2)
        mem.recluster_via_agglomerative_clustering(embedding_distance_threshold=0.4)
        dump("agg", mem, colors=False)

        mem = new_memory()
        fill(mem, scene, embs2)
        mem.recluster_via_combined(embedding_distance_threshold=0.4, eps=eps, min_points_per_cluster=min_pts)
        dump("comb", mem, colors=False)

        # _recluster_IoU with the pair measure swapped for the reference's axis-aligned calculate_3d_IoU (the
        # object-aligned one needs Open3D's Qhull boxes and Objectron, neither available here)
        saved = om_mod.calculate_obj_aligned_3d_IoU
        om_mod.calculate_obj_aligned_3d_IoU = aabb_iou
        try:
            mem = new_memory()
            fill(mem, scene, embs2)
            mem._recluster_IoU(IoU_threshold=0.05)
            dump("iou", mem, colors=False)
            for j, o in enumerate(mem.memory):
                out[f"iou_mean{j}"] = np.asarray(o.mean_emb)

            mem = new_memory()
            fill(mem, scene, embs2)
            mem.recluster_via_clustering_and_IoU(embedding_distance_threshold=0.4, eps=eps, min_points_per_cluster=min_pts,
                                                 IoU_threshold=0.05)
            dump("cliou", mem, colors=False)
        finally:
            om_mod.calculate_obj_aligned_3d_IoU = saved

    # process_image: a chair, a floor patch, a lamp that is too small, a second floor patch (merged into .floors)
    det = [("chair", blob(rng, (0.2, 0.1, 1.5), 600, 0.1)), ("floor", blob(rng, (0.0, -0.8, 2.0), 700, 0.3)),
           ("lamp", blob(rng, (-0.5, 0.4, 1.0), 100, 0.05)), ("floor", blob(rng, (0.8, -0.8, 2.5), 650, 0.3))]
    det_embs = rng.normal(size=(len(det), 16)).astype(np.float32)
    pose = np.array([0.4, -0.2, 1.1, 0.1, -0.3, 0.2, 0.9])
    pose[3:] /= np.linalg.norm(pose[3:])
    out["det_n"], out["det_embs"], out["det_pose"] = np.array(len(det)), det_embs, pose
    for j, (name, p) in enumerate(det):
        out[f"det_name{j}"], out[f"det_p{j}"] = np.array(name), p
    for tag, noise in (("proc", False), ("procnoise", True)):
        mem = new_memory()
        mem._get_object_info = types.MethodType(
            lambda self, *a, **k: ([n for n, _ in det], det_embs, [mk(p, np.clip(p * 0.1 + 0.5, 0, 1)) for _, p in det]), mem)
        np.random.seed(5)
        po = pose.copy()
        with contextlib.redirect_stdout(io.StringIO()):
            mem.process_image("rgb", "depth", po, True, add_noise=noise, min_points=500)
        dump(tag, mem)
        out[f"{tag}_pose_after"] = po
        out[f"{tag}_floor_names"] = np.array(mem.floors.names)
        out[f"{tag}_floor_p"] = np.asarray(mem.floors.pointcloud.points)
        out[f"{tag}_floor_e"] = np.stack([np.asarray(e) for e in mem.floors.embeddings])

    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes;", "recluster ->", int(out["recl_n"]), "objects, floor ->",
          int(out["floor_n"]), "agg/comb/iou/cliou ->", [[list(out[f"{t}_names{j}"]) for j in range(int(out[f"{t}_n"]))]
                                                          for t in ("agg", "comb", "iou", "cliou")],
          "objects, dbscan clusters", int(out["dbscan_labels"].max()) + 1,
          "iou", [out[f"iou_val{k}"].round(4).tolist() for k in range(len(pairs))])


if __name__ == "__main__":
    main()
