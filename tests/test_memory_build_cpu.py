"""CPU checks for the memory-build geometry: the oracle's restatement of Open3D's ClusterDBSCAN
loop against an independent implementation (scikit-learn's DBSCAN: same scan order, same border rule), the
order-independent formulation the CUDA kernels compute (csrc/dbscan.cu header) against that sequential loop, and the
golden fixture written by the reference's own ObjectMemory / IoU_ops code (tests/golden/make_build_golden.py)."""
import os

import numpy as np
import pytest

from oracle import o3d_core as oc

G = os.path.join(os.path.dirname(__file__), "golden", "build_golden.npz")


def _clouds():
    rng = np.random.default_rng(3)
    for trial in range(4):
        cs = rng.uniform(-1, 1, (5, 3))
        pts = np.concatenate([c + rng.normal(0, 0.05 + 0.04 * trial, (250, 3)) for c in cs] + [rng.uniform(-1.5, 1.5, (200, 3))])
        yield pts[rng.permutation(len(pts))]


def fixed_point_labels(points, eps, min_points):
    """What csrc/dbscan.cu computes, restated with plain loops: core mask, min-index union-find over core-core pairs,
    clusters ranked by their smallest core index, border points take the smallest adjacent cluster."""
    n = len(points)
    ptr, idx = oc.radius_neighbours(points, eps)
    core = (np.diff(ptr) >= min_points)
    parent = np.arange(n)

    def find(x):
        while parent[x] != x:
            parent[x] = parent[parent[x]]
            x = parent[x]
        return x
    for i in np.nonzero(core)[0]:
        for j in idx[ptr[i]:ptr[i + 1]]:
            if j < i and core[j]:
                a, b = find(i), find(j)
                if a != b:
                    parent[max(a, b)] = min(a, b)
    root = np.full(n, np.iinfo(np.int32).max, np.int64)
    for i in np.nonzero(core)[0]:
        root[i] = find(i)
    is_root = core & (root == np.arange(n))
    rank = np.cumsum(is_root) - is_root
    for i in np.nonzero(~core)[0]:
        nb = idx[ptr[i]:ptr[i + 1]]
        nb = nb[core[nb]]
        if len(nb):
            root[i] = root[nb].min()
    return np.where(root == np.iinfo(np.int32).max, -1, rank[np.minimum(root, n - 1)]).astype(np.int32)


@pytest.mark.parametrize("eps,min_points", [(0.05, 5), (0.1, 10), (0.08, 1), (0.15, 40)])
def test_oracle_dbscan_vs_sklearn_and_fixed_point(eps, min_points):
    from sklearn.cluster import DBSCAN
    for pts in _clouds():
        got = oc.cluster_dbscan(pts, eps, min_points)
        # sklearn includes d == eps, Open3D / nanoflann do not; no pair of these random points sits exactly on eps
        assert np.array_equal(got, DBSCAN(eps=eps, min_samples=min_points).fit(pts).labels_)
        assert np.array_equal(got, fixed_point_labels(pts, eps, min_points))


def test_oracle_dbscan_border_goes_to_first_cluster():
    """A border point within eps of core points of two clusters keeps the cluster that was discovered first."""
    a = np.array([[0, 0, 0], [0.1, 0, 0], [0.2, 0, 0], [0.3, 0, 0]], float)        # one cluster
    b = np.array([[0.9, 0, 0], [1.0, 0, 0], [1.1, 0, 0], [1.2, 0, 0]], float)        # another, 0.6 away
    mid = np.array([[0.6, 0, 0]], float)       # 0.3 from each: 3 neighbours with itself < 4, so a border point of both
    for order in ([0, 1, 2, 3, 4, 5, 6, 7, 8], [8, 4, 5, 6, 7, 0, 1, 2, 3], [7, 8, 0, 1, 2, 3, 4, 5, 6]):
        pts = np.concatenate([a, b, mid])[order]
        lab = oc.cluster_dbscan(pts, 0.31, 4)
        assert lab.max() == 1 and (lab >= 0).all()
        assert np.array_equal(lab, fixed_point_labels(pts, 0.31, 4))
        im = [i for i in range(9) if np.array_equal(pts[i], mid[0])][0]
        first_core = min(i for i in range(9) if i != im)
        assert lab[first_core] == 0 and lab[im] == 0          # the first cluster discovered reaches the border point first
    assert np.array_equal(oc.cluster_dbscan(np.zeros((0, 3)), 0.1, 3), np.zeros(0, np.int32))
    assert np.array_equal(oc.cluster_dbscan(np.zeros((4, 3)), 0.1, 5), np.full(4, -1))       # duplicates, too few
    assert np.array_equal(oc.cluster_dbscan(np.zeros((4, 3)), 0.1, 4), np.zeros(4))          # duplicates, one cluster


def test_build_golden_fixture_consistent():
    g = np.load(G)
    n = int(g["scene_n"])
    union = np.concatenate([g[f"scene_p{j}"] for j in range(n)])
    eps, mp = g["dbscan_eps_minpts"]
    assert np.array_equal(oc.cluster_dbscan(union, float(eps), int(mp)), g["dbscan_labels"])
    # the reference's recluster merged mug+cup, kept sofa, merged shelf+book+vase and dropped the all-noise object
    names = [list(g[f"recl_names{j}"]) for j in range(int(g["recl_n"]))]
    assert names == [["mug", "cup"], ["sofa"], ["shelf", "book", "vase"]]
    assert [int(g[f"recl_id{j}"]) for j in range(3)] == [0, 1, 2]
    assert np.array_equal(g["recl_p0"], np.concatenate([g["scene_p0"], g["scene_p1"]]))
    # embedding-driven variants, as the reference's own methods left the memory
    assert [list(g[f"agg_names{j}"]) for j in range(int(g["agg_n"]))] == [["sofa", "shelf", "book", "vase"], ["mug", "cup", "dust"]]
    assert [list(g[f"cliou_names{j}"]) for j in range(int(g["cliou_n"]))] == [["mug", "cup"], ["sofa"], ["shelf", "book", "vase"]]
    # process_image: the small lamp is skipped, both floor patches end up in .floors
    assert [list(g[f"proc_names{j}"]) for j in range(int(g["proc_n"]))] == [["chair"]]
    assert list(g["proc_floor_names"]) == ["floor"] and len(g["proc_floor_p"]) == 700 + 650


def test_iou_ops_module_has_the_reference_names():
    import importlib
    m = importlib.import_module("instance_based_loc_b200.utils.IoU_ops")
    for name in ("calculate_3d_IoU", "calculate_strict_overlap", "calculate_obj_aligned_3d_IoU"):
        assert callable(getattr(m, name))
    with pytest.raises(NotImplementedError):
        m.calculate_obj_aligned_3d_IoU(None, None)


def test_oracle_oriented_box_properties_and_golden():
    """The oracle's OrientedBoundingBox (Open3D 0.17's algorithm on scipy's Qhull; oracle/o3d_shim) encloses its points,
    has a right-handed orthonormal frame with extents in descending-variance order of the hull vertices, and reproduces the
    fixture that tests/golden/make_obb_golden.py wrote together with the reference's calculate_obj_aligned_3d_IoU values
    (the target for the not-yet-built GPU measure)."""
    from oracle import ref_loader
    o3d = ref_loader.open3d()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "obb_iou_golden.npz"))
    for k in range(int(g["n"])):
        for tag in ("a", "b"):
            pts = g[f"{tag}{k}"]
            bb = o3d.geometry.OrientedBoundingBox.create_from_points(o3d.utility.Vector3dVector(pts))
            assert np.allclose(bb.R.T @ bb.R, np.eye(3), atol=1e-12) and np.linalg.det(bb.R) > 0.999999
            local = (pts - bb.get_center()) @ bb.R
            assert (np.abs(local) <= bb.extent / 2 + 1e-9).all()
            assert np.allclose(np.asarray(bb.get_box_points()), g[f"{tag}{k}_corners"], atol=1e-12)
            assert np.allclose(bb.extent, g[f"{tag}{k}_extent"], atol=1e-12)
    iou = [float(g[f"iou{k}"]) for k in range(int(g["n"]))]
    assert iou[2] == 0.0 and abs(iou[3] - 1.0) < 1e-5 and all(0.0 <= v <= 1.0 + 1e-5 for v in iou)
    assert 0.25 < iou[0] < 0.45                      # a (1, 2, 3) box shifted by half its short side: 1/3 for exact boxes


@pytest.mark.skipif(not os.path.isdir("/root/reference/Objectron/objectron"), reason="reference tree not present")
def test_reference_obj_aligned_iou_matches_fixture():
    """Live in the build container: the unmodified reference function (its vendored Objectron + the oracle's boxes)."""
    import sys
    from oracle import ref_loader
    sys.path.insert(0, os.path.join(ref_loader.REF, "Objectron"))
    iou_mod = ref_loader.load("utils/IoU_ops.py")
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "obb_iou_golden.npz"))
    for k in range(int(g["n"])):
        assert float(iou_mod.calculate_obj_aligned_3d_IoU(g[f"a{k}"], g[f"b{k}"])) == float(g[f"iou{k}"])
