"""GPU parity tests for the memory-build geometry, through the C ABI: ``ibl_dbscan`` labels
bit-exact against the oracle's sequential restatement of Open3D's ClusterDBSCAN, ``ibl_bounds3`` / utils.IoU_ops against
values computed by the reference's own utils/IoU_ops.py, and ObjectMemory.process_image / remove_points_below_floor /
recluster_objects_with_dbscan against what the UNMODIFIED reference methods produced on the same inputs
(tests/golden/make_build_golden.py -> build_golden.npz).  The file sorts last so the hot-path suites run first."""
import os

import numpy as np
import pytest
import torch

from ibl_testutil import surface_cloud

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(__file__), "golden", "build_golden.npz")


@pytest.fixture(scope="module")
def K(cuda):
    from instance_based_loc_b200 import kernels
    return kernels


@pytest.fixture(scope="module")
def oc():
    from oracle import o3d_core
    o3d_core.lib()
    return o3d_core


@pytest.fixture(scope="module")
def g():
    return np.load(G)


def _pc(K, p, c=None):
    from instance_based_loc_b200.o3d.pointcloud import PointCloud
    return PointCloud.from_tensors(K.to_dev(np.asarray(p, np.float64).reshape(-1, 3)), None if c is None else K.to_dev(c))


def _labels(K, pts, eps, mp):
    lab, nc = _pc(K, pts).cluster_dbscan_device(eps, mp)
    lab = lab.cpu().numpy()
    assert nc == (int(lab.max()) + 1 if len(lab) else 0)
    return lab


def test_dbscan_vs_oracle(K, oc):
    rng = np.random.default_rng(3)
    for trial in range(4):
        cs = rng.uniform(-1, 1, (5, 3))
        pts = np.concatenate([c + rng.normal(0, 0.05 + 0.04 * trial, (250, 3)) for c in cs] + [rng.uniform(-1.5, 1.5, (200, 3))])
        pts = pts[rng.permutation(len(pts))]
        for eps, mp in ((0.05, 5), (0.1, 10), (0.08, 1), (0.15, 40), (0.4, 150)):
            assert np.array_equal(_labels(K, pts, eps, mp), oc.cluster_dbscan(pts, eps, mp)), (trial, eps, mp)


def test_dbscan_border_and_degenerate(K, oc):
    a = np.array([[0, 0, 0], [0.1, 0, 0], [0.2, 0, 0], [0.3, 0, 0]], float)
    b = np.array([[0.9, 0, 0], [1.0, 0, 0], [1.1, 0, 0], [1.2, 0, 0]], float)
    mid = np.array([[0.6, 0, 0]], float)                     # border point of both clusters: the first one keeps it
    for order in ([0, 1, 2, 3, 4, 5, 6, 7, 8], [8, 4, 5, 6, 7, 0, 1, 2, 3], [7, 8, 0, 1, 2, 3, 4, 5, 6]):
        pts = np.concatenate([a, b, mid])[order]
        assert np.array_equal(_labels(K, pts, 0.31, 4), oc.cluster_dbscan(pts, 0.31, 4))
    assert len(_labels(K, np.zeros((0, 3)), 0.1, 3)) == 0
    assert np.array_equal(_labels(K, np.zeros((4, 3)), 0.1, 5), np.full(4, -1))         # duplicates, too few: noise
    assert np.array_equal(_labels(K, np.zeros((4, 3)), 0.1, 4), np.zeros(4))            # duplicates, one cluster
    assert np.array_equal(_labels(K, np.array([[1.0, 2.0, 3.0]]), 0.1, 1), np.zeros(1))  # a single point is its own cluster
    # points exactly eps apart are NOT neighbours (nanoflann's d2 < r2): a lattice with spacing == eps has no pairs
    lat = np.stack(np.meshgrid(*[np.arange(4) * 0.25] * 3, indexing="ij"), -1).reshape(-1, 3)
    assert np.array_equal(_labels(K, lat, 0.25, 2), np.full(64, -1))
    assert np.array_equal(_labels(K, lat, 0.25, 2), oc.cluster_dbscan(lat, 0.25, 2))
    assert np.array_equal(_labels(K, lat, 0.2501, 2), np.zeros(64))


def test_dbscan_surface_cloud_full_size(K, oc):
    """Memory-build scale: 30 k surface points in five bodies (two of them touching) plus scattered outliers, with the
    reference's eps 0.2 / 300 points (object_memory.py:296) -- ~500 neighbours per point -- and two tighter settings."""
    rng = np.random.default_rng(11)
    offs = ([0, 0, 0], [1.3, 0, 0], [4, 0, 0], [4, 3, 0], [-4, 1, 2])
    parts = [surface_cloud(6000, 100 + k, scale=0.6)[0] + np.asarray(off, float) for k, off in enumerate(offs)]
    pts = np.concatenate(parts + [rng.uniform(-6, 6, (300, 3))])
    pts = np.ascontiguousarray(pts[rng.permutation(len(pts))])
    for eps, mp in ((0.2, 300), (0.1, 40), (0.05, 12)):
        want = oc.cluster_dbscan(pts, eps, mp)
        assert want.max() >= 1
        assert np.array_equal(_labels(K, pts, eps, mp), want), (eps, mp)


def test_dbscan_golden_labels(K, g):
    union = np.concatenate([g[f"scene_p{j}"] for j in range(int(g["scene_n"]))])
    eps, mp = g["dbscan_eps_minpts"]
    assert np.array_equal(_labels(K, union, float(eps), int(mp)), g["dbscan_labels"])
    from instance_based_loc_b200.o3d.pointcloud import PointCloud
    host = _pc(K, union).cluster_dbscan(eps=float(eps), min_points=int(mp), print_progress=False)
    assert isinstance(host, np.ndarray) and np.array_equal(np.array(host), g["dbscan_labels"])
    assert isinstance(PointCloud(), PointCloud)


def test_bounds_and_iou_golden(K, g):
    from instance_based_loc_b200.utils import IoU_ops
    rng = np.random.default_rng(5)
    big = rng.normal(size=(300001, 3)) * np.array([1.0, 1e6, 1e-6])
    lo, hi = K.bounds3(K.to_dev(big))
    assert np.array_equal(lo, big.min(axis=0)) and np.array_equal(hi, big.max(axis=0))
    pc = _pc(K, big)
    assert np.array_equal(pc.get_min_bound(), lo) and np.array_equal(pc.get_max_bound(), hi)
    with np.errstate(all="ignore"):
        for k in range(int(g["iou_n"])):
            a, b = _pc(K, g[f"iou_a{k}"]), _pc(K, g[f"iou_b{k}"])
            got = np.array([IoU_ops.calculate_3d_IoU(a, b), IoU_ops.calculate_strict_overlap(a, b)], np.float64)
            assert np.array_equal(got, g[f"iou_val{k}"], equal_nan=True), (k, got, g[f"iou_val{k}"])


def _memory(K, g, n=None, embs="scene_embs"):
    from instance_based_loc_b200.object_memory.object_memory import ObjectMemory
    om = ObjectMemory("cuda", None, None, 300.0, 300.0, get_embeddings_func=None, log_enabled=False,
                      dataset_floor_thickness=0.1)
    for j in range(int(g["scene_n"]) if n is None else n):
        p = g[f"scene_p{j}"]
        om.add_object(str(g[f"scene_name{j}"]), g[embs][j], _pc(K, p, np.clip(p * 0.1 + 0.5, 0, 1)))
    return om


def _check_memory(om, g, prefix, atol=0.0):
    assert len(om.memory) == int(g[f"{prefix}_n"])
    for j, o in enumerate(om.memory):
        assert list(o.names) == list(g[f"{prefix}_names{j}"]) and o.id == int(g[f"{prefix}_id{j}"])
        p, c = np.asarray(o.pointcloud.points), np.asarray(o.pointcloud.colors)
        assert p.shape == g[f"{prefix}_p{j}"].shape
        if atol == 0.0:
            assert np.array_equal(p, g[f"{prefix}_p{j}"])
        else:
            assert np.abs(p - g[f"{prefix}_p{j}"]).max() <= atol
        if f"{prefix}_c{j}" in g:
            assert np.array_equal(c, g[f"{prefix}_c{j}"])
        else:
            assert c.shape == p.shape
        assert np.array_equal(np.stack([np.asarray(e) for e in o.embeddings]), g[f"{prefix}_e{j}"])


def test_recluster_objects_with_dbscan_golden(K, g):
    om = _memory(K, g)
    eps, mp = g["dbscan_eps_minpts"]
    om.recluster_objects_with_dbscan(eps=float(eps), min_points_per_cluster=int(mp))
    _check_memory(om, g, "recl")


def test_remove_points_below_floor_golden(K, g):
    om = _memory(K, g, n=6)
    om.remove_points_below_floor()
    _check_memory(om, g, "floor")


@pytest.mark.parametrize("tag,noise", [("proc", False), ("procnoise", True)])
def test_process_image_golden(K, g, tag, noise, tmp_path):
    """transform_pointcloud is R p + t: BLAS order in the reference, ((R0 x + R1 y) + R2 z) + t in the kernel -> 1e-12."""
    om = _memory(K, g, n=0)
    n = int(g["det_n"])
    det = [(str(g[f"det_name{j}"]), g[f"det_p{j}"]) for j in range(n)]
    om._get_object_info = lambda *a, **k: ([nm for nm, _ in det], g["det_embs"],
                                           [_pc(K, p, np.clip(p * 0.1 + 0.5, 0, 1)) for _, p in det])
    np.random.seed(5)
    pose = g["det_pose"].copy()
    om.process_image("rgb", "depth", pose, True, add_noise=noise, min_points=500)
    assert np.array_equal(pose, g[f"{tag}_pose_after"])
    _check_memory(om, g, tag, atol=1e-12)
    assert list(om.floors.names) == list(g[f"{tag}_floor_names"])
    fp = np.asarray(om.floors.pointcloud.points)
    assert fp.shape == g[f"{tag}_floor_p"].shape and np.abs(fp - g[f"{tag}_floor_p"]).max() <= 1e-12
    assert np.array_equal(np.stack([np.asarray(e) for e in om.floors.embeddings]), g[f"{tag}_floor_e"])
    # ObjectMemory.save (:750-777): memory.txt, the two combined clouds, one directory per object and for the floors
    from instance_based_loc_b200.o3d import io as o3d_io
    om.save(str(tmp_path / "mem"))
    assert open(tmp_path / "mem" / "memory.txt").read() == repr(om)
    comb = o3d_io.read_point_cloud(str(tmp_path / "mem" / "combined_pointcloud_with_floor.ply"))
    want = np.concatenate([np.asarray(o.pointcloud.points) for o in om.memory] + [fp])
    assert np.array_equal(np.asarray(comb.points), want)
    one = o3d_io.read_point_cloud(str(tmp_path / "mem" / "objects" / "0" / "pointcloud.ply"))
    assert np.array_equal(np.asarray(one.points), np.asarray(om.memory[0].pointcloud.points))
    assert os.path.exists(tmp_path / "mem" / "floors" / "all_floors" / "info.pkl")
    assert os.path.exists(tmp_path / "mem" / "combined_pointcloud.ply")
    om.downsample_all_objects(0.05)                              # floors are down-sampled with the objects (:258-263)
    assert om.floors.pointcloud.n() < len(fp)


@pytest.mark.parametrize("tag", ["agg", "comb", "iou", "cliou"])
def test_embedding_and_iou_reclustering_golden(K, g, tag):
    """recluster_via_agglomerative_clustering / recluster_via_combined / _recluster_IoU / recluster_via_clustering_and_IoU
    against the unmodified reference methods (the IoU ones with the pair measure set to the axis-aligned calculate_3d_IoU
    on both sides: the object-aligned measure is not built and must raise, never fall back)."""
    from instance_based_loc_b200.utils import IoU_ops
    eps, mp = float(g["dbscan_eps_minpts"][0]), int(g["dbscan_eps_minpts"][1])
    om = _memory(K, g, embs="scene_embs2")
    if tag == "agg":
        om.recluster_via_agglomerative_clustering(embedding_distance_threshold=0.4)
    elif tag == "comb":
        om.recluster_via_combined(embedding_distance_threshold=0.4, eps=eps, min_points_per_cluster=mp)
    elif tag == "iou":
        om._recluster_IoU(IoU_threshold=0.05, iou_func=IoU_ops.calculate_3d_IoU)
        for j, o in enumerate(om.memory):
            assert np.array_equal(o.mean_emb, g[f"iou_mean{j}"])
        with pytest.raises(NotImplementedError):
            _memory(K, g, embs="scene_embs2")._recluster_IoU()
    else:
        om.recluster_via_clustering_and_IoU(embedding_distance_threshold=0.4, eps=eps, min_points_per_cluster=mp,
                                            IoU_threshold=0.05, iou_func=IoU_ops.calculate_3d_IoU)
    _check_memory(om, g, tag)


@pytest.mark.parametrize("m,seed,H,W,f", [(10, 3, 60, 80, 40.0), (40, 1002, 600, 600, 300.0), (25, 7, 97, 131, 60.0), (3, 5, 1, 1, 1.0)])
def test_raycast_kernel_vs_tensor_ops(K, m, seed, H, W, f):
    """ibl_raycast_scene (one CUDA kernel, thread per pixel) against the tensor-op ray caster on the CPU.
    Same first hit per pixel; depth and normal agree to rounding (the tensor ops go through BLAS dot products)."""
    from instance_based_loc_b200 import synth
    sc = synth.make_scene(m=m, seed=seed, D=16)
    T = synth.camera_pose(sc, np.random.default_rng(seed))
    b0, i0, n0 = [x.numpy() for x in synth.render(sc, T, H=H, W=W, f=f, want_geometry=True)]
    b1, i1, n1 = [x.cpu().numpy() for x in synth.render(sc, T, H=H, W=W, f=f, device="cuda", use_kernel=True, want_geometry=True)]
    same = i0 == i1
    assert same.mean() >= 0.9999                       # a pixel exactly on a silhouette may round either way
    assert np.array_equal(np.isinf(b0[same]), np.isinf(b1[same]))
    fin = same & np.isfinite(b0)
    assert fin.sum() == 0 or np.abs(b0 - b1)[fin].max() <= 1e-9
    assert np.abs(n0 - n1)[same].max() <= 1e-9
    if H > 1:
        assert (i1 >= 0).sum() > 0
        d, rgb, inst = synth.render(sc, T, H=H, W=W, f=f, device="cuda", use_kernel=True)
        assert d.shape == (H, W) and d.dtype == torch.float32 and rgb.shape == (H, W, 3) and rgb.dtype == torch.uint8
        assert np.array_equal(inst.cpu().numpy().reshape(-1), np.concatenate([sc.obj, [-1]])[i1])   # primitive -> object
