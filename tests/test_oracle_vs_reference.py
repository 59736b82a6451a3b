"""Build-container-only checks (skipped where /root/reference does not exist, e.g. on the GPU box): the UNMODIFIED
reference modules run on the oracle's open3d shim and are compared with the array-level oracle that the GPU parity
tests use.  This pins oracle/ref_port.py and the orchestration in oracle/ref_pipeline.py to the reference's own code;
the Open3D arithmetic underneath both stays PARITY UNPINNED (open3d is not installable here)."""
import contextlib
import io

import numpy as np
import pytest

from oracle import ref_loader

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="reference tree not present")


def test_reference_depth_utils_vs_port():
    import torch
    from oracle import ref_port
    du = ref_loader.load("utils/depth_utils.py")
    o3d = ref_loader.open3d()
    rng = np.random.default_rng(5)
    depth = rng.uniform(0.5, 3.0, (40, 56)).astype(np.float32)
    depth[rng.random(depth.shape) < 0.3] = 0
    rgb = rng.integers(0, 256, (40, 56, 3), dtype=np.uint8)
    masks = rng.random((3, 1, 40, 56)) < 0.5
    clouds = du.get_mask_coloured_pointclouds_from_depth(depth, rgb, torch.from_numpy(masks), 60.0, 60.0, None)
    port = ref_port.backproject_masks(depth, rgb, masks, 60.0, 60.0)
    for c, (p, col) in zip(clouds, port):
        assert np.array_equal(np.asarray(c.points), p) and np.array_equal(np.asarray(c.colors), col)
    pts = rng.normal(0, 0.05, (800, 3))
    pc = o3d.geometry.PointCloud()
    pc.points = o3d.utility.Vector3dVector(pts)
    pc.colors = o3d.utility.Vector3dVector(rng.random((800, 3)))
    ds = du.voxel_down_sample_with_colors(pc, 0.02)
    p, c, _, _ = ref_port.voxel_down_sample(pts, np.asarray(pc.colors), 0.02)
    assert np.array_equal(np.asarray(ds.points), p) and np.array_equal(np.asarray(ds.colors), c)
    pose = np.array([0.1, -0.2, 0.3, 0.1, 0.2, 0.3, 0.9])
    tp = du.transform_pointcloud(pc, pose.copy())
    np.testing.assert_allclose(np.asarray(tp.points), ref_port.transform_pointcloud(pts, pose.copy()), atol=1e-15)


@pytest.mark.parametrize("nd,m", [(1, 4), (2, 5), (3, 4), (5, 7)])
def test_reference_simvolume_vs_port(nd, m):
    from oracle import ref_port
    sv = ref_loader.load("utils/similarity_volume.py")
    S = np.random.default_rng(nd * 10 + m).uniform(-0.2, 0.95, (nd, m)).astype(np.float32)
    with contextlib.redirect_stderr(io.StringIO()):
        v = sv.SimVolume(S.copy())
        v.fast_construct_volume(min(nd, 3))
        ref_subs = [np.array(s) for s in v.subvolumes]
        a = v.get_top_indices_from_subvolumes(num_per_length=4)
    a = [[[int(x), int(y)] for x, y in assn] for assn in a]
    assert a == ref_port.simvolume_assignments(S, 4)
    assert len(ref_subs) == len(ref_port.simvolume_subvolumes(S)[1])


def test_reference_localise_vs_oracle_pipeline():
    """object_memory.py:852-1169, unmodified, fed synthetic detections through a stubbed _get_object_info, against
    oracle/ref_pipeline.localise_core on the same arrays and the same RANSAC draw stream."""
    from instance_based_loc_b200 import synth
    from oracle import ref_pipeline, ref_port, sampler
    om_mod, info_mod = ref_loader.load_object_memory()
    o3d = ref_loader.open3d()
    seed = 17
    scene = synth.make_scene(m=5, seed=seed, D=48, size=(0.35, 0.6))
    host = []
    for j in range(scene.m):
        p, c = synth.memory_cloud(scene, j, 0.03)
        p2, c2, _, _ = ref_port.voxel_down_sample(p, c, 0.03)
        host.append({"points": p2, "colors": c2, "embeddings": scene.memory_embs[j]})
    fr = synth.make_frames(scene, 1, H=135, W=180, f=90.0, seed=2, min_pixels=100)[0]
    cfg = {"radius_nb_points": 8, "radius": 0.05}
    det = ref_pipeline.detect_clouds(fr["depth"], fr["rgb"], fr["masks"], 90.0, 90.0, cfg)

    def mk(p, c):
        pc = o3d.geometry.PointCloud()
        pc.points = o3d.utility.Vector3dVector(p)
        pc.colors = o3d.utility.Vector3dVector(c)
        return pc

    mem = om_mod.ObjectMemory.__new__(om_mod.ObjectMemory)          # no detector set-up
    mem.memory = []
    for j, hm in enumerate(host):
        oi = info_mod.ObjectInfo(j, f"obj{j}", hm["embeddings"][0], mk(hm["points"], hm["colors"]), 10 ** 6)
        oi.embeddings = list(hm["embeddings"])
        oi._compute_means()
        mem.memory.append(oi)
    mem._get_object_info = lambda *a, **k: ([f"o{i}" for i in range(len(det))], np.array(fr["embs"], np.float32).copy(),
                                            [mk(p, c) for p, c in det])
    o3d.ransac_sampler = lambda C, H: (lambda h0, h1: sampler.samples(seed, h0, h1 - h0, C))
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf), contextlib.redirect_stderr(io.StringIO()):
        pose, assn = mem.localise("x.png", "x.npy", fpfh_global_dist_factor=1.5, fpfh_local_dist_factor=1.5,
                                  fpfh_voxel_size=0.05, perform_semantic_icp=False)
    rpose, rassn, rinfo = ref_pipeline.localise_core(det, fr["embs"], host, fpfh_global_dist_factor=1.5,
                                                     fpfh_local_dist_factor=1.5, fpfh_voxel_size=0.05, seed=seed)
    assert [[int(a), int(b)] for a, b in assn[0]] == rassn[0]
    np.testing.assert_allclose(pose, rpose, atol=1e-9)
    assert "Best assn" in buf.getvalue()
    assert not any("fallback" in r["detail"] for r in rinfo["registrations"])


@pytest.mark.parametrize("nd,m", [(2, 5), (3, 4), (4, 3)])
def test_reference_simvolume_legacy_api_vs_mirror_helpers(nd, m):
    """The legacy SimVolume surface of SURVEY 8(b).1 (construct_volume, construct_volume_choose_e, .subvolumes,
    get_top_indices; similarity_volume.py:30-96,156-194): the mirror's host helpers against the unmodified reference."""
    from instance_based_loc_b200.utils import similarity_volume as mine
    sv = ref_loader.load("utils/similarity_volume.py")
    S = np.random.default_rng(nd * 7 + m).uniform(-0.2, 0.95, (nd, m)).astype(np.float32)
    with contextlib.redirect_stderr(io.StringIO()), contextlib.redirect_stdout(io.StringIO()):
        v = sv.SimVolume(S.copy())
        vol, rep = v.construct_volume()
        v.fast_construct_volume(min(nd, 3))
        subs = [np.array(s) for s in v.subvolumes]
        chosen = [list(c) for c in v.chosen_objects]
        ce = v.construct_volume_choose_e(list(range(nd))[::-1])
        top = v.get_top_indices(np.array(rep, copy=True), 5)
    aug = np.ones((nd, m + 1), np.float16)
    aug[:, :-1] = S
    mvol = mine._outer_chain(list(aug))
    assert mvol.dtype == vol.dtype and np.array_equal(mvol, vol)
    mrep = np.where(mine._validity_mask(m, nd), mvol, np.float16(-np.inf))
    assert np.array_equal(mrep, rep)
    allowed = mine._validity_mask(m, min(nd, 3))
    for c, s in zip(chosen, subs):
        assert np.array_equal(np.where(allowed, mine._outer_chain([aug[i] for i in c]), np.float16(-np.inf)), s)
    assert np.array_equal(mine._outer_chain([aug[i] for i in range(nd)][::-1]), ce)
    got = mine.SimVolume.get_top_indices(None, np.array(mrep, copy=True), 5)
    assert [tuple(int(x) for x in a) for a, _ in got] == [tuple(int(x) for x in a) for a, _ in top]
