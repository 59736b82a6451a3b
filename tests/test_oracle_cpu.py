"""CPU tests of the oracle itself: pinned against the golden vectors generated from the reference's OWN numpy code
(tests/golden/make_golden.py) and against independent numpy / scipy / brute-force restatements for the Open3D part
(PARITY UNPINNED there: nothing in the reference pins Open3D's outputs)."""
import numpy as np
import pytest

from ibl_testutil import random_rigid, rot_angle_deg, surface_cloud


@pytest.fixture(scope="module")
def oc():
    from oracle import o3d_core
    o3d_core.lib()
    return o3d_core


# ------------------------------------------------------------------ pinned by the reference's own code
def test_backproject_port_matches_reference_golden(golden):
    from oracle import ref_port
    g = golden
    out = ref_port.backproject_masks(g["bp_depth"], g["bp_rgb"], g["bp_masks"], float(g["bp_fx"]), float(g["bp_fy"]))
    assert list(np.cumsum([0] + [len(p) for p, _ in out])) == list(g["bp_offsets"])
    assert np.array_equal(np.concatenate([p for p, _ in out]), g["bp_points"])
    assert np.array_equal(np.concatenate([c for _, c in out]), g["bp_colors"])
    p64, c64 = ref_port.backproject(g["bp64_depth"], g["bp_rgb"], float(g["bp_fx"]), float(g["bp_fy"]))
    assert np.array_equal(p64, g["bp64_points"]) and np.array_equal(c64, g["bp64_colors"])


def test_voxel_port_matches_reference_golden(golden):
    from oracle import ref_port
    g = golden
    p, c, _, vid = ref_port.voxel_down_sample(g["vox_pts"], g["vox_cols"], float(g["vox_size"]))
    assert np.array_equal(ref_port.voxel_indices(g["vox_pts"], float(g["vox_size"])), g["vox_keys"])
    assert np.array_equal(p, g["vox_out_pts"]) and np.array_equal(c, g["vox_out_cols"])
    assert vid.max() + 1 == len(p)


def _unflatten(lens, flat):
    out, o = [], 0
    for L in lens:
        out.append([[int(flat[o + 2 * t]), int(flat[o + 2 * t + 1])] for t in range(L)])
        o += 2 * L
    return out


def test_simvolume_port_matches_reference_golden(golden):
    from oracle import ref_port
    g = golden
    for ci in range(int(g["sv_ncases"])):
        S = g[f"sv{ci}_S"]
        _, subs = ref_port.simvolume_subvolumes(S)
        ref0 = g[f"sv{ci}_sub0"]
        # the reference's get_top_indices_from_subvolumes destroys entries (sets them to -inf) while searching:
        # compare where the stored volume is still finite, and the mask (-inf pattern) everywhere else
        fin = np.isfinite(ref0)
        assert np.array_equal(subs[0][fin], ref0[fin])
        assert np.all(np.isneginf(subs[0][np.isneginf(subs[0])]))
        assert ref_port.simvolume_assignments(S, 4) == _unflatten(g[f"sv{ci}_assn_len"], g[f"sv{ci}_assn_flat"]), ci


def test_simvolume_reference_main_smoke(golden):
    """similarity_volume.py:418-452 (cs[i,j] = i + j, 10 x 4): the known answer recorded in SURVEY.md 8(c)."""
    from oracle import ref_port
    cs = np.add.outer(np.arange(10), np.arange(4)).astype(np.float64)
    got = ref_port.simvolume_assignments(cs, 3)
    assert got == _unflatten(golden["svmain_assn_len"], golden["svmain_assn_flat"])
    assert got == [[[7, 3]], [[7, 3], [9, 2]], [[7, 2], [9, 3]], [[7, 3], [8, 2], [9, 1]], [[7, 2], [8, 3], [9, 1]],
                   [[7, 3], [8, 1], [9, 2]]]


def test_svd_transform_known_answers(golden):
    """test.py:1-60 of the reference (three point-triplet cases)."""
    from oracle import ref_port
    for k in range(3):
        T = ref_port.get_SVD_transform(golden[f"svd{k}_p"], golden[f"svd{k}_q"])
        np.testing.assert_allclose(T, golden[f"svd{k}_T"], atol=1e-12)
    np.testing.assert_allclose(golden["svd0_T"], [[0, -1, 0, 0], [1, 0, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]], atol=1e-12)
    np.testing.assert_allclose(golden["svd2_T"][:3, 3], [-1.5, 0.5, 2.5], atol=1e-9)


def test_simvolume_properties():
    """TestSimVolume's properties (similarity_volume.py:276-385): entry == fp16 product, repeats are -inf, the
    first chosen detection can never be unassigned."""
    from oracle import ref_port
    rng = np.random.default_rng(0)
    S = rng.uniform(0.1, 0.95, (4, 6)).astype(np.float32)
    chosen, subs = ref_port.simvolume_subvolumes(S)
    a = np.ones((4, 7), np.float16)
    a[:, :6] = S
    for ch, vol in zip(chosen, subs):
        assert vol[1, 2, 3] == np.float16(np.float16(a[ch[0], 1] * a[ch[1], 2]) * a[ch[2], 3])
        assert np.isneginf(vol[2, 2, 0]) and np.isneginf(vol[1, 0, 1]) and np.isneginf(vol[0, 3, 3])
        assert np.all(np.isneginf(vol[6]))                      # mask[[-1]*(s+1)] = -inf slab (Appendix B.3)
        assert np.isfinite(vol[0, 6, 6]) and np.isfinite(vol[3, 6, 1])


# ------------------------------------------------------------------ Open3D restatement: internal consistency
def test_knn_hybrid_vs_bruteforce(oc):
    p, _ = surface_cloud(700, 1)
    q = p[::5] + 0.01
    for radius, k in [(0.1, 30), (0.03, 30), (np.inf, 7)]:
        idx, d2, cnt = oc.knn_hybrid(p, q, radius, k)
        D = ((q[:, None, :] - p[None]) ** 2)
        D = (D[..., 0] + D[..., 1]) + D[..., 2]
        for i in range(len(q)):
            order = np.lexsort((np.arange(len(p)), D[i]))
            ok = order[D[i][order] < radius * radius][:k]
            assert cnt[i] == len(ok) and np.array_equal(idx[i, :cnt[i]], ok)
            assert np.array_equal(d2[i, :cnt[i]], D[i][ok])


def test_radius_count_and_nn1_vs_bruteforce(oc):
    p, _ = surface_cloud(900, 2)
    D = ((p[:, None, :] - p[None]) ** 2)
    D = (D[..., 0] + D[..., 1]) + D[..., 2]
    assert np.array_equal(oc.radius_count(p, 0.05), (D < 0.05 * 0.05).sum(1))
    q = p[:200] + 0.004
    idx, d2 = oc.nn1(p, q, 0.02)
    Dq = ((q[:, None, :] - p[None]) ** 2)
    Dq = (Dq[..., 0] + Dq[..., 1]) + Dq[..., 2]
    j = Dq.argmin(1)
    ok = Dq[np.arange(len(q)), j] < 0.02 ** 2
    assert np.array_equal(idx >= 0, ok) and np.array_equal(idx[ok], j[ok])


def test_fast_eigen_matches_eigh(oc):
    import ctypes
    rng = np.random.default_rng(3)
    for _ in range(200):
        A = rng.normal(size=(3, 3))
        C = A @ A.T * rng.uniform(1e-4, 10)
        cov6 = np.array([C[0, 0], C[0, 1], C[0, 2], C[1, 1], C[1, 2], C[2, 2]])
        out = np.zeros(3)
        oc.lib().orc_fast_eigen3x3(oc._p(cov6), oc._p(out))
        w, v = np.linalg.eigh(C)
        assert abs(abs(out @ v[:, 0]) - 1) < 1e-6
    out = np.zeros(3)
    oc.lib().orc_fast_eigen3x3(oc._p(np.array([3.0, 0, 0, 1.0, 0, 2.0])), oc._p(out))      # diagonal branch
    assert np.array_equal(out, [0, 1, 0])
    oc.lib().orc_fast_eigen3x3(oc._p(np.zeros(6)), oc._p(out))                              # zero matrix -> zero
    assert np.array_equal(out, [0, 0, 0])
    ctypes  # noqa: B018


def test_kabsch_matches_svd(oc):
    from oracle import ref_port
    rng = np.random.default_rng(4)
    for k in (3, 4, 50):
        s = rng.normal(size=(k, 3))
        T = random_rigid(k)
        t = s @ T[:3, :3].T + T[:3, 3] + rng.normal(0, 1e-3, (k, 3))
        np.testing.assert_allclose(oc.kabsch(s, t), ref_port.get_SVD_transform(s, t), atol=1e-9)
    # degenerate (collinear) triplet still returns a proper rotation
    s = np.array([[0, 0, 0], [1, 0, 0], [2, 0, 0.0]])
    R = oc.kabsch(s, s + 0.5)[:3, :3]
    np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-12)
    assert abs(np.linalg.det(R) - 1) < 1e-12


def test_fpfh_basic_properties(oc):
    p, _ = surface_cloud(1500, 5)
    n = oc.estimate_normals(p, 0.1, 30)
    np.testing.assert_allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-12)
    f, s = oc.compute_fpfh(p, n, 0.25, 100, return_spfh=True)
    assert f.shape == (33, 1500) and np.all(f >= 0)
    for blk in range(3):                                        # every 11-bin block: 100 (SPFH) + 100 (weighted sum)
        np.testing.assert_allclose(s[11 * blk:11 * blk + 11].sum(0), 100.0, atol=1e-9)
        np.testing.assert_allclose(f[11 * blk:11 * blk + 11].sum(0), 200.0, atol=1e-6)


def test_ransac_sequential_semantics(oc):
    """A.7: the result and the adaptive stop do not depend on how the stream is cut into blocks; est_k only shrinks."""
    tgt, _ = surface_cloud(2500, 7)
    T = random_rigid(8, 0.5, 0.2)
    sel = np.arange(0, 2500, 2)
    src = (np.linalg.inv(T) @ np.c_[tgt[sel], np.ones(len(sel))].T).T[:, :3]
    rng = np.random.default_rng(9)
    cor = np.stack([np.arange(len(sel)), sel], axis=1).astype(np.int32)[::3]
    bad = rng.random(len(cor)) < 0.6
    cor[bad, 1] = rng.integers(0, 2500, bad.sum())
    s = rng.integers(0, len(cor), (8000, 3), dtype=np.int32)
    a, ia = oc.ransac(src, tgt, cor, s, 0.075, block=8000)
    b, ib = oc.ransac(src, tgt, cor, s, 0.075, block=37)
    assert ia == ib and np.array_equal(a.transformation, b.transformation) and a.n_inliers == b.n_inliers
    assert ia["drawn"] < 8000 and rot_angle_deg(a.transformation[:3, :3], T[:3, :3]) < 3
    from oracle import sampler
    c, ic = oc.ransac(src, tgt, cor, lambda h0, h1: sampler.samples(5, h0, h1 - h0, len(cor)), 0.075, max_iteration=8000)
    d, id_ = oc.ransac(src, tgt, cor, sampler.samples(5, 0, 8000, len(cor)), 0.075)
    assert ic == id_ and np.array_equal(c.transformation, d.transformation)


def test_sampler_is_uniform_and_in_range():
    from oracle import sampler
    s = sampler.samples(123, 0, 200000, 97)
    assert s.shape == (200000, 3) and s.min() == 0 and s.max() == 96
    cnt = np.bincount(s.reshape(-1), minlength=97)
    assert cnt.min() > 0.9 * cnt.mean() and cnt.max() < 1.1 * cnt.mean()
    assert np.array_equal(sampler.samples(123, 1000, 10, 97), s[1000:1010])          # counter based: random access


def test_ply_roundtrip(tmp_path, golden):
    from oracle import ply_io
    p, c = golden["fix_points"], golden["fix_colors"]
    ply_io.write_ply(str(tmp_path / "a.ply"), p, c)
    p2, c2 = ply_io.read_ply(str(tmp_path / "a.ply"))
    assert np.array_equal(p, p2) and np.array_equal(c, c2)      # colours are k/255: exact through uchar


def test_oracle_pipeline_small():
    """the oracle's localise flow end to end on a tiny synthetic scene (also the CPU baseline of bench.py)."""
    from instance_based_loc_b200 import synth
    from oracle import ref_pipeline, ref_port
    scene = synth.make_scene(m=4, seed=3, D=32, size=(0.35, 0.6))
    host = []
    for j in range(scene.m):
        p, c = synth.memory_cloud(scene, j, 0.03)
        p2, c2, _, _ = ref_port.voxel_down_sample(p, c, 0.03)
        host.append({"points": p2, "colors": c2, "embeddings": scene.memory_embs[j]})
    fr = synth.make_frames(scene, 1, H=120, W=160, f=80.0, seed=1, min_pixels=80)[0]
    pose, assn, info = ref_pipeline.localise_frame(fr, host, 80.0, 80.0, seed=3, fpfh_global_dist_factor=1.5,
                                                   fpfh_local_dist_factor=1.5, fpfh_voxel_size=0.05)
    assert pose.shape == (7,) and abs(np.linalg.norm(pose[3:]) - 1) < 1e-9
    assert 1 <= len(info["assignments"]) <= 6 and len(info["registrations"]) == len(info["assignments"])
    assert info["assignments"][0][0][0] == 0                    # length-1 assignment of the largest detection first
    none = ref_pipeline.localise_core([], None, host)
    assert np.array_equal(none[0], [0, 0, 0, 0, 0, 0, 1]) and none[1] == [[], []]
