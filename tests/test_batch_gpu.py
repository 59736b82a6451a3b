"""Batched registration entry points (ibl_ransac_run_batch, ibl_icp_colored_batch, ibl_nn_score_batch) against the
single-problem calls they replace: every result field must be IDENTICAL -- the batch only changes how many problems share
a launch and a host round trip, never a problem's arithmetic, hypothesis stream or reduction order.  The single-problem
calls are the ones the oracle parity tests pin (test_kernels_gpu.py: test_ransac_run, test_colored_icp, test_nn_score)."""
import numpy as np
import pytest
import torch

from ibl_testutil import random_rigid, surface_cloud

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K(cuda):
    from instance_based_loc_b200 import kernels
    return kernels


def dev(a, dtype=torch.float64):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda()


def _case(seed, n_tgt, n_src, C, outlier):
    tgt, tcol = surface_cloud(n_tgt, seed)
    T = random_rigid(seed + 1, 0.6, 0.3)
    rng = np.random.default_rng(seed + 2)
    sel = np.sort(rng.choice(n_tgt, n_src, replace=False))
    src = (np.linalg.inv(T) @ np.c_[tgt[sel], np.ones(len(sel))].T).T[:, :3] + rng.normal(0, 0.001, (n_src, 3))
    ci = rng.choice(n_src, C, replace=False)
    cor = np.stack([ci, sel[ci]], axis=1).astype(np.int32)
    bad = rng.random(C) < outlier
    cor[bad, 1] = rng.integers(0, n_tgt, bad.sum())
    return dict(src=src, scol=tcol[sel], tgt=tgt, tcol=tcol, cor=cor, T=T)


@pytest.fixture(scope="module")
def problems(K):
    """five problems of different sizes / difficulty (one with too few correspondences, one tiny)."""
    from oracle import o3d_core as oc
    out = []
    for seed, n_tgt, n_src, C, outl in ((100, 6000, 3500, 1200, 0.7), (200, 9000, 5100, 900, 0.85), (300, 2500, 700, 300, 0.5),
                                        (400, 4000, 2000, 2, 0.0), (500, 5000, 513, 400, 0.9)):
        c = _case(seed, n_tgt, n_src, C, outl)
        tn = oc.estimate_normals(c["tgt"], 0.1, 30)
        P, N, Cc = dev(c["tgt"]), dev(tn), dev(c["tcol"])
        idx, _, cnt = K.knn_hybrid(K.Grid(P, K.knn_cell_for(P, 30, 0.15)), 0.15, 30, want_d2=False)
        grad = K.color_gradient(P, N, Cc, idx, cnt)
        g = K.Grid(P, 0.075, nnfield=0.075, nnfield_voxel=0.075 / 8)
        out.append(dict(c, d_src=dev(c["src"]), d_scol=dev(c["scol"]), d_tgt=P, d_tnrm=N, d_tcol=Cc, d_grad=grad, grid=g,
                        d_cor=dev(c["cor"], torch.int32), seed=seed))
    return out


def _specs(problems):
    return [dict(src=p["d_src"], src_col=p["d_scol"], tgt=p["d_tgt"], tgt_nrm=p["d_tnrm"], tgt_col=p["d_tcol"],
                 tgt_grad=p["d_grad"], grid_ransac=p["grid"], grid_icp=p["grid"], corres=p["d_cor"], seed=p["seed"])
            for p in problems]


@pytest.mark.parametrize("early_exit,H,block", [(True, 60000, 4096), (False, 20000, 1000), (True, 300000, 1 << 14)])
def test_ransac_batch_equals_single(K, problems, early_exit, H, block):
    arr = K.reg_problems(_specs(problems))
    got = K.ransac_run_batch(arr, H, 0.99, 0.075, edge_thr=0.9, use_checkers=True, early_exit=early_exit, block=block)
    for p, r in zip(problems, got):
        one = K.ransac_run(p["grid"], p["d_src"], p["d_tgt"], p["d_cor"], H, 0.99, 0.075, edge_thr=0.9, seed=p["seed"],
                           use_checkers=True, early_exit=early_exit, block=block)
        for k in ("best_hyp", "n_inliers", "validated", "drawn", "n_corres", "status", "fitness", "inlier_rmse", "err2"):
            assert r[k] == one[k], (k, r[k], one[k])
        assert np.array_equal(r["T"], one["T"])
    assert got[3]["status"] == 1 and np.array_equal(got[3]["T"], np.eye(4))      # 2 correspondences: empty result
    assert got[0]["fitness"] > 0.5


def test_ransac_batch_prune_switch(K, problems):
    """the exact early exit of the count pass changes no result field"""
    arr = K.reg_problems(_specs(problems))
    try:
        K.ransac_set_prune(False)
        a = K.ransac_run_batch(arr, 200000, 0.99, 0.075, early_exit=True, block=2048)
        K.ransac_set_prune(True)
        b = K.ransac_run_batch(arr, 200000, 0.99, 0.075, early_exit=True, block=2048)
    finally:
        K.ransac_set_prune(True)
    for x, y in zip(a, b):
        assert {k: v for k, v in x.items() if k != "T"} == {k: v for k, v in y.items() if k != "T"}
        assert np.array_equal(x["T"], y["T"])


def test_ransac_batch_checkers_off(K, problems):
    """configs[4]'s forced H_v = H: every hypothesis is scored."""
    sub = [problems[2], problems[4]]
    arr = K.reg_problems(_specs(sub))
    got = K.ransac_run_batch(arr, 3000, 0.99, 0.075, use_checkers=False, early_exit=False, block=1024)
    for p, r in zip(sub, got):
        one = K.ransac_run(p["grid"], p["d_src"], p["d_tgt"], p["d_cor"], 3000, 0.99, 0.075, seed=p["seed"],
                           use_checkers=False, early_exit=False, block=1024)
        assert r["validated"] == one["validated"] == 3000
        assert r["best_hyp"] == one["best_hyp"] and r["n_inliers"] == one["n_inliers"] and np.array_equal(r["T"], one["T"])
        assert r["inlier_rmse"] == one["inlier_rmse"]


def test_icp_colored_batch_equals_single(K, problems):
    live = [p for p in problems if len(p["cor"]) > 3]
    arr = K.reg_problems(_specs(live))
    inits = [p["T"] @ random_rigid(40 + k, 0.03, 0.01) for k, p in enumerate(live)]
    inits[1] = np.eye(4)                                          # a hopeless start: runs all 30 iterations or stalls
    got = K.icp_colored_batch(arr, np.stack(inits), 0.075)
    for p, T0, r in zip(live, inits, got):
        one = K.icp(p["grid"], p["d_src"], p["d_tgt"], 0.075, T0, mode=1, scol=p["d_scol"], tnrm=p["d_tnrm"], tcol=p["d_tcol"],
                    tgrad=p["d_grad"])
        for k in ("iterations", "n_inliers", "fitness", "inlier_rmse", "err2"):
            assert r[k] == one[k], (k, r[k], one[k])
        assert np.array_equal(r["T"], one["T"])
    assert len({r["iterations"] for r in got}) > 1                # problems converge after different numbers of steps


def test_nn_score_batch_equals_single(K, problems):
    g = problems[1]["grid"]
    srcs = [problems[0]["d_src"], problems[1]["d_src"], torch.zeros((0, 3), dtype=torch.float64, device="cuda"),
            problems[4]["d_src"], problems[1]["d_src"]]
    Ts = [problems[0]["T"], problems[1]["T"], np.eye(4), random_rigid(9, 0.2, 0.1), np.eye(4)]
    got = K.nn_score_batch(g, srcs, Ts, 0.02)
    for s, T, (n_in, err2) in zip(srcs, Ts, got):
        if s.shape[0] == 0:
            assert (n_in, err2) == (0, 0.0)
            continue
        a, b, _, _ = K.nn_score(g, s, 0.02, T=T)
        assert (n_in, err2) == (a, b)
    assert got[1][0] > 0.9 * srcs[1].shape[0]


def test_localise_batch_modes_agree(cuda):
    """localise_batch with the batched registration stage == the per-frame organisation == localise_arrays, frame by
    frame (poses bit-equal), including the RANSAC bookkeeping (drawn / validated) per registration."""
    from instance_based_loc_b200 import selfcheck, synth
    from instance_based_loc_b200.o3d import registration as reg
    scene = synth.make_scene(m=8, seed=23, D=96, size=(0.35, 0.75))
    om, _ = selfcheck.build_memory(scene, spacing=0.015, voxel=0.015, f=150.0)
    frames = synth.make_frames(scene, 6, H=225, W=300, f=150.0, seed=23, min_pixels=200)
    reg.ransac_seed = 23
    kw = dict(fpfh_global_dist_factor=1.5, fpfh_local_dist_factor=1.5, fpfh_voxel_size=0.05, perform_semantic_icp=False)
    frs = [{k: f[k] for k in ("rgb", "depth", "masks", "embs")} for f in frames]
    seq = []
    for f in frs:
        pose, assn = om.localise_arrays(f["rgb"], f["depth"], f["masks"], f["embs"], **kw)
        seq.append((pose, assn, om.last_localise_info))
    for workers, batched in ((1, True), (3, True), (3, False)):
        got = om.localise_batch(frs, workers=workers, batched_registration=batched, **kw)
        for (p0, a0, i0), (p1, a1, i1) in zip(seq, got):
            assert np.array_equal(p0, p1) and a0 == a1
            assert len(i0["registrations"]) == len(i1["registrations"])
            for g0, g1 in zip(i0["registrations"], i1["registrations"]):
                assert np.array_equal(g0["transform"], g1["transform"])
                assert g0["full_fitness"] == g1["full_fitness"] and g0["fitness"] == g1["fitness"]
                assert g0["ransac"].get("drawn") == g1["ransac"].get("drawn")
                assert g0["ransac"].get("validated") == g1["ransac"].get("validated")
                assert g0["icp"].get("iterations") == g1["icp"].get("iterations")


@pytest.mark.parametrize("na,nb", [(700, 900), (128, 128), (5000, 3000), (1, 300), (300, 1)])
def test_feature_match_packed_equals_generic(K, na, nb):
    """cached-pack matcher (both directions, memory-side centre / scale) == two ibl_feature_nn calls, including planted
    exact duplicates (lowest index wins) and rows far from the pack's centre."""
    rng = np.random.default_rng(na * 7 + nb)
    base = rng.gamma(2.0, 12.0, (40, 33))
    fb = base[rng.integers(0, 40, nb)] + rng.normal(0, 1.5, (nb, 33))
    fa = base[rng.integers(0, 40, na)] + rng.normal(0, 1.5, (na, 33))
    if nb > 10:
        fb[7] = fb[3]                                   # exact duplicates in the pack
        fa[0] = fb[3]
    if na > 10:
        fa[5] = fa[2]
        fa[9] *= 1.9                                    # outside the pack's own range, inside its spare bit
    A, B = dev(fa), dev(fb)
    pack = K.FeaturePack(B)
    ab, ba, flag = K.feature_match_packed(A, pack)
    assert int(flag.item()) == 0
    assert np.array_equal(ab.cpu().numpy(), K.feature_nn(A, B).cpu().numpy())
    assert np.array_equal(ba.cpu().numpy(), K.feature_nn(B, A).cpu().numpy())
    # a source far outside the pack's fp16 range raises the flag (the Python mirror then takes the general path)
    _, _, flag = K.feature_match_packed(A * 1e6, pack)
    assert int(flag.item()) != 0


def test_feature_match_packed_sorted_targets_ties_and_exhausted_pool(K):
    """The pack's target rows are sorted by norm inside the library (per-tile bands): ties between exact duplicates must still
    go to the lowest ORIGINAL index, in the chunked path, in the single-candidate shortcut and when the candidate pool runs
    dry (exact evaluation on the spot).  Many duplicated rows with very different norms, pool capped to a few chunks."""
    rng = np.random.default_rng(11)
    base = rng.gamma(2.0, 12.0, (25, 33)) * rng.uniform(0.2, 3.0, (25, 1))      # prototypes with a 15x spread of norms
    nb, na = 3000, 700
    pick_b = rng.integers(0, 25, nb)
    fb = base[pick_b] + (rng.random((nb, 1)) < 0.5) * rng.normal(0, 0.8, (nb, 33))   # half of the rows are EXACT copies of a prototype
    fa = base[rng.integers(0, 25, na)] + (rng.random((na, 1)) < 0.5) * rng.normal(0, 0.8, (na, 33))
    A, B = dev(fa), dev(fb)
    want_ab = K.feature_nn(A, B).cpu().numpy()
    want_ba = K.feature_nn(B, A).cpu().numpy()
    d = ((fa[:, None, :] - fb[None, :, :]) ** 2).sum(-1)
    assert np.array_equal(want_ab, d.argmin(1))             # numpy's argmin takes the first (lowest) index among exact ties
    pack = K.FeaturePack(B)
    try:
        for chunks in (0, 8):                               # default pool, then one that runs dry at once
            K.feature_nn_set_pool_chunks(chunks)
            ab, ba, flag = K.feature_match_packed(A, pack)
            assert int(flag.item()) == 0
            assert np.array_equal(ab.cpu().numpy(), want_ab), chunks
            assert np.array_equal(ba.cpu().numpy(), want_ba), chunks
    finally:
        K.feature_nn_set_pool_chunks(0)


@pytest.mark.parametrize("na,nb", [(700, 3000), (3000, 700), (130, 129), (5, 900)])
def test_feature_match_packed_mutual_subset(K, na, nb):
    """ibl_feature_match_packed_mutual: nn_ab equal to the full call; nn_ba equal to the full call on every pack row that
    occurs in nn_ab -- the only rows Open3D's mutual filter reads -- and -1 on the others; the filtered correspondences are
    therefore identical.  Exact duplicates on both sides, default pool and a pool that runs dry."""
    rng = np.random.default_rng(na + 3 * nb)
    base = rng.gamma(2.0, 12.0, (30, 33)) * rng.uniform(0.3, 2.5, (30, 1))
    fb = base[rng.integers(0, 30, nb)] + (rng.random((nb, 1)) < 0.6) * rng.normal(0, 0.8, (nb, 33))
    fa = base[rng.integers(0, 30, na)] + (rng.random((na, 1)) < 0.6) * rng.normal(0, 0.8, (na, 33))
    A, B = dev(fa), dev(fb)
    pack = K.FeaturePack(B)
    ab0, ba0, _ = K.feature_match_packed(A, pack)
    try:
        for chunks in (0, 8):
            K.feature_nn_set_pool_chunks(chunks)
            ab, ba, flag = K.feature_match_packed_mutual(A, pack)
            assert int(flag.item()) == 0
            ab_h, ba_h, ab0_h, ba0_h = ab.cpu().numpy(), ba.cpu().numpy(), ab0.cpu().numpy(), ba0.cpu().numpy()
            assert np.array_equal(ab_h, ab0_h), chunks
            chosen = np.zeros(nb, bool)
            chosen[ab0_h] = True
            assert np.array_equal(ba_h[chosen], ba0_h[chosen]), chunks
            assert np.all(ba_h[~chosen] == -1), chunks
            assert np.array_equal(K.mutual_corres(ab, ba, 3).cpu().numpy(), K.mutual_corres(ab0, ba0, 3).cpu().numpy())
    finally:
        K.feature_nn_set_pool_chunks(0)


def test_shared_knn_search_equals_separate_searches(K):
    """normals (Hybrid 0.1 / 30), colour gradients (0.15 / 30) and FPFH (0.25 / 100) derived from ONE 0.25 / 100 search are
    the neighbour lists of their own searches: same counts, same indices in the same order."""
    from instance_based_loc_b200.o3d.pointcloud import PointCloud
    p, c = surface_cloud(20000, 3)
    a, b = PointCloud(p), PointCloud(p)
    a.colors, b.colors = c, c
    big = a.knn(0.25, 100, want_d2=True)
    for r, k in ((0.1, 30), (0.15, 30), (0.25, 100), (0.02, 10)):
        idx, _, cnt = a.knn(r, k, want_d2=False)                      # derived
        b._grids.pop("knn_last", None)
        ridx, rd2, rcnt = b.knn(r, k, want_d2=True)                   # searched
        b._grids.pop("knn_last", None)
        assert np.array_equal(cnt.cpu().numpy(), rcnt.cpu().numpy())
        got, ref, cn = idx.cpu().numpy(), ridx.cpu().numpy(), rcnt.cpu().numpy()
        mask = np.arange(k)[None, :] < cn[:, None]
        assert np.array_equal(got[:, :k][mask], ref[mask])
    # and through the reference-shaped entry point: hint -> one search, identical normals and FPFH
    from instance_based_loc_b200.utils.fpfh_register import downsample_and_compute_fpfh
    from instance_based_loc_b200.o3d import KDTreeSearchParamHybrid
    from instance_based_loc_b200.o3d import registration as reg
    d1, f1 = downsample_and_compute_fpfh(a, 0.05)
    d2 = PointCloud(p)
    d2.estimate_normals(KDTreeSearchParamHybrid(radius=0.1, max_nn=30))
    d2._grids.pop("knn_last", None)
    f2 = reg.compute_fpfh_feature(d2, KDTreeSearchParamHybrid(radius=0.25, max_nn=100))
    assert np.array_equal(np.asarray(d1.normals), np.asarray(d2.normals))
    assert np.array_equal(f1.data, f2.data)


def test_remove_radius_outlier_many_equals_per_cloud(K):
    """one segmented grid + one count + one compaction for several clouds == the per-cloud filter (A.2), including clouds
    that touch or overlap each other, an empty one and a cloud without colours in the batch."""
    from instance_based_loc_b200.o3d.pointcloud import PointCloud, remove_radius_outlier_many
    rng = np.random.default_rng(4)
    clouds = []
    for k, n in enumerate((3000, 1200, 0, 5000, 40)):
        p, c = surface_cloud(max(n, 1), 10 + k, scale=0.4)
        p = p[:n] + np.array([0.25 * k, 0.0, 0.0])               # neighbours closer than the radius: clouds interpenetrate
        p = np.concatenate([p, rng.uniform(-1, 1, (n // 20, 3))]) if n else p
        pc = PointCloud(p)
        if n:
            pc.colors = np.concatenate([c[:n], rng.random((n // 20, 3))])
        clouds.append(pc)
    for nb in (8, 12):
        got = remove_radius_outlier_many(clouds, nb, 0.05)
        for c, g in zip(clouds, got):
            ref = c.remove_radius_outlier(nb, 0.05)[0] if c.n() else PointCloud()
            assert g.n() == ref.n()
            if ref.n():
                assert np.array_equal(np.asarray(g.points), np.asarray(ref.points))
                assert np.array_equal(np.asarray(g.colors), np.asarray(ref.colors))
    clouds[1] = PointCloud(np.asarray(clouds[1].points))              # no colours: the mixed case
    got = remove_radius_outlier_many(clouds, 8, 0.05)
    assert got[1].n() == clouds[1].remove_radius_outlier(8, 0.05)[0].n() and not got[1].has_colors() and got[0].has_colors()
