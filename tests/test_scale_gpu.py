"""GPU tests at BASELINE.json's FULL sizes (configs[2] "50k-point clouds, 1e5 hypotheses", configs[4] "1e4-1e6 hypotheses,
10k-200k points"), where the CPU oracle would take minutes: the results are checked through size-independent properties
and against exact brute force on random SUBSETS of the queries (torch f64 element-wise ops, one rounding per operation,
the same (dx*dx + dy*dy) + dz*dz order as the kernels).
"""
import numpy as np
import pytest
import torch

from ibl_testutil import dev, random_rigid, rot_angle_deg, surface_cloud

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K(cuda):
    from instance_based_loc_b200 import kernels
    return kernels


def h(t):
    return t.cpu().numpy()


def brute_d2(P, q):
    """[len(q), len(P)] squared distances in nanoflann's order, every operation rounded once."""
    dx = q[:, None, 0] - P[None, :, 0]
    dy = q[:, None, 1] - P[None, :, 1]
    dz = q[:, None, 2] - P[None, :, 2]
    return (dx * dx + dy * dy) + dz * dz


@pytest.mark.parametrize("n,radius,k", [(200_000, 0.25, 100), (200_000, 0.1, 30), (50_000, 0.02, 30)])
def test_knn_full_size(K, n, radius, k):
    p, _ = surface_cloud(n, 101, scale=1.5)
    P = dev(p)
    idx, d2, cnt = K.knn_hybrid(K.Grid(P, K.knn_cell_for(P, k, radius)), radius, k)
    valid = torch.arange(k, device=P.device)[None, :] < cnt[:, None]
    assert bool((idx[valid] >= 0).all()) and bool((idx[~valid] == -1).all())
    assert bool((d2[valid] < radius * radius).all()) and bool(torch.isinf(d2[~valid]).all())
    assert bool((idx[:, 0] == torch.arange(n, device=P.device)).all()) and bool((d2[:, 0] == 0).all())   # self first
    pad = torch.where(valid, d2, torch.full_like(d2, float("inf")))
    assert bool((pad[:, 1:] >= pad[:, :-1]).all())                                                      # ascending d2
    tie = (pad[:, 1:] == pad[:, :-1]) & valid[:, 1:]
    assert bool((idx[:, 1:][tie] > idx[:, :-1][tie]).all())                                             # ties by index
    # exact brute force on a random subset of the queries
    sel = torch.from_numpy(np.random.default_rng(1).choice(n, 300, replace=False)).to(P.device)
    D = brute_d2(P, P[sel])
    key = torch.where(D < radius * radius, D, torch.full_like(D, float("inf")))
    order = torch.argsort(key, dim=1, stable=True)[:, :k]                                               # (d2, index)
    want_d2 = torch.gather(key, 1, order)
    want_idx = torch.where(torch.isinf(want_d2), torch.full_like(order, -1), order)
    assert torch.equal(idx[sel].long(), want_idx)
    assert torch.equal(d2[sel], want_d2)
    assert torch.equal(cnt[sel].long(), (~torch.isinf(want_d2)).sum(1))


def test_radius_outlier_full_size(K):
    p, _ = surface_cloud(150_000, 102, scale=1.2)
    p = np.concatenate([p, np.random.default_rng(2).uniform(-2, 2, (3000, 3))])       # sparse clutter: the outliers
    P = dev(p)
    g_count, g_mask = K.Grid(P, 0.05), K.Grid(P, 0.05 * 1.001)
    cnt = K.radius_count(g_count, 0.05)
    for nb in (8, 12):
        keep = K.radius_outlier_mask(g_mask, 0.05, nb)                                # early-exit path, 3^3 cells
        assert torch.equal(keep.bool(), cnt > nb)
    assert 0 < int((cnt <= 8).sum()) < 4000
    sel = torch.from_numpy(np.random.default_rng(3).choice(len(p), 300, replace=False)).to(P.device)
    assert torch.equal(cnt[sel].long(), (brute_d2(P, P[sel]) < 0.05 * 0.05).sum(1))


def test_voxel_full_size(K):
    p, c = surface_cloud(200_000, 103, scale=1.5)
    P, C = dev(p), dev(c)
    op, oc_, _, vid, keys = K.voxel_downsample(P, C, None, 0.01, want_ids=True, want_keys=True)
    want_keys = torch.floor(P / 0.01).to(torch.int64)
    assert torch.equal(keys, want_keys)                                               # voxel indices bit-exact
    uniq, inv = torch.unique(want_keys, dim=0, return_inverse=True)
    assert op.shape[0] == uniq.shape[0]
    # first-occurrence order: voxel ids are assigned in order of first appearance
    first = torch.full((op.shape[0],), len(p), dtype=torch.int64, device=P.device)
    first.scatter_reduce_(0, vid.long(), torch.arange(len(p), device=P.device), reduce="amin")
    assert bool((first[1:] > first[:-1]).all())
    # every voxel's output point is the mean of its members (f64 sums in input order; tolerance covers torch's order)
    sums = torch.zeros_like(op).index_add_(0, vid.long(), P)
    counts = torch.zeros(op.shape[0], dtype=torch.float64, device=P.device).index_add_(0, vid.long(), torch.ones(len(p), dtype=torch.float64, device=P.device))
    assert torch.allclose(op, sums / counts[:, None], rtol=0, atol=1e-12)
    # members really are in their voxel
    assert torch.equal(torch.floor(P / 0.01).to(torch.int64), want_keys[first[vid.long()]])


@pytest.mark.parametrize("n_pts,H", [(50_000, 100_000), (10_000, 1_000_000)])
def test_ransac_full_size(K, n_pts, H):
    """configs[2] / configs[4]: fixed hypothesis count (no early exit), 30 % inlier correspondences."""
    tgt, _ = surface_cloud(n_pts, 104, scale=1.5)
    Tgt = random_rigid(105, 0.8, 0.4)
    rng = np.random.default_rng(106)
    sel = np.sort(rng.choice(n_pts, int(0.7 * n_pts), replace=False))
    src = (np.linalg.inv(Tgt) @ np.c_[tgt[sel], np.ones(len(sel))].T).T[:, :3] + rng.normal(0, 0.001, (len(sel), 3))
    C = 3000
    ci = rng.choice(len(sel), C, replace=False)
    cor = np.stack([ci, sel[ci]], axis=1).astype(np.int32)
    bad = rng.random(C) < 0.7
    cor[bad, 1] = rng.integers(0, n_pts, bad.sum())
    S, T_, Cd = dev(src), dev(tgt), dev(cor, torch.int32)
    g = K.Grid(T_, 0.075, nnfield=0.075, nnfield_voxel=0.075 / 8)
    r = K.ransac_run(g, S, T_, Cd, H, 0.99, 0.075, edge_thr=0.9, seed=7, early_exit=False, block=1 << 14)
    assert r["drawn"] == H and 0 < r["validated"] <= H
    assert rot_angle_deg(r["T"][:3, :3], Tgt[:3, :3]) < 1.0 and np.linalg.norm(r["T"][:3, 3] - Tgt[:3, 3]) < 0.02
    # the winner's score is what evaluate_registration computes for its transformation (count pass == exact pass == nn_score)
    n_in, err2, _, _ = K.nn_score(g, S, 0.075, T=r["T"])
    assert n_in == r["n_inliers"] and abs(err2 - r["err2"]) <= 1e-12 * max(1.0, err2)
    assert r["fitness"] == n_in / len(src) and r["fitness"] > 0.95
    # the exact per-hypothesis score of the winner, recomputed from its id through the public scoring entry point
    inl, err, cin, Ts = K.ransac_score(g, S, T_, Cd, torch.tensor([r["best_hyp"]], device=S.device), 0.075, seed=7)
    assert int(inl[0]) == r["n_inliers"] and np.array_equal(h(Ts).reshape(4, 4), r["T"])
    # sharding by hypothesis block: the best of two half-streams under the ordered rule is the same winner
    from instance_based_loc_b200 import dist
    halves = []
    for rank in range(2):
        h0, h1 = dist.hypothesis_block(H, rank, 2)
        halves.append(K.ransac_run(g, S, T_, Cd, h1, 0.99, 0.075, edge_thr=0.9, seed=7, early_exit=False, block=1 << 14, h0=h0))
        assert halves[-1]["drawn"] == h1 - h0
    assert halves[0]["validated"] + halves[1]["validated"] == r["validated"]
    best = min(halves, key=lambda x: (-x["fitness"], x["inlier_rmse"], x["best_hyp"]))
    assert best["best_hyp"] == r["best_hyp"] and np.array_equal(best["T"], r["T"])


def test_feature_nn_full_size(K):
    """50k x 50k FPFH-like rows with planted duplicates: exact argmin (lowest index on ties) on a random subset."""
    rng = np.random.default_rng(107)

    def rows(n):
        f = rng.gamma(0.4, 1.0, (n, 3, 11))
        return np.ascontiguousarray((200.0 * f / f.sum(-1, keepdims=True)).reshape(n, 33))

    fa, fb = rows(50_000), rows(50_000)
    fb[rng.integers(0, 50_000, 4000)] = fb[rng.integers(0, 50_000, 4000)]             # duplicate targets
    fa[rng.integers(0, 50_000, 2000)] = fb[rng.integers(0, 50_000, 2000)]             # queries that hit a target exactly
    A, B = dev(fa), dev(fb)
    got = K.feature_nn(A, B)
    sel = torch.from_numpy(rng.choice(50_000, 1500, replace=False)).to(A.device)
    want = torch.empty(len(sel), dtype=torch.int64, device=A.device)
    for i in range(0, len(sel), 250):
        a = A[sel[i:i + 250]]
        d = torch.zeros((a.shape[0], B.shape[0]), dtype=torch.float64, device=A.device)
        for j in range(33):                                                          # the oracle's order: j ascending
            t = a[:, j, None] - B[None, :, j]
            d = d + t * t
        m = d.min(1).values
        want[i:i + 250] = (d == m[:, None]).to(torch.uint8).argmax(1)                 # lowest index among exact minima
    assert torch.equal(got[sel].long(), want)


def test_icp_evaluate_consistency_full_size(K):
    """coloured ICP on a 60k-point pair: the reported fitness / rmse are evaluate_registration's for the returned T (up to
    rounding), the pose is recovered, and a second call from the converged pose stops at once (idempotence)."""
    from instance_based_loc_b200 import o3d
    from instance_based_loc_b200.o3d import registration as reg
    tgt, tcol = surface_cloud(60_000, 108, scale=1.5)
    Tgt = random_rigid(109, 0.03, 0.01)
    sel = np.sort(np.random.default_rng(110).choice(60_000, 45_000, replace=False))
    src = (np.linalg.inv(Tgt) @ np.c_[tgt[sel], np.ones(len(sel))].T).T[:, :3]
    src = src + np.random.default_rng(111).normal(0, 0.0005, src.shape)
    t = o3d.geometry.PointCloud(); t.points = o3d.utility.Vector3dVector(tgt); t.colors = o3d.utility.Vector3dVector(tcol)
    s = o3d.geometry.PointCloud(); s.points = o3d.utility.Vector3dVector(src); s.colors = o3d.utility.Vector3dVector(tcol[sel])
    t.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=0.1, max_nn=30))
    r = reg.registration_colored_icp(s, t, 0.03, np.eye(4))
    ev = reg.evaluate_registration(s, t, 0.03, r.transformation)
    # ICP scores the cloud it moved step by step, evaluate_registration moves the original by the composed matrix: the
    # two differ by rounding only (as in Open3D)
    assert abs(ev.fitness - r.fitness) <= 2 / len(src) and abs(ev.inlier_rmse - r.inlier_rmse) < 1e-9
    assert rot_angle_deg(r.transformation[:3, :3], Tgt[:3, :3]) < 0.1 and np.linalg.norm(r.transformation[:3, 3] - Tgt[:3, 3]) < 1e-3
    r2 = reg.registration_colored_icp(s, t, 0.03, r.transformation)
    assert r2.info["iterations"] <= 2 and abs(r2.fitness - r.fitness) < 1e-4
