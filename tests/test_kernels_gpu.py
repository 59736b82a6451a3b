"""GPU parity tests (run with -m gpu on the B200 box): every libibl.so entry point, called through the C ABI
(ctypes, instance_based_loc_b200.kernels), against the CPU oracle on the same seeded inputs and against the
golden vectors generated from the reference's own numpy code (tests/golden/make_golden.py).

Bars (BASELINE.json north_star): voxel indices, keep-masks, neighbour index sets, correspondences, inlier counts,
winning hypothesis -- bit-exact; FPFH <= 1e-4 relative; poses <= 0.1 deg / 1 mm.
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


@pytest.fixture(scope="module")
def oc():
    from oracle import o3d_core
    o3d_core.lib()
    return o3d_core


def h(t):
    return t.cpu().numpy()


# ------------------------------------------------------------------------------------------- a1/a2
def test_backproject_golden_f32(K, golden):
    g = golden
    masks = torch.from_numpy(g["bp_masks"][:, 0]).cuda()
    pts, cols, offs = K.backproject_masks(torch.from_numpy(g["bp_depth"]).cuda(), torch.from_numpy(g["bp_rgb"]).cuda(),
                                          masks, float(g["bp_fx"]), float(g["bp_fy"]))
    assert offs == list(g["bp_offsets"])
    assert np.array_equal(h(pts), g["bp_points"])          # bit-exact
    assert np.array_equal(h(cols), g["bp_colors"])


def test_backproject_golden_f64(K, golden):
    g = golden
    H, W = g["bp64_depth"].shape
    ones = torch.ones((1, H, W), dtype=torch.uint8, device="cuda")
    pts, cols, offs = K.backproject_masks(torch.from_numpy(g["bp64_depth"]).cuda(), torch.from_numpy(g["bp_rgb"]).cuda(),
                                          ones, float(g["bp_fx"]), float(g["bp_fy"]))
    assert np.array_equal(h(pts), g["bp64_points"])
    assert np.array_equal(h(cols), g["bp64_colors"])


@pytest.mark.parametrize("H,W", [(480, 640), (37, 53), (1, 7)])
def test_backproject_vs_port(K, H, W):
    from oracle import ref_port
    rng = np.random.default_rng(H * W)
    depth = rng.uniform(0.3, 5.0, (H, W)).astype(np.float32)
    depth[rng.random((H, W)) < 0.4] = 0
    rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    masks = rng.random((4, H, W)) < 0.3
    masks[3] = False                                        # an empty mask (ragged output)
    pts, cols, offs = K.backproject_masks(torch.from_numpy(depth).cuda(), torch.from_numpy(rgb).cuda(),
                                          torch.from_numpy(masks).cuda(), 525.0, 525.0)
    ref = ref_port.backproject_masks(depth, rgb, masks, 525.0, 525.0)
    assert offs == list(np.cumsum([0] + [len(p) for p, _ in ref]))
    assert offs[4] == offs[3]
    assert np.array_equal(h(pts), np.concatenate([p for p, _ in ref]))
    assert np.array_equal(h(cols), np.concatenate([c for _, c in ref]))


def test_backproject_no_masks(K):
    d = torch.zeros((4, 4), dtype=torch.float32, device="cuda")
    pts, cols, offs = K.backproject_masks(d, None, torch.zeros((0, 4, 4), dtype=torch.uint8, device="cuda"), 1.0, 1.0)
    assert offs == [0] and pts.shape[0] == 0


# ------------------------------------------------------------------------------------------- a3
def test_voxel_golden(K, golden):
    g = golden
    p, c, n, vid, keys = K.voxel_downsample(dev(g["vox_pts"]), dev(g["vox_cols"]), None, float(g["vox_size"]),
                                            want_ids=True, want_keys=True)
    assert np.array_equal(h(keys), g["vox_keys"])           # voxel indices bit-exact
    assert p.shape[0] == g["vox_out_pts"].shape[0]
    assert np.array_equal(h(p), g["vox_out_pts"])
    assert np.array_equal(h(c), g["vox_out_cols"])


@pytest.mark.parametrize("n,voxel", [(1, 0.01), (5000, 0.005), (20000, 0.05)])
def test_voxel_vs_port(K, n, voxel):
    from oracle import ref_port
    p, c = surface_cloud(n, n)
    nr = np.random.default_rng(n).normal(size=(n, 3))
    op, ocol, onrm, vid, keys = K.voxel_downsample(dev(p), dev(c), dev(nr), voxel, want_ids=True, want_keys=True)
    rp, rc, rn, rvid = ref_port.voxel_down_sample(p, c, voxel, nr)
    assert np.array_equal(h(keys), ref_port.voxel_indices(p, voxel))
    assert np.array_equal(h(vid), rvid)
    assert np.array_equal(h(op), rp) and np.array_equal(h(ocol), rc)
    np.testing.assert_allclose(h(onrm), rn, rtol=0, atol=1e-15)


# ------------------------------------------------------------------------------------------- a4
def test_transform_select_mean(K, oc):
    p, c = surface_cloud(7001, 3)
    T = random_rigid(5)
    assert np.array_equal(h(K.transform_points(T, dev(p))), oc.transform_points(T, p))
    keep = np.random.default_rng(1).random(len(p)) < 0.37
    a, b, _, oi = K.select_rows(torch.from_numpy(keep).cuda(), dev(p), dev(c), want_index=True)
    assert np.array_equal(h(a), p[keep]) and np.array_equal(h(b), c[keep])
    assert np.array_equal(h(oi), np.nonzero(keep)[0])
    assert np.array_equal(K.mean3(dev(p)), p.mean(0))          # numpy axis-0 mean is a serial row-by-row sum: bit-exact
    e, _, _, _ = K.select_rows(torch.zeros(len(p), dtype=torch.uint8, device="cuda"), dev(p))
    assert e.shape[0] == 0


# ------------------------------------------------------------------------------------------- a5 / A.1
@pytest.mark.parametrize("src", ["fixture", "synthetic"])
def test_radius_count_and_outlier(K, oc, golden, src):
    p = golden["fix_points"] if src == "fixture" else surface_cloud(12000, 11)[0]
    g = K.Grid(dev(p), 0.05)
    cnt = h(K.radius_count(g, 0.05))
    ref = oc.radius_count(p, 0.05)
    assert np.array_equal(cnt, ref)                          # bit-exact counts
    for nb in (8, 12):
        assert np.array_equal(h(K.radius_outlier_mask(g, 0.05, nb)).astype(bool), ref > nb)
    g2 = K.Grid(dev(p), 0.021)                               # a cell that does not divide the radius
    assert np.array_equal(h(K.radius_count(g2, 0.05)), ref)


@pytest.mark.parametrize("radius,k", [(0.1, 30), (0.25, 100), (0.02, 30), (float("inf"), 10)])
def test_knn_hybrid(K, oc, golden, radius, k):
    p = np.concatenate([golden["fix_points"][:2500], surface_cloud(2500, 5)[0] + golden["fix_points"].mean(0)])
    P = dev(p)
    cell = K.knn_cell_for(P, k, radius if np.isfinite(radius) else 1.0)
    idx, d2, cnt = K.knn_hybrid(K.Grid(P, cell), radius, k)
    ridx, rd2, rcnt = oc.knn_hybrid(p, p, radius, k)
    assert np.array_equal(h(cnt), rcnt)
    assert np.array_equal(h(idx), ridx)                      # same neighbours in the same (d2, index) order
    assert np.array_equal(h(d2), rd2)                        # bit-exact squared distances (inf padding)
    # external queries (not grid points), some far outside the box
    q = p[::7] + np.random.default_rng(0).normal(0, 0.05, p[::7].shape)
    q[:5] += 10.0
    idx, d2, cnt = K.knn_hybrid(K.Grid(P, cell), radius, k, queries=dev(q))
    ridx, rd2, rcnt = oc.knn_hybrid(p, q, radius, k)
    assert np.array_equal(h(cnt), rcnt) and np.array_equal(h(idx), ridx) and np.array_equal(h(d2), rd2)


def test_knn_tiny_clouds(K, oc):
    for n in (1, 2, 5):
        p = np.random.default_rng(n).normal(size=(n, 3))
        idx, d2, cnt = K.knn_hybrid(K.Grid(dev(p), 0.5), 10.0, 30)
        ridx, rd2, rcnt = oc.knn_hybrid(p, p, 10.0, 30)
        assert np.array_equal(h(idx), ridx) and np.array_equal(h(cnt), rcnt)


@pytest.mark.parametrize("max_d,ring,field", [(0.075, 8, 0), (0.02, 8, 0), (0.075, 0, 0), (0.075, 0, 0.075), (0.02, 0, 0.02),
                                                (0.02, 0, 0.075), (0.075, 0, 0.02)])
def test_nn_score(K, oc, max_d, ring, field):
    """ring search and the exact candidate field (field = the max_d the field was built for; a field built for a
    SMALLER max_d than the query must be ignored, one built for a larger one must give the same answer)."""
    tgt, _ = surface_cloud(9000, 21)
    src = (random_rigid(2, 0.05, 0.03) @ np.c_[surface_cloud(4000, 22)[0], np.ones(4000)].T).T[:, :3]
    src = np.concatenate([src, src[:300] + 0.06, src[:50] + 3.0, tgt[:100]])     # far, very far and exact-hit queries
    g = K.Grid(dev(tgt), max(max_d / 4, 1e-4), ring=ring, nnfield=field, nnfield_voxel=max(field / 4, 0.01))
    assert (g.nnfield_bytes > 0) == (field > 0)
    T = random_rigid(3, 0.02, 0.01)
    n_in, err2, nn, d2 = K.nn_score(g, dev(src), max_d, T=T, want_idx=True, want_d2=True)
    pts = oc.transform_points(T, src)
    ridx, rd2 = oc.nn1(tgt, pts, max_d)
    assert np.array_equal(h(nn), ridx)                       # correspondence set bit-exact
    assert n_in == int((ridx >= 0).sum())
    assert np.array_equal(h(d2)[ridx >= 0], rd2[ridx >= 0])
    assert abs(err2 - rd2[ridx >= 0].sum()) <= 1e-12 * max(1.0, err2)


# ------------------------------------------------------------------------------------------- a8
@pytest.fixture(scope="module")
def feat_case(K, oc, golden):
    """normals + SPFH + FPFH of one cloud on both sides (shared by the feature tests)."""
    p = np.concatenate([golden["fix_points"][:3000], surface_cloud(3000, 9)[0] + golden["fix_points"].mean(0)])
    P = dev(p)
    idx30, _, cnt30 = K.knn_hybrid(K.Grid(P, K.knn_cell_for(P, 30, 0.1)), 0.1, 30, want_d2=False)
    nrm = K.normals(P, idx30, cnt30)
    rn = oc.estimate_normals(p, 0.1, 30)
    idx, d2, cnt = K.knn_hybrid(K.Grid(P, K.knn_cell_for(P, 100, 0.25)), 0.25, 100)
    return dict(p=p, P=P, nrm=nrm, rn=rn, idx=idx, d2=d2, cnt=cnt)


def test_normals(feat_case):
    n, rn = h(feat_case["nrm"]), feat_case["rn"]
    # unoriented normals: the SIGN must match too (FPFH bins depend on it)
    np.testing.assert_allclose(n, rn, rtol=0, atol=1e-9)
    np.testing.assert_allclose(np.linalg.norm(n, axis=1), 1.0, atol=1e-12)


def test_normals_degenerate(K, oc):
    # collinear / coplanar / duplicate points exercise the FastEigen3x3 special branches
    p = np.zeros((40, 3))
    p[:10, 0] = np.arange(10) * 0.01                          # a line
    p[10:30, :2] = np.random.default_rng(0).uniform(0, 0.05, (20, 2)) + 1.0   # a plane patch z=0
    p[30:] = [5.0, 5.0, 5.0]                                  # duplicates -> zero covariance
    P = dev(p)
    idx, _, cnt = K.knn_hybrid(K.Grid(P, 0.05), 0.1, 30, want_d2=False)
    np.testing.assert_allclose(h(K.normals(P, idx, cnt)), oc.estimate_normals(p, 0.1, 30), rtol=0, atol=1e-9)


def test_spfh_fpfh(K, oc, feat_case):
    f = feat_case
    rn = f["rn"]
    # feed the ORACLE's normals to both sides so the comparison isolates SPFH/FPFH
    sp = K.spfh(f["P"], dev(rn), f["idx"], f["cnt"])
    fp = K.fpfh(f["idx"], f["d2"], f["cnt"], sp)
    rf, rs = oc.compute_fpfh(f["p"], rn, 0.25, 100, return_spfh=True)
    np.testing.assert_allclose(h(sp).T, rs, rtol=1e-12, atol=1e-9)
    got, ref = h(fp).T, rf
    scale = np.maximum(np.abs(ref), 1e-3 * np.abs(ref).max())
    assert np.max(np.abs(got - ref) / scale) <= 1e-4         # north_star: FPFH within 1e-4 relative
    # end to end with device normals as well
    sp2 = K.spfh(f["P"], f["nrm"], f["idx"], f["cnt"])
    fp2 = h(K.fpfh(f["idx"], f["d2"], f["cnt"], sp2)).T
    assert np.max(np.abs(fp2 - ref) / scale) <= 1e-4


def test_feature_nn_and_mutual(K, oc):
    rng = np.random.default_rng(4)
    fa = rng.uniform(0, 200, (33, 900)) * (rng.random((33, 900)) < 0.4)
    fb = np.concatenate([fa[:, ::3] + rng.normal(0, 2.0, fa[:, ::3].shape), rng.uniform(0, 200, (33, 500))], axis=1)
    fb[:, 7] = fb[:, 3]                                       # exact duplicate -> lowest index must win
    A, B = dev(np.ascontiguousarray(fa.T)), dev(np.ascontiguousarray(fb.T))
    ij, ji = K.feature_nn(A, B), K.feature_nn(B, A)
    assert np.array_equal(h(ij), oc.feature_nn(fa, fb))
    assert np.array_equal(h(ji), oc.feature_nn(fb, fa))
    cor = h(K.mutual_corres(ij, ji, 3))
    assert np.array_equal(cor, oc.match_features(fa, fb, True, 3))
    # survivor-list overflow of the fp32 filter: 12 identical targets -> exact slice rescan, lowest index wins
    fc = fb.copy()
    fc[:, 100:112] = fa[:, [5]]
    fc[:, 40] = fa[:, 5] + 1e-9                              # closer than fp32 can resolve, but not the nearest
    Cc = dev(np.ascontiguousarray(fc.T))
    got = h(K.feature_nn(A, Cc))
    assert np.array_equal(got, oc.feature_nn(fa, fc)) and got[5] == 100
    # large, FPFH-like (non-negative, three blocks summing to ~200) problem that spans several target slices
    big_a = rng.gamma(0.3, 1.0, (33, 5000)); big_b = rng.gamma(0.3, 1.0, (33, 9000))
    for blk in range(3):
        big_a[11 * blk:11 * blk + 11] *= 200.0 / big_a[11 * blk:11 * blk + 11].sum(0)
        big_b[11 * blk:11 * blk + 11] *= 200.0 / big_b[11 * blk:11 * blk + 11].sum(0)
    got = h(K.feature_nn(dev(np.ascontiguousarray(big_a.T)), dev(np.ascontiguousarray(big_b.T))))
    assert np.array_equal(got, oc.feature_nn(big_a, big_b))
    # fallback: fewer than 9 mutual matches -> all of corres_ij
    ij2 = torch.arange(20, dtype=torch.int32, device="cuda") % 4
    ji2 = torch.zeros(4, dtype=torch.int32, device="cuda")
    cor2 = h(K.mutual_corres(ij2, ji2, 3))
    assert cor2.shape == (20, 2) and np.array_equal(cor2[:, 1], h(ij2))


def _fpfh_like(rng, n):
    f = rng.gamma(0.4, 1.0, (n, 3, 11))
    return np.ascontiguousarray((200.0 * f / f.sum(-1, keepdims=True)).reshape(n, 33))


def test_feature_scores_tensor_core_band(K):
    """The tcgen05 tf32 scores s' = a32.b32 - |b32|^2/2 (csrc/match_tc.cu) stay inside the error band the matcher assumes:
    |(an - 2 s') - d_exact| <= 1.33e-5 an + 2.54e-5 bn  (TC_ERR_A / TC_ERR_B carry a further 1.15x margin)."""
    rng = np.random.default_rng(41)
    fa, fb = _fpfh_like(rng, 333), _fpfh_like(rng, 517)
    fb[100] = fa[7]
    S, mu, sc = K.feature_scores_debug(dev(fa), dev(fb))
    S, mu = h(S).astype(np.float64), h(mu)
    assert S.shape == (333, 517) and np.isfinite(S).all()
    # the filter works on rows centred on mu and scaled by the power of two sc (|value| <= 1024, for fp16), rounded to
    # fp32; neither changes the order of the distances
    assert sc == 2.0 ** round(np.log2(sc)) and 512 < np.abs((np.concatenate([fa, fb]) - mu) * sc).max() <= 1024
    a32 = ((fa - mu) * sc).astype(np.float32).astype(np.float64)
    an, bn = (a32 * a32).sum(1), (((fb - mu) * sc).astype(np.float32).astype(np.float64) ** 2).sum(1)
    assert an.mean() < 0.9 * sc * sc * (fa * fa).sum(1).mean()   # centring shrinks the norms the band scales with
    d_exact = sc * sc * ((fa[:, None, :] - fb[None, :, :]) ** 2).sum(-1)
    err = np.abs((an[:, None] - 2.0 * S) - d_exact)
    bound = 1.33e-5 * an[:, None] + 2.54e-5 * bn[None, :] + 3.5e-7 * (np.sqrt(an)[:, None] + np.sqrt(bn)[None, :]) + 6.2e-5
    assert (err <= bound).all()
    assert err.max() / (an.max() + bn.max()) < 5e-6          # measured ~6e-7 on B200: the band is conservative


def test_feature_nn_any_units(K, oc):
    """The fp16 filter rescales by a power of two, so features in any units match the oracle; at x1e25 the fp32 second
    filter overflows and must fall back to evaluating every candidate exactly, at x1e-25 its squares underflow."""
    rng = np.random.default_rng(47)
    fa, fb = _fpfh_like(rng, 300), _fpfh_like(rng, 420)
    fb[17] = fa[5]
    for unit in (1e6, 1e-6, 1e25, 1e-25):
        got = h(K.feature_nn(dev(fa * unit), dev(fb * unit)))
        assert np.array_equal(got, oc.feature_nn(np.ascontiguousarray((fa * unit).T), np.ascontiguousarray((fb * unit).T)))
        assert got[5] == 17


@pytest.mark.parametrize("na,nb", [(1, 1), (127, 129), (128, 128), (129, 127), (3, 1000), (1000, 3)])
def test_feature_nn_tile_edges(K, oc, na, nb):
    """Query / target counts around the 128-row tile size of the tensor-core filter (padding rows must never win)."""
    rng = np.random.default_rng(1000 * na + nb)
    fa, fb = _fpfh_like(rng, na), _fpfh_like(rng, nb)
    got = h(K.feature_nn(dev(fa), dev(fb)))
    assert np.array_equal(got, oc.feature_nn(np.ascontiguousarray(fa.T), np.ascontiguousarray(fb.T)))


def test_feature_nn_planar_degenerate(K, oc):
    """FPFH of planar patches: hundreds of IDENTICAL target rows (all inside the band) -> lowest index, like the oracle."""
    rng = np.random.default_rng(43)
    fa, fb = _fpfh_like(rng, 400), _fpfh_like(rng, 900)
    plane = _fpfh_like(rng, 1)[0]
    fb[200:700] = plane
    fb[650] = plane + 1e-7                                    # differs from its 499 neighbours far below fp32 resolution
    fa[::4] = plane
    fa[1::4] = plane * (1 + 1e-6)                             # exact nearest is row 650, with 499 rows inside the band
    K.feature_nn_stats(True)
    got = h(K.feature_nn(dev(fa), dev(fb)))
    assert np.array_equal(got, oc.feature_nn(np.ascontiguousarray(fa.T), np.ascontiguousarray(fb.T)))
    assert (got[::4] == 200).all()
    cand, _ = K.feature_nn_stats(True)
    # the same problem with the candidate pool capped at 3 chunks: most band candidates take the on-the-spot path
    K.feature_nn_set_pool_chunks(3)
    try:
        got2 = h(K.feature_nn(dev(fa), dev(fb)))
        cand2, ev2 = K.feature_nn_stats(True)
    finally:
        K.feature_nn_set_pool_chunks(0)
    assert np.array_equal(got2, got)
    assert cand2 > 200 * 499 and ev2 >= cand2 - 3 * 64 - 400   # everything beyond 3 chunks (and the singles) evaluated exactly


# ------------------------------------------------------------------------------------------- a9
@pytest.fixture(scope="module")
def reg_case():
    tgt, tcol = surface_cloud(6000, 31)
    Tgt = random_rigid(32, 0.6, 0.3)
    sel = np.sort(np.random.default_rng(33).choice(6000, 3500, replace=False))
    src = (np.linalg.inv(Tgt) @ np.c_[tgt[sel], np.ones(len(sel))].T).T[:, :3]
    src = src + np.random.default_rng(34).normal(0, 0.001, src.shape)
    rng = np.random.default_rng(35)
    C = 1200
    ci = rng.choice(len(sel), C, replace=False)
    cor = np.stack([ci, sel[ci]], axis=1).astype(np.int32)
    bad = rng.random(C) < 0.7                                 # 70 % outlier correspondences
    cor[bad, 1] = rng.integers(0, 6000, bad.sum())
    return dict(src=src, scol=tcol[sel], tgt=tgt, tcol=tcol, cor=cor, T=Tgt)


def test_sampler_matches_oracle(K, oc, reg_case):
    from oracle import sampler
    c = reg_case
    H, C = 5000, len(c["cor"])
    s = sampler.samples(12345, 100, H, C)
    assert s.min() >= 0 and s.max() < C
    ok_a, Ta = K.ransac_check(dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32), H, 0.9, 0.075,
                              samples=None, seed=12345, h0=100, want_T=True)
    ok_b, Tb = K.ransac_check(dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32), H, 0.9, 0.075,
                              samples=dev(s, torch.int32), want_T=True)
    assert np.array_equal(h(ok_a), h(ok_b)) and np.array_equal(h(Ta), h(Tb))


def test_ransac_check(K, oc, reg_case):
    c = reg_case
    rng = np.random.default_rng(1)
    H = 20000
    s = rng.integers(0, len(c["cor"]), (H, 3), dtype=np.int32)
    s[:50, 1] = s[:50, 0]                                     # degenerate triplets (repeated correspondence)
    ok, Ts = K.ransac_check(dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32), H, 0.9, 0.075,
                            samples=dev(s, torch.int32), want_T=True)
    rok, rT = oc.ransac_check(c["src"], c["tgt"], c["cor"], s, 0.9, 0.075, want_T=True)
    assert np.array_equal(h(ok).astype(bool), rok)            # checker decisions bit-exact
    assert rok.sum() > 10
    got = h(Ts).reshape(H, 4, 4)[rok]
    assert np.array_equal(got, rT[rok])                       # same + - * / sqrt sequence, FMA off: bit-exact
    ok2, _ = K.ransac_check(dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32), H, 0.9, 0.075,
                            samples=dev(s, torch.int32), use_checkers=False)
    assert h(ok2).all()


def test_ransac_score(K, oc, reg_case):
    c = reg_case
    rng = np.random.default_rng(2)
    s = rng.integers(0, len(c["cor"]), (4000, 3), dtype=np.int32)
    rok = oc.ransac_check(c["src"], c["tgt"], c["cor"], s, 0.9, 0.075)
    ids = np.nonzero(rok)[0][:40].astype(np.int64)
    g = K.Grid(dev(c["tgt"]), 0.075, nnfield=0.075, nnfield_voxel=0.01875)
    inl, err, cin, Ts = K.ransac_score(g, dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32),
                                       torch.from_numpy(ids).cuda(), 0.075, samples=dev(s, torch.int32))
    from scipy.spatial import cKDTree
    tree = cKDTree(c["tgt"])
    import ctypes
    for k, hid in enumerate(ids):
        T = oc.hyp_transform(c["src"], c["tgt"], c["cor"], s[hid])
        res = oc.score(oc.transform_points(T, c["src"]), c["tgt"], tree, 0.075, T)
        assert int(h(inl)[k]) == res.n_inliers                # inlier counts bit-exact
        assert abs(float(h(err)[k]) - res.err2) <= 1e-12 * max(1.0, res.err2)
        ci = oc.lib().orc_corres_inliers(oc._p(oc._f64(T)), oc._p(oc._f64(c["src"])), oc._p(oc._f64(c["tgt"])),
                                         oc._p(oc._i32(c["cor"])), len(c["cor"]), ctypes.c_double(0.075))
        assert int(h(cin)[k]) == ci


@pytest.mark.parametrize("field,voxel_div,max_d", [(0.075, 8, 0.075), (0.075, 4, 0.075), (0.075, 8, 0.05), (0.0, 0, 0.075)])
def test_ransac_count_equals_score(K, reg_case, field, voxel_div, max_d):
    """The count-only pass (first candidate within max_d; SURE voxels of the NN field when the query distance is the
    field's) must give the exact pass's inlier counts: 600 unchecked hypotheses, good and junk alike."""
    c = reg_case
    S, T_, C = dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32)
    g = (K.Grid(T_, 0.075, nnfield=field, nnfield_voxel=field / voxel_div) if field else K.Grid(T_, 0.075 / 4, ring=8))
    ids = torch.arange(600, dtype=torch.int64, device=S.device)
    inl_c, T_c = K.ransac_count(g, S, T_, C, ids, max_d, seed=99)
    inl_s, _, _, T_s = K.ransac_score(g, S, T_, C, ids, max_d, seed=99)
    assert torch.equal(inl_c, inl_s) and torch.equal(T_c, T_s)
    assert int(inl_s.max()) > 100 and int((inl_s == 0).sum()) >= 0


@pytest.mark.parametrize("early_exit,block,field", [(True, 4096, True), (True, 1000, False), (False, 4096, True)])
def test_ransac_run(K, oc, reg_case, early_exit, block, field):
    c = reg_case
    H = 30000
    s = np.random.default_rng(3).integers(0, len(c["cor"]), (H, 3), dtype=np.int32)
    g = K.Grid(dev(c["tgt"]), 0.075, nnfield=0.075) if field else K.Grid(dev(c["tgt"]), 0.075 / 4, ring=8)
    r = K.ransac_run(g, dev(c["src"]), dev(c["tgt"]), dev(c["cor"], torch.int32), H, 0.99, 0.075, edge_thr=0.9,
                     samples=dev(s, torch.int32), early_exit=early_exit, block=block)
    ref, info = oc.ransac(c["src"], c["tgt"], c["cor"], s, 0.075, 0.9, 0.99, early_exit=early_exit)
    assert r["best_hyp"] == info["best_hyp"]                  # winning hypothesis index
    assert r["n_inliers"] == ref.n_inliers
    assert r["validated"] == info["validated"] and r["drawn"] == info["drawn"]
    assert np.array_equal(r["T"], ref.transformation)
    assert abs(r["fitness"] - ref.fitness) == 0 and abs(r["inlier_rmse"] - ref.inlier_rmse) < 1e-12
    # and the pose is the right one
    assert rot_angle_deg(r["T"][:3, :3], c["T"][:3, :3]) < 2.0


def test_scratch_and_host_wait_knobs(K):
    """ibl_reserve_scratch grows the pool and keeps it; every host-wait mode gives the same results."""
    before = K.scratch_stats()["reserved"]
    K.reserve_scratch(64 << 20)
    torch.cuda.synchronize()
    assert K.scratch_stats()["reserved"] >= max(before, 64 << 20)
    p = dev(surface_cloud(4000, 77)[0])
    want = None
    try:
        for mode in ("auto", "spin", "yield", "block", "hybrid", "auto"):
            K.set_host_wait(mode)
            g = K.Grid(p, 0.03)
            idx, d2, cnt = K.knn_hybrid(g, 0.1, 30)
            got = (h(idx), h(cnt), K.mean3(p))
            if want is None:
                want = got
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    finally:
        K.set_host_wait("auto")


def test_ransac_too_few_correspondences(K, reg_case):
    c = reg_case
    g = K.Grid(dev(c["tgt"]), 0.02)
    r = K.ransac_run(g, dev(c["src"]), dev(c["tgt"]), dev(c["cor"][:2], torch.int32), 100, 0.99, 0.075)
    assert r["status"] == 1 and r["fitness"] == 0 and np.array_equal(r["T"], np.eye(4))


# ------------------------------------------------------------------------------------------- a10
def test_p2p_icp(K, oc, reg_case):
    c = reg_case
    init = c["T"] @ random_rigid(40, 0.03, 0.01)
    g = K.Grid(dev(c["tgt"]), 0.02 / 4, ring=8)
    r = K.icp(g, dev(c["src"]), dev(c["tgt"]), 0.02, init, mode=0)
    ref = oc.icp_p2p(c["src"], c["tgt"], 0.02, init)
    assert r["iterations"] == ref.iterations
    assert r["n_inliers"] == ref.n_inliers
    assert rot_angle_deg(r["T"][:3, :3], ref.transformation[:3, :3]) < 0.1
    assert np.linalg.norm(r["T"][:3, 3] - ref.transformation[:3, 3]) < 1e-3
    np.testing.assert_allclose(r["T"], ref.transformation, atol=1e-9)


def test_colored_icp(K, oc, reg_case):
    c = reg_case
    T0 = c["T"] @ random_rigid(41, 0.03, 0.01)
    tn = oc.estimate_normals(c["tgt"], 0.1, 30)
    P, N, Cc = dev(c["tgt"]), dev(tn), dev(c["tcol"])
    idx, _, cnt = K.knn_hybrid(K.Grid(P, K.knn_cell_for(P, 30, 0.15)), 0.15, 30, want_d2=False)
    grad = K.color_gradient(P, N, Cc, idx, cnt)
    rgrad = oc.color_gradient(c["tgt"], tn, c["tcol"], 0.075)
    np.testing.assert_allclose(h(grad), rgrad, rtol=1e-9, atol=1e-9)
    # one Gauss-Newton system on a fixed correspondence set
    pts = oc.transform_points(T0, c["src"])
    nn, _ = oc.nn1(c["tgt"], pts, 0.075)
    Tu, v27 = K.colored_update(dev(pts), dev(c["scol"]), dev(nn, torch.int32), P, N, Cc, dev(rgrad))
    cor = np.stack([np.nonzero(nn >= 0)[0], nn[nn >= 0]], axis=1).astype(np.int32)
    rT, JTJ, JTr = oc.colored_update(pts, c["scol"], c["tgt"], tn, c["tcol"], rgrad, cor)
    np.testing.assert_allclose(v27[:21], JTJ[np.triu_indices(6)], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(v27[21:], JTr, rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(Tu, rT, atol=1e-10)
    # the loop
    g = K.Grid(P, 0.075 / 4, ring=8)
    r = K.icp(g, dev(c["src"]), P, 0.075, T0, mode=1, scol=dev(c["scol"]), tnrm=N, tcol=Cc, tgrad=dev(rgrad))
    ref = oc.icp_colored(c["src"], c["scol"], c["tgt"], tn, c["tcol"], 0.075, T0)
    assert rot_angle_deg(r["T"][:3, :3], ref.transformation[:3, :3]) < 0.1     # north_star: 0.1 deg / 1 mm
    assert np.linalg.norm(r["T"][:3, 3] - ref.transformation[:3, 3]) < 1e-3
    assert r["n_inliers"] == ref.n_inliers


# ------------------------------------------------------------------------------------------- a6 / a7
@pytest.mark.parametrize("nd,m,D", [(7, 20, 768), (3, 5, 512), (1, 1, 64), (12, 33, 768)])
def test_cosine_segmax(K, nd, m, D):
    from oracle import ref_port
    rng = np.random.default_rng(nd * m)
    det = rng.normal(size=(nd, D)).astype(np.float32)
    mem = [rng.normal(size=(rng.integers(1, 9), D)).astype(np.float32) * 3 for _ in range(m)]
    seg = np.cumsum([0] + [len(x) for x in mem]).astype(np.int32)
    S = h(K.cosine_segmax(dev(det, torch.float32), dev(np.concatenate(mem), torch.float32), dev(seg, torch.int32)))
    ref = ref_port.closest_similarities(det, mem)
    np.testing.assert_allclose(S, ref, rtol=0, atol=2e-6)     # fp32 dot products, summation order differs from BLAS


def test_simvolume_golden(cuda, golden):
    from instance_based_loc_b200.utils.similarity_volume import SimVolume
    g = golden
    for ci in range(int(g["sv_ncases"])):
        S = g[f"sv{ci}_S"]
        sv = SimVolume(S)
        sv.fast_construct_volume(min(S.shape[0], 3))
        got = sv.get_top_indices_from_subvolumes(num_per_length=4)
        lens, flat = g[f"sv{ci}_assn_len"], g[f"sv{ci}_assn_flat"]
        exp, o = [], 0
        for L in lens:
            exp.append([[int(flat[o + 2 * t]), int(flat[o + 2 * t + 1])] for t in range(L)])
            o += 2 * L
        assert got == exp, (ci, S.shape, got, exp)


@pytest.mark.parametrize("nd,m", [(7, 51), (3, 40), (2, 20), (1, 9), (5, 3), (3, 130), (2, 300), (1, 64), (4, 120)])
def test_simvolume_topk_vs_port(K, nd, m):
    from oracle import ref_port
    rng = np.random.default_rng(nd + 100 * m)
    S = rng.uniform(-0.3, 0.98, (nd, m)).astype(np.float32)
    S[rng.random((nd, m)) < 0.1] = 0.5                        # many exact fp16 ties
    k = 4 * nd * 4
    flat, val = K.simvolume_topk(dev(S, torch.float32), k)
    chosen, subs = ref_port.simvolume_subvolumes(S)
    assert flat.shape[0] == len(chosen)
    for ci, vol in enumerate(subs):
        ridx, rval = ref_port.topk_first_index(vol, k)
        assert np.array_equal(h(flat)[ci], ridx), (ci, h(flat)[ci][:8], ridx[:8])
        assert np.array_equal(h(val)[ci].astype(np.float16), rval)
    from instance_based_loc_b200.utils.similarity_volume import SimVolume
    sv = SimVolume(S)
    sv.fast_construct_volume(min(nd, 3))
    assert sv.get_top_indices_from_subvolumes(4) == ref_port.simvolume_assignments(S, 4)


@pytest.mark.parametrize("nd,m,neg", [(7, 500, False), (7, 500, True), (5, 257, True), (3, 1000, False)])
def test_simvolume_pruning_changes_nothing(K, nd, m, neg):
    """configs[3] sizes (500-object memory): the exact pair pruning (row lower bound for the value passes, the k-th value
    for the index passes) gives the entries of the full walk, negative similarities (positive products of two negatives)
    and many fp16 ties included."""
    rng = np.random.default_rng(nd * 1000 + m)
    S = rng.uniform(-0.9 if neg else -0.2, 0.98, (nd, m)).astype(np.float32)
    S[rng.random((nd, m)) < 0.15] = 0.75
    k = 4 * nd * 4
    try:
        K.simvolume_set_prune(False)
        f0, v0 = K.simvolume_topk(dev(S, torch.float32), k)
        K.simvolume_set_prune(True)
        f1, v1 = K.simvolume_topk(dev(S, torch.float32), k)
    finally:
        K.simvolume_set_prune(True)
    assert np.array_equal(h(f0), h(f1)) and np.array_equal(h(v0), h(v1))
