"""oracle/o3d_core.py -- TEST INFRASTRUCTURE ONLY (PARITY UNPINNED, see oracle/__init__.py).

Array-level restatement of the Open3D 0.17.0 calls on the reference hot path
(utils/fpfh_register.py:86-150, utils/depth_utils.py:41,87, object_memory/object_memory.py:994),
following SURVEY.md Appendix A.  Neighbour searches use scipy's cKDTree to find candidates and
then re-rank with the exact d2 = (dx*dx + dy*dy) + dz*dz in f64 (nanoflann L2 accumulation
order), ascending (d2, index).  Scalar arithmetic lives in o3d_core.c / o3d_reg.c.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess

import numpy as np
from scipy.spatial import cKDTree

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    """Compile the C restatement into oracle/_build/liboracle.so (gcc, no FMA contraction)."""
    out_dir = os.path.join(_HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("o3d_core.c", "o3d_reg.c")]
    if not force and os.path.exists(so) and all(os.path.getmtime(so) >= os.path.getmtime(s) for s in srcs):
        return so
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
           "-fvisibility=hidden", "-o", so] + srcs + ["-lm"]
    subprocess.check_call(cmd)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "_build", "liboracle.so")
        if not os.path.exists(so):
            so = build()
        else:
            try:
                so = build()
            except Exception:
                pass
        _LIB = ctypes.CDLL(so)
        _LIB.orc_corres_inliers.restype = ctypes.c_int32
        _LIB.orc_ldlt_solve.restype = ctypes.c_int
        _LIB.orc_solve_update.restype = ctypes.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


# ------------------------------------------------------------------------------------------------
# A.1 searches
# ------------------------------------------------------------------------------------------------
def exact_d2(a, b):
    d = a - b
    return (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]


def knn_hybrid(points, queries, radius, max_nn, tree=None, workers=-1):
    """SearchHybrid for every query: (idx[nq,max_nn] int32, -1 padded; d2[nq,max_nn]; cnt[nq])."""
    points = _f64(points)
    queries = _f64(queries)
    n = len(points)
    nq = len(queries)
    idx_out = np.full((nq, max_nn), -1, np.int32)
    d2_out = np.full((nq, max_nn), np.inf, np.float64)
    cnt = np.zeros(nq, np.int32)
    if n == 0 or nq == 0:
        return idx_out, d2_out, cnt
    if tree is None:
        tree = cKDTree(points)
    kq = min(n, max_nn + 8)
    ub = radius * (1.0 + 1e-9) if np.isfinite(radius) else np.inf
    _, ci = tree.query(queries, k=kq, distance_upper_bound=ub, workers=workers)
    if kq == 1:
        ci = ci[:, None]
    valid = ci < n
    cis = np.where(valid, ci, 0)
    d2 = exact_d2(points[cis], queries[:, None, :])
    r2 = radius * radius
    d2 = np.where(valid & (d2 < r2), d2, np.inf)
    cikey = np.where(np.isfinite(d2), ci, np.iinfo(np.int64).max)
    order = np.lexsort((cikey, d2), axis=1)
    d2s = np.take_along_axis(d2, order, axis=1)[:, :max_nn]
    cs = np.take_along_axis(ci, order, axis=1)[:, :max_nn]
    ok = np.isfinite(d2s)
    w = d2s.shape[1]
    idx_out[:, :w] = np.where(ok, cs, -1)
    d2_out[:, :w] = d2s
    cnt[:] = ok.sum(axis=1)
    return idx_out, d2_out, cnt


def radius_count(points, radius, workers=-1):
    """#{j : d2(i,j) < r^2} including i, exact (A.2)."""
    points = _f64(points)
    n = len(points)
    if n == 0:
        return np.zeros(0, np.int64)
    tree = cKDTree(points)
    lo = tree.query_ball_point(points, radius * (1 - 1e-9), return_length=True, workers=workers)
    hi = tree.query_ball_point(points, radius * (1 + 1e-9), return_length=True, workers=workers)
    cnt = np.asarray(lo, np.int64).copy()
    amb = np.nonzero(np.asarray(hi) != cnt)[0]
    r2 = radius * radius
    for i in amb:
        cand = tree.query_ball_point(points[i], radius * (1 + 1e-9))
        cnt[i] = int(np.count_nonzero(exact_d2(points[cand], points[i]) < r2))
    return cnt


def radius_neighbours(points, radius, workers=-1):
    """CSR lists of {j : d2(i,j) < r^2} (self included, ascending index) for every point, exact (A.2)."""
    points = _f64(points)
    n = len(points)
    ptr = np.zeros(n + 1, np.int64)
    if n == 0:
        return ptr, np.zeros(0, np.int32)
    tree = cKDTree(points)
    cand = tree.query_ball_point(points, radius * (1 + 1e-9), workers=workers)
    r2 = radius * radius
    out = []
    for i in range(n):
        c = np.asarray(cand[i], np.int64)
        c = np.sort(c[exact_d2(points[c], points[i]) < r2])
        out.append(c.astype(np.int32))
        ptr[i + 1] = ptr[i] + len(c)
    return ptr, (np.concatenate(out) if out else np.zeros(0, np.int32))


def cluster_dbscan(points, eps, min_points):
    """PointCloud.cluster_dbscan (object_memory.py:315,511,632): int32 labels, -1 = noise (sequential C restatement)."""
    points = _f64(points)
    n = len(points)
    labels = np.zeros(n, np.int32)
    if n == 0:
        return labels
    ptr, idx = radius_neighbours(points, eps)
    idx = _i32(idx)
    lib().orc_dbscan(ctypes.c_int(n), _p(ptr), _p(idx), ctypes.c_int(int(min_points)), _p(labels))
    return labels


def remove_radius_outlier_mask(points, nb_points, radius):
    return radius_count(points, radius) > nb_points


def nn1(target, queries, max_d, tree=None, workers=-1):
    """SearchHybrid(q, max_d, 1): (idx int32 or -1, d2)."""
    target = _f64(target)
    queries = _f64(queries)
    if len(target) == 0 or len(queries) == 0:
        return np.full(len(queries), -1, np.int32), np.full(len(queries), np.inf)
    if tree is None:
        tree = cKDTree(target)
    k = min(2, len(target))
    _, ci = tree.query(queries, k=k, distance_upper_bound=max_d * (1 + 1e-9), workers=workers)
    if k == 1:
        ci = ci[:, None]
    valid = ci < len(target)
    cis = np.where(valid, ci, 0)
    d2 = exact_d2(target[cis], queries[:, None, :])
    d2 = np.where(valid, d2, np.inf)
    key = np.where(valid, ci, np.iinfo(np.int64).max)
    order = np.lexsort((key, d2), axis=1)[:, 0]
    rows = np.arange(len(queries))
    bd = d2[rows, order]
    bi = ci[rows, order]
    ok = bd < max_d * max_d
    return np.where(ok, bi, -1).astype(np.int32), np.where(ok, bd, np.inf)


# ------------------------------------------------------------------------------------------------
# A.3 / A.4 normals and FPFH
# ------------------------------------------------------------------------------------------------
def estimate_normals(points, radius, max_nn):
    points = _f64(points)
    n = len(points)
    idx, _, cnt = knn_hybrid(points, points, radius, max_nn)
    out = np.zeros((n, 3))
    lib().orc_normals(_p(points), n, _p(idx), _p(cnt), max_nn, _p(out))
    return out


def compute_fpfh(points, normals, radius, max_nn, return_spfh=False):
    """-> fpfh [33, n] float64 (Open3D Feature.data layout)."""
    points = _f64(points)
    normals = _f64(normals)
    n = len(points)
    idx, d2, cnt = knn_hybrid(points, points, radius, max_nn)
    d2 = _f64(np.where(np.isfinite(d2), d2, 0.0))
    spfh = np.zeros((33, n))
    fpfh = np.zeros((33, n))
    lib().orc_spfh(_p(points), _p(normals), n, _p(idx), _p(cnt), max_nn, _p(spfh))
    lib().orc_fpfh(n, _p(idx), _p(d2), _p(cnt), max_nn, _p(spfh), _p(fpfh))
    if return_spfh:
        return fpfh, spfh
    return fpfh


def feature_nn(fa, fb):
    """argmin_j |fa[:,i]-fb[:,j]|^2, exact f64, lowest index on ties."""
    fa = _f64(fa)
    fb = _f64(fb)
    nn = np.zeros(fa.shape[1], np.int32)
    lib().orc_feature_nn(_p(fa), fa.shape[1], _p(fb), fb.shape[1], _p(nn))
    return nn


def match_features(src_f, tgt_f, mutual_filter=True, ransac_n=3):
    """A.5 -> corres int32 [C,2]."""
    ns, nt = src_f.shape[1], tgt_f.shape[1]
    ij = feature_nn(src_f, tgt_f)
    corres_ij = np.stack([np.arange(ns, dtype=np.int32), ij], axis=1)
    if not mutual_filter:
        return corres_ij
    ji = feature_nn(tgt_f, src_f)
    keep = ji[ij] == np.arange(ns)
    mutual = corres_ij[keep]
    if len(mutual) >= ransac_n * 3:
        return mutual
    return corres_ij


# ------------------------------------------------------------------------------------------------
# A.6 - A.9 registration
# ------------------------------------------------------------------------------------------------
def transform_points(T, pts):
    pts = _f64(pts)
    T = _f64(T)
    out = np.empty_like(pts)
    lib().orc_transform_points(_p(T), _p(pts), len(pts), _p(out))
    return out


def kabsch(src, tgt):
    src = _f64(src)
    tgt = _f64(tgt)
    T = np.zeros((4, 4))
    lib().orc_kabsch(_p(src), _p(tgt), len(src), _p(T))
    return T


class RegResult:
    def __init__(self, T=None):
        self.transformation = np.eye(4) if T is None else np.array(T, dtype=np.float64)
        self.fitness = 0.0
        self.inlier_rmse = 0.0
        self.correspondence_set = np.zeros((0, 2), np.int32)
        self.n_inliers = 0
        self.err2 = 0.0

    def better_than(self, o):
        return self.fitness > o.fitness or (self.fitness == o.fitness and self.inlier_rmse < o.inlier_rmse)


def score(src_transformed, target, tree, max_d, T):
    """GetRegistrationResultAndCorrespondences (A.6)."""
    res = RegResult(T)
    if max_d <= 0 or len(src_transformed) == 0:
        return res
    idx, d2 = nn1(target, src_transformed, max_d, tree=tree)
    ok = idx >= 0
    ncor = int(ok.sum())
    res.n_inliers = ncor
    if ncor == 0:
        return res
    err2 = float(np.sum(d2[ok]))
    res.err2 = err2
    res.correspondence_set = np.stack([np.nonzero(ok)[0].astype(np.int32), idx[ok]], axis=1)
    res.fitness = ncor / len(src_transformed)
    res.inlier_rmse = math.sqrt(err2 / ncor)
    return res


def evaluate_registration(source, target, max_d, T=None):
    source = _f64(source)
    target = _f64(target)
    T = np.eye(4) if T is None else _f64(T)
    pts = source if np.array_equal(T, np.eye(4)) else transform_points(T, source)
    return score(pts, target, cKDTree(target) if len(target) else None, max_d, T)


def ransac_check(src, tgt, corres, samples, edge_thr, dist_thr, want_T=False):
    src, tgt = _f64(src), _f64(tgt)
    corres, samples = _i32(corres), _i32(samples)
    H = len(samples)
    ok = np.zeros(H, np.uint8)
    Ts = np.zeros((H, 4, 4)) if want_T else None
    lib().orc_ransac_check(_p(src), _p(tgt), _p(corres), _p(samples), ctypes.c_int64(H),
                           ctypes.c_double(edge_thr), ctypes.c_double(dist_thr), _p(ok),
                           _p(Ts) if want_T else None)
    return (ok.astype(bool), Ts) if want_T else ok.astype(bool)


def hyp_transform(src, tgt, corres, sample):
    s = _f64(src)[corres[sample, 0]]
    t = _f64(tgt)[corres[sample, 1]]
    return kabsch(s, t)


def ransac(src, tgt, corres, samples, max_d, edge_thr=0.9, confidence=0.99, early_exit=True,
           checkers=True, block=1 << 12, trace=None, max_iteration=None):
    """A.7 with sequential semantics over the host-provided sample stream.

    samples: int32 [max_iteration, 3] indices into corres, OR a callable (h0, h1) -> int32 [h1-h0, 3] together
    with max_iteration (the counter-based stream of oracle/sampler.py).  Returns (RegResult, info dict).
    """
    src, tgt = _f64(src), _f64(tgt)
    corres = _i32(corres)
    if callable(samples):
        H = int(max_iteration)
        get = samples
    else:
        samples = _i32(samples)
        H = len(samples)
        get = lambda a, b: samples[a:b]  # noqa: E731
    best = RegResult()
    best_h = -1
    info = {"drawn": 0, "validated": 0, "best_hyp": -1}
    if len(corres) < 3 or max_d <= 0 or H == 0:
        return best, info
    est_k = H
    h0 = 0
    L = lib()
    L.orc_grid_build.restype = ctypes.c_void_p
    grid = ctypes.c_void_p(L.orc_grid_build(_p(tgt), ctypes.c_int64(len(tgt)), ctypes.c_double(max_d)))
    block0 = block
    while h0 < min(H, est_k):
        h1 = min(H, h0 + block)
        blk = _i32(get(h0, h1))
        if checkers:
            ok = ransac_check(src, tgt, corres, blk, edge_thr, max_d)
        else:
            ok = np.ones(h1 - h0, bool)
        offs = np.nonzero(ok)[0]
        offs = offs[offs + h0 < est_k]
        if len(offs):
            # every validated hypothesis of the block is scored over the whole source (A.6) -- in parallel over the
            # hypotheses, like Open3D's OpenMP loop; the ordered best-update below consumes them in stream order
            sel = _i32(blk[offs])
            Ts = np.zeros((len(offs), 4, 4))
            L.orc_hyp_transforms(_p(src), _p(tgt), _p(corres), _p(sel), ctypes.c_int64(len(offs)), _p(Ts))
            cnt = np.zeros(len(offs), np.int64)
            err = np.zeros(len(offs))
            for a in range(0, len(offs), 128):                       # bounded scratch: 128 x n_src doubles per call
                b = min(a + 128, len(offs))
                L.orc_score_many(grid, _p(src), ctypes.c_int64(len(src)), _p(Ts[a:b]), ctypes.c_int64(b - a),
                                 ctypes.c_double(max_d), _p(cnt[a:b]), _p(err[a:b]))
        for k, off in enumerate(offs):
            h = h0 + int(off)
            if h >= est_k:
                break
            T = Ts[k]
            res = RegResult(T)
            res.n_inliers = int(cnt[k])
            if res.n_inliers:
                res.err2 = float(err[k])
                res.fitness = res.n_inliers / len(src)
                res.inlier_rmse = math.sqrt(res.err2 / res.n_inliers)
            info["validated"] += 1
            if trace is not None:
                trace.append((h, res.n_inliers, res.err2))
            if res.better_than(best):
                best = res
                best_h = h
                if early_exit:
                    inl = L.orc_corres_inliers(_p(_f64(T)), _p(src), _p(tgt), _p(corres), len(corres),
                                               ctypes.c_double(max_d))
                    rho = inl / len(corres)
                    den = math.log(1.0 - rho ** 3) if rho < 1.0 else -math.inf
                    if den == 0.0:
                        kk = math.inf
                    else:
                        kk = math.log(1.0 - confidence) / den
                    if kk < est_k:      # Registration.cpp: est_k_local_d < est_k_global ? ceil : keep
                        est_k = int(math.ceil(kk))
        h0 = h1
        block = min(block * 2, max(block0, 1 << 17))                  # fewer, larger blocks once the stream gets long
    L.orc_grid_free(grid)
    if best_h >= 0:
        # the winner's correspondence set (RegistrationResult.correspondence_set_), by the one-at-a-time scorer; its
        # count and err2 must be what the batched scorer said
        full = score(transform_points(best.transformation, src), tgt, cKDTree(tgt), max_d, best.transformation)
        assert full.n_inliers == best.n_inliers and full.err2 == best.err2, "batched and single scorers disagree"
        best = full
    info["drawn"] = int(min(H, max(est_k, 0)))
    info["best_hyp"] = best_h
    return best, info


def color_gradient(points, normals, colors, max_d):
    points, normals, colors = _f64(points), _f64(normals), _f64(colors)
    n = len(points)
    idx, _, cnt = knn_hybrid(points, points, max_d * 2.0, 30)
    g = np.zeros((n, 3))
    lib().orc_color_gradient(_p(points), _p(normals), _p(colors), n, _p(idx), _p(cnt), 30, _p(g))
    return g


def colored_update(src, scol, tgt, tnrm, tcol, tgrad, corres, lambda_geometric=0.968):
    corres = _i32(corres)
    JTJ = np.zeros((6, 6))
    JTr = np.zeros(6)
    lib().orc_colored_jtj(_p(_f64(src)), _p(_f64(scol)), _p(_f64(tgt)), _p(_f64(tnrm)), _p(_f64(tcol)),
                          _p(_f64(tgrad)), _p(corres), len(corres), ctypes.c_double(lambda_geometric),
                          _p(JTJ), _p(JTr))
    T = np.zeros((4, 4))
    lib().orc_solve_update(_p(JTJ), _p(JTr), _p(T))
    return T, JTJ, JTr


def p2p_update(src, tgt, corres):
    corres = _i32(corres)
    T = np.zeros((4, 4))
    lib().orc_kabsch_corres(_p(_f64(src)), _p(_f64(tgt)), _p(corres), len(corres), _p(T))
    return T


def icp(source, target, max_d, init, estimate, max_iteration=30, rel_fitness=1e-6, rel_rmse=1e-6):
    """RegistrationICP (A.9). estimate(pcd_points, corres) -> 4x4 update."""
    source, target = _f64(source), _f64(target)
    T = np.array(init, dtype=np.float64)
    tree = cKDTree(target)
    pcd = source.copy()
    if not np.array_equal(T, np.eye(4)):
        pcd = transform_points(T, pcd)
    res = score(pcd, target, tree, max_d, T)
    iters = 0
    for _ in range(max_iteration):
        upd = estimate(pcd, res.correspondence_set)
        T = upd @ T
        pcd = transform_points(upd, pcd)
        prev = res
        res = score(pcd, target, tree, max_d, T)
        iters += 1
        if abs(prev.fitness - res.fitness) < rel_fitness and abs(prev.inlier_rmse - res.inlier_rmse) < rel_rmse:
            break
    res.iterations = iters
    return res


def icp_p2p(source, target, max_d, init=None):
    init = np.eye(4) if init is None else init
    return icp(source, target, max_d, init, lambda pcd, cor: p2p_update(pcd, target, cor))


def icp_colored(source, scol, target, tnrm, tcol, max_d, init=None, lambda_geometric=0.968):
    init = np.eye(4) if init is None else init
    grad = color_gradient(target, tnrm, tcol, max_d)
    return icp(source, target, max_d, init,
               lambda pcd, cor: colored_update(pcd, scol, target, tnrm, tcol, grad, cor, lambda_geometric)[0])
