"""Torch-tensor wrappers over the C ABI (include/ibl.h).  PyTorch is only plumbing here: device
memory, the current CUDA stream and host<->device copies.  Every function takes / returns CUDA
tensors and raises if the CUDA library or a GPU is missing."""
from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_int64

import numpy as np
import torch

from ._lib import IblError, RegProblemT, RegResultT, ScoreT, check, lib

F64 = torch.float64


def _dev():
    if not torch.cuda.is_available():
        raise IblError("instance_based_loc_b200 needs a CUDA device (no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _stream():
    # raw handle of torch's current stream on the current device (torch.cuda.current_stream() costs ~15 us of
    # interpreter time per call, and there are ~140 calls per frame)
    return ctypes.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def to_dev(a, dtype=F64):
    """numpy / tensor -> contiguous CUDA tensor of dtype."""
    if isinstance(a, torch.Tensor):
        return a.to(device=_dev(), dtype=dtype).contiguous()
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(_dev())


def _T16(T):
    if T is None:
        return None
    arr = np.ascontiguousarray(np.asarray(T, dtype=np.float64).reshape(16))
    return (c_double * 16)(*arr.tolist())


# ---------------------------------------------------------------------------------------- clouds
def backproject_masks(depth, rgb, masks, fx, fy, want_colors=True):
    """depth [H,W] f32|f64 CUDA, rgb [H,W,3] u8 CUDA or None, masks [n,H,W] u8/bool CUDA.
    -> (pts [N,3] f64, cols [N,3] f64 | None, offsets list[int] of len n+1)."""
    H, W = depth.shape
    n = int(masks.shape[0])
    is64 = depth.dtype == torch.float64
    if not is64 and depth.dtype != torch.float32:
        raise IblError("depth must be float32 or float64")
    masks = masks.reshape(n, H * W)
    if masks.dtype == torch.bool:
        masks = masks.view(torch.uint8)
    masks = masks.contiguous()
    depth = depth.contiguous()
    hor = torch.from_numpy(np.linspace(-W / 2, W / 2, W, dtype=np.float32)).to(depth.device)
    ver = torch.from_numpy(np.linspace(H / 2, -H / 2, H, dtype=np.float32)).to(depth.device)
    offs = (c_int64 * (n + 1))()
    L = lib()
    args = (_p(depth), int(is64), _p(rgb) if rgb is not None else None, _p(masks), n, H, W, _p(hor), _p(ver),
            float(fx), float(fy))
    check(L.ibl_backproject_masks(*args, None, None, offs, _stream()), "backproject(count)")
    total = int(offs[n])
    pts = torch.empty((total, 3), dtype=F64, device=depth.device)
    cols = torch.empty((total, 3), dtype=F64, device=depth.device) if (rgb is not None and want_colors) else None
    if total:
        check(L.ibl_backproject_masks(*args, _p(pts), _p(cols), offs, _stream()), "backproject")
    return pts, cols, [int(o) for o in offs]


def voxel_downsample(pts, cols=None, nrm=None, voxel=0.01, want_ids=False, want_keys=False):
    n = pts.shape[0]
    dev = pts.device
    op = torch.empty_like(pts)
    oc = torch.empty_like(pts) if cols is not None else None
    on = torch.empty_like(pts) if nrm is not None else None
    vid = torch.empty(n, dtype=torch.int32, device=dev) if want_ids else None
    keys = torch.empty((n, 3), dtype=torch.int64, device=dev) if want_keys else None
    nv = check(lib().ibl_voxel_downsample(_p(pts), _p(cols), _p(nrm), n, float(voxel), _p(op), _p(oc), _p(on), _p(vid),
                                          _p(keys), _stream()), "voxel_downsample")
    return (op[:nv], oc[:nv] if oc is not None else None, on[:nv] if on is not None else None, vid, keys)


def select_rows(keep, a=None, b=None, c=None, want_index=False):
    n = keep.shape[0]
    dev = keep.device
    if keep.dtype == torch.bool:
        keep = keep.view(torch.uint8)
    oa = torch.empty_like(a) if a is not None else None
    ob = torch.empty_like(b) if b is not None else None
    oc = torch.empty_like(c) if c is not None else None
    oi = torch.empty(n, dtype=torch.int32, device=dev) if want_index else None
    k = check(lib().ibl_select_rows(_p(keep), n, _p(a), _p(b), _p(c), _p(oa), _p(ob), _p(oc), _p(oi), _stream()),
              "select_rows")
    cut = lambda t: t[:k] if t is not None else None  # noqa: E731
    return cut(oa), cut(ob), cut(oc), cut(oi)


def transform_points(T, pts, rotate_only=False, inplace=False):
    out = pts if inplace else torch.empty_like(pts)
    check(lib().ibl_transform_points(_T16(T), _p(pts), pts.shape[0], int(rotate_only), _p(out), _stream()), "transform")
    return out


def mean3(pts):
    m = (c_double * 3)()
    check(lib().ibl_mean3(_p(pts), pts.shape[0], m, _stream()), "mean3")
    return np.array([m[0], m[1], m[2]])


def mean3_many(clouds):
    """means of several [n,3] clouds with one launch and one synchronisation -> float64 [k,3]; row c equals mean3(clouds[c])."""
    k = len(clouds)
    out = np.zeros((k, 3))
    if k == 0:
        return out
    ptrs = (ctypes.c_void_p * k)(*[c.data_ptr() if c.shape[0] else None for c in clouds])
    ns = (c_int64 * k)(*[int(c.shape[0]) for c in clouds])
    check(lib().ibl_mean3_many(ptrs, ns, k, out.ctypes.data_as(ctypes.POINTER(c_double)), _stream()), "mean3_many")
    return out


# ------------------------------------------------------------------------------------------ grid
class Grid:
    """Owns an ibl_grid_t; keeps the point tensor alive."""

    def __init__(self, pts, cell, max_cells=0, ring=0, nnfield=0.0, nnfield_voxel=0.0, seg=None):
        """seg (int32 [n] CUDA, optional): per-point cloud id -> a SEGMENTED grid over a concatenation of clouds, on which
        radius queries only count neighbours of the query's own cloud (ibl_grid_create_segmented)."""
        self.pts = pts
        self.seg = seg
        self.h = ctypes.c_void_p()
        if seg is not None:
            check(lib().ibl_grid_create_segmented(_p(pts), pts.shape[0], _p(seg), float(cell), int(max_cells),
                                                  ctypes.byref(self.h), _stream()), "grid_create_segmented")
        else:
            check(lib().ibl_grid_create(_p(pts), pts.shape[0], float(cell), int(max_cells), ctypes.byref(self.h), _stream()),
                  "grid_create")
        if ring:
            check(lib().ibl_grid_build_ringfield(self.h, int(ring), _stream()), "ringfield")
        if nnfield:
            check(lib().ibl_grid_build_nnfield(self.h, float(nnfield), float(nnfield_voxel), _stream()), "nnfield")

    @property
    def nnfield_bytes(self):
        return int(lib().ibl_grid_nnfield_bytes(self.h))

    @property
    def cell(self):
        return lib().ibl_grid_cell(self.h)

    @property
    def occupied(self):
        return lib().ibl_grid_occupied(self.h)

    def __del__(self):
        try:
            if self.h:
                lib().ibl_grid_destroy(self.h)
                self.h = None
        except Exception:
            pass


def surface_spacing(pts, probe_cell):
    """Mean point spacing of a surface-like cloud from one coarse occupancy count: n points over ~occ * c0^2 of area.
    Only used to size search grids (performance, never results)."""
    n = pts.shape[0]
    if n == 0:
        return float(probe_cell)
    g0 = Grid(pts, max(probe_cell, 1e-4))
    occ = max(1, g0.occupied)
    c0 = g0.cell
    del g0
    return float(c0 * (occ / n) ** 0.5)


# cell = knn_cell_factor * R_k, R_k = radius holding k points of a surface.  The ring search stops at the first ring s with
# s * cell >= R_k.  With the two-pass kernel (distance histogram, then store only the bins that matter) the sort no longer
# depends on the cube size, so the balance is cell rows to walk (small cells) against candidates to measure (large cells).
# IBL_KNN_CELL_FACTOR overrides.
knn_cell_factor = float(os.environ.get("IBL_KNN_CELL_FACTOR", "0.9"))   # measured on B200 (tools/knn_probe.py, detected clouds of the bench): 0.8-1.0 best


def knn_cell_from_spacing(spacing, k, radius):
    """Grid cell for a hybrid k-NN search: k points of a surface lie within ~sqrt(k/pi) * spacing."""
    rk = (k / np.pi) ** 0.5 * spacing
    return float(min(max(rk * knn_cell_factor, 1e-4), radius))


def knn_cell_for(pts, k, radius):
    """Pick a grid cell for a k-NN search from the cloud's surface density."""
    if pts.shape[0] == 0:
        return radius
    return knn_cell_from_spacing(surface_spacing(pts, radius / 4.0), k, radius)


def radius_count(grid, radius):
    out = torch.empty(grid.pts.shape[0], dtype=torch.int32, device=grid.pts.device)
    check(lib().ibl_radius_count(grid.h, float(radius), _p(out), _stream()), "radius_count")
    return out


def radius_outlier_mask(grid, radius, nb_points):
    out = torch.empty(grid.pts.shape[0], dtype=torch.uint8, device=grid.pts.device)
    check(lib().ibl_radius_outlier_mask(grid.h, float(radius), int(nb_points), _p(out), _stream()), "radius_outlier")
    return out


def dbscan(grid, eps, min_points):
    """Open3D cluster_dbscan labels (int32[n] on the device, -1 = noise) and the number of clusters."""
    out = torch.empty(grid.pts.shape[0], dtype=torch.int32, device=grid.pts.device)
    nc = ctypes.c_int32(0)
    check(lib().ibl_dbscan(grid.h, float(eps), int(min_points), _p(out), ctypes.byref(nc), _stream()), "dbscan")
    return out, int(nc.value)


def bounds3(pts):
    """(min xyz, max xyz) of a device cloud as two host float64[3] arrays; bit-identical to numpy's min / max."""
    import numpy as np
    buf = (ctypes.c_double * 6)()
    check(lib().ibl_bounds3(_p(pts), pts.shape[0], buf, _stream()), "bounds3")
    a = np.array(buf[:], dtype=np.float64)
    return a[:3], a[3:]


def knn_hybrid(grid, radius, max_nn, queries=None, want_d2=True):
    nq = grid.pts.shape[0] if queries is None else queries.shape[0]
    dev = grid.pts.device
    idx = torch.empty((nq, max_nn), dtype=torch.int32, device=dev)
    d2 = torch.empty((nq, max_nn), dtype=F64, device=dev) if want_d2 else None
    cnt = torch.empty(nq, dtype=torch.int32, device=dev)
    check(lib().ibl_knn_hybrid(grid.h, _p(queries), nq, float(radius), int(max_nn), _p(idx), _p(d2), _p(cnt), _stream()),
          "knn_hybrid")
    return idx, d2, cnt


def nn_score(grid, src, max_d, T=None, want_idx=False, want_d2=False):
    n = src.shape[0]
    nn = torch.empty(n, dtype=torch.int32, device=src.device) if want_idx else None
    d2 = torch.empty(n, dtype=F64, device=src.device) if want_d2 else None
    res = ScoreT()
    check(lib().ibl_nn_score(grid.h, _p(src), n, _T16(T), float(max_d), _p(nn), _p(d2), ctypes.byref(res), _stream()),
          "nn_score")
    return int(res.n_inliers), float(res.err2), nn, d2


def nn_score_static(grid, src, max_d, out, T=None, nn=None, d2=None):
    """ibl_nn_score_static: nn_score with no allocation and no host sync.  ``out`` is a caller-owned device buffer of
    16 bytes (int64 n_inliers, float64 err2 -- e.g. torch.empty(2, dtype=torch.int64)); ``nn`` / ``d2`` optional caller-owned
    outputs.  At most IBL_NN_SCORE_STATIC_MAX_POINTS (2^20) source points; calls that overlap on the GPU race on the
    library's static scratch array."""
    assert out.is_cuda and out.numel() * out.element_size() >= 16 and out.is_contiguous()
    check(lib().ibl_nn_score_static(grid.h, _p(src), src.shape[0], _T16(T), float(max_d), _p(nn), _p(d2), _p(out), _stream()),
          "nn_score_static")


# -------------------------------------------------------------------------------------- features
def normals(pts, idx, cnt):
    out = torch.empty_like(pts)
    check(lib().ibl_normals(_p(pts), pts.shape[0], _p(idx), _p(cnt), idx.shape[1], _p(out), _stream()), "normals")
    return out


def spfh(pts, nrm, idx, cnt):
    out = torch.empty((pts.shape[0], 33), dtype=F64, device=pts.device)
    check(lib().ibl_spfh(_p(pts), _p(nrm), pts.shape[0], _p(idx), _p(cnt), idx.shape[1], _p(out), _stream()), "spfh")
    return out


def fpfh(idx, d2, cnt, spfh_t):
    out = torch.empty_like(spfh_t)
    check(lib().ibl_fpfh(spfh_t.shape[0], _p(idx), _p(d2), _p(cnt), idx.shape[1], _p(spfh_t), _p(out), _stream()), "fpfh")
    return out


def feature_nn(fa, fb):
    out = torch.empty(fa.shape[0], dtype=torch.int32, device=fa.device)
    check(lib().ibl_feature_nn(_p(fa), fa.shape[0], _p(fb), fb.shape[0], _p(out), _stream()), "feature_nn")
    return out


class FeaturePack:
    """Owns an ibl_feature_pack_t: the matcher's cached operands of one feature set (keeps the rows alive)."""

    def __init__(self, f):
        self.f = f
        self.h = ctypes.c_void_p()
        check(lib().ibl_feature_pack_create(_p(f), f.shape[0], ctypes.byref(self.h), _stream()), "feature_pack_create")

    def __del__(self):
        try:
            if self.h:
                lib().ibl_feature_pack_destroy(self.h)
                self.h = None
        except Exception:
            pass


def feature_match_packed(fa, pack):
    """Both search directions of a fresh set fa against a cached pack -> (nn_ab int32 [na], nn_ba int32 [pack rows],
    flag uint32 [1] device tensor: non-zero = fa outside the pack's fp16 range, results invalid)."""
    dev = fa.device
    ab = torch.empty(fa.shape[0], dtype=torch.int32, device=dev)
    ba = torch.empty(pack.f.shape[0], dtype=torch.int32, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    check(lib().ibl_feature_match_packed(_p(fa), fa.shape[0], pack.h, _p(ab), _p(ba), _p(flag), _stream()),
          "feature_match_packed")
    return ab, ba, flag


def feature_match_packed_mutual(fa, pack):
    """feature_match_packed for the mutual filter: nn_ba is computed only for the pack rows that occur in nn_ab (the only
    entries the filter reads) and is -1 elsewhere (ibl_feature_match_packed_mutual)."""
    dev = fa.device
    ab = torch.empty(fa.shape[0], dtype=torch.int32, device=dev)
    ba = torch.empty(pack.f.shape[0], dtype=torch.int32, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    check(lib().ibl_feature_match_packed_mutual(_p(fa), fa.shape[0], pack.h, _p(ab), _p(ba), _p(flag), _stream()),
          "feature_match_packed_mutual")
    return ab, ba, flag


def feature_scores_debug(fa, fb, swap_lbo_sbo=0):
    """Test hook: raw tcgen05 scores a32.b32 - |b32|^2/2 of the matcher's filter, float32 [na, nb], where a32 / b32 are
    the rows minus the returned centre, times the returned power-of-two scale.  -> (scores, centre float64 [33], scale)."""
    out = torch.empty((fa.shape[0], fb.shape[0]), dtype=torch.float32, device=fa.device)
    mu = torch.empty(34, dtype=torch.float64, device=fa.device)
    check(lib().ibl_feature_scores_debug(_p(fa), fa.shape[0], _p(fb), fb.shape[0], _p(out), _p(mu), int(swap_lbo_sbo),
                                         _stream()), "feature_scores_debug")
    return out, mu[:33], float(mu[33].item())


def feature_nn_stats(reset=True):
    """(band candidates, exact f64 evaluations) of feature_nn since the last reset."""
    v = (ctypes.c_int64 * 2)()
    check(lib().ibl_feature_nn_stats(v, int(reset)), "feature_nn_stats")
    return int(v[0]), int(v[1])


def feature_nn_wait_stats(start=False):
    """Wait clocks inside the tcgen05 matcher kernel (ibl_feature_nn_wait_stats).  start=True zeroes and enables the
    counters; start=False reads them (dict) and disables collection."""
    v = (ctypes.c_int64 * 8)()
    check(lib().ibl_feature_nn_wait_stats(v, 1 if start else 0), "feature_nn_wait_stats")
    if start:
        return None
    keys = ("ctas", "tiles", "mma_loop_clk", "mma_wait_accumulator_clk", "mma_wait_tile_clk", "epilogue_loop_clk",
            "epilogue_wait_accumulator_clk", "producer_wait_stage_clk")
    return dict(zip(keys, (int(x) for x in v)))


def feature_nn_set_pool_chunks(chunks=0):
    """Test hook: cap the matcher's candidate pool (0 = default) to force its pool-exhausted path."""
    check(lib().ibl_feature_nn_set_pool_chunks(int(chunks)), "feature_nn_set_pool_chunks")


def mutual_corres(ij, ji, ransac_n=3):
    cor = torch.empty((ij.shape[0], 2), dtype=torch.int32, device=ij.device)
    c = check(lib().ibl_mutual_corres(_p(ij), ij.shape[0], _p(ji), ji.shape[0], int(ransac_n), _p(cor), _stream()),
              "mutual_corres")
    return cor[:c]


# ---------------------------------------------------------------------------------------- ransac
def ransac_check(src, tgt, corres, H, edge_thr, dist_thr, samples=None, seed=0, h0=0, use_checkers=True, want_T=False):
    ok = torch.empty(H, dtype=torch.uint8, device=src.device)
    Ts = torch.zeros((H, 16), dtype=F64, device=src.device) if want_T else None
    check(lib().ibl_ransac_check(_p(src), _p(tgt), _p(corres), corres.shape[0], _p(samples), int(seed), int(h0), int(H),
                                 float(edge_thr), float(dist_thr), int(use_checkers), _p(ok), _p(Ts), _stream()),
          "ransac_check")
    return ok, Ts


def ransac_score(grid, src, tgt, corres, hyp_ids, max_d, samples=None, seed=0):
    nh = hyp_ids.shape[0]
    dev = src.device
    inl = torch.empty(nh, dtype=torch.int32, device=dev)
    err = torch.empty(nh, dtype=F64, device=dev)
    cin = torch.empty(nh, dtype=torch.int32, device=dev)
    Ts = torch.empty((nh, 16), dtype=F64, device=dev)
    check(lib().ibl_ransac_score(grid.h, _p(src), src.shape[0], _p(tgt), _p(corres), corres.shape[0], _p(samples),
                                 int(seed), _p(hyp_ids), nh, float(max_d), _p(inl), _p(err), _p(cin), _p(Ts), _stream()),
          "ransac_score")
    return inl, err, cin, Ts


def _reg_dict(r):
    return {"T": np.array(list(r.T)).reshape(4, 4), "fitness": r.fitness, "inlier_rmse": r.inlier_rmse,
            "n_inliers": r.n_inliers, "err2": r.err2, "best_hyp": r.best_hyp, "drawn": r.drawn,
            "validated": r.validated, "n_corres": r.n_corres, "iterations": r.iterations, "status": r.status}


def ransac_count(grid, src, tgt, corres, hyp_ids, max_d, samples=None, seed=0):
    """Inlier counts only (ibl_ransac_count) -> (inliers int32 [nh], T [nh,16])."""
    nh = hyp_ids.shape[0]
    inl = torch.empty(nh, dtype=torch.int32, device=src.device)
    Ts = torch.empty((nh, 16), dtype=F64, device=src.device)
    check(lib().ibl_ransac_count(grid.h, _p(src), src.shape[0], _p(tgt), _p(corres), corres.shape[0], _p(samples),
                                 int(seed), _p(hyp_ids), nh, float(max_d), _p(inl), _p(Ts), _stream()), "ransac_count")
    return inl, Ts


def ransac_run(grid, src, tgt, corres, max_iteration, confidence, max_d, edge_thr=0.9, samples=None, seed=0,
               use_checkers=True, early_exit=True, block=1 << 16, h0=0):
    """h0 > 0: only hypotheses [h0, max_iteration) of the sample stream (one rank's block, dist.hypothesis_block)."""
    r = RegResultT()
    check(lib().ibl_ransac_run_range(grid.h, _p(src), src.shape[0], _p(tgt), _p(corres), corres.shape[0], _p(samples),
                                     int(seed), int(h0), int(max_iteration), float(confidence), float(max_d),
                                     float(edge_thr), int(use_checkers), int(early_exit), int(block), ctypes.byref(r),
                                     _stream()), "ransac_run")
    return _reg_dict(r)


# ------------------------------------------------------------------------------------------- icp
def color_gradient(pts, nrm, col, idx, cnt):
    out = torch.empty_like(pts)
    check(lib().ibl_color_gradient(_p(pts), _p(nrm), _p(col), pts.shape[0], _p(idx), _p(cnt), idx.shape[1], _p(out),
                                   _stream()), "color_gradient")
    return out


def p2p_update(src, nn, tgt):
    T = (c_double * 16)()
    check(lib().ibl_p2p_update(_p(src), _p(nn), src.shape[0], _p(tgt), T, _stream()), "p2p_update")
    return np.array(list(T)).reshape(4, 4)


def colored_update(src, scol, nn, tgt, tnrm, tcol, tgrad, lambda_geometric=0.968):
    v = (c_double * 27)()
    T = (c_double * 16)()
    check(lib().ibl_colored_update(_p(src), _p(scol), _p(nn), src.shape[0], _p(tgt), _p(tnrm), _p(tcol), _p(tgrad),
                                   float(lambda_geometric), v, T, _stream()), "colored_update")
    return np.array(list(T)).reshape(4, 4), np.array(list(v))


def icp(grid, src, tgt, max_d, init=None, mode=0, scol=None, tnrm=None, tcol=None, tgrad=None, max_iteration=30,
        rel_fitness=1e-6, rel_rmse=1e-6, lambda_geometric=0.968):
    r = RegResultT()
    check(lib().ibl_icp(grid.h, _p(src), _p(scol), src.shape[0], _p(tgt), _p(tnrm), _p(tcol), _p(tgrad), int(mode),
                        float(max_d), _T16(init), int(max_iteration), float(rel_fitness), float(rel_rmse),
                        float(lambda_geometric), ctypes.byref(r), _stream()), "icp")
    return _reg_dict(r)


# --------------------------------------------------------------------------- batched registration
def reg_problems(specs):
    """specs: list of dicts {src, src_col, tgt, tgt_nrm, tgt_col, tgt_grad (CUDA f64 [n,3] or None), grid_ransac, grid_icp
    (Grid or None), corres (CUDA int32 [C,2] or None), seed} -> ctypes array of ibl_reg_problem_t.  The caller keeps the
    tensors alive."""
    arr = (RegProblemT * len(specs))()
    for q, sp in zip(arr, specs):
        src, cor = sp["src"], sp.get("corres")
        q.src, q.src_col, q.n_src = src.data_ptr(), (sp["src_col"].data_ptr() if sp.get("src_col") is not None else None), src.shape[0]
        tgt = sp["tgt"]
        q.tgt, q.n_tgt = tgt.data_ptr(), tgt.shape[0]
        for k in ("tgt_nrm", "tgt_col", "tgt_grad"):
            setattr(q, k, sp[k].data_ptr() if sp.get(k) is not None else None)
        q.grid_ransac = sp["grid_ransac"].h if sp.get("grid_ransac") is not None else None
        q.grid_icp = sp["grid_icp"].h if sp.get("grid_icp") is not None else None
        q.corres, q.n_corres = (cor.data_ptr(), cor.shape[0]) if cor is not None else (None, 0)
        q.seed = int(sp.get("seed", 0))
    return arr


def ransac_run_batch(problems, max_iteration, confidence, max_d, edge_thr=0.9, use_checkers=True, early_exit=True,
                     block=1 << 14):
    """ibl_ransac_run_batch over a reg_problems() array -> list of result dicts (one per problem, ransac_run's fields)."""
    n = len(problems)
    res = (RegResultT * n)()
    check(lib().ibl_ransac_run_batch(problems, n, int(max_iteration), float(confidence), float(max_d), float(edge_thr),
                                     int(use_checkers), int(early_exit), int(block), res, _stream()), "ransac_run_batch")
    return [_reg_dict(r) for r in res]


def ransac_set_prune(on=True):
    """Measurement switch of the batched RANSAC count pass (ibl_ransac_set_prune); results do not depend on it."""
    check(lib().ibl_ransac_set_prune(int(bool(on))), "ransac_set_prune")


def icp_colored_batch(problems, T_init, max_d, max_iteration=30, rel_fitness=1e-6, rel_rmse=1e-6, lambda_geometric=0.968):
    """ibl_icp_colored_batch: T_init host [n,4,4] -> list of result dicts."""
    n = len(problems)
    res = (RegResultT * n)()
    Ti = np.ascontiguousarray(np.asarray(T_init, np.float64).reshape(n, 16))
    check(lib().ibl_icp_colored_batch(problems, n, Ti.ctypes.data_as(ctypes.POINTER(c_double)), float(max_d), int(max_iteration),
                                      float(rel_fitness), float(rel_rmse), float(lambda_geometric), res, _stream()),
          "icp_colored_batch")
    return [_reg_dict(r) for r in res]


def nn_score_batch(grid, srcs, Ts, max_d):
    """evaluate_registration of many (source tensor, 4x4 transform) pairs against one target grid
    -> list of (n_inliers, err2)."""
    n = len(srcs)
    ptrs = (ctypes.c_void_p * n)(*[s.data_ptr() if s.shape[0] else None for s in srcs])
    ns = (c_int64 * n)(*[int(s.shape[0]) for s in srcs])
    Th = np.ascontiguousarray(np.asarray(Ts, np.float64).reshape(n, 16))
    res = (ScoreT * n)()
    check(lib().ibl_nn_score_batch(grid.h, ptrs, ns, Th.ctypes.data_as(ctypes.POINTER(c_double)), n, float(max_d), res,
                                   _stream()), "nn_score_batch")
    return [(int(r.n_inliers), float(r.err2)) for r in res]


# ------------------------------------------------------------------------------------ similarity
def cosine_segmax(det, mem, seg):
    nd, D = det.shape
    m = seg.shape[0] - 1
    S = torch.empty((nd, m), dtype=torch.float32, device=det.device)
    check(lib().ibl_cosine_segmax(_p(det), nd, _p(mem), mem.shape[0], D, _p(seg), m, _p(S), _stream()), "cosine_segmax")
    return S


def simvolume_set_prune(on=True):
    """Test / measurement switch of ibl_simvolume_topk's exact pair pruning (results do not depend on it)."""
    check(lib().ibl_simvolume_set_prune(int(bool(on))), "simvolume_set_prune")


def simvolume_topk(S, k):
    nd, m = S.shape
    s = min(nd, 3)
    ncomb = 1 if nd == 1 else int(round(np.prod([nd - i for i in range(s)]) / np.prod(range(1, s + 1))))
    flat = torch.empty((ncomb, k), dtype=torch.int64, device=S.device)
    val = torch.empty((ncomb, k), dtype=torch.float32, device=S.device)
    check(lib().ibl_simvolume_topk(_p(S), nd, m, int(k), _p(flat), _p(val), _stream()), "simvolume_topk")
    return flat, val


# ------------------------------------------------------------------------------- host wait policy
HOST_WAIT = {"auto": 0, "spin": 1, "yield": 2, "block": 3, "hybrid": 4}


def raycast_scene(T_wc, H, W, f, kind, half, R, t, room, height):
    """First hit per pixel of a synthetic scene: (depth t f64[H*W] (+inf = none), inst int32[H*W], normal f64[H*W,3])."""
    dev = _dev()
    P = int(H) * int(W)
    out_t = torch.empty(P, dtype=F64, device=dev)
    out_i = torch.empty(P, dtype=torch.int32, device=dev)
    out_n = torch.empty((P, 3), dtype=F64, device=dev)
    kind = to_dev(kind, torch.int32)
    half, R, t = to_dev(half), to_dev(R), to_dev(t)
    check(lib().ibl_raycast_scene(_T16(T_wc), int(H), int(W), float(f), _p(kind), _p(half), _p(R), _p(t), int(kind.shape[0]),
                                  float(room), float(height), _p(out_t), _p(out_i), _p(out_n), _stream()), "raycast_scene")
    return out_t, out_i, out_n


def set_host_wait(mode):
    """How host threads wait in stream synchronisations on the current device (ibl_set_host_wait)."""
    _dev()
    check(lib().ibl_set_host_wait(HOST_WAIT[mode] if isinstance(mode, str) else int(mode)), "set_host_wait")


def reserve_scratch(nbytes):
    """Grow the scratch pool for the current stream by nbytes (ibl_reserve_scratch)."""
    check(lib().ibl_reserve_scratch(int(nbytes), _stream()), "reserve_scratch")


def scratch_stats():
    """-> dict of the scratch pool's reserved / used bytes (now and high-water)."""
    v = (c_int64 * 4)()
    check(lib().ibl_scratch_stats(v), "scratch_stats")
    return {"reserved": int(v[0]), "reserved_high": int(v[1]), "used": int(v[2]), "used_high": int(v[3])}


# ------------------------------------------------------------------------------- instrumentation
def launch_count():
    return int(lib().ibl_launch_count())


def prof_enable(on=True):
    lib().ibl_prof_enable(int(bool(on)))


def prof_read():
    """-> {tag: {"ms", "launches", "work"}} accumulated since the last read (device-synchronising)."""
    L = lib()
    n = L.ibl_prof_num_tags()
    ms, la, wk = (c_double * n)(), (c_int64 * n)(), (c_double * n)()
    check(L.ibl_prof_read(ms, la, wk), "prof_read")
    return {L.ibl_prof_tag_name(i).decode(): {"ms": ms[i], "launches": int(la[i]), "work": wk[i]} for i in range(n)
            if la[i] > 0}
