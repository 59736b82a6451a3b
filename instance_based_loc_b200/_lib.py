"""ctypes binding of libibl.so (include/ibl.h).  Fails loudly when the CUDA library is missing:
there is NO CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import POINTER, c_char_p, c_double, c_int, c_int32, c_int64, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libibl.so")
_lib = None


class IblError(RuntimeError):
    pass


class ScoreT(ctypes.Structure):
    _fields_ = [("n_inliers", c_int64), ("err2", c_double)]


class RegResultT(ctypes.Structure):
    _fields_ = [("T", c_double * 16), ("fitness", c_double), ("inlier_rmse", c_double), ("n_inliers", c_int64),
                ("err2", c_double), ("best_hyp", c_int64), ("drawn", c_int64), ("validated", c_int64),
                ("n_corres", c_int64), ("iterations", c_int32), ("status", c_int32)]


class RegProblemT(ctypes.Structure):
    """ibl_reg_problem_t (include/ibl.h): one registration problem of a batch, device pointers."""
    _fields_ = [("src", c_void_p), ("src_col", c_void_p), ("n_src", c_int64),
                ("tgt", c_void_p), ("tgt_nrm", c_void_p), ("tgt_col", c_void_p), ("tgt_grad", c_void_p), ("n_tgt", c_int64),
                ("grid_ransac", c_void_p), ("grid_icp", c_void_p), ("corres", c_void_p), ("n_corres", c_int64),
                ("seed", c_uint64)]


def build(verbose: bool = False) -> str:
    """Compile libibl.so in-tree with nvcc for sm_100a (csrc/Makefile)."""
    out = subprocess.run(["make", "-j8", "-C", os.path.join(_HERE, "csrc")], capture_output=True, text=True)
    if out.returncode != 0:
        raise IblError("building libibl.so failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


_V = c_void_p
_PROTOS = {
    "ibl_last_error_string": (c_char_p, []),
    "ibl_version": (c_int, []),
    "ibl_backproject_masks": (c_int, [_V, c_int, _V, _V, c_int, c_int, c_int, _V, _V, c_double, c_double, _V, _V,
                                      POINTER(c_int64), _V]),
    "ibl_voxel_downsample": (c_int64, [_V, _V, _V, c_int64, c_double, _V, _V, _V, _V, _V, _V]),
    "ibl_select_rows": (c_int64, [_V, c_int64, _V, _V, _V, _V, _V, _V, _V, _V]),
    "ibl_transform_points": (c_int, [POINTER(c_double), _V, c_int64, c_int, _V, _V]),
    "ibl_mean3": (c_int, [_V, c_int64, POINTER(c_double), _V]),
    "ibl_mean3_many": (c_int, [POINTER(c_void_p), POINTER(c_int64), c_int32, POINTER(c_double), _V]),
    "ibl_grid_create": (c_int, [_V, c_int64, c_double, c_int64, POINTER(c_void_p), _V]),
    "ibl_grid_create_segmented": (c_int, [_V, c_int64, _V, c_double, c_int64, POINTER(c_void_p), _V]),
    "ibl_grid_destroy": (None, [_V]),
    "ibl_grid_cell": (c_double, [_V]),
    "ibl_grid_occupied": (c_int64, [_V]),
    "ibl_grid_build_ringfield": (c_int, [_V, c_int, _V]),
    "ibl_grid_build_nnfield": (c_int, [_V, c_double, c_double, _V]),
    "ibl_grid_nnfield_bytes": (c_int64, [_V]),
    "ibl_radius_count": (c_int, [_V, c_double, _V, _V]),
    "ibl_radius_outlier_mask": (c_int, [_V, c_double, c_int32, _V, _V]),
    "ibl_dbscan": (c_int, [_V, c_double, c_int32, _V, POINTER(c_int32), _V]),
    "ibl_bounds3": (c_int, [_V, c_int64, POINTER(c_double), _V]),
    "ibl_raycast_scene": (c_int, [POINTER(c_double), c_int32, c_int32, c_double, _V, _V, _V, _V, c_int32, c_double, c_double,
                                  _V, _V, _V, _V]),
    "ibl_knn_hybrid": (c_int, [_V, _V, c_int64, c_double, c_int32, _V, _V, _V, _V]),
    "ibl_nn_score": (c_int, [_V, _V, c_int64, POINTER(c_double), c_double, _V, _V, POINTER(ScoreT), _V]),
    "ibl_nn_score_static": (c_int, [_V, _V, c_int64, POINTER(c_double), c_double, _V, _V, _V, _V]),
    "ibl_normals": (c_int, [_V, c_int64, _V, _V, c_int32, _V, _V]),
    "ibl_spfh": (c_int, [_V, _V, c_int64, _V, _V, c_int32, _V, _V]),
    "ibl_fpfh": (c_int, [c_int64, _V, _V, _V, c_int32, _V, _V, _V]),
    "ibl_feature_nn": (c_int, [_V, c_int64, _V, c_int64, _V, _V]),
    "ibl_feature_pack_create": (c_int, [_V, c_int64, POINTER(c_void_p), _V]),
    "ibl_feature_pack_destroy": (None, [_V]),
    "ibl_feature_match_packed": (c_int, [_V, c_int64, _V, _V, _V, _V, _V]),
    "ibl_feature_match_packed_mutual": (c_int, [_V, c_int64, _V, _V, _V, _V, _V]),
    "ibl_feature_scores_debug": (c_int, [_V, c_int64, _V, c_int64, _V, _V, c_int, _V]),
    "ibl_feature_nn_stats": (c_int, [POINTER(c_int64), c_int]),
    "ibl_feature_nn_wait_stats": (c_int, [POINTER(c_int64), c_int]),
    "ibl_feature_nn_set_pool_chunks": (c_int, [c_int]),
    "ibl_mutual_corres": (c_int64, [_V, c_int64, _V, c_int64, c_int32, _V, _V]),
    "ibl_ransac_check": (c_int, [_V, _V, _V, c_int64, _V, c_uint64, c_int64, c_int64, c_double, c_double, c_int, _V,
                                 _V, _V]),
    "ibl_ransac_score": (c_int, [_V, _V, c_int64, _V, _V, c_int64, _V, c_uint64, _V, c_int64, c_double, _V, _V, _V,
                                 _V, _V]),
    "ibl_ransac_count": (c_int, [_V, _V, c_int64, _V, _V, c_int64, _V, c_uint64, _V, c_int64, c_double, _V, _V, _V]),
    "ibl_ransac_run": (c_int, [_V, _V, c_int64, _V, _V, c_int64, _V, c_uint64, c_int64, c_double, c_double, c_double,
                               c_int, c_int, c_int64, POINTER(RegResultT), _V]),
    "ibl_ransac_run_range": (c_int, [_V, _V, c_int64, _V, _V, c_int64, _V, c_uint64, c_int64, c_int64, c_double, c_double,
                                     c_double, c_int, c_int, c_int64, POINTER(RegResultT), _V]),
    "ibl_ransac_run_batch": (c_int, [POINTER(RegProblemT), c_int32, c_int64, c_double, c_double, c_double, c_int, c_int,
                                     c_int64, POINTER(RegResultT), _V]),
    "ibl_ransac_set_prune": (c_int, [c_int]),
    "ibl_icp_colored_batch": (c_int, [POINTER(RegProblemT), c_int32, POINTER(c_double), c_double, c_int, c_double, c_double,
                                      c_double, POINTER(RegResultT), _V]),
    "ibl_nn_score_batch": (c_int, [_V, POINTER(c_void_p), POINTER(c_int64), POINTER(c_double), c_int32, c_double,
                                   POINTER(ScoreT), _V]),
    "ibl_color_gradient": (c_int, [_V, _V, _V, c_int64, _V, _V, c_int32, _V, _V]),
    "ibl_p2p_update": (c_int, [_V, _V, c_int64, _V, POINTER(c_double), _V]),
    "ibl_colored_update": (c_int, [_V, _V, _V, c_int64, _V, _V, _V, _V, c_double, POINTER(c_double),
                                   POINTER(c_double), _V]),
    "ibl_icp": (c_int, [_V, _V, _V, c_int64, _V, _V, _V, _V, c_int, c_double, POINTER(c_double), c_int, c_double,
                        c_double, c_double, POINTER(RegResultT), _V]),
    "ibl_cosine_segmax": (c_int, [_V, c_int32, _V, c_int64, c_int32, _V, c_int32, _V, _V]),
    "ibl_simvolume_set_prune": (c_int, [c_int]),
    "ibl_simvolume_topk": (c_int, [_V, c_int32, c_int32, c_int32, _V, _V, _V]),
    "ibl_set_host_wait": (c_int, [c_int]),
    "ibl_reserve_scratch": (c_int, [c_int64, _V]),
    "ibl_scratch_stats": (c_int, [POINTER(c_int64)]),
    "ibl_launch_count": (c_int64, []),
    "ibl_prof_enable": (None, [c_int]),
    "ibl_prof_num_tags": (c_int, []),
    "ibl_prof_tag_name": (c_char_p, [c_int]),
    "ibl_prof_read": (c_int, [POINTER(c_double), POINTER(c_int64), POINTER(c_double)]),
}

EXPORTS = sorted(_PROTOS)


def lib():
    """Load libibl.so (no build attempt here: build() is explicit)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise IblError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(nvcc, sm_100a).  This package has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc, what=""):
    if rc < 0:
        msg = lib().ibl_last_error_string()
        raise IblError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")
    return rc
