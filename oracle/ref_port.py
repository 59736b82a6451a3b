"""oracle/ref_port.py -- TEST INFRASTRUCTURE ONLY.

Array-level restatement of the reference's own numpy code on the hot path.  Unlike the Open3D
part this IS pinned: tests/test_oracle_vs_reference.py runs the unmodified reference modules
(/root/reference/utils/{depth_utils,similarity_volume}.py) in the build container and compares,
and tests/golden/ holds vectors generated from them (tests/golden/make_golden.py).
"""
from __future__ import annotations

import itertools

import numpy as np


# a1: utils/depth_utils.py:46-90 (and :12-44).  NB w,h = depth.shape => w is ROWS.
def backproject(depth, rgb, fx, fy):
    """-> (points f64 [N,3], colors f64 [N,3] | None), row-major order, Z==0 dropped."""
    w, h = depth.shape
    hor = np.linspace(-h / 2, h / 2, h, dtype=np.float32)
    ver = np.linspace(w / 2, -w / 2, w, dtype=np.float32).reshape(-1, 1)
    hor = np.tile(hor, (w, 1))
    ver = np.tile(ver, (1, h))
    X = hor * depth / fx
    Y = ver * depth / fy
    Z = depth
    pts = np.stack([X, Y, Z], axis=2).reshape(-1, 3)
    valid = pts[:, 2] != 0
    pts = pts[valid]
    cols = None
    if rgb is not None:
        cols = (rgb.astype(np.float32) / 255.0).reshape(-1, 3)[valid].astype(np.float64)
    return pts.astype(np.float64), cols


# a2: utils/depth_utils.py:176-206
def backproject_masks(depth, rgb, masks, fx, fy):
    out = []
    for i in range(masks.shape[0]):
        m = np.asarray(masks[i]).squeeze()
        out.append(backproject(depth * m, rgb, fx, fy))
    return out


# a3: utils/depth_utils.py:211-265
def voxel_indices(points, voxel_size):
    return np.floor(np.asarray(points, np.float64) / voxel_size).astype(np.int64)


def voxel_down_sample(points, colors, voxel_size, normals=None):
    """First-occurrence voxel order; per-voxel sequential-sum mean (np.mean over a python list of
    rows reduces along axis 0 row by row)."""
    points = np.asarray(points, np.float64)
    vi = voxel_indices(points, voxel_size)
    _, first, inv = np.unique(vi, axis=0, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    order = np.argsort(first, kind="stable")          # unique-id -> rank by first occurrence
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    vid = rank[inv]
    nv = len(order)
    cnt = np.bincount(vid, minlength=nv).astype(np.float64)

    def seqmean(a):
        acc = np.zeros((nv, a.shape[1]))
        np.add.at(acc, vid, np.asarray(a, np.float64))
        return acc / cnt[:, None]

    p = seqmean(points)
    c = seqmean(colors) if colors is not None else None
    n = None
    if normals is not None:
        n = seqmean(normals)
        n = n / np.linalg.norm(n, axis=1, keepdims=True)
    return p, c, n, vid


# a4: utils/depth_utils.py:92-116
def transform_pointcloud(points, pose):
    from scipy.spatial.transform import Rotation
    pose = np.array(pose, dtype=np.float64)
    t = pose[:3]
    q = pose[3:] / np.linalg.norm(pose[3:])
    R = Rotation.from_quat(q).as_matrix()
    return (R @ np.asarray(points).T).T + t


# a6: object_memory/object_memory.py:922-936
def closest_similarities(detected_embs, memory_embs_list):
    det = np.array(detected_embs, dtype=np.float32)
    det = det / np.linalg.norm(det, axis=-1, keepdims=True)
    mem = [np.array([e / np.linalg.norm(e) for e in np.asarray(m, np.float32)]) for m in memory_embs_list]
    out = np.zeros((len(det), len(mem)), np.float32)
    for i, d in enumerate(det):
        for j, m in enumerate(mem):
            out[i, j] = np.max(np.dot(m, d))
    return out


# a7: utils/similarity_volume.py:13-18,102-164,213-270
def simvolume_subvolumes(S):
    """-> (chosen list, list of masked fp16 sub-volumes), vectorised; equals fast_construct_volume."""
    S = np.asarray(S)
    n, m = S.shape
    aug = np.ones((n, m + 1), np.float16)
    aug[:, :-1] = S
    if n == 1:
        v = aug[0].copy()
        v[-1] = -np.inf
        return [[0]], [v]
    s = min(n, 3)
    chosen = list(itertools.combinations(range(n), s))
    ar = np.arange(m + 1)
    subs = []
    for ch in chosen:
        a = aug[list(ch)]
        vol = np.multiply.outer(a[0], a[1]).astype(np.float16)
        if s == 3:
            vol = (vol[:, :, None] * a[2][None, None, :]).astype(np.float16)
            I, J, K = np.meshgrid(ar, ar, ar, indexing="ij", sparse=True)
            ok = ((I != J) | (I == m)) & ((I != K) | (I == m)) & ((J != K) | (J == m))
            ok = ok & (I != m)
        else:
            I, J = np.meshgrid(ar, ar, indexing="ij", sparse=True)
            ok = ((I != J) | (I == m)) & (I != m)
        if m < s:                  # no permutation of s distinct memory objects exists -> mask is all -inf
            ok = np.zeros_like(ok)
        vol = np.where(ok, vol, np.float16(-np.inf)).astype(np.float16)
        vol[np.isnan(vol)] = -np.inf
        subs.append(vol)
    return chosen, subs


def topk_first_index(vol, k):
    """k repeated argmax picks with first-flat-index tie-break -> (flat idx [k], vals [k])."""
    flat = vol.reshape(-1)
    finite = np.nonzero(flat > -np.inf)[0]
    vals = flat[finite].astype(np.float32)
    if len(finite) > k:
        kth = np.partition(vals, len(vals) - k)[len(vals) - k]
        sel = vals >= kth
        finite, vals = finite[sel], vals[sel]
    order = np.lexsort((finite, -vals))[:k]
    idx = list(finite[order])
    v = list(flat[finite[order]])
    while len(idx) < k:                       # everything is -inf: argmax keeps returning 0
        idx.append(0)
        v.append(np.float16(-np.inf))
    return np.array(idx, np.int64), np.array(v, np.float16)


def simvolume_assignments(S, num_per_length=4):
    S = np.asarray(S)
    n, m = S.shape
    chosen, subs = simvolume_subvolumes(S)
    k = num_per_length * n * 4
    assns, scored = [], []
    for ch, vol in zip(chosen, subs):
        idx, vals = topk_first_index(vol, k)
        for fi, v in zip(idx, vals):
            coords = np.unravel_index(int(fi), vol.shape)
            filt = [[int(i), int(c)] for i, c in zip(ch, coords) if c != m]
            if not filt:
                continue
            if filt not in assns:
                assns.append(filt)
                scored.append((filt, v))
    out = []
    for L in range(1, n + 1):
        cl = [f for f in scored if len(f[0]) == L]
        cl = sorted(cl, key=lambda x: x[-1], reverse=True)[:max(1, L)]
        out += [c[0] for c in cl]
    res = []
    for a in out:
        if a not in res:
            res.append(a)
    return res


# utils/fpfh_register.py:67-82 (get_SVD_transform) -- golden known answers live in test.py
def get_SVD_transform(p, q):
    p, q = np.asarray(p, np.float64), np.asarray(q, np.float64)
    up, uq = p.mean(0), q.mean(0)
    W = (q - uq).T @ (p - up)
    u, _, vh = np.linalg.svd(W)
    M = np.diag([1, 1, np.linalg.det(u) * np.linalg.det(vh)])
    R = u @ M @ vh
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = uq - R @ up
    return T
