"""GPU probe of the tcgen05 feature matcher (csrc/match_tc.cu).  Stages (each run as its own process under `timeout`):
  scores  : dumped tensor-core scores vs an f64 CPU reference of a32.b32 - |b32|^2/2, for both descriptor conventions
  nn      : ibl_feature_nn vs a brute-force f64 argmin (chunked torch), incl. exact duplicate rows (ties -> lowest index)
  time    : CUDA-event timing of ibl_feature_nn at a registration-sized problem (impl chosen by IBL_FEAT_IMPL)
Usage: python tools/tc_probe.py scores|nn|time"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from instance_based_loc_b200 import kernels as K  # noqa: E402


def fpfh_like(n, rng, dup=0):
    """Rows shaped like FPFH: three 11-bin histograms, each summing to ~200, many near-empty bins."""
    f = rng.gamma(0.4, 1.0, (n, 3, 11))
    f = 200.0 * f / f.sum(-1, keepdims=True)
    f = f.reshape(n, 33)
    if dup:
        src = rng.integers(0, n, dup)
        dst = rng.integers(0, n, dup)
        f[dst] = f[src]
    return np.ascontiguousarray(f)


def stage_scores():
    rng = np.random.default_rng(1)
    fa, fb = fpfh_like(200, rng), fpfh_like(300, rng)
    A, B = torch.from_numpy(fa).cuda(), torch.from_numpy(fb).cuda()
    for swap in (0, 1):
        S, mu, sc = K.feature_scores_debug(A, B, swap)
        S, mu = S.cpu().numpy().astype(np.float64), mu.cpu().numpy()
        a32 = ((fa - mu) * sc).astype(np.float32).astype(np.float64)
        b32 = ((fb - mu) * sc).astype(np.float32).astype(np.float64)
        ref = a32 @ b32.T - 0.5 * (b32 * b32).sum(1)[None, :]
        scale = (a32 * a32).sum(1)[:, None] + (b32 * b32).sum(1)[None, :]
        err = np.abs(S - ref) / scale
        print(f"scores swap={swap}: scale {sc}  max rel err {err.max():.3e}  mean {err.mean():.3e}  finite={np.isfinite(S).all()}  "
              f"mean |a-mu|^2 / |a|^2 = {(a32 * a32).sum() / (sc * sc * (fa * fa).sum()):.3f}", flush=True)


def real_fpfh():
    """Real FPFH rows: the bench's synthetic scene, a detected-cloud-sized union vs a memory union."""
    from instance_based_loc_b200 import o3d, synth
    from instance_based_loc_b200.utils.fpfh_register import downsample_and_compute_fpfh
    scene = synth.make_scene(m=40, seed=1002, D=768)
    out = []
    for objs, spacing in (((0, 1, 2), 0.012), ((0, 1, 2), 0.01)):
        pts = np.concatenate([synth.memory_cloud(scene, j, spacing)[0] for j in objs])
        cols = np.concatenate([synth.memory_cloud(scene, j, spacing)[1] for j in objs])
        pc = o3d.geometry.PointCloud()
        pc.points = o3d.utility.Vector3dVector(pts - pts.mean(0))
        pc.colors = o3d.utility.Vector3dVector(cols)
        _, f = downsample_and_compute_fpfh(pc, 0.05)
        out.append(f._dev)
    return out


def stage_real():
    A, B = real_fpfh()
    na, nb = A.shape[0], B.shape[0]
    K.feature_nn_stats(True)
    got = K.feature_nn(A, B).cpu().numpy()
    cand, ev = K.feature_nn_stats(True)
    want, dmin = brute_nn(A.cpu().numpy(), B.cpu().numpy())
    print(f"real na={na} nb={nb}: index match {int((got == want).sum())}/{na}; band candidates/query {cand / na:.2f}, "
          f"exact evaluations/query {ev / na:.2f}", flush=True)
    for _ in range(3):
        K.feature_nn(A, B)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        K.feature_nn(A, B)
    e1.record()
    torch.cuda.synchronize()
    print(f"real time: {e0.elapsed_time(e1) / 10:.3f} ms per call", flush=True)


def brute_nn(fa, fb, chunk=256):
    A, B = torch.from_numpy(fa).cuda(), torch.from_numpy(fb).cuda()
    idx = torch.empty(len(fa), dtype=torch.int64, device="cuda")
    dmin = torch.empty(len(fa), dtype=torch.float64, device="cuda")
    for i in range(0, len(fa), chunk):
        d = ((A[i:i + chunk, None, :] - B[None, :, :]) ** 2).sum(-1)
        m = d.min(1)
        dmin[i:i + chunk] = m.values
        # lowest index among exact minima
        idx[i:i + chunk] = (d == m.values[:, None]).to(torch.uint8).argmax(1)
    return idx.cpu().numpy(), dmin.cpu().numpy()


def stage_nn():
    rng = np.random.default_rng(2)
    for na, nb, dup in ((1, 1, 0), (5, 130, 0), (300, 257, 40), (5000, 7000, 800), (20000, 18000, 3000)):
        fa = fpfh_like(na, rng)
        fb = fpfh_like(nb, rng, dup)
        if dup:                       # also plant exact copies of target rows among the queries
            k = min(dup, na)
            fa[rng.integers(0, na, k)] = fb[rng.integers(0, nb, k)]
        got = K.feature_nn(torch.from_numpy(fa).cuda(), torch.from_numpy(fb).cuda()).cpu().numpy()
        want, dmin = brute_nn(fa, fb)
        same = int((got == want).sum())
        dgot = ((fa - fb[got]) ** 2).sum(1)
        worse = int((dgot > dmin * (1 + 1e-12) + 1e-9).sum())
        print(f"nn na={na} nb={nb} dup={dup}: index match {same}/{na}  strictly-worse distance {worse}  "
              f"higher-index-on-tie {int(((got != want) & (dgot <= dmin * (1 + 1e-12) + 1e-9) & (got > want)).sum())}", flush=True)


def stage_time():
    rng = np.random.default_rng(3)
    impl = os.environ.get("IBL_FEAT_IMPL", "tcgen05")
    for na, nb in ((18600, 18600), (50000, 50000)):
        A, B = torch.from_numpy(fpfh_like(na, rng)).cuda(), torch.from_numpy(fpfh_like(nb, rng)).cuda()
        for _ in range(3):
            K.feature_nn(A, B)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            K.feature_nn(A, B)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"time impl={impl} na={na} nb={nb}: {ms:.3f} ms  ({2 * 33 * na * nb / ms / 1e9:.1f} TFLOP/s of the 33-d distance matrix)", flush=True)


if __name__ == "__main__":
    {"scores": stage_scores, "nn": stage_nn, "time": stage_time, "real": stage_real}[sys.argv[1]]()
