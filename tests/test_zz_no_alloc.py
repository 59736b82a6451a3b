"""ibl_nn_score_static (include/ibl.h): the allocation-free variant of ibl_nn_score, called through the C ABI (ctypes) and
compared with ibl_nn_score on the same inputs.  Same kernels and reduction order, so the bar is the one test_kernels_gpu.py
holds for this row: correspondence indices and inlier counts bit-exact, squared distances and their sum EQUAL (== on the
float64 bits, tighter than the 1e-12 relative bar test_nn_score uses against the oracle).

Two calls in a row on DIFFERENT inputs -- the second smaller than the first, so the static scratch array still holds the
first call's partials beyond the second call's range -- each checked against the allocating entry point; then the size
limit.  No counterpart in the reference.

This file is named to be collected LAST (pytest -x stops at the first failure) and passes both alone in a
fresh process and after the other GPU tests."""
import ctypes

import numpy as np
import pytest
import torch

from ibl_testutil import dev, random_rigid, surface_cloud

pytestmark = pytest.mark.gpu

MAX_POINTS = 1 << 20      # IBL_NN_SCORE_STATIC_MAX_POINTS
IBL_E_RANGE = -4


@pytest.fixture(scope="module")
def K(cuda):
    from instance_based_loc_b200 import kernels
    return kernels


def _T16(T):
    return (ctypes.c_double * 16)(*np.asarray(T, np.float64).reshape(16).tolist())


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(K, L, g, src, T, max_d):
    from instance_based_loc_b200._lib import ScoreT
    n = src.shape[0]
    nn = torch.full((n,), -7, dtype=torch.int32, device=src.device)
    d2 = torch.full((n,), -7.0, dtype=torch.float64, device=src.device)
    res = ScoreT()
    rc = L.ibl_nn_score(g.h, _ptr(src), n, _T16(T), max_d, _ptr(nn), _ptr(d2), ctypes.byref(res), _stream())
    assert rc == 0, L.ibl_last_error_string()
    return int(res.n_inliers), float(res.err2), nn.cpu().numpy(), d2.cpu().numpy()


def _static(L, g, src, T, max_d, out):
    """outputs are caller-owned and pre-filled with garbage; the call itself allocates nothing and does not synchronise"""
    n = src.shape[0]
    nn = torch.full((n,), -7, dtype=torch.int32, device=src.device)
    d2 = torch.full((n,), -7.0, dtype=torch.float64, device=src.device)
    out.fill_(-1)
    rc = L.ibl_nn_score_static(g.h, _ptr(src), n, _T16(T), max_d, _ptr(nn), _ptr(d2), _ptr(out), _stream())
    assert rc == 0, L.ibl_last_error_string()
    torch.cuda.synchronize()
    raw = out.cpu().numpy()
    return int(raw[0]), float(raw[1:2].view(np.float64)[0]), nn.cpu().numpy(), d2.cpu().numpy()


def _case(n_src, seed):
    src = (random_rigid(seed, 0.05, 0.03) @ np.c_[surface_cloud(n_src, seed + 1)[0], np.ones(n_src)].T).T[:, :3]
    src = np.concatenate([src, src[:200] + 0.06, src[:40] + 3.0])          # far and very far queries: no neighbour
    return dev(src), random_rigid(seed + 2, 0.02, 0.01)


def test_static_entry_point_equals_nn_score_twice_in_a_row(K):
    from instance_based_loc_b200._lib import lib
    L = lib()
    tgt, _ = surface_cloud(9000, 21)
    max_d = 0.05
    g = K.Grid(dev(tgt), max_d / 4, ring=8)
    out = torch.empty(2, dtype=torch.int64, device="cuda")                 # one ibl_score_t {int64, double}
    seen = []
    for n_src, seed in ((20000, 31), (3000, 41)):                          # 2nd call uses fewer CTAs than the 1st left behind
        src, T = _case(n_src, seed)
        want = _existing(K, L, g, src, T, max_d)
        got = _static(L, g, src, T, max_d, out)
        assert got[0] == want[0] and got[0] > 0                            # inlier count, bit-exact
        assert np.array_equal(got[2], want[2])                             # correspondence indices, bit-exact
        assert (got[2] < 0).sum() >= 40                                    # the far queries found nothing
        assert np.array_equal(got[3].view(np.int64), want[3].view(np.int64))   # squared distances, same bits
        assert np.float64(got[1]).view(np.int64) == np.float64(want[1]).view(np.int64)   # their sum, same bits
        seen.append(got[:2])
    assert seen[0] != seen[1]                                              # the two calls really were different problems
    # the same problem again, without the optional outputs and with the identity transform spelled as NULL
    src, _ = _case(3000, 41)
    res = torch.empty(2, dtype=torch.int64, device="cuda")
    K.nn_score_static(g, src, max_d, res)
    a, b, _, _ = K.nn_score(g, src, max_d)
    raw = res.cpu().numpy()
    assert int(raw[0]) == a and raw[1:2].view(np.float64)[0] == b


def test_static_entry_point_edges_and_limit(K):
    from instance_based_loc_b200._lib import lib
    L = lib()
    tgt, _ = surface_cloud(2000, 5)
    g = K.Grid(dev(tgt), 0.01, ring=8)
    out = torch.full((2,), -1, dtype=torch.int64, device="cuda")
    src = dev(tgt[:10])
    # n = 0: the result is {0, 0.0}
    assert L.ibl_nn_score_static(g.h, _ptr(src), 0, None, 0.05, None, None, _ptr(out), _stream()) == 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0, 0]
    # exact hits: every source point is a target point
    assert L.ibl_nn_score_static(g.h, _ptr(src), 10, None, 0.05, None, None, _ptr(out), _stream()) == 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [10, 0]
    # exactly the limit: every slot of the static array is used
    big = dev(np.tile(tgt[:1024], (MAX_POINTS // 1024, 1)))
    assert big.shape[0] == MAX_POINTS
    assert L.ibl_nn_score_static(g.h, _ptr(big), MAX_POINTS, None, 0.05, None, None, _ptr(out), _stream()) == 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [MAX_POINTS, 0]
    a, b, _, _ = K.nn_score(g, big, 0.05)
    assert (a, b) == (MAX_POINTS, 0.0)
    assert L.ibl_nn_score_static(g.h, _ptr(src), 10, None, 0.05, None, None, _ptr(out), _stream()) == 0
    # beyond the static scratch: refused before anything is launched, with a message naming the limit
    rc = L.ibl_nn_score_static(g.h, _ptr(src), MAX_POINTS + 1, None, 0.05, None, None, _ptr(out), _stream())
    assert rc == IBL_E_RANGE
    msg = L.ibl_last_error_string().decode()
    assert "ibl_nn_score_static" in msg and str(MAX_POINTS) in msg
    # the result buffer is required
    assert L.ibl_nn_score_static(g.h, _ptr(src), 10, None, 0.05, None, None, None, _stream()) < 0
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [10, 0]                                   # and a refused call wrote nothing
