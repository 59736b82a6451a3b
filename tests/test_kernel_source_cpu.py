"""CPU check of CUDA kernels' ARITHMETIC without a GPU.  A thread-per-element kernel with no inter-thread communication
compiles as host C++ from the very same source text once the CUDA qualifiers are defined away and the _rn intrinsics are
mapped to plain IEEE operations (g++ -ffp-contract=off, which is what -fmad=false gives on the device); one "thread" at a
time then walks the elements.

 * ``raycast_kernel`` (csrc/raycast.cu) against the tensor-op ray caster of instance_based_loc_b200/synth.py
   (GPU run of the same kernel: tests/test_z_memory_build_gpu.py::test_raycast_kernel_vs_tensor_ops);
 * ``normals_kernel`` with ``fast_eigen3x3`` (csrc/features.cu) against the oracle's orc_normals, bit for bit
   (GPU run: tests/test_kernels_gpu.py::test_normals, test_normals_degenerate);
 * ``ransac_check_kernel`` (csrc/ransac.cu with rigid.cuh's Jacobi-Kabsch and ibl_common.cuh's ibl_xform) against the
   oracle's orc_ransac_check: pass mask and transforms bit for bit (GPU run: test_ransac_check);
 * ``color_gradient_kernel`` (csrc/icp.cu with rigid.cuh's LDLT solve) against orc_color_gradient, bit for bit
   (GPU run: test_colored_icp);
 * the four voxel kernels of csrc/voxel.cu (keys, segment heads, first occurrences, sequential means; the device sort and
   select replaced by std::stable_sort / a loop) against the REFERENCE's own voxel_down_sample_with_colors golden
   outputs, bit for bit (GPU run: test_voxel_golden, test_voxel_vs_port).

This is a test of the source, not a fallback: the host builds live in a temporary directory and no product module can
reach them."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PRELUDE = r'''
#include <cmath>
#include <cstdint>
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(x)
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
struct Dim1 { long long x; };
static Dim1 blockIdx{0}, threadIdx{0}, gridDim{1}, blockDim{1};
'''
DRIVER = r'''
extern "C" void host_raycast(const double *T, int H, int W, double f, const int32_t *kind, const double *half, const double *R,
                             const double *t, int m, double room, double height, double *ot, int32_t *oi, double *on) {
    RaycastCam cam;
    for (int i = 0; i < 3; ++i) { for (int j = 0; j < 3; ++j) cam.R[3 * i + j] = T[4 * i + j]; cam.o[i] = T[4 * i + 3]; }
    const double lw = W > 1 ? (double)W / (double)(W - 1) : 0.0, lh = H > 1 ? (double)H / (double)(H - 1) : 0.0;
    raycast_kernel(cam, H, W, f, lw, lh, kind, half, R, t, m, room, height, ot, oi, on);
}
'''


@pytest.fixture(scope="module")
def host_kernel(tmp_path_factory):
    src = open(os.path.join(ROOT, "instance_based_loc_b200", "csrc", "raycast.cu")).read()
    body = src[src.index("struct RaycastCam"):src.index("IBL_API int ibl_raycast_scene")]
    d = tmp_path_factory.mktemp("rc_host")
    cpp, so = str(d / "rc_host.cpp"), str(d / "librc_host.so")
    open(cpp, "w").write(PRELUDE + body + DRIVER)
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp])
    return ctypes.CDLL(so)


@pytest.mark.parametrize("m,seed,H,W,f", [(10, 3, 60, 80, 40.0), (40, 1002, 150, 150, 75.0), (25, 7, 97, 131, 60.0), (3, 5, 1, 1, 1.0)])
def test_raycast_kernel_source_vs_tensor_ops(host_kernel, m, seed, H, W, f):
    from instance_based_loc_b200 import synth
    sc = synth.make_scene(m=m, seed=seed, D=16)
    T = np.ascontiguousarray(synth.camera_pose(sc, np.random.default_rng(seed)))
    b0, i0, n0 = [x.numpy() for x in synth.render(sc, T, H=H, W=W, f=f, want_geometry=True)]
    P = H * W
    ot, oi, on = np.empty(P), np.empty(P, np.int32), np.empty((P, 3))
    kind, half, R, t = sc.kind.astype(np.int32), np.ascontiguousarray(sc.half), np.ascontiguousarray(sc.R), np.ascontiguousarray(sc.t)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    host_kernel.host_raycast(p(T), H, W, ctypes.c_double(f), p(kind), p(half), p(R), p(t), sc.n_prim, ctypes.c_double(sc.room),
                             ctypes.c_double(sc.height), p(ot), p(oi), p(on))
    same = i0 == oi
    assert same.mean() >= 0.9999
    assert np.array_equal(np.isinf(b0[same]), np.isinf(ot[same]))
    fin = same & np.isfinite(b0)
    assert fin.sum() == 0 or np.abs(b0 - ot)[fin].max() <= 1e-9
    assert np.abs(n0 - on)[same].max() <= 1e-9


NORMALS_DRIVER = r'''
extern "C" void host_normals(const double *pts, long long n, const int32_t *idx, const int32_t *cnt, int max_nn, double *out) {
    for (long long i = 0; i < n; ++i) { blockIdx.x = i; normals_kernel(pts, n, idx, cnt, max_nn, out); }   // one thread per launch
}
'''


@pytest.fixture(scope="module")
def host_normals(tmp_path_factory):
    src = open(os.path.join(ROOT, "instance_based_loc_b200", "csrc", "features.cu")).read()
    body = src[src.index("__device__ __forceinline__ void cross3"):src.index("IBL_API int ibl_normals")]
    d = tmp_path_factory.mktemp("nrm_host")
    cpp, so = str(d / "nrm_host.cpp"), str(d / "libnrm_host.so")
    open(cpp, "w").write(PRELUDE + "#include <cstddef>\n" + body + NORMALS_DRIVER)
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp])
    return ctypes.CDLL(so)


def test_normals_kernel_source_vs_oracle(host_normals):
    """FastEigen3x3 is branchy (largest cross product, degenerate covariances); the kernel source and the oracle's C must
    take the same branch everywhere: compared bit for bit on a surface cloud, a plane, a line and tiny neighbourhoods."""
    from ibl_testutil import surface_cloud
    from oracle import o3d_core as oc
    rng = np.random.default_rng(9)
    plane = np.c_[rng.uniform(-1, 1, (400, 2)), np.zeros(400)]
    line = np.c_[np.linspace(0, 1, 200), np.zeros(200), np.zeros(200)]
    sparse = rng.uniform(-5, 5, (50, 3))                                   # < 3 neighbours: identity covariance branch
    for pts, radius, k in ((surface_cloud(3000, 4)[0], 0.1, 30), (plane, 0.3, 30), (line, 0.05, 30), (sparse, 0.1, 30),
                           (np.round(surface_cloud(1500, 5)[0], 2), 0.1, 12)):    # snapped to a lattice: exact ties
        pts = np.ascontiguousarray(pts, np.float64)
        idx, _, cnt = oc.knn_hybrid(pts, pts, radius, k)
        idx, cnt = np.ascontiguousarray(idx, np.int32), np.ascontiguousarray(cnt, np.int32)
        want = np.zeros((len(pts), 3))
        oc.lib().orc_normals(pts.ctypes.data_as(ctypes.c_void_p), len(pts), idx.ctypes.data_as(ctypes.c_void_p),
                             cnt.ctypes.data_as(ctypes.c_void_p), k, want.ctypes.data_as(ctypes.c_void_p))
        got = np.full((len(pts), 3), np.nan)
        p = lambda a: a.ctypes.data_as(ctypes.c_void_p)      # noqa: E731
        host_normals.host_normals(p(pts), ctypes.c_longlong(len(pts)), p(idx), p(cnt), k, p(got))
        assert np.array_equal(got, want)


RANSAC_DRIVER = r'''
extern "C" void host_ransac_check(const double *src, const double *tgt, const int32_t *corres, long long n_corres,
                                  const int32_t *samples, long long H, double edge_thr, double dist_thr, int use_checkers,
                                  uint8_t *pass, double *T_out) {
    for (long long k = 0; k < H; ++k) {
        blockIdx.x = k;
        ransac_check_kernel(src, tgt, corres, n_corres, samples, 0ull, 0ll, H, edge_thr, dist_thr, use_checkers, pass, T_out);
    }
}
'''


@pytest.fixture(scope="module")
def host_ransac(tmp_path_factory):
    csrc = os.path.join(ROOT, "instance_based_loc_b200", "csrc")
    common = open(os.path.join(csrc, "ibl_common.cuh")).read()
    xform = common[common.index("// Transform one point with the Open3D/Eigen"):common.index("// Nearest neighbour of q within max_d")]
    src = open(os.path.join(csrc, "ransac.cu")).read()
    body = src[src.index("__device__ __forceinline__ void gather_sample"):src.index("IBL_API int ibl_ransac_check")]
    d = tmp_path_factory.mktemp("ransac_host")
    cpp, so = str(d / "ransac_host.cpp"), str(d / "libransac_host.so")
    stub = ("#include <cstddef>\n#include <cstdint>\n"
            f'#include "{os.path.join(csrc, "rigid.cuh")}"\n')
    open(cpp, "w").write(PRELUDE + stub + xform + body + RANSAC_DRIVER)
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp])
    return ctypes.CDLL(so)


def test_ransac_check_kernel_source_vs_oracle(host_ransac):
    """Edge-length test, Jacobi-Kabsch and distance test of one hypothesis per thread: pass mask and the 4x4 transform of
    every passing hypothesis bit-equal to the oracle's orc_ransac_check on the same host-seeded triplets."""
    from ibl_testutil import random_rigid, surface_cloud
    from oracle import o3d_core as oc
    rng = np.random.default_rng(21)
    src = surface_cloud(800, 6)[0]
    Tgt = random_rigid(3)
    tgt = src @ Tgt[:3, :3].T + Tgt[:3, 3] + rng.normal(0, 0.002, src.shape)
    C = 500
    corres = np.stack([rng.integers(0, len(src), C), rng.integers(0, len(tgt), C)], 1).astype(np.int32)
    corres[: C // 2, 1] = corres[: C // 2, 0]                          # half of them true matches
    samples = rng.integers(0, C, (4000, 3)).astype(np.int32)
    samples[:50, 1] = samples[:50, 0]                                   # degenerate triplets (a repeated correspondence)
    for edge_thr, dist_thr, use in ((0.9, 0.075, 1), (0.9, 0.02, 1), (0.0, 1e9, 1)):
        want_ok, want_T = oc.ransac_check(src, tgt, corres, samples, edge_thr, dist_thr, want_T=True)
        ok = np.zeros(len(samples), np.uint8)
        Ts = np.zeros((len(samples), 16))
        p = lambda a: a.ctypes.data_as(ctypes.c_void_p)      # noqa: E731
        s64, t64 = np.ascontiguousarray(src, np.float64), np.ascontiguousarray(tgt, np.float64)
        host_ransac.host_ransac_check(p(s64), p(t64), p(corres), ctypes.c_longlong(C), p(samples),
                                      ctypes.c_longlong(len(samples)), ctypes.c_double(edge_thr), ctypes.c_double(dist_thr),
                                      use, p(ok), p(Ts))
        assert np.array_equal(ok.astype(bool), want_ok)
        assert want_ok.sum() > 0
        assert np.array_equal(Ts.reshape(-1, 4, 4)[want_ok], want_T[want_ok])


GRAD_DRIVER = r'''
extern "C" void host_color_gradient(const double *pts, const double *nrm, const double *col, long long n, const int32_t *idx,
                                    const int32_t *cnt, int max_nn, double *grad) {
    for (long long k = 0; k < n; ++k) { blockIdx.x = k; color_gradient_kernel(pts, nrm, col, n, idx, cnt, max_nn, grad); }
}
'''


@pytest.fixture(scope="module")
def host_grad(tmp_path_factory):
    csrc = os.path.join(ROOT, "instance_based_loc_b200", "csrc")
    src = open(os.path.join(csrc, "icp.cu")).read()
    body = src[src.index("__global__ void __launch_bounds__(128) color_gradient_kernel"):src.index("IBL_API int ibl_color_gradient")]
    d = tmp_path_factory.mktemp("grad_host")
    cpp, so = str(d / "grad_host.cpp"), str(d / "libgrad_host.so")
    open(cpp, "w").write(PRELUDE + f'#include <cstddef>\n#include "{os.path.join(csrc, "rigid.cuh")}"\n' + body + GRAD_DRIVER)
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp])
    return ctypes.CDLL(so)


def test_color_gradient_kernel_source_vs_oracle(host_grad):
    """Coloured-ICP target colour gradients (A.8): 3x3 normal equations per point in neighbour-list order + LDLT solve,
    bit-equal to the oracle's orc_color_gradient (points with fewer than 4 neighbours keep a zero gradient)."""
    from ibl_testutil import surface_cloud
    from oracle import o3d_core as oc
    pts, col = surface_cloud(2500, 8)
    pts = np.ascontiguousarray(np.concatenate([pts, [[9.0, 9.0, 9.0]]]))          # an isolated point: nn < 4 branch
    col = np.ascontiguousarray(np.concatenate([col, [[0.5, 0.5, 0.5]]]))
    nrm = np.ascontiguousarray(oc.estimate_normals(pts, 0.1, 30))
    max_d = 0.075
    want = oc.color_gradient(pts, nrm, col, max_d)
    idx, _, cnt = oc.knn_hybrid(pts, pts, max_d * 2.0, 30)
    idx, cnt = np.ascontiguousarray(idx, np.int32), np.ascontiguousarray(cnt, np.int32)
    got = np.full_like(want, np.nan)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)          # noqa: E731
    host_grad.host_color_gradient(p(pts), p(nrm), p(col), ctypes.c_longlong(len(pts)), p(idx), p(cnt), 30, p(got))
    assert np.array_equal(got, want) and np.abs(want).max() > 0 and np.array_equal(want[-1], np.zeros(3))


VOXEL_DRIVER = r'''
#include <algorithm>
#include <numeric>
#include <vector>
// the host side of ibl_voxel_downsample (csrc/voxel.cu) with std::stable_sort / a loop in place of the cub device sort and
// select; the four kernels are the library's source text, one element per "launch"
extern "C" long long host_voxel(const double *pts, const double *cols, const double *nrm, long long n, double voxel,
                                double *out_pts, double *out_cols, double *out_nrm, int32_t *vox_id, int64_t *out_keys) {
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (long long i = 0; i < n; ++i)
        for (int a = 0; a < 3; ++a) { mn[a] = fmin(mn[a], pts[3 * i + a]); mx[a] = fmax(mx[a], pts[3 * i + a]); }
    VoxRange vr;
    unsigned long long r[3];
    for (int a = 0; a < 3; ++a) {
        double lo = floor(mn[a] / voxel), hi = floor(mx[a] / voxel);
        vr.kmin[a] = (long long)lo;
        r[a] = (unsigned long long)((long long)hi - (long long)lo + 1);
    }
    vr.r0 = r[0]; vr.r1 = r[1];
    std::vector<unsigned long long> key(n), key2(n);
    std::vector<int> idx(n), idx2(n);
    std::vector<uint8_t> flag(n);
    blockDim.x = 1;
    for (long long i = 0; i < n; ++i) { blockIdx.x = i; vox_key_kernel(pts, n, voxel, vr, key.data(), idx.data(), out_keys); }
    std::vector<int> perm(n);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return key[a] < key[b]; });      // radix sort is stable
    for (long long i = 0; i < n; ++i) { key2[i] = key[perm[i]]; idx2[i] = idx[perm[i]]; }
    for (long long i = 0; i < n; ++i) { blockIdx.x = i; vox_head_kernel(key2.data(), n, flag.data()); }
    std::vector<int> head_pos;
    for (long long i = 0; i < n; ++i) if (flag[i]) head_pos.push_back((int)i);                       // DeviceSelect::Flagged
    const int nseg = (int)head_pos.size();
    std::vector<int> first_idx(nseg), seg(nseg), seg2(nseg);
    for (int s = 0; s < nseg; ++s) { blockIdx.x = s; vox_first_kernel(head_pos.data(), nseg, idx2.data(), first_idx.data(), seg.data()); }
    std::vector<int> p2(nseg);
    std::iota(p2.begin(), p2.end(), 0);
    std::stable_sort(p2.begin(), p2.end(), [&](int a, int b) { return first_idx[a] < first_idx[b]; });
    for (int s = 0; s < nseg; ++s) seg2[s] = seg[p2[s]];
    for (int v = 0; v < nseg; ++v) {
        blockIdx.x = v;
        vox_mean_kernel(seg2.data(), head_pos.data(), nseg, n, idx2.data(), pts, cols, nrm, out_pts, cols ? out_cols : nullptr,
                        nrm ? out_nrm : nullptr, vox_id);
    }
    blockIdx.x = 0;
    return nseg;
}
'''


@pytest.fixture(scope="module")
def host_voxel(tmp_path_factory):
    src = open(os.path.join(ROOT, "instance_based_loc_b200", "csrc", "voxel.cu")).read()
    body = src[src.index("struct VoxRange"):src.index("IBL_API int64_t ibl_voxel_downsample")]
    d = tmp_path_factory.mktemp("vox_host")
    cpp, so = str(d / "vox_host.cpp"), str(d / "libvox_host.so")
    pre = PRELUDE + "#include <cstddef>\nstatic inline double __ddiv_rn(double a, double b) { return a / b; }\n"
    open(cpp, "w").write(pre + body + VOXEL_DRIVER)
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-o", so, cpp])
    L = ctypes.CDLL(so)
    L.host_voxel.restype = ctypes.c_longlong
    return L


def _voxel(L, pts, cols, nrm, voxel):
    p = lambda a: None if a is None else a.ctypes.data_as(ctypes.c_void_p)       # noqa: E731
    pts = np.ascontiguousarray(pts, np.float64)
    cols = None if cols is None else np.ascontiguousarray(cols, np.float64)
    nrm = None if nrm is None else np.ascontiguousarray(nrm, np.float64)
    n = len(pts)
    op, oc, on = np.empty_like(pts), np.empty_like(pts), np.empty_like(pts)
    vid, keys = np.empty(n, np.int32), np.empty((n, 3), np.int64)
    nv = L.host_voxel(p(pts), p(cols), p(nrm), ctypes.c_longlong(n), ctypes.c_double(voxel), p(op), p(oc), p(on), p(vid), p(keys))
    return op[:nv], oc[:nv], on[:nv], vid, keys


def test_voxel_kernels_source_vs_reference_golden(host_voxel, golden):
    """Row a3: voxel keys, first-occurrence order and the sequential means of the kernels' source text against the
    REFERENCE's own voxel_down_sample_with_colors outputs (tests/golden/make_golden.py), bit for bit; then normals and
    voxel ids against the numpy port on a second cloud."""
    from ibl_testutil import surface_cloud
    from oracle import ref_port
    g = golden
    op, oc, _, _, keys = _voxel(host_voxel, g["vox_pts"], g["vox_cols"], None, float(g["vox_size"]))
    assert np.array_equal(keys, g["vox_keys"])
    assert np.array_equal(op, g["vox_out_pts"]) and np.array_equal(oc, g["vox_out_cols"])
    p, c = surface_cloud(4000, 31)
    nr = np.random.default_rng(31).normal(size=p.shape)
    op, oc, on, vid, _ = _voxel(host_voxel, p, c, nr, 0.03)
    rp, rc, rn, rvid = ref_port.voxel_down_sample(p, c, 0.03, nr)
    assert np.array_equal(op, rp) and np.array_equal(oc, rc) and np.array_equal(vid, rvid)
    assert np.allclose(on, rn, rtol=0, atol=1e-15)
