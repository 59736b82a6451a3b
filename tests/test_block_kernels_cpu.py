"""CPU check of BLOCK-LEVEL CUDA kernels without a GPU: the two back-projection kernels of csrc/cloud.cu (row a1/a2) run
from their own source text as a 256-thread block of host threads; ``cub::BlockReduce`` / ``cub::BlockScan`` are replaced
by a barrier-based stand-in with the same interface (sum / exclusive prefix sum over the block, which is all the kernels
use them for), the device-wide scan between the two kernels by a host loop.  The result is compared bit for bit with the
golden vectors that tests/golden/make_golden.py produced by running the REFERENCE's own numpy
(utils/depth_utils.py:46-90,176-206) -- so the float32 / float64 rounding sequence of the kernel source is pinned against
the reference without a GPU.  GPU run of the same kernels: tests/test_kernels_gpu.py::test_backproject_*.
``seqsum3_kernel`` (the column means of row a4 / a12: double-buffered shared-memory tiles, three serial add chains, __syncthreads)
is run the same way and compared bit for bit with ``np.mean(points, axis=0)`` itself -- the call the reference makes
(object_memory/object_memory.py:1030-1031).
A test of the source text, not a fallback: the build lives in a temporary directory."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "instance_based_loc_b200", "csrc")

PRELUDE = r'''
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <pthread.h>
#include <algorithm>
using std::min;
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(x)
#define __shared__ static
struct float4 { float x, y, z, w; };
struct double2 { double x, y; };
struct uchar4 { unsigned char x, y, z, w; };
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }
struct Dim2 { long long x, y; };
static Dim2 blockIdx{0, 0}, gridDim{1, 1}, blockDim{256, 1};
static thread_local Dim2 threadIdx{0, 0};
static pthread_barrier_t g_bar;
static int g_vals[256];
static inline void __syncthreads() { pthread_barrier_wait(&g_bar); }
namespace cub {
template <class T, int N>
struct BlockReduce {
    struct TempStorage {};
    BlockReduce(TempStorage &) {}
    T Sum(T v) {
        g_vals[threadIdx.x] = v;
        pthread_barrier_wait(&g_bar);
        T s = 0;
        for (int i = 0; i < N; ++i) s += g_vals[i];
        pthread_barrier_wait(&g_bar);
        return threadIdx.x == 0 ? s : T(0);          // like cub: only thread 0 holds the block aggregate
    }
};
template <class T, int N>
struct BlockScan {
    struct TempStorage {};
    BlockScan(TempStorage &) {}
    void ExclusiveSum(T v, T &out) {
        g_vals[threadIdx.x] = v;
        pthread_barrier_wait(&g_bar);
        T s = 0;
        for (int i = 0; i < threadIdx.x; ++i) s += g_vals[i];
        pthread_barrier_wait(&g_bar);
        out = s;
    }
};
}  // namespace cub
template <class F>
static void run_block(F f) {
    struct Arg { F *f; int t; };
    pthread_barrier_init(&g_bar, nullptr, 256);
    static pthread_t th[256];
    static Arg args[256];
    for (int t = 0; t < 256; ++t) {
        args[t] = {&f, t};
        pthread_create(&th[t], nullptr, [](void *p) -> void * { Arg *a = (Arg *)p; threadIdx.x = a->t; (*a->f)(); return nullptr; }, &args[t]);
    }
    for (int t = 0; t < 256; ++t) pthread_join(th[t], nullptr);
    pthread_barrier_destroy(&g_bar);
}
'''

DRIVER = r'''
template <typename DT>
static long long backproject_host(const DT *depth, const uint8_t *rgb, const uint8_t *masks, int n_masks, int H, int W,
                                  const float *hor, const float *ver, double fx, double fy, double *out_pts, double *out_cols,
                                  int64_t *h_offsets) {
    const int64_t P = (int64_t)H * W;
    const int ntiles = (int)((P + 1023) / 1024);
    const bool vec = (P % 4 == 0) && ((uintptr_t)depth % 16 == 0) && ((uintptr_t)masks % 4 == 0);
    const size_t nc = (size_t)n_masks * ntiles;
    int *counts = new int[nc + 1];
    int64_t *offs = new int64_t[nc + 1];
    counts[nc] = 0;
    for (int mi = 0; mi < n_masks; ++mi)
        for (int tile = 0; tile < ntiles; ++tile) {
            blockIdx.x = tile; blockIdx.y = mi;
            run_block([&] { bp_count_kernel<DT>(depth, masks, P, ntiles, vec, counts); });
        }
    offs[0] = 0;                                           // cub::DeviceScan::ExclusiveSum in the library
    for (size_t i = 0; i < nc; ++i) offs[i + 1] = offs[i] + counts[i];
    if (out_pts)
        for (int mi = 0; mi < n_masks; ++mi)
            for (int tile = 0; tile < ntiles; ++tile) {
                blockIdx.x = tile; blockIdx.y = mi;
                run_block([&] { bp_write_kernel<DT>(depth, rgb, masks, P, W, ntiles, vec, hor, ver, fx, fy, offs, out_pts, rgb ? out_cols : nullptr); });
            }
    for (int i = 0; i <= n_masks; ++i) h_offsets[i] = offs[(size_t)i * ntiles];
    long long total = offs[nc];
    delete[] counts; delete[] offs;
    return total;
}
extern "C" void host_mean3(const double *pts, long long n, double *mean) {
    double sums[3] = {0, 0, 0};
    blockIdx.x = 0; blockIdx.y = 0;
    run_block([&] { seqsum3_kernel(pts, n, sums); });
    for (int a = 0; a < 3; ++a) mean[a] = sums[a] / (double)n;          // ibl_mean3's host division
}
extern "C" long long host_backproject(const void *depth, int is64, const uint8_t *rgb, const uint8_t *masks, int n_masks, int H,
                                      int W, const float *hor, const float *ver, double fx, double fy, double *out_pts,
                                      double *out_cols, int64_t *h_offsets) {
    if (is64) return backproject_host<double>((const double *)depth, rgb, masks, n_masks, H, W, hor, ver, fx, fy, out_pts, out_cols, h_offsets);
    return backproject_host<float>((const float *)depth, rgb, masks, n_masks, H, W, hor, ver, fx, fy, out_pts, out_cols, h_offsets);
}
'''


@pytest.fixture(scope="module")
def block_lib(tmp_path_factory):
    src = open(os.path.join(CSRC, "cloud.cu")).read()
    body = src[src.index("template <typename DT>\n__device__ __forceinline__ void load4"):src.index("template <typename DT>\nstatic int backproject_impl")]
    d = tmp_path_factory.mktemp("block_host")
    cpp, so = str(d / "block_host.cpp"), str(d / "libblock_host.so")
    mean = src[src.index("#define MEAN_TILE 1024"):src.index("IBL_API int ibl_mean3")]
    open(cpp, "w").write(PRELUDE + body + mean + DRIVER)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-pthread", "-o", so, cpp])
    L = ctypes.CDLL(so)
    L.host_backproject.restype = ctypes.c_longlong
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _run(L, depth, rgb, masks, fx, fy):
    H, W = depth.shape
    n = len(masks)
    depth = np.ascontiguousarray(depth)
    masks8 = np.ascontiguousarray(masks.reshape(n, H * W).astype(np.uint8))
    rgb = np.ascontiguousarray(rgb)
    hor = np.linspace(-W / 2, W / 2, W, dtype=np.float32)
    ver = np.linspace(H / 2, -H / 2, H, dtype=np.float32)
    offs = np.zeros(n + 1, np.int64)
    args = (_p(depth), int(depth.dtype == np.float64), _p(rgb), _p(masks8), n, H, W, _p(hor), _p(ver),
            ctypes.c_double(fx), ctypes.c_double(fy))
    total = L.host_backproject(*args, None, None, _p(offs))
    pts, cols = np.full((total, 3), np.nan), np.full((total, 3), np.nan)
    L.host_backproject(*args, _p(pts), _p(cols), _p(offs))
    return pts, cols, [int(o) for o in offs]


def test_backproject_kernel_source_vs_reference_golden(block_lib, golden):
    """The full golden frame: every mask, float32 depth, then the float64-depth variant with an all-ones mask -- the
    reference's own outputs (tests/golden/make_golden.py), bit for bit."""
    g = golden
    pts, cols, offs = _run(block_lib, g["bp_depth"], g["bp_rgb"], g["bp_masks"][:, 0], float(g["bp_fx"]), float(g["bp_fy"]))
    assert offs == list(g["bp_offsets"])
    assert np.array_equal(pts, g["bp_points"]) and np.array_equal(cols, g["bp_colors"])
    H, W = g["bp64_depth"].shape
    pts, cols, offs = _run(block_lib, g["bp64_depth"], g["bp_rgb"], np.ones((1, H, W), np.uint8), float(g["bp_fx"]), float(g["bp_fy"]))
    assert np.array_equal(pts, g["bp64_points"]) and np.array_equal(cols, g["bp64_colors"])


def test_backproject_kernel_source_ragged_sizes(block_lib):
    """H*W not a multiple of 4 (scalar tail path), an empty mask, NaN depth under a zero mask (NaN * 0 stays a point)."""
    from oracle import ref_port
    rng = np.random.default_rng(5)
    for H, W in ((37, 53), (1, 7)):
        depth = rng.uniform(0.3, 5.0, (H, W)).astype(np.float32)
        depth[rng.random((H, W)) < 0.4] = 0
        rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        masks = rng.random((3, H, W)) < 0.3
        masks[2] = False
        pts, cols, offs = _run(block_lib, depth, rgb, masks, 525.0, 525.0)
        ref = ref_port.backproject_masks(depth, rgb, masks, 525.0, 525.0)
        assert offs == list(np.cumsum([0] + [len(p) for p, _ in ref])) and offs[3] == offs[2]
        assert np.array_equal(pts, np.concatenate([p for p, _ in ref]))
        assert np.array_equal(cols, np.concatenate([q for _, q in ref]))


@pytest.mark.parametrize("n", [1, 2, 1023, 1024, 1025, 5000])
def test_mean_kernel_source_vs_numpy(block_lib, n):
    """np.mean over axis 0 of a C-contiguous [N,3] array adds the rows one by one; the kernel must give the same bits."""
    rng = np.random.default_rng(n)
    pts = np.ascontiguousarray(rng.normal(size=(n, 3)) * np.array([1.0, 1e3, 1e-3]) + np.array([2.0, -700.0, 0.01]))
    got = np.full(3, np.nan)
    block_lib.host_mean3(_p(pts), ctypes.c_longlong(n), _p(got))
    assert np.array_equal(got, np.mean(pts, axis=0))
