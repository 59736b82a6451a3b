"""CPU check of a CUDA kernel's ARITHMETIC without a GPU: the body of ``raycast_kernel`` (csrc/raycast.cu) is a
thread-per-pixel loop with no inter-thread communication, so the very same source text compiles as host C++ once the CUDA
qualifiers are defined away and the _rn intrinsics are mapped to plain IEEE operations (g++ -ffp-contract=off, which is
what -fmad=false gives on the device).  One "thread" then walks every pixel and the result is compared with the
tensor-op ray caster of instance_based_loc_b200/synth.py.  The GPU run of the same kernel is
tests/test_z_memory_build_gpu.py::test_raycast_kernel_vs_tensor_ops.  No product code is involved: nothing here is a
fallback, the host build lives in a temporary directory."""
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
struct Dim1 { int x; };
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
    host_kernel.host_raycast(p(T), H, W, ctypes.c_double(f), p(kind), p(half), p(R), p(t), m, ctypes.c_double(sc.room),
                             ctypes.c_double(sc.height), p(ot), p(oi), p(on))
    same = i0 == oi
    assert same.mean() >= 0.9999
    assert np.array_equal(np.isinf(b0[same]), np.isinf(ot[same]))
    fin = same & np.isfinite(b0)
    assert fin.sum() == 0 or np.abs(b0 - ot)[fin].max() <= 1e-9
    assert np.abs(n0 - on)[same].max() <= 1e-9
