"""CPU check of WARP-LEVEL CUDA kernels without a GPU: a 32-lane warp is emulated by 32 host threads that run the
kernel's own source text (csrc/*.cu, CUDA qualifiers defined away) and meet at a barrier inside every warp collective
(``__ballot_sync``, ``__shfl_xor_sync``); ``atomicCAS`` is the host's compare-exchange, so the lock-free union-find of
csrc/dbscan.cu runs under real concurrency (its union pass, which has no collectives, runs as 8 warps = 256 host threads).  The kernels keep their collectives warp-uniform (the same requirement the
full-mask intrinsics have on the device), which is what makes the barrier emulation valid.

 * ``radius_count_kernel`` (csrc/grid_query.cu) against the oracle's exact radius counts, counts and keep-mask variants;
 * ``knn_kernel`` (csrc/grid_query.cu: two-pass histogram k-NN with a per-warp shared-memory candidate buffer, warp scans
   and a warp bitonic sort) against the oracle's hybrid search: neighbour lists, squared distances and counts bit for bit,
   grid points and external queries, both buffer capacities;
 * ``ibl_nn1_thread`` (csrc/ibl_common.cuh: the per-thread exact nearest neighbour within max_d by Chebyshev ring
   expansion that nn_score / RANSAC scoring / ICP use when no candidate field is attached) against the oracle's nn1;
 * ``nn_score_kernel`` + ``score_final_kernel`` (csrc/grid_query.cu: evaluate_registration -- transform, exact NN, fixed
   shuffle-tree block reduction over 8 warps, __syncthreads) as a 256-thread block: correspondences and inlier count
   bit-exact, err2 to 1e-12 (the oracle sums in index order);
 * ``ransac_count_kernel`` / ``ransac_score_kernel`` (csrc/ransac.cu: one block per (source tile, hypothesis)) on the
   plain grid: per-hypothesis inlier counts of both kernels equal to the oracle's, err2 to 1e-12;
 * ``spfh_kernel`` / ``fpfh_kernel`` (csrc/features.cu: warp per point, bins across lanes, shared-memory staging) against
   the oracle's orc_spfh / orc_fpfh (SPFH to 1e-12, FPFH within the north-star 1e-4 relative);
 * ``dbscan_*`` kernels (csrc/dbscan.cu: init, union pass, roots, border pass, labels) against the oracle's sequential
   restatement of Open3D's ClusterDBSCAN, labels bit for bit.

The cell-sorted grid the kernels walk is built here with numpy in the layout of csrc/grid.cu (origin = min corner,
cell = floor((p - o) * inv), x fastest, stable order inside a cell, original index in the bits of .w).  A test of the
source text, not a fallback: the build lives in a temporary directory.  GPU runs of the same kernels:
tests/test_kernels_gpu.py::test_radius_count_and_outlier, tests/test_z_memory_build_gpu.py::test_dbscan_*."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "instance_based_loc_b200", "csrc")

PRELUDE = r'''
#include <algorithm>
#include <climits>
#include <cmath>
#include <math.h>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <pthread.h>
using std::max;
using std::min;
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(x)
typedef void *cudaStream_t;
struct uint2 { unsigned x, y; };
struct double4 { double x, y, z, w; };
struct ibl_grid;
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __dsub_rn(double a, double b) { return a - b; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline long long __double_as_longlong(double d) { long long r; memcpy(&r, &d, 8); return r; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int atomicCAS(int *a, int cmp, int val) {
    __atomic_compare_exchange_n(a, &cmp, val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return cmp;                                            // the old value, like the device intrinsic
}
struct Dim1 { long long x, y; };
static Dim1 blockIdx{0, 0}, gridDim{1, 1}, blockDim{32, 1};
static thread_local Dim1 threadIdx{0, 0};
// ---- one emulated warp: 32 host threads, a barrier inside every collective
static pthread_barrier_t g_bar;                            // the single emulated warp of run_warp
static unsigned long long g_slot[32];
static pthread_barrier_t g_wbar[8], g_bbar;                // run_block256: one barrier per warp + the block barrier
static unsigned long long g_wslot[8][32];
static thread_local pthread_barrier_t *t_bar = &g_bar;     // the barrier / exchange slots of the calling lane's warp
static thread_local unsigned long long *t_slot = g_slot;
#define LANE ((int)(threadIdx.x & 31))
static inline unsigned __ballot_sync(unsigned, bool p) {
    t_slot[LANE] = p ? 1ull : 0ull;
    pthread_barrier_wait(t_bar);
    unsigned r = 0;
    for (int i = 0; i < 32; ++i) r |= (unsigned)t_slot[i] << i;
    pthread_barrier_wait(t_bar);
    return r;
}
static inline int __shfl_xor_sync(unsigned, int v, int o) {
    t_slot[LANE] = (unsigned long long)(unsigned)v;
    pthread_barrier_wait(t_bar);
    int r = (int)(unsigned)t_slot[LANE ^ o];
    pthread_barrier_wait(t_bar);
    return r;
}
static inline int __shfl_sync(unsigned, int v, int src) {
    t_slot[LANE] = (unsigned long long)(unsigned)v;
    pthread_barrier_wait(t_bar);
    int r = (int)(unsigned)t_slot[src & 31];
    pthread_barrier_wait(t_bar);
    return r;
}
static inline int __shfl_up_sync(unsigned, int v, unsigned d) {
    t_slot[LANE] = (unsigned long long)(unsigned)v;
    pthread_barrier_wait(t_bar);
    int r = LANE >= (int)d ? (int)(unsigned)t_slot[LANE - d] : v;
    pthread_barrier_wait(t_bar);
    return r;
}
static inline double __shfl_sync(unsigned, double v, int src) {
    memcpy(&t_slot[LANE], &v, 8);
    pthread_barrier_wait(t_bar);
    double r;
    memcpy(&r, &t_slot[src & 31], 8);
    pthread_barrier_wait(t_bar);
    return r;
}
template <class T>
static inline T shfl_down_any(T v, int o) {
    static_assert(sizeof(T) <= 8, "");
    unsigned long long bits = 0;
    memcpy(&bits, &v, sizeof(T));
    t_slot[LANE] = bits;
    pthread_barrier_wait(t_bar);
    T r = v;
    if (LANE + o < 32) memcpy(&r, &t_slot[LANE + o], sizeof(T));
    pthread_barrier_wait(t_bar);
    return r;
}
static inline double __shfl_down_sync(unsigned, double v, int o) { return shfl_down_any(v, o); }
static inline int __shfl_down_sync(unsigned, int v, int o) { return shfl_down_any(v, o); }
static inline long long __shfl_down_sync(unsigned, long long v, int o) { return shfl_down_any(v, o); }
static inline void __syncthreads() { pthread_barrier_wait(&g_bbar); }
static inline double __ddiv_rn(double a, double b) { return a / b; }
static inline void __syncwarp() { pthread_barrier_wait(t_bar); }
// "shared memory": one block (= one warp) at a time, so a file-scope buffer stands in for the dynamic allocation
#define __shared__ static                                   /* one block at a time: a function-static array IS shared by the 32 lanes */
#define __align__(x)
alignas(16) unsigned char smem_raw[1 << 16];
static inline int atomicAdd(int *a, int v) { return __atomic_fetch_add(a, v, __ATOMIC_SEQ_CST); }
static inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
static inline double __longlong_as_double(long long v) { double r; memcpy(&r, &v, 8); return r; }
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
// a block of `nthreads` host threads for kernels WITHOUT collectives (nobody waits at the barrier): several warps at once
template <class F>
static void run_block_no_collectives(int nthreads, F f) {
    struct Arg { F *f; int t; };
    pthread_t *th = new pthread_t[nthreads];
    Arg *args = new Arg[nthreads];
    for (int t = 0; t < nthreads; ++t) {
        args[t] = {&f, t};
        pthread_create(&th[t], nullptr, [](void *p) -> void * { Arg *a = (Arg *)p; threadIdx.x = a->t; (*a->f)(); return nullptr; }, &args[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], nullptr);
    delete[] th; delete[] args;
}
// a 256-thread block: 8 warps with their own collective barriers, __syncthreads on the block barrier
template <class F>
static void run_block256(F f) {
    struct Arg { F *f; int t; };
    for (int w = 0; w < 8; ++w) pthread_barrier_init(&g_wbar[w], nullptr, 32);
    pthread_barrier_init(&g_bbar, nullptr, 256);
    static pthread_t th[256];
    static Arg args[256];
    for (int t = 0; t < 256; ++t) {
        args[t] = {&f, t};
        pthread_create(&th[t], nullptr, [](void *p) -> void * {
            Arg *a = (Arg *)p;
            threadIdx.x = a->t; t_bar = &g_wbar[a->t >> 5]; t_slot = g_wslot[a->t >> 5];
            (*a->f)();
            return nullptr; }, &args[t]);
    }
    for (int t = 0; t < 256; ++t) pthread_join(th[t], nullptr);
    for (int w = 0; w < 8; ++w) pthread_barrier_destroy(&g_wbar[w]);
    pthread_barrier_destroy(&g_bbar);
}
template <class F>
static void run_warp(F f) {
    struct Arg { F *f; int lane; };
    pthread_barrier_init(&g_bar, nullptr, 32);
    pthread_t th[32];
    Arg args[32];
    for (int l = 0; l < 32; ++l) {
        args[l] = {&f, l};
        pthread_create(&th[l], nullptr, [](void *p) -> void * { Arg *a = (Arg *)p; threadIdx.x = a->lane; (*a->f)(); return nullptr; }, &args[l]);
    }
    for (int l = 0; l < 32; ++l) pthread_join(th[l], nullptr);
    pthread_barrier_destroy(&g_bar);
}
'''

DRIVER = r'''
static GridView make_view(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims) {
    GridView g;
    memset(&g, 0, sizeof(g));
    g.ox = o[0]; g.oy = o[1]; g.oz = o[2]; g.cell = cell; g.inv_cell = 1.0 / cell;
    g.nx = dims[0]; g.ny = dims[1]; g.nz = dims[2];
    g.cell_start = cell_start; g.sorted = (const double4 *)sorted4;
    return g;
}
extern "C" void host_radius_count(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims,
                                  long long n, double radius, int nb_points, int32_t *count, uint8_t *keep) {
    GridView g = make_view(sorted4, cell_start, o, cell, dims);
    const int S = (int)ceil(radius * g.inv_cell + 1e-6);
    run_warp([&] { radius_count_kernel(g, n, radius * radius, S, nb_points, count, keep); });
}
extern "C" void host_knn(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims,
                         const double *queries, long long nq, double radius, int max_nn, int32_t *idx, double *d2, int32_t *cnt) {
    GridView g = make_view(sorted4, cell_start, o, cell, dims);
    const double r2 = std::isfinite(radius) ? radius * radius : INFINITY;
    if (max_nn <= 96) run_warp([&] { knn_kernel<256>(g, queries, nq, radius, r2, max_nn, idx, d2, cnt); });
    else run_warp([&] { knn_kernel<512>(g, queries, nq, radius, r2, max_nn, idx, d2, cnt); });
}
extern "C" void host_nn1(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims,
                         const double *queries, long long nq, double max_d, int32_t *idx, double *d2) {
    GridView g = make_view(sorted4, cell_start, o, cell, dims);          // no ring field, no candidate field: the ring walk
    for (long long i = 0; i < nq; ++i)
        idx[i] = ibl_nn1_thread(g, queries[3 * i], queries[3 * i + 1], queries[3 * i + 2], max_d, max_d * max_d, d2[i]);
}
extern "C" void host_nn_score(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims,
                              const double *src, long long n, const double *T16, double max_d, int32_t *nn_idx, double *nn_d2,
                              long long *n_inliers, double *err2) {
    GridView g = make_view(sorted4, cell_start, o, cell, dims);
    Mat16 T;
    for (int k = 0; k < 16; ++k) T.v[k] = T16[k];
    const int nb = (int)((n + 255) / 256);
    long long *pc = new long long[nb];
    double *pe = new double[nb];
    blockDim.x = 256;
    for (int b = 0; b < nb; ++b) {
        blockIdx.x = b;
        run_block256([&] { nn_score_kernel(g, src, n, T, 1, max_d, max_d * max_d, nn_idx, nn_d2, pc, pe); });
    }
    blockIdx.x = 0; blockDim.x = 32;
    ibl_score_t res;
    run_warp([&] { score_final_kernel(pc, pe, nb, &res); });
    *n_inliers = res.n_inliers; *err2 = res.err2;
    delete[] pc; delete[] pe;
}
extern "C" void host_ransac_score(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims,
                                  const double *src, long long n_src, const double *T_all, int nh, double max_d,
                                  int32_t *count_only, int32_t *score_cnt, double *score_err) {
    GridView g = make_view(sorted4, cell_start, o, cell, dims);
    const int ntiles = (int)((n_src + SCORE_TILE - 1) / SCORE_TILE);
    int *pc = new int[(size_t)nh * ntiles];
    double *pe = new double[(size_t)nh * ntiles];
    blockDim.x = SCORE_THREADS;
    for (int h = 0; h < nh; ++h) {
        count_only[h] = 0;
        for (int t = 0; t < ntiles; ++t) {
            blockIdx.x = t; blockIdx.y = h;
            run_block256([&] { ransac_count_kernel(g, src, n_src, T_all, max_d, max_d * max_d, count_only); });
            run_block256([&] { ransac_score_kernel(g, src, n_src, T_all, max_d, max_d * max_d, ntiles, pc, pe); });
        }
        score_cnt[h] = 0; score_err[h] = 0.0;
        for (int t = 0; t < ntiles; ++t) { score_cnt[h] += pc[(size_t)h * ntiles + t]; score_err[h] += pe[(size_t)h * ntiles + t]; }
    }
    blockIdx.x = 0; blockIdx.y = 0; blockDim.x = 32;
    delete[] pc; delete[] pe;
}
extern "C" void host_spfh_fpfh(const double *pts, const double *nrm, long long n, const int32_t *idx, const double *d2,
                               const int32_t *cnt, int max_nn, double *spfh, double *fpfh) {
    run_warp([&] { spfh_kernel(pts, nrm, n, idx, cnt, max_nn, spfh); });
    run_warp([&] { fpfh_kernel(n, idx, d2, cnt, max_nn, spfh, fpfh); });
}
extern "C" void host_dbscan(const double *sorted4, const int *cell_start, const double *o, double cell, const int *dims,
                            long long n, double eps, int min_points, const int32_t *count, int32_t *labels) {
    GridView g = make_view(sorted4, cell_start, o, cell, dims);
    const int S = (int)ceil(eps * g.inv_cell + 1e-6);
    int *parent = new int[n], *root = new int[n], *flag = new int[n], *rank = new int[n + 1];
    uint8_t *core = new uint8_t[n];
    blockDim.x = 1;                                        // thread-per-element kernels: one element per "launch"
    for (long long i = 0; i < n; ++i) { blockIdx.x = i; dbscan_init_kernel(count, n, min_points, parent, core); }
    blockIdx.x = 0; blockDim.x = 256;                      // the union pass has no collectives: 8 warps race on the union-find
    run_block_no_collectives(256, [&] { dbscan_walk_kernel<0>(g, n, eps * eps, S, core, parent, root); });
    blockDim.x = 1;
    for (long long i = 0; i < n; ++i) { blockIdx.x = i; dbscan_root_kernel(n, core, parent, root, flag); }
    blockIdx.x = 0; blockDim.x = 32;
    run_warp([&] { dbscan_walk_kernel<1>(g, n, eps * eps, S, core, parent, root); });
    rank[0] = 0;                                           // cub::DeviceScan::ExclusiveSum in the library
    for (long long i = 0; i < n; ++i) rank[i + 1] = rank[i] + flag[i];
    blockDim.x = 1;
    for (long long i = 0; i < n; ++i) { blockIdx.x = i; dbscan_label_kernel(n, root, rank, labels); }
    blockIdx.x = 0; blockDim.x = 32;
    delete[] parent; delete[] root; delete[] flag; delete[] rank; delete[] core;
}
'''


@pytest.fixture(scope="module")
def warp_lib(tmp_path_factory):
    common = open(os.path.join(CSRC, "ibl_common.cuh")).read()
    views = common[common.index("struct NNFieldView {"):common.index("struct ibl_grid {")] + \
        common[common.index("// device view passed by value to kernels"):common.index("void ibl_grid_free_field")] + \
        common[common.index("// exact nanoflann L2"):common.index("// bounding box: per-block partial min/max")]
    gq = open(os.path.join(CSRC, "grid_query.cu")).read()
    radius = gq[gq.index("__global__ void __launch_bounds__(256) radius_count_kernel"):gq.index("static int radius_launch")]
    knn = gq[gq.index("// hybrid k-NN\nstruct KnnBuf"):gq.index("IBL_API int ibl_knn_hybrid")]
    dyn = "extern __shared__ __align__(16) unsigned char smem_raw[];"          # dynamic shared memory -> the file-scope buffer
    assert dyn in knn
    knn = knn.replace(dyn, "")
    ft = open(os.path.join(CSRC, "features.cu")).read()
    feats = ft[ft.index("__device__ __forceinline__ void cross3"):ft.index("__device__ void eigvec0")] + \
        ft[ft.index("__device__ void pair_features"):ft.index("IBL_API int ibl_spfh")] + \
        ft[ft.index("#define FPFH_MAX_NN"):ft.index("IBL_API int ibl_fpfh")]
    score = gq[gq.index("struct Mat16 { double v[16]; };"):gq.index("IBL_API int ibl_nn_score")]
    score = "typedef struct { int64_t n_inliers; double err2; } ibl_score_t;\n" + score
    rs = open(os.path.join(CSRC, "ransac.cu")).read()
    rscore = rs[rs.index("#define SCORE_TILE 512"):rs.index("// one CTA per hypothesis: tile partials")]
    db = open(os.path.join(CSRC, "dbscan.cu")).read()
    dbscan = db[db.index("__device__ __forceinline__ int uf_find"):db.index("// stream-ordered scratch that is returned")]
    d = tmp_path_factory.mktemp("warp_host")
    cpp, so = str(d / "warp_host.cpp"), str(d / "libwarp_host.so")
    open(cpp, "w").write(PRELUDE + views + radius + knn + score + rscore + dbscan + feats + DRIVER)
    subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-pthread", "-o", so, cpp])
    return ctypes.CDLL(so)


def host_grid(pts, cell):
    """The layout of csrc/grid.cu: (sorted double[n][4] with the original index in the bits of .w, cell_start, origin, dims)."""
    pts = np.ascontiguousarray(pts, np.float64)
    o = pts.min(axis=0)
    dims = (np.floor((pts.max(axis=0) - o) / cell) + 1).astype(np.int64)
    inv = 1.0 / cell
    c = np.clip(np.floor((pts - o) * inv).astype(np.int64), 0, dims - 1)
    cid = (c[:, 2] * dims[1] + c[:, 1]) * dims[0] + c[:, 0]
    order = np.argsort(cid, kind="stable")
    ncells = int(dims.prod())
    cell_start = np.searchsorted(cid[order], np.arange(ncells + 1)).astype(np.int32)
    sorted4 = np.empty((len(pts), 4))
    sorted4[:, :3] = pts[order]
    sorted4[:, 3] = order.astype(np.int64).view(np.float64)
    return np.ascontiguousarray(sorted4), cell_start, np.ascontiguousarray(o), dims.astype(np.int32)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _cloud(seed, n_blob=120, n_noise=80):
    rng = np.random.default_rng(seed)
    cs = rng.uniform(-1, 1, (4, 3))
    pts = np.concatenate([c + rng.normal(0, 0.06, (n_blob, 3)) for c in cs] + [rng.uniform(-1.5, 1.5, (n_noise, 3))])
    return np.ascontiguousarray(pts[rng.permutation(len(pts))])


@pytest.mark.parametrize("radius,cell_factor", [(0.1, 1.001), (0.1, 0.5), (0.25, 1.7)])
def test_radius_count_kernel_source_vs_oracle(warp_lib, radius, cell_factor):
    from oracle import o3d_core as oc
    pts = _cloud(1)
    n = len(pts)
    sorted4, cell_start, o, dims = host_grid(pts, radius * cell_factor)
    want = oc.radius_count(pts, radius)
    count = np.full(n, -1, np.int32)
    warp_lib.host_radius_count(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(radius * cell_factor), _p(dims),
                               ctypes.c_longlong(n), ctypes.c_double(radius), 0, _p(count), None)
    assert np.array_equal(count, want)
    keep = np.full(n, 7, np.uint8)                                  # keep-mask variant: early exit, centre-first rows
    warp_lib.host_radius_count(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(radius * cell_factor), _p(dims),
                               ctypes.c_longlong(n), ctypes.c_double(radius), 8, None, _p(keep))
    assert np.array_equal(keep.astype(bool), want > 8)


@pytest.mark.parametrize("eps,min_points,seed", [(0.08, 5, 1), (0.12, 10, 2), (0.1, 1, 3), (0.2, 40, 4)])
def test_dbscan_kernels_source_vs_oracle(warp_lib, eps, min_points, seed):
    from oracle import o3d_core as oc
    pts = _cloud(seed)
    n = len(pts)
    sorted4, cell_start, o, dims = host_grid(pts, eps * 1.001)
    count = oc.radius_count(pts, eps).astype(np.int32)
    labels = np.full(n, -7, np.int32)
    warp_lib.host_dbscan(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(eps * 1.001), _p(dims), ctypes.c_longlong(n),
                         ctypes.c_double(eps), min_points, _p(count), _p(labels))
    want = oc.cluster_dbscan(pts, eps, min_points)
    assert np.array_equal(labels, want)
    assert want.max() >= 1 or min_points >= 40


@pytest.mark.parametrize("radius,k,cell,n_pts", [(0.1, 30, 0.05, 130), (0.25, 100, 0.08, 130), (np.inf, 30, 0.1, 130),
                                                  (0.05, 12, 0.05, 130), (0.3, 224, 0.2, 320)])
def test_knn_kernel_source_vs_oracle(warp_lib, radius, k, cell, n_pts):
    from ibl_testutil import surface_cloud
    from oracle import o3d_core as oc
    # Every warp collective is two pthread barriers here, and the register sort of <= 128 candidates is 300 shuffles per
    # query: small clouds.  The k = 224 case keeps 320 points, so that more than 128 candidates reach the shared-memory sort.
    pts = np.ascontiguousarray(surface_cloud(n_pts, 12)[0])
    n = len(pts)
    sorted4, cell_start, o, dims = host_grid(pts, cell)
    rng = np.random.default_rng(0)
    ext = np.ascontiguousarray(pts[::9] + rng.normal(0, 0.05, pts[::9].shape))
    ext[:3] += 10.0                                                  # far outside the grid box
    for queries in (None, ext):
        nq = n if queries is None else len(queries)
        idx = np.full((nq, k), -9, np.int32)
        d2 = np.full((nq, k), -9.0)
        cnt = np.full(nq, -9, np.int32)
        warp_lib.host_knn(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(cell), _p(dims),
                          None if queries is None else _p(queries), ctypes.c_longlong(nq), ctypes.c_double(radius), k,
                          _p(idx), _p(d2), _p(cnt))
        ridx, rd2, rcnt = oc.knn_hybrid(pts, pts if queries is None else queries, radius, k)
        assert np.array_equal(cnt, rcnt)
        assert np.array_equal(idx, ridx)
        assert np.array_equal(d2, rd2)


def test_spfh_fpfh_kernel_source_vs_oracle(warp_lib):
    from ibl_testutil import surface_cloud
    from oracle import o3d_core as oc
    pts = np.ascontiguousarray(surface_cloud(450, 14)[0])
    n, k, radius = len(pts), 100, 0.25
    nrm = np.ascontiguousarray(oc.estimate_normals(pts, 0.1, 30))
    idx, d2, cnt = oc.knn_hybrid(pts, pts, radius, k)
    idx, d2, cnt = np.ascontiguousarray(idx, np.int32), np.ascontiguousarray(d2), np.ascontiguousarray(cnt, np.int32)
    spfh = np.full((n, 33), np.nan)                                   # device layout: point-major double[n][33]
    fpfh = np.full((n, 33), np.nan)
    warp_lib.host_spfh_fpfh(_p(pts), _p(nrm), ctypes.c_longlong(n), _p(idx), _p(d2), _p(cnt), k, _p(spfh), _p(fpfh))
    rf, rs = oc.compute_fpfh(pts, nrm, radius, k, return_spfh=True)
    np.testing.assert_allclose(spfh.T, rs, rtol=1e-12, atol=1e-9)
    scale = np.maximum(np.abs(rf), 1e-3 * np.abs(rf).max())
    err = np.max(np.abs(fpfh.T - rf) / scale)
    assert err <= 1e-4, err


@pytest.mark.parametrize("max_d,cell", [(0.075, 0.075), (0.02, 0.05), (0.3, 0.1)])
def test_nn1_ring_walk_source_vs_oracle(warp_lib, max_d, cell):
    from ibl_testutil import random_rigid, surface_cloud
    from oracle import o3d_core as oc
    tgt = np.ascontiguousarray(surface_cloud(1500, 20)[0])
    T = random_rigid(4, max_angle=0.05, max_t=0.03)
    src = surface_cloud(900, 21)[0] @ T[:3, :3].T + T[:3, 3]
    src = np.ascontiguousarray(np.concatenate([src, tgt[:50], tgt[:5] + 7.0]))       # exact hits and far-away queries
    sorted4, cell_start, o, dims = host_grid(tgt, cell)
    idx = np.full(len(src), -9, np.int32)
    d2 = np.full(len(src), -9.0)
    warp_lib.host_nn1(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(cell), _p(dims), _p(src),
                      ctypes.c_longlong(len(src)), ctypes.c_double(max_d), _p(idx), _p(d2))
    ridx, rd2 = oc.nn1(tgt, src, max_d)
    assert np.array_equal(idx, ridx)
    hit = ridx >= 0
    assert hit.sum() > 50 and (~hit).sum() >= 5
    assert np.array_equal(d2[hit], rd2[hit])


@pytest.mark.parametrize("max_d", [0.02, 0.075])
def test_nn_score_kernels_source_vs_oracle(warp_lib, max_d):
    from ibl_testutil import random_rigid, surface_cloud
    from oracle import o3d_core as oc
    tgt = np.ascontiguousarray(surface_cloud(1500, 20)[0])
    T = np.ascontiguousarray(random_rigid(5, max_angle=0.03, max_t=0.01))
    src = np.ascontiguousarray(surface_cloud(1500, 20)[0][:700] + np.random.default_rng(1).normal(0, 0.004, (700, 3)))
    sorted4, cell_start, o, dims = host_grid(tgt, max_d)
    n = len(src)
    idx, d2 = np.full(n, -9, np.int32), np.full(n, -9.0)
    cnt, err = ctypes.c_longlong(-1), ctypes.c_double(-1.0)
    warp_lib.host_nn_score(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(max_d), _p(dims), _p(src), ctypes.c_longlong(n),
                           _p(T), ctypes.c_double(max_d), _p(idx), _p(d2), ctypes.byref(cnt), ctypes.byref(err))
    moved = oc.transform_points(T, src)
    ridx, rd2 = oc.nn1(tgt, moved, max_d)
    assert np.array_equal(idx, ridx)
    hit = ridx >= 0
    assert 50 < hit.sum() and np.array_equal(d2[hit], rd2[hit]) and np.isinf(d2[~hit]).all()
    assert cnt.value == int(hit.sum())
    assert abs(err.value - rd2[hit].sum()) <= 1e-12 * rd2[hit].sum()


def test_ransac_score_kernels_source_vs_oracle(warp_lib):
    from ibl_testutil import random_rigid, surface_cloud
    from oracle import o3d_core as oc
    tgt = np.ascontiguousarray(surface_cloud(1500, 20)[0])
    src = np.ascontiguousarray(surface_cloud(1500, 20)[0][:900] + np.random.default_rng(2).normal(0, 0.01, (900, 3)))
    max_d = 0.075
    # a good hypothesis, two mediocre ones and one that moves everything away (zero inliers)
    Ts = np.stack([np.eye(4), random_rigid(6, max_angle=0.05, max_t=0.02), random_rigid(7, max_angle=0.3, max_t=0.1),
                   np.array([[1, 0, 0, 50.0], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1.0]])])
    Ts = np.ascontiguousarray(Ts.reshape(len(Ts), 16))
    sorted4, cell_start, o, dims = host_grid(tgt, max_d)
    nh = len(Ts)
    c0, c1, e1 = np.full(nh, -9, np.int32), np.full(nh, -9, np.int32), np.full(nh, -9.0)
    warp_lib.host_ransac_score(_p(sorted4), _p(cell_start), _p(o), ctypes.c_double(max_d), _p(dims), _p(src),
                               ctypes.c_longlong(len(src)), _p(Ts), nh, ctypes.c_double(max_d), _p(c0), _p(c1), _p(e1))
    for h in range(nh):
        ridx, rd2 = oc.nn1(tgt, oc.transform_points(Ts[h].reshape(4, 4), src), max_d)
        hit = ridx >= 0
        assert c0[h] == c1[h] == int(hit.sum())
        assert abs(e1[h] - rd2[hit].sum()) <= 1e-12 * max(rd2[hit].sum(), 1e-300)
    assert c0[0] > c0[2] and c0[3] == 0
