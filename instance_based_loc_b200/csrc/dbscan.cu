// dbscan.cu -- Open3D PointCloud::ClusterDBSCAN on the cell-sorted grid (memory build, SURVEY.md section 8(f) row N2;
// object_memory/object_memory.py:315,511,632) and the axis-aligned bounds behind utils/IoU_ops.py:9-97.
//
// Open3D's sequential loop (precomputed radius neighbourhoods, scan in index order, expand from every unvisited core
// point) has an order-independent fixed point, which is what is computed here:
//   * i is a core point iff #{j : d2(i,j) < eps^2} (self included) >= min_points;
//   * clusters are the connected components of the core points under "within eps"; they are numbered in the order of
//     their smallest core index (the scan meets that point first);
//   * a non-core point takes the smallest cluster number among the core points within eps of it (the first cluster that
//     reaches it keeps it), else -1.
// Components come from a lock-free union-find that always hooks the larger root under the smaller one, so the root of a
// component IS its smallest core index and cluster numbers are a prefix sum over the root flags.  All integer work:
// labels are bit-exact against the sequential restatement in oracle/o3d_core.c.
//   union pass : warp per core point, whole x-rows of cells streamed with coalesced 32 B loads (same walk as
//                radius_count_kernel); a lane that sees a core neighbour with a smaller index unites the two
//   border pass: warp per non-core point, min over the roots of the core points within eps
#include <climits>
#include <cub/cub.cuh>

#include "ibl_common.cuh"

__device__ __forceinline__ int uf_find(int *parent, int x) {
    volatile int *vp = parent;
    int p = vp[x];
    while (p != x) {
        int gp = vp[p];
        if (gp != p) vp[x] = gp;          // path halving: x is not a root and never becomes one again; gp is an ancestor
        x = p;
        p = gp;
    }
    return x;
}

__device__ __forceinline__ void uf_unite(int *parent, int a, int b) {
    for (;;) {
        a = uf_find(parent, a);
        b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { int t = a; a = b; b = t; }                 // a > b: hook a under b
        if (atomicCAS(&parent[a], a, b) == a) return;           // a was still a root
    }
}

__global__ void dbscan_init_kernel(const int32_t *__restrict__ count, int64_t n, int32_t min_points, int *parent,
                                   uint8_t *core) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    parent[i] = (int)i;
    core[i] = count[i] >= min_points;
}

// MODE 0: unite every core point with its core neighbours of smaller index.
// MODE 1: root[i] of every NON-core point = min root over the core points within eps (INT_MAX: noise).
template <int MODE>
__global__ void __launch_bounds__(256) dbscan_walk_kernel(GridView g, int64_t n, double r2, int S, const uint8_t *__restrict__ core,
                                                          int *parent, int *__restrict__ root) {
    int lane = threadIdx.x & 31;
    int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t qi = warp; qi < n; qi += nwarps) {
        double4 q = g.sorted[qi];
        const int qo = (int)__double_as_longlong(q.w);
        const bool qcore = core[qo] != 0;
        if (MODE == 0 ? !qcore : qcore) continue;               // warp-uniform
        int cx = ibl_clampi(ibl_cell_coord(q.x, g.ox, g.inv_cell, g.nx), 0, g.nx - 1);
        int cy = ibl_clampi(ibl_cell_coord(q.y, g.oy, g.inv_cell, g.ny), 0, g.ny - 1);
        int cz = ibl_clampi(ibl_cell_coord(q.z, g.oz, g.inv_cell, g.nz), 0, g.nz - 1);
        int x0 = max(cx - S, 0), x1 = min(cx + S, g.nx - 1);
        int y0 = max(cy - S, 0), y1 = min(cy + S, g.ny - 1);
        int z0 = max(cz - S, 0), z1 = min(cz + S, g.nz - 1);
        int best = INT_MAX;
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                size_t rb = ((size_t)z * g.ny + y) * g.nx;
                int p0 = g.cell_start[rb + x0], p1 = g.cell_start[rb + x1 + 1];
                for (int p = p0 + lane; p < p1; p += 32) {
                    double4 t = g.sorted[p];
                    const int to = (int)__double_as_longlong(t.w);
                    if (!core[to]) continue;
                    if (MODE == 0 && to >= qo) continue;
                    if (!(ibl_d2(t.x, t.y, t.z, q.x, q.y, q.z) < r2)) continue;
                    if (MODE == 0) uf_unite(parent, qo, to);
                    else best = min(best, root[to]);
                }
            }
        if (MODE == 1) {
            for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
            if (lane == 0) root[qo] = best;
        }
    }
}

// core points: root = find(i), flag = (root == i); non-core: flag 0 (root is written by the border pass)
__global__ void dbscan_root_kernel(int64_t n, const uint8_t *__restrict__ core, int *parent, int *__restrict__ root,
                                   int *__restrict__ flag) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int f = 0;
    if (core[i]) {
        int r = uf_find(parent, (int)i);
        root[i] = r;
        f = (r == (int)i);
    }
    flag[i] = f;
}

__global__ void dbscan_label_kernel(int64_t n, const int *__restrict__ root, const int *__restrict__ rank,
                                    int32_t *__restrict__ labels) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int r = root[i];
    labels[i] = (r == INT_MAX) ? -1 : rank[r];
}

// stream-ordered scratch that is returned to the pool on every exit path (the IBL_CUDA / IBL_ARG macros return early)
struct DbscanScratch {
    cudaStream_t st;
    void *ptr[8];
    int n = 0;
    ~DbscanScratch() { for (int i = 0; i < n; ++i) ibl_free(ptr[i], st); }
    template <class T>
    cudaError_t get(T **out, size_t count) {
        cudaError_t e = ibl_alloc(out, count, st);
        if (e == cudaSuccess) ptr[n++] = *out;
        return e;
    }
};

IBL_API int ibl_dbscan(const ibl_grid_t *g, double eps, int32_t min_points, int32_t *labels, int32_t *h_nclusters,
                       void *stream) {
    IBL_ARG(g != nullptr && eps > 0 && (labels != nullptr || g->n == 0));
    cudaStream_t st = ibl_stream(stream);
    if (h_nclusters) *h_nclusters = 0;
    const int64_t n = g->n;
    if (n == 0) return IBL_OK;
    IBL_ARG(n < (int64_t)INT_MAX);
    DbscanScratch scr{st, {}};                                   // frees everything on every exit path
    int32_t *count = nullptr;
    int *parent = nullptr, *root = nullptr, *flag = nullptr, *rank = nullptr;
    uint8_t *core = nullptr, *tmp = nullptr;
    size_t tmp_bytes = 0;
    IBL_CUDA(scr.get(&count, n));
    IBL_CUDA(scr.get(&parent, n));
    IBL_CUDA(scr.get(&root, n));
    IBL_CUDA(scr.get(&flag, n));
    IBL_CUDA(scr.get(&rank, n + 1));
    IBL_CUDA(scr.get(&core, n));
    int rc = ibl_radius_count(g, eps, count, stream);
    if (rc != IBL_OK) return rc;
    const int S = (int)ceil(eps * g->inv_cell + 1e-6);
    const int tb = (int)((n + 255) / 256);
    int wb = (int)((n * 32 + 255) / 256);
    if (wb > IBL_SMS * 16) wb = IBL_SMS * 16;
    {
        IBL_PROF(IBL_PROF_MISC, st, 0.0);
        dbscan_init_kernel<<<tb, 256, 0, st>>>(count, n, min_points, parent, core);
        IBL_LAUNCH_CHECK();
        dbscan_walk_kernel<0><<<wb, 256, 0, st>>>(ibl_view(g), n, eps * eps, S, core, parent, root);
        IBL_LAUNCH_CHECK();
        dbscan_root_kernel<<<tb, 256, 0, st>>>(n, core, parent, root, flag);
        IBL_LAUNCH_CHECK();
        dbscan_walk_kernel<1><<<wb, 256, 0, st>>>(ibl_view(g), n, eps * eps, S, core, parent, root);
        IBL_LAUNCH_CHECK();
        IBL_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flag, rank, (int)n, st));
        IBL_CUDA(scr.get(&tmp, tmp_bytes));
        IBL_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flag, rank, (int)n, st));
        IBL_LAUNCH_CHECK();
        dbscan_label_kernel<<<tb, 256, 0, st>>>(n, root, rank, labels);
        IBL_LAUNCH_CHECK();
    }
    if (h_nclusters) {
        int last_rank = 0, last_flag = 0;
        IBL_CUDA(cudaMemcpyAsync(&last_rank, rank + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        IBL_CUDA(cudaMemcpyAsync(&last_flag, flag + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));
        *h_nclusters = last_rank + last_flag;
    }
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// Axis-aligned bounds of a cloud (pcd.min(axis=-1) / pcd.max(axis=-1), utils/IoU_ops.py:24-28,71-75): min and max are
// exact in any order, so a plain two-stage reduction is bit-identical to numpy.  NaNs are not expected in a cloud.
__global__ void __launch_bounds__(256) bounds3_kernel(const double *__restrict__ pts, int64_t n, double *__restrict__ part) {
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        for (int a = 0; a < 3; ++a) {
            double v = pts[3 * i + a];
            lo[a] = fmin(lo[a], v);
            hi[a] = fmax(hi[a], v);
        }
    __shared__ double s[8][6];
    for (int a = 0; a < 3; ++a)
        for (int o = 16; o; o >>= 1) {
            lo[a] = fmin(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmax(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0)
        for (int a = 0; a < 3; ++a) { s[w][a] = lo[a]; s[w][3 + a] = hi[a]; }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = s[0][threadIdx.x];
        for (int k = 1; k < 8; ++k) v = threadIdx.x < 3 ? fmin(v, s[k][threadIdx.x]) : fmax(v, s[k][threadIdx.x]);
        part[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

__global__ void bounds3_final_kernel(const double *__restrict__ part, int nb, double *__restrict__ out) {
    int a = threadIdx.x;
    if (a >= 6) return;
    double v = part[a];
    for (int k = 1; k < nb; ++k) v = a < 3 ? fmin(v, part[k * 6 + a]) : fmax(v, part[k * 6 + a]);
    out[a] = v;
}

IBL_API int ibl_bounds3(const double *pts, int64_t n, double *h_minmax6, void *stream) {
    IBL_ARG(h_minmax6 != nullptr && n >= 1 && pts != nullptr);
    cudaStream_t st = ibl_stream(stream);
    int nb = (int)((n + 255) / 256);
    if (nb > IBL_SMS * 4) nb = IBL_SMS * 4;
    double *part = nullptr;
    IBL_CUDA(ibl_alloc(&part, (size_t)(nb + 1) * 6, st));
    {
        IBL_PROF(IBL_PROF_MISC, st, 24.0 * n);
        bounds3_kernel<<<nb, 256, 0, st>>>(pts, n, part);
        IBL_LAUNCH_CHECK();
        bounds3_final_kernel<<<1, 32, 0, st>>>(part, nb, part + (size_t)nb * 6);
        IBL_LAUNCH_CHECK();
    }
    IBL_CUDA(cudaMemcpyAsync(h_minmax6, part + (size_t)nb * 6, 6 * sizeof(double), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(part, st);
    return IBL_OK;
}
