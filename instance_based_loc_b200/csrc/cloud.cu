// cloud.cu -- depth -> per-mask clouds, order-preserving selection, voxel down-sampling, rigid
// transforms (reference: utils/depth_utils.py:12-116,146-265).
#include <cub/cub.cuh>

#include <vector>

#include "ibl_common.cuh"

// ------------------------------------------------------------------------------------------------
// back-projection: each thread owns 4 consecutive pixels of one mask (float4 / uchar4 loads when
// H*W % 4 == 0); pass 1 counts, pass 2 scatters at scanned offsets (row-major order preserved).
template <typename DT>
__device__ __forceinline__ void load4(const DT *__restrict__ depth, const uint8_t *__restrict__ mask, int64_t p0,
                                      int64_t P, bool vec, DT d[4], uint8_t m[4]) {
    if (vec && p0 + 3 < P) {
        if constexpr (sizeof(DT) == 4) {
            float4 v = *reinterpret_cast<const float4 *>(depth + p0);
            d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
        } else {
            double2 a = *reinterpret_cast<const double2 *>(depth + p0);
            double2 b = *reinterpret_cast<const double2 *>(depth + p0 + 2);
            d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y;
        }
        uchar4 mm = *reinterpret_cast<const uchar4 *>(mask + p0);
        m[0] = mm.x; m[1] = mm.y; m[2] = mm.z; m[3] = mm.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            bool in = p0 + k < P;
            d[k] = in ? depth[p0 + k] : (DT)0;
            m[k] = in ? mask[p0 + k] : 0;
        }
    }
}

// depth * mask exactly as numpy: d*1 or d*0 (NaN*0 stays NaN and counts as != 0)
template <typename DT>
__device__ __forceinline__ DT masked(DT d, uint8_t m) { return m ? d : d * (DT)0; }

template <typename DT>
__global__ void __launch_bounds__(256) bp_count_kernel(const DT *__restrict__ depth, const uint8_t *__restrict__ masks,
                                                       int64_t P, int ntiles, bool vec, int *__restrict__ counts) {
    int mi = blockIdx.y, tile = blockIdx.x;
    int64_t p0 = ((int64_t)tile * 256 + threadIdx.x) * 4;
    DT d[4]; uint8_t m[4];
    load4<DT>(depth, masks + (size_t)mi * P, p0, P, vec, d, m);
    int c = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) c += (p0 + k < P) && (masked<DT>(d[k], m[k]) != (DT)0);
    typedef cub::BlockReduce<int, 256> BR;
    __shared__ typename BR::TempStorage ts;
    int tot = BR(ts).Sum(c);
    if (threadIdx.x == 0) counts[(size_t)mi * ntiles + tile] = tot;
}

template <typename DT>
__global__ void __launch_bounds__(256) bp_write_kernel(const DT *__restrict__ depth, const uint8_t *__restrict__ rgb,
                                                       const uint8_t *__restrict__ masks, int64_t P, int W, int ntiles,
                                                       bool vec, const float *__restrict__ hor,
                                                       const float *__restrict__ ver, double fx, double fy,
                                                       const int64_t *__restrict__ offs, double *__restrict__ out_pts,
                                                       double *__restrict__ out_cols) {
    int mi = blockIdx.y, tile = blockIdx.x;
    int64_t p0 = ((int64_t)tile * 256 + threadIdx.x) * 4;
    DT d[4]; uint8_t m[4];
    load4<DT>(depth, masks + (size_t)mi * P, p0, P, vec, d, m);
    bool ok[4];
    int c = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) { d[k] = masked<DT>(d[k], m[k]); ok[k] = (p0 + k < P) && (d[k] != (DT)0); c += ok[k]; }
    typedef cub::BlockScan<int, 256> BS;
    __shared__ typename BS::TempStorage ts;
    int pre;
    BS(ts).ExclusiveSum(c, pre);
    int64_t o = offs[(size_t)mi * ntiles + tile] + pre;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (!ok[k]) continue;
        int64_t p = p0 + k;
        int row = (int)(p / W), col = (int)(p - (int64_t)row * W);
        double X, Y, Z;
        if constexpr (sizeof(DT) == 4) {
            // numpy: (f32 * f32) / f32(fx)  -- two roundings in float32
            X = (double)__fdiv_rn(__fmul_rn(hor[col], d[k]), (float)fx);
            Y = (double)__fdiv_rn(__fmul_rn(ver[row], d[k]), (float)fy);
            Z = (double)d[k];
        } else {
            X = __ddiv_rn(__dmul_rn((double)hor[col], d[k]), fx);
            Y = __ddiv_rn(__dmul_rn((double)ver[row], d[k]), fy);
            Z = d[k];
        }
        out_pts[3 * o] = X; out_pts[3 * o + 1] = Y; out_pts[3 * o + 2] = Z;
        if (out_cols) {
#pragma unroll
            for (int a = 0; a < 3; ++a) out_cols[3 * o + a] = (double)__fdiv_rn((float)rgb[3 * p + a], 255.0f);
        }
        ++o;
    }
}

template <typename DT>
static int backproject_impl(const DT *depth, const uint8_t *rgb, const uint8_t *masks, int n_masks, int H, int W,
                            const float *hor, const float *ver, double fx, double fy, double *out_pts, double *out_cols,
                            int64_t *h_offsets, cudaStream_t st) {
    int64_t P = (int64_t)H * W;
    int ntiles = (int)((P + 1023) / 1024);
    bool vec = (P % 4 == 0) && ((uintptr_t)depth % 16 == 0) && ((uintptr_t)masks % 4 == 0);
    size_t nc = (size_t)n_masks * ntiles;
    int *counts = nullptr; int64_t *offs = nullptr;
    IBL_CUDA(ibl_alloc(&counts, nc + 1, st));
    IBL_CUDA(ibl_alloc(&offs, nc + 1, st));
    IBL_CUDA(cudaMemsetAsync(counts + nc, 0, sizeof(int), st));
    dim3 grid(ntiles, n_masks);
    {
        IBL_PROF(IBL_PROF_BACKPROJECT, st, (double)n_masks * P * (sizeof(DT) + 1));
        bp_count_kernel<DT><<<grid, 256, 0, st>>>(depth, masks, P, ntiles, vec, counts);
        IBL_LAUNCH_CHECK();
    }
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, counts, offs, (int)(nc + 1), st);
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tb ? tb : 1, st));
    IBL_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, counts, offs, (int)(nc + 1), st));
    if (out_pts) {
        IBL_PROF(IBL_PROF_BACKPROJECT, st, (double)n_masks * P * (sizeof(DT) + 1) + 3.0 * P);
        bp_write_kernel<DT><<<grid, 256, 0, st>>>(depth, rgb, masks, P, W, ntiles, vec, hor, ver, fx, fy, offs, out_pts,
                                                  rgb ? out_cols : nullptr);
        IBL_LAUNCH_CHECK();
    }
    // per-mask offsets = offs[mi*ntiles]
    int64_t *hall = (int64_t *)malloc(sizeof(int64_t) * (nc + 1));
    IBL_CUDA(cudaMemcpyAsync(hall, offs, sizeof(int64_t) * (nc + 1), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    for (int i = 0; i <= n_masks; ++i) h_offsets[i] = hall[(size_t)i * ntiles];
    free(hall);
    ibl_free(tmp, st); ibl_free(counts, st); ibl_free(offs, st);
    return IBL_OK;
}

IBL_API int ibl_backproject_masks(const void *depth, int depth_is_f64, const uint8_t *rgb, const uint8_t *masks,
                                  int n_masks, int H, int W, const float *hor, const float *ver, double fx, double fy,
                                  double *out_pts, double *out_cols, int64_t *h_offsets, void *stream) {
    IBL_ARG(h_offsets && n_masks >= 0 && H > 0 && W > 0);
    if (n_masks == 0) { h_offsets[0] = 0; return IBL_OK; }
    IBL_ARG(depth && masks && hor && ver);
    cudaStream_t st = ibl_stream(stream);
    if (depth_is_f64)
        return backproject_impl<double>((const double *)depth, rgb, masks, n_masks, H, W, hor, ver, fx, fy, out_pts,
                                        out_cols, h_offsets, st);
    return backproject_impl<float>((const float *)depth, rgb, masks, n_masks, H, W, hor, ver, fx, fy, out_pts, out_cols,
                                   h_offsets, st);
}

// ------------------------------------------------------------------------------------------------
// order-preserving selection
struct U8ToInt {
    __host__ __device__ int operator()(uint8_t v) const { return v ? 1 : 0; }
};

__global__ void select_scatter_kernel(const uint8_t *__restrict__ keep, const int *__restrict__ pos, int64_t n,
                                      const double *__restrict__ a, const double *__restrict__ b,
                                      const double *__restrict__ c, double *__restrict__ oa, double *__restrict__ ob,
                                      double *__restrict__ oc, int32_t *__restrict__ oidx) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n || !keep[i]) return;
    int64_t o = pos[i];
    if (a) { oa[3 * o] = a[3 * i]; oa[3 * o + 1] = a[3 * i + 1]; oa[3 * o + 2] = a[3 * i + 2]; }
    if (b) { ob[3 * o] = b[3 * i]; ob[3 * o + 1] = b[3 * i + 1]; ob[3 * o + 2] = b[3 * i + 2]; }
    if (c) { oc[3 * o] = c[3 * i]; oc[3 * o + 1] = c[3 * i + 1]; oc[3 * o + 2] = c[3 * i + 2]; }
    if (oidx) oidx[o] = (int32_t)i;
}

IBL_API int64_t ibl_select_rows(const uint8_t *keep, int64_t n, const double *a, const double *b, const double *c,
                                double *out_a, double *out_b, double *out_c, int32_t *out_index, void *stream) {
    IBL_ARG(n >= 0 && n < (1ll << 31));
    if (n == 0) return 0;
    IBL_ARG(keep != nullptr);
    cudaStream_t st = ibl_stream(stream);
    int *pos = nullptr;
    IBL_CUDA(ibl_alloc(&pos, (size_t)n + 1, st));
    cub::TransformInputIterator<int, U8ToInt, const uint8_t *> it(keep, U8ToInt());
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, it, pos, (int)n, st);
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tb ? tb : 1, st));
    IBL_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, it, pos, (int)n, st));
    int nb = (int)((n + 255) / 256);
    select_scatter_kernel<<<nb, 256, 0, st>>>(keep, pos, n, a, b, c, out_a, out_b, out_c, out_index);
    IBL_LAUNCH_CHECK();
    int last_pos = 0; uint8_t last_keep = 0;
    IBL_CUDA(cudaMemcpyAsync(&last_pos, pos + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(cudaMemcpyAsync(&last_keep, keep + (n - 1), 1, cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(tmp, st); ibl_free(pos, st);
    return (int64_t)last_pos + (last_keep ? 1 : 0);
}

// ------------------------------------------------------------------------------------------------
struct Mat16c { double v[16]; };
__global__ void transform_kernel(Mat16c T, const double *__restrict__ in, int64_t n, int rotate_only,
                                 double *__restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double x = in[3 * i], y = in[3 * i + 1], z = in[3 * i + 2], a, b, c;
    if (rotate_only) {
        a = __dadd_rn(__dadd_rn(__dmul_rn(T.v[0], x), __dmul_rn(T.v[1], y)), __dmul_rn(T.v[2], z));
        b = __dadd_rn(__dadd_rn(__dmul_rn(T.v[4], x), __dmul_rn(T.v[5], y)), __dmul_rn(T.v[6], z));
        c = __dadd_rn(__dadd_rn(__dmul_rn(T.v[8], x), __dmul_rn(T.v[9], y)), __dmul_rn(T.v[10], z));
    } else {
        ibl_xform(T.v, x, y, z, a, b, c);
    }
    out[3 * i] = a; out[3 * i + 1] = b; out[3 * i + 2] = c;
}
IBL_API int ibl_transform_points(const double *h_T16, const double *in, int64_t n, int rotate_only, double *out,
                                 void *stream) {
    IBL_ARG(h_T16 && n >= 0);
    if (n == 0) return IBL_OK;
    Mat16c T;
    memcpy(T.v, h_T16, sizeof(T.v));
    int nb = (int)((n + 255) / 256);
    transform_kernel<<<nb, 256, 0, ibl_stream(stream)>>>(T, in, n, rotate_only, out);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// Column means in EXACTLY numpy's order: np.mean(points, axis=0) on a C-contiguous [N,3] array adds the rows one by
// one (pairwise summation is only used along the contiguous axis), so the sum is a serial chain of N rounded adds
// per column.  One CTA: all threads stage a tile of rows in shared memory (coalesced), then three threads run the
// three dependent add chains.  ~8 cycles per row; N = 3e5 rows -> ~1.3 ms, N = 3e4 -> 0.13 ms.
#define MEAN_TILE 1024
__device__ __forceinline__ void seqsum3_block(const double *__restrict__ pts, int64_t n, double *__restrict__ out) {
    __shared__ double tile[2][MEAN_TILE * 3];
    double acc = 0.0;
    int64_t ntiles = (n + MEAN_TILE - 1) / MEAN_TILE;
    {
        int64_t cnt = min((int64_t)MEAN_TILE, n) * 3;
        for (int64_t e = threadIdx.x; e < cnt; e += blockDim.x) tile[0][e] = pts[e];
    }
    __syncthreads();
    for (int64_t t = 0; t < ntiles; ++t) {
        int cur = (int)(t & 1);
        int64_t base = t * MEAN_TILE;
        int rows = (int)min((int64_t)MEAN_TILE, n - base);
        if (threadIdx.x < 3) {                       // the serial chains (warp 0, lanes 0..2)
            const double *col = tile[cur] + threadIdx.x;
#pragma unroll 8
            for (int r = 0; r < rows; ++r) acc = __dadd_rn(acc, col[3 * r]);
        } else if (t + 1 < ntiles) {                 // everyone else prefetches the next tile meanwhile
            int64_t nb = (t + 1) * MEAN_TILE;
            int64_t cnt = min((int64_t)MEAN_TILE, n - nb) * 3;
            for (int64_t e = threadIdx.x - 3; e < cnt; e += blockDim.x - 3) tile[cur ^ 1][e] = pts[3 * nb + e];
        }
        __syncthreads();
    }
    if (threadIdx.x < 3) out[threadIdx.x] = acc;
}
__global__ void __launch_bounds__(256) seqsum3_kernel(const double *__restrict__ pts, int64_t n, double *__restrict__ out) {
    seqsum3_block(pts, n, out);
}
// the same for several clouds at once: CTA c sums cloud c (the chains of different clouds run side by side)
struct SeqSumJob { const double *pts; int64_t n; };
__global__ void __launch_bounds__(256) seqsum3_many_kernel(const SeqSumJob *__restrict__ jobs, double *__restrict__ out) {
    const SeqSumJob j = jobs[blockIdx.x];
    seqsum3_block(j.pts, j.n, out + 3 * (size_t)blockIdx.x);
}
IBL_API int ibl_mean3(const double *pts, int64_t n, double *h_mean, void *stream) {
    IBL_ARG(h_mean && n >= 0);
    h_mean[0] = h_mean[1] = h_mean[2] = 0.0;
    if (n == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    double *d = nullptr;
    IBL_CUDA(ibl_alloc(&d, 3, st));
    {
        IBL_PROF(IBL_PROF_MEAN, st, 24.0 * n);
        seqsum3_kernel<<<1, 256, 0, st>>>(pts, n, d);
        IBL_LAUNCH_CHECK();
    }
    double hs[3];
    IBL_CUDA(cudaMemcpyAsync(hs, d, sizeof(hs), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(d, st);
    for (int a = 0; a < 3; ++a) h_mean[a] = hs[a] / (double)n;
    return IBL_OK;
}

// compute_center / get_center of k clouds in ONE launch and ONE synchronisation (object_memory.py:1030-1031 takes the mean
// of every assignment's detected union): ptrs / ns are host arrays of device pointers and row counts, h_means host
// double[k][3].  Every mean equals ibl_mean3's.
IBL_API int ibl_mean3_many(const double *const *ptrs, const int64_t *ns, int32_t k, double *h_means, void *stream) {
    IBL_ARG(k >= 0 && (k == 0 || (ptrs && ns && h_means)));
    if (k == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    std::vector<SeqSumJob> jobs(k);
    double work = 0;
    for (int c = 0; c < k; ++c) {
        IBL_ARG(ns[c] >= 0 && (ns[c] == 0 || ptrs[c]));
        jobs[c].pts = ptrs[c]; jobs[c].n = ns[c];
        work += 24.0 * ns[c];
    }
    SeqSumJob *dj = nullptr; double *d = nullptr;
    IBL_CUDA(ibl_alloc(&dj, (size_t)k, st));
    IBL_CUDA(ibl_alloc(&d, (size_t)3 * k, st));
    IBL_CUDA(cudaMemcpyAsync(dj, jobs.data(), sizeof(SeqSumJob) * k, cudaMemcpyHostToDevice, st));
    {
        IBL_PROF(IBL_PROF_MEAN, st, work);
        seqsum3_many_kernel<<<(unsigned)k, 256, 0, st>>>(dj, d);
        IBL_LAUNCH_CHECK();
    }
    std::vector<double> hs((size_t)3 * k);
    IBL_CUDA(cudaMemcpyAsync(hs.data(), d, sizeof(double) * 3 * k, cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(dj, st); ibl_free(d, st);
    for (int c = 0; c < k; ++c)
        for (int a = 0; a < 3; ++a) h_means[3 * c + a] = ns[c] > 0 ? hs[3 * c + a] / (double)ns[c] : 0.0;
    return IBL_OK;
}
