// voxel.cu -- voxel_down_sample_with_colors (utils/depth_utils.py:211-265): key = floor(p/voxel)
// as int64 per axis, voxels emitted in first-occurrence order, per-voxel means accumulated in
// input order (== np.mean over the python list the reference builds).
#include <cub/cub.cuh>

#include "ibl_common.cuh"


struct VoxRange { long long kmin[3]; unsigned long long r0, r1; };

__global__ void vox_key_kernel(const double *__restrict__ pts, int64_t n, double voxel, VoxRange vr,
                               unsigned long long *__restrict__ key, int *__restrict__ idx, int64_t *__restrict__ out_keys) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    long long k[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        k[a] = (long long)floor(__ddiv_rn(pts[3 * i + a], voxel));
        if (out_keys) out_keys[3 * i + a] = k[a];
    }
    unsigned long long ux = (unsigned long long)(k[0] - vr.kmin[0]);
    unsigned long long uy = (unsigned long long)(k[1] - vr.kmin[1]);
    unsigned long long uz = (unsigned long long)(k[2] - vr.kmin[2]);
    key[i] = (uz * vr.r1 + uy) * vr.r0 + ux;
    idx[i] = (int)i;
}

__global__ void vox_head_kernel(const unsigned long long *__restrict__ key, int64_t n, uint8_t *__restrict__ flag) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (i == 0) || (key[i] != key[i - 1]);
}

__global__ void vox_first_kernel(const int *__restrict__ head_pos, int nseg, const int *__restrict__ idx_sorted,
                                 int *__restrict__ first_idx, int *__restrict__ seg_id) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nseg) { first_idx[s] = idx_sorted[head_pos[s]]; seg_id[s] = s; }
}

__global__ void vox_mean_kernel(const int *__restrict__ seg_by_rank, const int *__restrict__ head_pos, int nseg, int64_t n,
                                const int *__restrict__ idx_sorted, const double *__restrict__ pts,
                                const double *__restrict__ cols, const double *__restrict__ nrm,
                                double *__restrict__ out_pts, double *__restrict__ out_cols, double *__restrict__ out_nrm,
                                int32_t *__restrict__ vox_id) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nseg) return;
    int s = seg_by_rank[v];
    int p0 = head_pos[s], p1 = (s + 1 < nseg) ? head_pos[s + 1] : (int)n;
    double cnt = (double)(p1 - p0);
    double a[3] = {0, 0, 0}, c[3] = {0, 0, 0}, m[3] = {0, 0, 0};
    for (int p = p0; p < p1; ++p) {
        size_t i = (size_t)idx_sorted[p];
        if (p == p0) {
            for (int k = 0; k < 3; ++k) { a[k] = pts[3 * i + k]; if (cols) c[k] = cols[3 * i + k]; if (nrm) m[k] = nrm[3 * i + k]; }
        } else {
            for (int k = 0; k < 3; ++k) {
                a[k] = __dadd_rn(a[k], pts[3 * i + k]);
                if (cols) c[k] = __dadd_rn(c[k], cols[3 * i + k]);
                if (nrm) m[k] = __dadd_rn(m[k], nrm[3 * i + k]);
            }
        }
        if (vox_id) vox_id[i] = v;
    }
    for (int k = 0; k < 3; ++k) {
        out_pts[3 * (size_t)v + k] = __ddiv_rn(a[k], cnt);
        if (cols) out_cols[3 * (size_t)v + k] = __ddiv_rn(c[k], cnt);
    }
    if (nrm) {
        for (int k = 0; k < 3; ++k) m[k] = __ddiv_rn(m[k], cnt);
        double nn = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(m[0], m[0]), __dmul_rn(m[1], m[1])), __dmul_rn(m[2], m[2])));
        for (int k = 0; k < 3; ++k) out_nrm[3 * (size_t)v + k] = __ddiv_rn(m[k], nn);
    }
}

IBL_API int64_t ibl_voxel_downsample(const double *pts, const double *cols, const double *nrm, int64_t n, double voxel,
                                     double *out_pts, double *out_cols, double *out_nrm, int32_t *vox_id,
                                     int64_t *out_keys, void *stream) {
    IBL_ARG(n >= 0 && n < (1ll << 31) && voxel > 0);
    if (n == 0) return 0;
    IBL_ARG(pts && out_pts);
    cudaStream_t st = ibl_stream(stream);
    // bounding box -> key range (floor(p/voxel) is monotone in p)
    int nbb = (int)((n + 255) / 256);
    if (nbb > 4 * IBL_SMS) nbb = 4 * IBL_SMS;
    double *part = nullptr;
    IBL_CUDA(ibl_alloc(&part, (size_t)nbb * 6, st));
    bbox_kernel<<<nbb, 256, 0, st>>>(pts, n, part);
    IBL_LAUNCH_CHECK();
    double *hp = (double *)malloc(sizeof(double) * nbb * 6);
    IBL_CUDA(cudaMemcpyAsync(hp, part, sizeof(double) * nbb * 6, cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int b = 0; b < nbb; ++b)
        for (int a = 0; a < 3; ++a) { mn[a] = fmin(mn[a], hp[b * 6 + a]); mx[a] = fmax(mx[a], hp[b * 6 + 3 + a]); }
    free(hp);
    ibl_free(part, st);
    VoxRange vr;
    unsigned long long r[3];
    int bits = 0;
    for (int a = 0; a < 3; ++a) {
        if (!isfinite(mn[a]) || !isfinite(mx[a])) { ibl_set_error("non-finite coordinates"); return IBL_E_RANGE; }
        double lo = floor(mn[a] / voxel), hi = floor(mx[a] / voxel);
        if (fabs(lo) > 9e15 || fabs(hi) > 9e15 || hi - lo >= 2097152.0) { ibl_set_error("voxel key range too large"); return IBL_E_RANGE; }
        vr.kmin[a] = (long long)lo;
        r[a] = (unsigned long long)((long long)hi - (long long)lo + 1);
        int b = 0;
        while ((1ull << b) < r[a]) ++b;
        bits += b;
    }
    vr.r0 = r[0]; vr.r1 = r[1];
    int total_bits = 1;
    {
        // exact bit width of the largest packed key
        long double maxkey = ((long double)(r[2] - 1) * r[1] + (r[1] - 1)) * r[0] + (r[0] - 1);
        while (total_bits < 64 && (long double)(1ull << total_bits) <= maxkey) ++total_bits;
    }
    (void)bits;
    unsigned long long *key = nullptr, *key2 = nullptr;
    int *idx = nullptr, *idx2 = nullptr, *head_pos = nullptr, *first_idx = nullptr, *first2 = nullptr, *seg = nullptr, *seg2 = nullptr;
    uint8_t *flag = nullptr;
    int *d_nseg = nullptr;
    IBL_CUDA(ibl_alloc(&key, (size_t)n, st)); IBL_CUDA(ibl_alloc(&key2, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&idx, (size_t)n, st)); IBL_CUDA(ibl_alloc(&idx2, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&flag, (size_t)n, st)); IBL_CUDA(ibl_alloc(&head_pos, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&d_nseg, 1, st));
    int nb = (int)((n + 255) / 256);
    IBL_PROF(IBL_PROF_VOXEL, st, 72.0 * n);
    vox_key_kernel<<<nb, 256, 0, st>>>(pts, n, voxel, vr, key, idx, out_keys);
    IBL_LAUNCH_CHECK();
    size_t tb = 0, tb2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tb, key, key2, idx, idx2, (int)n, 0, total_bits, st);
    cub::CountingInputIterator<int> cit(0);
    cub::DeviceSelect::Flagged(nullptr, tb2, cit, flag, head_pos, d_nseg, (int)n, st);
    size_t tmax = tb > tb2 ? tb : tb2;
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tmax ? tmax : 1, st));
    IBL_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, key, key2, idx, idx2, (int)n, 0, total_bits, st));
    vox_head_kernel<<<nb, 256, 0, st>>>(key2, n, flag);
    IBL_LAUNCH_CHECK();
    IBL_CUDA(cub::DeviceSelect::Flagged(tmp, tb2, cit, flag, head_pos, d_nseg, (int)n, st));
    int nseg = 0;
    IBL_CUDA(cudaMemcpyAsync(&nseg, d_nseg, sizeof(int), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    IBL_CUDA(ibl_alloc(&first_idx, (size_t)nseg, st)); IBL_CUDA(ibl_alloc(&first2, (size_t)nseg, st));
    IBL_CUDA(ibl_alloc(&seg, (size_t)nseg, st)); IBL_CUDA(ibl_alloc(&seg2, (size_t)nseg, st));
    int nbs = (nseg + 255) / 256;
    vox_first_kernel<<<nbs, 256, 0, st>>>(head_pos, nseg, idx2, first_idx, seg);
    IBL_LAUNCH_CHECK();
    int ibits = 1;
    while ((1ll << ibits) < n) ++ibits;
    size_t tb3 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tb3, first_idx, first2, seg, seg2, nseg, 0, ibits, st);
    void *tmp3 = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp3, tb3 ? tb3 : 1, st));
    IBL_CUDA(cub::DeviceRadixSort::SortPairs(tmp3, tb3, first_idx, first2, seg, seg2, nseg, 0, ibits, st));
    vox_mean_kernel<<<nbs, 256, 0, st>>>(seg2, head_pos, nseg, n, idx2, pts, cols, nrm, out_pts, cols ? out_cols : nullptr,
                                         nrm ? out_nrm : nullptr, vox_id);
    IBL_LAUNCH_CHECK();
    ibl_free(tmp, st); ibl_free(tmp3, st);
    ibl_free(key, st); ibl_free(key2, st); ibl_free(idx, st); ibl_free(idx2, st); ibl_free(flag, st);
    ibl_free(head_pos, st); ibl_free(d_nseg, st); ibl_free(first_idx, st); ibl_free(first2, st);
    ibl_free(seg, st); ibl_free(seg2, st);
    return (int64_t)nseg;
}
