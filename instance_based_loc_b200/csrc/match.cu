// match.cu -- exact 33-d nearest-neighbour feature matching + mutual filter (Open3D
// RegistrationRANSACBasedOnFeatureMatching front half, utils/fpfh_register.py:110-112; A.5).
// Round-1 kernel: f64 brute force with shared-memory target tiles and partial-distance early-out;
// accumulation order j = 0..32 matches the oracle so the argmin (lowest index on ties) is exact.
#include <cub/cub.cuh>

#include "ibl_common.cuh"

#define FNN_TILE 64
#define FNN_THREADS 128

__global__ void __launch_bounds__(FNN_THREADS) feature_nn_kernel(const double *__restrict__ fa, int64_t na,
                                                                 const double *__restrict__ fb, int64_t nb,
                                                                 int32_t *__restrict__ nn) {
    __shared__ double tile[FNN_TILE * 33];
    int64_t i = blockIdx.x * (int64_t)FNN_THREADS + threadIdx.x;
    double a[33];
    bool live = i < na;
#pragma unroll
    for (int j = 0; j < 33; ++j) a[j] = live ? fa[i * 33 + j] : 0.0;
    double best = INFINITY;
    int bi = -1;
    for (int64_t t0 = 0; t0 < nb; t0 += FNN_TILE) {
        int tn = (int)min((int64_t)FNN_TILE, nb - t0);
        __syncthreads();
        for (int e = threadIdx.x; e < tn * 33; e += FNN_THREADS) tile[e] = fb[t0 * 33 + e];
        __syncthreads();
        if (!live) continue;
        for (int q = 0; q < tn; ++q) {
            const double *b = tile + q * 33;
            double s = 0.0;
#pragma unroll
            for (int j = 0; j < 11; ++j) { double d = __dsub_rn(a[j], b[j]); s = __dadd_rn(s, __dmul_rn(d, d)); }
            if (s >= best) continue;      // partial distance already not smaller: cannot win
#pragma unroll
            for (int j = 11; j < 22; ++j) { double d = __dsub_rn(a[j], b[j]); s = __dadd_rn(s, __dmul_rn(d, d)); }
            if (s >= best) continue;
#pragma unroll
            for (int j = 22; j < 33; ++j) { double d = __dsub_rn(a[j], b[j]); s = __dadd_rn(s, __dmul_rn(d, d)); }
            if (s < best) { best = s; bi = (int)(t0 + q); }
        }
    }
    if (live) nn[i] = bi;
}

IBL_API int ibl_feature_nn(const double *fa, int64_t na, const double *fb, int64_t nb, int32_t *nn, void *stream) {
    IBL_ARG(na >= 0 && nb >= 0);
    if (na == 0) return IBL_OK;
    IBL_ARG(fa && nn && (nb == 0 || fb));
    IBL_PROF(IBL_PROF_FEATURE_NN, ibl_stream(stream), 3.0 * 33.0 * (double)na * (double)nb);
    feature_nn_kernel<<<(unsigned)((na + FNN_THREADS - 1) / FNN_THREADS), FNN_THREADS, 0, ibl_stream(stream)>>>(fa, na, fb, nb, nn);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

__global__ void mutual_flag_kernel(const int32_t *__restrict__ ij, int64_t ns, const int32_t *__restrict__ ji, int64_t nt,
                                   uint8_t *__restrict__ flag) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= ns) return;
    int j = ij[i];
    flag[i] = (j >= 0 && j < nt && ji[j] == (int32_t)i);
}
__global__ void corres_fill_kernel(const int32_t *__restrict__ sel, int64_t c, const int32_t *__restrict__ ij,
                                   int32_t *__restrict__ corres) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= c) return;
    int i = sel ? sel[k] : (int)k;
    corres[2 * k] = i;
    corres[2 * k + 1] = ij[i];
}

IBL_API int64_t ibl_mutual_corres(const int32_t *ij, int64_t ns, const int32_t *ji, int64_t nt, int32_t ransac_n,
                                  int32_t *corres, void *stream) {
    IBL_ARG(ns >= 0 && nt >= 0 && ns < (1ll << 31));
    if (ns == 0) return 0;
    IBL_ARG(ij && ji && corres);
    cudaStream_t st = ibl_stream(stream);
    uint8_t *flag = nullptr; int32_t *sel = nullptr; int *dcount = nullptr;
    IBL_CUDA(ibl_alloc(&flag, (size_t)ns, st));
    IBL_CUDA(ibl_alloc(&sel, (size_t)ns, st));
    IBL_CUDA(ibl_alloc(&dcount, 1, st));
    int nb = (int)((ns + 255) / 256);
    mutual_flag_kernel<<<nb, 256, 0, st>>>(ij, ns, ji, nt, flag);
    IBL_LAUNCH_CHECK();
    cub::CountingInputIterator<int> cit(0);
    size_t tb = 0;
    cub::DeviceSelect::Flagged(nullptr, tb, cit, flag, sel, dcount, (int)ns, st);
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tb ? tb : 1, st));
    IBL_CUDA(cub::DeviceSelect::Flagged(tmp, tb, cit, flag, sel, dcount, (int)ns, st));
    int c = 0;
    IBL_CUDA(cudaMemcpyAsync(&c, dcount, sizeof(int), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(cudaStreamSynchronize(st));
    int64_t out_c;
    if (c >= ransac_n * 3) {
        corres_fill_kernel<<<(c + 255) / 256, 256, 0, st>>>(sel, c, ij, corres);
        out_c = c;
    } else {   // "Too few correspondences after mutual filter, fall back to original correspondences."
        corres_fill_kernel<<<nb, 256, 0, st>>>(nullptr, ns, ij, corres);
        out_c = ns;
    }
    IBL_LAUNCH_CHECK();
    ibl_free(tmp, st); ibl_free(flag, st); ibl_free(sel, st); ibl_free(dcount, st);
    return out_c;
}
