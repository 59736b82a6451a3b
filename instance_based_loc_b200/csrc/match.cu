// match.cu -- exact 33-d nearest-neighbour feature matching + mutual filter (Open3D
// RegistrationRANSACBasedOnFeatureMatching front half, utils/fpfh_register.py:110-112; A.5).
//
// The nearest-neighbour search itself is in match_tc.cu (tcgen05 fp16-split filter + fp32-difference filter + exact f64 re-evaluation, result identical
// to an f64 brute force with lowest-index tie-break); this file holds the entry point and the mutual filter.
#include <cub/cub.cuh>

#include "ibl_common.cuh"

int ibl_feature_nn_tc(const double *fa, int64_t na, const double *fb, int64_t nb, int32_t *nn, cudaStream_t st);

IBL_API int ibl_feature_nn(const double *fa, int64_t na, const double *fb, int64_t nb, int32_t *nn, void *stream) {
    IBL_ARG(na >= 0 && nb >= 0 && na < (1ll << 31) && nb < (1ll << 31));
    if (na == 0) return IBL_OK;
    IBL_ARG(fa && nn && (nb == 0 || fb));
    cudaStream_t st = ibl_stream(stream);
    if (nb == 0) { IBL_CUDA(cudaMemsetAsync(nn, 0xff, sizeof(int32_t) * (size_t)na, st)); return IBL_OK; }
    IBL_PROF(IBL_PROF_FEATURE_NN, st, 2.0 * 33.0 * (double)na * (double)nb);     // work unit: flop of the 33-d distance matrix
    return ibl_feature_nn_tc(fa, na, fb, nb, nn, st);                            // tcgen05 filter + exact f64 (match_tc.cu)
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
    IBL_CUDA(ibl_sync(st));
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
