// match.cu -- exact 33-d nearest-neighbour feature matching + mutual filter (Open3D
// RegistrationRANSACBasedOnFeatureMatching front half, utils/fpfh_register.py:110-112; A.5).
//
// Two stages, result identical to an f64 brute force with lowest-index tie-break:
//   filter : fp32 |a|^2 + |b|^2 - 2 a.b over all pairs (FFMA, two queries per thread, target tiles broadcast from
//            shared memory, targets split over CTAs so the grid fills the 148 SMs).
//   exact  : a target whose fp32 distance is within the rigorous fp32 error band of the running minimum (or of the
//            exact best so far) is re-evaluated on the spot in f64 with the oracle's accumulation order (j = 0..32,
//            one rounding per operation).  Everything outside the band provably cannot be, or tie, the nearest
//            neighbour.  Slices are scanned in ascending index order and merged in slice order, so exact ties
//            resolve to the lowest index.  (FPFH of planar patches produces thousands of near-identical rows: the
//            band test keeps that case cheap without any list that could overflow.)
#include <cub/cub.cuh>

#include "ibl_common.cuh"

#define FQ 2          // queries per thread
#define FT 128        // threads per CTA
#define FTILE 64      // targets per shared-memory tile
#define FROW 36       // floats per converted row: 33 features, |row|^2, 2 pad (144 B, float4 aligned)
#define FCAND 4

struct FSplitOut { double d; int idx; int pad; };   // exact best of one target slice for one query

// f64 [n][33] -> f32 [n][36] (+ squared norm of the ROUNDED row), max norm over rows via atomicMax on float bits
__global__ void feat_convert_kernel(const double *__restrict__ f, int64_t n, float *__restrict__ out,
                                    unsigned *__restrict__ maxnorm) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double s = 0.0;
    float *o = out + i * FROW;
    for (int j = 0; j < 33; ++j) {
        float v = (float)f[i * 33 + j];
        o[j] = v;
        s += (double)v * (double)v;
    }
    float nf = (float)s;
    o[33] = nf; o[34] = 0.f; o[35] = 0.f;
    atomicMax(maxnorm, __float_as_uint(nf));
}

// fp32 error of d32 = (|a|^2 + |b|^2) - 2 a.b against the exact f64 distance of the f64 rows:
//   input rounding 2.4e-7, norms 1.2e-7, 33-term FMA chain 2.0e-6, final adds 1.8e-7  (all x (|a|^2 + |b|^2))
// => |d32 - d| <= 2.5e-6 (|a|^2 + |b|^2).  FEAT_ERR keeps a 1.6x margin; eps = 2 * err.
#define FEAT_ERR 4.0e-6f
__device__ __forceinline__ float feat_err(float an, float bmax) { return FEAT_ERR * (an + bmax) + 5e-6f; }

// the oracle's exact distance: s = sum_j (a_j - b_j)^2, j ascending, one rounding per operation
__device__ __noinline__ double feat_exact(const double *__restrict__ a, const double *__restrict__ b) {
    double s = 0.0;
#pragma unroll 11
    for (int j = 0; j < 33; ++j) { double d = __dsub_rn(a[j], b[j]); s = __dadd_rn(s, __dmul_rn(d, d)); }
    return s;
}

__global__ void __launch_bounds__(FT) feat_filter_kernel(const float *__restrict__ A, int64_t na,
                                                         const float *__restrict__ B, int64_t nb,
                                                         const double *__restrict__ fa, const double *__restrict__ fb,
                                                         int nsplit, int64_t slice, const unsigned *__restrict__ bmax_bits,
                                                         FSplitOut *__restrict__ out) {
    __shared__ __align__(16) float sb[FTILE * FROW];
    const int split = blockIdx.y;
    const int64_t t_begin = (int64_t)split * slice, t_end = min(nb, t_begin + slice);
    const float bmax = __uint_as_float(*bmax_bits);
    float a[FQ][33], an[FQ], err[FQ], thr[FQ], best32[FQ];
    double bestd[FQ];
    int bi[FQ];
    int64_t q[FQ];
#pragma unroll
    for (int u = 0; u < FQ; ++u) {
        q[u] = ((int64_t)blockIdx.x * FT + threadIdx.x) * FQ + u;
        bool live = q[u] < na;
#pragma unroll
        for (int j = 0; j < 33; ++j) a[u][j] = live ? A[q[u] * FROW + j] : 0.f;
        an[u] = live ? A[q[u] * FROW + 33] : 0.f;
        err[u] = feat_err(an[u], bmax);
        best32[u] = INFINITY; thr[u] = live ? INFINITY : -INFINITY;
        bestd[u] = INFINITY; bi[u] = -1;
    }
    for (int64_t t0 = t_begin; t0 < t_end; t0 += FTILE) {
        int tn = (int)min((int64_t)FTILE, t_end - t0);
        __syncthreads();
        {
            const float4 *src = reinterpret_cast<const float4 *>(B + t0 * FROW);
            float4 *dst = reinterpret_cast<float4 *>(sb);
            for (int e = threadIdx.x; e < tn * (FROW / 4); e += FT) dst[e] = src[e];
            for (int e = tn * (FROW / 4) + threadIdx.x; e < FTILE * (FROW / 4); e += FT) dst[e] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();
        // rows past tn are zero padding; they are masked by the (t + v < tn) test below
        for (int t = 0; t < tn; t += 4) {
            float s[FQ][4];
#pragma unroll
            for (int u = 0; u < FQ; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) s[u][v] = 0.f;
            float bn[4];
#pragma unroll
            for (int j4 = 0; j4 < 9; ++j4) {
                float4 b4[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) b4[v] = *reinterpret_cast<const float4 *>(&sb[(t + v) * FROW + 4 * j4]);
                if (j4 < 8) {
#pragma unroll
                    for (int u = 0; u < FQ; ++u)
#pragma unroll
                        for (int v = 0; v < 4; ++v) {
                            s[u][v] = fmaf(a[u][4 * j4 + 0], b4[v].x, s[u][v]);
                            s[u][v] = fmaf(a[u][4 * j4 + 1], b4[v].y, s[u][v]);
                            s[u][v] = fmaf(a[u][4 * j4 + 2], b4[v].z, s[u][v]);
                            s[u][v] = fmaf(a[u][4 * j4 + 3], b4[v].w, s[u][v]);
                        }
                } else {
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        bn[v] = b4[v].y;
#pragma unroll
                        for (int u = 0; u < FQ; ++u) s[u][v] = fmaf(a[u][32], b4[v].x, s[u][v]);
                    }
                }
            }
#pragma unroll
            for (int v = 0; v < 4; ++v) {
#pragma unroll
                for (int u = 0; u < FQ; ++u) {
                    float d = (an[u] + bn[v]) - 2.f * s[u][v];
                    // thr = min(best32 + 2 err, exact best + err): only such a target can be, or tie, the exact NN
                    if (d <= thr[u] && t + v < tn) {
                        int64_t ti = t0 + t + v;
                        double de = feat_exact(fa + q[u] * 33, fb + ti * 33);
                        if (de < bestd[u]) { bestd[u] = de; bi[u] = (int)ti; }       // ascending scan: ties keep the lower index
                        best32[u] = fminf(best32[u], d);
                        thr[u] = fminf(best32[u] + 2.f * err[u], __double2float_ru(bestd[u]) + err[u]);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int u = 0; u < FQ; ++u) {
        if (q[u] < na) {
            FSplitOut o;
            o.d = bestd[u]; o.idx = bi[u]; o.pad = 0;
            out[q[u] * nsplit + split] = o;
        }
    }
}

__global__ void feat_merge_kernel(const FSplitOut *__restrict__ parts, int64_t na, int nsplit, int32_t *__restrict__ nn) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= na) return;
    double bestd = INFINITY;
    int bi = -1;
    for (int s = 0; s < nsplit; ++s) {                       // slices are in ascending index order
        FSplitOut p = parts[i * nsplit + s];
        if (p.idx >= 0 && p.d < bestd) { bestd = p.d; bi = p.idx; }
    }
    nn[i] = bi;
}

IBL_API int ibl_feature_nn(const double *fa, int64_t na, const double *fb, int64_t nb, int32_t *nn, void *stream) {
    IBL_ARG(na >= 0 && nb >= 0 && na < (1ll << 31) && nb < (1ll << 31));
    if (na == 0) return IBL_OK;
    IBL_ARG(fa && nn && (nb == 0 || fb));
    cudaStream_t st = ibl_stream(stream);
    if (nb == 0) { IBL_CUDA(cudaMemsetAsync(nn, 0xff, sizeof(int32_t) * (size_t)na, st)); return IBL_OK; }
    IBL_PROF(IBL_PROF_FEATURE_NN, st, 2.0 * 33.0 * (double)na * (double)nb);     // work unit: fp32 flop of the filter
    float *A = nullptr, *B = nullptr;
    unsigned *mx = nullptr;
    IBL_CUDA(ibl_alloc(&A, (size_t)na * FROW, st));
    IBL_CUDA(ibl_alloc(&B, (size_t)nb * FROW, st));
    IBL_CUDA(ibl_alloc(&mx, 2, st));
    IBL_CUDA(cudaMemsetAsync(mx, 0, 2 * sizeof(unsigned), st));
    feat_convert_kernel<<<(unsigned)((na + 127) / 128), 128, 0, st>>>(fa, na, A, mx);
    IBL_LAUNCH_CHECK();
    feat_convert_kernel<<<(unsigned)((nb + 127) / 128), 128, 0, st>>>(fb, nb, B, mx + 1);
    IBL_LAUNCH_CHECK();
    int qblocks = (int)((na + FT * FQ - 1) / (FT * FQ));
    int nsplit = (4 * IBL_SMS + qblocks - 1) / qblocks;          // ~4 CTAs per SM
    int64_t max_split = (nb + FTILE - 1) / FTILE;
    if (nsplit > max_split) nsplit = (int)max_split;
    if (nsplit > 64) nsplit = 64;
    if (nsplit < 1) nsplit = 1;
    int64_t slice = (nb + nsplit - 1) / nsplit;
    slice = (slice + FTILE - 1) / FTILE * FTILE;
    nsplit = (int)((nb + slice - 1) / slice);
    FSplitOut *parts = nullptr;
    IBL_CUDA(ibl_alloc(&parts, (size_t)na * nsplit, st));
    dim3 grid(qblocks, nsplit);
    feat_filter_kernel<<<grid, FT, 0, st>>>(A, na, B, nb, fa, fb, nsplit, slice, mx + 1, parts);
    IBL_LAUNCH_CHECK();
    feat_merge_kernel<<<(unsigned)((na + 127) / 128), 128, 0, st>>>(parts, na, nsplit, nn);
    IBL_LAUNCH_CHECK();
    ibl_free(A, st); ibl_free(B, st); ibl_free(mx, st); ibl_free(parts, st);
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
