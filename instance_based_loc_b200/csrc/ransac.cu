// ransac.cu -- RegistrationRANSACBasedOnCorrespondence (utils/fpfh_register.py:110-119; A.6/A.7).
//   check  : thread per hypothesis -- gather 3 correspondences, EdgeLength test first (needs no
//            transform), Jacobi-Kabsch, Distance test
//   score  : grid (source tiles x hypotheses); thread per source point does an exact 1-NN within
//            max_d on the target grid (ring field skips empty space); integer inlier counts and a
//            fixed-order err2 reduction (warp shuffles -> CTA -> tile partials)
//   count  : same grid as score, inlier counts only (first candidate within max_d ends a point's scan)
//   run    : host loop applying the ordered best-update / adaptive-stop rule; a chunk of validated hypotheses is
//            counted first and only those whose count reaches the running maximum get the exact score
#include <cub/cub.cuh>
#include <algorithm>
#include <vector>

#include "ibl_common.cuh"
#include "rigid.cuh"

__device__ __forceinline__ void gather_sample(const double *__restrict__ src, const double *__restrict__ tgt,
                                              const int32_t *__restrict__ corres, int64_t n_corres,
                                              const int32_t *__restrict__ samples, uint64_t seed, int64_t h, double *s,
                                              double *t) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        int c = samples ? samples[3 * h + j] : (int)ibl_sample(seed, (uint64_t)h, (uint32_t)j, (uint32_t)n_corres);
        size_t cs = (size_t)corres[2 * c], ct = (size_t)corres[2 * c + 1];
#pragma unroll
        for (int a = 0; a < 3; ++a) { s[3 * j + a] = src[3 * cs + a]; t[3 * j + a] = tgt[3 * ct + a]; }
    }
}

// one hypothesis: EdgeLength test (needs no transform), Jacobi-Kabsch, Distance test.  T is valid when true is returned.
__device__ __forceinline__ bool ransac_check_one(const double *__restrict__ src, const double *__restrict__ tgt,
                                                 const int32_t *__restrict__ corres, int64_t n_corres,
                                                 const int32_t *__restrict__ samples, uint64_t seed, int64_t h,
                                                 double edge_thr, double dist_thr, int use_checkers, double *T) {
    double s[9], t[9];
    gather_sample(src, tgt, corres, n_corres, samples, seed, h, s, t);
    if (use_checkers) {
        for (int i = 0; i < 3; ++i)
            for (int j = i + 1; j < 3; ++j) {
                double a[3] = {s[3 * i] - s[3 * j], s[3 * i + 1] - s[3 * j + 1], s[3 * i + 2] - s[3 * j + 2]};
                double b[3] = {t[3 * i] - t[3 * j], t[3 * i + 1] - t[3 * j + 1], t[3 * i + 2] - t[3 * j + 2]};
                double ds = sqrt(ibl_dot(a, a)), dt = sqrt(ibl_dot(b, b));
                if (ds < dt * edge_thr || dt < ds * edge_thr) return false;
            }
    }
    ibl_kabsch3(s, t, T);
    if (use_checkers) {
        for (int j = 0; j < 3; ++j) {
            double qx, qy, qz;
            ibl_xform(T, s[3 * j], s[3 * j + 1], s[3 * j + 2], qx, qy, qz);
            double d[3] = {t[3 * j] - qx, t[3 * j + 1] - qy, t[3 * j + 2] - qz};
            if (sqrt(ibl_dot(d, d)) > dist_thr) return false;
        }
    }
    return true;
}

__global__ void __launch_bounds__(128) ransac_check_kernel(const double *__restrict__ src, const double *__restrict__ tgt,
                                                           const int32_t *__restrict__ corres, int64_t n_corres,
                                                           const int32_t *__restrict__ samples, uint64_t seed, int64_t h0,
                                                           int64_t H, double edge_thr, double dist_thr, int use_checkers,
                                                           uint8_t *__restrict__ pass, double *__restrict__ T_out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= H) return;
    double T[16];
    bool ok = ransac_check_one(src, tgt, corres, n_corres, samples, seed, h0 + k, edge_thr, dist_thr, use_checkers, T);
    if (ok && T_out) {
#pragma unroll
        for (int e = 0; e < 16; ++e) T_out[16 * k + e] = T[e];
    }
    pass[k] = ok;
}

IBL_API int ibl_ransac_check(const double *src, const double *tgt, const int32_t *corres, int64_t n_corres,
                             const int32_t *samples, uint64_t seed, int64_t h0, int64_t H, double edge_thr,
                             double dist_thr, int use_checkers, uint8_t *pass, double *T_out, void *stream) {
    IBL_ARG(H >= 0 && n_corres >= 1 && n_corres < (1ll << 31));
    if (H == 0) return IBL_OK;
    IBL_ARG(src && tgt && corres && pass);
    IBL_PROF(IBL_PROF_RANSAC_CHECK, ibl_stream(stream), 157.0 * (double)H);
    ransac_check_kernel<<<(unsigned)((H + 127) / 128), 128, 0, ibl_stream(stream)>>>(
        src, tgt, corres, n_corres, samples, seed, h0, H, edge_thr, dist_thr, use_checkers, pass, T_out);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
__global__ void hyp_transform_kernel(const double *__restrict__ src, const double *__restrict__ tgt,
                                     const int32_t *__restrict__ corres, int64_t n_corres,
                                     const int32_t *__restrict__ samples, uint64_t seed,
                                     const int64_t *__restrict__ hyp_ids, int64_t nh, double *__restrict__ T_out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= nh) return;
    double s[9], t[9], T[16];
    gather_sample(src, tgt, corres, n_corres, samples, seed, hyp_ids[k], s, t);
    ibl_kabsch3(s, t, T);
#pragma unroll
    for (int e = 0; e < 16; ++e) T_out[16 * k + e] = T[e];
}

#define SCORE_TILE 512       // source points per CTA: small tiles keep the 148 SMs full when few hypotheses survive the checkers
#define SCORE_THREADS 256

__global__ void __launch_bounds__(SCORE_THREADS) ransac_score_kernel(GridView g, const double *__restrict__ src,
                                                                     int64_t n_src, const double *__restrict__ T_all,
                                                                     double max_d, double max_d2, int ntiles,
                                                                     int *__restrict__ part_cnt,
                                                                     double *__restrict__ part_err) {
    __shared__ double T[16];
    __shared__ double se[SCORE_THREADS / 32];
    __shared__ int sc[SCORE_THREADS / 32];
    int tile = blockIdx.x;
    int64_t hyp = blockIdx.y;
    if (threadIdx.x < 16) T[threadIdx.x] = T_all[16 * hyp + threadIdx.x];
    __syncthreads();
    int cnt = 0;
    double err = 0.0;
    int64_t base = (int64_t)tile * SCORE_TILE;
#pragma unroll 1
    for (int r = 0; r < SCORE_TILE / SCORE_THREADS; ++r) {
        int64_t i = base + r * SCORE_THREADS + threadIdx.x;
        if (i < n_src) {
            double x, y, z, bd;
            ibl_xform(T, src[3 * i], src[3 * i + 1], src[3 * i + 2], x, y, z);
            int j = ibl_nn1_thread(g, x, y, z, max_d, max_d2, bd);
            if (j >= 0) { ++cnt; err = __dadd_rn(err, bd); }
        }
    }
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) { err += __shfl_down_sync(0xffffffffu, err, o); cnt += __shfl_down_sync(0xffffffffu, cnt, o); }
    if (lane == 0) { se[w] = err; sc[w] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; int tc = 0;
        for (int k = 0; k < SCORE_THREADS / 32; ++k) { te += se[k]; tc += sc[k]; }
        part_cnt[hyp * ntiles + tile] = tc;
        part_err[hyp * ntiles + tile] = te;
    }
}

// Inlier COUNT only (same grid shape as ransac_score_kernel): "is any target within max_d" stops at the first candidate
// that qualifies, where the exact score has to finish the minimum over the whole candidate list.  Integer partials are
// added with atomicAdd (order-independent), so there is no finalize pass.  The counts equal ransac_score_kernel's.
__global__ void __launch_bounds__(SCORE_THREADS) ransac_count_kernel(GridView g, const double *__restrict__ src, int64_t n_src,
                                                                     const double *__restrict__ T_all, double max_d,
                                                                     double max_d2, int32_t *__restrict__ inliers) {
    __shared__ double T[16];
    __shared__ int sc[SCORE_THREADS / 32];
    int tile = blockIdx.x;
    int64_t hyp = blockIdx.y;
    if (threadIdx.x < 16) T[threadIdx.x] = T_all[16 * hyp + threadIdx.x];
    __syncthreads();
    int cnt = 0;
    int64_t base = (int64_t)tile * SCORE_TILE;
    const bool field = g.has_field && max_d <= g.f.max_d;
    const bool sure_ok = field && max_d == g.f.max_d;
#pragma unroll 1
    for (int r = 0; r < SCORE_TILE / SCORE_THREADS; ++r) {
        int64_t i = base + r * SCORE_THREADS + threadIdx.x;
        if (i < n_src) {
            double x, y, z, bd;
            ibl_xform(T, src[3 * i], src[3 * i + 1], src[3 * i + 2], x, y, z);
            if (field) cnt += ibl_any_within_field(g, x, y, z, max_d2, sure_ok) ? 1 : 0;
            else cnt += ibl_nn1_thread(g, x, y, z, max_d, max_d2, bd) >= 0 ? 1 : 0;
        }
    }
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    if (lane == 0) sc[w] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tc = 0;
        for (int k = 0; k < SCORE_THREADS / 32; ++k) tc += sc[k];
        if (tc) atomicAdd(&inliers[hyp], tc);
    }
}

// one CTA per hypothesis: tile partials (fixed order: thread-strided, warp tree, warps in order) + the correspondence
// inlier count (EvaluateInlierCorrespondenceRatio numerator).  The gathers are latency-bound, so a whole CTA per
// hypothesis rather than a warp.
#define FIN_THREADS 128
__global__ void __launch_bounds__(FIN_THREADS) ransac_finalize_kernel(const int *__restrict__ part_cnt,
                                                                      const double *__restrict__ part_err, int ntiles,
                                                                      int64_t nh, const double *__restrict__ T_all,
                                                                      const double *__restrict__ src,
                                                                      const double *__restrict__ tgt,
                                                                      const int32_t *__restrict__ corres, int64_t n_corres,
                                                                      double max_d, int32_t *__restrict__ inliers,
                                                                      double *__restrict__ err2,
                                                                      int32_t *__restrict__ corres_inliers) {
    __shared__ double se[FIN_THREADS / 32];
    __shared__ int sc[FIN_THREADS / 32], si[FIN_THREADS / 32];
    __shared__ double T[16];
    const int64_t hyp = blockIdx.x;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < 16) T[threadIdx.x] = T_all[16 * hyp + threadIdx.x];
    __syncthreads();
    double e = 0; int c = 0;
    for (int k = threadIdx.x; k < ntiles; k += FIN_THREADS) { e += part_err[hyp * ntiles + k]; c += part_cnt[hyp * ntiles + k]; }
    int ci = 0;
    if (corres_inliers) {
        const double md2 = max_d * max_d;
        for (int64_t k = threadIdx.x; k < n_corres; k += FIN_THREADS) {
            size_t cs = (size_t)corres[2 * k], ct = (size_t)corres[2 * k + 1];
            double x, y, z;
            ibl_xform(T, src[3 * cs], src[3 * cs + 1], src[3 * cs + 2], x, y, z);
            double d[3] = {x - tgt[3 * ct], y - tgt[3 * ct + 1], z - tgt[3 * ct + 2]};
            ci += ibl_dot(d, d) < md2;
        }
    }
    for (int o = 16; o; o >>= 1) {
        e += __shfl_down_sync(0xffffffffu, e, o);
        c += __shfl_down_sync(0xffffffffu, c, o);
        ci += __shfl_down_sync(0xffffffffu, ci, o);
    }
    if (lane == 0) { se[w] = e; sc[w] = c; si[w] = ci; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; int tc = 0, ti = 0;
        for (int k = 0; k < FIN_THREADS / 32; ++k) { te += se[k]; tc += sc[k]; ti += si[k]; }
        inliers[hyp] = tc; err2[hyp] = te;
        if (corres_inliers) corres_inliers[hyp] = ti;
    }
}

IBL_API int ibl_ransac_score(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                             const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                             const int64_t *hyp_ids, int64_t nh, double max_d, int32_t *inliers, double *err2,
                             int32_t *corres_inliers, double *T_out, void *stream) {
    IBL_ARG(tgt_grid && nh >= 0 && n_src >= 0 && max_d > 0 && nh <= 65535);
    if (nh == 0) return IBL_OK;
    IBL_ARG(src && tgt && corres && hyp_ids && inliers && err2 && T_out);
    cudaStream_t st = ibl_stream(stream);
    hyp_transform_kernel<<<(unsigned)((nh + 63) / 64), 64, 0, st>>>(src, tgt, corres, n_corres, samples, seed, hyp_ids, nh, T_out);
    IBL_LAUNCH_CHECK();
    int ntiles = (int)((n_src + SCORE_TILE - 1) / SCORE_TILE);
    if (ntiles < 1) ntiles = 1;
    int *pc = nullptr; double *pe = nullptr;
    IBL_CUDA(ibl_alloc(&pc, (size_t)nh * ntiles, st));
    IBL_CUDA(ibl_alloc(&pe, (size_t)nh * ntiles, st));
    dim3 grid(ntiles, (unsigned)nh);
    {
        IBL_PROF(IBL_PROF_RANSAC_SCORE, st, (double)nh * (24.0 * n_src + 128.0));
        ransac_score_kernel<<<grid, SCORE_THREADS, 0, st>>>(ibl_view(tgt_grid), src, n_src, T_out, max_d, max_d * max_d, ntiles, pc, pe);
        IBL_LAUNCH_CHECK();
    }
    ransac_finalize_kernel<<<(unsigned)nh, FIN_THREADS, 0, st>>>(pc, pe, ntiles, nh, T_out, src, tgt, corres, n_corres, max_d,
                                                                 inliers, err2, corres_inliers);
    IBL_LAUNCH_CHECK();
    ibl_free(pc, st); ibl_free(pe, st);
    return IBL_OK;
}

// Inlier counts only (the pass ibl_ransac_run uses on whole chunks): inliers[k] == ibl_ransac_score's inliers[k].
IBL_API int ibl_ransac_count(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                             const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                             const int64_t *hyp_ids, int64_t nh, double max_d, int32_t *inliers, double *T_out, void *stream) {
    IBL_ARG(tgt_grid && nh >= 0 && n_src >= 0 && max_d > 0 && nh <= 65535);
    if (nh == 0) return IBL_OK;
    IBL_ARG(src && tgt && corres && hyp_ids && inliers && T_out);
    cudaStream_t st = ibl_stream(stream);
    hyp_transform_kernel<<<(unsigned)((nh + 63) / 64), 64, 0, st>>>(src, tgt, corres, n_corres, samples, seed, hyp_ids, nh, T_out);
    IBL_LAUNCH_CHECK();
    IBL_CUDA(cudaMemsetAsync(inliers, 0, sizeof(int32_t) * nh, st));
    int ntiles = (int)((n_src + SCORE_TILE - 1) / SCORE_TILE);
    if (ntiles < 1) ntiles = 1;
    dim3 grid(ntiles, (unsigned)nh);
    IBL_PROF(IBL_PROF_RANSAC_SCORE, st, (double)nh * (24.0 * n_src + 128.0));
    ransac_count_kernel<<<grid, SCORE_THREADS, 0, st>>>(ibl_view(tgt_grid), src, n_src, T_out, max_d, max_d * max_d, inliers);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
struct IotaPlus {
    int64_t h0;
    __host__ __device__ int64_t operator()(int64_t k) const { return h0 + k; }
};

IBL_API int ibl_ransac_run(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                           const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                           int64_t max_iteration, double confidence, double max_d, double edge_thr, int use_checkers,
                           int early_exit, int64_t block, ibl_reg_result_t *h_result, void *stream) {
    return ibl_ransac_run_range(tgt_grid, src, n_src, tgt, corres, n_corres, samples, seed, 0, max_iteration, confidence,
                                max_d, edge_thr, use_checkers, early_exit, block, h_result, stream);
}

// Hypotheses [h_begin, max_iteration) of the same sample stream: the block one rank owns when a single registration
// is sharded by hypothesis over GPUs (dist.hypothesis_block).
IBL_API int ibl_ransac_run_range(const ibl_grid_t *tgt_grid, const double *src, int64_t n_src, const double *tgt,
                                 const int32_t *corres, int64_t n_corres, const int32_t *samples, uint64_t seed,
                                 int64_t h_begin, int64_t max_iteration, double confidence, double max_d, double edge_thr,
                                 int use_checkers, int early_exit, int64_t block, ibl_reg_result_t *h_result, void *stream) {
    IBL_ARG(h_result != nullptr && tgt_grid != nullptr && h_begin >= 0);
    memset(h_result, 0, sizeof(*h_result));
    for (int k = 0; k < 16; ++k) h_result->T[k] = (k % 5 == 0) ? 1.0 : 0.0;
    h_result->best_hyp = -1;
    h_result->n_corres = n_corres;
    if (n_corres < 3 || max_d <= 0 || max_iteration <= 0 || n_src <= 0) { h_result->status = 1; return IBL_OK; }
    cudaStream_t st = ibl_stream(stream);
    if (block <= 0) block = 1 << 14;
    // Blocks grow geometrically (x2 per round, capped): few host round trips for the common 1e5..1e6-draw case while
    // the work wasted past the adaptive stop stays bounded by one score chunk (survivors are scored in id order).
    const int64_t MAX_BLOCK = 1 << 20;
    const int64_t first_block = block;
    block = MAX_BLOCK;                     // buffers are sized for the largest block
    const int64_t SCORE_CHUNK = 512;
    uint8_t *pass = nullptr; int64_t *ids = nullptr; int *dcount = nullptr;
    int32_t *inl = nullptr, *inl2 = nullptr, *cinl = nullptr; double *err = nullptr, *Ts = nullptr; int64_t *sel_ids = nullptr;
    IBL_CUDA(ibl_alloc(&pass, (size_t)block, st));
    IBL_CUDA(ibl_alloc(&ids, (size_t)block, st));
    IBL_CUDA(ibl_alloc(&dcount, 1, st));
    IBL_CUDA(ibl_alloc(&inl, (size_t)SCORE_CHUNK, st));
    IBL_CUDA(ibl_alloc(&inl2, (size_t)SCORE_CHUNK, st));
    IBL_CUDA(ibl_alloc(&sel_ids, (size_t)SCORE_CHUNK, st));
    IBL_CUDA(ibl_alloc(&cinl, (size_t)SCORE_CHUNK, st));
    IBL_CUDA(ibl_alloc(&err, (size_t)SCORE_CHUNK, st));
    IBL_CUDA(ibl_alloc(&Ts, (size_t)SCORE_CHUNK * 16, st));
    size_t tb = 0;
    {
        cub::CountingInputIterator<int64_t> cit(0);
        cub::DeviceSelect::Flagged(nullptr, tb, cit, pass, ids, dcount, (int)block, st);
    }
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tb ? tb : 1, st));
    std::vector<int64_t> h_ids(SCORE_CHUNK), h_sel(SCORE_CHUNK), h_sel_ids(SCORE_CHUNK);
    std::vector<int32_t> h_inl(SCORE_CHUNK), h_inl2(SCORE_CHUNK), h_cinl(SCORE_CHUNK);
    std::vector<double> h_err(SCORE_CHUNK), h_T(SCORE_CHUNK * 16);
    int64_t est_k = max_iteration;
    double best_fit = 0.0, best_rmse = 0.0;
    int rc = IBL_OK;
    int64_t cur_block = std::min(first_block, MAX_BLOCK);
    for (int64_t h0 = h_begin; h0 < max_iteration && h0 < est_k;) {
        int64_t H = std::min(cur_block, std::min(max_iteration, est_k) - h0);
        if ((rc = ibl_ransac_check(src, tgt, corres, n_corres, samples, seed, h0, H, edge_thr, max_d, use_checkers, pass,
                                   nullptr, stream)) < 0) break;
        cub::CountingInputIterator<int64_t> cit(h0);
        IBL_CUDA(cub::DeviceSelect::Flagged(tmp, tb, cit, pass, ids, dcount, (int)H, st));
        int npass = 0;
        IBL_CUDA(cudaMemcpyAsync(&npass, dcount, sizeof(int), cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));
        bool stop = false;
        for (int64_t c0 = 0; c0 < npass && !stop; c0 += SCORE_CHUNK) {
            int64_t nc = std::min<int64_t>(SCORE_CHUNK, npass - c0);
            // stage 1: inlier counts of the whole chunk (cheap: first candidate within max_d ends a point's scan)
            if ((rc = ibl_ransac_count(tgt_grid, src, n_src, tgt, corres, n_corres, samples, seed, ids + c0, nc, max_d, inl, Ts,
                                       stream)) < 0) { stop = true; break; }
            IBL_CUDA(cudaMemcpyAsync(h_ids.data(), ids + c0, sizeof(int64_t) * nc, cudaMemcpyDeviceToHost, st));
            IBL_CUDA(cudaMemcpyAsync(h_inl.data(), inl, sizeof(int32_t) * nc, cudaMemcpyDeviceToHost, st));
            IBL_CUDA(ibl_sync(st));
            // A hypothesis can only replace the best if its count reaches the largest count seen before it (the ordered
            // rule is "fitness greater, or equal with a smaller rmse"; after any hypothesis the best count is the
            // running maximum).  Only those need the exact err2 / correspondence-inlier pass.
            int64_t nsel = 0;
            {
                int64_t run_max = h_result->best_hyp >= 0 ? h_result->n_inliers : 0;
                for (int64_t k = 0; k < nc; ++k) {
                    h_sel[k] = -1;
                    if (h_ids[k] >= est_k) break;
                    if (h_inl[k] > 0 && h_inl[k] >= run_max) { h_sel[k] = nsel; h_sel_ids[nsel++] = h_ids[k]; run_max = h_inl[k]; }
                }
            }
            if (nsel > 0) {                                // stage 2: exact score of the few contenders
                IBL_CUDA(cudaMemcpyAsync(sel_ids, h_sel_ids.data(), sizeof(int64_t) * nsel, cudaMemcpyHostToDevice, st));
                if ((rc = ibl_ransac_score(tgt_grid, src, n_src, tgt, corres, n_corres, samples, seed, sel_ids, nsel, max_d, inl2,
                                           err, cinl, Ts, stream)) < 0) { stop = true; break; }
                IBL_CUDA(cudaMemcpyAsync(h_inl2.data(), inl2, sizeof(int32_t) * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(cudaMemcpyAsync(h_cinl.data(), cinl, sizeof(int32_t) * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(cudaMemcpyAsync(h_err.data(), err, sizeof(double) * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(cudaMemcpyAsync(h_T.data(), Ts, sizeof(double) * 16 * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(ibl_sync(st));
            }
            for (int64_t k = 0; k < nc; ++k) {
                if (h_ids[k] >= est_k) { stop = true; break; }
                h_result->validated++;
                const int64_t j = h_sel[k];
                if (j < 0) continue;                       // cannot beat the best: fewer inliers
                if (h_inl2[j] != h_inl[k]) { ibl_set_error("ransac: count pass and exact pass disagree"); rc = IBL_E_CUDA; stop = true; break; }
                double fit = (double)h_inl2[j] / (double)n_src;
                double rmse = h_inl2[j] > 0 ? sqrt(h_err[j] / (double)h_inl2[j]) : 0.0;
                if (fit > best_fit || (fit == best_fit && rmse < best_rmse)) {
                    best_fit = fit; best_rmse = rmse;
                    h_result->fitness = fit; h_result->inlier_rmse = rmse;
                    h_result->n_inliers = h_inl2[j]; h_result->err2 = h_err[j];
                    h_result->best_hyp = h_ids[k];
                    memcpy(h_result->T, &h_T[16 * j], sizeof(double) * 16);
                    if (early_exit) {
                        double rho = (double)h_cinl[j] / (double)n_corres;
                        double den = rho < 1.0 ? log(1.0 - rho * rho * rho) : -INFINITY;
                        double kk = den == 0.0 ? INFINITY : log(1.0 - confidence) / den;
                        if (kk < (double)est_k) est_k = (int64_t)ceil(kk);
                    }
                }
            }
        }
        if (rc < 0) break;
        h0 += H;
        cur_block = std::min(cur_block * 2, MAX_BLOCK);
    }
    h_result->drawn = std::max<int64_t>(std::min(max_iteration, std::max<int64_t>(est_k, 0)) - h_begin, 0);
    ibl_free(tmp, st); ibl_free(pass, st); ibl_free(ids, st); ibl_free(dcount, st);
    ibl_free(inl, st); ibl_free(inl2, st); ibl_free(sel_ids, st); ibl_free(cinl, st); ibl_free(err, st); ibl_free(Ts, st);
    return rc;
}

// ================================================================================================
// Batched RANSAC: ALL registration problems of a step advance together (object_memory.py:1020-1108 runs <= 6 of these
// per frame, one after the other; a step of B frames holds <= 6 B independent problems).  Every round handles one block
// of hypotheses per active problem with ONE launch per stage for the whole batch and TWO host synchronisations (the
// inlier counts of the validated hypotheses; the exact scores of the contenders) -- instead of three per 512 validated
// hypotheses per problem.  The per-problem arithmetic, the hypothesis stream and the ordered best-update / adaptive-stop
// rule are those of ibl_ransac_run_range, so every result field is identical to running the problems one by one.
struct RbProb {
    GridView g;
    const double *src, *tgt;
    const int32_t *corres;
    int64_t n_src, n_corres;
    uint64_t seed;
    int64_t h0;         // first hypothesis of this round's block
    int32_t H;          // hypotheses in this round's block (0 = not active this round)
    int32_t ntiles;
    int32_t thr;        // inlier count of the problem's best hypothesis so far (0 = none yet): the pruning threshold
    int32_t nchunks, chunk_step;   // count pass: chunks of PRUNE_CHUNK source points, visited in steps of chunk_step
};
struct RbEntry { int32_t p, k; };   // validated hypothesis k (block-relative) of problem p

// grid (ceil(max H / 128), problems).  Validated hypotheses of ALL problems are appended to one list (order arbitrary, the
// host sorts): one atomicAdd per warp.
__global__ void __launch_bounds__(128) ransac_check_batch_kernel(const RbProb *__restrict__ probs, double edge_thr,
                                                                 double dist_thr, int use_checkers,
                                                                 RbEntry *__restrict__ entries, int *__restrict__ n_entries) {
    const RbProb *P = probs + blockIdx.y;
    const int H = P->H;
    if ((int64_t)blockIdx.x * 128 >= H) return;
    const int k = blockIdx.x * 128 + threadIdx.x;
    bool ok = false;
    if (k < H) {
        double T[16];
        ok = ransac_check_one(P->src, P->tgt, P->corres, P->n_corres, nullptr, P->seed, P->h0 + k, edge_thr, dist_thr,
                              use_checkers, T);
    }
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == 0) base = atomicAdd(n_entries, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (ok) { RbEntry e; e.p = (int)blockIdx.y; e.k = k; entries[base + __popc(m & ((1u << lane) - 1u))] = e; }
}

// The check in two kernels (checkers on).  The fused kernel above needs 140 registers for the inlined Jacobi-Kabsch, so only
// ~12 % of the warp slots are occupied while 99 % of the hypotheses die in the EdgeLength test, which needs six gathered
// points and a few square roots (ncu: latency-bound, L1 39 %, issue 23 %).  Kernel A does sample + gather + EdgeLength at
// full occupancy and appends the survivors to a list; kernel B solves the Kabsch and applies the Distance test for those
// only, one survivor per thread with every lane of a warp busy.  Same arithmetic per hypothesis, same set of validated
// hypotheses (the list order is arbitrary in both versions; the host sorts).
__global__ void __launch_bounds__(256) ransac_edge_batch_kernel(const RbProb *__restrict__ probs, double edge_thr,
                                                                RbEntry *__restrict__ edge_entries, int *__restrict__ n_edge) {
    const RbProb *P = probs + blockIdx.y;
    const int H = P->H;
    if ((int64_t)blockIdx.x * 256 >= H) return;
    const int k = blockIdx.x * 256 + threadIdx.x;
    bool ok = false;
    if (k < H) {
        double s[9], t[9];
        gather_sample(P->src, P->tgt, P->corres, P->n_corres, nullptr, P->seed, P->h0 + k, s, t);
        ok = true;
        for (int i = 0; i < 3; ++i)
            for (int j = i + 1; j < 3; ++j) {
                double a[3] = {s[3 * i] - s[3 * j], s[3 * i + 1] - s[3 * j + 1], s[3 * i + 2] - s[3 * j + 2]};
                double b[3] = {t[3 * i] - t[3 * j], t[3 * i + 1] - t[3 * j + 1], t[3 * i + 2] - t[3 * j + 2]};
                double ds = sqrt(ibl_dot(a, a)), dt = sqrt(ibl_dot(b, b));
                if (ds < dt * edge_thr || dt < ds * edge_thr) ok = false;
            }
    }
    const unsigned m = __ballot_sync(0xffffffffu, ok);
    if (m == 0) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == 0) base = atomicAdd(n_edge, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (ok) { RbEntry e; e.p = (int)blockIdx.y; e.k = k; edge_entries[base + __popc(m & ((1u << lane) - 1u))] = e; }
}
__global__ void __launch_bounds__(128) ransac_kabsch_batch_kernel(const RbProb *__restrict__ probs, double dist_thr,
                                                                  const RbEntry *__restrict__ edge_entries,
                                                                  const int *__restrict__ n_edge, RbEntry *__restrict__ entries,
                                                                  int *__restrict__ n_entries) {
    const int n = *n_edge;
    for (int base_i = blockIdx.x * 128; base_i < n; base_i += gridDim.x * 128) {      // uniform per warp
        const int i = base_i + threadIdx.x;
        bool ok = false;
        RbEntry ent = {0, 0};
        if (i < n) {
            ent = edge_entries[i];
            const RbProb *P = probs + ent.p;
            double s[9], t[9], T[16];
            gather_sample(P->src, P->tgt, P->corres, P->n_corres, nullptr, P->seed, P->h0 + ent.k, s, t);
            ibl_kabsch3(s, t, T);
            ok = true;
            for (int j = 0; j < 3; ++j) {
                double qx, qy, qz;
                ibl_xform(T, s[3 * j], s[3 * j + 1], s[3 * j + 2], qx, qy, qz);
                double d[3] = {t[3 * j] - qx, t[3 * j + 1] - qy, t[3 * j + 2] - qz};
                if (sqrt(ibl_dot(d, d)) > dist_thr) ok = false;
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (m == 0) continue;
        const int lane = threadIdx.x & 31;
        int base = 0;
        if (lane == 0) base = atomicAdd(n_entries, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (ok) entries[base + __popc(m & ((1u << lane) - 1u))] = ent;
    }
}

// Inlier count of the validated hypotheses, with an exact early exit.  Only a hypothesis whose count reaches the running
// maximum can change the RANSAC state, and the running maximum never falls below the count `thr` of the
// best hypothesis of the earlier rounds.  A CTA takes one hypothesis at a time (persistent grid over the entry list, whose
// length lives on the device: no host round trip between the check and the count), computes its transform, walks the
// source in chunks and stops as soon as
//     inliers so far + points not yet looked at  <  thr,
// i.e. when even a perfect remainder could not reach thr; such a hypothesis is reported as -1 ("below thr") and is never a
// contender.  Hypotheses that finish report their exact count.  With a best fitness of 0.9 a hypothesis that fits 10 % of
// the cloud is dropped after ~11 % of the points.  Chunks are visited with a stride coprime to their number, so the first
// chunks sample the whole cloud (the source is a concatenation of objects) rather than its first object.
#define PRUNE_CHUNK 1024
// FIELD: every problem's grid carries an NN candidate field that covers max_d (the product path).  The ring-search
// fallback then is not even compiled in, which halves the registers (128 -> 6x) and doubles the resident warps of a kernel
// whose time is dependent-load latency.
template <bool FIELD, int MINB>
__global__ void __launch_bounds__(SCORE_THREADS, MINB) ransac_count_batch_kernel(const RbProb *__restrict__ probs,
                                                                           const RbEntry *__restrict__ entries,
                                                                           const int *__restrict__ n_entries, double max_d,
                                                                           double max_d2, int32_t *__restrict__ inliers,
                                                                           int *__restrict__ next_entry) {
    __shared__ GridView g;
    __shared__ double T[16];
    __shared__ int sc[2][SCORE_THREADS / 32];
    __shared__ int s_e;
    const int n = *n_entries;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int cur_p = -1;
    // Entries are handed out through a device counter, not round-robin: a hypothesis costs anything between one chunk (pruned
    // at once) and the whole source, and with a static assignment the launch lasted as long as its unluckiest CTA.
#pragma unroll 1
    for (;;) {
        __syncthreads();                                   // the previous hypothesis' readers of g / T / sc / s_e are done
        if (threadIdx.x == 0) s_e = atomicAdd(next_entry, 1);
        __syncthreads();
        const int e = s_e;
        if (e >= n) break;
        const RbEntry ent = entries[e];
        const RbProb *P = probs + ent.p;
        if (ent.p != cur_p) {
            for (int k = threadIdx.x; k < (int)(sizeof(GridView) / 4); k += SCORE_THREADS)
                reinterpret_cast<int *>(&g)[k] = reinterpret_cast<const int *>(&P->g)[k];
            cur_p = ent.p;
        }
        if (threadIdx.x == 0) {
            double s[9], t[9], Tl[16];
            gather_sample(P->src, P->tgt, P->corres, P->n_corres, nullptr, P->seed, P->h0 + ent.k, s, t);
            ibl_kabsch3(s, t, Tl);
#pragma unroll
            for (int k = 0; k < 16; ++k) T[k] = Tl[k];
        }
        __syncthreads();
        const double *__restrict__ src = P->src;
        const int64_t n_src = P->n_src;
        const int thr = P->thr, nchunks = P->nchunks, step = P->chunk_step;
        const bool field = FIELD || (g.has_field && max_d <= g.f.max_d);
        const bool sure_ok = field && max_d == g.f.max_d;
        int total = 0;
        int64_t seen = 0;
        int chunk = 0;
        bool pruned = false;
#pragma unroll 1
        for (int it = 0; it < nchunks; ++it) {
            const int64_t base = (int64_t)chunk * PRUNE_CHUNK;
            int cnt = 0;
            if (FIELD) {
                // The four points a thread owns in this chunk go through the field lookup TOGETHER, level by level (source
                // point -> block pointer -> voxel entry -> first candidate -> its coordinates): the kernel's time is the
                // latency of that dependent chain (ncu: IPC 0.4, 11 warps waiting on a load per issue), and walking the
                // four chains side by side keeps four loads in flight per thread instead of one.  Same tests in the same
                // order per point as ibl_any_within_field, so every count is unchanged.
                constexpr int R = PRUNE_CHUNK / SCORE_THREADS;
                const NNFieldView &f = g.f;
                double qx[R], qy[R], qz[R];
                long long bidx[R];
                int vox[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int64_t i = base + r * SCORE_THREADS + threadIdx.x;
                    bidx[r] = -1; vox[r] = 0;
                    qx[r] = qy[r] = qz[r] = 0.0;
                    if (i < n_src) {
                        ibl_xform(T, src[3 * i], src[3 * i + 1], src[3 * i + 2], qx[r], qy[r], qz[r]);
                        const double fx = floor((qx[r] - f.ox) * f.inv_c), fy = floor((qy[r] - f.oy) * f.inv_c),
                                     fz = floor((qz[r] - f.oz) * f.inv_c);
                        if (fx >= 0.0 && fy >= 0.0 && fz >= 0.0 && fx < (double)f.nvx && fy < (double)f.nvy && fz < (double)f.nvz) {
                            const int ix = (int)fx, iy = (int)fy, iz = (int)fz;
                            bidx[r] = ((long long)(iz >> 3) * f.nby + (iy >> 3)) * f.nbx + (ix >> 3);
                            vox[r] = ((iz & 7) << 6) | ((iy & 7) << 3) | (ix & 7);
                        }
                    }
                }
                int blk[R];
#pragma unroll
                for (int r = 0; r < R; ++r) blk[r] = bidx[r] >= 0 ? f.block_ptr[bidx[r]] : -1;
                uint2 ent4[R];
#pragma unroll
                for (int r = 0; r < R; ++r) ent4[r] = blk[r] >= 0 ? f.entries[(size_t)blk[r] * 512 + vox[r]] : make_uint2(0u, 0u);
                int c0[R];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const bool sure = (ent4[r].y & 0x80000000u) && sure_ok;
                    const unsigned nc = ent4[r].y & 0x7fffffffu;
                    c0[r] = (!sure && nc > 0) ? f.cand[ent4[r].x] : -1;
                    if (sure) { ++cnt; ent4[r].y = 0u; }                  // decided: no candidates left to look at
                }
                double4 t0[R];
#pragma unroll
                for (int r = 0; r < R; ++r) t0[r] = c0[r] >= 0 ? g.sorted[c0[r]] : make_double4(0.0, 0.0, 0.0, 0.0);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (c0[r] < 0) continue;
                    bool hit = ibl_d2(t0[r].x, t0[r].y, t0[r].z, qx[r], qy[r], qz[r]) < max_d2;
                    const unsigned nc = ent4[r].y & 0x7fffffffu;
                    const int *c = f.cand + ent4[r].x;
                    for (unsigned k = 1; k < nc && !hit; ++k) {
                        const double4 t = g.sorted[c[k]];
                        hit = ibl_d2(t.x, t.y, t.z, qx[r], qy[r], qz[r]) < max_d2;
                    }
                    cnt += hit ? 1 : 0;
                }
            } else {
#pragma unroll 1
                for (int r = 0; r < PRUNE_CHUNK / SCORE_THREADS; ++r) {
                    int64_t i = base + r * SCORE_THREADS + threadIdx.x;
                    if (i < n_src) {
                        double x, y, z, bd;
                        ibl_xform(T, src[3 * i], src[3 * i + 1], src[3 * i + 2], x, y, z);
                        if (field) cnt += ibl_any_within_field(g, x, y, z, max_d2, sure_ok) ? 1 : 0;
                        else cnt += ibl_nn1_thread(g, x, y, z, max_d, max_d2, bd) >= 0 ? 1 : 0;
                    }
                }
            }
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (lane == 0) sc[it & 1][w] = cnt;
            __syncthreads();                               // one barrier per chunk: the two buffers alternate
#pragma unroll
            for (int k = 0; k < SCORE_THREADS / 32; ++k) total += sc[it & 1][k];
            seen += min((int64_t)PRUNE_CHUNK, n_src - base);
            if ((int64_t)total + (n_src - seen) < (int64_t)thr) { pruned = true; break; }   // uniform over the CTA
            chunk += step;
            if (chunk >= nchunks) chunk -= nchunks;
        }
        if (threadIdx.x == 0) inliers[e] = pruned ? -1 : total;
    }
}

// Tried and rejected (round 2): a WARP per hypothesis (no block barrier, no idle threads during the Kabsch solve).  A round
// holds only a few thousand validated hypotheses, whose cost ranges from one chunk (pruned) to the whole source: with 3552
// resident warps each warp gets about one hypothesis and the launch lasts as long as its single longest one -- 31 ms per
// bench step against 17 ms for the CTA-per-hypothesis kernel above, which finishes a long hypothesis eight times sooner.
// which contender does work item `item` belong to: prefix[c] <= item < prefix[c+1]
__device__ __forceinline__ int rb_find(const int64_t *__restrict__ prefix, int n, int64_t item) {
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (prefix[mid] <= item) lo = mid; else hi = mid;
    }
    return lo;
}

// transforms of the contenders (thread per contender)
__global__ void __launch_bounds__(64) hyp_transform_sel_kernel(const RbProb *__restrict__ probs, const RbEntry *__restrict__ sel,
                                                               int nsel, double *__restrict__ T_out) {
    const int c = blockIdx.x * 64 + threadIdx.x;
    if (c >= nsel) return;
    const RbProb *P = probs + sel[c].p;
    double s[9], t[9], T[16];
    gather_sample(P->src, P->tgt, P->corres, P->n_corres, nullptr, P->seed, P->h0 + sel[c].k, s, t);
    ibl_kabsch3(s, t, T);
#pragma unroll
    for (int e = 0; e < 16; ++e) T_out[(size_t)c * 16 + e] = T[e];
}

// exact score of the contenders: one CTA = (contender, source tile); tile partials go to part[item] in
// ransac_score_kernel's reduction order inside the CTA.
__global__ void __launch_bounds__(SCORE_THREADS) ransac_score_batch_kernel(const RbProb *__restrict__ probs,
                                                                           const int64_t *__restrict__ item_prefix, int nsel,
                                                                           const RbEntry *__restrict__ sel,
                                                                           const double *__restrict__ T_sel, double max_d,
                                                                           double max_d2, int *__restrict__ part_cnt,
                                                                           double *__restrict__ part_err) {
    __shared__ GridView g;
    __shared__ double T[16];
    __shared__ double se[SCORE_THREADS / 32];
    __shared__ int sc[SCORE_THREADS / 32];
    __shared__ int s_c;
    if (threadIdx.x == 0) s_c = rb_find(item_prefix, nsel, (int64_t)blockIdx.x);
    __syncthreads();
    const int c = s_c;
    const int64_t tile = (int64_t)blockIdx.x - item_prefix[c];
    const RbProb *P = probs + sel[c].p;
    for (int k = threadIdx.x; k < (int)(sizeof(GridView) / 4); k += SCORE_THREADS)
        reinterpret_cast<int *>(&g)[k] = reinterpret_cast<const int *>(&P->g)[k];
    if (threadIdx.x < 16) T[threadIdx.x] = T_sel[(size_t)c * 16 + threadIdx.x];
    __syncthreads();
    const double *__restrict__ src = P->src;
    const int64_t n_src = P->n_src;
    int cnt = 0;
    double err = 0.0;
    const int64_t base = tile * SCORE_TILE;
#pragma unroll 1
    for (int r = 0; r < SCORE_TILE / SCORE_THREADS; ++r) {
        int64_t i = base + r * SCORE_THREADS + threadIdx.x;
        if (i < n_src) {
            double x, y, z, bd;
            ibl_xform(T, src[3 * i], src[3 * i + 1], src[3 * i + 2], x, y, z);
            int j = ibl_nn1_thread(g, x, y, z, max_d, max_d2, bd);
            if (j >= 0) { ++cnt; err = __dadd_rn(err, bd); }
        }
    }
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) { err += __shfl_down_sync(0xffffffffu, err, o); cnt += __shfl_down_sync(0xffffffffu, cnt, o); }
    if (lane == 0) { se[w] = err; sc[w] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; int tc = 0;
        for (int k = 0; k < SCORE_THREADS / 32; ++k) { te += se[k]; tc += sc[k]; }
        part_cnt[blockIdx.x] = tc;
        part_err[blockIdx.x] = te;
    }
}

// one CTA per contender: tile partials in ransac_finalize_kernel's order + the correspondence inlier count
__global__ void __launch_bounds__(FIN_THREADS) ransac_finalize_batch_kernel(const RbProb *__restrict__ probs,
                                                                            const int64_t *__restrict__ item_prefix,
                                                                            const RbEntry *__restrict__ sel,
                                                                            const int *__restrict__ part_cnt,
                                                                            const double *__restrict__ part_err,
                                                                            const double *__restrict__ T_sel, double max_d,
                                                                            int32_t *__restrict__ inliers, double *__restrict__ err2,
                                                                            int32_t *__restrict__ corres_inliers) {
    __shared__ double se[FIN_THREADS / 32];
    __shared__ int sc[FIN_THREADS / 32], si[FIN_THREADS / 32];
    __shared__ double T[16];
    const int c = blockIdx.x;
    const RbProb *P = probs + sel[c].p;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x < 16) T[threadIdx.x] = T_sel[(size_t)c * 16 + threadIdx.x];
    __syncthreads();
    const int64_t p0 = item_prefix[c];
    const int ntiles = (int)(item_prefix[c + 1] - p0);
    double e = 0; int cn = 0;
    for (int k = threadIdx.x; k < ntiles; k += FIN_THREADS) { e += part_err[p0 + k]; cn += part_cnt[p0 + k]; }
    int ci = 0;
    {
        const double *__restrict__ src = P->src, *__restrict__ tgt = P->tgt;
        const int32_t *__restrict__ corres = P->corres;
        const double md2 = max_d * max_d;
        for (int64_t k = threadIdx.x; k < P->n_corres; k += FIN_THREADS) {
            size_t cs = (size_t)corres[2 * k], ct = (size_t)corres[2 * k + 1];
            double x, y, z;
            ibl_xform(T, src[3 * cs], src[3 * cs + 1], src[3 * cs + 2], x, y, z);
            double d[3] = {x - tgt[3 * ct], y - tgt[3 * ct + 1], z - tgt[3 * ct + 2]};
            ci += ibl_dot(d, d) < md2;
        }
    }
    for (int o = 16; o; o >>= 1) {
        e += __shfl_down_sync(0xffffffffu, e, o);
        cn += __shfl_down_sync(0xffffffffu, cn, o);
        ci += __shfl_down_sync(0xffffffffu, ci, o);
    }
    if (lane == 0) { se[w] = e; sc[w] = cn; si[w] = ci; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; int tc = 0, ti = 0;
        for (int k = 0; k < FIN_THREADS / 32; ++k) { te += se[k]; tc += sc[k]; ti += si[k]; }
        inliers[c] = tc; err2[c] = te; corres_inliers[c] = ti;
    }
}

// 1 (default): the count pass stops a hypothesis as soon as it cannot reach the best count so far (exact, see
// ransac_count_batch_kernel); 0: every validated hypothesis is counted over the whole source (measurement switch).
static int g_ransac_prune = 1;
// resident CTAs per SM the field-specialised count kernel is compiled for: 3 (80 registers, a few spills; measured 169 ms per
// bench pass) or 2 (114 registers, no spills; 231 ms): the extra warps matter more than the spills.
// IBL_RANSAC_COUNT_MINB overrides (measurement switch).
// the check as two kernels (1, default) or fused (0); IBL_RANSAC_CHECK_SPLIT overrides (measurement switch)
static int rb_check_split() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("IBL_RANSAC_CHECK_SPLIT"); v = (e && atoi(e) == 0) ? 0 : 1; }
    return v;
}
static int rb_count_minb() {
    static int v = 0;
    if (v == 0) { const char *e = getenv("IBL_RANSAC_COUNT_MINB"); v = (e && atoi(e) == 2) ? 2 : 3; }
    return v;
}
IBL_API int ibl_ransac_set_prune(int on) { g_ransac_prune = on ? 1 : 0; return IBL_OK; }

#define RB_MAX_BLOCK (1 << 20)     // hypotheses per problem per round, at most
#define RB_GROWTH 4                // block growth per round: 64 k, 256 k, 1 M, 1 M, ... (few rounds = few host round trips)
#define RB_OPT_ENTRIES (1 << 16)   // validated entries read back together with their number (more need a second copy)

IBL_API int ibl_ransac_run_batch(const ibl_reg_problem_t *problems, int32_t n_problems, int64_t max_iteration,
                                 double confidence, double max_d, double edge_thr, int use_checkers, int early_exit,
                                 int64_t first_block, ibl_reg_result_t *h_results, void *stream) {
    IBL_ARG(n_problems >= 0 && (n_problems == 0 || (problems && h_results)) && max_d > 0 && n_problems <= 65535);
    if (n_problems == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    const int P = n_problems;
    if (first_block <= 0) first_block = 1 << 16;
    struct Host { int64_t est_k, h0, cur_block; double best_fit, best_rmse; bool active; };
    std::vector<Host> hs(P);
    std::vector<RbProb> hp(P);
    for (int p = 0; p < P; ++p) {
        const ibl_reg_problem_t &q = problems[p];
        ibl_reg_result_t &r = h_results[p];
        memset(&r, 0, sizeof(r));
        for (int k = 0; k < 16; ++k) r.T[k] = (k % 5 == 0) ? 1.0 : 0.0;
        r.best_hyp = -1;
        r.n_corres = q.n_corres;
        hs[p] = Host{max_iteration, 0, std::min<int64_t>(first_block, RB_MAX_BLOCK), 0.0, 0.0, true};
        if (q.n_corres < 3 || max_iteration <= 0 || q.n_src <= 0 || !q.grid_ransac) { r.status = 1; hs[p].active = false; }
        IBL_ARG(!hs[p].active || (q.src && q.tgt && q.corres && q.n_corres < (1ll << 31)));
        RbProb &d = hp[p];
        memset(&d, 0, sizeof(d));
        if (hs[p].active) d.g = ibl_view(q.grid_ransac);
        d.src = q.src; d.tgt = q.tgt; d.corres = q.corres; d.n_src = q.n_src; d.n_corres = q.n_corres; d.seed = q.seed;
        d.ntiles = (int)std::max<int64_t>(1, (q.n_src + SCORE_TILE - 1) / SCORE_TILE);
        d.nchunks = (int)std::max<int64_t>(1, (q.n_src + PRUNE_CHUNK - 1) / PRUNE_CHUNK);
        // a stride near the golden section of the chunk count, made coprime to it: consecutive visits are far apart
        int step = std::max(1, (int)(0.6180339887 * d.nchunks));
        auto gcd = [](int a, int b) { while (b) { int t = a % b; a = b; b = t; } return a; };
        while (gcd(step, d.nchunks) != 1) ++step;
        d.chunk_step = d.nchunks == 1 ? 0 : step % d.nchunks;
    }
    bool all_field = true;
    for (int p = 0; p < P; ++p)
        if (hs[p].active && !(problems[p].grid_ransac->has_field && max_d <= problems[p].grid_ransac->field.max_d)) all_field = false;
    RbProb *d_probs = nullptr; int *d_n = nullptr, *d_ne = nullptr, *d_next = nullptr;
    IBL_CUDA(ibl_alloc(&d_probs, (size_t)P, st));
    IBL_CUDA(ibl_alloc(&d_n, 1, st));
    IBL_CUDA(ibl_alloc(&d_ne, 1, st));
    IBL_CUDA(ibl_alloc(&d_next, 1, st));
    std::vector<RbEntry> h_ent, sel;
    std::vector<int32_t> h_inl, h_inl2, h_cinl;
    std::vector<int64_t> sel_prefix;
    std::vector<double> h_err, h_T;
    struct Cand { int32_t k; int32_t inl; int sel; };
    std::vector<std::vector<Cand>> cands(P);
    int rc = IBL_OK;
    for (;;) {
        // ---- this round's block per problem
        int64_t cap = 0; int maxH = 0; bool any = false;
        for (int p = 0; p < P; ++p) {
            RbProb &d = hp[p];
            d.H = 0;
            Host &h = hs[p];
            if (h.active && !(h.h0 < max_iteration && h.h0 < h.est_k)) h.active = false;
            if (!h.active) continue;
            int64_t H = std::min(h.cur_block, std::min(max_iteration, h.est_k) - h.h0);
            d.H = (int32_t)H; d.h0 = h.h0;
            d.thr = (g_ransac_prune && h_results[p].best_hyp >= 0) ? (int32_t)h_results[p].n_inliers : 0;
            cap += H; maxH = std::max(maxH, (int)H); any = true;
        }
        if (!any) break;
        if (cap >= (1ll << 31)) { ibl_set_error("ransac batch: %lld hypotheses in one round", (long long)cap); rc = IBL_E_ARG; break; }
        RbEntry *d_ent = nullptr; int32_t *d_inl = nullptr;
        IBL_CUDA(ibl_alloc(&d_ent, (size_t)cap, st));       // every hypothesis of the round may validate (checkers off)
        IBL_CUDA(ibl_alloc(&d_inl, (size_t)cap, st));
        IBL_CUDA(cudaMemcpyAsync(d_probs, hp.data(), sizeof(RbProb) * P, cudaMemcpyHostToDevice, st));
        IBL_CUDA(cudaMemsetAsync(d_n, 0, sizeof(int), st));
        RbEntry *d_edge = nullptr;
        {
            IBL_PROF(IBL_PROF_RANSAC_CHECK, st, 157.0 * (double)cap);
            if (use_checkers && rb_check_split()) {
                IBL_CUDA(ibl_alloc(&d_edge, (size_t)cap, st));
                IBL_CUDA(cudaMemsetAsync(d_ne, 0, sizeof(int), st));
                ransac_edge_batch_kernel<<<dim3((unsigned)((maxH + 255) / 256), (unsigned)P), 256, 0, st>>>(d_probs, edge_thr, d_edge, d_ne);
                IBL_LAUNCH_CHECK();
                ransac_kabsch_batch_kernel<<<IBL_SMS * 4, 128, 0, st>>>(d_probs, max_d, d_edge, d_ne, d_ent, d_n);
                IBL_LAUNCH_CHECK();
            } else {
                ransac_check_batch_kernel<<<dim3((unsigned)((maxH + 127) / 128), (unsigned)P), 128, 0, st>>>(d_probs, edge_thr, max_d,
                                                                                                          use_checkers, d_ent, d_n);
                IBL_LAUNCH_CHECK();
            }
        }
        {
            // work figure: the un-pruned pass (every validated hypothesis over its whole source); the number of validated
            // hypotheses is only known on the device here, so it is filled in after the read-back
            IBL_PROF(IBL_PROF_RANSAC_SCORE, st, 0.0);
            IBL_CUDA(cudaMemsetAsync(d_next, 0, sizeof(int), st));
            if (all_field)
                if (rb_count_minb() == 2)
                    ransac_count_batch_kernel<true, 2><<<IBL_SMS * 4, SCORE_THREADS, 0, st>>>(d_probs, d_ent, d_n, max_d, max_d * max_d, d_inl, d_next);
                else
                    ransac_count_batch_kernel<true, 3><<<IBL_SMS * 6, SCORE_THREADS, 0, st>>>(d_probs, d_ent, d_n, max_d, max_d * max_d, d_inl, d_next);
            else
                ransac_count_batch_kernel<false, 1><<<IBL_SMS * 4, SCORE_THREADS, 0, st>>>(d_probs, d_ent, d_n, max_d, max_d * max_d, d_inl, d_next);
            IBL_LAUNCH_CHECK();
        }
        int nval = 0;
        const int64_t opt = std::min<int64_t>(cap, RB_OPT_ENTRIES);
        h_ent.resize((size_t)std::max<int64_t>(opt, 1)); h_inl.resize((size_t)std::max<int64_t>(opt, 1));
        IBL_CUDA(cudaMemcpyAsync(&nval, d_n, sizeof(int), cudaMemcpyDeviceToHost, st));
        IBL_CUDA(cudaMemcpyAsync(h_ent.data(), d_ent, sizeof(RbEntry) * opt, cudaMemcpyDeviceToHost, st));
        IBL_CUDA(cudaMemcpyAsync(h_inl.data(), d_inl, sizeof(int32_t) * opt, cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));                                                              // sync 1: counts
        if (nval > opt) {
            h_ent.resize((size_t)nval); h_inl.resize((size_t)nval);
            IBL_CUDA(cudaMemcpyAsync(h_ent.data() + opt, d_ent + opt, sizeof(RbEntry) * (nval - opt), cudaMemcpyDeviceToHost, st));
            IBL_CUDA(cudaMemcpyAsync(h_inl.data() + opt, d_inl + opt, sizeof(int32_t) * (nval - opt), cudaMemcpyDeviceToHost, st));
            IBL_CUDA(ibl_sync(st));
        }
        ibl_prof_add_work(IBL_PROF_RANSAC_SCORE, [&] { double w = 0; for (int e = 0; e < nval; ++e) w += 24.0 * hp[h_ent[e].p].n_src + 128.0; return w; }());
        if (nval > 0) {
            // ---- contenders: in id order, a hypothesis whose count reaches the running maximum before it
            for (int e = 0; e < nval; ++e) cands[h_ent[e].p].push_back(Cand{h_ent[e].k, h_inl[e], -1});
            sel.clear(); sel_prefix.assign(1, 0);
            for (int p = 0; p < P; ++p) {
                std::vector<Cand> &cv = cands[p];
                if (cv.empty()) continue;
                std::sort(cv.begin(), cv.end(), [](const Cand &a, const Cand &b) { return a.k < b.k; });
                int64_t run_max = h_results[p].best_hyp >= 0 ? h_results[p].n_inliers : 0;
                for (Cand &c : cv) {
                    if (hp[p].h0 + c.k >= hs[p].est_k) break;
                    if (c.inl > 0 && c.inl >= run_max) {
                        c.sel = (int)sel.size();
                        sel.push_back(RbEntry{p, c.k});
                        sel_prefix.push_back(sel_prefix.back() + hp[p].ntiles);
                        run_max = c.inl;
                    }
                }
            }
            const int nsel = (int)sel.size();
            if (nsel > 0) {
                RbEntry *d_sel = nullptr; int32_t *d_i2 = nullptr, *d_ci = nullptr; int64_t *d_spre = nullptr;
                int *d_pc = nullptr; double *d_pe = nullptr, *d_err = nullptr, *d_T = nullptr;
                const int64_t sitems = sel_prefix.back();
                IBL_CUDA(ibl_alloc(&d_sel, (size_t)nsel, st));
                IBL_CUDA(ibl_alloc(&d_spre, (size_t)nsel + 1, st));
                IBL_CUDA(ibl_alloc(&d_pc, (size_t)sitems, st)); IBL_CUDA(ibl_alloc(&d_pe, (size_t)sitems, st));
                IBL_CUDA(ibl_alloc(&d_i2, (size_t)nsel, st)); IBL_CUDA(ibl_alloc(&d_ci, (size_t)nsel, st));
                IBL_CUDA(ibl_alloc(&d_err, (size_t)nsel, st)); IBL_CUDA(ibl_alloc(&d_T, (size_t)nsel * 16, st));
                IBL_CUDA(cudaMemcpyAsync(d_sel, sel.data(), sizeof(RbEntry) * nsel, cudaMemcpyHostToDevice, st));
                IBL_CUDA(cudaMemcpyAsync(d_spre, sel_prefix.data(), sizeof(int64_t) * (nsel + 1), cudaMemcpyHostToDevice, st));
                hyp_transform_sel_kernel<<<(unsigned)((nsel + 63) / 64), 64, 0, st>>>(d_probs, d_sel, nsel, d_T);
                IBL_LAUNCH_CHECK();
                {
                    double work = 0;
                    for (int c = 0; c < nsel; ++c) work += 24.0 * hp[sel[c].p].n_src + 128.0;
                    IBL_PROF(IBL_PROF_RANSAC_SCORE, st, work);
                    ransac_score_batch_kernel<<<(unsigned)sitems, SCORE_THREADS, 0, st>>>(d_probs, d_spre, nsel, d_sel, d_T, max_d,
                                                                                        max_d * max_d, d_pc, d_pe);
                    IBL_LAUNCH_CHECK();
                }
                ransac_finalize_batch_kernel<<<(unsigned)nsel, FIN_THREADS, 0, st>>>(d_probs, d_spre, d_sel, d_pc, d_pe, d_T, max_d,
                                                                                   d_i2, d_err, d_ci);
                IBL_LAUNCH_CHECK();
                h_inl2.resize(nsel); h_cinl.resize(nsel); h_err.resize(nsel); h_T.resize((size_t)nsel * 16);
                IBL_CUDA(cudaMemcpyAsync(h_inl2.data(), d_i2, sizeof(int32_t) * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(cudaMemcpyAsync(h_cinl.data(), d_ci, sizeof(int32_t) * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(cudaMemcpyAsync(h_err.data(), d_err, sizeof(double) * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(cudaMemcpyAsync(h_T.data(), d_T, sizeof(double) * 16 * nsel, cudaMemcpyDeviceToHost, st));
                IBL_CUDA(ibl_sync(st));                                                      // sync 2: exact scores
                ibl_free(d_sel, st); ibl_free(d_spre, st); ibl_free(d_pc, st); ibl_free(d_pe, st);
                ibl_free(d_i2, st); ibl_free(d_ci, st); ibl_free(d_err, st); ibl_free(d_T, st);
            }
            // ---- the ordered rule, per problem (ibl_ransac_run_range's loop body)
            for (int p = 0; p < P && rc == IBL_OK; ++p) {
                ibl_reg_result_t &r = h_results[p];
                Host &h = hs[p];
                for (const Cand &c : cands[p]) {
                    const int64_t hid = hp[p].h0 + c.k;
                    if (hid >= h.est_k) break;
                    r.validated++;
                    const int j = c.sel;
                    if (j < 0) continue;
                    if (h_inl2[j] != c.inl) { ibl_set_error("ransac batch: count pass and exact pass disagree"); rc = IBL_E_CUDA; break; }
                    const double n_src = (double)hp[p].n_src;
                    double fit = (double)h_inl2[j] / n_src;
                    double rmse = h_inl2[j] > 0 ? sqrt(h_err[j] / (double)h_inl2[j]) : 0.0;
                    if (fit > h.best_fit || (fit == h.best_fit && rmse < h.best_rmse)) {
                        h.best_fit = fit; h.best_rmse = rmse;
                        r.fitness = fit; r.inlier_rmse = rmse; r.n_inliers = h_inl2[j]; r.err2 = h_err[j]; r.best_hyp = hid;
                        memcpy(r.T, &h_T[(size_t)j * 16], sizeof(double) * 16);
                        if (early_exit) {
                            double rho = (double)h_cinl[j] / (double)hp[p].n_corres;
                            double den = rho < 1.0 ? log(1.0 - rho * rho * rho) : -INFINITY;
                            double kk = den == 0.0 ? INFINITY : log(1.0 - confidence) / den;
                            if (kk < (double)h.est_k) h.est_k = (int64_t)ceil(kk);
                        }
                    }
                }
                cands[p].clear();
            }
        }
        ibl_free(d_ent, st); ibl_free(d_inl, st); ibl_free(d_edge, st);
        if (rc < 0) break;
        for (int p = 0; p < P; ++p) {
            if (hp[p].H == 0) continue;
            hs[p].h0 += hp[p].H;
            hs[p].cur_block = std::min<int64_t>(hs[p].cur_block * RB_GROWTH, RB_MAX_BLOCK);
        }
    }
    for (int p = 0; p < P; ++p)
        if (h_results[p].status == 0)
            h_results[p].drawn = std::max<int64_t>(std::min(max_iteration, std::max<int64_t>(hs[p].est_k, 0)), 0);
    ibl_free(d_probs, st); ibl_free(d_n, st); ibl_free(d_ne, st); ibl_free(d_next, st);
    return rc;
}
