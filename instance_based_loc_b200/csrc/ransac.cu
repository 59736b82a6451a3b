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

__global__ void __launch_bounds__(128) ransac_check_kernel(const double *__restrict__ src, const double *__restrict__ tgt,
                                                           const int32_t *__restrict__ corres, int64_t n_corres,
                                                           const int32_t *__restrict__ samples, uint64_t seed, int64_t h0,
                                                           int64_t H, double edge_thr, double dist_thr, int use_checkers,
                                                           uint8_t *__restrict__ pass, double *__restrict__ T_out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= H) return;
    int64_t h = h0 + k;
    double s[9], t[9], T[16];
    gather_sample(src, tgt, corres, n_corres, samples, seed, h, s, t);
    bool ok = true;
    if (use_checkers) {
        for (int i = 0; i < 3 && ok; ++i)
            for (int j = i + 1; j < 3; ++j) {
                double a[3] = {s[3 * i] - s[3 * j], s[3 * i + 1] - s[3 * j + 1], s[3 * i + 2] - s[3 * j + 2]};
                double b[3] = {t[3 * i] - t[3 * j], t[3 * i + 1] - t[3 * j + 1], t[3 * i + 2] - t[3 * j + 2]};
                double ds = sqrt(ibl_dot(a, a)), dt = sqrt(ibl_dot(b, b));
                if (ds < dt * edge_thr || dt < ds * edge_thr) { ok = false; break; }
            }
    }
    if (ok) {
        ibl_kabsch3(s, t, T);
        if (use_checkers) {
            for (int j = 0; j < 3; ++j) {
                double qx, qy, qz;
                ibl_xform(T, s[3 * j], s[3 * j + 1], s[3 * j + 2], qx, qy, qz);
                double d[3] = {t[3 * j] - qx, t[3 * j + 1] - qy, t[3 * j + 2] - qz};
                if (sqrt(ibl_dot(d, d)) > dist_thr) { ok = false; break; }
            }
        }
        if (ok && T_out) {
#pragma unroll
            for (int e = 0; e < 16; ++e) T_out[16 * k + e] = T[e];
        }
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
