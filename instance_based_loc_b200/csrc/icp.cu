// icp.cu -- RegistrationICP loop, point-to-point and coloured estimators, colour-gradient
// initialisation (Open3D Registration.cpp / ColoredICP.cpp as called from
// utils/fpfh_register.py:132-141).
// Normal equations are reduced with warp shuffles (fixed tree) -> CTA -> per-CTA partials that the
// host adds in index order: deterministic run to run.
#include <algorithm>
#include <vector>

#include "ibl_common.cuh"
#include "rigid.cuh"

// ------------------------------------------------------------------------------------------------
// generic K-wide deterministic reduction: f(i, v[K]) fills the contribution of element i
template <int K, class F>
__global__ void __launch_bounds__(256) reduceK_kernel(int64_t n, F f, double *__restrict__ part) {
    double acc[K];
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double v[K];
        if (f(i, v)) {
#pragma unroll
            for (int k = 0; k < K; ++k) acc[k] += v[k];
        }
    }
    __shared__ double sm[8][K];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double x = acc[k];
        for (int o = 16; o; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
        if (lane == 0) sm[w][k] = x;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        double t = 0;
        for (int ww = 0; ww < 8; ++ww) t += sm[ww][k];
        part[(size_t)blockIdx.x * K + k] = t;
    }
}

template <int K, class F>
static int reduceK(int64_t n, F f, double *h_out, cudaStream_t st) {
    for (int k = 0; k < K; ++k) h_out[k] = 0.0;
    if (n == 0) return IBL_OK;
    int nb = (int)((n + 1023) / 1024);
    if (nb > IBL_SMS * 2) nb = IBL_SMS * 2;
    double *part = nullptr;
    IBL_CUDA(ibl_alloc(&part, (size_t)nb * K, st));
    {
        IBL_PROF(IBL_PROF_ICP_REDUCE, st, (double)n * (28.0 + 8.0 * K));
        reduceK_kernel<K, F><<<nb, 256, 0, st>>>(n, f, part);
        IBL_LAUNCH_CHECK();
    }
    std::vector<double> hp((size_t)nb * K);
    IBL_CUDA(cudaMemcpyAsync(hp.data(), part, sizeof(double) * nb * K, cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(part, st);
    for (int b = 0; b < nb; ++b)
        for (int k = 0; k < K; ++k) h_out[k] += hp[(size_t)b * K + k];
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
struct P2PSums {   // count, sum s(3), sum t(3)
    const double *src; const int32_t *nn; const double *tgt;
    __device__ bool operator()(int64_t i, double *v) const {
        int j = nn[i];
        if (j < 0) return false;
        v[0] = 1.0;
        for (int a = 0; a < 3; ++a) { v[1 + a] = src[3 * i + a]; v[4 + a] = tgt[3 * (size_t)j + a]; }
        return true;
    }
};
struct P2PCov {    // sum (t-mt)(s-ms)^T
    const double *src; const int32_t *nn; const double *tgt; double ms[3], mt[3];
    __device__ bool operator()(int64_t i, double *v) const {
        int j = nn[i];
        if (j < 0) return false;
        double ds[3], dt[3];
        for (int a = 0; a < 3; ++a) { ds[a] = src[3 * i + a] - ms[a]; dt[a] = tgt[3 * (size_t)j + a] - mt[a]; }
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) v[3 * a + b] = dt[a] * ds[b];
        return true;
    }
};

IBL_API int ibl_p2p_update(const double *src, const int32_t *nn_idx, int64_t n, const double *tgt, double *h_T16,
                           void *stream) {
    IBL_ARG(h_T16 && n >= 0);
    for (int k = 0; k < 16; ++k) h_T16[k] = (k % 5 == 0) ? 1.0 : 0.0;
    if (n == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    double s7[7];
    P2PSums f1{src, nn_idx, tgt};
    int rc = reduceK<7>(n, f1, s7, st);
    if (rc < 0) return rc;
    if (s7[0] == 0.0) return IBL_OK;   // corres.empty() -> Identity
    P2PCov f2{src, nn_idx, tgt, {0, 0, 0}, {0, 0, 0}};
    for (int a = 0; a < 3; ++a) { f2.ms[a] = s7[1 + a] / s7[0]; f2.mt[a] = s7[4 + a] / s7[0]; }
    double c9[9];
    rc = reduceK<9>(n, f2, c9, st);
    if (rc < 0) return rc;
    double sg[3][3];
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) sg[a][b] = c9[3 * a + b] / s7[0];
    ibl_kabsch_finish(sg, f2.ms, f2.mt, h_T16);
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// A.8 colour gradient: thread per target point, rows accumulated in neighbour-list order
__global__ void __launch_bounds__(128) color_gradient_kernel(const double *__restrict__ pts, const double *__restrict__ nrm,
                                                             const double *__restrict__ col, int64_t n,
                                                             const int32_t *__restrict__ nbr_idx,
                                                             const int32_t *__restrict__ nbr_cnt, int max_nn,
                                                             double *__restrict__ grad) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    double g[3] = {0, 0, 0};
    int nn = nbr_cnt[k];
    if (nn >= 4) {
        double vt[3] = {pts[3 * k], pts[3 * k + 1], pts[3 * k + 2]};
        double nt[3] = {nrm[3 * k], nrm[3 * k + 1], nrm[3 * k + 2]};
        double it = (col[3 * k] + col[3 * k + 1] + col[3 * k + 2]) / 3.0;
        double AtA[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Atb[3] = {0, 0, 0};
        for (int i = 1; i < nn; ++i) {
            size_t adj = (size_t)nbr_idx[k * max_nn + i];
            double va[3] = {pts[3 * adj], pts[3 * adj + 1], pts[3 * adj + 2]};
            double d[3] = {va[0] - vt[0], va[1] - vt[1], va[2] - vt[2]};
            double pr = ibl_dot(d, nt);
            double row[3] = {(va[0] - pr * nt[0]) - vt[0], (va[1] - pr * nt[1]) - vt[1], (va[2] - pr * nt[2]) - vt[2]};
            double ia = (col[3 * adj] + col[3 * adj + 1] + col[3 * adj + 2]) / 3.0;
            double bi = ia - it;
            for (int a = 0; a < 3; ++a) {
                for (int b = 0; b < 3; ++b) AtA[3 * a + b] += row[a] * row[b];
                Atb[a] += row[a] * bi;
            }
        }
        double w = (double)(nn - 1);
        double row[3] = {w * nt[0], w * nt[1], w * nt[2]};
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) AtA[3 * a + b] += row[a] * row[b];
        double x[3];
        if (ibl_ldlt_solve<3>(AtA, Atb, x)) { g[0] = x[0]; g[1] = x[1]; g[2] = x[2]; }
    }
    grad[3 * k] = g[0]; grad[3 * k + 1] = g[1]; grad[3 * k + 2] = g[2];
}

IBL_API int ibl_color_gradient(const double *pts, const double *nrm, const double *col, int64_t n, const int32_t *nbr_idx,
                               const int32_t *nbr_cnt, int32_t max_nn, double *grad, void *stream) {
    IBL_ARG(n >= 0);
    if (n == 0) return IBL_OK;
    IBL_ARG(pts && nrm && col && nbr_idx && nbr_cnt && grad);
    IBL_PROF(IBL_PROF_COLOR_GRAD, ibl_stream(stream), (double)n * (72.0 + 52.0 * max_nn + 24.0));
    color_gradient_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ibl_stream(stream)>>>(pts, nrm, col, n, nbr_idx, nbr_cnt, max_nn, grad);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// one correspondence (source point vs, its colour row, target index ct) -> 21 upper-triangle JTJ + 6 JTr
struct ColoredTerms {
    const double *tgt, *tnrm, *tcol, *tgrad; double sg, sp;
    __device__ __forceinline__ void operator()(const double *vs, double is, size_t ct, double *v) const {
        double vt[3] = {tgt[3 * ct], tgt[3 * ct + 1], tgt[3 * ct + 2]};
        double nt[3] = {tnrm[3 * ct], tnrm[3 * ct + 1], tnrm[3 * ct + 2]};
        double J[2][6], r[2], cr[3];
        ibl_cross(vs, nt, cr);
        for (int a = 0; a < 3; ++a) { J[0][a] = sg * cr[a]; J[0][3 + a] = sg * nt[a]; }
        double d[3] = {vs[0] - vt[0], vs[1] - vt[1], vs[2] - vt[2]};
        double dn = ibl_dot(d, nt);
        r[0] = sg * dn;
        double vp[3] = {vs[0] - dn * nt[0], vs[1] - dn * nt[1], vs[2] - dn * nt[2]};
        double it = (tcol[3 * ct] + tcol[3 * ct + 1] + tcol[3 * ct + 2]) / 3.0;
        double dit[3] = {tgrad[3 * ct], tgrad[3 * ct + 1], tgrad[3 * ct + 2]};
        double dv[3] = {vp[0] - vt[0], vp[1] - vt[1], vp[2] - vt[2]};
        double is0 = ibl_dot(dit, dv) + it;
        double dm[3];
        for (int b = 0; b < 3; ++b) {
            double s = 0.0;
            for (int a = 0; a < 3; ++a) s += dit[a] * ((a == b ? 1.0 : 0.0) - nt[a] * nt[b]);
            dm[b] = -s;
        }
        ibl_cross(vs, dm, cr);
        for (int a = 0; a < 3; ++a) { J[1][a] = sp * cr[a]; J[1][3 + a] = sp * dm[a]; }
        r[1] = sp * (is - is0);
        int o = 0;
        for (int a = 0; a < 6; ++a)
            for (int b = a; b < 6; ++b) v[o++] = J[0][a] * J[0][b] + J[1][a] * J[1][b];
        for (int a = 0; a < 6; ++a) v[21 + a] = J[0][a] * r[0] + J[1][a] * r[1];
    }
};
struct ColoredJTJ {   // reduceK functor over a stored correspondence array
    const double *src, *scol; const int32_t *nn; ColoredTerms t;
    __device__ bool operator()(int64_t i, double *v) const {
        int j = nn[i];
        if (j < 0) return false;
        double vs[3] = {src[3 * i], src[3 * i + 1], src[3 * i + 2]};
        double is = (scol[3 * i] + scol[3 * i + 1] + scol[3 * i + 2]) / 3.0;
        t(vs, is, (size_t)j, v);
        return true;
    }
};

static void solve_update_host(const double *v27, double *T16) {
    double A[36], nb[6], x[6];
    int o = 0;
    for (int a = 0; a < 6; ++a)
        for (int b = a; b < 6; ++b) { A[6 * a + b] = v27[o]; A[6 * b + a] = v27[o]; ++o; }
    for (int a = 0; a < 6; ++a) nb[a] = -v27[21 + a];
    for (int k = 0; k < 16; ++k) T16[k] = (k % 5 == 0) ? 1.0 : 0.0;
    if (!ibl_ldlt_solve<6>(A, nb, x)) return;
    double cx = cos(x[0]), sx = sin(x[0]), cy = cos(x[1]), sy = sin(x[1]), cz = cos(x[2]), sz = sin(x[2]);
    T16[0] = cz * cy; T16[1] = cz * sy * sx - sz * cx; T16[2] = cz * sy * cx + sz * sx; T16[3] = x[3];
    T16[4] = sz * cy; T16[5] = sz * sy * sx + cz * cx; T16[6] = sz * sy * cx - cz * sx; T16[7] = x[4];
    T16[8] = -sy;     T16[9] = cy * sx;                T16[10] = cy * cx;               T16[11] = x[5];
}

IBL_API int ibl_colored_update(const double *src, const double *scol, const int32_t *nn_idx, int64_t n, const double *tgt,
                               const double *tnrm, const double *tcol, const double *tgrad, double lambda_geometric,
                               double *h_out27, double *h_T16, void *stream) {
    IBL_ARG(n >= 0 && h_out27);
    ColoredJTJ f{src, scol, nn_idx, {tgt, tnrm, tcol, tgrad, sqrt(lambda_geometric), sqrt(1.0 - lambda_geometric)}};
    int rc = reduceK<27>(n, f, h_out27, ibl_stream(stream));
    if (rc < 0) return rc;
    if (h_T16) solve_update_host(h_out27, h_T16);
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
static void matmul4(const double *A, const double *B, double *C) {
    double t[16];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0;
            for (int k = 0; k < 4; ++k) s += A[4 * i + k] * B[4 * k + j];
            t[4 * i + j] = s;
        }
    memcpy(C, t, sizeof(t));
}
static bool is_identity(const double *T) {
    for (int k = 0; k < 16; ++k) if (T[k] != ((k % 5 == 0) ? 1.0 : 0.0)) return false;
    return true;
}

// ------------------------------------------------------------------------------------------------
// One coloured-ICP iteration in ONE kernel (A.8 loop body), and the loop's host rule moved onto the device so that
// iterations can be queued back to back.  Every thread moves its source point by the update of the previous solve, finds
// its nearest target within max_d (exact NN field / ring search), and contributes to BOTH the registration score of the
// moved cloud (inlier count, err2 -- reduced in nn_score_kernel / score_final_kernel's order, so fitness and rmse are
// bit-identical to evaluate_registration) and the normal equations of the NEXT solve.  The last CTA to finish adds the
// per-CTA partials in a fixed order, applies the convergence rule, solves the 6x6 system and composes the transformation
// (IcpState).  Kernels launched after convergence return at once, so the host queues a batch of iterations and looks at
// the state once per batch: 2-4 host round trips per registration instead of two per iteration.
#define STEP_W 29
struct IcpState {
    double T[16];            // accumulated transformation (what RegistrationResult reports)
    double U[16];            // update the next step applies to the moved cloud
    double fit, rmse, err2;
    long long n_inliers;
    int steps;               // steps done; step 0 scores the initial pose, step s >= 1 is ICP iteration s
    int done;                // converged, or max_iteration reached
    int has_U;
    unsigned arrived;        // CTAs that have written their partials in the running step
};

__device__ void icp_solve_update(const double *v27, double *T16) {
    double A[36], nb[6], x[6];
    int o = 0;
    for (int a = 0; a < 6; ++a)
        for (int b = a; b < 6; ++b) { A[6 * a + b] = v27[o]; A[6 * b + a] = v27[o]; ++o; }
    for (int a = 0; a < 6; ++a) nb[a] = -v27[21 + a];
    for (int k = 0; k < 16; ++k) T16[k] = (k % 5 == 0) ? 1.0 : 0.0;
    if (!ibl_ldlt_solve<6>(A, nb, x)) return;
    double cx = cos(x[0]), sx = sin(x[0]), cy = cos(x[1]), sy = sin(x[1]), cz = cos(x[2]), sz = sin(x[2]);
    T16[0] = cz * cy; T16[1] = cz * sy * sx - sz * cx; T16[2] = cz * sy * cx + sz * sx; T16[3] = x[3];
    T16[4] = sz * cy; T16[5] = sz * sy * sx + cz * cx; T16[6] = sz * sy * cx - cz * sx; T16[7] = x[4];
    T16[8] = -sy;     T16[9] = cy * sx;                T16[10] = cy * cx;               T16[11] = x[5];
}

// The body of one CTA (256 threads) of a coloured-ICP step: block `blk` of the problem's `nb` blocks.  in == nullptr: read
// the problem's own source on step 0 and the moved cloud `out` afterwards (batched launches: the choice is per problem).
__device__ __forceinline__ void colored_step_block(const GridView &g, const double *__restrict__ in, const double *__restrict__ src0,
                                                   double *__restrict__ out, const double *__restrict__ scol, int64_t n, IcpState *S,
                                                   const ColoredTerms &terms, double max_d, double max_d2,
                                                   int32_t *__restrict__ nn, double *part, int nb, int blk,
                                                   int max_iteration, double rel_fitness, double rel_rmse) {
    __shared__ double sU[16];
    __shared__ int s_flag[3];
    if (threadIdx.x < 16) sU[threadIdx.x] = S->U[threadIdx.x];
    if (threadIdx.x == 0) { s_flag[0] = S->done; s_flag[1] = S->has_U; s_flag[2] = S->steps; }
    __syncthreads();
    if (s_flag[0]) return;                                // converged in an earlier step of this batch
    const int has_U = s_flag[1];
    if (in == nullptr) in = s_flag[2] == 0 ? src0 : out;
    int64_t i = blk * (int64_t)blockDim.x + threadIdx.x;
    int hit = 0;
    double e = 0.0;
    double v[27];
#pragma unroll
    for (int k = 0; k < 27; ++k) v[k] = 0.0;
    if (i < n) {
        double p[3] = {in[3 * i], in[3 * i + 1], in[3 * i + 2]};
        if (has_U) { double a, b, c; ibl_xform(sU, p[0], p[1], p[2], a, b, c); p[0] = a; p[1] = b; p[2] = c; }
        out[3 * i] = p[0]; out[3 * i + 1] = p[1]; out[3 * i + 2] = p[2];
        double bd;
        int j = ibl_nn1_thread(g, p[0], p[1], p[2], max_d, max_d2, bd);
        if (nn) nn[i] = j;
        if (j >= 0) {
            hit = 1; e = bd;
            double is = (scol[3 * i] + scol[3 * i + 1] + scol[3 * i + 2]) / 3.0;
            terms(p, is, (size_t)j, v);
        }
    }
    __shared__ double se[8];
    __shared__ int sc[8];
    __shared__ double sm[8][27];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) { e += __shfl_down_sync(0xffffffffu, e, o); hit += __shfl_down_sync(0xffffffffu, hit, o); }
    if (lane == 0) { se[w] = e; sc[w] = hit; }
#pragma unroll
    for (int k = 0; k < 27; ++k) {
        double x = v[k];
        for (int o = 16; o; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
        if (lane == 0) sm[w][k] = x;
    }
    __syncthreads();
    double *row = part + (size_t)blk * STEP_W;
    if (threadIdx.x == 0) {
        double te = 0; long long tc = 0;
        for (int k = 0; k < 8; ++k) { te += se[k]; tc += sc[k]; }
        row[0] = (double)tc;
        row[1] = te;
    }
    if (threadIdx.x < 27) {
        double t = 0;
        for (int ww = 0; ww < 8; ++ww) t += sm[ww][threadIdx.x];
        row[2 + threadIdx.x] = t;
    }
    // ---- last CTA: fixed-order reduction of the partials, convergence rule, next update ----
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_flag[0] = (atomicAdd(&S->arrived, 1u) == (unsigned)(nb - 1));
    __syncthreads();
    if (!s_flag[0]) return;
    __threadfence();
    __shared__ double s_tot[2 + 27];
    if (w == 0) {                                          // score_final_kernel's order: lane-strided sums, shuffle tree
        double le = 0; long long lc = 0;
        for (int k = lane; k < nb; k += 32) { le += __ldcg(part + (size_t)k * STEP_W + 1); lc += (long long)__ldcg(part + (size_t)k * STEP_W); }
        for (int o = 16; o; o >>= 1) { le += __shfl_down_sync(0xffffffffu, le, o); lc += __shfl_down_sync(0xffffffffu, lc, o); }
        if (lane == 0) { s_tot[0] = (double)lc; s_tot[1] = le; }
    } else if (threadIdx.x < 32 + 27) {                    // normal equations: CTA partials in index order
        int k = threadIdx.x - 32;
        double t = 0;
        for (int b = 0; b < nb; ++b) t += __ldcg(part + (size_t)b * STEP_W + 2 + k);
        s_tot[2 + k] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const long long c = (long long)s_tot[0];
        const double err2 = s_tot[1];
        const double fit = (double)c / (double)n;
        const double rmse = c > 0 ? sqrt(err2 / (double)c) : 0.0;
        const int steps = S->steps;
        int done = 0;
        if (steps >= 1 && fabs(S->fit - fit) < rel_fitness && fabs(S->rmse - rmse) < rel_rmse) done = 1;
        if (steps >= max_iteration) done = 1;
        S->fit = fit; S->rmse = rmse; S->err2 = err2; S->n_inliers = c; S->steps = steps + 1;
        if (!done) {
            double U[16], T[16], R[16];
            icp_solve_update(s_tot + 2, U);
            for (int k = 0; k < 16; ++k) T[k] = S->T[k];
            for (int a = 0; a < 4; ++a)
                for (int b = 0; b < 4; ++b) {
                    double t = 0;
                    for (int k = 0; k < 4; ++k) t += U[4 * a + k] * T[4 * k + b];
                    R[4 * a + b] = t;
                }
            bool ident = true;
            for (int k = 0; k < 16; ++k) { S->T[k] = R[k]; S->U[k] = U[k]; ident = ident && (U[k] == ((k % 5 == 0) ? 1.0 : 0.0)); }
            S->has_U = ident ? 0 : 1;
        }
        S->arrived = 0;
        S->done = done;
    }
}

__global__ void __launch_bounds__(256) colored_step_kernel(GridView g, const double *__restrict__ in, double *__restrict__ out,
                                                           const double *__restrict__ scol, int64_t n, IcpState *S,
                                                           ColoredTerms terms, double max_d, double max_d2,
                                                           int32_t *__restrict__ nn, double *part, int nb,
                                                           int max_iteration, double rel_fitness, double rel_rmse) {
    colored_step_block(g, in, in, out, scol, n, S, terms, max_d, max_d2, nn, part, nb, (int)blockIdx.x, max_iteration, rel_fitness,
                       rel_rmse);
}

// ---- batched: all coloured-ICP problems of a step advance by one iteration per launch ----
struct IcpProb {
    GridView g;
    const double *src; double *cur; const double *scol;
    int64_t n;
    ColoredTerms terms;
    IcpState *S;
    double *part;
    int nb;
};

__global__ void __launch_bounds__(256) colored_step_batch_kernel(const IcpProb *__restrict__ probs, int nprob,
                                                                 const int64_t *__restrict__ blk_prefix, double max_d,
                                                                 double max_d2, int max_iteration, double rel_fitness,
                                                                 double rel_rmse) {
    __shared__ GridView g;
    __shared__ int s_p;
    if (threadIdx.x == 0) {
        int lo = 0, hi = nprob;
        const int64_t item = (int64_t)blockIdx.x;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (blk_prefix[mid] <= item) lo = mid; else hi = mid; }
        s_p = lo;
    }
    __syncthreads();
    const IcpProb *P = probs + s_p;
    for (int k = threadIdx.x; k < (int)(sizeof(GridView) / 4); k += 256)
        reinterpret_cast<int *>(&g)[k] = reinterpret_cast<const int *>(&P->g)[k];
    __syncthreads();
    colored_step_block(g, nullptr, P->src, P->cur, P->scol, P->n, P->S, P->terms, max_d, max_d2, nullptr, P->part, P->nb,
                       (int)((int64_t)blockIdx.x - blk_prefix[s_p]), max_iteration, rel_fitness, rel_rmse);
}

static int icp_colored_fused(const ibl_grid_t *tgt_grid, const double *src, const double *scol, int64_t n_src,
                             const double *tgt, const double *tnrm, const double *tcol, const double *tgrad, double max_d,
                             double *T, int max_iteration, double rel_fitness, double rel_rmse, double lambda_geometric,
                             ibl_reg_result_t *h_result, cudaStream_t st) {
    double *pcd = nullptr, *part = nullptr; int32_t *nn = nullptr; IcpState *dS = nullptr;
    const int nb = (int)((n_src + 255) / 256);
    IBL_CUDA(ibl_alloc(&pcd, (size_t)n_src * 3, st));
    IBL_CUDA(ibl_alloc(&nn, (size_t)n_src, st));
    IBL_CUDA(ibl_alloc(&part, (size_t)nb * STEP_W, st));
    IBL_CUDA(ibl_alloc(&dS, 1, st));
    IcpState hS;
    memset(&hS, 0, sizeof(hS));
    memcpy(hS.T, T, sizeof(hS.T));
    memcpy(hS.U, T, sizeof(hS.U));                         // step 0 moves the source by the initial transformation
    hS.has_U = is_identity(T) ? 0 : 1;
    IBL_CUDA(cudaMemcpyAsync(dS, &hS, sizeof(hS), cudaMemcpyHostToDevice, st));
    ColoredTerms terms{tgt, tnrm, tcol, tgrad, sqrt(lambda_geometric), sqrt(1.0 - lambda_geometric)};
    const GridView gv = ibl_view(tgt_grid);
    auto launch = [&](const double *in) -> int {
        IBL_PROF(IBL_PROF_ICP_REDUCE, st, (double)n_src * (24.0 + 24.0 + 24.0 + 4.0 + 96.0) + 8.0 * STEP_W * nb);
        colored_step_kernel<<<nb, 256, 0, st>>>(gv, in, pcd, scol, n_src, dS, terms, max_d, max_d * max_d, nn, part, nb,
                                                max_iteration, rel_fitness, rel_rmse);
        IBL_LAUNCH_CHECK();
        return IBL_OK;
    };
    int rc = launch(src);                                  // step 0: score of the initial pose + first normal equations
    if (rc < 0) return rc;
    int queued = 0;
    for (;;) {
        int batch = std::min(queued == 0 ? 7 : 8, max_iteration - queued);
        for (int k = 0; k < batch; ++k) if ((rc = launch(pcd)) < 0) return rc;
        queued += batch;
        IBL_CUDA(cudaMemcpyAsync(&hS, dS, sizeof(hS), cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));
        if (hS.done || queued >= max_iteration) break;
    }
    memcpy(T, hS.T, sizeof(hS.T));
    memcpy(h_result->T, hS.T, sizeof(double) * 16);
    h_result->fitness = hS.fit; h_result->inlier_rmse = hS.rmse;
    h_result->n_inliers = hS.n_inliers; h_result->err2 = hS.err2; h_result->iterations = hS.steps - 1;
    ibl_free(pcd, st); ibl_free(nn, st); ibl_free(part, st); ibl_free(dS, st);
    return IBL_OK;
}

IBL_API int ibl_icp(const ibl_grid_t *tgt_grid, const double *src, const double *scol, int64_t n_src, const double *tgt,
                    const double *tnrm, const double *tcol, const double *tgrad, int mode, double max_d,
                    const double *h_T_init, int max_iteration, double rel_fitness, double rel_rmse, double lambda_geometric,
                    ibl_reg_result_t *h_result, void *stream) {
    IBL_ARG(tgt_grid && h_result && n_src >= 0 && max_d > 0 && (mode == 0 || (scol && tnrm && tcol && tgrad)));
    cudaStream_t st = ibl_stream(stream);
    memset(h_result, 0, sizeof(*h_result));
    double T[16];
    for (int k = 0; k < 16; ++k) T[k] = h_T_init ? h_T_init[k] : ((k % 5 == 0) ? 1.0 : 0.0);
    memcpy(h_result->T, T, sizeof(T));
    h_result->best_hyp = -1;
    if (n_src == 0) { h_result->status = 1; return IBL_OK; }
    if (mode == 1 && tgt_grid->n > 0)
        return icp_colored_fused(tgt_grid, src, scol, n_src, tgt, tnrm, tcol, tgrad, max_d, T, max_iteration, rel_fitness,
                                 rel_rmse, lambda_geometric, h_result, st);
    double *pcd = nullptr; int32_t *nn = nullptr;
    IBL_CUDA(ibl_alloc(&pcd, (size_t)n_src * 3, st));
    IBL_CUDA(ibl_alloc(&nn, (size_t)n_src, st));
    int rc;
    if (is_identity(T)) {
        IBL_CUDA(cudaMemcpyAsync(pcd, src, sizeof(double) * 3 * n_src, cudaMemcpyDeviceToDevice, st));
    } else if ((rc = ibl_transform_points(T, src, n_src, 0, pcd, stream)) < 0) return rc;
    ibl_score_t sc;
    if ((rc = ibl_nn_score(tgt_grid, pcd, n_src, nullptr, max_d, nn, nullptr, &sc, stream)) < 0) return rc;
    double fit = (double)sc.n_inliers / (double)n_src;
    double rmse = sc.n_inliers > 0 ? sqrt(sc.err2 / (double)sc.n_inliers) : 0.0;
    int it = 0;
    for (; it < max_iteration;) {
        double upd[16], v27[27];
        if (mode == 0) rc = ibl_p2p_update(pcd, nn, n_src, tgt, upd, stream);
        else rc = ibl_colored_update(pcd, scol, nn, n_src, tgt, tnrm, tcol, tgrad, lambda_geometric, v27, upd, stream);
        if (rc < 0) return rc;
        matmul4(upd, T, T);
        if ((rc = ibl_transform_points(upd, pcd, n_src, 0, pcd, stream)) < 0) return rc;
        double pf = fit, pr = rmse;
        if ((rc = ibl_nn_score(tgt_grid, pcd, n_src, nullptr, max_d, nn, nullptr, &sc, stream)) < 0) return rc;
        fit = (double)sc.n_inliers / (double)n_src;
        rmse = sc.n_inliers > 0 ? sqrt(sc.err2 / (double)sc.n_inliers) : 0.0;
        ++it;
        if (fabs(pf - fit) < rel_fitness && fabs(pr - rmse) < rel_rmse) break;
    }
    memcpy(h_result->T, T, sizeof(T));
    h_result->fitness = fit; h_result->inlier_rmse = rmse;
    h_result->n_inliers = sc.n_inliers; h_result->err2 = sc.err2; h_result->iterations = it;
    ibl_free(pcd, st); ibl_free(nn, st);
    return IBL_OK;
}

// All coloured-ICP problems of a step (object_memory.py:1020-1108: one per assignment and frame) iterate together: one
// launch per ICP iteration for the whole batch, the per-problem convergence rule and 6x6 solve on the device (the last CTA
// of every problem), and one read-back of the problem states after 12 iterations and one at the end -- instead of 2-4 host round trips
// per problem.  The per-problem arithmetic and reduction order are colored_step_kernel's, so results equal ibl_icp(mode 1).
IBL_API int ibl_icp_colored_batch(const ibl_reg_problem_t *problems, int32_t n_problems, const double *h_T_init, double max_d,
                                  int max_iteration, double rel_fitness, double rel_rmse, double lambda_geometric,
                                  ibl_reg_result_t *h_results, void *stream) {
    IBL_ARG(n_problems >= 0 && (n_problems == 0 || (problems && h_results && h_T_init)) && max_d > 0 && max_iteration >= 0);
    if (n_problems == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    const int P = n_problems;
    std::vector<IcpState> hS(P);
    std::vector<IcpProb> hp(P);
    std::vector<int> live;
    int64_t npts = 0, nblk = 0;
    for (int p = 0; p < P; ++p) {
        const ibl_reg_problem_t &q = problems[p];
        ibl_reg_result_t &r = h_results[p];
        memset(&r, 0, sizeof(r));
        memcpy(r.T, h_T_init + 16 * (size_t)p, sizeof(double) * 16);
        r.best_hyp = -1;
        if (q.n_src <= 0) { r.status = 1; continue; }
        IBL_ARG(q.grid_icp && q.src && q.src_col && q.tgt && q.tgt_nrm && q.tgt_col && q.tgt_grad && q.n_tgt > 0);
        live.push_back(p);
        npts += q.n_src; nblk += (q.n_src + 255) / 256;
    }
    if (live.empty()) return IBL_OK;
    double *cur = nullptr, *part = nullptr; IcpState *dS = nullptr; IcpProb *d_probs = nullptr; int64_t *d_prefix = nullptr;
    IBL_CUDA(ibl_alloc(&cur, (size_t)npts * 3, st));
    IBL_CUDA(ibl_alloc(&part, (size_t)nblk * STEP_W, st));
    IBL_CUDA(ibl_alloc(&dS, (size_t)P, st));
    IBL_CUDA(ibl_alloc(&d_probs, (size_t)P, st));
    IBL_CUDA(ibl_alloc(&d_prefix, (size_t)P + 1, st));
    memset(hS.data(), 0, sizeof(IcpState) * P);
    {
        int64_t po = 0, bo = 0;
        for (int p : live) {
            const ibl_reg_problem_t &q = problems[p];
            IcpProb &d = hp[p];
            d.g = ibl_view(q.grid_icp);
            d.src = q.src; d.cur = cur + 3 * po; d.scol = q.src_col; d.n = q.n_src;
            d.terms = ColoredTerms{q.tgt, q.tgt_nrm, q.tgt_col, q.tgt_grad, sqrt(lambda_geometric), sqrt(1.0 - lambda_geometric)};
            d.S = dS + p; d.nb = (int)((q.n_src + 255) / 256); d.part = part + (size_t)bo * STEP_W;
            po += q.n_src; bo += d.nb;
            memcpy(hS[p].T, h_T_init + 16 * (size_t)p, sizeof(double) * 16);
            memcpy(hS[p].U, h_T_init + 16 * (size_t)p, sizeof(double) * 16);      // step 0 moves the source by the initial pose
            hS[p].has_U = is_identity(hS[p].T) ? 0 : 1;
        }
    }
    IBL_CUDA(cudaMemcpyAsync(dS, hS.data(), sizeof(IcpState) * P, cudaMemcpyHostToDevice, st));
    std::vector<IcpProb> table;
    std::vector<int64_t> prefix;
    int queued = -1;                                       // ICP iterations queued so far (step 0 = the initial score)
    while (!live.empty()) {
        table.clear(); prefix.assign(1, 0);
        double work = 0;
        for (int p : live) {
            table.push_back(hp[p]);
            prefix.push_back(prefix.back() + hp[p].nb);
            work += (double)hp[p].n * (24.0 + 24.0 + 24.0 + 4.0 + 96.0) + 8.0 * STEP_W * hp[p].nb;
        }
        const int L = (int)live.size();
        IBL_CUDA(cudaMemcpyAsync(d_probs, table.data(), sizeof(IcpProb) * L, cudaMemcpyHostToDevice, st));
        IBL_CUDA(cudaMemcpyAsync(d_prefix, prefix.data(), sizeof(int64_t) * (L + 1), cudaMemcpyHostToDevice, st));
        // the first read-back after 12 iterations (coloured ICP usually converges in 5-15), then everything that is left:
        // kernels of a converged problem return at once, so queueing past convergence costs a few empty launches, while
        // every read-back is a host round trip for the whole batch
        const int batch = queued < 0 ? std::min(13, max_iteration + 1) : max_iteration - queued;
        for (int k = 0; k < batch; ++k) {
            IBL_PROF(IBL_PROF_ICP_REDUCE, st, work);
            colored_step_batch_kernel<<<(unsigned)prefix.back(), 256, 0, st>>>(d_probs, L, d_prefix, max_d, max_d * max_d,
                                                                             max_iteration, rel_fitness, rel_rmse);
            IBL_LAUNCH_CHECK();
        }
        queued = queued < 0 ? batch - 1 : queued + batch;
        IBL_CUDA(cudaMemcpyAsync(hS.data(), dS, sizeof(IcpState) * P, cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));
        std::vector<int> next;
        for (int p : live) {
            if (hS[p].done || queued >= max_iteration) {
                ibl_reg_result_t &r = h_results[p];
                memcpy(r.T, hS[p].T, sizeof(double) * 16);
                r.fitness = hS[p].fit; r.inlier_rmse = hS[p].rmse; r.n_inliers = hS[p].n_inliers; r.err2 = hS[p].err2;
                r.iterations = hS[p].steps - 1;
            } else {
                next.push_back(p);
            }
        }
        live.swap(next);
    }
    ibl_free(cur, st); ibl_free(part, st); ibl_free(dS, st); ibl_free(d_probs, st); ibl_free(d_prefix, st);
    return IBL_OK;
}
