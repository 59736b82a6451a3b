// features.cu -- normals (FastEigen3x3), SPFH / FPFH and the exact 33-d feature matcher
// (Open3D EstimateNormals.cpp / Feature.cpp / Registration.cpp as called from
// utils/fpfh_register.py:86-119).  Device feature layout is POINT-MAJOR
// double[n][33] (one 264 B row per point: coalesced writes, contiguous gathers); the Python
// mirror exposes the Open3D [33][n] view.
#include <cub/cub.cuh>

#include "ibl_common.cuh"

// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cross3(const double *a, const double *b, double *o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

__device__ void eigvec0(const double *A, double ev, double *out) {
    double r0[3] = {A[0] - ev, A[1], A[2]};
    double r1[3] = {A[1], A[3] - ev, A[4]};
    double r2[3] = {A[2], A[4], A[5] - ev};
    double c01[3], c02[3], c12[3];
    cross3(r0, r1, c01); cross3(r0, r2, c02); cross3(r1, r2, c12);
    double d0 = dot3(c01, c01), d1 = dot3(c02, c02), d2 = dot3(c12, c12);
    double dmax = d0; int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    double s;
    if (imax == 0) { s = sqrt(d0); out[0] = c01[0] / s; out[1] = c01[1] / s; out[2] = c01[2] / s; }
    else if (imax == 1) { s = sqrt(d1); out[0] = c02[0] / s; out[1] = c02[1] / s; out[2] = c02[2] / s; }
    else { s = sqrt(d2); out[0] = c12[0] / s; out[1] = c12[1] / s; out[2] = c12[2] / s; }
}

__device__ void eigvec1(const double *A, const double *e0, double ev1, double *out) {
    double U[3], V[3];
    if (fabs(e0[0]) > fabs(e0[1])) {
        double inv = 1.0 / sqrt(e0[0] * e0[0] + e0[2] * e0[2]);
        U[0] = -e0[2] * inv; U[1] = 0.0; U[2] = e0[0] * inv;
    } else {
        double inv = 1.0 / sqrt(e0[1] * e0[1] + e0[2] * e0[2]);
        U[0] = 0.0; U[1] = e0[2] * inv; U[2] = -e0[1] * inv;
    }
    cross3(e0, U, V);
    double AU[3] = {A[0] * U[0] + A[1] * U[1] + A[2] * U[2], A[1] * U[0] + A[3] * U[1] + A[4] * U[2],
                    A[2] * U[0] + A[4] * U[1] + A[5] * U[2]};
    double AV[3] = {A[0] * V[0] + A[1] * V[1] + A[2] * V[2], A[1] * V[0] + A[3] * V[1] + A[4] * V[2],
                    A[2] * V[0] + A[4] * V[1] + A[5] * V[2]};
    double m00 = dot3(U, AU) - ev1, m01 = dot3(U, AV), m11 = dot3(V, AV) - ev1;
    double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
    if (a00 >= a11) {
        double mx = a00 > a01 ? a00 : a01;
        if (mx > 0) {
            if (a00 >= a01) { m01 /= m00; m00 = 1.0 / sqrt(1.0 + m01 * m01); m01 *= m00; }
            else            { m00 /= m01; m01 = 1.0 / sqrt(1.0 + m00 * m00); m00 *= m01; }
            for (int k = 0; k < 3; ++k) out[k] = m01 * U[k] - m00 * V[k];
        } else { out[0] = U[0]; out[1] = U[1]; out[2] = U[2]; }
    } else {
        double mx = a11 > a01 ? a11 : a01;
        if (mx > 0) {
            if (a11 >= a01) { m01 /= m11; m11 = 1.0 / sqrt(1.0 + m01 * m01); m01 *= m11; }
            else            { m11 /= m01; m01 = 1.0 / sqrt(1.0 + m11 * m11); m11 *= m01; }
            for (int k = 0; k < 3; ++k) out[k] = m11 * U[k] - m01 * V[k];
        } else { out[0] = U[0]; out[1] = U[1]; out[2] = U[2]; }
    }
}

// cov: a00 a01 a02 a11 a12 a22 -> eigenvector of the smallest eigenvalue (A.3)
__device__ void fast_eigen3x3(const double *cov, double *out) {
    double A[6];
    double mx = cov[0];
    for (int k = 1; k < 6; ++k) if (cov[k] > mx) mx = cov[k];
    if (mx == 0.0) { out[0] = out[1] = out[2] = 0.0; return; }
    for (int k = 0; k < 6; ++k) A[k] = cov[k] / mx;
    double norm = A[1] * A[1] + A[2] * A[2] + A[4] * A[4];
    if (norm > 0) {
        double q = (A[0] + A[3] + A[5]) / 3.0;
        double b00 = A[0] - q, b11 = A[3] - q, b22 = A[5] - q;
        double p = sqrt((b00 * b00 + b11 * b11 + b22 * b22 + norm * 2.0) / 6.0);
        double c00 = b11 * b22 - A[4] * A[4];
        double c01 = A[1] * b22 - A[4] * A[2];
        double c02 = A[1] * A[4] - b11 * A[2];
        double det = (b00 * c00 - A[1] * c01 + A[2] * c02) / (p * p * p);
        double half_det = det * 0.5;
        if (half_det < -1.0) half_det = -1.0;
        if (half_det > 1.0) half_det = 1.0;
        double angle = acos(half_det) / 3.0;
        const double two_thirds_pi = 2.09439510239319549;
        double beta2 = cos(angle) * 2.0;
        double beta0 = cos(angle + two_thirds_pi) * 2.0;
        double beta1 = -(beta0 + beta2);
        double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
        double v0[3], v1[3], v2[3];
        if (half_det >= 0) {
            eigvec0(A, e2, v2);
            if (e2 < e0 && e2 < e1) { out[0] = v2[0]; out[1] = v2[1]; out[2] = v2[2]; return; }
            eigvec1(A, v2, e1, v1);
            if (e1 < e0 && e1 < e2) { out[0] = v1[0]; out[1] = v1[1]; out[2] = v1[2]; return; }
            cross3(v1, v2, out);
        } else {
            eigvec0(A, e0, v0);
            if (e0 < e1 && e0 < e2) { out[0] = v0[0]; out[1] = v0[1]; out[2] = v0[2]; return; }
            eigvec1(A, v0, e1, v1);
            if (e1 < e0 && e1 < e2) { out[0] = v1[0]; out[1] = v1[1]; out[2] = v1[2]; return; }
            cross3(v0, v1, out);
        }
    } else {
        if (cov[0] < cov[3] && cov[0] < cov[5]) { out[0] = 1; out[1] = 0; out[2] = 0; }
        else if (cov[3] < cov[0] && cov[3] < cov[5]) { out[0] = 0; out[1] = 1; out[2] = 0; }
        else { out[0] = 0; out[1] = 0; out[2] = 1; }
    }
}

// thread per point; neighbour rows are read in list order so the 9 cumulants round identically
__global__ void __launch_bounds__(128) normals_kernel(const double *__restrict__ pts, int64_t n,
                                                      const int32_t *__restrict__ nbr_idx,
                                                      const int32_t *__restrict__ nbr_cnt, int max_nn,
                                                      double *__restrict__ normals) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int cnt = nbr_cnt[i];
    double cov[6];
    if (cnt >= 3) {
        double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        const int32_t *row = nbr_idx + i * max_nn;
        for (int k = 0; k < cnt; ++k) {
            const double *p = pts + 3 * (size_t)row[k];
            double x = p[0], y = p[1], z = p[2];
            c[0] += x; c[1] += y; c[2] += z;
            c[3] += x * x; c[4] += x * y; c[5] += x * z;
            c[6] += y * y; c[7] += y * z; c[8] += z * z;
        }
        double dc = (double)cnt;
        for (int k = 0; k < 9; ++k) c[k] /= dc;
        cov[0] = c[3] - c[0] * c[0]; cov[1] = c[4] - c[0] * c[1]; cov[2] = c[5] - c[0] * c[2];
        cov[3] = c[6] - c[1] * c[1]; cov[4] = c[7] - c[1] * c[2]; cov[5] = c[8] - c[2] * c[2];
    } else {
        cov[0] = 1; cov[1] = 0; cov[2] = 0; cov[3] = 1; cov[4] = 0; cov[5] = 1;
    }
    double nv[3];
    fast_eigen3x3(cov, nv);
    if (sqrt(nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2]) == 0.0) { nv[0] = 0; nv[1] = 0; nv[2] = 1; }
    normals[3 * i] = nv[0]; normals[3 * i + 1] = nv[1]; normals[3 * i + 2] = nv[2];
}

IBL_API int ibl_normals(const double *pts, int64_t n, const int32_t *nbr_idx, const int32_t *nbr_cnt, int32_t max_nn,
                        double *normals, void *stream) {
    IBL_ARG(n >= 0 && max_nn >= 1);
    if (n == 0) return IBL_OK;
    IBL_ARG(pts && nbr_idx && nbr_cnt && normals);
    IBL_PROF(IBL_PROF_NORMALS, ibl_stream(stream), (double)n * (24.0 + 28.0 * max_nn + 24.0));
    normals_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ibl_stream(stream)>>>(pts, n, nbr_idx, nbr_cnt, max_nn, normals);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
__device__ void pair_features(const double *p1, const double *n1, const double *p2, const double *n2, double *f) {
    double d[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    f[0] = f[1] = f[2] = 0.0;
    f[3] = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (f[3] == 0.0) return;
    double na[3] = {n1[0], n1[1], n1[2]}, nb[3] = {n2[0], n2[1], n2[2]};
    double angle1 = dot3(n1, d) / f[3];
    double angle2 = dot3(n2, d) / f[3];
    // Open3D compares acos(|angle1|) > acos(|angle2|).  acos falls with slope <= -1 on [0, 1], so when the arguments differ
    // by more than 1e-12 the two arc cosines differ by more than that too -- far above any rounding of either library's
    // acos -- and the comparison of the arguments decides; only a near-tie pays for the two f64 arc cosines (25 % of this
    // kernel's instructions).  (Arguments above 1 give NaN in both forms of the test: NaN > x is false, as is x < NaN.)
    const double c1 = fabs(angle1), c2 = fabs(angle2);
    const bool swap = fabs(c1 - c2) > 1e-12 && c1 <= 1.0 && c2 <= 1.0 ? (c1 < c2) : (acos(c1) > acos(c2));
    if (swap) {
        for (int k = 0; k < 3; ++k) { na[k] = n2[k]; nb[k] = n1[k]; d[k] *= -1.0; }
        f[2] = -angle2;
    } else {
        f[2] = angle1;
    }
    double v[3], w[3];
    cross3(d, na, v);
    double vn = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (vn == 0.0) { f[0] = f[1] = f[2] = f[3] = 0.0; return; }
    v[0] /= vn; v[1] /= vn; v[2] /= vn;
    cross3(na, v, w);
    f[1] = dot3(v, nb);
    f[0] = atan2(dot3(w, nb), dot3(na, nb));
}

__device__ __forceinline__ int clampbin(double x) {
    int h = (int)floor(x);
    if (h < 0) h = 0;
    if (h >= 11) h = 10;
    return h;
}

#define IBL_PI 3.14159265358979323846

// warp per point: lanes take neighbours, bins counted with shared-memory integer atomics, then
// each bin is rebuilt by repeated addition of inc (== the reference's += sequence, bit-exact)
__global__ void __launch_bounds__(128) spfh_kernel(const double *__restrict__ pts, const double *__restrict__ nrm,
                                                   int64_t n, const int32_t *__restrict__ nbr_idx,
                                                   const int32_t *__restrict__ nbr_cnt, int max_nn,
                                                   double *__restrict__ spfh) {
    __shared__ int bins[4][36];
    int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < n; i += nwarps) {
        int cnt = nbr_cnt[i];
        for (int t = lane; t < 36; t += 32) bins[wib][t] = 0;
        __syncwarp();
        if (cnt > 1) {
            double p1[3] = {pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]};
            double n1[3] = {nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]};
            for (int k = 1 + lane; k < cnt; k += 32) {
                size_t j = (size_t)nbr_idx[i * max_nn + k];
                double p2[3] = {pts[3 * j], pts[3 * j + 1], pts[3 * j + 2]};
                double n2[3] = {nrm[3 * j], nrm[3 * j + 1], nrm[3 * j + 2]};
                double f[4];
                pair_features(p1, n1, p2, n2, f);
                int h0 = clampbin(11 * (f[0] + IBL_PI) / (2.0 * IBL_PI));
                int h1 = clampbin(11 * (f[1] + 1.0) * 0.5);
                int h2 = clampbin(11 * (f[2] + 1.0) * 0.5);
                atomicAdd(&bins[wib][h0], 1);
                atomicAdd(&bins[wib][11 + h1], 1);
                atomicAdd(&bins[wib][22 + h2], 1);
            }
        }
        __syncwarp();
        double inc = cnt > 1 ? 100.0 / (double)(cnt - 1) : 0.0;
        for (int t = lane; t < 33; t += 32) {
            int c = bins[wib][t];
            double v = 0.0;
            for (int r = 0; r < c; ++r) v = __dadd_rn(v, inc);
            spfh[i * 33 + t] = v;
        }
        __syncwarp();
    }
}

IBL_API int ibl_spfh(const double *pts, const double *nrm, int64_t n, const int32_t *nbr_idx, const int32_t *nbr_cnt,
                     int32_t max_nn, double *spfh, void *stream) {
    IBL_ARG(n >= 0 && max_nn >= 1);
    if (n == 0) return IBL_OK;
    IBL_ARG(pts && nrm && nbr_idx && nbr_cnt && spfh);
    int64_t blocks = (n + 3) / 4;
    if (blocks > IBL_SMS * 32) blocks = IBL_SMS * 32;
    IBL_PROF(IBL_PROF_SPFH, ibl_stream(stream), (double)n * (48.0 + 52.0 * max_nn + 264.0));
    spfh_kernel<<<(unsigned)blocks, 128, 0, ibl_stream(stream)>>>(pts, nrm, n, nbr_idx, nbr_cnt, max_nn, spfh);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// warp per point: lane j owns bin j of every neighbour's SPFH; the three block sums are accumulated by lanes 0..2 in
// the reference's (k outer, j inner) order from values staged in shared memory.  Bin 32 does not fit the 32 lanes: its
// quotients are computed for ALL neighbours of the point in a lane-parallel pre-pass (lane = neighbour), so the serial
// per-neighbour loop issues one f64 division per lane instead of two (a divergent `if (lane == 0)` division costs the
// whole warp its issue slots).  Every quotient and every sum is the same operation in the same order as before.
#define FPFH_MAX_NN 224
// s / dist, correctly rounded, for s >= 0.  SPFH histograms are sparse, and a ZERO numerator sends __ddiv_rn down its
// slow path (the fast path requires a normal numerator); with one zero lane per warp every iteration paid for the ~100
// instruction special-case routine (ncu: the slow-path call executed once per neighbour, 47 % of the kernel's
// instructions).  +0 / dist is +0 for dist > 0, so zero lanes divide 1.0 instead and select 0 afterwards.
__device__ __forceinline__ double fpfh_quot(double s, double dist) {
    const bool zero = (s == 0.0);
    double num = zero ? 1.0 : s;
#if defined(__CUDA_ARCH__)
    asm volatile("" : "+d"(num));        // opaque: otherwise the compiler folds the two selects away and divides s itself
#endif
    const double q = __ddiv_rn(num, dist);
    return zero ? 0.0 : q;
}
__global__ void __launch_bounds__(128) fpfh_kernel(int64_t n, const int32_t *__restrict__ nbr_idx,
                                                   const double *__restrict__ nbr_d2,
                                                   const int32_t *__restrict__ nbr_cnt, int max_nn,
                                                   const double *__restrict__ spfh, double *__restrict__ fpfh) {
    __shared__ double vals[4][32];
    __shared__ double v32s[4][FPFH_MAX_NN];
    int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < n; i += nwarps) {
        int cnt = nbr_cnt[i];
        for (int k = 1 + lane; k < cnt; k += 32) {
            double dist = nbr_d2[i * max_nn + k];
            size_t q = (size_t)nbr_idx[i * max_nn + k];
            v32s[wib][k] = (dist == 0.0) ? 0.0 : fpfh_quot(spfh[q * 33 + 32], dist);
        }
        __syncwarp();
        double acc = 0.0, acc32 = 0.0, bsum = 0.0;
        // Two-deep software pipeline over the neighbours: the distance / index of neighbour k+2 and the SPFH row of neighbour
        // k+1 are in flight while neighbour k is processed (ncu: 45 % of the stall samples sat on the first use of these
        // three dependent loads).  Slots past the end read row 0, which is valid memory and never used.
        const int64_t nb0 = i * (int64_t)max_nn;
        double d_n1 = 1 < cnt ? nbr_d2[nb0 + 1] : 0.0, d_n2 = 2 < cnt ? nbr_d2[nb0 + 2] : 0.0;
        int q_n1 = 1 < cnt ? nbr_idx[nb0 + 1] : 0, q_n2 = 2 < cnt ? nbr_idx[nb0 + 2] : 0;
        double s_n1 = spfh[(size_t)q_n1 * 33 + lane];
        for (int k = 1; k < cnt; ++k) {
            const double dist = d_n1, sk = s_n1;
            d_n1 = d_n2; q_n1 = q_n2;
            s_n1 = spfh[(size_t)q_n1 * 33 + lane];
            d_n2 = k + 2 < cnt ? nbr_d2[nb0 + k + 2] : 0.0;           // after the shift: d_n1 / q_n1 / s_n1 = neighbour k+1, d_n2 / q_n2 = k+2
            q_n2 = k + 2 < cnt ? nbr_idx[nb0 + k + 2] : 0;
            if (dist == 0.0) continue;           // uniform across the warp
            double v = fpfh_quot(sk, dist);
            acc = __dadd_rn(acc, v);
            vals[wib][lane] = v;
            __syncwarp();
            if (lane < 3) {
#pragma unroll
                for (int t = 0; t < 10; ++t) bsum = __dadd_rn(bsum, vals[wib][lane * 11 + t]);
                double last = v32s[wib][k];
                if (lane < 2) last = vals[wib][lane * 11 + 10];
                bsum = __dadd_rn(bsum, last);
                if (lane == 2) acc32 = __dadd_rn(acc32, last);
            }
            __syncwarp();
        }
        if (lane < 3) bsum = (bsum != 0.0) ? __ddiv_rn(100.0, bsum) : bsum;
        double s0 = __shfl_sync(0xffffffffu, bsum, 0), s1 = __shfl_sync(0xffffffffu, bsum, 1),
               s2 = __shfl_sync(0xffffffffu, bsum, 2);
        if (cnt > 1) {
            double sc = lane < 11 ? s0 : (lane < 22 ? s1 : s2);
            fpfh[i * 33 + lane] = __dadd_rn(__dmul_rn(acc, sc), spfh[i * 33 + lane]);
            if (lane == 2) fpfh[i * 33 + 32] = __dadd_rn(__dmul_rn(acc32, s2), spfh[i * 33 + 32]);
        } else {
            fpfh[i * 33 + lane] = 0.0;
            if (lane == 0) fpfh[i * 33 + 32] = 0.0;
        }
        __syncwarp();                            // v32s is rewritten by the next point's pre-pass
    }
}

IBL_API int ibl_fpfh(int64_t n, const int32_t *nbr_idx, const double *nbr_d2, const int32_t *nbr_cnt, int32_t max_nn,
                     const double *spfh, double *fpfh, void *stream) {
    IBL_ARG(n >= 0 && max_nn >= 1 && max_nn <= FPFH_MAX_NN);
    if (n == 0) return IBL_OK;
    IBL_ARG(nbr_idx && nbr_d2 && nbr_cnt && spfh && fpfh);
    int64_t blocks = (n + 3) / 4;
    if (blocks > IBL_SMS * 32) blocks = IBL_SMS * 32;
    IBL_PROF(IBL_PROF_FPFH, ibl_stream(stream), (double)n * (max_nn * (264.0 + 12.0) + 528.0));
    fpfh_kernel<<<(unsigned)blocks, 128, 0, ibl_stream(stream)>>>(n, nbr_idx, nbr_d2, nbr_cnt, max_nn, spfh, fpfh);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}
