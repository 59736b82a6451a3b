// rigid.cuh -- rigid estimation shared by RANSAC and ICP (host + device).
// Eigen::umeyama(with_scaling=false) is restated as a Jacobi-Kabsch that uses only + - * / sqrt
// (SURVEY.md A.7/A.9): eigen-decompose sigma^T sigma by cyclic Jacobi, U from sigma*V by
// Gram-Schmidt, third axes by cross products (this is the det-corrected Kabsch rotation).
// Built with -fmad=false (device) and -ffp-contract=off (host) so both sides round identically.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define IBL_HD __host__ __device__ __forceinline__
#else
#define IBL_HD inline
#endif

IBL_HD void ibl_cross(const double *a, const double *b, double *o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
IBL_HD double ibl_dot(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

IBL_HD void ibl_jacobi3(double m[3][3], double v[3][3]) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 32; ++sweep) {
        if (m[0][1] == 0.0 && m[0][2] == 0.0 && m[1][2] == 0.0) break;
        for (int e = 0; e < 3; ++e) {
            int p = (e == 2) ? 1 : 0, q = (e == 0) ? 1 : 2;
            double apq = m[p][q];
            if (apq == 0.0) continue;
            double dsum = fabs(m[p][p]) + fabs(m[q][q]);
            if (dsum + fabs(apq) == dsum) { m[p][q] = 0.0; m[q][p] = 0.0; continue; }
            double theta = (m[q][q] - m[p][p]) / (2.0 * apq);
            double t = 1.0 / (fabs(theta) + sqrt(theta * theta + 1.0));
            if (theta < 0.0) t = -t;
            double c = 1.0 / sqrt(t * t + 1.0);
            double s = t * c;
            int r = 3 - p - q;
            double app = m[p][p], aqq = m[q][q];
            double arp = m[r][p], arq = m[r][q];
            m[p][p] = app - t * apq;
            m[q][q] = aqq + t * apq;
            m[p][q] = 0.0; m[q][p] = 0.0;
            double nrp = c * arp - s * arq;
            double nrq = s * arp + c * arq;
            m[r][p] = nrp; m[p][r] = nrp;
            m[r][q] = nrq; m[q][r] = nrq;
            for (int k = 0; k < 3; ++k) {
                double vkp = v[k][p], vkq = v[k][q];
                v[k][p] = c * vkp - s * vkq;
                v[k][q] = s * vkp + c * vkq;
            }
        }
    }
}

// sigma = E[(t-mt)(s-ms)^T] (3x3), ms / mt means  ->  row-major 4x4 T
IBL_HD void ibl_kabsch_finish(const double sg[3][3], const double *ms, const double *mt, double *T) {
    double m[3][3], v[3][3];
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) m[a][b] = sg[0][a] * sg[0][b] + sg[1][a] * sg[1][b] + sg[2][a] * sg[2][b];
    ibl_jacobi3(m, v);
    double lam[3] = {m[0][0], m[1][1], m[2][2]};
    int i0 = 0;
    if (lam[1] > lam[i0]) i0 = 1;
    if (lam[2] > lam[i0]) i0 = 2;
    int i1 = -1;
    for (int a = 0; a < 3; ++a) {
        if (a == i0) continue;
        if (i1 < 0 || lam[a] > lam[i1]) i1 = a;
    }
    double v0[3] = {v[0][i0], v[1][i0], v[2][i0]};
    double v1[3] = {v[0][i1], v[1][i1], v[2][i1]};
    double v2[3];
    ibl_cross(v0, v1, v2);
    double R[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    double a0[3], a1[3], u0[3], u1[3], u2[3];
    for (int a = 0; a < 3; ++a) a0[a] = sg[a][0] * v0[0] + sg[a][1] * v0[1] + sg[a][2] * v0[2];
    double n0 = sqrt(ibl_dot(a0, a0));
    if (n0 > 0.0) {
        for (int a = 0; a < 3; ++a) u0[a] = a0[a] / n0;
        for (int a = 0; a < 3; ++a) a1[a] = sg[a][0] * v1[0] + sg[a][1] * v1[1] + sg[a][2] * v1[2];
        double pr = ibl_dot(u0, a1);
        for (int a = 0; a < 3; ++a) a1[a] = a1[a] - pr * u0[a];
        double n1 = sqrt(ibl_dot(a1, a1));
        if (n1 > 0.0) {
            for (int a = 0; a < 3; ++a) u1[a] = a1[a] / n1;
        } else if (fabs(u0[0]) > fabs(u0[1])) {
            double inv = 1.0 / sqrt(u0[0] * u0[0] + u0[2] * u0[2]);
            u1[0] = -u0[2] * inv; u1[1] = 0.0; u1[2] = u0[0] * inv;
        } else {
            double inv = 1.0 / sqrt(u0[1] * u0[1] + u0[2] * u0[2]);
            u1[0] = 0.0; u1[1] = u0[2] * inv; u1[2] = -u0[1] * inv;
        }
        ibl_cross(u0, u1, u2);
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) R[a][b] = (u0[a] * v0[b] + u1[a] * v1[b]) + u2[a] * v2[b];
    }
    for (int a = 0; a < 3; ++a) {
        double rs = (R[a][0] * ms[0] + R[a][1] * ms[1]) + R[a][2] * ms[2];
        T[4 * a + 0] = R[a][0]; T[4 * a + 1] = R[a][1]; T[4 * a + 2] = R[a][2];
        T[4 * a + 3] = mt[a] - rs;
    }
    T[12] = 0; T[13] = 0; T[14] = 0; T[15] = 1;
}

// three gathered pairs (s, t row-major [3][3])
IBL_HD void ibl_kabsch3(const double *s, const double *t, double *T) {
    double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
    for (int i = 0; i < 3; ++i)
        for (int a = 0; a < 3; ++a) { ms[a] += s[3 * i + a]; mt[a] += t[3 * i + a]; }
    for (int a = 0; a < 3; ++a) { ms[a] /= 3.0; mt[a] /= 3.0; }
    double sg[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < 3; ++i) {
        double ds[3] = {s[3 * i] - ms[0], s[3 * i + 1] - ms[1], s[3 * i + 2] - ms[2]};
        double dt[3] = {t[3 * i] - mt[0], t[3 * i + 1] - mt[1], t[3 * i + 2] - mt[2]};
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b) sg[a][b] += dt[a] * ds[b];
    }
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) sg[a][b] /= 3.0;
    ibl_kabsch_finish(sg, ms, mt, T);
}

// unpivoted LDL^T solve of a symmetric N x N system; 0 on zero / non-finite pivot
template <int N>
IBL_HD int ibl_ldlt_solve(const double *A, const double *b, double *x) {
    double L[N][N], D[N], y[N];
    for (int j = 0; j < N; ++j) {
        double d = A[j * N + j];
        for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k] * D[k];
        if (!(d != 0.0) || !isfinite(d)) return 0;
        D[j] = d;
        for (int i = j + 1; i < N; ++i) {
            double s = A[i * N + j];
            for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k] * D[k];
            L[i][j] = s / d;
        }
    }
    for (int i = 0; i < N; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= L[i][k] * y[k]; y[i] = s; }
    for (int i = 0; i < N; ++i) y[i] /= D[i];
    for (int i = N - 1; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < N; ++k) s -= L[k][i] * x[k]; x[i] = s; }
    for (int i = 0; i < N; ++i) if (!isfinite(x[i])) return 0;
    return 1;
}

// counter-based hypothesis sampler (splitmix64 finaliser + multiply-shift range reduction):
// sample j (0..2) of hypothesis h over C correspondences.  Restated in oracle/sampler.py.
IBL_HD uint32_t ibl_sample(uint64_t seed, uint64_t h, uint32_t j, uint32_t C) {
    uint64_t z = seed + (3ull * h + j + 1ull) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (uint32_t)(((z >> 32) * (uint64_t)C) >> 32);
}
