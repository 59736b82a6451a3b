/*
 * oracle/o3d_reg.c -- TEST INFRASTRUCTURE ONLY (see o3d_core.c header; PARITY UNPINNED).
 *
 * Registration-side arithmetic of Open3D 0.17.0 as used by the reference
 * (utils/fpfh_register.py:110-141): 3-point / N-point rigid estimation (Eigen::umeyama without
 * scaling), the RANSAC correspondence checkers, coloured-ICP gradient initialisation and
 * Gauss-Newton normal equations.  SURVEY.md Appendix A.6-A.9.
 *
 * Eigen's JacobiSVD is third-party and not restated instruction by instruction; the rigid
 * estimate is DEFINED here as: eigen-decompose sigma^T sigma by cyclic Jacobi, U from sigma*V
 * (Gram-Schmidt), third axes by cross products (== Kabsch with det correction).  It uses only
 * + - * / sqrt so a device implementation with FMA contraction disabled is bit-identical.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

static void cross3(const double *a, const double *b, double *o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
static double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

/* cyclic Jacobi on symmetric 3x3 m (full storage); v receives eigenvectors as columns */
static void jacobi3(double m[3][3], double v[3][3]) {
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) v[i][j] = (i == j) ? 1.0 : 0.0;
    static const int PQ[3][2] = {{0, 1}, {0, 2}, {1, 2}};
    for (int sweep = 0; sweep < 32; ++sweep) {
        if (m[0][1] == 0.0 && m[0][2] == 0.0 && m[1][2] == 0.0) break;
        for (int e = 0; e < 3; ++e) {
            int p = PQ[e][0], q = PQ[e][1];
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

/* src/tgt: gathered point pairs [k][3]; T: row-major 4x4 */
static void kabsch(const double *src, const double *tgt, int k, double *T) {
    double ms[3] = {0, 0, 0}, mt[3] = {0, 0, 0};
    for (int i = 0; i < k; ++i) for (int a = 0; a < 3; ++a) { ms[a] += src[3 * i + a]; mt[a] += tgt[3 * i + a]; }
    for (int a = 0; a < 3; ++a) { ms[a] /= (double)k; mt[a] /= (double)k; }
    double sg[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int i = 0; i < k; ++i) {
        double ds[3] = {src[3 * i] - ms[0], src[3 * i + 1] - ms[1], src[3 * i + 2] - ms[2]};
        double dt[3] = {tgt[3 * i] - mt[0], tgt[3 * i + 1] - mt[1], tgt[3 * i + 2] - mt[2]};
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) sg[a][b] += dt[a] * ds[b];
    }
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) sg[a][b] /= (double)k;
    double m[3][3], v[3][3];
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b)
        m[a][b] = sg[0][a] * sg[0][b] + sg[1][a] * sg[1][b] + sg[2][a] * sg[2][b];
    jacobi3(m, v);
    double lam[3] = {m[0][0], m[1][1], m[2][2]};
    int i0 = 0;
    if (lam[1] > lam[i0]) i0 = 1;
    if (lam[2] > lam[i0]) i0 = 2;
    int i1 = -1;
    for (int a = 0; a < 3; ++a) { if (a == i0) continue; if (i1 < 0 || lam[a] > lam[i1]) i1 = a; }
    double v0[3] = {v[0][i0], v[1][i0], v[2][i0]};
    double v1[3] = {v[0][i1], v[1][i1], v[2][i1]};
    double v2[3];
    cross3(v0, v1, v2);
    double R[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    double a0[3], a1[3], u0[3], u1[3], u2[3];
    for (int a = 0; a < 3; ++a) a0[a] = sg[a][0] * v0[0] + sg[a][1] * v0[1] + sg[a][2] * v0[2];
    double n0 = sqrt(dot3(a0, a0));
    if (n0 > 0.0) {
        for (int a = 0; a < 3; ++a) u0[a] = a0[a] / n0;
        for (int a = 0; a < 3; ++a) a1[a] = sg[a][0] * v1[0] + sg[a][1] * v1[1] + sg[a][2] * v1[2];
        double pr = dot3(u0, a1);
        for (int a = 0; a < 3; ++a) a1[a] = a1[a] - pr * u0[a];
        double n1 = sqrt(dot3(a1, a1));
        if (n1 > 0.0) {
            for (int a = 0; a < 3; ++a) u1[a] = a1[a] / n1;
        } else {
            /* collinear sample: any unit vector orthogonal to u0 */
            if (fabs(u0[0]) > fabs(u0[1])) {
                double inv = 1.0 / sqrt(u0[0] * u0[0] + u0[2] * u0[2]);
                u1[0] = -u0[2] * inv; u1[1] = 0.0; u1[2] = u0[0] * inv;
            } else {
                double inv = 1.0 / sqrt(u0[1] * u0[1] + u0[2] * u0[2]);
                u1[0] = 0.0; u1[1] = u0[2] * inv; u1[2] = -u0[1] * inv;
            }
        }
        cross3(u0, u1, u2);
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b)
            R[a][b] = (u0[a] * v0[b] + u1[a] * v1[b]) + u2[a] * v2[b];
    }
    for (int a = 0; a < 3; ++a) {
        double rs = (R[a][0] * ms[0] + R[a][1] * ms[1]) + R[a][2] * ms[2];
        T[4 * a + 0] = R[a][0]; T[4 * a + 1] = R[a][1]; T[4 * a + 2] = R[a][2];
        T[4 * a + 3] = mt[a] - rs;
    }
    T[12] = 0; T[13] = 0; T[14] = 0; T[15] = 1;
}

ORC_API void orc_kabsch(const double *src, const double *tgt, int k, double *T16) { kabsch(src, tgt, k, T16); }

/* rigid estimate from a correspondence list (TransformationEstimationPointToPoint, A.9) */
ORC_API void orc_kabsch_corres(const double *src, const double *tgt, const int32_t *corres, int c, double *T16) {
    if (c == 0) {
        for (int i = 0; i < 16; ++i) T16[i] = (i % 5 == 0) ? 1.0 : 0.0;
        return;
    }
    double *s = (double *)malloc(sizeof(double) * 3 * (size_t)c);
    double *t = (double *)malloc(sizeof(double) * 3 * (size_t)c);
    for (int i = 0; i < c; ++i) for (int a = 0; a < 3; ++a) {
        s[3 * (size_t)i + a] = src[3 * (size_t)corres[2 * i] + a];
        t[3 * (size_t)i + a] = tgt[3 * (size_t)corres[2 * i + 1] + a];
    }
    kabsch(s, t, c, T16);
    free(s); free(t);
}

static void xform(const double *T, const double *p, double *o) {
    for (int a = 0; a < 3; ++a) o[a] = ((T[4 * a] * p[0] + T[4 * a + 1] * p[1]) + T[4 * a + 2] * p[2]) + T[4 * a + 3];
}

ORC_API void orc_transform_points(const double *T, const double *pts, int n, double *out) {
#pragma omp parallel for
    for (int i = 0; i < n; ++i) xform(T, pts + 3 * (size_t)i, out + 3 * (size_t)i);
}

/* A.7 hypothesis generation + cheap checkers.  samples[H][3] index into corres[C][2].
 * pass[h] = 1 iff EdgeLength(edge_thr) and Distance(dist_thr) both pass. T_out[h][16] always written. */
ORC_API void orc_ransac_check(const double *src, const double *tgt, const int32_t *corres,
                              const int32_t *samples, int64_t H, double edge_thr, double dist_thr,
                              uint8_t *pass, double *T_out /* may be NULL */) {
#pragma omp parallel for schedule(static, 4096)
    for (int64_t h = 0; h < H; ++h) {
        double s[9], t[9], T[16];
        for (int j = 0; j < 3; ++j) {
            int c = samples[3 * h + j];
            for (int a = 0; a < 3; ++a) {
                s[3 * j + a] = src[3 * (size_t)corres[2 * c] + a];
                t[3 * j + a] = tgt[3 * (size_t)corres[2 * c + 1] + a];
            }
        }
        kabsch(s, t, 3, T);
        if (T_out) memcpy(T_out + 16 * h, T, sizeof(T));
        int ok = 1;
        for (int i = 0; i < 3 && ok; ++i) for (int j = i + 1; j < 3; ++j) {
            double a[3] = {s[3 * i] - s[3 * j], s[3 * i + 1] - s[3 * j + 1], s[3 * i + 2] - s[3 * j + 2]};
            double b[3] = {t[3 * i] - t[3 * j], t[3 * i + 1] - t[3 * j + 1], t[3 * i + 2] - t[3 * j + 2]};
            double ds = sqrt(dot3(a, a)), dt = sqrt(dot3(b, b));
            if (ds < dt * edge_thr || dt < ds * edge_thr) { ok = 0; break; }
        }
        if (ok) for (int j = 0; j < 3; ++j) {
            double q[3];
            xform(T, s + 3 * j, q);
            double d[3] = {t[3 * j] - q[0], t[3 * j + 1] - q[1], t[3 * j + 2] - q[2]};
            if (sqrt(dot3(d, d)) > dist_thr) { ok = 0; break; }
        }
        pass[h] = (uint8_t)ok;
    }
}

/* EvaluateInlierCorrespondenceRatio numerator: #{c : |T s_c - t_c|^2 < max_d^2} (squaredNorm) */
ORC_API int32_t orc_corres_inliers(const double *T, const double *src, const double *tgt,
                                   const int32_t *corres, int c, double max_d) {
    int32_t cnt = 0;
    for (int i = 0; i < c; ++i) {
        double q[3];
        xform(T, src + 3 * (size_t)corres[2 * i], q);
        const double *t = tgt + 3 * (size_t)corres[2 * i + 1];
        double d[3] = {q[0] - t[0], q[1] - t[1], q[2] - t[2]};
        if (dot3(d, d) < max_d * max_d) ++cnt;
    }
    return cnt;
}

/* unpivoted LDL^T solve of a symmetric n x n system (n<=6). returns 0 on zero/NaN pivot. */
static int ldlt_solve(int n, const double *A, const double *b, double *x) {
    double L[6][6], D[6], y[6];
    for (int j = 0; j < n; ++j) {
        double d = A[j * n + j];
        for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k] * D[k];
        if (!(d != 0.0) || !isfinite(d)) return 0;
        D[j] = d;
        for (int i = j + 1; i < n; ++i) {
            double s = A[i * n + j];
            for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k] * D[k];
            L[i][j] = s / d;
        }
    }
    for (int i = 0; i < n; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= L[i][k] * y[k]; y[i] = s; }
    for (int i = 0; i < n; ++i) y[i] /= D[i];
    for (int i = n - 1; i >= 0; --i) { double s = y[i]; for (int k = i + 1; k < n; ++k) s -= L[k][i] * x[k]; x[i] = s; }
    for (int i = 0; i < n; ++i) if (!isfinite(x[i])) return 0;
    return 1;
}

ORC_API int orc_ldlt_solve(int n, const double *A, const double *b, double *x) { return ldlt_solve(n, A, b, x); }

/* A.8 InitializePointCloudForColoredICP: per target point colour gradient */
ORC_API void orc_color_gradient(const double *pts, const double *nrm, const double *col, int n,
                                const int32_t *nbr_idx, const int32_t *nbr_cnt, int max_nn, double *grad) {
#pragma omp parallel for schedule(dynamic, 256)
    for (int k = 0; k < n; ++k) {
        double *g = grad + 3 * (size_t)k;
        g[0] = g[1] = g[2] = 0.0;
        int nn = nbr_cnt[k];
        if (nn < 4) continue;
        const double *vt = pts + 3 * (size_t)k, *nt = nrm + 3 * (size_t)k;
        double it = (col[3 * (size_t)k] + col[3 * (size_t)k + 1] + col[3 * (size_t)k + 2]) / 3.0;
        double AtA[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, Atb[3] = {0, 0, 0};
        for (int i = 1; i < nn; ++i) {
            int adj = nbr_idx[(size_t)k * max_nn + i];
            const double *va = pts + 3 * (size_t)adj;
            double d[3] = {va[0] - vt[0], va[1] - vt[1], va[2] - vt[2]};
            double pr = dot3(d, nt);
            double row[3] = {(va[0] - pr * nt[0]) - vt[0], (va[1] - pr * nt[1]) - vt[1], (va[2] - pr * nt[2]) - vt[2]};
            double ia = (col[3 * (size_t)adj] + col[3 * (size_t)adj + 1] + col[3 * (size_t)adj + 2]) / 3.0;
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
        if (ldlt_solve(3, AtA, Atb, x)) { g[0] = x[0]; g[1] = x[1]; g[2] = x[2]; }
    }
}

/* A.8 coloured-ICP normal equations over a correspondence set (source already transformed).
 * out27: JTJ upper triangle row-major (21) then JTr (6). Sequential accumulation. */
ORC_API void orc_colored_jtj(const double *src, const double *scol, const double *tgt, const double *tnrm,
                             const double *tcol, const double *tgrad, const int32_t *corres, int c,
                             double lambda_geometric, double *JTJ36, double *JTr6) {
    double sg = sqrt(lambda_geometric), sp = sqrt(1.0 - lambda_geometric);
    memset(JTJ36, 0, 36 * sizeof(double));
    memset(JTr6, 0, 6 * sizeof(double));
    for (int i = 0; i < c; ++i) {
        int cs = corres[2 * i], ct = corres[2 * i + 1];
        const double *vs = src + 3 * (size_t)cs, *vt = tgt + 3 * (size_t)ct, *nt = tnrm + 3 * (size_t)ct;
        double J[2][6], r[2], cr[3];
        cross3(vs, nt, cr);
        for (int a = 0; a < 3; ++a) { J[0][a] = sg * cr[a]; J[0][3 + a] = sg * nt[a]; }
        double d[3] = {vs[0] - vt[0], vs[1] - vt[1], vs[2] - vt[2]};
        double dn = dot3(d, nt);
        r[0] = sg * dn;
        double vp[3] = {vs[0] - dn * nt[0], vs[1] - dn * nt[1], vs[2] - dn * nt[2]};
        double is = (scol[3 * (size_t)cs] + scol[3 * (size_t)cs + 1] + scol[3 * (size_t)cs + 2]) / 3.0;
        double it = (tcol[3 * (size_t)ct] + tcol[3 * (size_t)ct + 1] + tcol[3 * (size_t)ct + 2]) / 3.0;
        const double *dit = tgrad + 3 * (size_t)ct;
        double dv[3] = {vp[0] - vt[0], vp[1] - vt[1], vp[2] - vt[2]};
        double is0 = dot3(dit, dv) + it;
        /* ditM = -dit^T (I - nt nt^T) */
        double dm[3];
        for (int b = 0; b < 3; ++b) {
            double s = 0.0;
            for (int a = 0; a < 3; ++a) s += dit[a] * ((a == b ? 1.0 : 0.0) - nt[a] * nt[b]);
            dm[b] = -s;
        }
        cross3(vs, dm, cr);
        for (int a = 0; a < 3; ++a) { J[1][a] = sp * cr[a]; J[1][3 + a] = sp * dm[a]; }
        r[1] = sp * (is - is0);
        for (int e = 0; e < 2; ++e) for (int a = 0; a < 6; ++a) {
            for (int b = 0; b < 6; ++b) JTJ36[6 * a + b] += J[e][a] * J[e][b];
            JTr6[a] += J[e][a] * r[e];
        }
    }
}

/* SolveJacobianSystemAndObtainExtrinsicMatrix + TransformVector6dToMatrix4d */
ORC_API int orc_solve_update(const double *JTJ36, const double *JTr6, double *T16) {
    double nb[6], x[6];
    for (int i = 0; i < 6; ++i) nb[i] = -JTr6[i];
    for (int i = 0; i < 16; ++i) T16[i] = (i % 5 == 0) ? 1.0 : 0.0;
    if (!ldlt_solve(6, JTJ36, nb, x)) return 0;
    double cx = cos(x[0]), sx = sin(x[0]), cy = cos(x[1]), sy = sin(x[1]), cz = cos(x[2]), sz = sin(x[2]);
    /* R = Rz * Ry * Rx */
    T16[0] = cz * cy; T16[1] = cz * sy * sx - sz * cx; T16[2] = cz * sy * cx + sz * sx; T16[3] = x[3];
    T16[4] = sz * cy; T16[5] = sz * sy * sx + cz * cx; T16[6] = sz * sy * cx - cz * sx; T16[7] = x[4];
    T16[8] = -sy;     T16[9] = cy * sx;                T16[10] = cy * cx;               T16[11] = x[5];
    return 1;
}

/* ------------------------------------------------------------------------------------------------
 * A.6 scoring of MANY hypotheses at once (GetRegistrationResultAndCorrespondences per hypothesis).  Open3D 0.17 runs the
 * RANSAC iterations under `#pragma omp parallel` (Registration.cpp); this is the same parallelism: one hypothesis per
 * thread, every source point looked up in a uniform grid over the target (cell = max_d: the 27 surrounding cells hold
 * every point within max_d).  Per hypothesis: the number of source points with a target strictly within max_d and the
 * sum of their squared nearest distances, added in numpy's pairwise order (o3d_core.py `score` uses np.sum), so the
 * figures are bit-identical to the one-at-a-time path. */
static double pairwise_sum(const double *a, int64_t n) {
    if (n < 8) {
        double res = 0.;
        for (int64_t i = 0; i < n; ++i) res += a[i];
        return res;
    } else if (n <= 128) {
        double r[8];
        for (int k = 0; k < 8; ++k) r[k] = a[k];
        int64_t i;
        for (i = 8; i < n - (n % 8); i += 8)
            for (int k = 0; k < 8; ++k) r[k] += a[i + k];
        double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
        for (; i < n; ++i) res += a[i];
        return res;
    } else {
        int64_t n2 = n / 2;
        n2 -= n2 % 8;
        return pairwise_sum(a, n2) + pairwise_sum(a + n2, n - n2);
    }
}

typedef struct {
    int64_t n_tgt, dim[3], ncell;
    double lo[3], cell;
    const double *tgt;
    int32_t *start, *order;
} orc_grid_t;

/* uniform grid over the target (cells of about max_d / 4) for orc_score_many; the caller keeps `tgt` alive */
ORC_API void *orc_grid_build(const double *tgt, int64_t n_tgt, double max_d) {
    orc_grid_t *g = (orc_grid_t *)calloc(1, sizeof(orc_grid_t));
    g->n_tgt = n_tgt; g->tgt = tgt;
    double hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int a = 0; a < 3; ++a) g->lo[a] = INFINITY;
    for (int64_t i = 0; i < n_tgt; ++i)
        for (int a = 0; a < 3; ++a) { double v = tgt[3 * i + a]; if (v < g->lo[a]) g->lo[a] = v; if (v > hi[a]) hi[a] = v; }
    g->cell = max_d / 4.0;
    if (n_tgt == 0) { g->dim[0] = g->dim[1] = g->dim[2] = 1; g->lo[0] = g->lo[1] = g->lo[2] = 0.0; }
    else for (;;) {
        for (int a = 0; a < 3; ++a) g->dim[a] = (int64_t)floor((hi[a] - g->lo[a]) / g->cell) + 1;
        if ((double)g->dim[0] * (double)g->dim[1] * (double)g->dim[2] <= 6.4e7) break;
        g->cell *= 1.5;
    }
    g->ncell = g->dim[0] * g->dim[1] * g->dim[2];
    g->start = (int32_t *)calloc((size_t)g->ncell + 1, sizeof(int32_t));
    g->order = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n_tgt ? n_tgt : 1));
    int32_t *cidx = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n_tgt ? n_tgt : 1));
    for (int64_t i = 0; i < n_tgt; ++i) {
        int64_t c[3];
        for (int a = 0; a < 3; ++a) {
            c[a] = (int64_t)floor((tgt[3 * i + a] - g->lo[a]) / g->cell);
            if (c[a] >= g->dim[a]) c[a] = g->dim[a] - 1;
            if (c[a] < 0) c[a] = 0;
        }
        cidx[i] = (int32_t)((c[2] * g->dim[1] + c[1]) * g->dim[0] + c[0]);
        g->start[cidx[i] + 1]++;
    }
    for (int64_t c = 0; c < g->ncell; ++c) g->start[c + 1] += g->start[c];
    int32_t *fill = (int32_t *)malloc(sizeof(int32_t) * (size_t)g->ncell);
    memcpy(fill, g->start, sizeof(int32_t) * (size_t)g->ncell);
    for (int64_t i = 0; i < n_tgt; ++i) g->order[fill[cidx[i]]++] = (int32_t)i;      /* ascending index inside a cell */
    free(fill); free(cidx);
    return g;
}

ORC_API void orc_grid_free(void *h) {
    orc_grid_t *g = (orc_grid_t *)h;
    if (!g) return;
    free(g->start); free(g->order); free(g);
}

/* squared distance to the nearest target strictly within max_d of q (lowest index on exact ties), or -1.  Cells are
 * searched ring by ring: a target in Chebyshev ring s lies farther than s - 1 cells away, so the search ends before ring s
 * once the best distance is below (s - 1) * cell -- a good hypothesis finds its neighbour in the first rings, a bad one
 * only walks empty cells. */
static double grid_nn_d2(const orc_grid_t *g, const double *q, double max_d, double md2) {
    const double cell = g->cell;
    const int reach = (int)ceil(max_d / cell) + 1;     /* +1: a point on a cell face may round to either side */
    int64_t c[3];
    for (int a = 0; a < 3; ++a) {
        double f = floor((q[a] - g->lo[a]) / cell);
        if (f < -(double)reach - 1 || f > (double)(g->dim[a] + reach)) return -1.0;
        c[a] = (int64_t)f;
    }
    const int64_t *dim = g->dim;
    double best = INFINITY;
    int32_t bi = -1;
    for (int s = 0; s <= reach; ++s) {
        if (s >= 2) { double gg = (double)(s - 1) * cell * (1.0 - 1e-9); if (best < gg * gg) break; }
        for (int64_t z = c[2] - s; z <= c[2] + s; ++z) {
            if (z < 0 || z >= dim[2]) continue;
            const int zedge = (z == c[2] - s) || (z == c[2] + s);
            for (int64_t y = c[1] - s; y <= c[1] + s; ++y) {
                if (y < 0 || y >= dim[1]) continue;
                const int edge = zedge || (y == c[1] - s) || (y == c[1] + s);
                const int64_t row = (z * dim[1] + y) * dim[0];
                for (int part = 0; part < (edge || s == 0 ? 1 : 2); ++part) {       /* full x-run on the shell's faces */
                    int64_t x0, x1;
                    if (edge || s == 0) { x0 = c[0] - s; x1 = c[0] + s; }
                    else if (part == 0) { x0 = x1 = c[0] - s; }
                    else { x0 = x1 = c[0] + s; }
                    if (x0 < 0) x0 = 0;
                    if (x1 >= dim[0]) x1 = dim[0] - 1;
                    if (x0 > x1) continue;
                    for (int32_t k = g->start[row + x0]; k < g->start[row + x1 + 1]; ++k) {
                        const double *t = g->tgt + 3 * (size_t)g->order[k];
                        double dx = t[0] - q[0], dy = t[1] - q[1], dz = t[2] - q[2];
                        double d2 = (dx * dx + dy * dy) + dz * dz;
                        if (d2 < best || (d2 == best && g->order[k] < bi)) { best = d2; bi = g->order[k]; }
                    }
                }
            }
        }
    }
    return (bi >= 0 && best < md2) ? best : -1.0;
}

/* nh hypotheses Ts[nh][16] scored over the whole source: tasks are (hypothesis, chunk of the source), so a handful of
 * hypotheses still uses every core; per hypothesis the inlier distances are gathered in source order and added in
 * numpy's pairwise order. */
ORC_API void orc_score_many(void *grid, const double *src, int64_t n_src, const double *Ts, int64_t nh, double max_d,
                            int64_t *out_cnt, double *out_err2) {
    const orc_grid_t *g = (const orc_grid_t *)grid;
    for (int64_t h = 0; h < nh; ++h) { out_cnt[h] = 0; out_err2[h] = 0.0; }
    if (!g || g->n_tgt == 0 || n_src == 0 || nh == 0 || !(max_d > 0)) return;
    const double md2 = max_d * max_d;
    const int64_t CH = 4096;
    const int64_t nchunk = (n_src + CH - 1) / CH;
    double *d2 = (double *)malloc(sizeof(double) * (size_t)nh * (size_t)n_src);
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t task = 0; task < nh * nchunk; ++task) {
        const int64_t h = task / nchunk, c0 = (task % nchunk) * CH;
        const int64_t c1 = c0 + CH < n_src ? c0 + CH : n_src;
        const double *T = Ts + 16 * h;
        for (int64_t i = c0; i < c1; ++i) {
            double q[3];
            xform(T, src + 3 * i, q);
            d2[h * n_src + i] = grid_nn_d2(g, q, max_d, md2);
        }
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t h = 0; h < nh; ++h) {
        double *row = d2 + h * n_src;
        int64_t cnt = 0;
        for (int64_t i = 0; i < n_src; ++i) if (row[i] >= 0.0) row[cnt++] = row[i];
        out_cnt[h] = cnt;
        out_err2[h] = pairwise_sum(row, cnt);
    }
    free(d2);
}

/* transforms of hypotheses given by samples[nh][3] (indices into corres) */
ORC_API void orc_hyp_transforms(const double *src, const double *tgt, const int32_t *corres, const int32_t *samples, int64_t nh,
                                double *T_out) {
#pragma omp parallel for
    for (int64_t h = 0; h < nh; ++h) {
        double s[9], t[9];
        for (int j = 0; j < 3; ++j) {
            int c = samples[3 * h + j];
            for (int a = 0; a < 3; ++a) {
                s[3 * j + a] = src[3 * (size_t)corres[2 * c] + a];
                t[3 * j + a] = tgt[3 * (size_t)corres[2 * c + 1] + a];
            }
        }
        kabsch(s, t, 3, T_out + 16 * h);
    }
}
