/*
 * oracle/o3d_core.c -- TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline).
 *
 * CPU restatement (plain C, IEEE f64, no FMA contraction) of the Open3D 0.17.0 arithmetic
 * that the reference's hot path delegates to (reference call sites:
 * utils/fpfh_register.py:86-150, utils/depth_utils.py:41,87, object_memory/object_memory.py:994).
 * Open3D 0.17.0 (pip dependency, environment.yml:229) is NOT vendored under /root/reference and
 * is not installable here, so this file restates its published algorithm (SURVEY.md Appendix A;
 * upstream cpp/open3d/geometry/EstimateNormals.cpp, pipelines/registration/{Feature,Registration,
 * ColoredICP,CorrespondenceChecker}.cpp, utility/Eigen.cpp).  PARITY UNPINNED: no golden vector
 * in the reference pins these outputs.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libibl.so) never links or calls it.
 *
 * Neighbour lists are supplied by the Python side (scipy cKDTree + exact re-ranking), see
 * oracle/o3d_core.py.  All arrays are row-major; points are double[n][3].
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ---------------------------------------------------------------------------------------------
 * A.3  FastEigen3x3 (utility/Eigen.cpp / EstimateNormals.cpp, after Eberly's robust 3x3 solver)
 * ------------------------------------------------------------------------------------------- */
static void cross3(const double *a, const double *b, double *o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
static double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

/* A is symmetric, stored as a00 a01 a02 a11 a12 a22 */
static void eigvec0(const double *A, double ev, double *out) {
    double r0[3] = {A[0] - ev, A[1], A[2]};
    double r1[3] = {A[1], A[3] - ev, A[4]};
    double r2[3] = {A[2], A[4], A[5] - ev};
    double c01[3], c02[3], c12[3];
    cross3(r0, r1, c01);
    cross3(r0, r2, c02);
    cross3(r1, r2, c12);
    double d0 = dot3(c01, c01), d1 = dot3(c02, c02), d2 = dot3(c12, c12);
    double dmax = d0;
    int imax = 0;
    if (d1 > dmax) { dmax = d1; imax = 1; }
    if (d2 > dmax) { imax = 2; }
    const double *c = imax == 0 ? c01 : (imax == 1 ? c02 : c12);
    double d = imax == 0 ? d0 : (imax == 1 ? d1 : d2);
    double s = sqrt(d);
    out[0] = c[0] / s; out[1] = c[1] / s; out[2] = c[2] / s;
}

static void eigvec1(const double *A, const double *e0, double ev1, double *out) {
    double U[3], V[3];
    if (fabs(e0[0]) > fabs(e0[1])) {
        double inv = 1.0 / sqrt(e0[0] * e0[0] + e0[2] * e0[2]);
        U[0] = -e0[2] * inv; U[1] = 0.0; U[2] = e0[0] * inv;
    } else {
        double inv = 1.0 / sqrt(e0[1] * e0[1] + e0[2] * e0[2]);
        U[0] = 0.0; U[1] = e0[2] * inv; U[2] = -e0[1] * inv;
    }
    cross3(e0, U, V);
    double AU[3] = {A[0] * U[0] + A[1] * U[1] + A[2] * U[2],
                    A[1] * U[0] + A[3] * U[1] + A[4] * U[2],
                    A[2] * U[0] + A[4] * U[1] + A[5] * U[2]};
    double AV[3] = {A[0] * V[0] + A[1] * V[1] + A[2] * V[2],
                    A[1] * V[0] + A[3] * V[1] + A[4] * V[2],
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

/* cov: a00 a01 a02 a11 a12 a22 ; out: eigenvector of the smallest eigenvalue (unoriented) */
static void fast_eigen3x3(const double *cov, double *out) {
    double A[6];
    double mx = cov[0];
    /* maxCoeff over the full symmetric matrix == max over the 6 unique entries */
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
            if (e2 < e0 && e2 < e1) { memcpy(out, v2, 24); return; }
            eigvec1(A, v2, e1, v1);
            if (e1 < e0 && e1 < e2) { memcpy(out, v1, 24); return; }
            cross3(v1, v2, out);
            return;
        } else {
            eigvec0(A, e0, v0);
            if (e0 < e1 && e0 < e2) { memcpy(out, v0, 24); return; }
            eigvec1(A, v0, e1, v1);
            if (e1 < e0 && e1 < e2) { memcpy(out, v1, 24); return; }
            cross3(v0, v1, out);
            return;
        }
    } else {
        /* diagonal matrix (A*max_coeff restores cov; comparisons are scale invariant) */
        if (cov[0] < cov[3] && cov[0] < cov[5]) { out[0] = 1; out[1] = 0; out[2] = 0; }
        else if (cov[3] < cov[0] && cov[3] < cov[5]) { out[0] = 0; out[1] = 1; out[2] = 0; }
        else { out[0] = 0; out[1] = 0; out[2] = 1; }
    }
}

ORC_API void orc_fast_eigen3x3(const double *cov6, double *out3) { fast_eigen3x3(cov6, out3); }

/* A.3 EstimateNormals: nbr_idx[n][max_nn] (ascending d2, self first), nbr_cnt[n] */
ORC_API void orc_normals(const double *pts, int n, const int32_t *nbr_idx, const int32_t *nbr_cnt,
                         int max_nn, double *normals) {
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < n; ++i) {
        int cnt = nbr_cnt[i];
        double cov[6];
        if (cnt >= 3) {
            double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
            for (int k = 0; k < cnt; ++k) {
                const double *p = pts + 3 * (size_t)nbr_idx[(size_t)i * max_nn + k];
                c[0] += p[0]; c[1] += p[1]; c[2] += p[2];
                c[3] += p[0] * p[0]; c[4] += p[0] * p[1]; c[5] += p[0] * p[2];
                c[6] += p[1] * p[1]; c[7] += p[1] * p[2]; c[8] += p[2] * p[2];
            }
            for (int k = 0; k < 9; ++k) c[k] /= (double)cnt;
            cov[0] = c[3] - c[0] * c[0];
            cov[1] = c[4] - c[0] * c[1];
            cov[2] = c[5] - c[0] * c[2];
            cov[3] = c[6] - c[1] * c[1];
            cov[4] = c[7] - c[1] * c[2];
            cov[5] = c[8] - c[2] * c[2];
        } else {
            cov[0] = 1; cov[1] = 0; cov[2] = 0; cov[3] = 1; cov[4] = 0; cov[5] = 1;
        }
        double nv[3];
        fast_eigen3x3(cov, nv);
        if (sqrt(nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2]) == 0.0) { nv[0] = 0; nv[1] = 0; nv[2] = 1; }
        normals[3 * (size_t)i + 0] = nv[0];
        normals[3 * (size_t)i + 1] = nv[1];
        normals[3 * (size_t)i + 2] = nv[2];
    }
}

/* ---------------------------------------------------------------------------------------------
 * A.4  FPFH (pipelines/registration/Feature.cpp)
 * ------------------------------------------------------------------------------------------- */
static void pair_features(const double *p1, const double *n1, const double *p2, const double *n2, double *f) {
    double d[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    f[0] = f[1] = f[2] = 0.0;
    f[3] = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (f[3] == 0.0) { return; }
    const double *na = n1, *nb = n2;
    double angle1 = dot3(n1, d) / f[3];
    double angle2 = dot3(n2, d) / f[3];
    if (acos(fabs(angle1)) > acos(fabs(angle2))) {
        na = n2; nb = n1;
        d[0] *= -1.0; d[1] *= -1.0; d[2] *= -1.0;
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

ORC_API void orc_pair_features(const double *p1, const double *n1, const double *p2, const double *n2, double *f4) {
    pair_features(p1, n1, p2, n2, f4);
}

static int clampbin(double x) {
    int h = (int)floor(x);
    if (h < 0) h = 0;
    if (h >= 11) h = 10;
    return h;
}

/* spfh is [33][n] (Open3D Feature layout: dim x points) */
ORC_API void orc_spfh(const double *pts, const double *nrm, int n, const int32_t *nbr_idx,
                      const int32_t *nbr_cnt, int max_nn, double *spfh) {
    memset(spfh, 0, sizeof(double) * 33 * (size_t)n);
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < n; ++i) {
        int cnt = nbr_cnt[i];
        if (cnt <= 1) continue;
        double inc = 100.0 / (double)(cnt - 1);
        for (int k = 1; k < cnt; ++k) {
            int j = nbr_idx[(size_t)i * max_nn + k];
            double f[4];
            pair_features(pts + 3 * (size_t)i, nrm + 3 * (size_t)i, pts + 3 * (size_t)j, nrm + 3 * (size_t)j, f);
            int h0 = clampbin(11 * (f[0] + M_PI) / (2.0 * M_PI));
            int h1 = clampbin(11 * (f[1] + 1.0) * 0.5);
            int h2 = clampbin(11 * (f[2] + 1.0) * 0.5);
            spfh[(size_t)h0 * n + i] += inc;
            spfh[(size_t)(h1 + 11) * n + i] += inc;
            spfh[(size_t)(h2 + 22) * n + i] += inc;
        }
    }
}

ORC_API void orc_fpfh(int n, const int32_t *nbr_idx, const double *nbr_d2, const int32_t *nbr_cnt,
                      int max_nn, const double *spfh, double *fpfh) {
    memset(fpfh, 0, sizeof(double) * 33 * (size_t)n);
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < n; ++i) {
        int cnt = nbr_cnt[i];
        if (cnt <= 1) continue;
        double sum[3] = {0, 0, 0};
        double acc[33];
        for (int j = 0; j < 33; ++j) acc[j] = 0.0;
        for (int k = 1; k < cnt; ++k) {
            double dist = nbr_d2[(size_t)i * max_nn + k];
            if (dist == 0.0) continue;
            int q = nbr_idx[(size_t)i * max_nn + k];
            for (int j = 0; j < 33; ++j) {
                double val = spfh[(size_t)j * n + q] / dist;
                sum[j / 11] += val;
                acc[j] += val;
            }
        }
        for (int j = 0; j < 3; ++j) if (sum[j] != 0.0) sum[j] = 100.0 / sum[j];
        for (int j = 0; j < 33; ++j) {
            acc[j] *= sum[j / 11];
            acc[j] += spfh[(size_t)j * n + i];
            fpfh[(size_t)j * n + i] = acc[j];
        }
    }
}

/* A.5 exact brute-force 33-d nearest neighbour: for every column of fa find argmin over fb columns.
 * d2 = sum_j (a_j-b_j)^2 accumulated j=0..32 in order; ties -> lowest index. fa:[33][na] fb:[33][nb] */
typedef struct { double norm; int32_t idx; } nrm_idx_t;
static int cmp_nrm(const void *x, const void *y) {
    const nrm_idx_t *a = (const nrm_idx_t *)x, *b = (const nrm_idx_t *)y;
    if (a->norm < b->norm) return -1;
    if (a->norm > b->norm) return 1;
    return (a->idx > b->idx) - (a->idx < b->idx);
}

/* exact distance in the defining order (j ascending), abandoned as soon as the partial sum EXCEEDS `best`: the terms are
 * non-negative and rounded addition is monotone, so such a sum cannot end at or below best.  Returns INFINITY then. */
static inline double feat_d2_bounded(const double *a, const double *b, double best) {
    double s = 0.0;
    for (int j = 0; j < 32; j += 8) {
        for (int k = 0; k < 8; ++k) { double d = a[j + k] - b[j + k]; s += d * d; }
        if (s > best) return INFINITY;
    }
    double d = a[32] - b[32];
    return s + d * d;
}

ORC_API void orc_feature_nn(const double *fa, int na, const double *fb, int nb, int32_t *nn) {
    /* transpose for locality */
    double *ta = (double *)malloc(sizeof(double) * 33 * (size_t)(na ? na : 1));
    double *tb = (double *)malloc(sizeof(double) * 33 * (size_t)(nb ? nb : 1));
    for (int j = 0; j < 33; ++j) {
        for (int i = 0; i < na; ++i) ta[(size_t)i * 33 + j] = fa[(size_t)j * na + i];
        for (int i = 0; i < nb; ++i) tb[(size_t)i * 33 + j] = fb[(size_t)j * nb + i];
    }
    /* Search order only (the answer is still the exact argmin, lowest index on ties): targets sorted by Euclidean norm;
     * a query starts at the targets whose norm is closest to its own and walks outwards in both directions, and a
     * direction is closed once (|a| - |b|)^2, a lower bound of |a - b|^2, exceeds the best distance so far. */
    nrm_idx_t *ord = (nrm_idx_t *)malloc(sizeof(nrm_idx_t) * (size_t)(nb ? nb : 1));
    for (int q = 0; q < nb; ++q) {
        double s = 0.0;
        for (int j = 0; j < 33; ++j) s += tb[(size_t)q * 33 + j] * tb[(size_t)q * 33 + j];
        ord[q].norm = sqrt(s); ord[q].idx = q;
    }
    qsort(ord, (size_t)nb, sizeof(nrm_idx_t), cmp_nrm);
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < na; ++i) {
        const double *a = ta + (size_t)i * 33;
        double an = 0.0;
        for (int j = 0; j < 33; ++j) an += a[j] * a[j];
        an = sqrt(an);
        int lo = 0, hi = nb;                                /* first target with norm >= an */
        while (lo < hi) { int mid = (lo + hi) >> 1; if (ord[mid].norm < an) lo = mid + 1; else hi = mid; }
        int up = lo, dn = lo - 1;
        double best = INFINITY;
        int bi = -1;
        while (up < nb || dn >= 0) {
            for (int side = 0; side < 2; ++side) {
                int *pos = side ? &dn : &up;
                if (side ? (*pos < 0) : (*pos >= nb)) continue;
                double gap = fabs(ord[*pos].norm - an);
                gap -= 1e-9 * (ord[*pos].norm + an);        /* the norms themselves carry rounding error */
                if (gap > 0.0 && gap * gap > best) { *pos = side ? -1 : nb; continue; }    /* nothing closer further out */
                int q = ord[*pos].idx;
                double s = feat_d2_bounded(a, tb + (size_t)q * 33, best);
                if (s < best || (s == best && q < bi)) { best = s; bi = q; }
                *pos += side ? -1 : 1;
            }
        }
        nn[i] = bi;
    }
    free(ta); free(tb); free(ord);
}

/* ---------------------------------------------------------------------------------------------
 * PointCloud::ClusterDBSCAN (upstream cpp/open3d/geometry/PointCloudCluster.cpp; reference call sites
 * object_memory/object_memory.py:315,511,632 -- memory build).
 * The sequential loop as published: radius neighbourhoods (d2 < eps^2, self included) are precomputed, labels start
 * at -2 (unvisited), points are scanned in index order; an unvisited point with fewer than min_points neighbours
 * becomes noise (-1), otherwise it seeds cluster `cluster_label` and the frontier set is expanded: a noise point met
 * from a core point is relabelled (border) and not expanded, an unvisited one is labelled and, if it is itself a core
 * point, its unvisited neighbours join the frontier.  Open3D keeps the frontier in an unordered_set; the labels do not
 * depend on the order it is drained in, here it is a stack with an "already queued/visited" flag per point.
 * nbr_ptr: int64[n+1] CSR offsets, nbr_idx: int32 neighbour indices (self included).  Returns the cluster count.
 * ------------------------------------------------------------------------------------------- */
ORC_API int orc_dbscan(int n, const int64_t *nbr_ptr, const int32_t *nbr_idx, int min_points, int32_t *labels) {
    int32_t *stack = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    uint8_t *seen = (uint8_t *)malloc((size_t)(n > 0 ? n : 1));   /* nbs_visited or already in nbs_next, per cluster */
    int32_t *touched = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; ++i) labels[i] = -2;
    memset(seen, 0, (size_t)(n > 0 ? n : 1));
    int cluster_label = 0;
    for (int idx = 0; idx < n; ++idx) {
        if (labels[idx] != -2) continue;
        int64_t b = nbr_ptr[idx], e = nbr_ptr[idx + 1];
        if (e - b < min_points) { labels[idx] = -1; continue; }
        int top = 0, nt = 0;
        seen[idx] = 1; touched[nt++] = idx;
        for (int64_t k = b; k < e; ++k) {
            int q = nbr_idx[k];
            if (!seen[q]) { seen[q] = 1; touched[nt++] = q; stack[top++] = q; }
        }
        labels[idx] = cluster_label;
        while (top > 0) {
            int nb = stack[--top];
            if (labels[nb] == -1) labels[nb] = cluster_label;
            if (labels[nb] != -2) continue;
            labels[nb] = cluster_label;
            int64_t b2 = nbr_ptr[nb], e2 = nbr_ptr[nb + 1];
            if (e2 - b2 >= min_points)
                for (int64_t k = b2; k < e2; ++k) {
                    int q = nbr_idx[k];
                    if (!seen[q]) { seen[q] = 1; touched[nt++] = q; stack[top++] = q; }
                }
        }
        for (int k = 0; k < nt; ++k) seen[touched[k]] = 0;
        cluster_label++;
    }
    free(stack); free(seen); free(touched);
    return cluster_label;
}
