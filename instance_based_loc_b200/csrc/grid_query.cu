// grid_query.cu -- exact neighbour queries on the cell-sorted grid.
//   radius count  : warp per query, contiguous x-rows of cells streamed with coalesced 32 B loads
//   hybrid k-NN   : warp per query, Chebyshev ring expansion, shared-memory candidate buffer,
//                   warp bitonic sort on (d2 bits, index) -> ascending (d2, index) like the oracle
//   1-NN + score  : thread per query (ring field skips empty space), deterministic block reduce
#include <mutex>
#include <vector>

#include "ibl_common.cuh"

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) radius_count_kernel(GridView g, int64_t n, double r2, int S, int32_t nb_points,
                                                           int32_t *__restrict__ count, uint8_t *__restrict__ keep) {
    int lane = threadIdx.x & 31;
    int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t qi = warp; qi < n; qi += nwarps) {
        double4 q = g.sorted[qi];
        int cx = ibl_clampi(ibl_cell_coord(q.x, g.ox, g.inv_cell, g.nx), 0, g.nx - 1);
        int cy = ibl_clampi(ibl_cell_coord(q.y, g.oy, g.inv_cell, g.ny), 0, g.ny - 1);
        int cz = ibl_clampi(ibl_cell_coord(q.z, g.oz, g.inv_cell, g.nz), 0, g.nz - 1);
        int x0 = max(cx - S, 0), x1 = min(cx + S, g.nx - 1);
        int y0 = max(cy - S, 0), y1 = min(cy + S, g.ny - 1);
        int z0 = max(cz - S, 0), z1 = min(cz + S, g.nz - 1);
        // The count is warp-uniform (ballot + popc per batch of 32 candidates).  When only the keep mask is wanted the scan
        // stops as soon as the count exceeds nb_points -- with ~140 neighbours inside the radius and nb_points = 8 that is
        // after the first row or two, so rows are visited centre first (dz, dy = 0, -1, +1, -2, +2, ...).
        int c = 0;
        const bool early = (count == nullptr);
        bool done = false;
        for (int kz = 0; kz <= 2 * S && !done; ++kz) {
            int z = cz + ((kz & 1) ? -((kz + 1) >> 1) : (kz >> 1));
            if (z < z0 || z > z1) continue;
            for (int ky = 0; ky <= 2 * S && !done; ++ky) {
                int y = cy + ((ky & 1) ? -((ky + 1) >> 1) : (ky >> 1));
                if (y < y0 || y > y1) continue;
                size_t rb = ((size_t)z * g.ny + y) * g.nx;
                int p0 = g.cell_start[rb + x0], p1 = g.cell_start[rb + x1 + 1];
                for (int pb = p0; pb < p1; pb += 32) {
                    int p = pb + lane;
                    bool in = false;
                    if (p < p1) {
                        double4 t = g.sorted[p];
                        in = ibl_d2(t.x, t.y, t.z, q.x, q.y, q.z) < r2;
                        // segmented grid: only points of the query's own cloud count
                        if (g.segmented) in = in && ((__double_as_longlong(t.w) >> 32) == (__double_as_longlong(q.w) >> 32));
                    }
                    c += __popc(__ballot_sync(0xffffffffu, in));
                    if (early && c > nb_points) { done = true; break; }
                }
            }
        }
        if (lane == 0) {
            int oi = (int)__double_as_longlong(q.w);
            if (count) count[oi] = c;
            if (keep) keep[oi] = c > nb_points;
        }
    }
}

static int radius_launch(const ibl_grid_t *g, double radius, int32_t nb, int32_t *count, uint8_t *keep, void *stream) {
    IBL_ARG(g != nullptr && radius > 0);
    if (g->n == 0) return IBL_OK;
    int S = (int)ceil(radius * g->inv_cell + 1e-6);
    int64_t warps_needed = g->n;
    int blocks = (int)((warps_needed * 32 + 255) / 256);
    int maxb = IBL_SMS * 16;
    if (blocks > maxb) blocks = maxb;
    IBL_PROF(IBL_PROF_RADIUS, ibl_stream(stream), 33.0 * g->n);
    radius_count_kernel<<<blocks, 256, 0, ibl_stream(stream)>>>(ibl_view(g), g->n, radius * radius, S, nb, count, keep);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}
IBL_API int ibl_radius_count(const ibl_grid_t *g, double radius, int32_t *count, void *stream) {
    return radius_launch(g, radius, 0, count, nullptr, stream);
}
IBL_API int ibl_radius_outlier_mask(const ibl_grid_t *g, double radius, int32_t nb_points, uint8_t *keep, void *stream) {
    return radius_launch(g, radius, nb_points, nullptr, keep, stream);
}

// ------------------------------------------------------------------------------------------------
// hybrid k-NN
struct KnnBuf {
    unsigned long long *key;   // d2 bit pattern (non-negative doubles order like uint64)
    int *val;                  // original index
};

__device__ __forceinline__ bool knn_less(unsigned long long ka, int va, unsigned long long kb, int vb) {
    return ka < kb || (ka == kb && va < vb);
}

// warp-cooperative bitonic sort of buf[0..n2) (n2 power of two >= 32)
__device__ void knn_sort(KnnBuf b, int n2, int lane) {
    for (int k = 2; k <= n2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = lane; t < (n2 >> 1); t += 32) {
                int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                int p = i | j;
                bool up = (i & k) == 0;
                unsigned long long ka = b.key[i], kb = b.key[p];
                int va = b.val[i], vb = b.val[p];
                bool gt = knn_less(kb, vb, ka, va);
                if (gt == up) { b.key[i] = kb; b.val[i] = vb; b.key[p] = ka; b.val[p] = va; }
            }
            __syncwarp();
        }
    }
}

// Bitonic sort of 128 (key, val) pairs held in REGISTERS, four per lane (pair i = 32 r + lane sits in slot r of its lane):
// the 25 stages whose partner is in another lane exchange through three 32-bit shuffles, the 3 stages with stride 32 / 64
// stay inside the lane.  Ascending (key, val); same comparison as knn_less, so the result equals knn_sort's.  The shared-
// memory version costs ~2900 instructions per query at 128 entries (31 % of the k-NN kernel, ncu), this one ~1300.
__device__ __forceinline__ void knn_cex(unsigned long long &ka, int &va, unsigned long long &kb, int &vb, bool a_keeps_min) {
    const bool b_less = knn_less(kb, vb, ka, va);
    if (b_less == a_keeps_min) { unsigned long long tk = ka; ka = kb; kb = tk; int tv = va; va = vb; vb = tv; }
}
__device__ __forceinline__ void knn_sort128_regs(unsigned long long (&key)[4], int (&val)[4], int lane) {
#pragma unroll
    for (int k = 2; k <= 128; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= 32) {
                // partner slot r ^ (j / 32) of the same lane; the pair (r, r | jr) with r & jr == 0 is sorted up iff (i & k) == 0
                const int jr = j >> 5;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if ((r & jr) == 0) {
                        const bool up = (((r << 5) | lane) & k) == 0;
                        knn_cex(key[r], val[r], key[r | jr], val[r | jr], up);
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int i = (r << 5) | lane;
                    const bool up = (i & k) == 0, lower = (lane & j) == 0;
                    const int olo = __shfl_xor_sync(0xffffffffu, (int)(unsigned)(key[r] & 0xffffffffull), j);
                    const int ohi = __shfl_xor_sync(0xffffffffu, (int)(unsigned)(key[r] >> 32), j);
                    const int ov = __shfl_xor_sync(0xffffffffu, val[r], j);
                    const unsigned long long ok = ((unsigned long long)(unsigned)ohi << 32) | (unsigned long long)(unsigned)olo;
                    const bool keep_min = lower == up;      // the lower index of a pair keeps the minimum when sorting up
                    const bool other_less = knn_less(ok, ov, key[r], val[r]);
                    if (other_less == keep_min) { key[r] = ok; val[r] = ov; }
                }
            }
        }
    }
}

__device__ __forceinline__ int next_pow2_ge32(int c) {
    int n2 = 32;
    while (n2 < c) n2 <<= 1;
    return n2;
}

// Two passes per query (one warp each).
//   pass 1 : Chebyshev ring expansion that only COUNTS: every candidate within the radius goes into a histogram of d2
//            (128 bins of equal width cell^2 / 8 in SQUARED distance: float conversion, one multiplication, truncation).
//            After ring s every point closer than s*cell has been seen, so as soon as the bins below that bound hold max_nn
//            points the search stops, and the first bin j* whose cumulative count reaches max_nn is a distance bound that
//            provably contains the max_nn nearest (and every tie with the last one).
//   pass 2 : the cells of the cube again, but only the part of it that the bound can reach: a row of cells whose rectangle
//            in (y, z) lies farther than the bound is skipped and the others are clipped in x; only candidates in bins
//            <= j* are stored -- about max_nn + one bin's population -- then ONE warp bitonic sort on (d2 bits, index)
//            gives the oracle's ascending (d2, index) order.
// Both passes first compute the (first, last) positions of up to 32 cell segments in PARALLEL (one lane per segment: two
// loads of cell_start each), then walk the non-empty ones: the old per-row loop paid two dependent global loads per row of
// cells before it could touch a point (ncu: 25 % of the samples sat on the first use of a candidate, IPC 2.2), and most
// rows of a cube around a surface are empty.
// If the stored set overflows the buffer (duplicates, a degenerate bin) a compaction -- sort, keep the best max_nn,
// remember the threshold -- takes over.
#define KNN_NB 128
// histogram bin of a squared distance.  The only properties used are that the map is monotone non-decreasing in d2
// (conversion to float, multiplication by a positive constant and truncation all are) and that both passes use the same map.
__device__ __forceinline__ int knn_bin(double d2, float inv_w2) {
    return (int)fminf(__fmul_rn((float)d2, inv_w2), (float)(KNN_NB - 1));
}
// positions [p0, p1) in g.sorted of the cells xa..xb of row (y, z); empty if the row or the range is outside the grid
__device__ __forceinline__ void knn_segment(const GridView &g, int z, int y, int xa, int xb, int &p0, int &p1) {
    p0 = p1 = 0;
    xa = max(xa, 0); xb = min(xb, g.nx - 1);
    if (z < 0 || z >= g.nz || y < 0 || y >= g.ny || xa > xb) return;
    const size_t rb = ((size_t)z * g.ny + y) * g.nx;
    p0 = g.cell_start[rb + xa];
    p1 = g.cell_start[rb + xb + 1];
}
// distance from q to the interval of coordinates that cell c of an axis can hold (widened by 1e-6 cell; the two boundary
// cells are open-ended, because points are clamped into them)
__device__ __forceinline__ double knn_axis_gap(double q, double o, double cell, int c, int n) {
    const double lo = o + ((double)c - 1e-6) * cell, hi = o + ((double)c + 1.0 + 1e-6) * cell;
    double d = 0.0;
    if (c > 0 && q < lo) d = lo - q;
    if (c < n - 1 && q > hi) d = q - hi;
    return d;
}
template <int CAP>
__global__ void __launch_bounds__(128) knn_kernel(GridView g, const double *__restrict__ queries, int64_t nq,
                                                  double radius, double r2, int max_nn, int32_t *__restrict__ idx,
                                                  double *__restrict__ d2out, int32_t *__restrict__ cnt) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    KnnBuf b;
    b.key = reinterpret_cast<unsigned long long *>(smem_raw) + (size_t)wib * CAP;
    b.val = reinterpret_cast<int *>(smem_raw + sizeof(unsigned long long) * CAP * wpb) + (size_t)wib * CAP;
    int *hist = reinterpret_cast<int *>(smem_raw + (sizeof(unsigned long long) + sizeof(int)) * CAP * wpb) + (size_t)wib * KNN_NB;
    int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int s_rad = isfinite(radius) ? (int)ceil(radius * g.inv_cell + 1e-6) : 0x3fffffff;
    const float inv_w2 = (float)(8.0 * g.inv_cell * g.inv_cell);   // bins of cell^2 / 8 in SQUARED distance: 128 bins reach 4 cells
    for (int64_t qi = warp; qi < nq; qi += nwarps) {
        double qx, qy, qz;
        int64_t row;
        if (queries == nullptr) {
            double4 q = g.sorted[qi];
            qx = q.x; qy = q.y; qz = q.z;
            row = (int64_t)__double_as_longlong(q.w);
        } else {
            qx = queries[3 * qi]; qy = queries[3 * qi + 1]; qz = queries[3 * qi + 2];
            row = qi;
        }
        const int cx = ibl_cell_coord(qx, g.ox, g.inv_cell, g.nx);
        const int cy = ibl_cell_coord(qy, g.oy, g.inv_cell, g.ny);
        const int cz = ibl_cell_coord(qz, g.oz, g.inv_cell, g.nz);
        int s_begin;                                       // rings needed to reach the grid box from an outside query
        {
            int kx = ibl_clampi(cx, 0, g.nx - 1), ky = ibl_clampi(cy, 0, g.ny - 1), kz = ibl_clampi(cz, 0, g.nz - 1);
            s_begin = max(max(abs(cx - kx), abs(cy - ky)), abs(cz - kz));
        }
        // ---------------- pass 1: count ----------------
        for (int t = lane; t < KNN_NB; t += 32) hist[t] = 0;
        __syncwarp();
        int s_end = s_begin, jstar = KNN_NB;               // jstar == KNN_NB: no bin bound (the radius / the whole grid decides)
        for (int s = s_begin;; ++s) {
            // ring s = the cells at Chebyshev distance s.  Slot j = (row of the (2s+1)^2 rows in (z, y), segment 0 / 1): a
            // row on the shell's faces in z or y is ONE segment cx-s .. cx+s; an inner row is the two single cells cx-s, cx+s.
            // (rows outside the grid are not enumerated: a query far outside the box would otherwise walk empty slots)
            const int z0 = max(cz - s, 0), z1 = min(cz + s, g.nz - 1), y0 = max(cy - s, 0), y1 = min(cy + s, g.ny - 1);
            const int Wy = max(y1 - y0 + 1, 0), Wz = max(z1 - z0 + 1, 0);
            const long long nslots = 2ll * Wy * Wz;
            for (long long j0 = 0; j0 < nslots; j0 += 32) {
                const long long j = j0 + lane;
                int p0 = 0, p1 = 0;
                if (j < nslots) {
                    const int rw = (int)(j >> 1), seg = (int)(j & 1);
                    const int z = z0 + rw / Wy, y = y0 + rw % Wy;
                    const bool edge = z == cz - s || z == cz + s || y == cy - s || y == cy + s;
                    if (edge) { if (seg == 0) knn_segment(g, z, y, cx - s, cx + s, p0, p1); }
                    else knn_segment(g, z, y, seg ? cx + s : cx - s, seg ? cx + s : cx - s, p0, p1);
                }
                unsigned live = __ballot_sync(0xffffffffu, p1 > p0);
                while (live) {
                    const int src = __ffs(live) - 1;
                    live &= live - 1;
                    const int a = __shfl_sync(0xffffffffu, p0, src), e = __shfl_sync(0xffffffffu, p1, src);
                    for (int p = a + lane; p < e; p += 32) {
                        double4 t = g.sorted[p];
                        double d2 = ibl_d2(t.x, t.y, t.z, qx, qy, qz);
                        if (d2 < r2) atomicAdd(&hist[knn_bin(d2, inv_w2)], 1);
                    }
                }
            }
            __syncwarp();
            s_end = s;
            bool covers = (cx - s <= 0) && (cx + s >= g.nx - 1) && (cy - s <= 0) && (cy + s >= g.ny - 1) &&
                          (cz - s <= 0) && (cz + s >= g.nz - 1);
            if (covers || s >= s_rad) break;
            double gr = ((double)s - 1e-6) * g.cell;       // every point closer than this has been seen
            if (gr > 0) {
                // knn_bin is monotone in d2, so every candidate in a bin below knn_bin(gr^2) is closer than gr (the last bin
                // is open-ended and is never below anything)
                int jm = knn_bin(gr * gr, inv_w2) - 1;
                if (jm >= 0) {
                    int h0 = 4 * lane <= jm ? hist[4 * lane] : 0, h1 = 4 * lane + 1 <= jm ? hist[4 * lane + 1] : 0;
                    int h2 = 4 * lane + 2 <= jm ? hist[4 * lane + 2] : 0, h3 = 4 * lane + 3 <= jm ? hist[4 * lane + 3] : 0;
                    int own = h0 + h1 + h2 + h3, incl = own;
                    for (int o = 1; o < 32; o <<= 1) { int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
                    int total = __shfl_sync(0xffffffffu, incl, 31);
                    if (total >= max_nn) {
                        int before = incl - own;
                        unsigned m = __ballot_sync(0xffffffffu, before < max_nn && incl >= max_nn);
                        int src_lane = __ffs(m) - 1;
                        int j = 4 * lane;
                        if (before + h0 < max_nn) { j = 4 * lane + 1; if (before + h0 + h1 < max_nn) { j = 4 * lane + 2; if (before + h0 + h1 + h2 < max_nn) j = 4 * lane + 3; } }
                        jstar = __shfl_sync(0xffffffffu, j, src_lane);
                        break;
                    }
                }
            }
        }
        // ---------------- pass 2: store the candidates of bins <= jstar from the reachable part of the cube ----------------
        int count = 0;
        bool have_tau = false;
        unsigned long long tau_k = ~0ull;
        int tau_v = 0x7fffffff;
        {
            const int s = s_end;
            const int z0 = max(cz - s, 0), z1 = min(cz + s, g.nz - 1), y0 = max(cy - s, 0), y1 = min(cy + s, g.ny - 1);
            const int Wy = max(y1 - y0 + 1, 0), Wz = max(z1 - z0 + 1, 0);
            const long long nrows = (long long)Wy * Wz;
            // every stored candidate has knn_bin(d2) <= jstar, hence d2 below the upper edge of bin jstar (taken 1e-5 wide of
            // the float arithmetic in knn_bin) and below r2: nothing farther than ub2 can be stored
            double ub2 = r2;
            if (jstar < KNN_NB - 1) ub2 = fmin(ub2, ((double)(jstar + 1) / (double)inv_w2) * (1.0 + 1e-5));
            for (long long j0 = 0; j0 < nrows; j0 += 32) {
                const long long j = j0 + lane;
                int p0 = 0, p1 = 0;
                if (j < nrows) {
                    const int z = z0 + (int)(j / Wy), y = y0 + (int)(j % Wy);
                    int xa = cx - s, xb = cx + s;
                    bool reach = true;
                    if (isfinite(ub2)) {
                        const double gy = knn_axis_gap(qy, g.oy, g.cell, y, g.ny), gz = knn_axis_gap(qz, g.oz, g.cell, z, g.nz);
                        const double rest = ub2 - (gy * gy + gz * gz);
                        if (rest < 0.0) reach = false;
                        else {
                            const double dxm = sqrt(rest) + 1e-6 * g.cell;
                            xa = max(xa, ibl_cell_coord(qx - dxm, g.ox, g.inv_cell, g.nx));
                            xb = min(xb, ibl_cell_coord(qx + dxm, g.ox, g.inv_cell, g.nx));
                            // points are clamped into the boundary cells: a range that ends outside the grid includes them
                            if (xb < 0) xb = 0;
                            if (xa > g.nx - 1) xa = g.nx - 1;
                        }
                    }
                    if (reach) knn_segment(g, z, y, xa, xb, p0, p1);
                }
                unsigned live = __ballot_sync(0xffffffffu, p1 > p0);
                while (live) {
                    const int src = __ffs(live) - 1;
                    live &= live - 1;
                    const int a = __shfl_sync(0xffffffffu, p0, src), e = __shfl_sync(0xffffffffu, p1, src);
                    for (int pb = a; pb < e; pb += 32) {
                        if (count + 32 > CAP) {   // overflow: keep the max_nn best, remember the threshold
                            int n2 = next_pow2_ge32(count);
                            for (int t = count + lane; t < n2; t += 32) { b.key[t] = ~0ull; b.val[t] = 0x7fffffff; }
                            __syncwarp();
                            knn_sort(b, n2, lane);
                            if (count >= max_nn) {
                                count = max_nn;
                                have_tau = true;
                                tau_k = b.key[max_nn - 1];
                                tau_v = b.val[max_nn - 1];
                            }
                            __syncwarp();
                        }
                        int p = pb + lane;
                        bool ok = false;
                        unsigned long long kk = 0;
                        int vv = 0;
                        if (p < e) {
                            double4 t = g.sorted[p];
                            double d2 = ibl_d2(t.x, t.y, t.z, qx, qy, qz);
                            kk = (unsigned long long)__double_as_longlong(d2);
                            vv = (int)__double_as_longlong(t.w);
                            ok = d2 < r2 && knn_bin(d2, inv_w2) <= jstar &&
                                 (!have_tau || knn_less(kk, vv, tau_k, tau_v));
                        }
                        unsigned m = __ballot_sync(0xffffffffu, ok);
                        if (ok) {
                            int pos = count + __popc(m & lt_mask);
                            b.key[pos] = kk;
                            b.val[pos] = vv;
                        }
                        count += __popc(m);
                        __syncwarp();
                    }
                }
            }
        }
        // final ordering
        int outc = min(count, max_nn);
        if (count <= 128) {
            // the usual case (max_nn + one bin's population): sort in registers, write straight from them
            __syncwarp();
            unsigned long long key[4];
            int val[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int t = (r << 5) | lane;
                key[r] = t < count ? b.key[t] : ~0ull;
                val[r] = t < count ? b.val[t] : 0x7fffffff;
            }
            knn_sort128_regs(key, val, lane);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int t = (r << 5) | lane;
                if (t < max_nn) {
                    const bool v = t < outc;
                    idx[row * max_nn + t] = v ? val[r] : -1;
                    if (d2out) d2out[row * max_nn + t] = v ? __longlong_as_double((long long)key[r]) : INFINITY;
                }
            }
            for (int t = 128 + lane; t < max_nn; t += 32) {
                idx[row * max_nn + t] = -1;
                if (d2out) d2out[row * max_nn + t] = INFINITY;
            }
        } else {
            int n2 = next_pow2_ge32(count);
            for (int t = count + lane; t < n2; t += 32) { b.key[t] = ~0ull; b.val[t] = 0x7fffffff; }
            __syncwarp();
            knn_sort(b, n2, lane);
            for (int t = lane; t < max_nn; t += 32) {
                bool v = t < outc;
                idx[row * max_nn + t] = v ? b.val[t] : -1;
                if (d2out) d2out[row * max_nn + t] = v ? __longlong_as_double((long long)b.key[t]) : INFINITY;
            }
        }
        if (lane == 0) cnt[row] = outc;
        __syncwarp();
    }
}

IBL_API int ibl_knn_hybrid(const ibl_grid_t *g, const double *queries, int64_t nq, double radius, int32_t max_nn,
                           int32_t *idx, double *d2, int32_t *cnt, void *stream) {
    IBL_ARG(g != nullptr && max_nn >= 1 && max_nn <= 224 && idx != nullptr && cnt != nullptr && radius > 0);
    if (queries == nullptr) nq = g->n;
    if (nq == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    const int threads = 128, wpb = threads / 32;
    int64_t blocks64 = (nq + wpb - 1) / wpb;
    int blocks = (int)(blocks64 > IBL_SMS * 32 ? IBL_SMS * 32 : blocks64);
    double r2 = isfinite(radius) ? radius * radius : INFINITY;
    IBL_PROF(IBL_PROF_KNN, st, (double)nq * (32.0 + 4.0 + max_nn * (d2 ? 12.0 : 4.0)));
    if (max_nn <= 96) {
        constexpr int CAP = 256;
        size_t sm = (size_t)wpb * (CAP * 12 + KNN_NB * 4);
        knn_kernel<CAP><<<blocks, threads, sm, st>>>(ibl_view(g), queries, nq, radius, r2, max_nn, idx, d2, cnt);
    } else {
        constexpr int CAP = 512;
        size_t sm = (size_t)wpb * (CAP * 12 + KNN_NB * 4);
        knn_kernel<CAP><<<blocks, threads, sm, st>>>(ibl_view(g), queries, nq, radius, r2, max_nn, idx, d2, cnt);
    }
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// 1-NN + registration score
struct Mat16 { double v[16]; };

__global__ void __launch_bounds__(256) nn_score_kernel(GridView g, const double *__restrict__ src, int64_t n, Mat16 T,
                                                       int has_T, double max_d, double max_d2,
                                                       int32_t *__restrict__ nn_idx, double *__restrict__ nn_d2,
                                                       long long *__restrict__ part_cnt, double *__restrict__ part_err) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int hit = 0;
    double e = 0.0;
    if (i < n) {
        double x = src[3 * i], y = src[3 * i + 1], z = src[3 * i + 2];
        if (has_T) { double a, b, c; ibl_xform(T.v, x, y, z, a, b, c); x = a; y = b; z = c; }
        double bd;
        int j = ibl_nn1_thread(g, x, y, z, max_d, max_d2, bd);
        if (nn_idx) nn_idx[i] = j;
        if (nn_d2) nn_d2[i] = j >= 0 ? bd : INFINITY;
        if (j >= 0) { hit = 1; e = bd; }
    }
    // deterministic block reduction (fixed shuffle tree, then warp 0 in index order)
    __shared__ double se[8];
    __shared__ int sc[8];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) { e += __shfl_down_sync(0xffffffffu, e, o); hit += __shfl_down_sync(0xffffffffu, hit, o); }
    if (lane == 0) { se[w] = e; sc[w] = hit; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; long long tc = 0;
        for (int k = 0; k < 8; ++k) { te += se[k]; tc += sc[k]; }
        part_cnt[blockIdx.x] = tc;
        part_err[blockIdx.x] = te;
    }
}

__global__ void score_final_kernel(const long long *part_cnt, const double *part_err, int nb, ibl_score_t *out) {
    // single warp, fixed order: lane-strided partial sums then shuffle tree
    int lane = threadIdx.x;
    double e = 0; long long c = 0;
    for (int k = lane; k < nb; k += 32) { e += part_err[k]; c += part_cnt[k]; }
    for (int o = 16; o; o >>= 1) { e += __shfl_down_sync(0xffffffffu, e, o); c += __shfl_down_sync(0xffffffffu, c, o); }
    if (lane == 0) { out->n_inliers = c; out->err2 = e; }
}

IBL_API int ibl_nn_score(const ibl_grid_t *g, const double *src, int64_t n, const double *h_T16, double max_d,
                         int32_t *nn_idx, double *nn_d2, ibl_score_t *h_result, void *stream) {
    IBL_ARG(g != nullptr && n >= 0 && max_d > 0);
    cudaStream_t st = ibl_stream(stream);
    ibl_score_t res = {0, 0.0};
    if (n > 0 && g->n > 0) {
        Mat16 T;
        int has_T = h_T16 != nullptr;
        for (int k = 0; k < 16; ++k) T.v[k] = has_T ? h_T16[k] : ((k % 5 == 0) ? 1.0 : 0.0);
        int nb = (int)((n + 255) / 256);
        long long *pc = nullptr; double *pe = nullptr; ibl_score_t *dres = nullptr;
        IBL_CUDA(ibl_alloc(&pc, (size_t)nb, st));
        IBL_CUDA(ibl_alloc(&pe, (size_t)nb, st));
        IBL_CUDA(ibl_alloc(&dres, 1, st));
        {
            IBL_PROF(IBL_PROF_NN_SCORE, st, (double)n * (24.0 + (nn_idx ? 4.0 : 0.0) + (nn_d2 ? 8.0 : 0.0)));
            nn_score_kernel<<<nb, 256, 0, st>>>(ibl_view(g), src, n, T, has_T, max_d, max_d * max_d, nn_idx, nn_d2, pc, pe);
            IBL_LAUNCH_CHECK();
        }
        score_final_kernel<<<1, 32, 0, st>>>(pc, pe, nb, dres);
        IBL_LAUNCH_CHECK();
        if (h_result) {
            IBL_CUDA(cudaMemcpyAsync(&res, dres, sizeof(res), cudaMemcpyDeviceToHost, st));
            IBL_CUDA(ibl_sync(st));
        }
        ibl_free(pc, st); ibl_free(pe, st); ibl_free(dres, st);
    } else if (n > 0 && nn_idx) {
        IBL_CUDA(cudaMemsetAsync(nn_idx, 0xff, sizeof(int32_t) * (size_t)n, st));
    }
    if (h_result) *h_result = res;
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// evaluate_registration for many (source, transform) pairs against one target grid in ONE launch: CTA `b` of query q is
// nn_score_kernel's block b (same per-CTA reduction), and one warp per query adds the CTA partials in score_final_kernel's
// order, so every result equals ibl_nn_score's.
struct NnsQuery { const double *src; int64_t n; double T[16]; int has_T; int pad; };

__global__ void __launch_bounds__(256) nn_score_batch_kernel(GridView g, const NnsQuery *__restrict__ qs, int nq,
                                                             const int64_t *__restrict__ blk_prefix, double max_d, double max_d2,
                                                             long long *__restrict__ part_cnt, double *__restrict__ part_err) {
    __shared__ int s_q;
    __shared__ double T[16];
    if (threadIdx.x == 0) {
        int lo = 0, hi = nq;
        const int64_t item = (int64_t)blockIdx.x;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (blk_prefix[mid] <= item) lo = mid; else hi = mid; }
        s_q = lo;
    }
    __syncthreads();
    const NnsQuery *Q = qs + s_q;
    if (threadIdx.x < 16) T[threadIdx.x] = Q->T[threadIdx.x];
    __syncthreads();
    const int64_t i = ((int64_t)blockIdx.x - blk_prefix[s_q]) * 256 + threadIdx.x;
    int hit = 0;
    double e = 0.0;
    if (i < Q->n) {
        const double *__restrict__ src = Q->src;
        double x = src[3 * i], y = src[3 * i + 1], z = src[3 * i + 2];
        if (Q->has_T) { double a, b, c; ibl_xform(T, x, y, z, a, b, c); x = a; y = b; z = c; }
        double bd;
        int j = ibl_nn1_thread(g, x, y, z, max_d, max_d2, bd);
        if (j >= 0) { hit = 1; e = bd; }
    }
    __shared__ double se[8];
    __shared__ int sc[8];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int o = 16; o; o >>= 1) { e += __shfl_down_sync(0xffffffffu, e, o); hit += __shfl_down_sync(0xffffffffu, hit, o); }
    if (lane == 0) { se[w] = e; sc[w] = hit; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double te = 0; long long tc = 0;
        for (int k = 0; k < 8; ++k) { te += se[k]; tc += sc[k]; }
        part_cnt[blockIdx.x] = tc;
        part_err[blockIdx.x] = te;
    }
}

__global__ void score_final_batch_kernel(const long long *part_cnt, const double *part_err, const int64_t *blk_prefix, ibl_score_t *out) {
    const int q = blockIdx.x, lane = threadIdx.x;
    const int64_t b0 = blk_prefix[q];
    const int nb = (int)(blk_prefix[q + 1] - b0);
    double e = 0; long long c = 0;
    for (int k = lane; k < nb; k += 32) { e += part_err[b0 + k]; c += part_cnt[b0 + k]; }
    for (int o = 16; o; o >>= 1) { e += __shfl_down_sync(0xffffffffu, e, o); c += __shfl_down_sync(0xffffffffu, c, o); }
    if (lane == 0) { out[q].n_inliers = c; out[q].err2 = e; }
}

IBL_API int ibl_nn_score_batch(const ibl_grid_t *g, const double *const *srcs, const int64_t *ns, const double *h_T16,
                               int32_t n_queries, double max_d, ibl_score_t *h_results, void *stream) {
    IBL_ARG(g != nullptr && n_queries >= 0 && max_d > 0 && (n_queries == 0 || (srcs && ns && h_results)));
    if (n_queries == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    std::vector<NnsQuery> hq(n_queries);
    std::vector<int64_t> prefix(n_queries + 1, 0);
    double work = 0;
    for (int q = 0; q < n_queries; ++q) {
        IBL_ARG(ns[q] >= 0 && (ns[q] == 0 || srcs[q]));
        hq[q].src = srcs[q]; hq[q].n = g->n > 0 ? ns[q] : 0; hq[q].has_T = h_T16 != nullptr; hq[q].pad = 0;
        for (int k = 0; k < 16; ++k) hq[q].T[k] = h_T16 ? h_T16[16 * (size_t)q + k] : ((k % 5 == 0) ? 1.0 : 0.0);
        prefix[q + 1] = prefix[q] + (hq[q].n + 255) / 256;
        work += (double)hq[q].n * 24.0;
        h_results[q].n_inliers = 0; h_results[q].err2 = 0.0;
    }
    const int64_t nb = prefix[n_queries];
    if (nb == 0) return IBL_OK;
    NnsQuery *dq = nullptr; int64_t *dpre = nullptr; long long *pc = nullptr; double *pe = nullptr; ibl_score_t *dres = nullptr;
    IBL_CUDA(ibl_alloc(&dq, (size_t)n_queries, st));
    IBL_CUDA(ibl_alloc(&dpre, (size_t)n_queries + 1, st));
    IBL_CUDA(ibl_alloc(&pc, (size_t)nb, st));
    IBL_CUDA(ibl_alloc(&pe, (size_t)nb, st));
    IBL_CUDA(ibl_alloc(&dres, (size_t)n_queries, st));
    IBL_CUDA(cudaMemcpyAsync(dq, hq.data(), sizeof(NnsQuery) * n_queries, cudaMemcpyHostToDevice, st));
    IBL_CUDA(cudaMemcpyAsync(dpre, prefix.data(), sizeof(int64_t) * (n_queries + 1), cudaMemcpyHostToDevice, st));
    {
        IBL_PROF(IBL_PROF_NN_SCORE, st, work);
        nn_score_batch_kernel<<<(unsigned)nb, 256, 0, st>>>(ibl_view(g), dq, n_queries, dpre, max_d, max_d * max_d, pc, pe);
        IBL_LAUNCH_CHECK();
    }
    score_final_batch_kernel<<<(unsigned)n_queries, 32, 0, st>>>(pc, pe, dpre, dres);
    IBL_LAUNCH_CHECK();
    IBL_CUDA(cudaMemcpyAsync(h_results, dres, sizeof(ibl_score_t) * n_queries, cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(dq, st); ibl_free(dpre, st); ibl_free(pc, st); ibl_free(pe, st); ibl_free(dres, st);
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// ibl_nn_score_static: ibl_nn_score for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
//
// The per-CTA partials that ibl_nn_score keeps in stream-ordered scratch live here in ONE file-scope device array of a
// fixed size, held by every GPU that loads the library until exit: IBL_NN_SCORE_STATIC_MAX_POINTS (2^20) source points
// -> 4096 CTAs x (int64 count + double squared error) = 64 KiB.  The same two kernels run, in the same order, so the
// result equals ibl_nn_score's bit for bit.  The result goes to a caller-owned DEVICE ibl_score_t: nothing is read back,
// nothing synchronises, nothing is allocated or freed.
//
// In host code `g_nns_static` is only the host-side stand-in of the symbol; the device address is looked up ONCE per
// device (with that device current) by cudaGetSymbolAddress and cached.
//
// Every call shares the array and the array keeps the last call's contents.  Nothing in it has to start at zero: a call
// with nb CTAs overwrites partials [0, nb) before its final kernel reads exactly those, so stale values are never read.
// Calls that OVERLAP on the GPU (different streams, or graph branches without an edge between them) race on the array;
// the caller serialises them.
struct NnsStatic {
    long long cnt[IBL_NN_SCORE_STATIC_MAX_POINTS / 256];
    double err[IBL_NN_SCORE_STATIC_MAX_POINTS / 256];
};
static __device__ __align__(256) NnsStatic g_nns_static;

#define NNS_MAX_DEVICES 64
static NnsStatic *g_nns_static_ptr[NNS_MAX_DEVICES];      // device address of g_nns_static, per device ordinal
static std::mutex g_nns_static_mu;

static int nns_static_address(NnsStatic **out) {
    int dev = -1;
    IBL_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NNS_MAX_DEVICES) {
        ibl_set_error("ibl_nn_score_static: device ordinal %d outside [0, %d)", dev, NNS_MAX_DEVICES);
        return IBL_E_RANGE;
    }
    std::lock_guard<std::mutex> lk(g_nns_static_mu);
    if (g_nns_static_ptr[dev] == nullptr) {
        NnsStatic *p = nullptr;
        IBL_CUDA(cudaGetSymbolAddress((void **)&p, g_nns_static));
        if (p == nullptr) {
            ibl_set_error("ibl_nn_score_static: cudaGetSymbolAddress returned a null address on device %d", dev);
            return IBL_E_CUDA;
        }
        g_nns_static_ptr[dev] = p;
    }
    *out = g_nns_static_ptr[dev];
    return IBL_OK;
}

IBL_API int ibl_nn_score_static(const ibl_grid_t *g, const double *src, int64_t n, const double *h_T16, double max_d,
                                int32_t *nn_idx, double *nn_d2, ibl_score_t *d_result, void *stream) {
    IBL_ARG(g != nullptr && n >= 0 && max_d > 0 && d_result != nullptr && (n == 0 || src != nullptr));
    if (n > IBL_NN_SCORE_STATIC_MAX_POINTS) {
        ibl_set_error("ibl_nn_score_static: n = %lld source points exceeds the static scratch limit of %d "
                      "(IBL_NN_SCORE_STATIC_MAX_POINTS); use ibl_nn_score", (long long)n, IBL_NN_SCORE_STATIC_MAX_POINTS);
        return IBL_E_RANGE;
    }
    cudaStream_t st = ibl_stream(stream);
    if (n == 0 || g->n == 0) {        // {0, 0.0} is all-zero bytes; nn_idx = -1, nn_d2 untouched, as in ibl_nn_score
        IBL_CUDA(cudaMemsetAsync(d_result, 0, sizeof(ibl_score_t), st));
        if (n > 0 && nn_idx) IBL_CUDA(cudaMemsetAsync(nn_idx, 0xff, sizeof(int32_t) * (size_t)n, st));
        return IBL_OK;
    }
    NnsStatic *scratch = nullptr;
    int rc = nns_static_address(&scratch);
    if (rc != IBL_OK) return rc;
    Mat16 T;
    const int has_T = h_T16 != nullptr;
    for (int k = 0; k < 16; ++k) T.v[k] = has_T ? h_T16[k] : ((k % 5 == 0) ? 1.0 : 0.0);
    const int nb = (int)((n + 255) / 256);
    nn_score_kernel<<<nb, 256, 0, st>>>(ibl_view(g), src, n, T, has_T, max_d, max_d * max_d, nn_idx, nn_d2,
                                        scratch->cnt, scratch->err);
    IBL_LAUNCH_CHECK();
    score_final_kernel<<<1, 32, 0, st>>>(scratch->cnt, scratch->err, nb, d_result);
    IBL_LAUNCH_CHECK();
    return IBL_OK;
}
