// ibl_common.cuh -- shared device/host helpers for libibl.so (sm_100a).
// All f64 arithmetic that decides a count or an index uses explicit _rn intrinsics so the result
// does not depend on FMA contraction (the library is also built with -fmad=false).
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/ibl.h"

#define IBL_API extern "C" __attribute__((visibility("default")))

void ibl_set_error(const char *fmt, ...);

#define IBL_CUDA(call)                                                                          \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            ibl_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return IBL_E_CUDA;                                                                  \
        }                                                                                       \
    } while (0)

#define IBL_ARG(cond)                                                            \
    do {                                                                         \
        if (!(cond)) {                                                           \
            ibl_set_error("%s:%d bad argument: %s", __FILE__, __LINE__, #cond);  \
            return IBL_E_ARG;                                                    \
        }                                                                        \
    } while (0)

extern unsigned long long g_ibl_launches;
cudaError_t ibl_sync(cudaStream_t st);   // host wait on a stream under the ibl_set_host_wait policy (prof.cu)
#define IBL_LAUNCH_CHECK()                              \
    do {                                                \
        __sync_fetch_and_add(&g_ibl_launches, 1ull);    \
        IBL_CUDA(cudaGetLastError());                   \
    } while (0)

// optional per-kernel event timing (prof.cu); a scope brackets the launches that follow it in the same block
enum IblProfTag { IBL_PROF_BACKPROJECT = 0, IBL_PROF_RADIUS, IBL_PROF_KNN, IBL_PROF_NORMALS, IBL_PROF_SPFH, IBL_PROF_FPFH,
                  IBL_PROF_FEATURE_NN, IBL_PROF_RANSAC_CHECK, IBL_PROF_RANSAC_SCORE, IBL_PROF_NN_SCORE, IBL_PROF_ICP_REDUCE,
                  IBL_PROF_COLOR_GRAD, IBL_PROF_GRID_BUILD, IBL_PROF_SIMVOL, IBL_PROF_COSINE, IBL_PROF_VOXEL, IBL_PROF_MEAN,
                  IBL_PROF_MISC, IBL_PROF_FEATURE_GEMM /* nested inside FEATURE_NN: the tcgen05 kernel alone */, IBL_PROF_COUNT };
int ibl_prof_begin(int tag, cudaStream_t st, double bytes);
void ibl_prof_end(int slot, cudaStream_t st);
void ibl_prof_add_work(int tag, double bytes);
struct IblProfScope {
    int slot; cudaStream_t st;
    IblProfScope(int tag, cudaStream_t s, double bytes) : slot(ibl_prof_begin(tag, s, bytes)), st(s) {}
    ~IblProfScope() { ibl_prof_end(slot, st); }
};
#define IBL_PROF(tag, st, bytes) IblProfScope prof_scope__((tag), (st), (double)(bytes))

static inline cudaStream_t ibl_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// number of SMs on B200; grids of grid-stride kernels are sized in multiples of this
#define IBL_SMS 148

// stream-ordered scratch allocation (cudaMallocAsync pool: no device sync on free)
// The default cudaMallocAsync pool releases unused memory back to the driver at every synchronisation (release
// threshold 0), which makes the next allocation a cuMemMap.  Keep the memory in the pool instead.
static inline void ibl_pool_init() {
    static thread_local int done_dev = -1;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev == done_dev) return;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
        // Several frames run on several streams.  By default the pool may satisfy an allocation on stream B with memory
        // whose free is still pending on stream A by making B WAIT for A ("internal dependencies"): that silently chains
        // independent frames together (sporadic 20-50 % throughput drops at ~50 % GPU utilisation were traced to it).
        // Grow the pool instead; opportunistic reuse (the free has already completed) stays on.
        int off = getenv("IBL_POOL_INTERNAL_DEPS") ? 1 : 0;    // experiment switch; the product setting is 0
        cudaMemPoolSetAttribute(pool, cudaMemPoolReuseAllowInternalDependencies, &off);
    }
    done_dev = dev;
}
template <class T>
static inline cudaError_t ibl_alloc(T **p, size_t count, cudaStream_t st) {
    ibl_pool_init();
    if (count == 0) count = 1;
    return cudaMallocAsync(reinterpret_cast<void **>(p), count * sizeof(T), st);
}
static inline void ibl_free(void *p, cudaStream_t st) {
    if (p) cudaFreeAsync(p, st);
}

// ------------------------------------------------------------------------------------------------
// Exact nearest-neighbour candidate field (nnfield.cu): a sparse two-level voxel table over the target, dilated by
// max_d.  Voxel v (side c, centre c_v, half diagonal h) lists every target point p with
//     |p - c_v| <= min(D(v) + 2h, max_d + h),   D(v) = min_p |p - c_v|,
// which is a superset of the possible answers of "nearest neighbour within max_d" for every query inside v
// (NN distance <= D + h, so the NN lies within D + 2h of the centre).  A query is one table lookup + one short,
// convergent loop over the list: no ring walk, identical result (lowest original index wins exact ties).
struct NNFieldView {
    double ox, oy, oz, inv_c, max_d;
    int nvx, nvy, nvz, nbx, nby, nbz;
    const int *block_ptr;        // [nbx*nby*nbz] -> active block id or -1   (blocks are 8x8x8 voxels)
    const uint2 *entries;        // [nactive*512] (start, count) into cand
    const int *cand;             // positions into grid.sorted
};

struct ibl_grid {
    int64_t n;            // points
    double ox, oy, oz;    // origin (min corner)
    double cell, inv_cell;
    int nx, ny, nz;
    int64_t ncells;
    int64_t n_occupied;
    const double *pts;    // caller's points (original order), not owned
    int *cell_start;      // [ncells+1]
    double4 *sorted;      // cell-sorted (x, y, z, bits(orig index))
    int *order;           // sorted position -> original index
    uint8_t *ring;        // optional Chebyshev ring-distance field [ncells]
    int ring_cap;
    cudaStream_t stream;  // stream the grid was built on (frees are ordered on it)
    int segmented;        // sorted[].w carries (segment << 32 | original index): radius queries stay inside a segment
    // optional candidate field
    int has_field;
    NNFieldView field;
    int *f_block_ptr; uint2 *f_entries; int *f_cand;
    int64_t f_nactive, f_ncand;
};

// device view passed by value to kernels
struct GridView {
    double ox, oy, oz, cell, inv_cell;
    int nx, ny, nz;
    const int *cell_start;
    const double4 *sorted;
    const uint8_t *ring;
    int ring_cap;
    int segmented;
    int has_field;
    NNFieldView f;
};

void ibl_grid_free_field(ibl_grid *g);

static inline GridView ibl_view(const ibl_grid *g) {
    GridView v;
    v.ox = g->ox; v.oy = g->oy; v.oz = g->oz; v.cell = g->cell; v.inv_cell = g->inv_cell;
    v.nx = g->nx; v.ny = g->ny; v.nz = g->nz;
    v.cell_start = g->cell_start; v.sorted = g->sorted; v.ring = g->ring; v.ring_cap = g->ring_cap;
    v.segmented = g->segmented;
    v.has_field = g->has_field; v.f = g->field;
    return v;
}

// exact nanoflann L2: (dx*dx + dy*dy) + dz*dz, every op rounded once
__device__ __forceinline__ double ibl_d2(double ax, double ay, double az, double bx, double by, double bz) {
    double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by), dz = __dsub_rn(az, bz);
    return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

__device__ __forceinline__ int ibl_cell_coord(double p, double o, double inv, int /*n*/) {
    double c = floor((p - o) * inv);
    c = fmin(fmax(c, -1.0e9), 1.0e9);
    return (int)c;
}

__device__ __forceinline__ int ibl_clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Transform one point with the Open3D/Eigen (non-FMA) order ((T0*x + T1*y) + T2*z) + T3
__device__ __forceinline__ void ibl_xform(const double *T, double x, double y, double z, double &ox, double &oy,
                                          double &oz) {
    ox = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[0], x), __dmul_rn(T[1], y)), __dmul_rn(T[2], z)), T[3]);
    oy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[4], x), __dmul_rn(T[5], y)), __dmul_rn(T[6], z)), T[7]);
    oz = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[8], x), __dmul_rn(T[9], y)), __dmul_rn(T[10], z)), T[11]);
}

// Nearest neighbour of q within max_d (strict d2 < max_d2) by expanding Chebyshev rings.
// Returns original index or -1; best_d2 out.  Single-thread version (one query per thread).
__device__ __forceinline__ int ibl_nn1_field(const GridView &g, double qx, double qy, double qz, double max_d2,
                                             double &best_d2) {
    const NNFieldView &f = g.f;
    double best = max_d2;
    long long best_key = -1;
    double fx = floor((qx - f.ox) * f.inv_c), fy = floor((qy - f.oy) * f.inv_c), fz = floor((qz - f.oz) * f.inv_c);
    if (fx >= 0.0 && fy >= 0.0 && fz >= 0.0 && fx < (double)f.nvx && fy < (double)f.nvy && fz < (double)f.nvz) {
        int ix = (int)fx, iy = (int)fy, iz = (int)fz;
        int blk = f.block_ptr[((size_t)(iz >> 3) * f.nby + (iy >> 3)) * f.nbx + (ix >> 3)];
        if (blk >= 0) {
            uint2 e = f.entries[(size_t)blk * 512 + (((iz & 7) << 6) | ((iy & 7) << 3) | (ix & 7))];
            const int *c = f.cand + e.x;
            const unsigned n = e.y & 0x7fffffffu;          // top bit: NF_SURE (count-only queries)
            for (unsigned k = 0; k < n; ++k) {
                double4 t = g.sorted[c[k]];
                double d2 = ibl_d2(t.x, t.y, t.z, qx, qy, qz);
                long long oi = (long long)__double_as_longlong(t.w);
                if (d2 < best || (d2 == best && best_key >= 0 && oi < best_key)) { best = d2; best_key = oi; }
            }
        }
    }
    best_d2 = best;
    return (int)best_key;
}

// Is ANY target strictly within max_d of q?  (== "the nearest neighbour is within max_d", i.e. q is an inlier.)  With the
// candidate field the voxel's list holds every target that can be the nearest neighbour of a point of the voxel, so the
// scan stops at the first candidate inside max_d -- usually the first one -- instead of finishing the exact minimum.
// sure_ok: the query's max_d IS the max_d the field was built for (the SURE bit is only valid for that distance).
__device__ __forceinline__ bool ibl_any_within_field(const GridView &g, double qx, double qy, double qz, double max_d2,
                                                     bool sure_ok) {
    const NNFieldView &f = g.f;
    double fx = floor((qx - f.ox) * f.inv_c), fy = floor((qy - f.oy) * f.inv_c), fz = floor((qz - f.oz) * f.inv_c);
    if (!(fx >= 0.0 && fy >= 0.0 && fz >= 0.0 && fx < (double)f.nvx && fy < (double)f.nvy && fz < (double)f.nvz)) return false;
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    int blk = f.block_ptr[((size_t)(iz >> 3) * f.nby + (iy >> 3)) * f.nbx + (ix >> 3)];
    if (blk < 0) return false;
    uint2 e = f.entries[(size_t)blk * 512 + (((iz & 7) << 6) | ((iy & 7) << 3) | (ix & 7))];
    if ((e.y & 0x80000000u) && sure_ok) return true;       // whole voxel within max_d of a target (nnfield.cu: NF_SURE)
    const int *c = f.cand + e.x;
    const unsigned n = e.y & 0x7fffffffu;
    for (unsigned k = 0; k < n; ++k) {
        double4 t = g.sorted[c[k]];
        if (ibl_d2(t.x, t.y, t.z, qx, qy, qz) < max_d2) return true;
    }
    return false;
}

__device__ __forceinline__ int ibl_nn1_thread(const GridView &g, double qx, double qy, double qz, double max_d,
                                              double max_d2, double &best_d2) {
    if (g.has_field && max_d <= g.f.max_d) return ibl_nn1_field(g, qx, qy, qz, max_d2, best_d2);
    int cx = ibl_cell_coord(qx, g.ox, g.inv_cell, g.nx);
    int cy = ibl_cell_coord(qy, g.oy, g.inv_cell, g.ny);
    int cz = ibl_cell_coord(qz, g.oz, g.inv_cell, g.nz);
    double best = max_d2;
    long long best_key = -1;   // original index
    int s_max = (int)ceil(max_d * g.inv_cell) + 1;
    int s0 = 0;
    // quick reject / ring skip through the ring-distance field
    if (g.ring != nullptr) {
        int kx = ibl_clampi(cx, 0, g.nx - 1), ky = ibl_clampi(cy, 0, g.ny - 1), kz = ibl_clampi(cz, 0, g.nz - 1);
        int out = max(max(abs(cx - kx), abs(cy - ky)), abs(cz - kz));   // rings to reach the grid box
        int rf = g.ring[((size_t)kz * g.ny + ky) * g.nx + kx];          // rings from clamped cell to data
        // nearest occupied cell is at Chebyshev distance >= max(out, rf - out) from (cx,cy,cz)
        int lb = max(out, rf - out);
        if (rf >= g.ring_cap && out == 0) lb = max(lb, g.ring_cap);
        s0 = max(lb, 0);
    } else {
        int kx = ibl_clampi(cx, 0, g.nx - 1), ky = ibl_clampi(cy, 0, g.ny - 1), kz = ibl_clampi(cz, 0, g.nz - 1);
        s0 = max(max(abs(cx - kx), abs(cy - ky)), abs(cz - kz));
    }
    for (int s = s0; s <= s_max; ++s) {
        // everything in rings >= s is farther than (s-1)*cell (minus fp margin)
        if (s > 0) {
            double gr = ((double)(s - 1) - 1e-6) * g.cell;
            if (gr > 0 && gr * gr >= best) break;
        }
        int z0 = max(cz - s, 0), z1 = min(cz + s, g.nz - 1);
        int y0 = max(cy - s, 0), y1 = min(cy + s, g.ny - 1);
        for (int z = z0; z <= z1; ++z) {
            bool zedge = (z == cz - s) || (z == cz + s);
            for (int y = y0; y <= y1; ++y) {
                bool edge = zedge || (y == cy - s) || (y == cy + s);
                size_t rowbase = ((size_t)z * g.ny + y) * g.nx;
                int nseg = edge ? 1 : (s == 0 ? 1 : 2);
                for (int seg = 0; seg < nseg; ++seg) {
                    int xa, xb;
                    if (edge || s == 0) { xa = cx - s; xb = cx + s; }
                    else if (seg == 0) { xa = xb = cx - s; }
                    else { xa = xb = cx + s; }
                    xa = max(xa, 0); xb = min(xb, g.nx - 1);
                    if (xa > xb) continue;
                    int p0 = g.cell_start[rowbase + xa], p1 = g.cell_start[rowbase + xb + 1];
                    for (int p = p0; p < p1; ++p) {
                        double4 t = g.sorted[p];
                        double d2 = ibl_d2(t.x, t.y, t.z, qx, qy, qz);
                        long long oi = (long long)__double_as_longlong(t.w);
                        if (d2 < best || (d2 == best && best_key >= 0 && oi < best_key)) { best = d2; best_key = oi; }
                    }
                }
            }
        }
    }
    best_d2 = best;
    return (int)best_key;
}

// ------------------------------------------------------------------------------------------------
// bounding box: per-block partial min/max (fixed tree), finished on the host
static __global__ void bbox_kernel(const double *__restrict__ pts, int64_t n, double *__restrict__ part) {
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            double v = pts[3 * i + a];
            mn[a] = fmin(mn[a], v);
            mx[a] = fmax(mx[a], v);
        }
    }
    __shared__ double sm[6][32];
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o; o >>= 1) {
            mn[a] = fmin(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmax(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
        if (lane == 0) { sm[a][w] = mn[a]; sm[3 + a][w] = mx[a]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        int nw = blockDim.x >> 5;
        double v = sm[threadIdx.x][0];
        for (int k = 1; k < nw; ++k) v = threadIdx.x < 3 ? fmin(v, sm[threadIdx.x][k]) : fmax(v, sm[threadIdx.x][k]);
        part[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

