// grid.cu -- cell-sorted uniform grid + exact radius / k-NN / 1-NN queries (replaces the
// KDTreeFlann constructions and searches Open3D performs at every reference call site:
// utils/depth_utils.py:41,87, utils/fpfh_register.py:91-97,110-141,146-148).
#include <cub/cub.cuh>
#include <stdarg.h>

#include "ibl_common.cuh"

static thread_local char g_err[512] = "";
void ibl_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
IBL_API const char *ibl_last_error_string(void) { return g_err; }
IBL_API int ibl_version(void) { return 100; }

// ------------------------------------------------------------------------------------------------
__global__ void cellid_kernel(const double *__restrict__ pts, int64_t n, double ox, double oy, double oz, double inv,
                              int nx, int ny, int nz, uint32_t *__restrict__ cid, int *__restrict__ ord) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int cx = ibl_clampi(ibl_cell_coord(pts[3 * i], ox, inv, nx), 0, nx - 1);
    int cy = ibl_clampi(ibl_cell_coord(pts[3 * i + 1], oy, inv, ny), 0, ny - 1);
    int cz = ibl_clampi(ibl_cell_coord(pts[3 * i + 2], oz, inv, nz), 0, nz - 1);
    cid[i] = (uint32_t)(((size_t)cz * ny + cy) * nx + cx);
    ord[i] = (int)i;
}

__global__ void gather_sorted_kernel(const double *__restrict__ pts, const int *__restrict__ ord, int64_t n,
                                     const uint32_t *__restrict__ cid_sorted, double4 *__restrict__ sorted,
                                     int *__restrict__ cell_count, const int32_t *__restrict__ seg) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int o = ord[i];
    double4 v;
    v.x = pts[3 * (size_t)o]; v.y = pts[3 * (size_t)o + 1]; v.z = pts[3 * (size_t)o + 2];
    v.w = __longlong_as_double(seg ? (((long long)seg[o] << 32) | (long long)(unsigned)o) : (long long)o);
    sorted[i] = v;
    atomicAdd(&cell_count[cid_sorted[i]], 1);   // integer histogram: order independent
}

__global__ void count_occupied_kernel(const int *__restrict__ cell_start, int64_t ncells, unsigned long long *out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int occ = (i < ncells) && (cell_start[i + 1] > cell_start[i]);
    unsigned b = __ballot_sync(0xffffffffu, occ);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(out, (unsigned long long)__popc(b));
}

static int grid_create_impl(const double *pts, int64_t n, double cell, int64_t max_cells, const int32_t *seg, ibl_grid_t **out,
                            void *stream);
IBL_API int ibl_grid_create(const double *pts, int64_t n, double cell, int64_t max_cells, ibl_grid_t **out,
                            void *stream) {
    return grid_create_impl(pts, n, cell, max_cells, nullptr, out, stream);
}
// One grid over the concatenation of several clouds; seg int32[n] (device) names the cloud of every point.  Radius
// queries on such a grid (ibl_radius_count / ibl_radius_outlier_mask) only count neighbours of the query's own cloud, so
// one build and one query replace a build and a query per cloud (the per-mask outlier filters of a frame,
// utils/depth_utils.py:85-88 and object_memory.py:992-998).  Other queries are not defined on a segmented grid.
IBL_API int ibl_grid_create_segmented(const double *pts, int64_t n, const int32_t *seg, double cell, int64_t max_cells,
                                      ibl_grid_t **out, void *stream) {
    IBL_ARG(seg != nullptr || n == 0);
    return grid_create_impl(pts, n, cell, max_cells, seg, out, stream);
}
static int grid_create_impl(const double *pts, int64_t n, double cell, int64_t max_cells, const int32_t *seg, ibl_grid_t **out,
                            void *stream) {
    IBL_ARG(out != nullptr && n >= 0 && cell > 0 && n < (1ll << 31));
    cudaStream_t st = ibl_stream(stream);
    if (max_cells <= 0) max_cells = 1ll << 24;
    ibl_grid *g = new ibl_grid();
    memset(g, 0, sizeof(*g));
    g->n = n; g->pts = pts; g->stream = st; g->segmented = seg != nullptr;
    double mn[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
    if (n > 0) {
        int nb = (int)((n + 255) / 256);
        if (nb > 4 * IBL_SMS) nb = 4 * IBL_SMS;
        double *part = nullptr;
        IBL_CUDA(ibl_alloc(&part, (size_t)nb * 6, st));
        bbox_kernel<<<nb, 256, 0, st>>>(pts, n, part);
        IBL_LAUNCH_CHECK();
        double *hp = (double *)malloc(sizeof(double) * nb * 6);
        IBL_CUDA(cudaMemcpyAsync(hp, part, sizeof(double) * nb * 6, cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));
        for (int a = 0; a < 3; ++a) { mn[a] = INFINITY; mx[a] = -INFINITY; }
        for (int b = 0; b < nb; ++b)
            for (int a = 0; a < 3; ++a) { mn[a] = fmin(mn[a], hp[b * 6 + a]); mx[a] = fmax(mx[a], hp[b * 6 + 3 + a]); }
        free(hp);
        ibl_free(part, st);
        for (int a = 0; a < 3; ++a)
            if (!isfinite(mn[a]) || !isfinite(mx[a])) { delete g; ibl_set_error("non-finite point coordinates"); return IBL_E_RANGE; }
    }
    double c = cell;
    int64_t d[3];
    for (;;) {
        for (int a = 0; a < 3; ++a) d[a] = (int64_t)floor((mx[a] - mn[a]) / c) + 1;
        if (d[0] * d[1] * d[2] <= max_cells && d[0] < (1 << 20) && d[1] < (1 << 20) && d[2] < (1 << 20)) break;
        c *= 1.25;
    }
    g->cell = c; g->inv_cell = 1.0 / c;
    g->ox = mn[0]; g->oy = mn[1]; g->oz = mn[2];
    g->nx = (int)d[0]; g->ny = (int)d[1]; g->nz = (int)d[2];
    g->ncells = d[0] * d[1] * d[2];
    IBL_CUDA(ibl_alloc(&g->cell_start, (size_t)g->ncells + 1, st));
    IBL_CUDA(ibl_alloc(&g->sorted, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&g->order, (size_t)n, st));
    IBL_CUDA(cudaMemsetAsync(g->cell_start, 0, sizeof(int) * ((size_t)g->ncells + 1), st));
    if (n > 0) {
        uint32_t *cid = nullptr, *cid2 = nullptr;
        int *ord = nullptr;
        IBL_CUDA(ibl_alloc(&cid, (size_t)n, st));
        IBL_CUDA(ibl_alloc(&cid2, (size_t)n, st));
        IBL_CUDA(ibl_alloc(&ord, (size_t)n, st));
        int nb = (int)((n + 255) / 256);
        IBL_PROF(IBL_PROF_GRID_BUILD, st, 72.0 * n + 8.0 * g->ncells);
        cellid_kernel<<<nb, 256, 0, st>>>(pts, n, g->ox, g->oy, g->oz, g->inv_cell, g->nx, g->ny, g->nz, cid, ord);
        IBL_LAUNCH_CHECK();
        int bits = 1;
        while ((1ll << bits) < g->ncells) ++bits;
        size_t tmp_bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, cid, cid2, ord, g->order, (int)n, 0, bits, st);
        void *tmp = nullptr;
        IBL_CUDA(cudaMallocAsync(&tmp, tmp_bytes ? tmp_bytes : 1, st));
        IBL_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, cid, cid2, ord, g->order, (int)n, 0, bits, st));
        // histogram into cell_start[c] then exclusive scan in place
        gather_sorted_kernel<<<nb, 256, 0, st>>>(pts, g->order, n, cid2, g->sorted, g->cell_start, seg);
        IBL_LAUNCH_CHECK();
        ibl_free(tmp, st);
        size_t tmp2 = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmp2, g->cell_start, g->cell_start, (int)(g->ncells + 1), st);
        IBL_CUDA(cudaMallocAsync(&tmp, tmp2 ? tmp2 : 1, st));
        IBL_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp2, g->cell_start, g->cell_start, (int)(g->ncells + 1), st));
        ibl_free(tmp, st);
        ibl_free(cid, st); ibl_free(cid2, st); ibl_free(ord, st);
        g->n_occupied = -1;                                // counted on demand (ibl_grid_occupied): saves a host round trip per build
    }
    *out = g;
    return IBL_OK;
}

IBL_API void ibl_grid_destroy(ibl_grid_t *g) {
    if (!g) return;
    // stream-ordered frees on the creating stream: queries already enqueued there finish first.  (cudaFree on a
    // cudaMallocAsync allocation does NOT synchronise, so it must not be used while kernels may be in flight.)
    cudaStream_t st = g->stream;
    ibl_grid_free_field(g);
    ibl_free(g->cell_start, st);
    ibl_free(g->sorted, st);
    ibl_free(g->order, st);
    ibl_free(g->ring, st);
    delete g;
}
IBL_API double ibl_grid_cell(const ibl_grid_t *g) { return g ? g->cell : 0.0; }
// Number of non-empty cells.  Only the surface-spacing probe asks for it, so it is counted here, on the grid's stream,
// the first time it is asked for (one small kernel and one host round trip) instead of in every ibl_grid_create.
IBL_API int64_t ibl_grid_occupied(const ibl_grid_t *gc) {
    ibl_grid *g = const_cast<ibl_grid *>(gc);
    if (!g) return 0;
    if (g->n_occupied >= 0) return g->n_occupied;
    cudaStream_t st = g->stream;
    unsigned long long *occ = nullptr, hocc = 0;
    if (ibl_alloc(&occ, 1, st) != cudaSuccess) return -1;
    cudaMemsetAsync(occ, 0, 8, st);
    int64_t nbc = (g->ncells + 255) / 256;
    count_occupied_kernel<<<(unsigned)nbc, 256, 0, st>>>(g->cell_start, g->ncells, occ);
    __sync_fetch_and_add(&g_ibl_launches, 1ull);
    cudaMemcpyAsync(&hocc, occ, 8, cudaMemcpyDeviceToHost, st);
    ibl_sync(st);
    ibl_free(occ, st);
    g->n_occupied = (int64_t)hocc;
    return g->n_occupied;
}

// ------------------------------------------------------------------------------------------------
// Chebyshev ring-distance field by level-set dilation
__global__ void ring_init_kernel(const int *__restrict__ cell_start, int64_t ncells, uint8_t *ring, uint8_t cap) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < ncells) ring[i] = (cell_start[i + 1] > cell_start[i]) ? 0 : cap;
}
__global__ void ring_step_kernel(uint8_t *ring, int nx, int ny, int nz, uint8_t level, uint8_t cap) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int64_t ncells = (int64_t)nx * ny * nz;
    if (i >= ncells || ring[i] != cap) return;
    int x = (int)(i % nx), y = (int)((i / nx) % ny), z = (int)(i / ((int64_t)nx * ny));
    uint8_t want = level - 1;
    for (int dz = -1; dz <= 1; ++dz) {
        int zz = z + dz; if (zz < 0 || zz >= nz) continue;
        for (int dy = -1; dy <= 1; ++dy) {
            int yy = y + dy; if (yy < 0 || yy >= ny) continue;
            for (int dx = -1; dx <= 1; ++dx) {
                int xx = x + dx; if (xx < 0 || xx >= nx) continue;
                if (ring[((size_t)zz * ny + yy) * nx + xx] == want) { ring[i] = level; return; }
            }
        }
    }
}
IBL_API int ibl_grid_build_ringfield(ibl_grid_t *g, int max_ring, void *stream) {
    IBL_ARG(g != nullptr && max_ring >= 1 && max_ring <= 255);
    cudaStream_t st = ibl_stream(stream);
    if (g->ring) { ibl_free(g->ring, st); g->ring = nullptr; }
    IBL_CUDA(ibl_alloc(&g->ring, (size_t)g->ncells, st));
    g->ring_cap = max_ring;
    int64_t nb = (g->ncells + 255) / 256;
    ring_init_kernel<<<(unsigned)nb, 256, 0, st>>>(g->cell_start, g->ncells, g->ring, (uint8_t)max_ring);
    IBL_LAUNCH_CHECK();
    // level L cells only read level L-1 values, which are final before the launch: race free
    for (int L = 1; L < max_ring; ++L) {
        ring_step_kernel<<<(unsigned)nb, 256, 0, st>>>(g->ring, g->nx, g->ny, g->nz, (uint8_t)L, (uint8_t)max_ring);
        IBL_LAUNCH_CHECK();
    }
    return IBL_OK;
}
