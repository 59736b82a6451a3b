// nnfield.cu -- builds the exact nearest-neighbour candidate field of a grid (see NNFieldView in ibl_common.cuh).
// It replaces the KD-tree that Open3D builds for GetRegistrationResultAndCorrespondences
// (Registration.cpp; reference call sites utils/fpfh_register.py:110-141) for the query pattern of RANSAC
// validation and ICP: millions of "nearest target point within max_d" queries against ONE fixed target.
//
// Build (all on the device, two host read-backs for allocation sizes):
//   mark    : every point flags the 8x8x8-voxel blocks its reach sphere (radius R = max_d + h) touches
//   pass A  : every (point, voxel centre within R) pair does atomicMin on the voxel's squared centre distance
//   pass B  : the same pairs count the candidates  |p - c_v| <= min(D + 2h, max_d + h)
//   scan    : exclusive sum of the counts -> (start, count) entries
//   pass C  : the same pairs write the candidate positions (order inside a list is irrelevant: the query
//             compares (d2, original index), so the answer does not depend on it)
#include <cub/cub.cuh>

#include "ibl_common.cuh"

struct FieldGeom {
    double ox, oy, oz, c, inv_c, R, R2, h, max_d;
    int nvx, nvy, nvz, nbx, nby, nbz;
};

__device__ __forceinline__ int fg_vox(double p, double o, double inv) { return (int)floor((p - o) * inv); }

__global__ void nf_mark_kernel(const double4 *__restrict__ sorted, int64_t n, FieldGeom G, int *__restrict__ flag) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double4 p = sorted[i];
    int x0 = max(fg_vox(p.x - G.R, G.ox, G.inv_c), 0) >> 3, x1 = min(fg_vox(p.x + G.R, G.ox, G.inv_c), G.nvx - 1) >> 3;
    int y0 = max(fg_vox(p.y - G.R, G.oy, G.inv_c), 0) >> 3, y1 = min(fg_vox(p.y + G.R, G.oy, G.inv_c), G.nvy - 1) >> 3;
    int z0 = max(fg_vox(p.z - G.R, G.oz, G.inv_c), 0) >> 3, z1 = min(fg_vox(p.z + G.R, G.oz, G.inv_c), G.nvz - 1) >> 3;
    for (int z = z0; z <= z1; ++z)
        for (int y = y0; y <= y1; ++y)
            for (int x = x0; x <= x1; ++x) flag[((size_t)z * G.nby + y) * G.nbx + x] = 1;
}

__global__ void nf_blockptr_kernel(const int *__restrict__ flag, const int *__restrict__ scan, int64_t nb,
                                   int *__restrict__ block_ptr) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < nb) block_ptr[i] = flag[i] ? scan[i] : -1;
}

// MODE 0: atomicMin of squared centre distance; 1: count candidates; 2: write candidates
template <int MODE>
__global__ void __launch_bounds__(256) nf_splat_kernel(const double4 *__restrict__ sorted, int64_t n, FieldGeom G,
                                                       const int *__restrict__ block_ptr,
                                                       unsigned long long *__restrict__ dmin2,
                                                       unsigned int *__restrict__ cnt, const uint2 *__restrict__ entries,
                                                       int *__restrict__ cand) {
    int lane = threadIdx.x & 31;
    int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t i = warp; i < n; i += nwarps) {
        double4 p = sorted[i];
        int x0 = max(fg_vox(p.x - G.R, G.ox, G.inv_c), 0), x1 = min(fg_vox(p.x + G.R, G.ox, G.inv_c), G.nvx - 1);
        int y0 = max(fg_vox(p.y - G.R, G.oy, G.inv_c), 0), y1 = min(fg_vox(p.y + G.R, G.oy, G.inv_c), G.nvy - 1);
        int z0 = max(fg_vox(p.z - G.R, G.oz, G.inv_c), 0), z1 = min(fg_vox(p.z + G.R, G.oz, G.inv_c), G.nvz - 1);
        int wx = x1 - x0 + 1, wy = y1 - y0 + 1, wz = z1 - z0 + 1;
        if (wx <= 0 || wy <= 0 || wz <= 0) continue;
        int tot = wx * wy * wz;
        for (int t = lane; t < tot; t += 32) {
            int ix = x0 + t % wx, iy = y0 + (t / wx) % wy, iz = z0 + t / (wx * wy);
            double cx = G.ox + ((double)ix + 0.5) * G.c, cy = G.oy + ((double)iy + 0.5) * G.c, cz = G.oz + ((double)iz + 0.5) * G.c;
            double dx = p.x - cx, dy = p.y - cy, dz = p.z - cz;
            double d2 = dx * dx + dy * dy + dz * dz;
            if (d2 > G.R2) continue;
            int blk = block_ptr[((size_t)(iz >> 3) * G.nby + (iy >> 3)) * G.nbx + (ix >> 3)];
            if (blk < 0) continue;     // cannot happen (mark covers the same voxels); defensive
            size_t v = (size_t)blk * 512 + (((iz & 7) << 6) | ((iy & 7) << 3) | (ix & 7));
            if (MODE == 0) {
                atomicMin(&dmin2[v], (unsigned long long)__double_as_longlong(d2));
            } else {
                double D = sqrt(__longlong_as_double((long long)dmin2[v]));
                double lim = fmin(D + 2.0 * G.h, G.max_d + G.h);
                if (d2 <= lim * lim) {
                    if (MODE == 1) atomicAdd(&cnt[v], 1u);
                    else cand[entries[v].x + atomicAdd(&cnt[v], 1u)] = (int)i;
                }
            }
        }
    }
}

// entries[v] = (start, count | SURE bit).  SURE: every point of the voxel is closer than max_d to the target nearest to the
// voxel centre (|q - c| <= h, |p* - c| = D, so |q - p*| <= D + h), with a 1e-9 relative margin that dwarfs the rounding of
// the query's squared distance -- an inlier COUNT needs no candidate evaluation there (ibl_any_within_field).
#define NF_SURE 0x80000000u
__global__ void nf_entries_kernel(const unsigned int *__restrict__ cnt, const unsigned int *__restrict__ start, int64_t nv,
                                  const unsigned long long *__restrict__ dmin2, double h, double max_d,
                                  uint2 *__restrict__ entries) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nv) return;
    unsigned c = cnt[i];
    if (c) {
        double D = sqrt(__longlong_as_double((long long)dmin2[i]));
        if ((D + h) * (1.0 + 1e-9) < max_d) c |= NF_SURE;
    }
    entries[i] = make_uint2(start[i], c);
}

static void nf_free(ibl_grid *g) {
    cudaStream_t st = g->stream;
    ibl_free(g->f_block_ptr, st); ibl_free(g->f_entries, st); ibl_free(g->f_cand, st);
    g->f_block_ptr = nullptr; g->f_entries = nullptr; g->f_cand = nullptr;
    g->has_field = 0; g->f_nactive = 0; g->f_ncand = 0;
}
void ibl_grid_free_field(ibl_grid *g) { nf_free(g); }

IBL_API int ibl_grid_build_nnfield(ibl_grid_t *g, double max_d, double voxel, void *stream) {
    IBL_ARG(g != nullptr && max_d > 0 && voxel >= 0);
    cudaStream_t st = ibl_stream(stream);
    if (g->has_field) nf_free(g);
    if (g->n == 0) return IBL_OK;
    FieldGeom G;
    double c = voxel > 0 ? voxel : max_d / 4.0;
    // bounding box of the points = grid origin .. origin + dims*cell (a superset is fine)
    double ext[3] = {g->nx * g->cell, g->ny * g->cell, g->nz * g->cell};
    for (;;) {
        G.c = c; G.inv_c = 1.0 / c;
        G.h = c * 0.8660254037844387 * 1.000001 + 1e-9;      // half diagonal + margin for query-side rounding
        G.max_d = max_d; G.R = max_d + G.h; G.R2 = G.R * G.R;
        double pad = G.R + c;
        G.ox = g->ox - pad; G.oy = g->oy - pad; G.oz = g->oz - pad;
        double nvx = ceil((ext[0] + 2 * pad) / c), nvy = ceil((ext[1] + 2 * pad) / c), nvz = ceil((ext[2] + 2 * pad) / c);
        double nblocks = ceil(nvx / 8) * ceil(nvy / 8) * ceil(nvz / 8);
        if (nblocks <= (double)(1 << 26) && nvx < 2e9 && nvy < 2e9 && nvz < 2e9) {
            G.nvx = (int)nvx; G.nvy = (int)nvy; G.nvz = (int)nvz;
            G.nbx = (G.nvx + 7) / 8; G.nby = (G.nvy + 7) / 8; G.nbz = (G.nvz + 7) / 8;
            break;
        }
        c *= 1.25;
    }
    int64_t nb = (int64_t)G.nbx * G.nby * G.nbz;
    IBL_PROF(IBL_PROF_GRID_BUILD, st, 0.0);
    int *flag = nullptr, *scan = nullptr;
    IBL_CUDA(ibl_alloc(&flag, (size_t)nb + 1, st));
    IBL_CUDA(ibl_alloc(&scan, (size_t)nb + 1, st));
    IBL_CUDA(ibl_alloc(&g->f_block_ptr, (size_t)nb, st));
    IBL_CUDA(cudaMemsetAsync(flag, 0, sizeof(int) * ((size_t)nb + 1), st));
    int nbp = (int)((g->n + 255) / 256);
    nf_mark_kernel<<<nbp, 256, 0, st>>>(g->sorted, g->n, G, flag);
    IBL_LAUNCH_CHECK();
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, flag, scan, (int)(nb + 1), st);
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tb ? tb : 1, st));
    IBL_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, flag, scan, (int)(nb + 1), st));
    nf_blockptr_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, st>>>(flag, scan, nb, g->f_block_ptr);
    IBL_LAUNCH_CHECK();
    int nactive = 0;
    IBL_CUDA(cudaMemcpyAsync(&nactive, scan + nb, sizeof(int), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(tmp, st); ibl_free(flag, st); ibl_free(scan, st);
    int64_t nv = (int64_t)nactive * 512;
    if (nv >= (1ll << 31)) { nf_free(g); ibl_set_error("nnfield: too many active voxels"); return IBL_E_RANGE; }
    unsigned long long *dmin2 = nullptr;
    unsigned int *cnt = nullptr, *start = nullptr;
    IBL_CUDA(ibl_alloc(&dmin2, (size_t)nv, st));
    IBL_CUDA(ibl_alloc(&cnt, (size_t)nv + 1, st));
    IBL_CUDA(ibl_alloc(&start, (size_t)nv + 1, st));
    IBL_CUDA(ibl_alloc(&g->f_entries, (size_t)nv, st));
    IBL_CUDA(cudaMemsetAsync(dmin2, 0x7f, sizeof(unsigned long long) * (size_t)nv, st));   // 0x7f7f.. = huge double
    IBL_CUDA(cudaMemsetAsync(cnt, 0, sizeof(unsigned int) * ((size_t)nv + 1), st));
    int64_t warps = g->n;
    int blocks = (int)((warps * 32 + 255) / 256);
    if (blocks > IBL_SMS * 16) blocks = IBL_SMS * 16;
    nf_splat_kernel<0><<<blocks, 256, 0, st>>>(g->sorted, g->n, G, g->f_block_ptr, dmin2, cnt, nullptr, nullptr);
    IBL_LAUNCH_CHECK();
    nf_splat_kernel<1><<<blocks, 256, 0, st>>>(g->sorted, g->n, G, g->f_block_ptr, dmin2, cnt, nullptr, nullptr);
    IBL_LAUNCH_CHECK();
    tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt, start, (int)(nv + 1), st);
    IBL_CUDA(cudaMallocAsync(&tmp, tb ? tb : 1, st));
    IBL_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tb, cnt, start, (int)(nv + 1), st));
    nf_entries_kernel<<<(unsigned)((nv + 255) / 256), 256, 0, st>>>(cnt, start, nv, dmin2, G.h, max_d, g->f_entries);
    IBL_LAUNCH_CHECK();
    unsigned int total = 0;
    IBL_CUDA(cudaMemcpyAsync(&total, start + nv, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    IBL_CUDA(ibl_sync(st));
    ibl_free(tmp, st);
    if (total >= (1u << 31)) { ibl_free(dmin2, st); ibl_free(cnt, st); ibl_free(start, st); nf_free(g); ibl_set_error("nnfield: too many candidates"); return IBL_E_RANGE; }
    IBL_CUDA(ibl_alloc(&g->f_cand, (size_t)total, st));
    IBL_CUDA(cudaMemsetAsync(cnt, 0, sizeof(unsigned int) * ((size_t)nv + 1), st));
    nf_splat_kernel<2><<<blocks, 256, 0, st>>>(g->sorted, g->n, G, g->f_block_ptr, dmin2, cnt, g->f_entries, g->f_cand);
    IBL_LAUNCH_CHECK();
    ibl_free(dmin2, st); ibl_free(cnt, st); ibl_free(start, st);
    NNFieldView &f = g->field;
    f.ox = G.ox; f.oy = G.oy; f.oz = G.oz; f.inv_c = G.inv_c; f.max_d = max_d;
    f.nvx = G.nvx; f.nvy = G.nvy; f.nvz = G.nvz; f.nbx = G.nbx; f.nby = G.nby; f.nbz = G.nbz;
    f.block_ptr = g->f_block_ptr; f.entries = g->f_entries; f.cand = g->f_cand;
    g->f_nactive = nactive; g->f_ncand = total;
    g->has_field = 1;
    return IBL_OK;
}

IBL_API int64_t ibl_grid_nnfield_bytes(const ibl_grid_t *g) {
    if (!g || !g->has_field) return 0;
    return (int64_t)g->field.nbx * g->field.nby * g->field.nbz * 4 + g->f_nactive * 512 * 8 + g->f_ncand * 4;
}
