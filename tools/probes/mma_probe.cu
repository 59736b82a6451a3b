// mma_probe.cu -- how many clocks does one tcgen05.mma.kind::f16 128 x 128 x 16 take, as a function of the SHARED-MEMORY
// LAYOUT of its operands?
//
// ncu of the matcher's GEMM (csrc/match_tc.cu, profiles/r02_source_top.txt) shows the tensor pipe busy for ~100 % of the
// SM's active cycles while only ~28 % of the dense rate is delivered: every MMA takes ~220 clk instead of the 64 clk floor
// (M 128 x N 128 / 256).  The kernel's operands are no-swizzle K-major with LBO = 128 B (K-adjacent core matrices) and
// SBO = 1792 B (8-row groups) -- all sixteen 8-row groups of a K chunk start at the same bank.  This probe issues R rounds
// of 14 back-to-back MMAs (two accumulators x K = 112) from one thread per CTA, one CTA per SM, commits once and waits, for
// several (layout type, LBO, SBO, K-step) combinations.  Operand CONTENTS do not matter for timing (zeros).
//
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o mma_probe mma_probe.cu      run: ./mma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ull << 46) | ((uint64_t)(layout & 7) << 61);
}
#define IDESC ((1u << 4) | ((128u >> 3) << 17) | ((128u >> 4) << 24))

__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t ad, uint64_t bd, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
}

struct Variant { const char *name; uint32_t layout, lbo, sbo, kin, kstep_in, kstep_out, tile_bytes; };   // MMA k: (k / kin) * kstep_out + (k % kin) * kstep_in

__global__ void __launch_bounds__(64, 1) mma_kernel(Variant v, int rounds, int same_b, unsigned long long *clk) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (uint32_t i = threadIdx.x; i < 3 * v.tile_bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = slot;
    if (warp == 0 && lane == 0) {
        const uint32_t sA0 = smem_u32(smem), sA1 = sA0 + v.tile_bytes, sB = sA0 + 2 * v.tile_bytes;
        unsigned long long t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
            const uint32_t acc = tm + (r & 1) * 256;
            for (int h = 0; h < 2; ++h) {
#pragma unroll
                for (int k = 0; k < 7; ++k) {
                    const uint32_t ko = (k / v.kin) * v.kstep_out + (k % v.kin) * v.kstep_in;
                    uint64_t ad = smem_desc((h ? sA1 : sA0) + ko, v.lbo, v.sbo, v.layout);
                    uint64_t bd = smem_desc((same_b ? sB : (h ? sA0 : sA1)) + ko, v.lbo, v.sbo, v.layout);
                    mma_f16(acc + h * 128, ad, bd, k > 0);
                }
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                     ::"r"(smem_u32(&bar)) : "memory");
        unsigned long long t1 = clock64();
        clk[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
    }
}


// ---- part 2: the same MMA stream with the matcher's other traffic running beside it ------------------------------------------
// warp 0 lane 0: MMAs (kernel's layout).  copy != 0: warp 1 lane 0 streams 28 KB bulk copies global -> shared (a buffer the MMAs
// do not read), `copy` in flight, like the matcher's producer.  readers: warps 2 .. 2+readers-1 read accumulator columns back
// with tcgen05.ld.32x32b.x32 in a loop, like the epilogue.  Both run until the MMA thread has finished.
#define LD_X32(r, taddr)                                                                                                   \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                  \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),     \
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),         \
                   "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),        \
                   "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                      \
                 : "r"(taddr) : "memory")

__global__ void __launch_bounds__(320, 1) mma_mix_kernel(int rounds, int copy, int readers, const uint8_t *gsrc,
                                                         unsigned long long *clk, unsigned *sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar, cbar[4];
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t TB = 28672;
    for (uint32_t i = threadIdx.x; i < 7 * TB / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) {
        done = 0;
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        for (int k = 0; k < 4; ++k) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&cbar[k])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = slot;
    if (warp == 0) {
        if (lane == 0) {
            const uint32_t sA0 = smem_u32(smem), sA1 = sA0 + TB, sB = sA0 + 2 * TB;
            unsigned long long t0 = clock64();
            for (int r = 0; r < rounds; ++r) {
                const uint32_t acc = tm + (r & 1) * 256;
                for (int h = 0; h < 2; ++h) {
#pragma unroll
                    for (int k = 0; k < 7; ++k)
                        mma_f16(acc + h * 128, smem_desc((h ? sA1 : sA0) + k * 256, 128, 1792, 0), smem_desc(sB + k * 256, 128, 1792, 0), k > 0);
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
            asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                         ::"r"(smem_u32(&bar)) : "memory");
            clk[blockIdx.x] = clock64() - t0;
            done = 1;
        }
    } else if (warp == 1) {
        if (lane == 0 && copy > 0) {
            // `copy` stages of 28 KB; a stage is re-issued as soon as its previous copy has landed
            unsigned ph[4] = {0, 0, 0, 0};
            const uint8_t *src = gsrc + (size_t)blockIdx.x * 4 * TB;
            for (int k = 0; k < copy; ++k) {
                asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(&cbar[k])), "r"(TB) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(smem) + (3 + k) * TB), "l"(src + (size_t)k * TB), "r"(TB), "r"(smem_u32(&cbar[k])) : "memory");
            }
            unsigned long long n = 0;
            while (!done) {
                for (int k = 0; k < copy && !done; ++k) {
                    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                                 ::"r"(smem_u32(&cbar[k])), "r"(ph[k]) : "memory");
                    ph[k] ^= 1; ++n;
                    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(&cbar[k])), "r"(TB) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(smem_u32(smem) + (3 + k) * TB), "l"(src + (size_t)k * TB), "r"(TB), "r"(smem_u32(&cbar[k])) : "memory");
                }
            }
            for (int k = 0; k < copy; ++k)      // drain: no copy may be in flight when the CTA exits
                asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
                             ::"r"(smem_u32(&cbar[k])), "r"(ph[k]) : "memory");
            clk[gridDim.x + blockIdx.x] = n;
        }
    } else if (warp - 2 < readers) {
        const int w = warp - 2;
        const uint32_t lane_base = tm + ((uint32_t)((w & 3) * 32) << 16) + (uint32_t)((w >> 2) * 128);
        uint32_t a[32];
        unsigned acc = 0, n = 0;
        while (!done) {
            for (int c = 0; c < 4; ++c) {
                LD_X32(a, lane_base + (n & 1) * 256 + c * 32);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                acc ^= a[0] ^ a[31];
            }
            ++n;
        }
        if (acc == 0x12345678u) sink[0] = acc;
        if (lane == 0) clk[2 * gridDim.x + blockIdx.x * 8 + w] = n;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
    }
}

int main() {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    unsigned long long *clk;
    cudaMalloc(&clk, sizeof(*clk) * sms);
    const int rounds = 2000;
    // K-major tiles of 128 rows x 112 halves (K padded to 128 for the 64 B / 128 B swizzles, whose atoms are wider)
    Variant vs[] = {
        {"none  LBO 128   SBO 1792  (matcher today: 8-row groups 1792 B apart, K chunks adjacent)", 0, 128, 1792, 1, 0, 256, 28672},
        {"none  LBO 2048  SBO 128   (8-row groups adjacent, K chunks 2 KB apart)", 0, 2048, 128, 1, 0, 4096, 28672},
        {"none  LBO 128   SBO 1920  (8-row groups 1792 + 128 B apart: rotates the banks)", 0, 128, 1920, 1, 0, 256, 30720},
        {"none  LBO 2064  SBO 128   (K chunks 2 KB + 16 B apart)", 0, 2064, 128, 1, 0, 4128, 28896},
        {"32B   swizzle   SBO 256   (atom = 8 rows x 32 B, K chunk = one atom column, 4 KB apart)", 6, 4096, 256, 1, 0, 4096, 28672},
        {"64B   swizzle   SBO 512   (atom = 8 rows x 64 B; K step 32 B inside the atom)", 4, 8192, 512, 2, 32, 8192, 32768},
        {"128B  swizzle   SBO 1024  (atom = 8 rows x 128 B; K step 32 B inside the atom)", 2, 16384, 1024, 4, 32, 16384, 32768},
    };
    for (int same_b = 1; same_b >= 0; --same_b) {
        for (const Variant &v : vs) {
            size_t smem = 3 * (size_t)v.tile_bytes + 1024;
            cudaFuncSetAttribute(mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            mma_kernel<<<sms, 64, smem>>>(v, 10, same_b, clk);
            mma_kernel<<<sms, 64, smem>>>(v, rounds, same_b, clk);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("{\"variant\": \"%s\", \"error\": \"%s\"}\n", v.name, cudaGetErrorString(e)); return 1; }
            unsigned long long h[256], mx = 0, sum = 0;
            cudaMemcpy(h, clk, sizeof(*clk) * sms, cudaMemcpyDeviceToHost);
            for (int i = 0; i < sms; ++i) { sum += h[i]; if (h[i] > mx) mx = h[i]; }
            double per = (double)sum / sms / ((double)rounds * 14);
            printf("{\"variant\": \"%s\", \"b_operand\": \"%s\", \"rounds\": %d, \"clk_per_mma_avg\": %.1f, \"clk_per_mma_max\": %.1f, "
                   "\"frac_of_64clk_floor\": %.3f}\n", v.name, same_b ? "one tile for both accumulators" : "two tiles",
                   rounds, per, (double)mx / ((double)rounds * 14), 64.0 / per);
            fflush(stdout);
        }
    }
    // part 2
    {
        const int rounds2 = 2000;
        uint8_t *gsrc; unsigned *sink; unsigned long long *clk2;
        cudaMalloc(&gsrc, (size_t)sms * 4 * 28672); cudaMemset(gsrc, 0, (size_t)sms * 4 * 28672);
        cudaMalloc(&sink, 4); cudaMalloc(&clk2, sizeof(*clk2) * sms * 10);
        size_t smem = 7 * 28672 + 1024;
        cudaFuncSetAttribute(mma_mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        const int cfgs[][2] = {{0, 0}, {4, 0}, {0, 4}, {0, 8}, {4, 8}, {2, 8}, {1, 8}};
        for (auto &c : cfgs) {
            cudaMemset(clk2, 0, sizeof(*clk2) * sms * 10);
            mma_mix_kernel<<<sms, 320, smem>>>(rounds2, c[0], c[1], gsrc, clk2, sink);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("{\"mix\": [%d, %d], \"error\": \"%s\"}\n", c[0], c[1], cudaGetErrorString(e)); return 1; }
            static unsigned long long h[148 * 10 + 64];
            cudaMemcpy(h, clk2, sizeof(*clk2) * sms * 10, cudaMemcpyDeviceToHost);
            double sum = 0, copies = 0, reads = 0;
            for (int i = 0; i < sms; ++i) { sum += h[i]; copies += h[sms + i]; }
            for (int i = 0; i < sms * 8; ++i) reads += h[2 * sms + i];
            double per = sum / sms / (rounds2 * 14.0);
            double clk_total = sum / sms;
            printf("{\"mix\": \"MMAs + %d bulk copies of 28 KB in flight + %d TMEM-reading warps\", \"clk_per_mma\": %.1f, "
                   "\"copy_bytes_per_clk_per_sm\": %.1f, \"tmem_read_bytes_per_clk_per_sm\": %.1f}\n", c[0], c[1], per,
                   copies / sms * 28672.0 / clk_total, reads / sms * (4 * 32 * 32 * 4.0) / clk_total);
            fflush(stdout);
        }
    }
    return 0;
}
