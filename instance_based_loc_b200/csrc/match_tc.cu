// match_tc.cu -- tensor-core (tcgen05, kind::f16) filter for the exact 33-d FPFH nearest-neighbour matcher
// (Open3D RegistrationRANSACBasedOnFeatureMatching front half, utils/fpfh_register.py:110-112).
//
// The answer is still the f64 brute-force argmin with lowest-index ties (match.cu: feat_exact); the tensor cores only
// decide which pairs need that exact evaluation.  For a query a and a target b the kernel computes the score
//     s'(a,b) = a32.b32 - |b32|^2 / 2            (argmax_b s' == argmin_b |a32 - b32|^2)
// as ONE fp16 GEMM of depth 112 (kind::f16, fp32 accumulate) using a 3-way split of the centred, power-of-two-scaled fp32
// row  x32 = xh + xl  (xh = RN_fp16(x32), xl = RN_fp16(x32 - xh); 11-bit significands, |x32| <= 1024):
//     query row  : [ ah(33) | al(33) | ah(33) | 1024  1024  1024 | 0 .. ]
//     target row : [ bh(33) | bh(33) | bl(33) |  -h0   -h1   -h2 | 0 .. ]   h/1024 = |b32|^2/2048 = h0 + h1 + h2 (fp16 pieces)
// so every product is exact in fp32 and only the al.bl term is dropped.
// ONE pass over the score matrix: every epilogue thread owns a query row, keeps the running maximum m of its scores and
// records every target b with s'(q,b) >= m - E(q) at the moment it streams by (m only grows, so the recorded set is a
// superset of the final band [smax - E, smax]).  E is a rigorous bound on |(|a32|^2 - 2 s') - d_exact| (see feat_tc_err),
// so the exact nearest neighbour, and every exact tie with it, is inside the band.  feat_tc_eval_kernel then puts the
// recorded candidates through a second, ~10^3 times tighter filter on fp32 DIFFERENCES (feat_tc_err32) and evaluates the
// few that survive exactly in f64.
//
// Kernel shape (one CTA per SM, 192 threads): warps 0-3 = epilogue (TMEM lane quarter = warp id), warp 4 = bulk-copy
// producer, warp 5 = MMA issuer.  Operand tiles are pre-packed in global memory in the exact shared-memory image of the
// no-swizzle K-major UMMA layout (8x16B core matrices, LBO = 128 B along K, SBO = 1792 B along rows), so one
// cp.async.bulk brings in a whole 128 x 112 tile (28 KB).  The query tile stays resident; target tiles stream through a
// 4-stage ring; two 128-column TMEM accumulators overlap the MMA of tile t+1 with the epilogue of tile t, whose TMEM reads
// are software-pipelined over two register buffers.  The kernel is bound by reading every fp32 score out of TMEM once
// (64 B/clk/SM: 1024 clk per 128 x 128 tile against 448 clk of MMAs).
#include <algorithm>
#include <cmath>

#include <cuda_fp16.h>
#include <cub/cub.cuh>

#include "ibl_common.cuh"

#define TC_M 128                    // rows per tile (queries and targets)
#define TC_K 112                    // GEMM depth: 3 x 33 + 3 + 10 pad  (7 MMAs of K = 16, fp16 operands)
#define TC_KCH (TC_K / 8)           // 16-byte chunks (8 halves) per row
#define TC_LBO 128                  // bytes between K-adjacent core matrices
#define TC_SBO (TC_KCH * 128)       // bytes between 8-row groups
#define TC_TILE_BYTES (16 * TC_SBO) // 28,672
#define TC_STAGES 4
#define TC_RANGE 1024.0             // rows are scaled by a power of two so that |value| <= TC_RANGE (fp16 range / precision)
#define TC_QT 2                      // query tiles per CTA: 2 x 4 epilogue warps, two warps per scheduler
#define TC_THREADS (32 * (4 * TC_QT + 2))
#define TC_CHUNK 64                 // band candidates per pool chunk (rows chain as many chunks as they need)
#define TC_SMEM (TC_TILE_BYTES * (TC_QT + TC_STAGES) + 256)

// per (query, split) row: winner of the pool-exhausted fallback (d, idx), number of band candidates, first pool chunk
struct FSplitOutTC { double d; int idx; int cnt; int first; int pad; };
// candidate pool: chunk c holds entries[c * TC_CHUNK ..), belongs to row[c], has n[c] entries and continues in next[c]
// Chunks [0, base) are owned statically: chunk r is the FIRST chunk of split-row r (no allocation); the rest of the pool is
// handed out through `counter`.  base = rows x splits, or 0 when a test hook has capped the pool below that.
struct TcPool { uint32_t *entries; int *row, *n, *next; int *counter; int cap; int base; };

// ------------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        // the suspend-time hint lets the warp sleep in hardware instead of re-issuing the test: ten warps share four
        // schedulers here, and a spinning waiter takes issue slots from the single MMA-issuing thread
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity), "r"(2000u) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, 128 x 128 x 16, fp16 in / f32 accumulate
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float *v) {
    uint32_t *r = reinterpret_cast<uint32_t *>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// The same load without the wait: the registers are NOT valid until tc_ld_wait32 has been executed on them.  The wait takes
// the 32 registers as read-write operands, so every use of the values depends on it and cannot be scheduled above it.
__device__ __forceinline__ void tc_ld32_issue(uint32_t taddr, float *v) {
    uint32_t *r = reinterpret_cast<uint32_t *>(v);
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld_wait32(float *v) {
    uint32_t *r = reinterpret_cast<uint32_t *>(v);
    asm volatile("tcgen05.wait::ld.sync.aligned;"
        : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
          "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
          "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
          "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
        :: "memory");
}
// no-swizzle K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout, version 1)
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ull << 46);
}
// cute::UMMA::InstrDescriptor: c = f32 (1 << 4), a = b = f16 (format 0 at bits 7 and 10), K-major both, N >> 3 at 17, M >> 4 at 24
#define TC_IDESC ((1u << 4) | ((128u >> 3) << 17) | ((128u >> 4) << 24))

// ------------------------------------------------------------------------------------------------ operand packing
__device__ __forceinline__ size_t tc_row_offset(int64_t row) {   // byte offset of (row, k = 0) inside the packed image
    int64_t tile = row / TC_M;
    int r = (int)(row % TC_M);
    return (size_t)tile * TC_TILE_BYTES + (size_t)(r >> 3) * TC_SBO + (size_t)(r & 7) * 16;
}

// Centre of the filter's coordinate system: the mean of up to 4096 evenly strided target rows (one CTA, fixed order).
// ANY centre is valid -- |a - b| does not depend on it -- it only shrinks |a|^2 + |b|^2, which scales the error band;
// FPFH rows of a scene cluster around a few typical histograms, so the gain is large.
#define TC_MU_GROUPS 31
__global__ void __launch_bounds__(33 * TC_MU_GROUPS) feat_tc_center_kernel(const double *__restrict__ f, int64_t n, double *__restrict__ mu) {
    __shared__ double part[TC_MU_GROUPS][33];
    const int j = threadIdx.x % 33, g = threadIdx.x / 33;
    const int64_t S = n < 4096 ? n : 4096;
    const int64_t stride = n / S;
    double s = 0.0;
    for (int64_t k = g; k < S; k += TC_MU_GROUPS) s += f[(k * stride) * 33 + j];
    part[g][j] = s;
    __syncthreads();
    if (g == 0) {
        double t = 0.0;
        for (int k = 0; k < TC_MU_GROUPS; ++k) t += part[k][j];
        mu[j] = t / (double)S;
    }
}

// Largest |f - mu| over all rows of one feature set (atomicMax on the bits of a non-negative float, rounded up).  The pack
// kernel turns it into a power-of-two scale that brings every value into [-TC_RANGE, TC_RANGE]: scaling by 2^e is exact,
// so the filter's geometry is unchanged, and fp16 (range 65504, 11-bit significand) is then safe for any input units.
__global__ void __launch_bounds__(256) feat_tc_range_kernel(const double *__restrict__ f, int64_t n, const double *__restrict__ mu,
                                                            unsigned *__restrict__ maxabs_bits) {
    float m = 0.f;
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < n * 33; e += (int64_t)gridDim.x * blockDim.x)
        m = fmaxf(m, __double2float_ru(fabs(f[e] - mu[e % 33])));
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(maxabs_bits, __float_as_uint(m));
}
__device__ __forceinline__ double tc_scale(const unsigned *maxabs_bits) {
    float m = __uint_as_float(*maxabs_bits);
    if (!(m > 0.f) || !isfinite(m)) return 1.0;
    int x;
    frexp((double)m, &x);                                 // m <= 2^x
    return ldexp(1.0, 10 - x);                            // m * scale <= 2^10 = TC_RANGE
}

// role 0 = query rows, role 1 = target rows.  Rows in [n, padded) are written too: zeros for queries; for targets a
// score below anything a real row can reach (|a.b| <= 33 * 2^20, h <= 33 * 2^19), so a padding row is never a maximum and
// never inside a band.  One thread builds one row in registers and stores it as 14 uint4 (one per 16-byte K chunk).
// The norm term travels as  query 1024 x target -(h / 1024)  in three fp16 pieces (exact power-of-two rescaling).
__global__ void __launch_bounds__(128) feat_tc_pack_kernel(const double *__restrict__ f, int64_t n, int64_t n_pad, int role,
                                                           const double *__restrict__ mu, const unsigned *__restrict__ maxabs_bits,
                                                           uint8_t *__restrict__ img, float *__restrict__ norm32,
                                                           unsigned *__restrict__ maxnorm) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    unsigned short row[TC_K];
#pragma unroll
    for (int k = 0; k < TC_K; ++k) row[k] = 0;
    if (i < n) {
        const double sc = tc_scale(maxabs_bits);
        double s = 0.0;
#pragma unroll
        for (int j = 0; j < 33; ++j) {
            float v = (float)((f[i * 33 + j] - mu[j]) * sc);   // centred and scaled: distances scale by sc^2, exactly
            __half vh = __float2half_rn(v);
            __half vl = __float2half_rn(v - __half2float(vh));  // v - vh is exact in fp32
            s += (double)v * (double)v;
            row[j] = __half_as_ushort(vh);
            row[33 + j] = __half_as_ushort(role == 0 ? vl : vh);
            row[66 + j] = __half_as_ushort(role == 0 ? vh : vl);
        }
        float nf = __double2float_ru(s);
        if (norm32) norm32[i] = nf;
        atomicMax(maxnorm, __float_as_uint(nf));
        if (role == 0) {
            row[99] = row[100] = row[101] = __half_as_ushort(__float2half_rn(1024.f));
        } else {
            double hq = 0.5 * s / 1024.0;
            __half h0 = __double2half(hq);
            __half h1 = __double2half(hq - (double)__half2float(h0));
            __half h2 = __double2half(hq - (double)__half2float(h0) - (double)__half2float(h1));
            row[99] = __half_as_ushort(__hneg(h0)); row[100] = __half_as_ushort(__hneg(h1)); row[101] = __half_as_ushort(__hneg(h2));
        }
    } else if (role == 1) {
        row[99] = row[100] = row[101] = __half_as_ushort(__float2half_rn(-60000.f));   // score -1.8e8
    }
    uint4 *dst = reinterpret_cast<uint4 *>(img + tc_row_offset(i));
#pragma unroll
    for (int c = 0; c < TC_KCH; ++c) {
        uint4 q;
        q.x = row[8 * c] | ((unsigned)row[8 * c + 1] << 16); q.y = row[8 * c + 2] | ((unsigned)row[8 * c + 3] << 16);
        q.z = row[8 * c + 4] | ((unsigned)row[8 * c + 5] << 16); q.w = row[8 * c + 6] | ((unsigned)row[8 * c + 7] << 16);
        dst[c * (TC_LBO / 16)] = q;
    }
}

// fp32 copies of the RAW rows (stride TC_R32 floats, 16-byte aligned) and an upper bound of their Euclidean norms, for
// the second filter in feat_tc_eval_kernel.
#define TC_R32 36
__global__ void __launch_bounds__(256) feat_tc_rows32_kernel(const double *__restrict__ f, int64_t n, float *__restrict__ rows32,
                                                             float *__restrict__ nrm) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    double s = 0.0;
    float *o = rows32 + i * TC_R32;
#pragma unroll
    for (int j = 0; j < 33; ++j) {
        double v = f[i * 33 + j];
        s += v * v;
        o[j] = __double2float_rn(v);
    }
    o[33] = o[34] = o[35] = 0.f;
    nrm[i] = __double2float_ru(sqrt(s) * (1.0 + 1e-12));
}

// ------------------------------------------------------------------------------------------------ error band
// All quantities are those of the centred, power-of-two-scaled rows (|value| <= 1024); scaling multiplies every distance
// by the same exact factor, so order and ties are unchanged.
// D32 := an32 - 2 s' (real arithmetic on the computed fp32 values) against the exact f64 distance d of the (scaled) f64 rows:
//   input rounding f64 -> f32          2.4e-7 (an + bn)
//   an32 rounding                      0.6e-7  an
//   3-term fp16 split (11-bit significands, like tf32): representation + dropped al.bl
//                                      3 * 2^-22 |a_j b_j| per term  ->  7.2e-7 (an + bn)   in D
//   fp16 subnormal floor of the low parts: |x - (xh + xl)| <= 2^-25 absolute  ->  3.5e-7 (sqrt an + sqrt bn)   in D
//   norm pieces h/1024 = h0 + h1 + h2 in fp16, times 1024:  residual <= 2^-25 * 1024  ->  6.2e-5   in D
//   accumulation: 102 terms, each addition modelled as TRUNCATED at fp32 (1 ulp of the running magnitude, the most
//   pessimistic behaviour a tensor-core accumulator can have), running magnitude <= (an + 2 bn) / 2
//                                      2 * 102 * 2^-23 * (an + 2 bn) / 2  =  1.22e-5 an + 2.43e-5 bn
// => |D32 - d| <= 1.33e-5 an + 2.54e-5 bn + 3.5e-7 (sqrt an + sqrt bn) + 6.2e-5.  TC_ERR_A / TC_ERR_B keep a 1.15x margin.
// In score space the band is E (s' = -D/2, and nearest-neighbour candidates are those with D32 <= min D32 + 2E).
#define TC_ERR_A 1.53e-5f
#define TC_ERR_B 2.92e-5f
// The bound as a sum of a query part and a target part, every operation rounded up:
//     |D32 - d|  <=  err_a(an) + err_b(bn),   err_a = A an + 4e-7 sqrt(an),   err_b = B bn + 4e-7 sqrt(bn) + 1e-4.
// It is monotone in bn, so a target tile whose rows all have |b|^2 <= bn_tile can use err_b(bn_tile) instead of err_b(max over
// ALL targets): with the targets of a cached pack sorted by norm, most tiles get a band several times narrower (the median
// centred FPFH norm of a memory cloud is ~9x below the maximum).
__device__ __forceinline__ float feat_tc_err_a(float an) {
    return __fadd_ru(__fmul_ru(TC_ERR_A, an), __fmul_ru(4.0e-7f, __fsqrt_ru(an)));
}
__device__ __forceinline__ float feat_tc_err_b(float bn) {
    return __fadd_ru(__fadd_ru(__fmul_ru(TC_ERR_B, bn), __fmul_ru(4.0e-7f, __fsqrt_ru(bn))), 1.0e-4f);
}
// monotone float <-> int key so that atomicMax(int) orders floats
__device__ __forceinline__ int f2key(float f) { int b = __float_as_int(f); return b >= 0 ? b : b ^ 0x7fffffff; }
__device__ __forceinline__ float key2f(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__device__ __noinline__ double feat_tc_exact(const double *__restrict__ a, const double *__restrict__ b) {
    double s = 0.0;      // the oracle's order: j ascending, one rounding per operation
#pragma unroll 11
    for (int j = 0; j < 33; ++j) { double d = __dsub_rn(a[j], b[j]); s = __dadd_rn(s, __dmul_rn(d, d)); }
    return s;
}

__device__ unsigned long long g_tc_stats[2];   // band candidates, exact f64 evaluations (instrumentation; see ibl_feature_nn_stats)
// Who waits for whom inside feat_tc_kernel (instrumentation, off unless ibl_feature_nn_wait_stats(.., 1) switched it on):
// [0] CTAs, [1] target tiles streamed, [2] MMA thread: clocks in its tile loop, [3] of those waiting for a free accumulator
// (epilogue-bound), [4] waiting for a target tile in shared memory (copy-bound), [5] epilogue warp 0: clocks in its tile
// loop, [6] of those waiting for an accumulator (MMA-bound), [7] producer: clocks waiting for a free stage.
__device__ unsigned long long g_tc_wait[8];
__device__ int g_tc_wait_on;

// A band candidate of this thread's query goes into the row's chain of pool chunks (one atomicAdd per TC_CHUNK
// candidates; the chain itself is thread-private) for feat_tc_eval_kernel.  Only if the pool is exhausted is it
// evaluated on the spot.
struct TcRowState { int cnt, cur, ccnt, first, n_eval, bi, exhausted; double bestd; };
// slow path of tc_push: the row needs a (new) chunk, or the pool has run dry
__device__ __noinline__ TcRowState tc_push_slow(TcRowState st, const TcPool pool, int row, int t,
                                                const double *__restrict__ fa_q, const double *__restrict__ fb,
                                                const int *__restrict__ perm) {
    if (!st.exhausted) {
        int c = pool.base + atomicAdd(pool.counter, 1);
        if (c >= pool.cap) {
            st.exhausted = 1;                             // a thread asks the pool at most once after it runs dry
        } else {
            if (st.cur >= 0) { pool.n[st.cur] = TC_CHUNK; pool.next[st.cur] = c; } else st.first = c;
            pool.row[c] = row; pool.next[c] = -1;
            st.cur = c; st.ccnt = 0;
        }
    }
    if (st.exhausted) {                                   // exact evaluation right here (correct, just slow)
        const int to = perm ? perm[t] : t;              // the target's ORIGINAL index: exact rows and ties go by it
        double de = feat_tc_exact(fa_q, fb + (size_t)to * 33);
        if (st.bi < 0 || de < st.bestd || (de == st.bestd && to < st.bi)) { st.bestd = de; st.bi = to; }
        ++st.n_eval;
        return st;
    }
    pool.entries[(size_t)st.cur * TC_CHUNK + st.ccnt] = (uint32_t)t;
    ++st.ccnt;
    return st;                                            // by value: the row state stays in registers on the fast path
}
__device__ __forceinline__ void tc_push(TcRowState &st, const TcPool &pool, int row, int t, const double *__restrict__ fa_q,
                                        const double *__restrict__ fb, const int *__restrict__ perm) {
    ++st.cnt;
    if (st.cur >= 0 && st.ccnt < TC_CHUNK) {              // fast path: one store
        pool.entries[(size_t)st.cur * TC_CHUNK + st.ccnt] = (uint32_t)t;
        ++st.ccnt;
    } else {
        st = tc_push_slow(st, pool, row, t, fa_q, fb, perm);
    }
}

// One chunk of 32 scores of this thread's query row, columns tb .. tb+31.  With S the true score of a column and s' the
// computed one, |S - s'| <= eh for every column of this tile (eh = half the tile's band).  L = max over the columns seen so
// far of (s' - their eh) is a lower bound of the best TRUE score, so the best column -- and every exact tie with it -- has
// s' >= L - eh when it streams by: those are recorded.  L only grows, so the recorded set is a superset of what the final L
// would select; the surplus only costs exact evaluations later.  With one eh for all tiles this is the plain running
// maximum m = L + eh and the band [m - 2 eh, m].  The chunk maximum is a tree of independent FMNMX, not a 32-deep chain.
__device__ __forceinline__ void tc_band_chunk(const float *v, float &L, float eh, int tb, TcRowState &rs, const TcPool &pool,
                                              int row, const double *__restrict__ fa_q, const double *__restrict__ fb,
                                              const int *__restrict__ perm) {
    if (!(eh > -INFINITY)) return;                        // a row past the end (eh = -inf) records nothing, whatever its scores are
    float t[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) t[c] = fmaxf(v[2 * c], v[2 * c + 1]);
#pragma unroll
    for (int w = 8; w; w >>= 1) {
#pragma unroll
        for (int c = 0; c < w; ++c) t[c] = fmaxf(t[c], t[c + w]);
    }
    const float cm = t[0];
    L = fmaxf(L, __fsub_rd(cm, eh));                      // eh = -inf for rows past the end: L = r = +inf, nothing is recorded
    const float r = __fsub_rd(L, eh);
    if (cm >= r) {
        // v[c] >= r  <=>  the sign bit of fl(v[c] - r) is clear (subtraction of finite floats has the exact sign, denormals
        // are kept).  A funnel shift collects the 32 sign bits, column 0 ending at bit 31: two instructions per column
        // instead of compare + select + or.
        unsigned below = 0;
#pragma unroll
        for (int c = 0; c < 32; ++c) below = __funnelshift_l(__float_as_uint(__fsub_rn(v[c], r)), below, 1);
        unsigned mask = ~below;
        while (mask) {
            const int c = __clz(mask);                    // ascending column order
            mask &= ~(0x80000000u >> c);
            tc_push(rs, pool, row, tb + c, fa_q, fb, perm);
        }
    }
}

// MODE 0: dump s' to dump[q][ld] (debug).  MODE 1: running row maximum + band candidates in one pass.
// grid = (ceil(query tiles / TC_QT), splits); split y owns the target tiles y, y + nsplit, y + 2 nsplit, ... and visits them
// with a stride coprime to their number (tc_tile_of).  With the targets of a cached pack sorted by norm, a contiguous range
// walked in order made the scores of a row climb steadily -- every new record is a band candidate -- while interleaved splits
// and a scattered visiting order show every row a tile from its own norm range within the first few tiles, so its lower
// bound L is close to final early on.  For unsorted targets the order is irrelevant.
// A CTA keeps TC_QT = 2 query tiles resident and gives each its own pair of TMEM accumulators and its own four epilogue
// warps (a warp may only read the TMEM lanes 32 (warp % 4) .. +31, i.e. a thread is one query row): every streamed target
// tile feeds 2 x 7 MMAs, and the eight epilogue warps -- two per scheduler -- overlap each other's dependent-issue latency.
// tools/probes/tmem_probe.cu (profiles/r02_tmem_probe.jsonl) measured the TMEM read-back itself at 450 B/clk/SM with four
// reading warps and 720 with eight (90-145 clk per 128 x 128 tile, far below the 448 clk of 7 MMAs), and the same loop with
// a running maximum over the values at 330 clk (four warps) against 190 clk (eight): the epilogue is bound by its ALU work
// issued from one warp per scheduler, not by TMEM.
template <int MODE>
__global__ void __launch_bounds__(TC_THREADS, 1)
feat_tc_kernel(const uint8_t *__restrict__ Aimg, int64_t na, int n_qtiles, const uint8_t *__restrict__ Bimg, int64_t nb, int n_btiles,
               int tiles_per_split, int nsplit, const float *__restrict__ an32, const unsigned *__restrict__ bmax_bits,
               const double *__restrict__ fa, const double *__restrict__ fb,
               FSplitOutTC *__restrict__ out, TcPool pool, float *__restrict__ dump, int64_t dump_ld, int swap_lbo_sbo,
               const float *__restrict__ tile_eb, const int *__restrict__ perm, const int *__restrict__ n_live_dev,
               const int *__restrict__ qmap) {
    // n_live_dev / qmap (both or neither): the query image holds a device-side number of rows, gathered from a larger set;
    // row q of the image is row qmap[q] of an32 / fa.  na is then only an upper bound (grid and scratch were sized by it),
    // and a CTA whose query tiles are all past the live rows leaves at once.
    if (n_live_dev && (int64_t)blockIdx.x * TC_QT * TC_M >= (int64_t)*n_live_dev) return;
    if (n_live_dev) na = min(na, (int64_t)*n_live_dev);
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sA = smem_u32(smem);                   // TC_QT query tiles
    const uint32_t sB0 = sA + TC_QT * TC_TILE_BYTES;
    uint8_t *tail = smem + TC_TILE_BYTES * (TC_QT + TC_STAGES);
    const uint32_t bar0 = smem_u32(tail);                 // 8-byte barriers: a_full, full[4], empty[4], tfull[2], tempty[2]
    const uint32_t BAR_A = bar0, BAR_FULL = bar0 + 8, BAR_EMPTY = bar0 + 8 + 8 * TC_STAGES,
                   BAR_TFULL = bar0 + 8 + 16 * TC_STAGES, BAR_TEMPTY = BAR_TFULL + 16;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(tail + 128);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt0 = blockIdx.x * TC_QT, split = blockIdx.y;
    const int nq_here = min(TC_QT, n_qtiles - qt0);       // query tiles this CTA really has (the last CTA may have one)
    (void)tiles_per_split;
    const int ntiles = split < n_btiles ? (n_btiles - split + nsplit - 1) / nsplit : 0;
    // visiting stride: ~0.618 ntiles, made coprime to ntiles (then it -> (it * stride) % ntiles is a permutation)
    int vstride = max(1, (int)(0.6180339887f * (float)ntiles));
    while (vstride > 1) {
        int x = vstride, y = ntiles;
        while (y) { const int r = x % y; x = y; y = r; }
        if (x == 1) break;
        --vstride;
    }
    // every role walks the same sequence: vcur = (it * vstride) % ntiles, kept incrementally; tile = split + nsplit * vcur
#define TC_NEXT_TILE(vcur) do { (vcur) += vstride; if ((vcur) >= ntiles) (vcur) -= ntiles; } while (0)
    constexpr int W_PROD = 4 * TC_QT, W_MMA = 4 * TC_QT + 1;
    const bool wait_on = (MODE == 1) && g_tc_wait_on != 0;

    if (threadIdx.x == 0) {
        mbar_init(BAR_A, 1);
        for (int s = 0; s < TC_STAGES; ++s) { mbar_init(BAR_FULL + 8 * s, 1); mbar_init(BAR_EMPTY + 8 * s, 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(BAR_TFULL + 8 * s, 1); mbar_init(BAR_TEMPTY + 8 * s, 128 * TC_QT); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == W_MMA) {                                  // TMEM: 512 columns = 2 buffers x TC_QT accumulators x 128 columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32((const void *)tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == W_PROD) {
        // ===== producer: the query tiles once, then the target tiles of this split through the ring =====
        if (lane == 0 && ntiles > 0) {
            mbar_expect_tx(BAR_A, (uint32_t)nq_here * TC_TILE_BYTES);
            for (int h = 0; h < nq_here; ++h)
                bulk_g2s(sA + h * TC_TILE_BYTES, Aimg + (size_t)(qt0 + h) * TC_TILE_BYTES, TC_TILE_BYTES, BAR_A);
            long long w_empty = 0;
            int vcur = 0;
            for (int it = 0; it < ntiles; ++it) {
                int s = it % TC_STAGES;
                const long long c0 = wait_on ? clock64() : 0;
                mbar_wait(BAR_EMPTY + 8 * s, ((it / TC_STAGES) & 1) ^ 1);
                if (wait_on) w_empty += clock64() - c0;
                mbar_expect_tx(BAR_FULL + 8 * s, TC_TILE_BYTES);
                bulk_g2s(sB0 + s * TC_TILE_BYTES, Bimg + (size_t)(split + nsplit * vcur) * TC_TILE_BYTES, TC_TILE_BYTES, BAR_FULL + 8 * s);
                TC_NEXT_TILE(vcur);
            }
            if (wait_on) atomicAdd(&g_tc_wait[7], (unsigned long long)w_empty);
        }
    } else if (warp == W_MMA) {
        // ===== MMA issuer: one thread, 7 tcgen05.mma (128 x 128 x 16) per (query tile, target tile) =====
        if (lane == 0 && ntiles > 0) {
            const uint32_t lbo = swap_lbo_sbo ? TC_SBO : TC_LBO, sbo = swap_lbo_sbo ? TC_LBO : TC_SBO;
            mbar_wait(BAR_A, 0);
            const uint64_t a_desc0 = tc_smem_desc(sA, lbo, sbo), b_desc0 = tc_smem_desc(sB0, lbo, sbo);
            long long w_tempty = 0, w_full = 0;
            const long long loop0 = wait_on ? clock64() : 0;
            for (int it = 0; it < ntiles; ++it) {
                int s = it % TC_STAGES, as = it & 1;
                const long long c0 = wait_on ? clock64() : 0;
                mbar_wait(BAR_TEMPTY + 8 * as, ((it >> 1) & 1) ^ 1);
                const long long c1 = wait_on ? clock64() : 0;
                mbar_wait(BAR_FULL + 8 * s, (it / TC_STAGES) & 1);
                if (wait_on) { w_tempty += c1 - c0; w_full += clock64() - c1; }
                tc_fence_after();
                // descriptors differ from the precomputed ones only in the 16-byte-unit start address (low 14 bits, no carry
                // out: shared memory is < 256 KB): one 64-bit add each instead of rebuilding them -- this thread shares its
                // scheduler with two epilogue warps and every instruction it does not issue is an MMA issued sooner
                const uint64_t b_off = (uint64_t)((s * TC_TILE_BYTES) >> 4);
                for (int h = 0; h < nq_here; ++h) {
                    const uint64_t a_off = (uint64_t)((h * TC_TILE_BYTES) >> 4);
#pragma unroll
                    for (int k = 0; k < TC_K / 16; ++k) {      // one MMA consumes K = 16 halves = two 16-byte chunks
                        const uint64_t k_off = (uint64_t)((k * 2 * TC_LBO) >> 4);
                        tc_mma_f16(tmem_base + as * (128 * TC_QT) + h * 128, a_desc0 + a_off + k_off, b_desc0 + b_off + k_off,
                                   TC_IDESC, k > 0);
                    }
                }
                tc_commit(BAR_EMPTY + 8 * s);             // smem stage free once these MMAs have read it
                tc_commit(BAR_TFULL + 8 * as);            // accumulators ready for the epilogue
            }
            if (wait_on) {
                atomicAdd(&g_tc_wait[0], 1ull);
                atomicAdd(&g_tc_wait[1], (unsigned long long)ntiles);
                atomicAdd(&g_tc_wait[2], (unsigned long long)(clock64() - loop0));
                atomicAdd(&g_tc_wait[3], (unsigned long long)w_tempty);
                atomicAdd(&g_tc_wait[4], (unsigned long long)w_full);
            }
        }
    } else {
        // ===== epilogue: thread = one query row of query tile h = warp / 4; TMEM lane = 32 (warp % 4) + lane =====
        const int h = warp >> 2;
        const int64_t q = (int64_t)(qt0 + h) * TC_M + (warp & 3) * 32 + lane;
        const bool live = h < nq_here && q < na;
        const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(h * 128);
        if (MODE == 0) {
            int vcur = 0;
            for (int it = 0; it < ntiles; ++it) {
                int as = it & 1;
                mbar_wait(BAR_TFULL + 8 * as, (it >> 1) & 1);
                tc_fence_after();
                const int64_t t0 = (int64_t)(split + nsplit * vcur) * TC_M;
                TC_NEXT_TILE(vcur);
#pragma unroll 1
                for (int ch = 0; ch < 4; ++ch) {
                    float v[32];
                    tc_ld32(lane_base + as * (128 * TC_QT) + ch * 32, v);
                    if (live) {
#pragma unroll
                        for (int c = 0; c < 32; ++c) { int64_t t = t0 + ch * 32 + c; if (t < nb) dump[q * dump_ld + t] = v[c]; }
                    }
                }
                tc_fence_before();
                mbar_arrive(BAR_TEMPTY + 8 * as);
            }
        } else {
            // Single pass: running maximum + band recording (tc_band_chunk).  The TMEM reads are software-pipelined over two
            // register buffers -- the load of chunk c+1 is in flight while chunk c is processed -- and the accumulator is handed
            // back to the MMA warp as soon as its last chunk has been READ, before that chunk is processed.
            // tile_eb (cached packs): the target part of the band per TARGET TILE, from the largest norm inside the tile;
            // otherwise one value from the largest target norm of all
            float L = -INFINITY;
            const int64_t qo = (live && qmap) ? (int64_t)qmap[q] : q;      // the row's index in an32 / fa
            const float ea = live ? feat_tc_err_a(an32[qo]) : -INFINITY;
            const float eb_all = feat_tc_err_b(__uint_as_float(*bmax_bits));
            TcRowState rs;
            rs.cnt = 0; rs.cur = -1; rs.ccnt = 0; rs.first = -1; rs.n_eval = 0; rs.bi = -1; rs.exhausted = 0; rs.bestd = INFINITY;
            const int row = (int)(q * nsplit + split);
            const double *fa_q = fa + qo * 33;
            if (live && row < pool.base) {                   // this row's first chunk is chunk `row`: no pool traffic for it
                rs.cur = row; rs.first = row;
                pool.row[row] = row; pool.next[row] = -1;
            }
            float va[32], vb[32];
            long long w_tfull = 0;
            const long long eloop0 = wait_on ? clock64() : 0;
            if (ntiles > 0) {
                mbar_wait(BAR_TFULL, 0);
                if (wait_on) w_tfull += clock64() - eloop0;
                tc_fence_after();
                tc_ld32_issue(lane_base, va);
            }
            int vcur = 0;
#pragma unroll 1
            for (int it = 0; it < ntiles; ++it) {
                const int as = it & 1;
                const uint32_t acc = lane_base + as * (128 * TC_QT);
                const int tile = split + nsplit * vcur;
                TC_NEXT_TILE(vcur);
                const int tb = tile * TC_M;                         // padding rows score far below any band
                const float eh = __fmul_ru(0.5f, __fadd_ru(ea, tile_eb ? tile_eb[tile] : eb_all));
                tc_ld_wait32(va); tc_ld32_issue(acc + 32, vb);
                tc_band_chunk(va, L, eh, tb, rs, pool, row, fa_q, fb, perm);
                tc_ld_wait32(vb); tc_ld32_issue(acc + 64, va);
                tc_band_chunk(vb, L, eh, tb + 32, rs, pool, row, fa_q, fb, perm);
                tc_ld_wait32(va); tc_ld32_issue(acc + 96, vb);
                tc_band_chunk(va, L, eh, tb + 64, rs, pool, row, fa_q, fb, perm);
                tc_ld_wait32(vb);
                tc_fence_before();
                mbar_arrive(BAR_TEMPTY + 8 * as);
                if (it + 1 < ntiles) {
                    const long long c0 = wait_on ? clock64() : 0;
                    mbar_wait(BAR_TFULL + 8 * (as ^ 1), ((it + 1) >> 1) & 1);
                    if (wait_on) w_tfull += clock64() - c0;
                    tc_fence_after();
                    tc_ld32_issue(lane_base + (as ^ 1) * (128 * TC_QT), va);
                }
                tc_band_chunk(vb, L, eh, tb + 96, rs, pool, row, fa_q, fb, perm);
            }
            if (wait_on && threadIdx.x == 0) {
                atomicAdd(&g_tc_wait[5], (unsigned long long)(clock64() - eloop0));
                atomicAdd(&g_tc_wait[6], (unsigned long long)w_tfull);
            }
            if (live) {
                if (rs.cur >= 0) pool.n[rs.cur] = rs.ccnt;
                FSplitOutTC o;
                o.d = rs.bestd; o.idx = rs.bi; o.cnt = rs.cnt; o.first = rs.first; o.pad = 0;
                out[row] = o;
                if (rs.cnt) atomicAdd(&g_tc_stats[0], (unsigned long long)rs.cnt);
                if (rs.n_eval) atomicAdd(&g_tc_stats[1], (unsigned long long)rs.n_eval);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == W_MMA) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
    }
#undef TC_NEXT_TILE
}

struct TcChunkRes { double d; int idx; int pad; };

// Second filter, on fp32 DIFFERENCES.  s = sum_j fl(fl(a32_j - b32_j)^2) in fp32 (j ascending, every operation rounded to
// nearest, no FMA) against the exact d = sum_j (a_j - b_j)^2, with u = 2^-24:
//   input rounding + the subtraction      |delta_j - x_j| <= 2.0000002 u (|a_j| + |b_j|)   =>  ||delta - x|| <= H := 2.0000002 u (|a| + |b|)
//   squares and the 33-term sum           |s - sum delta_j^2| <= 36 u sum delta_j^2,  sum delta_j^2 <= (1 + 37 u) s
//   sum delta^2 against d                 |sum delta^2 - d| <= 2 sqrt(d) H + H^2,  sqrt(d) <= sqrt(sum delta^2) + H
// => |s - d| <= 37 u s + 2.00001 sqrt(s) H + 3 H^2; the kernel uses 40 u s + 2.1 sqrt(s) H + 3.1 H^2 + 1e-30, rounded up.  The
// error scales with |a| |a - b| instead of the GEMM filter's |a|^2, i.e. it is ~10^3 times tighter for near neighbours.
__device__ __forceinline__ float feat_tc_err32(float s, float na, float nb) {
    const float H = __fmul_ru(1.1920932e-7f, __fadd_ru(na, nb));          // 2.0000002 * 2^-24, rounded up
    float e = __fmul_ru(2.3841864e-6f, s);                                 // 40 u s
    e = __fadd_ru(e, __fmul_ru(2.1f, __fmul_ru(__fsqrt_ru(s), H)));
    e = __fadd_ru(e, __fmul_ru(3.1f, __fmul_ru(H, H)));
    return __fadd_ru(e, 1e-30f);
}
__device__ __forceinline__ float feat_tc_d32(const float *__restrict__ a, const float *__restrict__ b) {
    float s = 0.f;
#pragma unroll
    for (int j4 = 0; j4 < 32; j4 += 4) {
        const float4 x = *reinterpret_cast<const float4 *>(a + j4), y = *reinterpret_cast<const float4 *>(b + j4);
        float d;
        d = __fsub_rn(x.x, y.x); s = __fadd_rn(s, __fmul_rn(d, d));
        d = __fsub_rn(x.y, y.y); s = __fadd_rn(s, __fmul_rn(d, d));
        d = __fsub_rn(x.z, y.z); s = __fadd_rn(s, __fmul_rn(d, d));
        d = __fsub_rn(x.w, y.w); s = __fadd_rn(s, __fmul_rn(d, d));
    }
    const float d = __fsub_rn(a[32], b[32]);
    return __fadd_rn(s, __fmul_rn(d, d));
}

// EIGHT LANES per pool chunk (<= 64 candidates, up to eight per lane), four chunks per warp.  Phase 1: the fp32 difference
// distance of every candidate with its error radius; U = the smallest upper bound in the chunk.  Phase 2: only candidates
// whose lower bound does not exceed U -- the only ones that can be the chunk's nearest or tie with it -- get the oracle's f64
// evaluation, reduced in the oracle's order (distance, then lower index) into res[chunk].
// A chunk holds 8-9 candidates on average (every split-row owns at least one chunk) and about one survivor, and the kernel's
// time is the latency of its gathers: with a whole warp per chunk a quarter of the lanes worked in phase 1 and one in
// thirty in phase 2; four chunks side by side put four times as many gathers in flight per warp.
// The band provably holds the exact nearest neighbour, so the chunk of a row with ONE candidate needs no work at all here:
// d = -1 marks "not evaluated"; the merge evaluates it only if another split of the same query also has candidates.
template <int TC_EVAL_GS>                                // lanes per chunk
__global__ void __launch_bounds__(256) feat_tc_eval_kernel(const FSplitOutTC *__restrict__ parts, TcPool pool, int nsplit,
                                                           TcChunkRes *__restrict__ res, const double *__restrict__ fa,
                                                           const double *__restrict__ fb, const float *__restrict__ a32,
                                                           const float *__restrict__ b32, const float *__restrict__ na32,
                                                           const float *__restrict__ nb32, const int *__restrict__ perm,
                                                           const int *__restrict__ n_live_dev, const int *__restrict__ qmap) {
    // pool entries, b32 and nb32 are in the TARGET IMAGE's row order; perm (may be NULL = identity) maps such a position to
    // the target's original index, which addresses fb and decides exact ties.  n_live_dev / qmap: see feat_tc_kernel (a32,
    // na32 and fa are indexed through qmap; the static chunks of rows past the live ones were never written and are skipped)
    constexpr int TC_EVAL_R = TC_CHUNK / TC_EVAL_GS;      // candidates per lane
    const int lane = threadIdx.x & 31, grp = lane / TC_EVAL_GS, gl = lane % TC_EVAL_GS;
    constexpr int GPW = 32 / TC_EVAL_GS;                  // chunks per warp
    const int nwarps = gridDim.x * (blockDim.x >> 5);
    const int nchunks = min(pool.base + *pool.counter, pool.cap);
    unsigned long long n_eval = 0;
    for (int c0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * GPW; c0 < nchunks; c0 += nwarps * GPW) {
        const int c = c0 + grp;
        bool valid = c < nchunks;
        if (valid && n_live_dev && c < pool.base && c / nsplit >= *n_live_dev) valid = false;   // chunk c < base is split-row c
        const int row = valid ? pool.row[c] : 0;
        int n = valid ? pool.n[c] : 0;
        const uint32_t *e = pool.entries + (size_t)c * TC_CHUNK;
        bool single = false;
        if (valid && parts[row].cnt == 1) {
            if (gl == 0) { TcChunkRes r; r.d = -1.0; r.idx = perm ? perm[e[0]] : (int)e[0]; r.pad = 0; res[c] = r; }
            single = true;
            n = 0;
        }
        const int64_t q = (valid && qmap) ? (int64_t)qmap[row / nsplit] : (int64_t)(row / nsplit);
        const float *aq = a32 + q * TC_R32;
        const float naq = n > 0 ? na32[q] : 0.f;
        int t[TC_EVAL_R];
        float lo[TC_EVAL_R];
        float U = INFINITY;
#pragma unroll
        for (int r = 0; r < TC_EVAL_R; ++r) {
            t[r] = -1; lo[r] = INFINITY;
            const int k = r * TC_EVAL_GS + gl;
            if (k < n) {
                t[r] = (int)e[k];
                const float s = feat_tc_d32(aq, b32 + (size_t)t[r] * TC_R32);
                const float E2 = feat_tc_err32(s, naq, nb32[t[r]]);
                if (isfinite(s) && isfinite(E2)) {
                    lo[r] = __fsub_rd(s, E2);
                    U = fminf(U, __fadd_ru(s, E2));
                } else {
                    lo[r] = -INFINITY;                     // units beyond fp32's range: no bound, always evaluated exactly
                }
            }
        }
        for (int o = TC_EVAL_GS / 2; o; o >>= 1) U = fminf(U, __shfl_xor_sync(0xffffffffu, U, o));
        const double *a = fa + (size_t)q * 33;
        double bestd = INFINITY;
        int bi = -1;
#pragma unroll
        for (int r = 0; r < TC_EVAL_R; ++r) {
            if (t[r] >= 0 && lo[r] <= U) {
                const int to = perm ? perm[t[r]] : t[r];
                double de = feat_tc_exact(a, fb + (size_t)to * 33);
                if (bi < 0 || de < bestd || (de == bestd && to < bi)) { bestd = de; bi = to; }
                ++n_eval;
            }
        }
        for (int o = TC_EVAL_GS / 2; o; o >>= 1) {
            double od = __shfl_xor_sync(0xffffffffu, bestd, o);
            int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (oi >= 0 && (bi < 0 || od < bestd || (od == bestd && oi < bi))) { bestd = od; bi = oi; }
        }
        if (gl == 0 && valid && !single) { TcChunkRes r; r.d = bestd; r.idx = bi; r.pad = 0; res[c] = r; }
    }
    if (n_eval) atomicAdd(&g_tc_stats[1], n_eval);
}

// Per query: combine the splits.  Each split row contributes its chain of chunk results and, if the pool ran dry, the
// winner of its on-the-spot evaluations.
__global__ void feat_tc_merge_kernel(const FSplitOutTC *__restrict__ parts, int64_t na, int nsplit, TcPool pool,
                                     const TcChunkRes *__restrict__ res, const double *__restrict__ fa,
                                     const double *__restrict__ fb, int32_t *__restrict__ nn, const int *__restrict__ perm,
                                     const int *__restrict__ n_live_dev, const int *__restrict__ qmap) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= na || (n_live_dev && i >= *n_live_dev)) return;
    const int64_t io = qmap ? (int64_t)qmap[i] : i;       // row i of the query image is row io of fa / nn
    long long total = 0;
    int only = -1;
    for (int s = 0; s < nsplit; ++s) {
        const FSplitOutTC p = parts[i * nsplit + s];
        total += p.cnt;
        if (p.cnt == 1) {
            if (p.first >= 0) { const int t = (int)pool.entries[(size_t)p.first * TC_CHUNK]; only = perm ? perm[t] : t; }
            else only = p.idx;                            // evaluated on the spot (pool exhausted): already an original index
        }
    }
    if (total == 0) { nn[io] = -1; return; }
    if (total == 1) { nn[io] = only; return; }            // the single band candidate IS the exact nearest neighbour
    double bestd = INFINITY;
    int bi = -1;
    unsigned long long extra = 0;
    for (int s = 0; s < nsplit; ++s) {
        const FSplitOutTC p = parts[i * nsplit + s];
        if (p.cnt == 0) continue;
        if (p.idx >= 0 && (bi < 0 || p.d < bestd || (p.d == bestd && p.idx < bi))) { bestd = p.d; bi = p.idx; }
        for (int c = p.first; c >= 0; c = pool.next[c]) {
            TcChunkRes r = res[c];
            if (r.idx < 0) continue;
            double d = r.d;
            if (d < 0.0) { d = feat_tc_exact(fa + io * 33, fb + (size_t)r.idx * 33); ++extra; }
            if (bi < 0 || d < bestd || (d == bestd && r.idx < bi)) { bestd = d; bi = r.idx; }
        }
    }
    if (extra) atomicAdd(&g_tc_stats[1], extra);
    nn[io] = bi;
}
struct TcPlan { int qtiles, btiles, nsplit, tiles_per_split; int64_t na_pad, nb_pad; };
// grid = query tiles x splits of the target tiles, one CTA per SM at a time.  The split count is chosen so that the CTAs fill
// whole waves of 148: cost(ns) = waves x (tiles per CTA + a fixed per-CTA cost of ~6 tile times for TMEM allocation, the
// query-tile load and the pipeline fill); more splits also mean more partial rows to merge, so ties go to the smaller ns.
static TcPlan tc_plan(int64_t na, int64_t nb) {
    TcPlan p;
    p.qtiles = (int)((na + TC_M - 1) / TC_M);
    p.btiles = (int)((nb + TC_M - 1) / TC_M);
    p.na_pad = (int64_t)p.qtiles * TC_M;
    p.nb_pad = (int64_t)p.btiles * TC_M;
    int best_ns = 1;
    long long best_cost = -1;
    for (int ns = 1; ns <= 32 && ns <= p.btiles; ++ns) {
        int tps = (p.btiles + ns - 1) / ns;
        int real_ns = (p.btiles + tps - 1) / tps;
        if (real_ns != ns) continue;
        if (ns > 1 && tps < 8) break;
        long long waves = ((long long)((p.qtiles + TC_QT - 1) / TC_QT) * ns + IBL_SMS - 1) / IBL_SMS;
        long long cost = waves * (tps + 6);
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_ns = ns; }
    }
    p.tiles_per_split = (p.btiles + best_ns - 1) / best_ns;
    p.nsplit = (p.btiles + p.tiles_per_split - 1) / p.tiles_per_split;
    return p;
}
static int tc_configure() {
    // the attribute is per function and per context; setting it again is cheap
    IBL_CUDA(cudaFuncSetAttribute(feat_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
    IBL_CUDA(cudaFuncSetAttribute(feat_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM));
    return IBL_OK;
}

static int g_tc_pool_cap_override = 0;
// Test hook: cap the candidate pool at `chunks` chunks of 64 candidates (0 restores the default sizing), so a test can
// force the pool-exhausted path, where candidates are evaluated on the spot inside pass 2.
IBL_API int ibl_feature_nn_set_pool_chunks(int chunks) {
    IBL_ARG(chunks >= 0);
    g_tc_pool_cap_override = chunks;
    return IBL_OK;
}

// The search proper on packed operands: one-pass tcgen05 score GEMM with band recording, the fp32-difference filter with
// exact f64 evaluation of the survivors, and the per-query merge.  A / B are the packed tile images of the queries (role 0)
// and targets (role 1) in ONE common centred / scaled coordinate system, an the queries' squared norms in it, bmax_bits the
// float bits of max |b|^2; fa / fb the raw f64 rows, a32 / b32 / na32 / nb32 their fp32 copies and norm bounds.
static int tc_search(const uint8_t *A, int64_t na, const uint8_t *B, int64_t nb, const float *an, const unsigned *bmax_bits,
                     const double *fa, const double *fb, const float *a32, const float *b32, const float *na32,
                     const float *nb32, int32_t *nn, cudaStream_t st, const float *tile_eb = nullptr, const int *perm = nullptr,
                     const int *n_live_dev = nullptr, const int *qmap = nullptr) {
    // B, b32, nb32 are in the target image's row order; tile_eb / perm (both or neither): a cached pack whose target rows are
    // sorted by norm -- per-tile band terms and the map from image row to original row (fb is in original order)
    TcPlan p = tc_plan(na, nb);
    FSplitOutTC *parts = nullptr;
    IBL_CUDA(ibl_alloc(&parts, (size_t)na * p.nsplit, st));
    // candidate pool: one chunk for every row that has a candidate at all + 256k chunks (16 M candidates) of headroom
    TcPool pool;
    const int64_t nrows = na * p.nsplit;
    pool.cap = (int)std::min<int64_t>(nrows + (1 << 18), (int64_t)1 << 24);
    if (g_tc_pool_cap_override > 0) pool.cap = g_tc_pool_cap_override;
    pool.base = nrows <= pool.cap ? (int)nrows : 0;
    int *pool_meta = nullptr;
    TcChunkRes *res = nullptr;
    IBL_CUDA(ibl_alloc(&pool.entries, (size_t)pool.cap * TC_CHUNK, st));
    IBL_CUDA(ibl_alloc(&pool_meta, (size_t)pool.cap * 3 + 1, st));
    IBL_CUDA(ibl_alloc(&res, (size_t)pool.cap, st));
    pool.row = pool_meta; pool.n = pool_meta + pool.cap; pool.next = pool_meta + 2 * (size_t)pool.cap;
    pool.counter = pool_meta + 3 * (size_t)pool.cap;
    IBL_CUDA(cudaMemsetAsync(pool.counter, 0, sizeof(int), st));
    dim3 grid((p.qtiles + TC_QT - 1) / TC_QT, p.nsplit);
    {
        // nested scope: the tcgen05 kernel alone, with the flops it ISSUES (depth 112 over the padded tiles)
        IBL_PROF(IBL_PROF_FEATURE_GEMM, st, 2.0 * TC_K * (double)p.na_pad * (double)p.nb_pad);
        feat_tc_kernel<1><<<grid, TC_THREADS, TC_SMEM, st>>>(A, na, p.qtiles, B, nb, p.btiles, p.tiles_per_split, p.nsplit, an, bmax_bits,
                                                            fa, fb, parts, pool, nullptr, 0, 0, tile_eb, perm, n_live_dev, qmap);
        IBL_LAUNCH_CHECK();
    }
    {
        // the chunk count lives on the device (no host sync): a fixed grid strides over it
        static int gs = 0;
        if (gs == 0) { const char *e = getenv("IBL_TC_EVAL_GS"); gs = e ? atoi(e) : 8; }
        if (gs == 4) feat_tc_eval_kernel<4><<<8 * IBL_SMS, 256, 0, st>>>(parts, pool, p.nsplit, res, fa, fb, a32, b32, na32, nb32, perm, n_live_dev, qmap);
        else if (gs == 16) feat_tc_eval_kernel<16><<<8 * IBL_SMS, 256, 0, st>>>(parts, pool, p.nsplit, res, fa, fb, a32, b32, na32, nb32, perm, n_live_dev, qmap);
        else feat_tc_eval_kernel<8><<<8 * IBL_SMS, 256, 0, st>>>(parts, pool, p.nsplit, res, fa, fb, a32, b32, na32, nb32, perm, n_live_dev, qmap);
        IBL_LAUNCH_CHECK();
    }
    feat_tc_merge_kernel<<<(unsigned)((na + 127) / 128), 128, 0, st>>>(parts, na, p.nsplit, pool, res, fa, fb, nn, perm, n_live_dev, qmap);
    IBL_LAUNCH_CHECK();
    ibl_free(pool.entries, st); ibl_free(pool_meta, st); ibl_free(res, st); ibl_free(parts, st);
    return IBL_OK;
}

// exact 33-d nearest neighbour of every row of fa among the rows of fb (tensor-core filter + exact f64 re-evaluation)
int ibl_feature_nn_tc(const double *fa, int64_t na, const double *fb, int64_t nb, int32_t *nn, cudaStream_t st) {
    int rc = tc_configure();
    if (rc < 0) return rc;
    TcPlan p = tc_plan(na, nb);
    uint8_t *A = nullptr, *B = nullptr;
    float *an = nullptr;
    unsigned *mx = nullptr;
    IBL_CUDA(ibl_alloc(&A, (size_t)p.qtiles * TC_TILE_BYTES, st));
    IBL_CUDA(ibl_alloc(&B, (size_t)p.btiles * TC_TILE_BYTES, st));
    IBL_CUDA(ibl_alloc(&an, (size_t)p.na_pad, st));
    IBL_CUDA(ibl_alloc(&mx, 3, st));                       // max |a|^2, max |b|^2, max |f - mu| (float bits)
    double *mu = nullptr;
    IBL_CUDA(ibl_alloc(&mu, 33, st));
    IBL_CUDA(cudaMemsetAsync(mx, 0, 3 * sizeof(unsigned), st));
    feat_tc_center_kernel<<<1, 33 * TC_MU_GROUPS, 0, st>>>(fb, nb, mu);
    IBL_LAUNCH_CHECK();
    feat_tc_range_kernel<<<(unsigned)std::min<int64_t>((na * 33 + 255) / 256, 2 * IBL_SMS), 256, 0, st>>>(fa, na, mu, mx + 2);
    IBL_LAUNCH_CHECK();
    feat_tc_range_kernel<<<(unsigned)std::min<int64_t>((nb * 33 + 255) / 256, 2 * IBL_SMS), 256, 0, st>>>(fb, nb, mu, mx + 2);
    IBL_LAUNCH_CHECK();
    feat_tc_pack_kernel<<<(unsigned)(p.na_pad / 128), 128, 0, st>>>(fa, na, p.na_pad, 0, mu, mx + 2, A, an, mx);
    IBL_LAUNCH_CHECK();
    feat_tc_pack_kernel<<<(unsigned)(p.nb_pad / 128), 128, 0, st>>>(fb, nb, p.nb_pad, 1, mu, mx + 2, B, nullptr, mx + 1);
    IBL_LAUNCH_CHECK();
    float *a32 = nullptr, *b32 = nullptr, *na32 = nullptr, *nb32 = nullptr;      // second filter (feat_tc_eval_kernel)
    IBL_CUDA(ibl_alloc(&a32, (size_t)na * TC_R32, st));
    IBL_CUDA(ibl_alloc(&b32, (size_t)nb * TC_R32, st));
    IBL_CUDA(ibl_alloc(&na32, (size_t)na, st));
    IBL_CUDA(ibl_alloc(&nb32, (size_t)nb, st));
    feat_tc_rows32_kernel<<<(unsigned)((na + 255) / 256), 256, 0, st>>>(fa, na, a32, na32);
    IBL_LAUNCH_CHECK();
    feat_tc_rows32_kernel<<<(unsigned)((nb + 255) / 256), 256, 0, st>>>(fb, nb, b32, nb32);
    IBL_LAUNCH_CHECK();
    rc = tc_search(A, na, B, nb, an, mx + 1, fa, fb, a32, b32, na32, nb32, nn, st);
    ibl_free(a32, st); ibl_free(b32, st); ibl_free(na32, st); ibl_free(nb32, st);
    ibl_free(mu, st); ibl_free(A, st); ibl_free(B, st); ibl_free(an, st); ibl_free(mx, st);
    return rc;
}

// ------------------------------------------------------------------------------------------------ cached packs
// Everything the matcher derives from ONE feature set, built once: the filter's centre (its mean) and power-of-two scale
// (chosen with one spare bit, so rows of the other side may be twice as far from the centre), the packed query-role and
// target-role tile images, the squared norms, and the fp32 rows / norm bounds of the second filter.  A memory-side
// feature set is matched against a new source cloud every frame (fpfh_register.py:110-112 in both directions): with its
// pack cached, a registration reads the memory-side f64 rows only for the few exact evaluations.
struct ibl_feature_pack {
    int64_t n, n_pad;
    const double *f;           // the caller's rows [n][33] (kept alive by the caller)
    double *mu;                // [33]
    unsigned *mx;              // [0] max |row|^2 as query, [1] max |row|^2 as target, [2] 2 * max |f - mu| (float bits)
    uint8_t *img0, *img1;      // role 0 (query) tile image in ORIGINAL row order / role 1 (target) image SORTED by norm
    float *an32;               // [n_pad] squared norms of the scaled rows (original order)
    float *rows32, *nrm32;     // second filter, original order (the pack as the query side)
    float *rows32s, *nrm32s;   // second filter, sorted order (the pack as the target side)
    int *perm;                 // [n_pad] sorted position -> original row (identity on the padding rows)
    float *tile_eb;            // [n_pad / 128] target part of the band of each sorted target tile (feat_tc_err_b of its largest norm)
    cudaStream_t stream;
};

__global__ void feat_tc_iota_kernel(int *v, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int)i;
}
// Row i of the sorted target image = row perm[i] of the image in original order (padding rows stay where they are), with its
// fp32 row and norm bound; every tile keeps the largest scaled norm of its rows (float bits, atomicMax).
__global__ void __launch_bounds__(128) feat_tc_sort_rows_kernel(int64_t n, int64_t n_pad, int *__restrict__ perm,
                                                                const uint8_t *__restrict__ img1o, uint8_t *__restrict__ img1s,
                                                                const float *__restrict__ rows32, const float *__restrict__ nrm32,
                                                                float *__restrict__ rows32s, float *__restrict__ nrm32s,
                                                                const float *__restrict__ an32, unsigned *__restrict__ tile_bn) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    if (i >= n) perm[i] = (int)i;
    const int64_t src = perm[i];
    const uint4 *s = reinterpret_cast<const uint4 *>(img1o + tc_row_offset(src));
    uint4 *d = reinterpret_cast<uint4 *>(img1s + tc_row_offset(i));
#pragma unroll
    for (int c = 0; c < TC_KCH; ++c) d[c * (TC_LBO / 16)] = s[c * (TC_LBO / 16)];
    if (i < n) {
        const float4 *rs = reinterpret_cast<const float4 *>(rows32 + src * TC_R32);
        float4 *rd = reinterpret_cast<float4 *>(rows32s + i * TC_R32);
#pragma unroll
        for (int c = 0; c < TC_R32 / 4; ++c) rd[c] = rs[c];
        nrm32s[i] = nrm32[src];
        atomicMax(tile_bn + i / TC_M, __float_as_uint(an32[src]));
    }
}
__global__ void feat_tc_tile_eb_kernel(const unsigned *__restrict__ tile_bn, float *__restrict__ tile_eb, int ntiles) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < ntiles) tile_eb[t] = feat_tc_err_b(__uint_as_float(tile_bn[t]));
}

__global__ void feat_tc_headroom_kernel(unsigned *maxabs_bits) {
    float m = __uint_as_float(*maxabs_bits);
    *maxabs_bits = __float_as_uint(__fmul_ru(m, 2.0f));
}

// Both roles, the norms and the fp32 rows of a feature set from ONE read of its f64 rows, in the coordinate system of
// another set's pack (centre mu, scale from maxabs_bits).  flag is raised if a scaled value leaves fp16's comfortable
// range (|v| > 16384): the caller then falls back to ibl_feature_nn.
__global__ void __launch_bounds__(128) feat_tc_pack_both_kernel(const double *__restrict__ f, int64_t n, int64_t n_pad,
                                                                const double *__restrict__ mu, const unsigned *__restrict__ maxabs_bits,
                                                                uint8_t *__restrict__ img0, uint8_t *__restrict__ img1,
                                                                float *__restrict__ an32, unsigned *__restrict__ maxnorm0,
                                                                unsigned *__restrict__ maxnorm1, float *__restrict__ rows32,
                                                                float *__restrict__ nrm32, unsigned *__restrict__ flag) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    unsigned short r0[TC_K], r1[TC_K];
#pragma unroll
    for (int k = 0; k < TC_K; ++k) { r0[k] = 0; r1[k] = 0; }
    if (i < n) {
        const double sc = tc_scale(maxabs_bits);
        double s = 0.0, sraw = 0.0;
        bool over = false;
        float *o32 = rows32 + i * TC_R32;
#pragma unroll
        for (int j = 0; j < 33; ++j) {
            const double raw = f[i * 33 + j];
            sraw += raw * raw;
            o32[j] = __double2float_rn(raw);
            float v = (float)((raw - mu[j]) * sc);
            over |= !(fabsf(v) <= 16384.f);
            __half vh = __float2half_rn(v);
            __half vl = __float2half_rn(v - __half2float(vh));
            s += (double)v * (double)v;
            r0[j] = __half_as_ushort(vh); r0[33 + j] = __half_as_ushort(vl); r0[66 + j] = __half_as_ushort(vh);
            r1[j] = __half_as_ushort(vh); r1[33 + j] = __half_as_ushort(vh); r1[66 + j] = __half_as_ushort(vl);
        }
        o32[33] = o32[34] = o32[35] = 0.f;
        nrm32[i] = __double2float_ru(sqrt(sraw) * (1.0 + 1e-12));
        float nf = __double2float_ru(s);
        an32[i] = nf;
        atomicMax(maxnorm0, __float_as_uint(nf));
        atomicMax(maxnorm1, __float_as_uint(nf));
        if (over) atomicOr(flag, 1u);
        r0[99] = r0[100] = r0[101] = __half_as_ushort(__float2half_rn(1024.f));
        double hq = 0.5 * s / 1024.0;
        __half h0 = __double2half(hq);
        __half h1 = __double2half(hq - (double)__half2float(h0));
        __half h2 = __double2half(hq - (double)__half2float(h0) - (double)__half2float(h1));
        r1[99] = __half_as_ushort(__hneg(h0)); r1[100] = __half_as_ushort(__hneg(h1)); r1[101] = __half_as_ushort(__hneg(h2));
    } else {
        r1[99] = r1[100] = r1[101] = __half_as_ushort(__float2half_rn(-60000.f));   // a padding target never scores
        an32[i] = 0.f;
    }
    uint4 *d0 = reinterpret_cast<uint4 *>(img0 + tc_row_offset(i));
    uint4 *d1 = reinterpret_cast<uint4 *>(img1 + tc_row_offset(i));
#pragma unroll
    for (int c = 0; c < TC_KCH; ++c) {
        uint4 q;
        q.x = r0[8 * c] | ((unsigned)r0[8 * c + 1] << 16); q.y = r0[8 * c + 2] | ((unsigned)r0[8 * c + 3] << 16);
        q.z = r0[8 * c + 4] | ((unsigned)r0[8 * c + 5] << 16); q.w = r0[8 * c + 6] | ((unsigned)r0[8 * c + 7] << 16);
        d0[c * (TC_LBO / 16)] = q;
        q.x = r1[8 * c] | ((unsigned)r1[8 * c + 1] << 16); q.y = r1[8 * c + 2] | ((unsigned)r1[8 * c + 3] << 16);
        q.z = r1[8 * c + 4] | ((unsigned)r1[8 * c + 5] << 16); q.w = r1[8 * c + 6] | ((unsigned)r1[8 * c + 7] << 16);
        d1[c * (TC_LBO / 16)] = q;
    }
}

IBL_API int ibl_feature_pack_create(const double *f, int64_t n, ibl_feature_pack_t **out, void *stream) {
    IBL_ARG(out && n > 0 && f && n < (1ll << 31));
    cudaStream_t st = ibl_stream(stream);
    int rc = tc_configure();
    if (rc < 0) return rc;
    ibl_feature_pack *pk = (ibl_feature_pack *)calloc(1, sizeof(ibl_feature_pack));
    pk->n = n; pk->n_pad = (n + TC_M - 1) / TC_M * TC_M; pk->f = f; pk->stream = st;
    const size_t img = (size_t)(pk->n_pad / TC_M) * TC_TILE_BYTES;
    IBL_CUDA(ibl_alloc(&pk->mu, 33, st));
    IBL_CUDA(ibl_alloc(&pk->mx, 4, st));
    IBL_CUDA(ibl_alloc(&pk->img0, img, st));
    IBL_CUDA(ibl_alloc(&pk->img1, img, st));
    IBL_CUDA(ibl_alloc(&pk->an32, (size_t)pk->n_pad, st));
    IBL_CUDA(ibl_alloc(&pk->rows32, (size_t)n * TC_R32, st));
    IBL_CUDA(ibl_alloc(&pk->nrm32, (size_t)n, st));
    IBL_CUDA(cudaMemsetAsync(pk->mx, 0, 4 * sizeof(unsigned), st));
    feat_tc_center_kernel<<<1, 33 * TC_MU_GROUPS, 0, st>>>(f, n, pk->mu);
    IBL_LAUNCH_CHECK();
    feat_tc_range_kernel<<<(unsigned)std::min<int64_t>((n * 33 + 255) / 256, 2 * IBL_SMS), 256, 0, st>>>(f, n, pk->mu, pk->mx + 2);
    IBL_LAUNCH_CHECK();
    feat_tc_headroom_kernel<<<1, 1, 0, st>>>(pk->mx + 2);
    IBL_LAUNCH_CHECK();
    uint8_t *img1o = nullptr;                              // target image in original order: only the source of the sorted one
    IBL_CUDA(ibl_alloc(&img1o, img, st));
    feat_tc_pack_both_kernel<<<(unsigned)(pk->n_pad / 128), 128, 0, st>>>(f, n, pk->n_pad, pk->mu, pk->mx + 2, pk->img0, img1o,
                                                                         pk->an32, pk->mx, pk->mx + 1, pk->rows32, pk->nrm32,
                                                                         pk->mx + 3);
    IBL_LAUNCH_CHECK();
    // target role: rows sorted by their scaled norm (non-negative floats order like their bit patterns), so that the rows of a
    // tile have similar norms and the tile's band term is tight
    const int ntiles = (int)(pk->n_pad / TC_M);
    IBL_CUDA(ibl_alloc(&pk->rows32s, (size_t)n * TC_R32, st));
    IBL_CUDA(ibl_alloc(&pk->nrm32s, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&pk->perm, (size_t)pk->n_pad, st));
    IBL_CUDA(ibl_alloc(&pk->tile_eb, (size_t)ntiles, st));
    unsigned *keys_out = nullptr, *tile_bn = nullptr; int *iota = nullptr;
    IBL_CUDA(ibl_alloc(&keys_out, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&iota, (size_t)n, st));
    IBL_CUDA(ibl_alloc(&tile_bn, (size_t)ntiles, st));
    IBL_CUDA(cudaMemsetAsync(tile_bn, 0, sizeof(unsigned) * ntiles, st));
    feat_tc_iota_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(iota, n);
    IBL_LAUNCH_CHECK();
    size_t tmp_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, reinterpret_cast<const unsigned *>(pk->an32), keys_out, iota, pk->perm, (int)n, 0, 32, st);
    void *tmp = nullptr;
    IBL_CUDA(cudaMallocAsync(&tmp, tmp_bytes ? tmp_bytes : 1, st));
    IBL_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, reinterpret_cast<const unsigned *>(pk->an32), keys_out, iota, pk->perm, (int)n, 0, 32, st));
    feat_tc_sort_rows_kernel<<<(unsigned)(pk->n_pad / 128), 128, 0, st>>>(n, pk->n_pad, pk->perm, img1o, pk->img1, pk->rows32, pk->nrm32,
                                                                         pk->rows32s, pk->nrm32s, pk->an32, tile_bn);
    IBL_LAUNCH_CHECK();
    feat_tc_tile_eb_kernel<<<(unsigned)((ntiles + 127) / 128), 128, 0, st>>>(tile_bn, pk->tile_eb, ntiles);
    IBL_LAUNCH_CHECK();
    ibl_free(tmp, st); ibl_free(keys_out, st); ibl_free(iota, st); ibl_free(tile_bn, st); ibl_free(img1o, st);
    *out = pk;
    return IBL_OK;
}

IBL_API void ibl_feature_pack_destroy(ibl_feature_pack_t *pk) {
    if (!pk) return;
    cudaStream_t st = pk->stream;
    ibl_free(pk->mu, st); ibl_free(pk->mx, st); ibl_free(pk->img0, st); ibl_free(pk->img1, st); ibl_free(pk->an32, st);
    ibl_free(pk->rows32, st); ibl_free(pk->nrm32, st); ibl_free(pk->rows32s, st); ibl_free(pk->nrm32s, st);
    ibl_free(pk->perm, st); ibl_free(pk->tile_eb, st);
    free(pk);
}

// Both directions of A.5 for a fresh feature set fa against a cached pack: nn_ab[i] = nearest row of the pack for row i of
// fa, nn_ba[j] = nearest row of fa for row j of the pack.  fa is read ONCE (feat_tc_pack_both_kernel, in the pack's
// coordinate system), the pack's rows only for exact evaluations.  d_flag (device uint32, optional) is OR-ed with 1 when fa
// does not fit the pack's fp16 range -- the results are then unreliable and the caller must use ibl_feature_nn.
IBL_API int ibl_feature_match_packed(const double *fa, int64_t na, const ibl_feature_pack_t *pk, int32_t *nn_ab, int32_t *nn_ba,
                                     uint32_t *d_flag, void *stream) {
    IBL_ARG(pk && na > 0 && fa && nn_ab && nn_ba && na < (1ll << 31));
    cudaStream_t st = ibl_stream(stream);
    int rc = tc_configure();
    if (rc < 0) return rc;
    const int64_t na_pad = (na + TC_M - 1) / TC_M * TC_M;
    const size_t img = (size_t)(na_pad / TC_M) * TC_TILE_BYTES;
    uint8_t *A0 = nullptr, *A1 = nullptr; float *an = nullptr, *a32 = nullptr, *na32 = nullptr; unsigned *mx = nullptr;
    IBL_CUDA(ibl_alloc(&A0, img, st));
    IBL_CUDA(ibl_alloc(&A1, img, st));
    IBL_CUDA(ibl_alloc(&an, (size_t)na_pad, st));
    IBL_CUDA(ibl_alloc(&a32, (size_t)na * TC_R32, st));
    IBL_CUDA(ibl_alloc(&na32, (size_t)na, st));
    IBL_CUDA(ibl_alloc(&mx, 3, st));
    IBL_CUDA(cudaMemsetAsync(mx, 0, 3 * sizeof(unsigned), st));
    {
        IBL_PROF(IBL_PROF_FEATURE_NN, st, 2.0 * 2.0 * 33.0 * (double)na * (double)pk->n);    // both directions
        feat_tc_pack_both_kernel<<<(unsigned)(na_pad / 128), 128, 0, st>>>(fa, na, na_pad, pk->mu, pk->mx + 2, A0, A1, an, mx, mx + 1,
                                                                          a32, na32, d_flag ? d_flag : mx + 2);
        IBL_LAUNCH_CHECK();
        rc = tc_search(A0, na, pk->img1, pk->n, an, pk->mx + 1, fa, pk->f, a32, pk->rows32s, na32, pk->nrm32s, nn_ab, st,
                       pk->tile_eb, pk->perm);
        if (rc == IBL_OK)
            rc = tc_search(pk->img0, pk->n, A1, na, pk->an32, mx + 1, pk->f, fa, pk->rows32, a32, pk->nrm32, na32, nn_ba, st);
    }
    ibl_free(A0, st); ibl_free(A1, st); ibl_free(an, st); ibl_free(a32, st); ibl_free(na32, st); ibl_free(mx, st);
    return rc;
}

// ---- mutual matching: the reverse direction only where the mutual filter reads it ----------------------------------------
// Open3D's mutual filter keeps (i, j = nn_ab[i]) iff nn_ba[j] == i (CorrespondencesFromFeatures): nn_ba is read ONLY at the
// targets some source row chose.  On the bench a source of ~15 k rows reaches a few thousand distinct rows of a ~37 k-row
// memory side, so the reverse search runs for those rows only: ONE kernel claims the chosen rows, compacts them (device-side
// count) and gathers their query-role image rows into whole tiles, then the search runs with the live row count on the device.
// One kernel: the first source row that names pack row j claims it (atomic exchange on a zeroed flag), takes the next slot of
// the compact list (device counter) and copies j's query-role image row there.  The list's ORDER depends on scheduling; no
// result does -- a row's search is independent of its slot, and results are scattered back through the list.  Slots past
// the count keep the zeros the image was cleared to: recycled scratch may hold fp16 Inf / NaN patterns, and an infinite score
// defeats the dead rows' "threshold = +inf" rule (inf >= inf) -- found as an illegal address under four streams.
__global__ void __launch_bounds__(128) feat_tc_claim_rows_kernel(const int32_t *__restrict__ nn_ab, int64_t na, int64_t nb,
                                                                 int *__restrict__ claimed, int *__restrict__ J, int *__restrict__ nJ,
                                                                 const uint8_t *__restrict__ img, uint8_t *__restrict__ out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= na) return;
    const int j = nn_ab[i];
    if (j < 0 || j >= nb) return;
    if (atomicExch(claimed + j, 1) != 0) return;
    const int k = atomicAdd(nJ, 1);
    J[k] = j;
    const uint4 *s = reinterpret_cast<const uint4 *>(img + tc_row_offset(j));
    uint4 *d = reinterpret_cast<uint4 *>(out + tc_row_offset(k));
#pragma unroll
    for (int c = 0; c < TC_KCH; ++c) d[c * (TC_LBO / 16)] = s[c * (TC_LBO / 16)];
}

// ibl_feature_match_packed for callers that apply the mutual filter: nn_ab as there; nn_ba[j] is the nearest row of fa for
// every pack row j that occurs in nn_ab and -1 for every other row (the filter never reads those).
IBL_API int ibl_feature_match_packed_mutual(const double *fa, int64_t na, const ibl_feature_pack_t *pk, int32_t *nn_ab,
                                            int32_t *nn_ba, uint32_t *d_flag, void *stream) {
    IBL_ARG(pk && na > 0 && fa && nn_ab && nn_ba && na < (1ll << 31));
    cudaStream_t st = ibl_stream(stream);
    int rc = tc_configure();
    if (rc < 0) return rc;
    const int64_t na_pad = (na + TC_M - 1) / TC_M * TC_M;
    const size_t img = (size_t)(na_pad / TC_M) * TC_TILE_BYTES;
    const int64_t nj_max = std::min<int64_t>(na, pk->n);           // at most one chosen pack row per source row
    const int64_t nj_pad = (nj_max + TC_M - 1) / TC_M * TC_M;
    uint8_t *A0 = nullptr, *A1 = nullptr, *Q = nullptr; float *an = nullptr, *a32 = nullptr, *na32 = nullptr;
    unsigned *mx = nullptr; int *J = nullptr, *nJ = nullptr, *claimed = nullptr;
    IBL_CUDA(ibl_alloc(&A0, img, st));
    IBL_CUDA(ibl_alloc(&A1, img, st));
    IBL_CUDA(ibl_alloc(&an, (size_t)na_pad, st));
    IBL_CUDA(ibl_alloc(&a32, (size_t)na * TC_R32, st));
    IBL_CUDA(ibl_alloc(&na32, (size_t)na, st));
    IBL_CUDA(ibl_alloc(&mx, 3, st));
    IBL_CUDA(ibl_alloc(&Q, (size_t)(nj_pad / TC_M) * TC_TILE_BYTES, st));
    IBL_CUDA(ibl_alloc(&claimed, (size_t)pk->n, st));
    IBL_CUDA(ibl_alloc(&J, (size_t)nj_pad, st));
    IBL_CUDA(ibl_alloc(&nJ, 1, st));
    IBL_CUDA(cudaMemsetAsync(mx, 0, 3 * sizeof(unsigned), st));
    IBL_CUDA(cudaMemsetAsync(claimed, 0, sizeof(int) * (size_t)pk->n, st));
    IBL_CUDA(cudaMemsetAsync(nJ, 0, sizeof(int), st));
    IBL_CUDA(cudaMemsetAsync(Q, 0, (size_t)(nj_pad / TC_M) * TC_TILE_BYTES, st));
    IBL_CUDA(cudaMemsetAsync(nn_ba, 0xff, sizeof(int32_t) * (size_t)pk->n, st));
    {
        // work figure: the forward direction only.  The reverse direction's row count lives on the device, so its flops are
        // NOT counted although its time is inside the scope: the reported rate is a lower bound.
        IBL_PROF(IBL_PROF_FEATURE_NN, st, 2.0 * 33.0 * (double)na * (double)pk->n);
        feat_tc_pack_both_kernel<<<(unsigned)(na_pad / 128), 128, 0, st>>>(fa, na, na_pad, pk->mu, pk->mx + 2, A0, A1, an, mx, mx + 1,
                                                                          a32, na32, d_flag ? d_flag : mx + 2);
        IBL_LAUNCH_CHECK();
        rc = tc_search(A0, na, pk->img1, pk->n, an, pk->mx + 1, fa, pk->f, a32, pk->rows32s, na32, pk->nrm32s, nn_ab, st,
                       pk->tile_eb, pk->perm);
        if (rc == IBL_OK) {
            feat_tc_claim_rows_kernel<<<(unsigned)((na + 127) / 128), 128, 0, st>>>(nn_ab, na, pk->n, claimed, J, nJ, pk->img0, Q);
            IBL_LAUNCH_CHECK();
            rc = tc_search(Q, nj_max, A1, na, pk->an32, mx + 1, pk->f, fa, pk->rows32, a32, pk->nrm32, na32, nn_ba, st,
                           nullptr, nullptr, nJ, J);
        }
    }
    ibl_free(A0, st); ibl_free(A1, st); ibl_free(an, st); ibl_free(a32, st); ibl_free(na32, st); ibl_free(mx, st);
    ibl_free(Q, st); ibl_free(claimed, st); ibl_free(J, st); ibl_free(nJ, st);
    return rc;
}

// Debug / test hook: the raw tensor-core scores s'(a_i, b_j) = a32.b32 - |b32|^2/2 as float32 [na][nb] (row-major), where
// a32 / b32 are the rows after subtracting the filter's centre and multiplying by its power-of-two scale (centre and
// scale are written to center_out, device double[34], if not NULL).
// swap_lbo_sbo = 1 exchanges the two stride fields of the shared-memory descriptors (layout probe; 0 is the product setting).
IBL_API int ibl_feature_scores_debug(const double *fa, int64_t na, const double *fb, int64_t nb, float *scores,
                                     double *center_out, int swap_lbo_sbo, void *stream) {
    IBL_ARG(fa && fb && scores && na > 0 && nb > 0);
    cudaStream_t st = ibl_stream(stream);
    int rc = tc_configure();
    if (rc < 0) return rc;
    TcPlan p = tc_plan(na, nb);
    uint8_t *A = nullptr, *B = nullptr;
    float *an = nullptr;
    unsigned *mx = nullptr;
    IBL_CUDA(ibl_alloc(&A, (size_t)p.qtiles * TC_TILE_BYTES, st));
    IBL_CUDA(ibl_alloc(&B, (size_t)p.btiles * TC_TILE_BYTES, st));
    IBL_CUDA(ibl_alloc(&an, (size_t)p.na_pad, st));
    IBL_CUDA(ibl_alloc(&mx, 3, st));                       // max |a|^2, max |b|^2, max |f - mu| (float bits)
    IBL_CUDA(cudaMemsetAsync(mx, 0, 3 * sizeof(unsigned), st));
    double *mu = center_out;
    if (!mu) IBL_CUDA(ibl_alloc(&mu, 33, st));
    feat_tc_center_kernel<<<1, 33 * TC_MU_GROUPS, 0, st>>>(fb, nb, mu);
    IBL_LAUNCH_CHECK();
    feat_tc_range_kernel<<<(unsigned)std::min<int64_t>((na * 33 + 255) / 256, 2 * IBL_SMS), 256, 0, st>>>(fa, na, mu, mx + 2);
    IBL_LAUNCH_CHECK();
    feat_tc_range_kernel<<<(unsigned)std::min<int64_t>((nb * 33 + 255) / 256, 2 * IBL_SMS), 256, 0, st>>>(fb, nb, mu, mx + 2);
    IBL_LAUNCH_CHECK();
    feat_tc_pack_kernel<<<(unsigned)(p.na_pad / 128), 128, 0, st>>>(fa, na, p.na_pad, 0, mu, mx + 2, A, an, mx);
    IBL_LAUNCH_CHECK();
    feat_tc_pack_kernel<<<(unsigned)(p.nb_pad / 128), 128, 0, st>>>(fb, nb, p.nb_pad, 1, mu, mx + 2, B, nullptr, mx + 1);
    IBL_LAUNCH_CHECK();
    dim3 grid((p.qtiles + TC_QT - 1) / TC_QT, p.nsplit);
    feat_tc_kernel<0><<<grid, TC_THREADS, TC_SMEM, st>>>(A, na, p.qtiles, B, nb, p.btiles, p.tiles_per_split, p.nsplit, an, mx + 1,
                                                        fa, fb, nullptr, TcPool{}, scores, nb, swap_lbo_sbo, nullptr, nullptr, nullptr, nullptr);
    IBL_LAUNCH_CHECK();
    if (center_out) {                                     // debug hook only: read the scale back and publish it
        unsigned bits = 0;
        IBL_CUDA(cudaMemcpyAsync(&bits, mx + 2, sizeof(bits), cudaMemcpyDeviceToHost, st));
        IBL_CUDA(ibl_sync(st));
        float m;
        memcpy(&m, &bits, sizeof(m));
        double sc = 1.0;
        if (m > 0.f && std::isfinite(m)) { int x; frexp((double)m, &x); sc = ldexp(1.0, 10 - x); }
        IBL_CUDA(cudaMemcpyAsync(center_out + 33, &sc, sizeof(sc), cudaMemcpyHostToDevice, st));
        IBL_CUDA(ibl_sync(st));
    } else {
        ibl_free(mu, st);
    }
    ibl_free(A, st); ibl_free(B, st); ibl_free(an, st); ibl_free(mx, st);
    return IBL_OK;
}

// Instrumentation: the wait clocks of feat_tc_kernel (g_tc_wait above).  mode 1: zero the counters and switch collection on;
// mode 0: read them into h_out8 and switch collection off.  Syncs.
IBL_API int ibl_feature_nn_wait_stats(int64_t *h_out8, int mode) {
    IBL_CUDA(cudaDeviceSynchronize());
    unsigned long long v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int on = mode ? 1 : 0;
    if (mode) {
        IBL_CUDA(cudaMemcpyToSymbol(g_tc_wait, v, sizeof(v)));
    } else {
        IBL_ARG(h_out8 != nullptr);
        IBL_CUDA(cudaMemcpyFromSymbol(v, g_tc_wait, sizeof(v)));
        for (int k = 0; k < 8; ++k) h_out8[k] = (int64_t)v[k];
    }
    IBL_CUDA(cudaMemcpyToSymbol(g_tc_wait_on, &on, sizeof(on)));
    return IBL_OK;
}

// Instrumentation: {band candidates, exact f64 evaluations} of ibl_feature_nn accumulated since the last reset.  Syncs.
IBL_API int ibl_feature_nn_stats(int64_t *h_out2, int reset) {
    IBL_ARG(h_out2 != nullptr);
    unsigned long long v[2] = {0, 0};
    IBL_CUDA(cudaDeviceSynchronize());
    IBL_CUDA(cudaMemcpyFromSymbol(v, g_tc_stats, sizeof(v)));
    h_out2[0] = (int64_t)v[0]; h_out2[1] = (int64_t)v[1];
    if (reset) { unsigned long long z[2] = {0, 0}; IBL_CUDA(cudaMemcpyToSymbol(g_tc_stats, z, sizeof(z))); }
    return IBL_OK;
}
