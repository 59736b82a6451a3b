// simvol.cu -- embedding similarity (object_memory/object_memory.py:922-936) and the SimVolume
// assignment search (utils/similarity_volume.py:13-18,102-164,213-270).
//
// SimVolume never materialises the (m+1)^s fp16 volume: entries are recomputed on the fly
// (two fp16 roundings, left to right) and the k best per sub-volume, ordered by (value desc, flat
// index asc) exactly like the reference's repeated argmax, are found by a 5-digit radix select
// over the key (value-hi8, value-lo8, i, j, k) followed by one collect pass and a CTA bitonic sort.
#include <cuda_fp16.h>

#include "ibl_common.cuh"

// ------------------------------------------------------------------------------------------------
// cosine seg-max.  CTA per memory object; warps stride its stored embeddings; 8 detections per pass.
__global__ void __launch_bounds__(128) l2norm_rows_kernel(const float *__restrict__ x, int n, int D, float *__restrict__ out) {
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (row >= n) return;
    float s = 0.f;
    for (int d = lane; d < D; d += 32) { float v = x[(size_t)row * D + d]; s += v * v; }
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    float nrm = sqrtf(s);
    for (int d = lane; d < D; d += 32) out[(size_t)row * D + d] = x[(size_t)row * D + d] / nrm;
}

#define CS_DET 8
__global__ void __launch_bounds__(128) cosine_segmax_kernel(const float *__restrict__ detn, int nd,
                                                            const float *__restrict__ mem, int D,
                                                            const int32_t *__restrict__ seg, float *__restrict__ S, int m) {
    __shared__ float wmax[4][CS_DET];
    int j = blockIdx.x;
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int e0 = seg[j], e1 = seg[j + 1];
    for (int i0 = 0; i0 < nd; i0 += CS_DET) {
        float best[CS_DET];
#pragma unroll
        for (int t = 0; t < CS_DET; ++t) best[t] = -INFINITY;
        for (int e = e0 + w; e < e1; e += 4) {
            const float *row = mem + (size_t)e * D;
            float ss = 0.f;
            for (int d = lane; d < D; d += 32) { float v = row[d]; ss += v * v; }
            for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            float nrm = sqrtf(ss);
            float acc[CS_DET];
#pragma unroll
            for (int t = 0; t < CS_DET; ++t) acc[t] = 0.f;
            for (int d = lane; d < D; d += 32) {
                float v = row[d] / nrm;
#pragma unroll
                for (int t = 0; t < CS_DET; ++t)
                    if (i0 + t < nd) acc[t] += v * detn[(size_t)(i0 + t) * D + d];
            }
#pragma unroll
            for (int t = 0; t < CS_DET; ++t) {
                float a = acc[t];
                for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
                best[t] = fmaxf(best[t], a);
            }
        }
        if (lane == 0) {
#pragma unroll
            for (int t = 0; t < CS_DET; ++t) wmax[w][t] = best[t];
        }
        __syncthreads();
        if (threadIdx.x < CS_DET && i0 + threadIdx.x < nd) {
            float b = fmaxf(fmaxf(wmax[0][threadIdx.x], wmax[1][threadIdx.x]), fmaxf(wmax[2][threadIdx.x], wmax[3][threadIdx.x]));
            S[(size_t)(i0 + threadIdx.x) * m + j] = b;
        }
        __syncthreads();
    }
}

IBL_API int ibl_cosine_segmax(const float *det, int32_t nd, const float *mem, int64_t E, int32_t D, const int32_t *seg,
                              int32_t m, float *S, void *stream) {
    IBL_ARG(nd >= 0 && m >= 0 && D > 0 && E >= 0);
    if (nd == 0 || m == 0) return IBL_OK;
    IBL_ARG(det && mem && seg && S);
    cudaStream_t st = ibl_stream(stream);
    float *detn = nullptr;
    IBL_CUDA(ibl_alloc(&detn, (size_t)nd * D, st));
    l2norm_rows_kernel<<<(nd + 3) / 4, 128, 0, st>>>(det, nd, D, detn);
    IBL_LAUNCH_CHECK();
    IBL_PROF(IBL_PROF_COSINE, st, 4.0 * D * ((double)E + nd) + 4.0 * nd * m);
    cosine_segmax_kernel<<<m, 128, 0, st>>>(detn, nd, mem, D, seg, S, m);
    IBL_LAUNCH_CHECK();
    ibl_free(detn, st);
    return IBL_OK;
}

// ------------------------------------------------------------------------------------------------
// SimVolume top-k
#define SV_MAXBINS 1024
struct SvState {
    unsigned long long total;   // valid entries
    int k_eff;                  // min(k, total)
    int remaining;              // picks still to place at the current digit
    int sel[5];                 // selected digit per level (hi8, lo8, i, j, k)
    int out_count;
};

__device__ __forceinline__ unsigned ord16(__half h) {
    unsigned b = __half_as_ushort(h);
    return (b & 0x8000u) ? (~b & 0xffffu) : (b | 0x8000u);
}
__device__ __forceinline__ float ord16_to_float(unsigned o) {
    unsigned b = (o & 0x8000u) ? (o & 0x7fffu) : (~o & 0xffffu);
    return __half2float(__ushort_as_half((unsigned short)b));
}

// sub-volume geometry: s = number of chosen detections (1..3); axes beyond s have extent 1
struct SvGeom { int nd, m, s, n0, n1, n2, ncomb; };

__device__ __forceinline__ void comb_rows(int nd, int s, int c, int *r) {
    // lexicographic combination #c of s rows out of nd
    if (s == 1) { r[0] = 0; r[1] = r[2] = -1; return; }
    int idx = 0;
    if (s == 2) {
        for (int a = 0; a < nd; ++a) for (int b = a + 1; b < nd; ++b) { if (idx == c) { r[0] = a; r[1] = b; r[2] = -1; return; } ++idx; }
    } else {
        for (int a = 0; a < nd; ++a) for (int b = a + 1; b < nd; ++b) for (int d = b + 1; d < nd; ++d) {
            if (idx == c) { r[0] = a; r[1] = b; r[2] = d; return; } ++idx; }
    }
    r[0] = r[1] = r[2] = 0;
}

// A lower bound on the k-th largest entry of every sub-volume whose FIRST detection is row d: the entries (i, unassigned,
// unassigned) are valid for every memory object i and equal row_d[i] exactly (x 1.0 in fp16 changes nothing), so the k-th
// largest value of the row is reached by at least k entries.  One CTA per detection row: bitonic sort of the row's ordered
// fp16 keys.  lb[d] = 0 (no bound) when the row has fewer than k entries.
__global__ void __launch_bounds__(256) sv_lower_bound_kernel(const float *__restrict__ S, int m, int k, unsigned *__restrict__ lb) {
    __shared__ unsigned key[SV_MAXBINS];
    const int d = blockIdx.x;
    int n2 = 32;
    while (n2 < m) n2 <<= 1;
    for (int t = threadIdx.x; t < n2; t += blockDim.x) key[t] = t < m ? ord16(__float2half_rn(S[(size_t)d * m + t])) : 0u;
    __syncthreads();
    for (int kk = 2; kk <= n2; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
                int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                int p = i | j;
                bool desc = (i & kk) == 0;
                unsigned a = key[i], b = key[p];
                if ((a < b) == desc) { key[i] = b; key[p] = a; }
            }
            __syncthreads();
        }
    if (threadIdx.x == 0) lb[d] = m >= k ? key[k - 1] : 0u;
}

// level 0..4 histogram passes and level 5 = collect
__global__ void __launch_bounds__(256) sv_pass_kernel(const float *__restrict__ S, SvGeom G, int level,
                                                      const SvState *states, int *__restrict__ hist, int bins,
                                                      unsigned long long *__restrict__ out_keys, SvState *wstates,
                                                      int kcap, const unsigned *__restrict__ row_lb) {
    __shared__ __half row0[SV_MAXBINS], row1[SV_MAXBINS], row2[SV_MAXBINS];
    __shared__ int sh[SV_MAXBINS];
    __shared__ unsigned long long stotal;
    int sub = blockIdx.y;
    int r[3];
    comb_rows(G.nd, G.s, sub, r);
    int m = G.m;
    for (int t = threadIdx.x; t <= m; t += blockDim.x) {
        row0[t] = t < m ? __float2half_rn(S[(size_t)r[0] * m + t]) : __float2half_rn(1.0f);
        row1[t] = (G.s >= 2 && t < m) ? __float2half_rn(S[(size_t)r[1] * m + t]) : __float2half_rn(1.0f);
        row2[t] = (G.s >= 3 && t < m) ? __float2half_rn(S[(size_t)r[2] * m + t]) : __float2half_rn(1.0f);
    }
    for (int t = threadIdx.x; t < bins; t += blockDim.x) sh[t] = 0;
    if (threadIdx.x == 0) stotal = 0;
    __syncthreads();
    // extreme values of the last axis (its "unassigned" entry 1.0 included): with them a pair (i, j) bounds every entry
    // of its k-line -- fp16 multiplication is monotone in each factor, so max_k f16(p * c_k) is f16(p * max c) for
    // p >= 0 and f16(p * min c) for p < 0
    __shared__ float s_ext[2];
    if (threadIdx.x < 32) {
        float mx = -INFINITY, mn = INFINITY;
        for (int t = threadIdx.x; t < G.n2; t += 32) { float v = __half2float(row2[t < m ? t : m]); mx = fmaxf(mx, v); mn = fminf(mn, v); }
        if (G.s < 3) { mx = 1.0f; mn = 1.0f; }
        for (int o2 = 16; o2; o2 >>= 1) { mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o2)); mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o2)); }
        if (threadIdx.x == 0) { s_ext[0] = mx; s_ext[1] = mn; }
    }
    __syncthreads();
    const __half c_max = __float2half_rn(s_ext[0]), c_min = __float2half_rn(s_ext[1]);
    SvState stt = states[sub];
    const bool all_invalid = m < G.s;
    int lane = threadIdx.x & 31;
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int nwarps = (gridDim.x * blockDim.x) >> 5;
    int npairs = m * G.n1;   // i in [0,m) (first detection can never be unassigned), j in [0,n1)
    unsigned vstar = ((unsigned)stt.sel[0] << 8) | (unsigned)stt.sel[1];
    // Entries below `floor_o` cannot matter at this level: levels 0-1 look for the value of the k-th largest entry, which is
    // at least the row bound (sv_lower_bound_kernel); levels 2-5 only touch entries equal to (or, collecting, at least)
    // that value.  A pair whose k-line cannot reach the floor is skipped whole -- at m = 500 that leaves a few hundred of
    // the 250 000 pairs.  (The count of ALL valid entries is only needed when there may be fewer than k of them, i.e. when
    // there is no row bound; then nothing is skipped at level 0.)
    const unsigned row_floor = row_lb ? row_lb[r[0]] : 0u;
    const unsigned floor_o = level <= 1 ? row_floor : vstar;
    unsigned long long local_total = 0;
    for (int pi = warp; pi < npairs && !all_invalid; pi += nwarps) {
        int i = pi / G.n1, j = pi - i * G.n1;
        if (G.s >= 2 && j != m && j == i) continue;                 // repeated memory object
        if (level >= 3 && level <= 4) { if (i != stt.sel[2]) continue; }
        if (level == 4 && j != stt.sel[3]) continue;
        __half p = (G.s >= 2) ? __hmul(row0[i], row1[j]) : row0[i];
        if (floor_o) {
            const __half best = (G.s >= 3) ? __hmax(__hmul(p, c_max), __hmul(p, c_min)) : p;
            if (!__hisnan(best) && ord16(best) < floor_o) continue;
        }
        for (int k0 = 0; k0 < G.n2; k0 += 32) {
            int k = k0 + lane;
            bool valid = k < G.n2;
            unsigned o = 0;
            if (valid && G.s >= 3) {
                valid = (k == m) || (k != i && k != j);
                __half v = __hmul(p, row2[k]);
                o = ord16(v);
                if (__hisnan(v) || __hisinf(v) && (__half_as_ushort(v) & 0x8000u)) valid = false;
            } else if (valid) {
                o = ord16(p);
                if (__hisnan(p) || __hisinf(p) && (__half_as_ushort(p) & 0x8000u)) valid = false;
            }
            int digit = -1;
            if (valid) {
                switch (level) {
                    case 0: digit = (int)(o >> 8); break;
                    case 1: if ((int)(o >> 8) == stt.sel[0]) digit = (int)(o & 0xffu); break;
                    case 2: if (o == vstar) digit = i; break;
                    case 3: if (o == vstar) digit = j; break;
                    case 4: if (o == vstar) digit = k; break;
                    default: break;
                }
            }
            if (level == 0) local_total += valid;
            if (level <= 4) {
                unsigned act = __ballot_sync(0xffffffffu, digit >= 0);
                if (digit >= 0) {
                    unsigned peers = __match_any_sync(act, digit);
                    if ((int)(__ffs(peers) - 1) == lane) atomicAdd(&sh[digit], __popc(peers));
                }
            } else if (valid) {   // collect: strictly better value, or equal value and flat <= flat*
                long long flat = ((long long)i * G.n1 + j) * G.n2 + k;
                long long fstar = ((long long)stt.sel[2] * G.n1 + stt.sel[3]) * G.n2 + stt.sel[4];
                if (stt.k_eff > 0 && (o > vstar || (o == vstar && flat <= fstar))) {
                    int pos = atomicAdd(&wstates[sub].out_count, 1);
                    if (pos < kcap)
                        out_keys[(size_t)sub * kcap + pos] = ((unsigned long long)o << 40) | (unsigned long long)((1ll << 40) - 1 - flat);
                }
            }
        }
    }
    if (level == 0) {
        for (int o2 = 16; o2; o2 >>= 1) local_total += __shfl_xor_sync(0xffffffffu, local_total, o2);
        if (lane == 0 && local_total) atomicAdd(&stotal, local_total);
    }
    __syncthreads();
    if (level <= 4)
        for (int t = threadIdx.x; t < bins; t += blockDim.x) if (sh[t]) atomicAdd(&hist[(size_t)sub * SV_MAXBINS + t], sh[t]);
    if (level == 0 && threadIdx.x == 0 && stotal) atomicAdd(&wstates[sub].total, stotal);
}

// one thread per sub-volume: walk the histogram from the best bin and fix this digit
__global__ void sv_select_kernel(int *__restrict__ hist, int bins, int level, int k, SvState *__restrict__ states, int ncomb,
                                 int row_bounded) {
    int sub = blockIdx.x * blockDim.x + threadIdx.x;
    if (sub >= ncomb) return;
    SvState st = states[sub];
    int *h = hist + (size_t)sub * SV_MAXBINS;
    if (level == 0) {
        if (st.total < (unsigned long long)k && row_bounded) st.total = (unsigned long long)k;   // pruned count: >= k entries exist
        st.k_eff = (int)(st.total < (unsigned long long)k ? st.total : (unsigned long long)k);
        st.remaining = st.k_eff;
        st.out_count = 0;
    }
    int sel = 0;
    if (st.k_eff > 0) {
        int rem = st.remaining;
        if (level <= 1) {          // larger digit = better value
            for (int b = bins - 1; b >= 0; --b) { if (h[b] >= rem) { sel = b; break; } rem -= h[b]; }
        } else {                   // smaller index = better (first flat index wins ties)
            for (int b = 0; b < bins; ++b) { if (h[b] >= rem) { sel = b; break; } rem -= h[b]; }
        }
        st.remaining = rem;
    }
    st.sel[level] = sel;
    states[sub] = st;
    for (int b = 0; b < bins; ++b) h[b] = 0;
}

// CTA per sub-volume: bitonic sort (descending key) of the collected entries, decode
__global__ void __launch_bounds__(256) sv_sort_kernel(const unsigned long long *__restrict__ keys, int kcap, int k,
                                                      const SvState *__restrict__ states, int64_t *__restrict__ out_flat,
                                                      float *__restrict__ out_val) {
    __shared__ unsigned long long sk[1024];
    int sub = blockIdx.x;
    int ke = states[sub].k_eff;
    int n2 = 32;
    while (n2 < ke) n2 <<= 1;
    for (int t = threadIdx.x; t < n2; t += blockDim.x) sk[t] = t < ke ? keys[(size_t)sub * kcap + t] : 0ull;
    __syncthreads();
    for (int kk = 2; kk <= n2; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < (n2 >> 1); t += blockDim.x) {
                int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                int p = i | j;
                bool desc = (i & kk) == 0;
                unsigned long long a = sk[i], b = sk[p];
                if ((a < b) == desc) { sk[i] = b; sk[p] = a; }
            }
            __syncthreads();
        }
    for (int t = threadIdx.x; t < k; t += blockDim.x) {
        if (t < ke) {
            unsigned long long key = sk[t];
            out_flat[(size_t)sub * k + t] = (int64_t)(((1ll << 40) - 1) - (long long)(key & ((1ull << 40) - 1)));
            out_val[(size_t)sub * k + t] = ord16_to_float((unsigned)(key >> 40));
        } else {   // everything left is -inf: np.argmax keeps returning flat index 0
            out_flat[(size_t)sub * k + t] = 0;
            out_val[(size_t)sub * k + t] = -INFINITY;
        }
    }
}

// 1 (default): pairs whose k-line cannot reach the current floor are skipped (exact); 0: every pass walks the whole
// sub-volume (measurement / test switch)
static int g_sv_prune = 1;
IBL_API int ibl_simvolume_set_prune(int on) { g_sv_prune = on ? 1 : 0; return IBL_OK; }

IBL_API int ibl_simvolume_topk(const float *S, int32_t nd, int32_t m, int32_t k, int64_t *out_flat, float *out_val,
                               void *stream) {
    IBL_ARG(S && out_flat && out_val && nd >= 1 && m >= 1 && k >= 1 && k <= 1024 && m + 1 <= SV_MAXBINS);
    cudaStream_t st = ibl_stream(stream);
    SvGeom G;
    G.nd = nd; G.m = m; G.s = nd < 3 ? nd : 3;
    G.n0 = m + 1; G.n1 = G.s >= 2 ? m + 1 : 1; G.n2 = G.s >= 3 ? m + 1 : 1;
    G.ncomb = G.s == 1 ? 1 : (G.s == 2 ? nd * (nd - 1) / 2 : nd * (nd - 1) * (nd - 2) / 6);
    SvState *states = nullptr; int *hist = nullptr; unsigned long long *keys = nullptr;
    IBL_CUDA(ibl_alloc(&states, (size_t)G.ncomb, st));
    IBL_CUDA(ibl_alloc(&hist, (size_t)G.ncomb * SV_MAXBINS, st));
    IBL_CUDA(ibl_alloc(&keys, (size_t)G.ncomb * k, st));
    IBL_CUDA(cudaMemsetAsync(states, 0, sizeof(SvState) * G.ncomb, st));
    IBL_CUDA(cudaMemsetAsync(hist, 0, sizeof(int) * (size_t)G.ncomb * SV_MAXBINS, st));
    long long npairs = (long long)m * G.n1;
    int bx = (int)((npairs + 7) / 8);            // 8 warps per CTA, one pair per warp-iteration
    int cap = (IBL_SMS * 8 + G.ncomb - 1) / G.ncomb;
    if (bx > cap) bx = cap;
    if (bx < 1) bx = 1;
    dim3 grid(bx, G.ncomb);
    IBL_PROF(IBL_PROF_SIMVOL, st, 6.0 * (double)G.ncomb * m * G.n1 * G.n2);
    // row bounds (a pruning floor) exist when every row holds at least k values
    unsigned *row_lb = nullptr;
    const int row_bounded = (m >= k && g_sv_prune) ? 1 : 0;
    if (row_bounded) {
        IBL_CUDA(ibl_alloc(&row_lb, (size_t)nd, st));
        sv_lower_bound_kernel<<<nd, 256, 0, st>>>(S, m, k, row_lb);
        IBL_LAUNCH_CHECK();
    }
    for (int level = 0; level <= 4; ++level) {
        int bins = level <= 1 ? 256 : m + 1;
        sv_pass_kernel<<<grid, 256, 0, st>>>(S, G, level, states, hist, bins, keys, states, k, g_sv_prune ? row_lb : nullptr);
        IBL_LAUNCH_CHECK();
        sv_select_kernel<<<(G.ncomb + 63) / 64, 64, 0, st>>>(hist, bins, level, k, states, G.ncomb, row_bounded);
        IBL_LAUNCH_CHECK();
    }
    sv_pass_kernel<<<grid, 256, 0, st>>>(S, G, 5, states, hist, 1, keys, states, k, g_sv_prune ? row_lb : nullptr);
    IBL_LAUNCH_CHECK();
    ibl_free(row_lb, st);
    sv_sort_kernel<<<G.ncomb, 256, 0, st>>>(keys, k, k, states, out_flat, out_val);
    IBL_LAUNCH_CHECK();
    ibl_free(states, st); ibl_free(hist, st); ibl_free(keys, st);
    return IBL_OK;
}
