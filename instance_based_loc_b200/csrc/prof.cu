// prof.cu -- launch counter and optional per-kernel CUDA-event timing (bench.py's roofline object is measured
// live with these events on the launching stream; see include/ibl.h "instrumentation").
#include <algorithm>
#include <mutex>
#include <stdlib.h>
#include <sys/prctl.h>
#include <time.h>
#include <vector>

#include "ibl_common.cuh"

unsigned long long g_ibl_launches = 0;
static int g_prof_on = 0;
static std::mutex g_prof_mu;
struct ProfRec { int tag; cudaEvent_t e0, e1; double bytes; };
static std::vector<ProfRec> g_recs;
static std::vector<cudaEvent_t> g_pool;

static const char *kTagNames[IBL_PROF_COUNT] = {
    "backproject", "radius_count", "knn", "normals", "spfh", "fpfh", "feature_nn", "ransac_check", "ransac_score",
    "nn_score", "icp_reduce", "color_gradient", "grid_build", "simvolume", "cosine_segmax", "voxel", "mean3", "misc",
    "feature_gemm"};

static cudaEvent_t get_event() {
    if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

int ibl_prof_begin(int tag, cudaStream_t st, double bytes) {
    if (!g_prof_on) return -1;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    ProfRec r;
    r.tag = tag; r.bytes = bytes; r.e0 = get_event(); r.e1 = get_event();
    cudaEventRecord(r.e0, st);
    g_recs.push_back(r);
    return (int)g_recs.size() - 1;
}
// work of a launch whose size was only known on the device when it was queued (added to the tag's total at the next read)
static double g_extra_work[IBL_PROF_COUNT];
void ibl_prof_add_work(int tag, double bytes) {
    if (!g_prof_on || tag < 0 || tag >= IBL_PROF_COUNT) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_extra_work[tag] += bytes;
}
void ibl_prof_end(int slot, cudaStream_t st) {
    if (slot < 0) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (slot < (int)g_recs.size()) cudaEventRecord(g_recs[slot].e1, st);
}

// mode 4 ("hybrid"): this library's own stream waits poll cudaStreamQuery for g_spin_us microseconds -- the ICP / RANSAC
// round trips are 20-40 us, so they still complete at spin latency -- and then sleep in short steps, so the long waits
// (feature matching, k-NN: 0.3-2 ms) stop burning a core each.
static int g_hybrid = 0;
static int g_spin_us = 60;
cudaError_t ibl_sync(cudaStream_t st) {
    if (!g_hybrid) return cudaStreamSynchronize(st);
    timespec t0, t;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (;;) {
        cudaError_t e = cudaStreamQuery(st);
        if (e != cudaErrorNotReady) return e;
        clock_gettime(CLOCK_MONOTONIC, &t);
        long us = (t.tv_sec - t0.tv_sec) * 1000000L + (t.tv_nsec - t0.tv_nsec) / 1000L;
        if (us < g_spin_us) { __builtin_ia32_pause(); continue; }
        timespec nap = {0, us < 400 ? 20000L : 50000L};
        nanosleep(&nap, nullptr);
    }
}

IBL_API int ibl_set_host_wait(int mode) {
    IBL_ARG(mode >= 0 && mode <= 4);
    static const unsigned flags[5] = {cudaDeviceScheduleAuto, cudaDeviceScheduleSpin, cudaDeviceScheduleYield,
                                      cudaDeviceScheduleBlockingSync, cudaDeviceScheduleAuto};
    IBL_CUDA(cudaSetDeviceFlags(flags[mode]));
    g_hybrid = mode == 4;
    if (g_hybrid) {
        prctl(PR_SET_TIMERSLACK, 1000UL, 0, 0, 0);          // this thread; worker threads set theirs on their first wait
        if (const char *e = getenv("IBL_WAIT_SPIN_US")) g_spin_us = atoi(e);
    }
    return IBL_OK;
}

// Grow the stream-ordered scratch pool by `bytes` now (allocate + free on `stream`): the pool keeps what it has ever
// held (release threshold = max), so a frame stream that reserves its working set once never waits for the driver to
// map memory in the middle of a frame.  Memory freed on one stream is reused by another only opportunistically, so
// every frame stream reserves for itself.
IBL_API int ibl_reserve_scratch(int64_t bytes, void *stream) {
    IBL_ARG(bytes >= 0);
    if (bytes == 0) return IBL_OK;
    cudaStream_t st = ibl_stream(stream);
    // several blocks rather than one: the allocator splits blocks but does not merge across separate mappings
    const int64_t piece = 512ll << 20;
    std::vector<void *> held;
    for (int64_t left = bytes; left > 0; left -= piece) {
        unsigned char *p = nullptr;
        IBL_CUDA(ibl_alloc(&p, (size_t)std::min(left, piece), st));
        held.push_back(p);
    }
    for (void *p : held) ibl_free(p, st);
    return IBL_OK;
}

// {reserved now, reserved high-water, used now, used high-water} bytes of the device's default stream-ordered pool
IBL_API int ibl_scratch_stats(int64_t *h_out4) {
    IBL_ARG(h_out4 != nullptr);
    int dev = 0;
    cudaMemPool_t pool;
    IBL_CUDA(cudaGetDevice(&dev));
    IBL_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    const cudaMemPoolAttr attrs[4] = {cudaMemPoolAttrReservedMemCurrent, cudaMemPoolAttrReservedMemHigh,
                                      cudaMemPoolAttrUsedMemCurrent, cudaMemPoolAttrUsedMemHigh};
    for (int k = 0; k < 4; ++k) {
        unsigned long long v = 0;
        IBL_CUDA(cudaMemPoolGetAttribute(pool, attrs[k], &v));
        h_out4[k] = (int64_t)v;
    }
    return IBL_OK;
}

IBL_API int64_t ibl_launch_count(void) { return (int64_t)__sync_fetch_and_add(&g_ibl_launches, 0ull); }
IBL_API int ibl_prof_num_tags(void) { return IBL_PROF_COUNT; }
IBL_API const char *ibl_prof_tag_name(int tag) { return (tag >= 0 && tag < IBL_PROF_COUNT) ? kTagNames[tag] : ""; }
IBL_API void ibl_prof_enable(int on) { g_prof_on = on; }

// ms / launches / bytes: arrays of ibl_prof_num_tags() entries, accumulated since the last read.  Syncs the device.
IBL_API int ibl_prof_read(double *ms, int64_t *launches, double *bytes) {
    IBL_ARG(ms && launches && bytes);
    IBL_CUDA(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (int t = 0; t < IBL_PROF_COUNT; ++t) { ms[t] = 0; launches[t] = 0; bytes[t] = g_extra_work[t]; g_extra_work[t] = 0; }
    for (auto &r : g_recs) {
        float f = 0.f;
        if (cudaEventElapsedTime(&f, r.e0, r.e1) == cudaSuccess) { ms[r.tag] += f; launches[r.tag] += 1; bytes[r.tag] += r.bytes; }
        g_pool.push_back(r.e0); g_pool.push_back(r.e1);
    }
    g_recs.clear();
    (void)cudaGetLastError();
    return IBL_OK;
}
