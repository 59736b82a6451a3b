// prof.cu -- launch counter and optional per-kernel CUDA-event timing (bench.py's roofline object is measured
// live with these events on the launching stream; see include/ibl.h "instrumentation").
#include <mutex>
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
    "nn_score", "icp_reduce", "color_gradient", "grid_build", "simvolume", "cosine_segmax", "voxel", "mean3", "misc"};

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
void ibl_prof_end(int slot, cudaStream_t st) {
    if (slot < 0) return;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (slot < (int)g_recs.size()) cudaEventRecord(g_recs[slot].e1, st);
}

IBL_API int ibl_set_host_wait(int mode) {
    IBL_ARG(mode >= 0 && mode <= 3);
    static const unsigned flags[4] = {cudaDeviceScheduleAuto, cudaDeviceScheduleSpin, cudaDeviceScheduleYield,
                                      cudaDeviceScheduleBlockingSync};
    IBL_CUDA(cudaSetDeviceFlags(flags[mode]));
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
    for (int t = 0; t < IBL_PROF_COUNT; ++t) { ms[t] = 0; launches[t] = 0; bytes[t] = 0; }
    for (auto &r : g_recs) {
        float f = 0.f;
        if (cudaEventElapsedTime(&f, r.e0, r.e1) == cudaSuccess) { ms[r.tag] += f; launches[r.tag] += 1; bytes[r.tag] += r.bytes; }
        g_pool.push_back(r.e0); g_pool.push_back(r.e1);
    }
    g_recs.clear();
    (void)cudaGetLastError();
    return IBL_OK;
}
