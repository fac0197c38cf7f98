// Library-wide plumbing: version, thread-local error text, launch accounting, per-stage timers.
#include <atomic>
#include <vector>

#include "common.cuh"

namespace die {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

char* err_buf() { return g_err; }
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// Optional per-stage device timing: while enabled every kernel launch is bracketed by two CUDA
// events recorded on the kernel's own stream; die_profile_end() synchronises and sums them.
static const char* const kStageNames[ST_COUNT] = {
    "route_md5",   "kv_get",      "kv_contains", "kv_delete",  "kv_purge", "kv_set_precheck",
    "kv_set_window", "kv_set_claim", "kv_set_finalize", "kv_set_evict", "kv_maintenance", "miss_scan",
    "pack_hits",   "model",       "unpack_store", "pack_ragged", "unpack_ragged", "bucket_by_shard"};

struct ProfRec {
    int stage;
    cudaEvent_t e0, e1;
};
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::vector<cudaEvent_t> g_pool;

static cudaEvent_t get_event() {
    if (!g_pool.empty()) {
        cudaEvent_t e = g_pool.back();
        g_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

void prof_before(int stage, cudaStream_t s) {
    if (!g_prof_on) return;
    ProfRec r{stage, get_event(), get_event()};
    cudaEventRecord(r.e0, s);
    g_prof.push_back(r);
}
void prof_after(int, cudaStream_t s) {
    if (!g_prof_on) return;
    cudaEventRecord(g_prof.back().e1, s);
}

}  // namespace die

using namespace die;

extern "C" int32_t die_abi_version(void) { return DIE_ABI_VERSION; }
extern "C" const char* die_last_error(void) { return die::g_err; }
extern "C" uint64_t die_launch_count(void) { return die::g_launches.load(std::memory_order_relaxed); }
extern "C" die_status_t die_set_device(int32_t device) {
    DIE_CUDA(cudaSetDevice(device));
    return DIE_OK;
}

extern "C" die_status_t die_profile_begin(void) {
    g_prof.clear();
    g_prof_on = true;
    return DIE_OK;
}

extern "C" die_status_t die_profile_end(int32_t cap, double* ms_out, uint64_t* count_out, int32_t* n_stages) {
    g_prof_on = false;
    DIE_CUDA(cudaDeviceSynchronize());
    double ms[ST_COUNT] = {0};
    uint64_t cnt[ST_COUNT] = {0};
    for (const ProfRec& r : g_prof) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r.e0, r.e1) == cudaSuccess) {
            ms[r.stage] += t;
            cnt[r.stage] += 1;
        }
        g_pool.push_back(r.e0);
        g_pool.push_back(r.e1);
    }
    g_prof.clear();
    for (int i = 0; i < ST_COUNT && i < cap; ++i) {
        if (ms_out) ms_out[i] = ms[i];
        if (count_out) count_out[i] = cnt[i];
    }
    if (n_stages) *n_stages = ST_COUNT;
    return DIE_OK;
}

extern "C" const char* die_profile_stage_name(int32_t stage) {
    return stage >= 0 && stage < ST_COUNT ? kStageNames[stage] : "";
}
