// Library-wide plumbing: version, thread-local error text, launch accounting, per-stage timers.
#include <algorithm>
#include <atomic>
#include <vector>

#include "common.cuh"

namespace die {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

char* err_buf() { return g_err; }
static std::atomic<unsigned long long> g_replays{0};
void count_replay() { g_replays.fetch_add(1, std::memory_order_relaxed); }
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// Optional per-stage device timing: while enabled every kernel launch is bracketed by two CUDA
// events recorded on the kernel's own stream; die_profile_end() synchronises and sums them.
static const char* const kStageNames[ST_COUNT] = {
    "route_md5",   "kv_get",      "kv_contains", "kv_delete",  "kv_purge", "kv_set_precheck",
    "kv_set_window", "kv_set_claim", "kv_set_finalize", "kv_set_evict", "kv_maintenance", "miss_scan",
    "pack_hits",   "model",       "unpack_store", "pack_ragged", "unpack_ragged", "bucket_by_shard", "xchg_dispatch",
    "xchg_barrier", "xchg_compact", "xchg_return", "xchg_scatter", "fused_step"};

struct ProfRec {
    int stage;
    cudaEvent_t e0, e1;
    cudaStream_t stream;
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

static unsigned long long* g_trace = nullptr;
static unsigned long long g_trace_cap = 0;
unsigned long long* trace_buffer() { return g_trace; }

bool prof_active() { return g_prof_on; }
bool pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("DIE_PDL");
        on = e ? (atoi(e) != 0) : 1;
    }
    return on != 0;
}
void prof_before(int stage, cudaStream_t s) {
    if (!g_prof_on) return;
    ProfRec r{stage, get_event(), get_event(), s};
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
extern "C" uint64_t die_replay_count(void) { return die::g_replays.load(std::memory_order_relaxed); }
extern "C" die_status_t die_set_device(int32_t device) {
    DIE_CUDA(cudaSetDevice(device));
    return DIE_OK;
}

extern "C" die_status_t die_profile_begin(void) {
    g_prof.clear();
    g_prof_on = true;
    return DIE_OK;
}

// Diagnostics: the recorded launches as a timeline (start / end in microseconds after the first recorded
// launch began, and a small integer per distinct stream).  Call before die_profile_end, which frees the records.
extern "C" die_status_t die_profile_timeline(int32_t cap, int32_t* stage_out, int32_t* stream_out, double* start_us,
                                             double* end_us, int32_t* n_out) {
    g_prof_on = false;
    DIE_CUDA(cudaDeviceSynchronize());
    std::vector<cudaStream_t> streams;
    int n = 0;
    for (const ProfRec& r : g_prof) {
        if (n >= cap) break;
        float t0 = 0.f, t1 = 0.f;
        if (cudaEventElapsedTime(&t0, g_prof.front().e0, r.e0) != cudaSuccess) continue;
        if (cudaEventElapsedTime(&t1, g_prof.front().e0, r.e1) != cudaSuccess) continue;
        int si = 0;
        for (; si < (int)streams.size(); ++si)
            if (streams[si] == r.stream) break;
        if (si == (int)streams.size()) streams.push_back(r.stream);
        stage_out[n] = r.stage;
        stream_out[n] = si;
        start_us[n] = 1e3 * t0;
        end_us[n] = 1e3 * t1;
        ++n;
    }
    if (n_out) *n_out = n;
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

// Diagnostics: kernels replayed from a graph log (tag, %globaltimer ns) at their start into a device buffer.
extern "C" die_status_t die_debug_trace_enable(uint64_t n_records) {
    if (g_trace) cudaFree(g_trace);
    g_trace = nullptr;
    g_trace_cap = n_records;
    if (n_records == 0) return DIE_OK;
    DIE_CUDA(cudaMalloc(&g_trace, (2 + 2 * n_records) * sizeof(unsigned long long)));
    const unsigned long long head[2] = {0ull, n_records};
    DIE_CUDA(cudaMemcpy(g_trace, head, sizeof(head), cudaMemcpyHostToDevice));
    return DIE_OK;
}

extern "C" die_status_t die_debug_trace_read_sync(uint64_t cap, uint64_t* tag_out, uint64_t* ns_out, uint64_t* n_out) {
    if (!g_trace) DIE_FAIL(DIE_ERR_INVALID, "die_debug_trace_read_sync: tracing is not enabled");
    DIE_CUDA(cudaDeviceSynchronize());
    unsigned long long head[2];
    DIE_CUDA(cudaMemcpy(head, g_trace, sizeof(head), cudaMemcpyDeviceToHost));
    const uint64_t n = std::min<uint64_t>(std::min<uint64_t>(head[0], head[1]), cap);
    std::vector<unsigned long long> buf(2 * n);
    if (n) DIE_CUDA(cudaMemcpy(buf.data(), g_trace + 2, 2 * n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < n; ++i) {
        tag_out[i] = buf[2 * i];
        ns_out[i] = buf[2 * i + 1];
    }
    if (n_out) *n_out = n;
    const unsigned long long zero = 0ull;
    DIE_CUDA(cudaMemcpy(g_trace, &zero, sizeof(zero), cudaMemcpyHostToDevice));
    return DIE_OK;
}

extern "C" const char* die_profile_stage_name(int32_t stage) {
    return stage >= 0 && stage < ST_COUNT ? kStageNames[stage] : "";
}

// ---- calibration kernel: random-sector read ceiling ---------------------------------------------------------
// The table kernels are bound by random 32/64-byte accesses, not by streaming bandwidth.  This kernel issues
// `per_thread` independent random reads of `bytes_each` (16, 32 or 64) per thread over a large buffer and
// nothing else, to give bench.py the denominator "what the memory system delivers for random sectors".
namespace die {
__global__ void __launch_bounds__(256)
random_read_kernel(const uint4* __restrict__ buf, unsigned long long n_vec16, int per_thread, int vec_per_access,
                   unsigned long long seed, unsigned long long* __restrict__ sink) {
    unsigned long long x = (blockIdx.x * 256ull + threadIdx.x) * 0x9E3779B97F4A7C15ull + seed;
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (int it = 0; it < per_thread; it += 8) {
        uint4 v[8];
        unsigned long long idx[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            x ^= x >> 12; x ^= x << 25; x ^= x >> 27;   // xorshift64*
            idx[k] = ((x * 0x2545F4914F6CDD1Dull) % (n_vec16 / vec_per_access)) * vec_per_access;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            v[k] = __ldcg(buf + idx[k]);
            for (int j = 1; j < vec_per_access; ++j) {
                const uint4 w = __ldcg(buf + idx[k] + j);
                v[k].x ^= w.x; v[k].y ^= w.y; v[k].z ^= w.z; v[k].w ^= w.w;
            }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) { acc.x ^= v[k].x; acc.y ^= v[k].y; acc.z ^= v[k].z; acc.w ^= v[k].w; }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x12345678u) *sink = acc.x;  // keep the loads alive
}
}  // namespace die

extern "C" die_status_t die_debug_random_read(const void* buf, uint64_t bytes, int32_t n_threads, int32_t per_thread,
                                              int32_t bytes_each, uint64_t seed, uint64_t* sink, die_stream_t stream) {
    if (!buf || bytes < 4096 || n_threads <= 0 || per_thread <= 0 || (bytes_each != 16 && bytes_each != 32 && bytes_each != 64))
        DIE_FAIL(DIE_ERR_INVALID, "die_debug_random_read: bad arguments");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_LAUNCH(ST_BUCKET, s, random_read_kernel<<<(n_threads + 255) / 256, 256, 0, s>>>(
        static_cast<const uint4*>(buf), bytes / 16, (per_thread + 7) / 8 * 8, bytes_each / 16, seed,
        reinterpret_cast<unsigned long long*>(sink)));
    return DIE_OK;
}
