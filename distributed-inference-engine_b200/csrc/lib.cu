// Library-wide plumbing: version, thread-local error text, launch accounting.
#include <atomic>

#include "common.cuh"

namespace die {

static thread_local char g_err[512] = "";
static std::atomic<unsigned long long> g_launches{0};

char* err_buf() { return g_err; }
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace die

extern "C" int32_t die_abi_version(void) { return DIE_ABI_VERSION; }
extern "C" const char* die_last_error(void) { return die::g_err; }
extern "C" uint64_t die_launch_count(void) { return die::g_launches.load(std::memory_order_relaxed); }
extern "C" die_status_t die_set_device(int32_t device) {
    DIE_CUDA(cudaSetDevice(device));
    return DIE_OK;
}
