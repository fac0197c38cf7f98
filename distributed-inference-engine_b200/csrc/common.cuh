// Shared host/device helpers for libdie.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/die.h"

namespace die {

// ---- error plumbing ---------------------------------------------------------------------
enum Stage {
    ST_ROUTE = 0, ST_KV_GET, ST_KV_CONTAINS, ST_KV_DELETE, ST_KV_PURGE, ST_KV_PRECHECK, ST_KV_WINDOW, ST_KV_CLAIM,
    ST_KV_FINALIZE, ST_KV_EVICT, ST_KV_MAINT, ST_MISS_SCAN, ST_PACK_HITS, ST_MODEL, ST_UNPACK_STORE, ST_PACK,
    ST_UNPACK, ST_BUCKET, ST_XCHG, ST_XCHG_BARRIER, ST_XCHG_COMPACT, ST_XCHG_RETURN, ST_XCHG_SCATTER, ST_FUSED_STEP, ST_COUNT
};
void prof_before(int stage, cudaStream_t s);
void prof_after(int stage, cudaStream_t s);
bool prof_active();                   // per-launch events are being recorded: replays from graphs are off
unsigned long long* trace_buffer();   // device log of replayed kernels' start times, or nullptr (diagnostics)
char* err_buf();                       // thread-local, 512 bytes
void count_replay();                   // graph replays launched (tests check that the replay path ran)
void count_launch(unsigned n = 1);     // gpu_launches accounting for bench.py

#define DIE_FAIL(code, ...)                                   \
    do {                                                      \
        snprintf(::die::err_buf(), 512, __VA_ARGS__);         \
        return (code);                                        \
    } while (0)

#define DIE_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess)                                                          \
            DIE_FAIL(DIE_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,          \
                     cudaGetErrorString(e__));                                           \
    } while (0)

#define DIE_LAUNCH_CHECK()                                                               \
    do {                                                                                 \
        ::die::count_launch();                                                           \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess)                                                          \
            DIE_FAIL(DIE_ERR_CUDA, "%s:%d kernel launch -> %s", __FILE__, __LINE__,      \
                     cudaGetErrorString(e__));                                           \
    } while (0)

// Launch a kernel, optionally bracketed by CUDA events on its own stream (die_profile_begin/end).
#define DIE_LAUNCH(stage, stream, ...)             \
    do {                                           \
        ::die::prof_before((stage), (stream));     \
        __VA_ARGS__;                               \
        ::die::prof_after((stage), (stream));      \
        DIE_LAUNCH_CHECK();                        \
    } while (0)

// ---- programmatic dependent launch ---------------------------------------------------------------------
// The step is a chain of short kernels; with this attribute a kernel's CTAs are scheduled while the previous
// kernel of the stream drains, run their prologue (argument / descriptor loads, independent operand fetches)
// and block in pdl_wait() until that kernel has completed and its writes are visible.  Every CTA of such a
// kernel calls pdl_wait() before it touches anything a predecessor wrote.  DIE_PDL=0 launches them normally.
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

#define DIE_TRY(expr)                        \
    do {                                     \
        die_status_t s__ = (expr);           \
        if (s__ != DIE_OK) return s__;       \
    } while (0)

inline int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- device helpers ---------------------------------------------------------------------
struct __align__(16) u128 {
    unsigned long long lo, hi;
};

#ifdef __CUDACC__
__device__ __forceinline__ u128 ld128(const void* p) {
    u128 r;
    asm volatile("ld.global.v2.u64 {%0, %1}, [%2];" : "=l"(r.lo), "=l"(r.hi) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st128(void* p, u128 v) {
    asm volatile("st.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v.lo), "l"(v.hi) : "memory");
}
// 128-bit compare-and-swap (ATOMG.E.CAS.128 on sm_100a).
__device__ __forceinline__ u128 cas128(void* addr, u128 cmp, u128 val) {
    u128 old;
    asm volatile(
        "{\n\t.reg .b128 c, v, o;\n\t"
        "mov.b128 c, {%3, %4};\n\t"
        "mov.b128 v, {%5, %6};\n\t"
        "atom.global.cas.b128 o, [%2], c, v;\n\t"
        "mov.b128 {%0, %1}, o;\n\t}"
        : "=l"(old.lo), "=l"(old.hi)
        : "l"(addr), "l"(cmp.lo), "l"(cmp.hi), "l"(val.lo), "l"(val.hi)
        : "memory");
    return old;
}
// streaming 128-bit accesses that do not allocate in L1 (read-once / write-once payload)
__device__ __forceinline__ int4 ldg_stream(const int4* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(int4* p, const int4& v) {
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

// block-wide exclusive scan of one unsigned per thread (1024 threads); returns the block total
__device__ __forceinline__ unsigned block_scan_1024(unsigned cnt, unsigned& excl, unsigned* s_warp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        unsigned w = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += v;
        }
        s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    excl = (warp ? s_warp[warp - 1] : 0u) + (incl - cnt);
    return s_warp[31];
}
#endif

}  // namespace die
