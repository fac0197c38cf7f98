// The composed hot path as ONE persistent kernel per group of flushed batches (SURVEY.md call-stack (E)):
// route -> probe -> pack -> infer -> unpack -> fill, batch after batch, with grid-wide barriers where the
// reference's "applied in request order" contract needs one and nowhere else.
//
// A 4096-request batch moves ~5 MB (< 1 us of HBM time); what a step costs is its chain of dependent DRAM
// round trips plus, in the three-kernel form (pipeline.cu), two kernel boundaries and a serial "last CTA" tail.
// Here the chain is all that is left:
//
//   grid = n_req request CTAs + n_chunks window CTAs, all co-resident (cooperative launch), 256 threads each
//   request CTA c owns requests [c*R, c*R+R) of every batch (R = 32..96):
//     its payload rows ARE its packed tile: cp.async straight into the SWIZZLE_128B A tile of a tcgen05 GEMM
//     (batcher.py:211's gather never touches HBM), prefetched one batch ahead; W stays in shared memory
//     for the whole launch; results leave TMEM for out[q] and the cached row (batcher.py:225-227) directly
//   window CTA k owns chunk k (1024 records) of the LRU log behind the head (kvstore.py:82-93)
//
//   per batch:  A  probe (kvstore.py:136-164), one request per thread            | window CTAs idle
//               -- barrier 1: every get of the batch precedes every set (call-stack (E))
//               B  insert the misses (kvstore.py:104-134) | hit rows -> out | MMA | bitmap of current log records
//               -- barrier 2: every insert is in, so E = size - max_size (_evict_if_needed) is known to all
//               C  GEMM epilogue: out[q], cached row                             | evict the E oldest records
//               -- barrier 3: the next batch's probes see this batch's evictions (request CTAs arrive before
//                  their epilogue and wait after it, so the epilogue overlaps the evictions)
//
// MD5 of all the launch's keys (model_registry.py:159) is a prologue over every CTA.  Results are identical to the
// three-kernel sequence (same stamps, same claim / evict rules); tests/test_gpu_pipeline.py compares both with the
// oracle and with each other.
#include "model_internal.cuh"
#include "pipeline_device.cuh"
#include "pipeline_internal.cuh"
#include "route_device.cuh"
#include "tc_common.cuh"

namespace die {

constexpr int PS_THREADS = 256;
constexpr int PS_M = 128;           // MMA tile rows (rows >= the CTA's request count are never read back)
constexpr int PS_MAX_R = 96;        // requests per request CTA (staging R x (N+4) floats must fit the A tile)
constexpr int PS_MAX_CHUNKS = 64;   // window CTAs

// Device-resident control block of the persistent kernel (one per pipeline, zero-initialised).
struct StepGrid {
    unsigned int bar_count;   // arrivals at all barriers so far (monotone: barrier k of a launch completes at
                              // bar_base + (k+1) * n_cta; nothing is ever reset while CTAs may still be polling)
    unsigned int bar_base;    // bar_count when the current launch began (written by the previous launch's last act)
    unsigned int abort;       // a barrier timed out: every later barrier falls through, status is raised
    unsigned int pad;
    unsigned long long dup_tag;  // == t_set of the batch in which a key occurred twice among the misses
    unsigned long long hnew;     // LRU-log position just past the newest victim of the current batch
    unsigned int wcount[PS_MAX_CHUNKS];  // current records per window chunk of the current batch
    // %globaltimer marks of the last batch, CTA 0 (diagnostics): [0] start, [4] probes issued, [1] barrier 1 passed,
    // [5] inserts done, [2] barrier 2 passed, [6] epilogue done, [3] barrier 3 passed
    // ... and the same marks of every CTA ([7]: request CTAs, accumulators ready / window CTAs, bitmap done)
    unsigned long long ts[8];
    unsigned long long cta_ts[160][8];
};

struct StepParams {
    StepRef sr;                 // first batch of the launch (stream launch: values; replay: descriptor + index)
    int n_steps;                // batches this launch processes
    int B;                      // requests per batch (upper bound when n_req_dev is set)
    int R;                      // requests per request CTA (multiple of 32)
    int n_req;                  // request CTAs
    int n_chunks;               // window CTAs
    int K, N;                   // linear layer: Y[., N] = X[., K] @ W[N, K]^T + bias
    int from_keys;              // 1: hash the keys in the prologue into `digests`; 0: `digests` are inputs
    const uint32_t* n_req_dev;  // optional device-side batch size (owner pass of the exchange)
    uint4* digests;             // [n_steps][B]
    const float* W;
    const float* bias;
    float* arena;               // [n_rows, N] cached result rows
    StepGrid* g;
    KvMirrorArgs mirror;        // m.m != nullptr: snapshot the table counters for the host before the first batch
    // Owner pass of the exchange, several steps in one launch: digests and n_req_dev are rings of `lanes` per-step
    // buffers, step s uses lane (lane0 + s) % lanes (0 lanes: one buffer per step of the launch, in order).
    int lanes, lane0;
    int dbg;                    // diagnostics (DIE_STEP_DEBUG): 1 = barriers only, no work; 2 = no L2 prefetch of the next
                                // batch's slots (timing experiments)
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned int* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ps_timer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Split grid barrier on one monotone counter.  arrive(): the whole CTA has finished what the barrier protects --
// one release-add by the CTA's last thread (a thread no phase puts on its critical path).  wait(): that thread polls
// until the count reaches `target`, then releases the CTA.  No thread "completes" the barrier and nothing is reset,
// so an arrival is visible to the pollers after a single L2 round trip.
constexpr int PS_BAR_THREAD = PS_THREADS - 1;
__device__ int g_ps_bar_mode = 1;  // 1: release-add + acquire polls (one fence per side; measured 1.2 us vs 1.7 us per barrier); DIE_STEP_BAR_MODE
__device__ __forceinline__ void grid_arrive(StepGrid* g) {
    tc::fence_before_thread_sync();
    __syncthreads();
    if (threadIdx.x == PS_BAR_THREAD) {
        if (g_ps_bar_mode == 0) {
            __threadfence();  // cumulative: everything the CTA wrote before the bar.sync above
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&g->bar_count) : "memory");
        } else if (g_ps_bar_mode == 1) {
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&g->bar_count) : "memory");
        } else {
            __threadfence();
            asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(&g->bar_count) : "memory");
        }
    }
}
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned int* p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ int g_ps_dbg = 0;  // DIE_STEP_DEBUG as seen by device helpers (timing experiments)
__device__ int g_ps_poll_ns = 0;  // pause between polls of the barrier counter (tuning knob, DIE_STEP_POLL_NS)
__device__ __forceinline__ void grid_wait(StepGrid* g, unsigned target, KvState* st) {
    if (threadIdx.x == PS_BAR_THREAD) {
        const unsigned long long t0 = ps_timer();
        unsigned spins = 0;
        const int pause = g_ps_poll_ns;
        const int mode = g_ps_bar_mode;
        while ((int)((mode == 1 ? ld_acquire_u32(&g->bar_count) : ld_relaxed_u32(&g->bar_count)) - target) < 0) {
            if (pause) __nanosleep(pause);
            if ((++spins & 1023u) == 0u) {
                if (ld_acquire_u32(&g->abort)) break;
                if (ps_timer() - t0 > 2000000000ull) {  // 2 s: a CTA of the grid is gone; fail loudly, never hang
                    atomicExch(&g->abort, 1u);
                    st->status = DIE_ERR_CUDA;
                    break;
                }
            }
        }
        if (mode != 1) asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    __syncthreads();
    tc::fence_after_thread_sync();
}

__device__ __forceinline__ void ps_cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(tc::smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}

// Rows [q0, q0+nq) of the batch's payload -> rows [0, nq) of a K-major SWIZZLE_128B A tile (blocks of
// [128 rows x 32 tf32]).  This is PACK: the CTA's requests, in request order, become the rows of its GEMM tile.
__device__ __forceinline__ void ps_load_tile(uint8_t* a_base, const float* __restrict__ payload, int q0, int nq, int K) {
    const int chunks_per_row = K / 4;
    const int total = nq * chunks_per_row;
    for (int e = threadIdx.x; e < total; e += PS_THREADS) {
        const int r = e / chunks_per_row, ck = e - r * chunks_per_row;
        ps_cp_async16(a_base + (size_t)(ck >> 3) * (PS_M * 128) + tc::sw128_offset((uint32_t)r, (uint32_t)(ck & 7)),
                      payload + (size_t)(q0 + r) * K + ck * 4);
    }
}

constexpr int PS_PER = WCHUNK / PS_THREADS;     // log records per thread of a window CTA
constexpr uint32_t PS_LOG_NONE = 0xffffffffu;   // no record (beyond t_set, or never written)
constexpr uint32_t PS_LOG_LATE = 0xfffffffeu;   // record of this batch's own gets: read after barrier 1

__device__ __forceinline__ void prefetch_l2(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// Window CTA, phase A (idle otherwise): the chunk's log entries into registers and the slots they name towards L2,
// so that the stamp reads after barrier 1 are L2 hits instead of random DRAM round trips.  Positions at or beyond
// t_get are being written by this batch's probes right now: they are read after the barrier.
__device__ __forceinline__ void ps_window_preload(const KvView& kv, unsigned long long base, unsigned long long t_get,
                                                  uint32_t (&slot)[PS_PER]) {
#pragma unroll
    for (int k = 0; k < PS_PER; ++k) {
        const unsigned long long p = base + (unsigned long long)(k * PS_THREADS + (int)threadIdx.x);
        slot[k] = p < t_get ? __ldcg(&kv.log[p & kv.rmask]) : PS_LOG_LATE;
    }
#pragma unroll
    for (int k = 0; k < PS_PER; ++k)
        if ((unsigned long long)slot[k] <= kv.mask) prefetch_l2(&kv.slots[slot[k]].stamp);
}

// Window CTA, phase B: bitmap of the chunk's log records that are still current (= entries of the reference's
// _access_order, kvstore.py:71-78) among positions [head + chunk*1024, ...) below t_set.  Word w of s_mask covers
// records [32w, 32w+32) of the chunk, s_pref[w] = current records in the words before w.  Returns the chunk's count.
__device__ __forceinline__ unsigned ps_window_bitmap(const KvView& kv, unsigned long long base, unsigned long long t_set,
                                                     uint32_t (&slot)[PS_PER], unsigned* s_mask, unsigned* s_pref,
                                                     unsigned* s_cnt) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long stamp[PS_PER];
    if (tid == 0) *s_cnt = 0u;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PS_PER; ++k) {
        const unsigned long long p = base + (unsigned long long)(k * PS_THREADS + tid);
        if (slot[k] == PS_LOG_LATE) slot[k] = p < t_set ? __ldcg(&kv.log[p & kv.rmask]) : PS_LOG_NONE;
    }
#pragma unroll
    for (int k = 0; k < PS_PER; ++k)
        stamp[k] = (unsigned long long)slot[k] <= kv.mask ? ld_u64_cg(&kv.slots[slot[k]].stamp) : 0ull;
    unsigned cnt = 0;
#pragma unroll
    for (int k = 0; k < PS_PER; ++k) {
        const unsigned long long p = base + (unsigned long long)(k * PS_THREADS + tid);
        const bool cur = p < t_set && stamp[k] == p;
        // a current record may be evicted in phase C and its successor slot looked at one phase later: bring it to L2
        if (cur) prefetch_l2(&kv.slots[((unsigned long long)slot[k] + 1ull) & kv.mask]);
        const unsigned m = __ballot_sync(0xffffffffu, cur);
        if (lane == 0) {
            s_mask[k * (PS_THREADS / 32) + warp] = m;
            cnt += __popc(m);
        }
    }
    if (lane == 0 && cnt) atomicAdd(s_cnt, cnt);
    __syncthreads();
    if (warp == 0) {
        const unsigned v = __popc(s_mask[lane]);
        unsigned incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        s_pref[lane] = incl - v;
    }
    __syncthreads();
    return *s_cnt;
}

// Window CTA, phase C: evict the chunk's current records whose rank among all current records of the window is
// below e (oldest first, kvstore.py:89-91).
//   - victims become tombstones here; ps_window_relax turns those whose successor slot is EMPTY into EMPTY during the
//     next batch's phase A (off this phase's critical path);
//   - the victim's result row goes back to the free queue at position free_tail + rank: ranks are unique, so the
//     evictions need no atomic, and the counters (size, evictions, head) are settled by one thread after
//     barrier 3 (ps_settle_evictions) so that nobody reads them while they change;
//   - a thread's four records are handled together: their loads are in flight side by side (two round trips in all).
__device__ __forceinline__ void ps_window_evict(const KvView& kv, StepGrid* g, int chunk, int n_chunks,
                                                unsigned long long base, unsigned long long tail0, unsigned long long t_set,
                                                unsigned long long e, const uint32_t (&slot)[PS_PER], const unsigned* s_mask,
                                                const unsigned* s_pref, uint32_t (&evicted)[PS_PER]) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int k = 0; k < PS_PER; ++k) evicted[k] = PS_LOG_NONE;
    if (e == 0) return;  // uniform for the grid
    // tail0 = free_tail as every window CTA read it in phase B (nothing pushes between that read and barrier 3 except
    // these evictions); the queue's tail moves past them here, before any push of the next batch can happen
    if (chunk == 0 && tid == 0) kv.st->free_tail = tail0 + e;
    // current records in the chunks before this one / in the whole window: two loads per lane, side by side
    const int lane_ = tid & 31;
    const unsigned w0 = lane_ < n_chunks ? __ldcg(&g->wcount[lane_]) : 0u;
    const unsigned w1 = lane_ + 32 < n_chunks ? __ldcg(&g->wcount[lane_ + 32]) : 0u;
    unsigned bsum = (lane_ < chunk ? w0 : 0u) + (lane_ + 32 < chunk ? w1 : 0u), tsum = w0 + w1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        bsum += __shfl_xor_sync(0xffffffffu, bsum, o);
        tsum += __shfl_xor_sync(0xffffffffu, tsum, o);
    }
    const unsigned long long before = bsum, total = tsum;
    constexpr int PER = PS_PER;
    bool victim[PER];
    unsigned long long rank[PER];
    uint32_t sidx[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const int rec = k * PS_THREADS + tid;  // record index inside the chunk
        const int w = rec >> 5, b = rec & 31;
        const unsigned word = s_mask[w];
        victim[k] = before < e && ((word >> b) & 1u);
        rank[k] = 0;
        if (victim[k]) {
            rank[k] = before + s_pref[w] + __popc(word & ((1u << b) - 1u));
            victim[k] = rank[k] < e;
        }
        sidx[k] = victim[k] ? slot[k] : 0u;  // the log entry read for the bitmap
    }
    unsigned long long row[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k)
        if (victim[k]) row[k] = ld_u64_cg(&kv.slots[sidx[k]].value);  // same sector as the stamp read for the bitmap
    unsigned n_tomb = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        if (!victim[k]) continue;
        Slot* sl = &kv.slots[sidx[k]];
        st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});  // ps_window_relax turns it into EMPTY later where that is allowed
        sl->stamp = 0ull;
        n_tomb += 1u;
        evicted[k] = sidx[k];
        if (kv.free_rows) kv.free_rows[(tail0 + rank[k]) & kv.row_mask] = (uint32_t)row[k];
        if (rank[k] == e - 1) g->hnew = base + (unsigned long long)(k * PS_THREADS + tid) + 1;  // the newest victim
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n_tomb += __shfl_xor_sync(0xffffffffu, n_tomb, o);
    if ((tid & 31) == 0 && n_tomb) atomicAdd(&kv.st->n_tomb, (unsigned long long)n_tomb);
    // window too short for this log (fewer than e current records in it): finish the walk sequentially, exactly
    if (chunk == 0 && tid == 0 && total < e) {
        unsigned long long p = base + (unsigned long long)n_chunks * WCHUNK, r = total;
        while (r < e && p < t_set) {
            const uint32_t si = kv.log[p & kv.rmask];
            if ((unsigned long long)si <= kv.mask && ld_u64_cg(&kv.slots[si].stamp) == p) {
                Slot* sl = &kv.slots[si];
                const unsigned long long v = ld_u64_cg(&sl->value);
                const bool tomb = !fp_empty(ld_fp(&kv.slots[((unsigned long long)si + 1ull) & kv.mask]));
                st128(sl, tomb ? u128{FP_TOMB_WORD, FP_TOMB_WORD} : u128{0ull, 0ull});
                sl->stamp = 0ull;
                if (tomb) atomicAdd(&kv.st->n_tomb, 1ull);
                if (kv.free_rows) kv.free_rows[(tail0 + r) & kv.row_mask] = (uint32_t)v;
                ++r;
            }
            ++p;
        }
        if (r < e) kv.st->status = DIE_ERR_INVALID;  // unreachable while max_batch <= max_size
        g->hnew = p;
    }
}

// Window CTA, next batch's phase A (idle otherwise): an evicted slot whose successor is EMPTY holds no probe chain up
// (a chain through it would stop at the successor anyway), so it goes back to EMPTY instead of staying a tombstone --
// most evicted slots do at a live load of 1/8, which is what keeps table rebuilds thousands of batches apart.  Done
// here rather than in the eviction itself because the successor is a random DRAM line and phase C is on the step's
// critical path.  Probes of the next batch may run meanwhile: tombstone or EMPTY, they stop at the same slot with the
// same answer, and nobody inserts in phase A.
__device__ __forceinline__ void ps_window_relax(const KvView& kv, uint32_t (&evicted)[PS_PER]) {
    u128 next[PS_PER];
#pragma unroll
    for (int k = 0; k < PS_PER; ++k)
        if (evicted[k] != PS_LOG_NONE) next[k] = ld_fp(&kv.slots[((unsigned long long)evicted[k] + 1ull) & kv.mask]);
    unsigned relaxed = 0;
#pragma unroll
    for (int k = 0; k < PS_PER; ++k) {
        if (evicted[k] == PS_LOG_NONE) continue;
        if (fp_empty(next[k])) {
            st128(&kv.slots[evicted[k]], u128{0ull, 0ull});
            ++relaxed;
        }
        evicted[k] = PS_LOG_NONE;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) relaxed += __shfl_xor_sync(0xffffffffu, relaxed, o);
    if ((threadIdx.x & 31) == 0 && relaxed) atomicAdd(&kv.st->n_tomb, (unsigned long long)(-(long long)relaxed));
}

// After barrier 3, one thread: the batch's evictions leave the counters (kvstore.py:87-93).
__device__ __forceinline__ void ps_settle_evictions(const KvView& kv, const StepGrid* g, unsigned long long e) {
    if (e == 0) return;
    KvState* st = kv.st;
    atomicAdd(&st->size, (unsigned long long)(-(long long)e));  // atomic: the next batch's probes may delete expired entries
    st->evictions += e;
    st->head = ld_u64_cg(&g->hnew);
}

// Request CTA's private stock of result rows: the inserts of a batch take their rows from shared memory, and one idle
// warp tops the stock up to R handles with a single pop on the table's free queue per batch, in phase C, where nobody
// waits for it.  (In phase C the queue only grows -- evicted rows are appended beyond its tail -- and it holds at least
// max_batch handles that were free before the phase, see release_row.)
__device__ __forceinline__ void ps_stash_refill(const KvView& kv, int* s_stash, int* s_nstash, int R) {
    const int lane = (int)lane_id();
    const int have = *s_nstash;
    __syncwarp();
    const int need = R - have;
    if (need > 0) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&kv.st->free_head, (unsigned long long)need);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (int i = lane; i < need; i += 32) s_stash[have + i] = (int)kv.free_rows[(base + (unsigned long long)i) & kv.row_mask];
        __syncwarp();
        if (lane == 0) *s_nstash = R;
    }
}
__device__ __forceinline__ void ps_stash_return(const KvView& kv, const int* s_stash, int* s_nstash) {
    const int lane = (int)lane_id();
    const int have = *s_nstash;
    __syncwarp();
    if (have > 0) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&kv.st->free_tail, (unsigned long long)have);
        base = __shfl_sync(0xffffffffu, base, 0);
        for (int i = lane; i < have; i += 32) kv.free_rows[(base + (unsigned long long)i) & kv.row_mask] = (uint32_t)s_stash[i];
        __syncwarp();
        if (lane == 0) *s_nstash = 0;
    }
}

__device__ __forceinline__ void warp_count_add(unsigned long long* ctr, bool pred, bool negative = false) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m && lane_id() == 0)
        atomicAdd(ctr, negative ? (unsigned long long)(-(long long)__popc(m)) : (unsigned long long)__popc(m));
}

// The requests request CTA `cta` owns in a step: [q0, q0 + nq).  The batch -- its size may only exist on the device
// (owner pass of the exchange, sized for owner_max but holding about half of it) -- is spread evenly over all request
// CTAs instead of filling the first ones with R rows each.
__device__ __forceinline__ void ps_span(const StepParams& P, int lane_idx, int cta, int& Bs, int& q0, int& nq) {
    Bs = P.B;
    if (P.n_req_dev) Bs = min(Bs, (int)__ldcg(P.n_req_dev + (P.lanes ? lane_idx : 0)));
    const int Rs = min(P.R, max(4, ((Bs + P.n_req - 1) / P.n_req + 3) & ~3));
    q0 = cta * Rs;
    nq = max(0, min(Rs, Bs - q0));
}

__global__ void __launch_bounds__(PS_THREADS, 1)
pipe_step_kernel(const __grid_constant__ StepParams P, const __grid_constant__ QueueDesc qv, int use_qv) {
    extern __shared__ uint8_t ps_smem_raw[];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_tmem;
    __shared__ unsigned s_mask[WCHUNK / 32];
    __shared__ unsigned s_cnt;
    __shared__ unsigned long long s_value[PS_MAX_R];
    __shared__ int s_slot[PS_MAX_R];
    __shared__ int s_row[PS_MAX_R];
    __shared__ uint8_t s_code[PS_MAX_R];
    __shared__ unsigned s_pref[WCHUNK / 32];
    __shared__ int s_stash[2 * PS_MAX_R];  // result-row handles this CTA holds for its next misses
    __shared__ int s_nstash;
    __shared__ __align__(16) float s_bias[256];

    StepRef sr = P.sr;
    if (use_qv) sr.q = &qv;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int cta = (int)blockIdx.x;
    const bool is_req = cta < P.n_req;
    const int chunk = cta - P.n_req;
    const unsigned n_cta = gridDim.x;
    StepGrid* g = P.g;
    unsigned target = ld_acquire_u32(&g->bar_base);  // barrier k of this launch completes at bar_base + (k+1) * n_cta
    const int K = P.K, N = P.N, R = P.R;
    const int KB = K / 32;
    const int vout = N / 4;

    uint8_t* smem = ps_smem_raw + ((1024u - (tc::smem_u32(ps_smem_raw) & 1023u)) & 1023u);
    uint8_t* b_base = smem;                                    // W: KB blocks of [N rows x 128 B]
    uint8_t* a_base0 = smem + (size_t)N * K * 4;               // two A tiles: KB blocks of [128 rows x 128 B] each
    const size_t a_bytes = (size_t)PS_M * K * 4;

    const StepArgs a_first = step_resolve(sr);
    const KvView kv = a_first.kv;
    KvState* st = kv.st;
    if (cta == 0 && tid == 0 && P.mirror.m)
        kv_mirror_write(P.mirror.st, P.mirror.m, P.mirror.ops_seen, P.mirror.gen, P.mirror.seq);
    trace_mark(sr, TR_GET);

    // ---- prologue: TMEM + weights (request CTAs), MD5 of every key of the launch (all CTAs) -------------------
    const uint32_t ncols = N <= 32 ? 32u : (N <= 64 ? 64u : (N <= 128 ? 128u : 256u));
    if (tid < N) s_bias[tid] = P.bias ? P.bias[tid] : 0.f;  // N <= 256 = the CTA's thread count
    if (tid == 0) s_nstash = 0;
    __syncthreads();
    if (is_req) {
        if (warp == 3 && !(P.dbg & 1)) ps_stash_refill(kv, s_stash, &s_nstash, R);  // nothing pushes before the first batch
        if (warp == 0) tc::tmem_alloc(&s_tmem, ncols);
        if (tid == 32) tc::mbar_init(&s_bar, 1);
    }
    if (P.from_keys) {
        const int cps = (P.B + ROUTE_THREADS - 1) / ROUTE_THREADS;
        const int total = P.n_steps * cps;
        uint8_t* stage = a_base0;  // the A tiles are not in use yet
        for (int task = cta; task < total; task += (int)n_cta) {
            const int s = task / cps, c = task - s * cps;
            StepRef r = sr;
            r.i += s;
            const StepArgs a = step_resolve(r);
            const Md5State h = cta_hash_keys_at(a.key_bytes, a.key_off, P.B, c * ROUTE_THREADS, stage);
            const int i = c * ROUTE_THREADS + tid;
            if (i < P.B) P.digests[(size_t)s * P.B + i] = make_uint4(h.a, h.b, h.c, h.d);
            __syncthreads();  // the staging buffer is reused by the next chunk
        }
        grid_arrive(g);
        target += n_cta;
        grid_wait(g, target, st);
    }
    if (is_req) {
        const int chunks_per_row = K / 4;
        for (int e = tid; e < N * chunks_per_row; e += PS_THREADS) {
            const int r = e / chunks_per_row, ck = e - r * chunks_per_row;
            ps_cp_async16(b_base + (size_t)(ck >> 3) * ((size_t)N * 128) + tc::sw128_offset((uint32_t)r, (uint32_t)(ck & 7)),
                          P.W + (size_t)r * K + ck * 4);
        }
        int B0, q00, nq0;
        ps_span(P, P.lanes ? P.lane0 % P.lanes : 0, cta, B0, q00, nq0);
        ps_load_tile(a_base0, a_first.payload, q00, nq0, K);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
#define PS_MARK(i)                                             \
    do {                                                       \
        if (tid == 0) {                                        \
            const unsigned long long t__ = ps_timer();         \
            if (cta < 160) g->cta_ts[cta][i] = t__;            \
            if (cta == 0) g->ts[i] = t__;                      \
        }                                                      \
    } while (0)
    unsigned mma_phase = 0;
    u128 key_next{0ull, 0ull};   // this thread's key of the next batch, fetched one batch ahead
    bool have_next = false;
    uint32_t wslot[PS_PER];      // window CTAs: the chunk's log entries of the current batch
    uint32_t wevicted[PS_PER];   // ... and the slots it evicted (relaxed to EMPTY where possible in the next phase A)
#pragma unroll
    for (int k = 0; k < PS_PER; ++k) wevicted[k] = PS_LOG_NONE;
    unsigned long long cur_head = 0;  // window CTAs: LRU head, tracked locally (st->head is settled by one thread)
    if (!is_req) cur_head = ld_u64_cg(&st->head);

    for (int s = 0; s < P.n_steps; ++s) {
        StepRef r = sr;
        r.i += s;
        const StepArgs a = step_resolve(r);
        const int lane_s = P.lanes ? (P.lane0 + s) % P.lanes : s;       // which per-step digest / count buffer
        const int lane_n = P.lanes ? (P.lane0 + s + 1) % P.lanes : s + 1;
        const bool next_ready = s + 1 < P.n_steps && !(P.dbg & 1);
        int Bs, q0, nq, Bn = 0, q0n = 0, nqn = 0;
        ps_span(P, lane_s, cta, Bs, q0, nq);
        if (!is_req || (P.dbg & 1)) nq = 0;
        if (next_ready) ps_span(P, lane_n, cta, Bn, q0n, nqn);
        uint8_t* a_cur = a_base0 + (size_t)(s & 1) * a_bytes;
        const uint4* dig = P.digests + (size_t)lane_s * P.B;
        const uint4* dig_next = P.digests + (size_t)lane_n * P.B;
        u128 key{0ull, 0ull};
        int code = 3;  // 3 = no request in this row
        int my_row = -1;  // result row popped for this thread's miss
        PS_MARK(0);

        // ================= phase A: probe =================
        if (is_req) {
            if (next_ready) {  // PACK of the next batch: its rows travel while this batch is processed
                StepRef r2 = sr;
                r2.i += s + 1;
                ps_load_tile(a_base0 + (size_t)((s + 1) & 1) * a_bytes, step_resolve(r2).payload, q0n, nqn, K);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            if (tid < R) {
                const bool active = tid < nq;
                SlotLine home{};
                if (active) {  // the key was fetched during the previous batch (its home slot prefetched to L2)
                    key = have_next ? key_next : fp_of(__ldcg(dig + q0 + tid));
                    home = ld_slot_line(&kv.slots[key.lo & kv.mask]);
                }
                have_next = next_ready && tid < nqn;
                if (have_next) key_next = fp_of(__ldcg(dig_next + q0n + tid));
                if (active) {
                    const int q = q0 + tid;
                    unsigned long long value;
                    long long slot;
                    code = kv_get_one_from<false>(kv, key, home, a.now, a.t_get + (unsigned long long)q, value, slot);
                    a.code[q] = code == 1 ? (uint8_t)1 : (uint8_t)0;  // callers see 1 = served from the cache
                    s_value[tid] = value;
                }
                s_code[tid] = (uint8_t)code;
                {   // a miss will insert in phase B: it takes its result row from the CTA's stock now
                    const bool want = active && code != 1;
                    const unsigned m = __ballot_sync(0xffffffffu, want);
                    if (m) {
                        const int leader = __ffs(m) - 1;
                        int top = 0;
                        if (lane == leader) top = atomicSub(&s_nstash, __popc(m));
                        top = __shfl_sync(0xffffffffu, top, leader);
                        if (want) my_row = s_stash[top - 1 - __popc(m & ((1u << lane) - 1u))];
                    }
                }
                if (have_next && !(P.dbg & 2)) {  // next batch's home slot and where its probe chain continues -> L2
                    const Slot* nh = &kv.slots[key_next.lo & kv.mask];
                    prefetch_l2(nh);
                    prefetch_l2(&nh->stamp);
                    prefetch_l2(&kv.slots[(key_next.lo + 2ull) & kv.mask]);
                }
                warp_count_add(&st->hit_part[cta & 15][0], active && code == 1);
                warp_count_add(&st->miss_part[cta & 15][0], active && code != 1);
                const bool deleted = active && code == 2;  // expired: this op deleted the entry (kvstore.py:153-156)
                if (__any_sync(0xffffffffu, deleted)) {
                    warp_count_add(&st->size, deleted, true);
                    warp_count_add(&st->n_tomb, deleted);
                }
            }
            PS_MARK(4);
            asm volatile("cp.async.wait_group 1;" ::: "memory");  // this batch's tile (and W) have landed
            tc::fence_proxy_async_smem();
        } else if (!(P.dbg & 1)) {
            ps_window_preload(kv, cur_head + (unsigned long long)chunk * WCHUNK, a.t_get, wslot);
            ps_window_relax(kv, wevicted);
        }
        grid_arrive(g);
        target += n_cta;
        grid_wait(g, target, st);
        PS_MARK(1);

        // ================= phase B: insert the misses | hit rows | MMA | victim bitmap =================
        unsigned long long wbase = 0, wtail = 0;
        if (is_req) {
            if (tid == 4 * 32 && nq > 0) {  // INFER: one thread issues the tile's MMAs, accumulators in TMEM
                const uint32_t idesc = tc::instr_desc(tc::FMT_TF32, PS_M, (uint32_t)N);
                const uint32_t a_s = tc::smem_u32(a_cur), b_s = tc::smem_u32(b_base);
                for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint64_t ad = tc::smem_desc_k_major_sw128(a_s + kb * (PS_M * 128) + j * 32);
                        const uint64_t bd = tc::smem_desc_k_major_sw128(b_s + kb * (N * 128) + j * 32);
                        tc::mma_tf32(s_tmem, ad, bd, idesc, (kb | j) != 0);
                    }
                }
                tc::mma_commit(&s_bar);
            }
            if (tid < R) {
                const bool miss = tid < nq && code != 1;
                bool claimed = false, from_empty = false;
                if (miss) {  // KVCache.set of a key the get above found absent (kvstore.py:104-134)
                    const int q = q0 + tid;
                    const long long sidx = kv_claim_absent(kv, key, claimed, from_empty);
                    Slot* sl = &kv.slots[sidx];
                    const unsigned long long stamp = a.t_set + (unsigned long long)q;
                    atomicMax(&sl->stamp, stamp);
                    int row = -1;
                    if (claimed) {
                        row = my_row;
                        sl->value = (unsigned long long)row;
                        sl->created_at = a.now;
                        sl->ttl = kv.default_ttl;
                        sl->last_accessed = a.now;
                        sl->access_count = 0u;
                    } else {
                        g->dup_tag = a.t_set;  // same key twice among the misses: the later op owns the cached row
                    }
                    kv.log[stamp & kv.rmask] = (uint32_t)sidx;
                    s_slot[tid] = (int)sidx;
                    s_row[tid] = row;
                }
                if (miss && !claimed) s_stash[atomicAdd(&s_nstash, 1)] = my_row;  // a duplicate's spare row goes back
                warp_count_add(&st->size, miss && claimed);
                warp_count_add(&st->n_empty_claimed, miss && claimed && from_empty);
                warp_count_add(&st->n_tomb, miss && claimed && !from_empty, true);
            } else if (warp >= 4) {  // hits: cached result row -> out[q]
                __syncwarp();
                const int4* arena = reinterpret_cast<const int4*>(P.arena);
                int4* out = reinterpret_cast<int4*>(a.out);
                for (int t = warp - 4; t < nq; t += 4) {
                    if (s_code[t] != 1) continue;
                    const long long row = (long long)s_value[t];
                    for (int c = lane; c < vout; c += 32)
                        stg_stream(out + (long long)(q0 + t) * vout + c, ldg_stream(arena + row * vout + c));
                }
            }
        } else if (!(P.dbg & 1)) {
            wbase = cur_head + (unsigned long long)chunk * WCHUNK;
            wtail = ld_u64_cg(&st->free_tail);  // stable from barrier 1 to barrier 3 but for the evictions themselves
            const unsigned cnt = ps_window_bitmap(kv, wbase, a.t_set, wslot, s_mask, s_pref, &s_cnt);
            if (tid == 0) g->wcount[chunk] = cnt;
            PS_MARK(7);
        }
        PS_MARK(5);
        grid_arrive(g);
        target += n_cta;
        grid_wait(g, target, st);
        PS_MARK(2);

        // ================= phase C: epilogue (UNPACK) | evictions =================
        if (is_req) {
            const bool any_dup = ld_u64_cg(&g->dup_tag) == a.t_set;
            grid_arrive(g);  // barrier 3 protects the evictions: nothing it orders is written below
            target += n_cta;
            // warp 7 carries the barrier thread (its release fence waits for the CTA's outstanding writes): the epilogue
            // runs on warps 0-6 and synchronises among them only
            if (warp == 3 && !(P.dbg & 1)) ps_stash_refill(kv, s_stash, &s_nstash, R);  // after barrier 2: takes and returns are done
            if (nq > 0 && warp < 7) {
                if (tid < nq && s_code[tid] != 1) {  // which cached row this op's result goes to (-1: none)
                    int row = s_row[tid];
                    if (any_dup) {  // only the last op on a slot (largest stamp) owns its value
                        const Slot* sl = &kv.slots[s_slot[tid]];
                        row = ld_u64_cg(&sl->stamp) == a.t_set + (unsigned long long)(q0 + tid) ? (int)ld_u64_cg(&sl->value) : -1;
                    }
                    s_row[tid] = row;
                }
                tc::mbar_wait(&s_bar, mma_phase & 1u);
                ++mma_phase;
                tc::fence_after_thread_sync();
                PS_MARK(7);
                float* stage = reinterpret_cast<float*>(a_cur);  // the MMAs have consumed the tile
                const int pitch = N + 4;
                if (warp < 4 && warp * 32 < nq) {
                    const int row = warp * 32 + lane;
                    for (int c0 = 0; c0 < N; c0 += 32) {
                        uint32_t v[32];
                        tc::tmem_ld_32x32(s_tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            if (c0 + j < N) {
                                float4 o;
                                const float4 bv = *reinterpret_cast<const float4*>(s_bias + c0 + j);
                                o.x = __uint_as_float(v[j + 0]) + bv.x;
                                o.y = __uint_as_float(v[j + 1]) + bv.y;
                                o.z = __uint_as_float(v[j + 2]) + bv.z;
                                o.w = __uint_as_float(v[j + 3]) + bv.w;
                                *reinterpret_cast<float4*>(stage + (size_t)row * pitch + c0 + j) = o;
                            }
                        }
                    }
                }
                tc::fence_before_thread_sync();
                asm volatile("bar.sync 1, 224;" ::: "memory");
                int4* out = reinterpret_cast<int4*>(a.out);
                int4* arena = reinterpret_cast<int4*>(P.arena);
                for (int e = tid; e < nq * vout; e += PS_THREADS - 32) {
                    const int rr = e / vout, c = e - rr * vout;
                    if (s_code[rr] == 1) continue;
                    const int4 o = *reinterpret_cast<const int4*>(stage + (size_t)rr * pitch + c * 4);
                    stg_stream(out + (long long)(q0 + rr) * vout + c, o);
                    const int row = s_row[rr];
                    if (row >= 0) stg_stream(arena + (long long)row * vout + c, o);
                }
            }
            PS_MARK(6);
            grid_wait(g, target, st);
        } else {
            // every insert is in: _evict_if_needed's count for the whole batch (size is not written in this phase)
            const unsigned long long size = ld_u64_cg(&st->size);
            const unsigned long long e = size > kv.max_size && !(P.dbg & 1) ? size - kv.max_size : 0ull;
            ps_window_evict(kv, g, chunk, P.n_chunks, wbase, wtail, a.t_set, e, wslot, s_mask, s_pref, wevicted);
            grid_arrive(g);
            target += n_cta;
            grid_wait(g, target, st);
            if (e > 0) cur_head = ld_u64_cg(&g->hnew);  // the head every window CTA uses for the next batch
            if (chunk == 0 && tid == 0) ps_settle_evictions(kv, g, e);  // before this CTA arrives at the next barrier
        }
        PS_MARK(3);
    }
    if (!is_req && !(P.dbg & 1)) ps_window_relax(kv, wevicted);  // after the last barrier nothing else runs on the table
    if (is_req && warp == 3) ps_stash_return(kv, s_stash, &s_nstash);  // the unused stock goes back to the table's queue
    // every CTA has passed the last barrier once CTA 0 has: the next launch starts counting from here
    if (cta == 0 && tid == PS_BAR_THREAD) g->bar_base = target;
    trace_mark(sr, TR_STEP_END);
    if (is_req && warp == 0) {
        tc::fence_after_thread_sync();
        tc::tmem_dealloc(s_tmem, ncols);
    }
}

// ---- host side ------------------------------------------------------------------------------------------
static size_t ps_smem_bytes(int K, int N) { return (size_t)N * K * 4 + 2 * (size_t)PS_M * K * 4 + 1024; }

die_status_t pipe_step_grid_alloc(void** out) {
    StepGrid* g = nullptr;
    DIE_CUDA(cudaMalloc(&g, sizeof(StepGrid)));
    DIE_CUDA(cudaMemset(g, 0, sizeof(StepGrid)));
    *out = g;
    return DIE_OK;
}

die_status_t pipe_step_marks(void* grid, unsigned long long out[8]) {
    DIE_CUDA(cudaMemcpy(out, &static_cast<StepGrid*>(grid)->ts[0], 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return DIE_OK;
}
die_status_t pipe_step_cta_marks(void* grid, int n_cta, unsigned long long* out) {
    if (n_cta > 160) n_cta = 160;
    DIE_CUDA(cudaMemcpy(out, &static_cast<StepGrid*>(grid)->cta_ts[0][0], (size_t)n_cta * 8 * sizeof(unsigned long long),
                        cudaMemcpyDeviceToHost));
    return DIE_OK;
}

// Shape of the persistent launch for batches of up to B requests, or false when this (B, K, N, window) does not
// fit one co-resident grid -- the caller then uses the three-kernel sequence.
bool pipe_step_shape(int B, int K, int N, int n_chunks, int* R_out, int* n_req_out) {
    static int enabled = -1, max_ctas = 0, forced_r = 0;
    if (enabled < 0) {
        const char* e = getenv("DIE_PERSISTENT");
        enabled = e ? atoi(e) : 1;
        if (const char* r = getenv("DIE_STEP_R")) forced_r = atoi(r);
    }
    if (!enabled || B <= 0) return false;
    if (K % 32 != 0 || N % 16 != 0 || N < 16 || N > 256 || n_chunks < 1 || n_chunks > PS_MAX_CHUNKS) return false;
    const size_t smem = ps_smem_bytes(K, N);
    if (smem > 220 * 1024 || (size_t)2 * PS_M * K * 4 < (size_t)ROUTE_STAGE_BYTES + 32) return false;
    if (max_ctas == 0) {
        if (cudaFuncSetAttribute(pipe_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024) != cudaSuccess) {
            cudaGetLastError();
            max_ctas = -1;
        } else {
            max_ctas = num_sms();  // one CTA per SM: the tiles take most of an SM's shared memory
        }
    }
    if (max_ctas < 0) return false;
    if (forced_r) {  // DIE_STEP_R: exactly R requests per request CTA (experiments)
        const int n_req = ceil_div(B, forced_r);
        if (forced_r > PS_MAX_R || (size_t)forced_r * (N + 4) * 4 > (size_t)PS_M * K * 4 || n_req + n_chunks > max_ctas) return false;
        *R_out = forced_r;
        *n_req_out = n_req;
        return true;
    }
    // The fewest rows per request CTA (a multiple of 32) whose grid fits beside the window CTAs.  A grid sized for
    // owner_max (2 x the usual batch) therefore leaves SMs free: wanted, since the sharded step's copy and barrier
    // kernels run beside this one and a cooperative grid that needs every SM waits for each of them to go idle.
    // The batch itself is spread evenly over the request CTAs (ps_span), whatever R allows.
    for (int R = 32; R <= PS_MAX_R; R += 32) {
        if ((size_t)R * (N + 4) * 4 > (size_t)PS_M * K * 4) break;  // epilogue staging lives in the consumed A tile
        const int n_req = ceil_div(B, R);
        if (n_req + n_chunks <= max_ctas) {
            *R_out = R;
            *n_req_out = n_req;
            return true;
        }
    }
    return false;
}

die_status_t pipe_step_launch(const StepRef& sr, const QueueDesc* qv, int n_steps, int B, int n_chunks, int K, int N,
                              bool from_keys, const uint32_t* n_req_dev, uint32_t* digests, const float* W,
                              const float* bias, float* arena, void* grid, const KvMirrorArgs* mirror, cudaStream_t s,
                              const StepLanes* lanes) {
    StepParams P;
    memset(&P, 0, sizeof(P));
    if (!pipe_step_shape(B, K, N, n_chunks, &P.R, &P.n_req))
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "pipe_step_launch: shape B=%d K=%d N=%d does not fit the persistent step kernel", B, K, N);
    P.sr = sr;
    P.n_steps = n_steps;
    P.B = B;
    P.n_chunks = n_chunks;
    P.K = K;
    P.N = N;
    P.from_keys = from_keys ? 1 : 0;
    P.n_req_dev = n_req_dev;
    P.digests = reinterpret_cast<uint4*>(digests);
    P.W = W;
    P.bias = bias;
    P.arena = arena;
    P.g = static_cast<StepGrid*>(grid);
    if (mirror) P.mirror = *mirror;
    if (lanes) {
        P.lanes = lanes->lanes;
        P.lane0 = lanes->lane0;
    }
    static int dbg = -1;
    if (dbg < 0) {
        dbg = getenv("DIE_STEP_DEBUG") ? atoi(getenv("DIE_STEP_DEBUG")) : 0;
        cudaMemcpyToSymbol(g_ps_dbg, &dbg, sizeof(dbg));
        if (const char* e = getenv("DIE_STEP_POLL_NS")) {
            const int ns = atoi(e);
            cudaMemcpyToSymbol(g_ps_poll_ns, &ns, sizeof(ns));
        }
        if (const char* e = getenv("DIE_STEP_BAR_MODE")) {
            const int m = atoi(e);
            cudaMemcpyToSymbol(g_ps_bar_mode, &m, sizeof(m));
        }
    }
    P.dbg = dbg;
    QueueDesc q;
    memset(&q, 0, sizeof(q));
    if (qv) q = *qv;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(P.n_req + n_chunks);
    cfg.blockDim = dim3(PS_THREADS);
    cfg.dynamicSmemBytes = ps_smem_bytes(K, N);
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;  // the grid barriers need every CTA resident
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    DIE_LAUNCH(ST_FUSED_STEP, s, cudaLaunchKernelEx(&cfg, pipe_step_kernel, P, q, qv ? 1 : 0));
    return DIE_OK;
}

}  // namespace die
