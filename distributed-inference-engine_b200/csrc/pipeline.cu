// Composed hot path for one owner shard: route -> get -> pack misses -> model -> unpack -> set.
// This is SURVEY.md call-stack (E) -- the coordinator the reference's README.md:56-60 describes
// but never wrote.  Cached values are result rows in `arena`; the table hands out row handles.
//
// Fast sequence (max_batch <= kv_max_size, so a batch can never need more victims than the table
// holds and the sequential replay of kvstore.cu is unreachable) -- 3 launches per flushed batch:
//   1 pipe_get_kernel         MD5 of the keys (model_registry.py:159; hoisted into pipe_hash_all_kernel when
//                             a queue is replayed) + KVCache.get per request (kvstore.py:136-164); every CTA
//                             ranks its own misses in request order (batcher.py:211), the last CTA to finish
//                             scans the per-CTA counts and resets the batch's control block
//   2 pipe_claim_pack_kernel  CTAs side by side: insert the missing keys (kvstore.py:104-134) | PACK payload ->
//                             packed[rank], hits: cached result row -> out | bitmap of the still-current LRU-log
//                             records in the victim window (kvstore.py:82-93)
//   3 linear_tf32_kernel<SCATTER> (model_tc.cu)  packed -> GEMM -> UNPACK in the epilogue: out[q] and the cached
//                             row (batcher.py:225-227); extra CTAs evict the batch's LRU victims.
//     echo / fp32 models: model kernel + pipe_unpack_evict_kernel instead (4 launches)
// General sequence (tiny caches): the stand-alone kernels of kvstore.cu, 9 launches.
//
// Queues of batches (die_pipeline_run_device_many / run_host_many) are processed in groups of graph_steps
// steps replayed from one CUDA graph: see StepRef / QueueDesc in pipeline_device.cuh and replay_fast() below.
#include <cuda_bf16.h>

#include "model_internal.cuh"
#include "pipeline_device.cuh"
#include <vector>

#include "pipeline_internal.cuh"
#include "route_device.cuh"

struct die_pipeline {
    die_pipeline_config_t cfg;
    die_kv* kv;
    bool fast;
    uint64_t n_rows;
    float* arena;        // [n_rows, dim_out]
    float* W;            // [dim_out, dim_in]
    float* bias;         // [dim_out]
    uint32_t* digests;   // [B,4]
    int32_t* shard;      // [B]
    uint8_t* code;       // [B]
    uint64_t* value;     // [B] result-row handle of a hit
    int32_t* miss_rank;  // [B] (general sequence)
    uint8_t* lrank;      // [B] fast sequence: rank inside the get CTA, see PipeBufs
    int32_t* cta_miss;
    int32_t* cta_base;
    int32_t* miss_req;   // [B] inverse of miss_rank
    int32_t* set_slot;   // [B]
    int32_t* set_row;    // [B]
    uint32_t* wmask;     // victim window: validity bitmap, 1024 log records per chunk
    uint32_t* wcount;    // current records per chunk
    int n_chunks;
    PipeCtrl* ctrl;
    float* packed_in;    // [B, dim_in]
    float* packed_out;   // [B, dim_out]
    // DIE_MODEL_MLP_BF16: bf16 weights (W1 [H, D], W2 [D, H]) and bf16 staging of the packed misses
    void* mlp_w1;
    void* mlp_w2;
    float* mlp_b1;
    float* mlp_b2;
    void* mlp_x;         // [B, D] bf16
    void* mlp_h;         // [B, H] bf16
    void* mlp_y;         // [B, D] bf16
    // staging for the host-buffer entry points (NBUF sets: copy-in, compute and copy-out of consecutive
    // batches overlap in die_pipeline_run_host_many)
    static constexpr int NBUF = 3;
    uint8_t* d_key_bytes;   // NBUF x (max_key_bytes + 16)
    int64_t* d_key_off;     // NBUF x (B + 1)
    float* d_payload;       // NBUF x B x dim_in
    float* d_out;           // NBUF x B x dim_out
    uint8_t* d_hit;         // NBUF x B
    size_t key_stage_stride;
    cudaStream_t s_in, s_out;
    cudaEvent_t ev_in[NBUF], ev_done[NBUF], ev_out[NBUF];
    // captured replays of K steps (see replay_fast)
    struct Replay {
        int B, K;
        bool from_keys;
        const uint32_t* digests;
        const uint32_t* n_req_dev;
        const die::QueueDesc* qd;
        cudaGraphExec_t exec;
    };
    static constexpr int MAX_GRAPHS = 8;
    Replay graphs[MAX_GRAPHS];
    int n_graphs;
    int graph_steps;          // K; 0 = always launch from the stream
    cudaStream_t cap_stream;  // capture only: the caller's stream may be the legacy stream, which cannot capture
    die::QueueDesc* qdesc;    // device
    uint32_t* digests_all;    // [graph_steps][max_batch][4]: the digests of all batches of a replay
    // staging of die_pipeline_run_host_many's grouped form: NBUF sets of graph_steps batches each, so that one
    // copy per array moves a whole group and one replay computes it
    uint8_t* g_keys;
    int64_t* g_off;
    float* g_pay;
    float* g_out;
    uint8_t* g_hit;
    int g_steps;  // batches per set the buffers were sized for (0: not allocated)
    // the persistent step kernel (pipe_step.cu): TF32 linear model, batches that fit one co-resident grid
    bool persistent;
    void* sgrid;  // its device-resident control block
};

namespace die {

static_assert(PT == ROUTE_THREADS, "the fused get kernel stages keys with the route kernel's CTA shape");

// block-wide exclusive scan for PT threads
__device__ __forceinline__ unsigned block_scan_pt(unsigned cnt, unsigned& excl, unsigned* s_warp) {
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
        unsigned w = lane < PT / 32 ? s_warp[lane] : 0u;
#pragma unroll
        for (int o = 1; o < PT / 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += v;
        }
        if (lane < PT / 32) s_warp[lane] = w;
    }
    __syncthreads();
    excl = (warp ? s_warp[warp - 1] : 0u) + (incl - cnt);
    return s_warp[PT / 32 - 1];
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ unsigned ld_u8_cg(const uint8_t* p) {
    unsigned r;
    asm volatile("ld.global.cg.u8 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}

// Tail of the get kernel, run by the last CTA.  Every CTA has ranked its own misses (request order inside the
// CTA) and published their count; the tail only scans those counts -- one value per CTA, not one per request --
// and resets the batch's control block.  rank of miss q = cta_base[q / PT] + lrank[q]  (batcher.py:211 order).
__device__ void pipe_rank_tail(const KvView& kv, const PipeBufs& pb, int B, int n_cta) {
    __shared__ unsigned s_warp[PT / 32];
    const int tid = threadIdx.x;
    unsigned base = 0;
    for (int c0 = 0; c0 < n_cta; c0 += PT) {
        const int c = c0 + tid;
        const unsigned cnt = c < n_cta ? (unsigned)__ldcg(pb.cta_miss + c) : 0u;
        unsigned excl;
        const unsigned total = block_scan_pt(cnt, excl, s_warp);
        if (c < n_cta) pb.cta_base[c] = (int32_t)(base + excl);
        base += total;
        __syncthreads();
    }
    KvState* st = kv.st;
    if (tid == 0) {
        const unsigned long long size = ld_u64_cg(&st->size);
        pb.ctrl->ts[3] = globaltimer_ns();  // scan done
        pb.ctrl->n_miss = base;
        // upper bound of the evictions this batch's inserts can need (duplicate keys insert only once)
        pb.ctrl->e_ub = size + base > kv.max_size ? size + base - kv.max_size : 0ull;
        pb.ctrl->e_final = 0ull;
        pb.ctrl->ticket_get = 0u;
        pb.ctrl->ticket_claim = 0u;
        pb.ctrl->any_dup = 0u;
        atomicAdd(&st->hits, (unsigned long long)B - base);  // the batch's counters, added once (kvstore.py:150,163)
        atomicAdd(&st->misses, (unsigned long long)base);
    }
    for (int c = tid; c < pb.n_chunks; c += PT) pb.wcount[c] = 0u;
}

// Victim window, stage 1 (CTAs next to the claim / pack CTAs of the second kernel): chunk c of 1024 log
// records starting at the LRU head gets a bitmap of its records that are still current (= entries that are
// in the reference's _access_order, kvstore.py:71-78) and their count.  Chunks are independent; the order
// (oldest first, kvstore.py:89-91) is restored in stage 2 from the per-chunk counts.  Only pre-batch entries
// and this batch's hits are eligible (positions < t_set); concurrent inserts carry stamps >= t_set.
__device__ void pipe_window_chunk(const KvView& kv, const PipeBufs& pb, int chunk, unsigned long long t_set) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long head = ld_u64_cg(&kv.st->head);
    if (chunk == 0 && tid == 0) pb.ctrl->whead = head;
    constexpr int PER = WCHUNK / PT;  // records per thread, strided so that a warp covers 32 consecutive records
    uint32_t slot[PER];
    unsigned long long stamp[PER];
    const unsigned long long base = head + (unsigned long long)chunk * WCHUNK;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const unsigned long long p = base + (unsigned long long)(k * PT + tid);
        slot[k] = p < t_set ? __ldcg(&kv.log[p & kv.rmask]) : 0xffffffffu;
    }
#pragma unroll
    for (int k = 0; k < PER; ++k)
        stamp[k] = (unsigned long long)slot[k] <= kv.mask ? __ldcg(&kv.slots[slot[k]].stamp) : 0ull;
    unsigned cnt = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const unsigned long long p = base + (unsigned long long)(k * PT + tid);
        const unsigned m = __ballot_sync(0xffffffffu, p < t_set && stamp[k] == p);
        if (lane == 0) {
            pb.wmask[chunk * 32 + k * (PT / 32) + warp] = m;  // word w covers records [32w, 32w+32) of the chunk
            cnt += __popc(m);
        }
    }
    if (lane == 0 && cnt) atomicAdd(&pb.wcount[chunk], cnt);
}

template <bool FROM_KEYS>
__global__ void __launch_bounds__(PT)
pipe_get_kernel(StepRef sr, PipeBufs pb, uint4* __restrict__ digests, int B, const uint32_t* __restrict__ n_req_dev) {
    __shared__ __align__(16) uint8_t stage[FROM_KEYS ? ROUTE_STAGE_BYTES + 32 : 16];
    __shared__ bool s_last;
    const StepArgs a = step_resolve(sr);
    trace_mark(sr, TR_GET);
    const KvView& kv = a.kv;
    const uint8_t* __restrict__ key_bytes = a.key_bytes;
    const int64_t* __restrict__ key_off = a.key_off;
    const double now = a.now;
    const unsigned long long t_get = a.t_get;
    pb.code = a.code;
    // What this kernel reads BEFORE griddepcontrol.wait, and why that is ordered: the step descriptor (written by
    // queue_desc_set_kernel, which is followed by a full kernel boundary -- the graph launch -- never by a PDL
    // edge) and, when FROM_KEYS, the key bytes (host-to-device copies or another stream behind an event, never the
    // same-stream predecessor kernel).  Everything a predecessor kernel may have written -- the device-side batch
    // size, digests, table, control block -- is read after the wait.
    const int i = blockIdx.x * PT + threadIdx.x;
    u128 key{0ull, 0ull};
    Md5State st{};
    if (FROM_KEYS) st = cta_hash_keys(key_bytes, key_off, B, stage);  // rows past a device-side count are hashed and dropped
    pdl_wait();
    if (n_req_dev) B = min(B, (int)*n_req_dev);  // batch size known only on the device (after an exchange)
    const bool active = i < B;
    if (FROM_KEYS && active) {
        const uint4 d = make_uint4(st.a, st.b, st.c, st.d);
        digests[i] = d;
        key = fp_of(d);
    }
    if (!FROM_KEYS && active) key = fp_of(digests[i]);
    int code = 0;
    if (i == 0) pb.ctrl->ts[0] = pb.ctrl->ts[1] = globaltimer_ns();  // predecessor done, keys hashed (CTA 0)
    if (active) {
        unsigned long long value;
        long long s;
        code = kv_get_one<false>(kv, key, now, t_get + (unsigned long long)i, value, s);
        pb.code[i] = (uint8_t)code;
        pb.value[i] = value;
    }
    {   // rank of this miss among the misses of the CTA
        __shared__ unsigned s_wmiss[PT / 32];
        const bool miss = active && code != 1;
        const unsigned m = __ballot_sync(0xffffffffu, miss);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        if (lane == 0) s_wmiss[warp] = __popc(m);
        __syncthreads();
        unsigned r = __popc(m & ((1u << lane) - 1u)), tot = 0;
#pragma unroll
        for (int w = 0; w < PT / 32; ++w) {
            r += w < warp ? s_wmiss[w] : 0u;
            tot += s_wmiss[w];
        }
        if (active) pb.lrank[i] = (uint8_t)r;
        if (threadIdx.x == 0) pb.cta_miss[blockIdx.x] = (int32_t)tot;
    }
    const bool deleted = active && code == 2;  // rare: only expired entries
    if (__any_sync(0xffffffffu, deleted)) {
        warp_sub(&kv.st->size, deleted);
        warp_add(&kv.st->n_tomb, deleted);
    }
    // the last CTA to get here has every CTA's results in L2 and runs the batch-wide tail
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&pb.ctrl->ticket_get, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x == 0) pb.ctrl->ts[2] = globaltimer_ns();  // all CTAs probed, tail starts
    pipe_rank_tail(kv, pb, B, (int)gridDim.x);
    if (sr.q && threadIdx.x == 0) trace_mark_q(sr.q, sr.i, TR_GET_TAIL);
}

// ROUTE of all K batches of a replay in one launch (grid.y = step): hashing a batch's keys does not depend on
// the batches before it, so it leaves the step's critical path; the get kernels then start from the digests.
__global__ void __launch_bounds__(PT)
pipe_hash_all_kernel(const QueueDesc* __restrict__ q, uint4* __restrict__ digests, int B) {
    __shared__ __align__(16) uint8_t stage[ROUTE_STAGE_BYTES + 32];
    StepRef sr;
    sr.q = q;
    sr.i = (int)blockIdx.y;
    const StepArgs a = step_resolve(sr);
    const Md5State st = cta_hash_keys(a.key_bytes, a.key_off, B, stage);
    const int i = blockIdx.x * PT + threadIdx.x;
    if (i < B) digests[(size_t)blockIdx.y * B + i] = make_uint4(st.a, st.b, st.c, st.d);
}

// CTAs [0, g_claim): KVCache.set of every miss (all known absent).  Last n_chunks CTAs: LRU victim window.
// CTAs in between: PACK + hit copy.
__global__ void __launch_bounds__(PT)
pipe_claim_pack_kernel(StepRef sr, PipeBufs pb, const uint4* __restrict__ digests, int B,
                       const uint32_t* __restrict__ n_req_dev, int g_claim, const int4* __restrict__ arena, int vin,
                       int vout, int4* __restrict__ packed_in) {
    const StepArgs a = step_resolve(sr);
    pdl_wait();
    trace_mark(sr, TR_CLAIM);
    const KvView& kv = a.kv;
    const double now = a.now;
    const unsigned long long t_set = a.t_set;
    const int4* __restrict__ payload = reinterpret_cast<const int4*>(a.payload);
    int4* __restrict__ out = reinterpret_cast<int4*>(a.out);
    pb.code = a.code;
    if (n_req_dev) B = min(B, (int)*n_req_dev);
    if ((int)blockIdx.x < g_claim) {
        __shared__ bool s_last;
        const int i = blockIdx.x * PT + threadIdx.x;
        const bool active = i < B && pb.code[i] != 1;
        bool claimed = false, from_empty = false;
        if (active) {
            const u128 key = fp_of(digests[i]);
            const long long s = kv_claim_absent(kv, key, claimed, from_empty);
            Slot* sl = &kv.slots[s];
            const unsigned long long stamp = t_set + (unsigned long long)i;
            atomicMax(&sl->stamp, stamp);
            int32_t row = -1;
            if (claimed) {
                row = (int32_t)acquire_row(kv);
                sl->value = (unsigned long long)row;
                sl->created_at = now;
                sl->ttl = kv.default_ttl;
                sl->last_accessed = now;
                sl->access_count = 0u;
            } else {
                pb.ctrl->any_dup = 1u;  // same key twice in the batch: the later op owns the cached row
            }
            kv.log[stamp & kv.rmask] = (uint32_t)s;
            pb.set_slot[i] = (int32_t)s;
            pb.set_row[i] = row;
        }
        block_add(&kv.st->size, active && claimed);
        block_add(&kv.st->n_empty_claimed, active && claimed && from_empty);
        block_sub(&kv.st->n_tomb, active && claimed && !from_empty);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicAdd(&pb.ctrl->ticket_claim, 1u) == (unsigned)g_claim - 1;
        __syncthreads();
        if (s_last && threadIdx.x == 0) {
            __threadfence();
            const unsigned long long size = ld_u64_cg(&kv.st->size);
            pb.ctrl->e_final = size > kv.max_size ? size - kv.max_size : 0ull;  // _evict_if_needed count
        }
        return;
    }
    const int first_window = (int)gridDim.x - pb.n_chunks;
    if ((int)blockIdx.x >= first_window) {
        pipe_window_chunk(kv, pb, (int)blockIdx.x - first_window, t_set);
        return;
    }
    const int vmax = vin > vout ? vin : vout;
    const long long total = (long long)B * vmax;
    const long long stride = (long long)(first_window - g_claim) * PT;
    for (long long v = (long long)(blockIdx.x - g_claim) * PT + threadIdx.x; v < total; v += stride) {
        const long long q = v / vmax;
        const int c = (int)(v - q * vmax);
        if (pb.code[q] == 1) {
            if (c < vout) stg_stream(out + q * vout + c, ldg_stream(arena + (long long)pb.value[q] * vout + c));
        } else {
            const int rank = pb.cta_base[q / PT] + (int)pb.lrank[q];
            if (c < vin) stg_stream(packed_in + (long long)rank * vin + c, ldg_stream(payload + q * vin + c));
            if (c == 0) pb.miss_req[rank] = (int32_t)q;  // rank -> request, for the fused GEMM epilogue
        }
    }
}

// CTAs [0, g_evict): evict the e_final least-recently-used entries.  Rest: UNPACK + store the cached rows.
__global__ void __launch_bounds__(PT)
pipe_unpack_evict_kernel(StepRef sr, PipeBufs pb, const int4* __restrict__ packed_out, int B,
                         const uint32_t* __restrict__ n_req_dev, int g_evict, int vout, int4* __restrict__ arena) {
    const StepArgs a = step_resolve(sr);
    pdl_wait();
    trace_mark(sr, TR_UNPACK);
    const KvView& kv = a.kv;
    const unsigned long long t_set = a.t_set;
    int4* __restrict__ out = reinterpret_cast<int4*>(a.out);
    pb.code = a.code;
    if (n_req_dev) B = min(B, (int)*n_req_dev);
    if ((int)blockIdx.x < g_evict) {
        pipe_evict_cta(kv, pb, (int)blockIdx.x, t_set);
        return;
    }
    const bool any_dup = pb.ctrl->any_dup != 0u;
    const long long total = (long long)B * vout;
    const long long stride = (long long)(gridDim.x - g_evict) * PT;
    for (long long v = (long long)(blockIdx.x - g_evict) * PT + threadIdx.x; v < total; v += stride) {
        const long long q = v / vout;
        const int c = (int)(v - q * vout);
        if (pb.code[q] == 1) continue;
        const int4 y = ldg_stream(packed_out + (long long)(pb.cta_base[q / PT] + (int)pb.lrank[q]) * vout + c);
        stg_stream(out + v, y);
        if (c == 0) pb.code[q] = 0;  // callers see 1 = served from the cache, 0 = computed (2 -> 0)
        long long row = pb.set_row[q];
        if (any_dup) {
            const Slot* sl = &kv.slots[pb.set_slot[q]];
            row = sl->stamp == t_set + (unsigned long long)q ? (long long)sl->value : -1;
        }
        if (row >= 0) stg_stream(arena + row * vout + c, y);
    }
}

// ---- general sequence (tiny caches), built from the stand-alone kernels ------------------------------
__global__ void __launch_bounds__(1024)
miss_scan_kernel(const uint8_t* __restrict__ code, int B, int32_t* __restrict__ rank, uint32_t* __restrict__ n_miss) {
    __shared__ unsigned s_warp[32];
    unsigned base = 0;
    for (int pos = 0; pos < B; pos += 1024 * 4) {
        const int q0 = pos + threadIdx.x * 4;
        bool miss[4];
        unsigned cnt = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            miss[k] = q0 + k < B && code[q0 + k] != 1;
            cnt += miss[k];
        }
        unsigned excl;
        const unsigned total = block_scan_1024(cnt, excl, s_warp);
        unsigned r = base + excl;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (q0 + k < B) rank[q0 + k] = miss[k] ? (int32_t)r++ : -1;
        base += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_miss = base;
}

__global__ void __launch_bounds__(256)
pack_hits_kernel(const int4* __restrict__ payload, const uint8_t* __restrict__ code,
                 const int32_t* __restrict__ rank, const unsigned long long* __restrict__ value,
                 const int4* __restrict__ arena, long long B, int vin, int vout, int4* __restrict__ packed_in,
                 int4* __restrict__ out) {
    const int vmax = vin > vout ? vin : vout;
    const long long total = B * vmax;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long q = v / vmax;
        const int c = (int)(v - q * vmax);
        if (code[q] == 1) {
            if (c < vout) stg_stream(out + q * vout + c, ldg_stream(arena + (long long)value[q] * vout + c));
        } else if (c < vin) {
            stg_stream(packed_in + (long long)rank[q] * vin + c, ldg_stream(payload + q * vin + c));
        }
    }
}

__global__ void __launch_bounds__(256)
echo_rows_kernel(const int4* __restrict__ in, int4* __restrict__ out, int vrow, const uint32_t* __restrict__ n_rows) {
    const long long total = (long long)*n_rows * vrow;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256)
        stg_stream(out + v, ldg_stream(in + v));
}

__global__ void __launch_bounds__(256)
unpack_store_kernel(KvView kv, const int4* __restrict__ packed_out, uint8_t* __restrict__ code,
                    const int32_t* __restrict__ rank, const int32_t* __restrict__ set_slot, unsigned long long t_set,
                    long long B, int vout, int4* __restrict__ out, int4* __restrict__ arena) {
    const long long total = B * vout;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long q = v / vout;
        const int c = (int)(v - q * vout);
        if (code[q] == 1) continue;
        const int4 y = ldg_stream(packed_out + (long long)rank[q] * vout + c);
        stg_stream(out + v, y);
        if (c == 0) code[q] = 0;  // callers see 1 = served from the cache, 0 = computed (2 -> 0)
        const int32_t s = set_slot[q];
        if (s >= 0) {
            const Slot* sl = &kv.slots[s];
            if (sl->stamp == t_set + (unsigned long long)q) stg_stream(arena + (long long)sl->value * vout + c, y);
        }
    }
}

static int grid_256(long long items) {
    const long long blocks = (items + 255) / 256;
    const long long cap = (long long)num_sms() * 16;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

// fp32 rows <-> bf16 rows around the MLP (the interface keeps fp32 payloads and results; the model computes in bf16).
// n_rows_dev: only the packed misses are converted.
__global__ void __launch_bounds__(256)
rows_f32_to_bf16_kernel(const float4* __restrict__ in, uint2* __restrict__ out, long long vec_per_row, long long max_rows,
                        const uint32_t* __restrict__ n_rows_dev) {
    const long long n = (n_rows_dev ? min((long long)*n_rows_dev, max_rows) : max_rows) * vec_per_row;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const float4 v = __ldg(in + i);
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        out[i] = make_uint2(*reinterpret_cast<const unsigned*>(&lo), *reinterpret_cast<const unsigned*>(&hi));
    }
}
__global__ void __launch_bounds__(256)
rows_bf16_to_f32_kernel(const uint2* __restrict__ in, float4* __restrict__ out, long long vec_per_row, long long max_rows,
                        const uint32_t* __restrict__ n_rows_dev) {
    const long long n = (n_rows_dev ? min((long long)*n_rows_dev, max_rows) : max_rows) * vec_per_row;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const uint2 v = in[i];
        const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.x));
        const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.y));
        out[i] = make_float4(lo.x, lo.y, hi.x, hi.y);
    }
}

static die_status_t run_model(die_pipeline* p, int B, cudaStream_t s) {
    const die_pipeline_config_t& c = p->cfg;
    const uint32_t* n_miss = &p->ctrl->n_miss;
    if (c.model == DIE_MODEL_ECHO) {
        const int vin = c.dim_in / 4;
        DIE_LAUNCH(ST_MODEL, s, echo_rows_kernel<<<grid_256((long long)B * vin), 256, 0, s>>>(
            reinterpret_cast<const int4*>(p->packed_in), reinterpret_cast<int4*>(p->packed_out), vin, n_miss));
        return DIE_OK;
    }
    if (c.model == DIE_MODEL_LINEAR_FP32)
        return linear_f32_launch(p->packed_in, p->W, p->bias, p->packed_out, B, c.dim_in, c.dim_out, n_miss, s);
    if (c.model == DIE_MODEL_MLP_BF16) {
        // The GEMMs run over all B rows of the staging buffers (the miss count stays on the device): rows past the
        // misses hold an earlier batch's values and their results are never read.
        const long long vec = c.dim_in / 4;
        const int grid = (int)std::min<long long>(((long long)B * vec + 255) / 256, (long long)num_sms() * 8);
        DIE_LAUNCH(ST_MODEL, s, rows_f32_to_bf16_kernel<<<grid, 256, 0, s>>>(
            reinterpret_cast<const float4*>(p->packed_in), static_cast<uint2*>(p->mlp_x), vec, B, n_miss));
        DIE_TRY(mlp_bf16_launch(p->mlp_x, p->mlp_w1, p->mlp_b1, p->mlp_w2, p->mlp_b2, p->mlp_h, p->mlp_y, B, c.dim_in, c.hidden, s));
        DIE_LAUNCH(ST_MODEL, s, rows_bf16_to_f32_kernel<<<grid, 256, 0, s>>>(
            static_cast<const uint2*>(p->mlp_y), reinterpret_cast<float4*>(p->packed_out), vec, B, n_miss));
        return DIE_OK;
    }
    return linear_tf32_launch(p->packed_in, p->W, p->bias, p->packed_out, B, c.dim_in, c.dim_out, n_miss, s);
}

static PipeBufs bufs_of(die_pipeline* p, uint8_t* code) {
    PipeBufs pb;
    pb.code = code;
    pb.value = reinterpret_cast<unsigned long long*>(p->value);
    pb.lrank = p->lrank;
    pb.cta_miss = p->cta_miss;
    pb.cta_base = p->cta_base;
    pb.miss_req = p->miss_req;
    pb.set_slot = p->set_slot;
    pb.set_row = p->set_row;
    pb.wmask = p->wmask;
    pb.wcount = p->wcount;
    pb.n_chunks = p->n_chunks;
    pb.ctrl = p->ctrl;
    return pb;
}

// The launches of one step.  from_keys: hash the keys in the get kernel (digests are written to `digests`);
// otherwise `digests` are inputs.
static die_status_t launch_fast(die_pipeline* p, const StepRef& sr, bool from_keys, uint32_t* digests, int B,
                                const uint32_t* n_req_dev, cudaStream_t s) {
    const die_pipeline_config_t& c = p->cfg;
    const int vin = c.dim_in / 4, vout = c.dim_out / 4;
    int ps_r, ps_n;
    if (p->persistent && pipe_step_shape(B, c.dim_in, c.dim_out, p->n_chunks, &ps_r, &ps_n))  // the whole step in one launch
        return pipe_step_launch(sr, nullptr, 1, B, p->n_chunks, c.dim_in, c.dim_out, from_keys, n_req_dev, digests, p->W,
                                p->bias, p->arena, p->sgrid, nullptr, s);
    const PipeBufs pb = bufs_of(p, nullptr);  // the kernels take the hit codes' address from the step
    const int g = ceil_div(B, PT);
    uint4* dg = reinterpret_cast<uint4*>(digests);
    if (from_keys)
        DIE_LAUNCH(ST_KV_GET, s, launch_pdl(pipe_get_kernel<true>, dim3(g), dim3(PT), 0, s, sr, pb, dg, B, n_req_dev));
    else
        DIE_LAUNCH(ST_KV_GET, s, launch_pdl(pipe_get_kernel<false>, dim3(g), dim3(PT), 0, s, sr, pb, dg, B, n_req_dev));
    const int g_pack = grid_256((long long)B * std::max(vin, vout));
    DIE_LAUNCH(ST_KV_CLAIM, s, launch_pdl(pipe_claim_pack_kernel, dim3(g + g_pack + p->n_chunks), dim3(PT), 0, s, sr, pb,
                                          static_cast<const uint4*>(dg), B, n_req_dev, g,
                                          reinterpret_cast<const int4*>(p->arena), vin, vout,
                                          reinterpret_cast<int4*>(p->packed_in)));
    if (c.model == DIE_MODEL_LINEAR_TF32) {
        // third and last launch: tensor-core GEMM whose epilogue is the UNPACK, plus the eviction CTAs
        TcScatter sc;
        sc.sr = sr;
        sc.miss_req = p->miss_req;
        sc.set_row = p->set_row;
        sc.set_slot = p->set_slot;
        sc.pb = pb;
        sc.any_dup = &p->ctrl->any_dup;
        sc.arena = p->arena;
        sc.tiles = 0;
        return linear_tf32_scatter_launch(p->packed_in, p->W, p->bias, B, c.dim_in, c.dim_out, &p->ctrl->n_miss, sc, s);
    }
    // echo is the identity (fake_model.py:52): the unpack kernel reads the packed batch directly
    const float* result = p->packed_in;
    if (c.model != DIE_MODEL_ECHO) {
        DIE_TRY(run_model(p, B, s));
        result = p->packed_out;
    }
    const int g_unpack = grid_256((long long)B * vout);
    const int g_evict = p->n_chunks * (WCHUNK / PT);
    DIE_LAUNCH(ST_UNPACK_STORE, s, launch_pdl(pipe_unpack_evict_kernel, dim3(g_evict + g_unpack), dim3(PT), 0, s, sr, pb,
                                              reinterpret_cast<const int4*>(result), B, n_req_dev, g_evict, vout,
                                              reinterpret_cast<int4*>(p->arena)));
    return DIE_OK;
}

static die_status_t run_fast(die_pipeline* p, const uint8_t* key_bytes, const int64_t* key_off, uint32_t* digests,
                             int B, const uint32_t* n_req_dev, const float* payload, double now, float* out,
                             uint8_t* hit_out, cudaStream_t s) {
    die_kv* kv = p->kv;
    DIE_TRY(kv_reserve(kv, (uint64_t)B, false, s));
    DIE_TRY(kv_reserve(kv, (uint64_t)B, true, s));
    StepRef sr;
    sr.q = nullptr;
    sr.i = 0;
    sr.a.kv = view_of(kv);
    sr.a.key_bytes = key_bytes;
    sr.a.key_off = key_off;
    sr.a.payload = payload;
    sr.a.out = out;
    sr.a.code = hit_out ? hit_out : p->code;
    sr.a.now = now;
    sr.a.t_get = kv->tail;
    sr.a.t_set = kv->tail + (unsigned long long)B;
    kv->tail += 2ull * (unsigned long long)B;
    return launch_fast(p, sr, key_off != nullptr, digests, B, n_req_dev, s);
}

// ---- replay of K steps from a CUDA graph -----------------------------------------------------------------
// The host cost of a step is its launches; a queue of batches pays it once per K steps instead: the K steps
// are captured once (their arguments come from the device-resident QueueDesc, see pipeline_device.cuh) and
// replayed with one graph launch after a one-thread kernel has rewritten the descriptor.
__global__ void queue_desc_set_kernel(QueueDesc* dst, QueueDesc v, KvMirrorArgs m) {
    *dst = v;
    kv_mirror_write(m.st, m.m, m.ops_seen, m.gen, m.seq);  // state before this replay -> the host's bounds
}

static bool graph_fits(const die_pipeline* p, int B, int K) {
    const die_kv* kv = p->kv;
    // the conservative host bounds of kv_reserve are taken for K steps at once: keep that a small part of
    // the LRU ring and of the EMPTY-slot reserve, or maintenance would run at every replay
    return p->fast && K > 0 && (uint64_t)2 * K * B <= kv->ring_cap / 8 && (uint64_t)K * B <= kv->capacity / 32;
}

static die_status_t graph_get(die_pipeline* p, bool from_keys, uint32_t* digests, int B, const uint32_t* n_req_dev,
                              QueueDesc* qd, int K, cudaGraphExec_t* out) {
    for (int g = 0; g < p->n_graphs; ++g) {
        const die_pipeline::Replay& r = p->graphs[g];
        if (r.B == B && r.K == K && r.from_keys == from_keys && r.digests == digests && r.n_req_dev == n_req_dev && r.qd == qd) {
            *out = r.exec;
            return DIE_OK;
        }
    }
    if (p->n_graphs == die_pipeline::MAX_GRAPHS) {  // recycle the oldest
        cudaGraphExecDestroy(p->graphs[0].exec);
        for (int g = 1; g < p->n_graphs; ++g) p->graphs[g - 1] = p->graphs[g];
        p->n_graphs -= 1;
    }
    if (!p->cap_stream) DIE_CUDA(cudaStreamCreateWithFlags(&p->cap_stream, cudaStreamNonBlocking));
    if (from_keys && !p->digests_all)  // before the capture starts: no allocation inside one
        DIE_CUDA(cudaMalloc(&p->digests_all, (size_t)std::max(p->graph_steps, K) * p->cfg.max_batch * 16));
    cudaGraph_t graph = nullptr;
    DIE_CUDA(cudaStreamBeginCapture(p->cap_stream, cudaStreamCaptureModeThreadLocal));
    die_status_t st = DIE_OK;
    if (from_keys) {  // one hash launch for the whole replay
        if (p->digests_all) {
            pipe_hash_all_kernel<<<dim3(ceil_div(B, PT), K), PT, 0, p->cap_stream>>>(qd, reinterpret_cast<uint4*>(p->digests_all), B);
            if (cudaGetLastError() != cudaSuccess) st = DIE_ERR_CUDA;
        }
    }
    for (int i = 0; i < K && st == DIE_OK; ++i) {
        StepRef sr;
        memset(&sr, 0, sizeof(sr));
        sr.q = qd;
        sr.i = i;
        if (from_keys && p->digests_all)
            st = launch_fast(p, sr, false, p->digests_all + (size_t)i * B * 4, B, n_req_dev, p->cap_stream);
        else
            st = launch_fast(p, sr, from_keys, digests, B, n_req_dev, p->cap_stream);
    }
    const cudaError_t e = cudaStreamEndCapture(p->cap_stream, &graph);
    if (st != DIE_OK) {
        if (graph) cudaGraphDestroy(graph);
        return st;
    }
    DIE_CUDA(e);
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    DIE_CUDA(ei);
    die_pipeline::Replay& r = p->graphs[p->n_graphs++];
    r.B = B;
    r.K = K;
    r.from_keys = from_keys;
    r.digests = digests;
    r.n_req_dev = n_req_dev;
    r.qd = qd;
    r.exec = exec;
    *out = exec;
    return DIE_OK;
}

// Run K steps described by `d` (everything but kv / tail0 filled in by the caller) as one replay.
static die_status_t replay_fast(die_pipeline* p, QueueDesc d, QueueDesc* qd, bool from_keys, uint32_t* digests, int B,
                                const uint32_t* n_req_dev, int K, cudaStream_t s) {
    die_kv* kv = p->kv;
    int ps_r, ps_n;
    if (p->persistent && from_keys && K <= p->graph_steps &&
        pipe_step_shape(B, p->cfg.dim_in, p->cfg.dim_out, p->n_chunks, &ps_r, &ps_n)) {
        // One persistent launch processes the K batches; its arguments travel in the launch itself, so there is
        // neither a graph nor a descriptor kernel: host cost = one launch per K batches.
        const KvMirrorArgs mirror = kv_mirror_args(kv);  // before this group's ops are reserved
        DIE_TRY(kv_reserve(kv, (uint64_t)2 * K * B, false, s));
        DIE_TRY(kv_reserve(kv, (uint64_t)K * B, true, s));
        d.kv = view_of(kv);
        d.batch = B;
        d.trace = trace_buffer();
        d.tail0 = kv->tail;
        kv->tail += 2ull * (unsigned long long)K * (unsigned long long)B;
        StepRef sr;
        memset(&sr, 0, sizeof(sr));
        DIE_TRY(pipe_step_launch(sr, &d, K, B, p->n_chunks, p->cfg.dim_in, p->cfg.dim_out, true, n_req_dev,
                                 p->digests_all, p->W, p->bias, p->arena, p->sgrid, &mirror, s));
        count_replay();
        return DIE_OK;
    }
    cudaGraphExec_t exec;
    DIE_TRY(graph_get(p, from_keys, digests, B, n_req_dev, qd, K, &exec));
    const KvMirrorArgs mirror = kv_mirror_args(kv);  // before this replay's ops are reserved
    DIE_TRY(kv_reserve(kv, (uint64_t)2 * K * B, false, s));
    DIE_TRY(kv_reserve(kv, (uint64_t)K * B, true, s));
    d.kv = view_of(kv);
    d.batch = B;
    d.trace = trace_buffer();
    d.tail0 = kv->tail;
    kv->tail += 2ull * (unsigned long long)K * (unsigned long long)B;
    queue_desc_set_kernel<<<1, 1, 0, s>>>(qd, d, mirror);
    DIE_LAUNCH_CHECK();
    DIE_CUDA(cudaGraphLaunch(exec, s));
    count_replay();
    count_launch((unsigned)(K * pipeline_launches_per_step(p)) + (from_keys ? 1u : 0u));
    return DIE_OK;
}

bool pipeline_graph_fits(const die_pipeline* p, int B_max, int K) { return graph_fits(p, B_max, K) && !prof_active(); }

die_status_t pipeline_owner_reserve(die_pipeline* p, int B_max, int K, cudaStream_t s, KvView* view,
                                    unsigned long long* tail0, KvMirrorArgs* mirror) {
    die_kv* kv = p->kv;
    if (mirror) *mirror = kv_mirror_args(kv);
    DIE_TRY(kv_reserve(kv, (uint64_t)2 * K * B_max, false, s));
    DIE_TRY(kv_reserve(kv, (uint64_t)K * B_max, true, s));
    *view = view_of(kv);
    *tail0 = kv->tail;
    kv->tail += 2ull * (unsigned long long)K * (unsigned long long)B_max;
    return DIE_OK;
}

die_status_t pipeline_owner_launch(die_pipeline* p, const StepRef& sr, uint32_t* digests, int B_max,
                                   const uint32_t* n_req_dev, cudaStream_t s) {
    return launch_fast(p, sr, false, digests, B_max, n_req_dev, s);
}

bool pipeline_owner_group_supported(const die_pipeline* p, int B_max) {
    int r, n;
    // One persistent launch per group of 16 owner passes (exchange.cu's software-pipelined replays).  On by default;
    // DIE_XCHG_OWNER_GROUP=0 (read at die_xchg_create) keeps one launch per step inside one graph per group.
    const int on = getenv("DIE_XCHG_OWNER_GROUP") ? atoi(getenv("DIE_XCHG_OWNER_GROUP")) : 1;
    return on && p->persistent && pipe_step_shape(B_max, p->cfg.dim_in, p->cfg.dim_out, p->n_chunks, &r, &n);
}

die_status_t pipeline_owner_launch_group(die_pipeline* p, const QueueDesc* qd_dev, int first, int K, uint32_t* digests_ring,
                                         int B_max, const uint32_t* n_dev_ring, const StepLanes& lanes, cudaStream_t s) {
    StepRef sr;
    memset(&sr, 0, sizeof(sr));
    sr.q = qd_dev;  // device-resident descriptor, rewritten before every replay
    sr.i = first;
    return pipe_step_launch(sr, nullptr, K, B_max, p->n_chunks, p->cfg.dim_in, p->cfg.dim_out, false, n_dev_ring,
                            digests_ring, p->W, p->bias, p->arena, p->sgrid, nullptr, s, &lanes);
}

int pipeline_launches_per_step(const die_pipeline* p) {
    if (p->cfg.model == DIE_MODEL_MLP_BF16) return 7;  // get, claim + pack, to bf16, two GEMMs, to fp32, unpack + evict
    return p->persistent ? 1 : (p->cfg.model == DIE_MODEL_LINEAR_FP32 ? 4 : 3);
}

static die_status_t run_general(die_pipeline* p, const uint32_t* digests, int B, const float* payload, double now,
                                float* out, uint8_t* hit_out, cudaStream_t s) {
    const die_pipeline_config_t& c = p->cfg;
    const int vin = c.dim_in / 4, vout = c.dim_out / 4;
    uint8_t* code = hit_out ? hit_out : p->code;
    uint32_t* n_miss = &p->ctrl->n_miss;
    DIE_TRY(kv_get_batch_impl(p->kv, digests, B, now, code, p->value, nullptr, s));
    DIE_LAUNCH(ST_MISS_SCAN, s, miss_scan_kernel<<<1, 1024, 0, s>>>(code, B, p->miss_rank, n_miss));
    DIE_LAUNCH(ST_PACK_HITS, s, pack_hits_kernel<<<grid_256((long long)B * std::max(vin, vout)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(payload), code, p->miss_rank,
        reinterpret_cast<const unsigned long long*>(p->value), reinterpret_cast<const int4*>(p->arena), B, vin, vout,
        reinterpret_cast<int4*>(p->packed_in), reinterpret_cast<int4*>(out)));
    DIE_TRY(run_model(p, B, s));
    const unsigned long long t_set = p->kv->tail;
    DIE_TRY(kv_set_batch_impl(p->kv, digests, nullptr, nullptr, B, now, 0, code, 1, n_miss, p->set_slot, nullptr, s));
    DIE_LAUNCH(ST_UNPACK_STORE, s, unpack_store_kernel<<<grid_256((long long)B * vout), 256, 0, s>>>(
        view_of(p->kv), reinterpret_cast<const int4*>(p->packed_out), code, p->miss_rank, p->set_slot, t_set, B, vout,
        reinterpret_cast<int4*>(out), reinterpret_cast<int4*>(p->arena)));
    return DIE_OK;
}

}  // namespace die

using namespace die;

extern "C" die_status_t die_pipeline_create(const die_pipeline_config_t* cfg, die_pipeline_t** out) {
    if (!cfg || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: null argument");
    if (cfg->max_batch <= 0 || cfg->dim_in <= 0 || cfg->dim_out <= 0 || (cfg->dim_in & 3) || (cfg->dim_out & 3))
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: max_batch > 0 and dims multiple of 4 fp32 required");
    if (cfg->model == DIE_MODEL_ECHO && cfg->dim_in != cfg->dim_out)
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: echo needs dim_in == dim_out");
    if (cfg->model < DIE_MODEL_ECHO || cfg->model > DIE_MODEL_MLP_BF16)
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: unknown model %d", cfg->model);
    if (cfg->model == DIE_MODEL_MLP_BF16 && (cfg->dim_in != cfg->dim_out || cfg->hidden <= 0 || (cfg->hidden & 7) || (cfg->dim_in & 7)))
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: the MLP needs dim_in == dim_out and hidden, both multiples of 8");
    if (cfg->model != DIE_MODEL_ECHO && (!cfg->weight || !cfg->bias))
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: linear model needs weight and bias");
    die_pipeline* p = new die_pipeline();
    memset(p, 0, sizeof(*p));
    p->cfg = *cfg;
    p->cfg.weight = nullptr;
    p->cfg.bias = nullptr;
    const uint64_t B = (uint64_t)cfg->max_batch;
    p->fast = B <= cfg->kv_max_size;
    p->n_rows = cfg->kv_max_size + 2 * B + 256;  // spare rows: see release_row (kv_internal.cuh) and pipe_step.cu
    DIE_TRY(kv_create_impl(cfg->kv_max_size, cfg->default_ttl, p->n_rows, B, &p->kv));
    if ((uint64_t)cfg->max_batch > p->kv->max_chunk)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_pipeline_create: max_batch %d exceeds the table's batch limit %llu",
                 cfg->max_batch, (unsigned long long)p->kv->max_chunk);
    DIE_CUDA(cudaMalloc(&p->arena, p->n_rows * cfg->dim_out * sizeof(float)));
    if (cfg->model == DIE_MODEL_MLP_BF16) {
        const size_t D = (size_t)cfg->dim_in, H = (size_t)cfg->hidden;
        std::vector<__nv_bfloat16> w(2 * D * H);
        for (size_t i = 0; i < 2 * D * H; ++i) w[i] = __float2bfloat16_rn(cfg->weight[i]);
        DIE_CUDA(cudaMalloc(&p->mlp_w1, H * D * 2));
        DIE_CUDA(cudaMalloc(&p->mlp_w2, D * H * 2));
        DIE_CUDA(cudaMemcpy(p->mlp_w1, w.data(), H * D * 2, cudaMemcpyHostToDevice));
        DIE_CUDA(cudaMemcpy(p->mlp_w2, w.data() + H * D, D * H * 2, cudaMemcpyHostToDevice));
        DIE_CUDA(cudaMalloc(&p->mlp_b1, H * sizeof(float)));
        DIE_CUDA(cudaMalloc(&p->mlp_b2, D * sizeof(float)));
        DIE_CUDA(cudaMemcpy(p->mlp_b1, cfg->bias, H * sizeof(float), cudaMemcpyHostToDevice));
        DIE_CUDA(cudaMemcpy(p->mlp_b2, cfg->bias + H, D * sizeof(float), cudaMemcpyHostToDevice));
        DIE_CUDA(cudaMalloc(&p->mlp_x, B * D * 2));
        DIE_CUDA(cudaMalloc(&p->mlp_h, B * H * 2));
        DIE_CUDA(cudaMalloc(&p->mlp_y, B * D * 2));
        DIE_CUDA(cudaMemset(p->mlp_x, 0, B * D * 2));
    } else if (cfg->model != DIE_MODEL_ECHO) {
        DIE_CUDA(cudaMalloc(&p->W, (size_t)cfg->dim_out * cfg->dim_in * sizeof(float)));
        DIE_CUDA(cudaMalloc(&p->bias, (size_t)cfg->dim_out * sizeof(float)));
        DIE_CUDA(cudaMemcpy(p->W, cfg->weight, (size_t)cfg->dim_out * cfg->dim_in * sizeof(float),
                            cudaMemcpyHostToDevice));
        DIE_CUDA(cudaMemcpy(p->bias, cfg->bias, (size_t)cfg->dim_out * sizeof(float), cudaMemcpyHostToDevice));
    }
    DIE_CUDA(cudaMalloc(&p->digests, B * 16));
    DIE_CUDA(cudaMalloc(&p->shard, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->code, B));
    DIE_CUDA(cudaMalloc(&p->value, B * sizeof(uint64_t)));
    DIE_CUDA(cudaMalloc(&p->miss_rank, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->miss_req, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->lrank, B));
    DIE_CUDA(cudaMalloc(&p->cta_miss, (B / PT + 1) * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->cta_base, (B / PT + 1) * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->set_slot, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->set_row, B * sizeof(int32_t)));
    p->n_chunks = (int)((4 * B + WCHUNK - 1) / WCHUNK);  // window = 4 log records per request of the largest batch
    DIE_CUDA(cudaMalloc(&p->wmask, (size_t)p->n_chunks * 32 * sizeof(uint32_t)));
    DIE_CUDA(cudaMalloc(&p->wcount, (size_t)p->n_chunks * sizeof(uint32_t)));
    DIE_CUDA(cudaMalloc(&p->ctrl, sizeof(PipeCtrl)));
    DIE_CUDA(cudaMemset(p->ctrl, 0, sizeof(PipeCtrl)));
    DIE_CUDA(cudaMalloc(&p->packed_in, B * cfg->dim_in * sizeof(float)));
    DIE_CUDA(cudaMalloc(&p->packed_out, B * cfg->dim_out * sizeof(float)));
    const size_t key_bytes = cfg->max_key_bytes > 0 ? (size_t)cfg->max_key_bytes : B * 64;
    p->cfg.max_key_bytes = (int32_t)key_bytes;
    constexpr int NB = die_pipeline::NBUF;
    p->key_stage_stride = (key_bytes + 16 + 255) / 256 * 256;
    DIE_CUDA(cudaMalloc(&p->d_key_bytes, NB * p->key_stage_stride));
    DIE_CUDA(cudaMalloc(&p->d_key_off, NB * (B + 1) * sizeof(int64_t)));
    DIE_CUDA(cudaMalloc(&p->d_payload, NB * B * cfg->dim_in * sizeof(float)));
    DIE_CUDA(cudaMalloc(&p->d_out, NB * B * cfg->dim_out * sizeof(float)));
    DIE_CUDA(cudaMalloc(&p->d_hit, NB * B));
    DIE_CUDA(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
    DIE_CUDA(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < NB; ++i) {
        DIE_CUDA(cudaEventCreateWithFlags(&p->ev_in[i], cudaEventDisableTiming));
        DIE_CUDA(cudaEventCreateWithFlags(&p->ev_done[i], cudaEventDisableTiming));
        DIE_CUDA(cudaEventCreateWithFlags(&p->ev_out[i], cudaEventDisableTiming));
    }
    DIE_CUDA(cudaMalloc(&p->qdesc, sizeof(QueueDesc)));
    p->graph_steps = 16;
    if (const char* e = getenv("DIE_GRAPH_STEPS")) p->graph_steps = atoi(e);  // 0: every step launched from the stream
    int ps_r, ps_n;
    p->persistent = p->fast && cfg->model == DIE_MODEL_LINEAR_TF32 &&
                    pipe_step_shape(cfg->max_batch, cfg->dim_in, cfg->dim_out, p->n_chunks, &ps_r, &ps_n);
    if (p->persistent) {
        DIE_TRY(pipe_step_grid_alloc(&p->sgrid));
        if (p->graph_steps > 0) DIE_CUDA(cudaMalloc(&p->digests_all, (size_t)p->graph_steps * B * 16));
    }
    *out = p;
    return DIE_OK;
}

extern "C" die_status_t die_pipeline_destroy(die_pipeline_t* p) {
    if (!p) return DIE_OK;
    die_kv_destroy(p->kv);
    void* bufs[] = {p->mlp_w1, p->mlp_w2, p->mlp_b1, p->mlp_b2, p->mlp_x, p->mlp_h, p->mlp_y, p->arena, p->W, p->bias, p->digests, p->shard, p->code, p->value, p->miss_rank, p->miss_req, p->lrank, p->cta_miss, p->cta_base,
                    p->set_slot, p->set_row, p->wmask, p->wcount, p->ctrl, p->packed_in, p->packed_out, p->d_key_bytes,
                    p->d_key_off, p->d_payload, p->d_out, p->d_hit};
    for (void* b : bufs) cudaFree(b);
    cudaFree(p->qdesc);
    cudaFree(p->sgrid);
    cudaFree(p->digests_all);
    cudaFree(p->g_keys); cudaFree(p->g_off); cudaFree(p->g_pay); cudaFree(p->g_out); cudaFree(p->g_hit);
    for (int g = 0; g < p->n_graphs; ++g) cudaGraphExecDestroy(p->graphs[g].exec);
    if (p->cap_stream) cudaStreamDestroy(p->cap_stream);
    if (p->s_in) cudaStreamDestroy(p->s_in);
    if (p->s_out) cudaStreamDestroy(p->s_out);
    for (int i = 0; i < die_pipeline::NBUF; ++i) {
        if (p->ev_in[i]) cudaEventDestroy(p->ev_in[i]);
        if (p->ev_done[i]) cudaEventDestroy(p->ev_done[i]);
        if (p->ev_out[i]) cudaEventDestroy(p->ev_out[i]);
    }
    delete p;
    return DIE_OK;
}

extern "C" die_kv_t* die_pipeline_kv(die_pipeline_t* p) { return p ? p->kv : nullptr; }

extern "C" die_status_t die_pipeline_debug_marks_sync(die_pipeline_t* p, uint64_t out[8]) {
    if (!p || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_debug_marks_sync: null argument");
    DIE_CUDA(cudaDeviceSynchronize());
    if (p->persistent) {  // marks of the last batch of the persistent kernel: [0] start, [1..3] after barriers 1..3
        unsigned long long t[8];
        DIE_TRY(pipe_step_marks(p->sgrid, t));
        for (int i = 0; i < 8; ++i) out[i] = t[i];
        return DIE_OK;
    }
    PipeCtrl h;
    DIE_CUDA(cudaMemcpy(&h, p->ctrl, sizeof(h), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 8; ++i) out[i] = h.ts[i];
    return DIE_OK;
}

extern "C" die_status_t die_pipeline_debug_cta_marks_sync(die_pipeline_t* p, int32_t n_cta, uint64_t* out) {
    if (!p || !out || n_cta < 1) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_debug_cta_marks_sync: bad arguments");
    if (!p->persistent) DIE_FAIL(DIE_ERR_UNSUPPORTED, "die_pipeline_debug_cta_marks_sync: this pipeline does not use the persistent step kernel");
    DIE_CUDA(cudaDeviceSynchronize());
    return pipe_step_cta_marks(p->sgrid, n_cta, reinterpret_cast<unsigned long long*>(out));
}

extern "C" die_status_t die_pipeline_run_owner_device(die_pipeline_t* p, const uint32_t* digests, int32_t B,
                                                      const float* payload, double now, float* out,
                                                      uint8_t* hit_out, die_stream_t stream) {
    if (!p || B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_owner_device: bad B");
    if (B == 0) return DIE_OK;
    if (!digests || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_owner_device: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (p->fast) return run_fast(p, nullptr, nullptr, const_cast<uint32_t*>(digests), B, nullptr, payload, now, out, hit_out, s);
    return run_general(p, digests, B, payload, now, out, hit_out, s);
}

// Owner-side pass whose request count lives on the device (*n_req_dev <= B_max): what a rank runs on the
// requests its peers just wrote into its receive buffer, without reading the count back to the host.
extern "C" die_status_t die_pipeline_run_owner_device_n(die_pipeline_t* p, const uint32_t* digests, int32_t B_max,
                                                        const uint32_t* n_req_dev, const float* payload, double now,
                                                        float* out, uint8_t* hit_out, die_stream_t stream) {
    if (!p || B_max <= 0 || B_max > p->cfg.max_batch || !n_req_dev)
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_owner_device_n: bad arguments");
    if (!digests || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_owner_device_n: null buffer");
    if (!p->fast) DIE_FAIL(DIE_ERR_UNSUPPORTED, "die_pipeline_run_owner_device_n: needs max_batch <= kv_max_size");
    return run_fast(p, nullptr, nullptr, const_cast<uint32_t*>(digests), B_max, n_req_dev, payload, now, out, hit_out,
                    static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_pipeline_run_device(die_pipeline_t* p, const uint8_t* key_bytes, const int64_t* key_off,
                                                int32_t B, const float* payload, double now, float* out,
                                                uint8_t* hit_out, uint32_t* digest_out, die_stream_t stream) {
    if (!p || B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device: bad B");
    if (B == 0) return DIE_OK;
    if (!key_off || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    uint32_t* dg = digest_out ? digest_out : p->digests;
    if (p->fast) return run_fast(p, key_bytes, key_off, dg, B, nullptr, payload, now, out, hit_out, s);
    DIE_TRY(die_route_md5(key_bytes, key_off, B, 1, nullptr, nullptr, 0, 0, p->shard, dg, nullptr, stream));
    return run_general(p, dg, B, payload, now, out, hit_out, s);
}

extern "C" die_status_t die_pipeline_run_host(die_pipeline_t* p, const uint8_t* key_bytes, const int64_t* key_off,
                                              int32_t B, const float* payload, double now, float* out,
                                              uint8_t* hit_out, die_stream_t stream) {
    if (!p || B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_host: bad B");
    if (B == 0) return DIE_OK;
    if (!key_off || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_host: null buffer");
    const int64_t nbytes = key_off[B] - key_off[0];
    if (key_off[0] != 0 || nbytes < 0 || nbytes > p->cfg.max_key_bytes)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_pipeline_run_host: key bytes %lld outside [0, %d] (key_off[0] must be 0)",
                 (long long)nbytes, p->cfg.max_key_bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const die_pipeline_config_t& c = p->cfg;
    if (nbytes) DIE_CUDA(cudaMemcpyAsync(p->d_key_bytes, key_bytes, (size_t)nbytes, cudaMemcpyHostToDevice, s));
    DIE_CUDA(cudaMemcpyAsync(p->d_key_off, key_off, (size_t)(B + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    DIE_CUDA(cudaMemcpyAsync(p->d_payload, payload, (size_t)B * c.dim_in * sizeof(float), cudaMemcpyHostToDevice, s));
    DIE_TRY(die_pipeline_run_device(p, p->d_key_bytes, p->d_key_off, B, p->d_payload, now, p->d_out, p->d_hit, nullptr,
                                    stream));
    DIE_CUDA(cudaMemcpyAsync(out, p->d_out, (size_t)B * c.dim_out * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (hit_out) DIE_CUDA(cudaMemcpyAsync(hit_out, p->d_hit, (size_t)B, cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    return DIE_OK;
}

// A queue of flushed batches submitted in one call: batch j reads key_bytes + j*key_bytes_stride,
// key_off + j*key_off_stride (0 = every batch has the same offsets), payload + j*B*dim_in and writes
// out + (j % out_ring)*B*dim_out, hit_out + (j % out_ring)*B.  Batches are applied strictly in order.
extern "C" die_status_t die_pipeline_run_device_many(die_pipeline_t* p, int32_t n_batches, const uint8_t* key_bytes,
                                                     int64_t key_bytes_stride, const int64_t* key_off,
                                                     int64_t key_off_stride, int32_t B, const float* payload,
                                                     double now0, double now_step, float* out, uint8_t* hit_out,
                                                     int32_t out_ring, die_stream_t stream) {
    if (!p || n_batches < 0 || out_ring < 1) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device_many: bad arguments");
    if (B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device_many: bad B");
    if (B == 0 || n_batches == 0) return DIE_OK;
    if (!key_off || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device_many: null buffer");
    const size_t in_row = (size_t)B * p->cfg.dim_in, out_row = (size_t)B * p->cfg.dim_out;
    int32_t j = 0;
    const int K = p->graph_steps;
    if (graph_fits(p, B, K) && (!prof_active() || p->persistent)) {  // one persistent launch per group can be event-timed
        QueueDesc d;
        memset(&d, 0, sizeof(d));
        d.key_bytes = key_bytes;
        d.key_off = key_off;
        d.payload = payload;
        d.out = out;
        d.hit = hit_out ? hit_out : p->code;
        d.key_bytes_stride = key_bytes_stride;
        d.key_off_stride = key_off_stride;
        d.payload_stride = (long long)in_row;
        d.out_stride = (long long)out_row;
        d.hit_stride = hit_out ? B : 0;
        d.ring = out_ring;
        d.now0 = now0;
        d.now_step = now_step;
        for (; n_batches - j >= K; j += K) {
            d.j0 = j;
            DIE_TRY(replay_fast(p, d, p->qdesc, true, p->digests, B, nullptr, K, static_cast<cudaStream_t>(stream)));
        }
    }
    for (; j < n_batches; ++j) {
        const int r = j % out_ring;
        DIE_TRY(die_pipeline_run_device(p, key_bytes + (size_t)j * key_bytes_stride, key_off + (size_t)j * key_off_stride,
                                        B, payload + (size_t)j * in_row, now0 + j * now_step, out + (size_t)r * out_row,
                                        hit_out ? hit_out + (size_t)r * B : nullptr, nullptr, stream));
    }
    return DIE_OK;
}

// The same queue through HOST buffers (pinned): copy-in of batch j+1, the kernels of batch j and the copy-out of
// batch j-1 run on three streams over NBUF staging sets.  out [n, B, dim_out] and hit_out [n, B] are all written.
// Synchronises before returning.
extern "C" die_status_t die_pipeline_run_host_many(die_pipeline_t* p, int32_t n_batches, const uint8_t* key_bytes,
                                                   int64_t key_bytes_stride, const int64_t* key_off,
                                                   int64_t key_off_stride, int32_t B, const float* payload,
                                                   double now0, double now_step, float* out, uint8_t* hit_out,
                                                   die_stream_t stream) {
    if (!p || n_batches < 0 || B <= 0 || B > p->cfg.max_batch)
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_host_many: bad arguments");
    if (!key_off || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_host_many: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const die_pipeline_config_t& c = p->cfg;
    constexpr int NB = die_pipeline::NBUF;
    const size_t in_row = (size_t)B * c.dim_in, out_row = (size_t)B * c.dim_out;
    for (int32_t j = 0; j < n_batches; ++j) {
        const int64_t* ko = key_off + (size_t)j * key_off_stride;
        const int64_t nbytes = ko[B] - ko[0];
        if (ko[0] != 0 || nbytes < 0 || nbytes > c.max_key_bytes)
            DIE_FAIL(DIE_ERR_TOO_LARGE, "die_pipeline_run_host_many: batch %d has %lld key bytes (limit %d)", j,
                     (long long)nbytes, c.max_key_bytes);
    }
    // Grouped form: G consecutive batches are contiguous in the caller's arrays, so one copy per array moves a
    // group, one graph replay computes it, and the copy engines work on the neighbouring groups meanwhile.
    // Groups of fewer batches would shorten the call's first copy-in and last copy-out, which nothing overlaps, but
    // measured the other way round on B200 (C2, 128 batches per call: groups of 2 / 4 / 8 / 16 -> 52.5 / 61.0 / 67.4 /
    // 70.4 M requests/s): per-copy overheads outweigh the shorter ramp.  DIE_HOST_GROUP overrides.
    int G = p->graph_steps;
    if (const char* e = getenv("DIE_HOST_GROUP")) G = std::max(2, std::min(G, atoi(e)));
    if (n_batches >= 2 && G >= 2 && pipeline_graph_fits(p, B, std::min<int>(G, n_batches)) &&
        (n_batches == 1 || (size_t)key_bytes_stride <= p->key_stage_stride) &&
        (key_off_stride == 0 || key_off_stride == B + 1)) {
        if (p->g_steps < G) {
            const size_t mb = (size_t)c.max_batch;
            cudaFree(p->g_keys); cudaFree(p->g_off); cudaFree(p->g_pay); cudaFree(p->g_out); cudaFree(p->g_hit);
            p->g_steps = 0;
            DIE_CUDA(cudaMalloc(&p->g_keys, (size_t)NB * G * p->key_stage_stride));
            DIE_CUDA(cudaMalloc(&p->g_off, (size_t)NB * G * (mb + 1) * sizeof(int64_t)));
            DIE_CUDA(cudaMalloc(&p->g_pay, (size_t)NB * G * mb * c.dim_in * sizeof(float)));
            DIE_CUDA(cudaMalloc(&p->g_out, (size_t)NB * G * mb * c.dim_out * sizeof(float)));
            DIE_CUDA(cudaMalloc(&p->g_hit, (size_t)NB * G * mb));
            p->g_steps = G;
        }
        const size_t mb = (size_t)c.max_batch;
        // Nothing overlaps the call's first copy-in and its last copy-out (and the kernels between them): with uniform
        // groups that is two whole groups of PCIe time per call (C2: 1.9 of 7.4 ms).  The first and the last G batches are
        // therefore split G/4 + 3G/4 and 3G/4 + G/4; everything between moves in full groups (fewest copies and replays).
        const bool ramp = G >= 8 && n_batches >= 3 * G;
        auto group_at = [&](int32_t j) {
            const int rem = n_batches - j;
            if (ramp && j == 0) return G / 4;
            if (ramp && j == G / 4) return G - G / 4;
            if (ramp && rem == G) return G - G / 4;
            return std::min<int>(G, rem);
        };
        // DIE_HOST_TRACE=1 (diagnostics): timed events around every group's copy-in, kernels and copy-out, printed
        // relative to the call's start after the call has drained
        static const bool host_trace = getenv("DIE_HOST_TRACE") && atoi(getenv("DIE_HOST_TRACE"));
        std::vector<cudaEvent_t> tev;
        auto mark = [&](cudaStream_t st) {
            if (!host_trace) return;
            cudaEvent_t e;
            cudaEventCreate(&e);
            cudaEventRecord(e, st);
            tev.push_back(e);
        };
        mark(s);
        int set = 0;
        for (int32_t j = 0, K = 0; j < n_batches; j += K, ++set) {
            K = group_at(j);
            const int b = set % NB;
            uint8_t* dk = p->g_keys + (size_t)b * G * p->key_stage_stride;
            int64_t* dko = p->g_off + (size_t)b * G * (mb + 1);
            float* dp = p->g_pay + (size_t)b * G * mb * c.dim_in;
            float* dout = p->g_out + (size_t)b * G * mb * c.dim_out;
            uint8_t* dh = p->g_hit + (size_t)b * G * mb;
            if (set >= NB) {  // set b is free once group set-NB has been computed (inputs) and copied out (outputs)
                DIE_CUDA(cudaStreamWaitEvent(p->s_in, p->ev_done[b], 0));
                DIE_CUDA(cudaStreamWaitEvent(s, p->ev_out[b], 0));
            }
            const int64_t* ko_last = key_off + (size_t)(j + K - 1) * key_off_stride;
            const size_t key_span = (size_t)(K - 1) * key_bytes_stride + (size_t)(ko_last[B] - ko_last[0]);
            mark(p->s_in);
            if (key_span)
                DIE_CUDA(cudaMemcpyAsync(dk, key_bytes + (size_t)j * key_bytes_stride, key_span, cudaMemcpyHostToDevice, p->s_in));
            DIE_CUDA(cudaMemcpyAsync(dko, key_off + (size_t)j * key_off_stride,
                                     (size_t)(key_off_stride ? K : 1) * (B + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, p->s_in));
            DIE_CUDA(cudaMemcpyAsync(dp, payload + (size_t)j * in_row, (size_t)K * in_row * sizeof(float), cudaMemcpyHostToDevice, p->s_in));
            DIE_CUDA(cudaEventRecord(p->ev_in[b], p->s_in));
            mark(p->s_in);
            DIE_CUDA(cudaStreamWaitEvent(s, p->ev_in[b], 0));
            mark(s);
            QueueDesc d;
            memset(&d, 0, sizeof(d));
            d.key_bytes = dk;
            d.key_off = dko;
            d.payload = dp;
            d.out = dout;
            d.hit = dh;
            d.key_bytes_stride = key_bytes_stride;
            d.key_off_stride = key_off_stride;
            d.payload_stride = (long long)in_row;
            d.out_stride = (long long)out_row;
            d.hit_stride = B;
            d.ring = K;
            d.j0 = 0;
            d.jt = j;
            d.now0 = now0;
            d.now_step = now_step;
            if (K >= 2) {
                DIE_TRY(replay_fast(p, d, p->qdesc, true, p->digests, B, nullptr, K, s));
            } else {
                DIE_TRY(die_pipeline_run_device(p, dk, dko, B, dp, now0 + j * now_step, dout, dh, nullptr, stream));
            }
            DIE_CUDA(cudaEventRecord(p->ev_done[b], s));
            mark(s);
            DIE_CUDA(cudaStreamWaitEvent(p->s_out, p->ev_done[b], 0));
            mark(p->s_out);
            DIE_CUDA(cudaMemcpyAsync(out + (size_t)j * out_row, dout, (size_t)K * out_row * sizeof(float), cudaMemcpyDeviceToHost, p->s_out));
            if (hit_out) DIE_CUDA(cudaMemcpyAsync(hit_out + (size_t)j * B, dh, (size_t)K * B, cudaMemcpyDeviceToHost, p->s_out));
            DIE_CUDA(cudaEventRecord(p->ev_out[b], p->s_out));
            mark(p->s_out);
        }
        DIE_CUDA(cudaStreamSynchronize(p->s_out));
        DIE_CUDA(cudaStreamSynchronize(s));
        if (host_trace) {  // per group: copy-in [start, end], kernels [start, end], copy-out [start, end] in ms
            for (size_t g = 0; 1 + 6 * g + 5 < tev.size(); ++g) {
                float t[6];
                for (int k = 0; k < 6; ++k) cudaEventElapsedTime(&t[k], tev[0], tev[1 + 6 * g + k]);
                fprintf(stderr, "group %2zu  in %.3f-%.3f  run %.3f-%.3f  out %.3f-%.3f\n", g, t[0], t[1], t[2], t[3], t[4], t[5]);
            }
            for (cudaEvent_t e : tev) cudaEventDestroy(e);
        }
        return DIE_OK;
    }
    for (int32_t j = 0; j < n_batches; ++j) {
        const int b = j % NB;
        const int64_t* ko = key_off + (size_t)j * key_off_stride;
        const int64_t nbytes = ko[B] - ko[0];
        if (ko[0] != 0 || nbytes < 0 || nbytes > c.max_key_bytes)
            DIE_FAIL(DIE_ERR_TOO_LARGE, "die_pipeline_run_host_many: batch %d has %lld key bytes (limit %d)", j,
                     (long long)nbytes, c.max_key_bytes);
        uint8_t* dk = p->d_key_bytes + (size_t)b * p->key_stage_stride;
        int64_t* dko = p->d_key_off + (size_t)b * (B + 1);
        float* dp = p->d_payload + (size_t)b * ((size_t)c.max_batch * c.dim_in);
        float* dout = p->d_out + (size_t)b * ((size_t)c.max_batch * c.dim_out);
        uint8_t* dh = p->d_hit + (size_t)b * c.max_batch;
        // staging set b is free once batch j-NB has been computed (inputs) and copied out (outputs)
        if (j >= NB) {
            DIE_CUDA(cudaStreamWaitEvent(p->s_in, p->ev_done[b], 0));
            DIE_CUDA(cudaStreamWaitEvent(s, p->ev_out[b], 0));
        }
        if (nbytes) DIE_CUDA(cudaMemcpyAsync(dk, key_bytes + (size_t)j * key_bytes_stride, (size_t)nbytes, cudaMemcpyHostToDevice, p->s_in));
        DIE_CUDA(cudaMemcpyAsync(dko, ko, (size_t)(B + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, p->s_in));
        DIE_CUDA(cudaMemcpyAsync(dp, payload + (size_t)j * in_row, in_row * sizeof(float), cudaMemcpyHostToDevice, p->s_in));
        DIE_CUDA(cudaEventRecord(p->ev_in[b], p->s_in));
        DIE_CUDA(cudaStreamWaitEvent(s, p->ev_in[b], 0));
        DIE_TRY(die_pipeline_run_device(p, dk, dko, B, dp, now0 + j * now_step, dout, dh, nullptr, stream));
        DIE_CUDA(cudaEventRecord(p->ev_done[b], s));
        DIE_CUDA(cudaStreamWaitEvent(p->s_out, p->ev_done[b], 0));
        DIE_CUDA(cudaMemcpyAsync(out + (size_t)j * out_row, dout, out_row * sizeof(float), cudaMemcpyDeviceToHost, p->s_out));
        if (hit_out) DIE_CUDA(cudaMemcpyAsync(hit_out + (size_t)j * B, dh, (size_t)B, cudaMemcpyDeviceToHost, p->s_out));
        DIE_CUDA(cudaEventRecord(p->ev_out[b], p->s_out));
    }
    DIE_CUDA(cudaStreamSynchronize(p->s_out));
    DIE_CUDA(cudaStreamSynchronize(s));
    return DIE_OK;
}
