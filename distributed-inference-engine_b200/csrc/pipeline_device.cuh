// Device-side pieces of the fused pipeline that more than one translation unit needs (pipeline.cu and the
// tensor-core GEMM whose extra CTAs evict): per-batch control block, scratch pointers, eviction stage.
#pragma once
#include "kv_internal.cuh"

struct PipeCtrl {
    unsigned int n_miss, ticket_get, ticket_claim, any_dup;
    unsigned long long e_ub, e_final;
    unsigned long long whead;  // LRU-log position the victim window of this batch starts at
    unsigned long long ts[8];  // %globaltimer marks of the last get kernel (ns): diagnostics only
};

namespace die {

// ---- what changes from one flushed batch to the next ---------------------------------------------------
// A step's kernels take a StepRef.  Launched from a stream, the values travel in the launch itself (`a`).
// Replayed from a CUDA graph, the launch arguments are frozen, so node i of the graph derives its values
// from a device-resident QueueDesc the host rewrites before every replay (`q`, `i`): the batches of a queue
// sit at regular strides, the injected clock advances by a fixed step and every batch consumes the same
// number of LRU stamps, so one descriptor describes all K steps of a replay.
struct StepArgs {
    KvView kv;
    const uint8_t* key_bytes;
    const int64_t* key_off;
    const float* payload;
    float* out;
    uint8_t* code;  // hit / miss per request
    double now;
    unsigned long long t_get, t_set;  // first LRU stamp of the batch's gets / of its sets
};

constexpr int XCHG_SETS = 48;  // buffer sets of the sharded step (exchange.cu): three groups of 16 steps in flight

struct QueueDesc {
    KvView kv;
    const uint8_t* key_bytes;
    const int64_t* key_off;
    const float* payload;
    float* out;
    uint8_t* hit;
    long long key_bytes_stride, key_off_stride, payload_stride, out_stride, hit_stride;  // elements per batch
    long long j0;  // queue index of the replay's first batch (for the buffers)
    long long jt;  // added to j0 for the clock: the batches of a staged group are a window of a longer queue
    int ring;      // out / hit slots, reused round-robin
    int pay_ring;  // payload slots, reused round-robin (0: one per batch of the queue)
    int batch;     // stamps per get (= per set) pass of one step
    double now0, now_step;
    unsigned long long tail0;  // LRU stamp of the replay's first op
    unsigned int epoch0[XCHG_SETS];  // exchange only: barrier epoch of each buffer set before the replay
    unsigned long long* trace; // diagnostics (die_debug_trace_enable): [0] cursor, [1] capacity, then (tag, ns) pairs
};

struct StepRef {
    const QueueDesc* q;  // nullptr: use `a`
    int i;               // step index inside the replay
    StepArgs a;
};

#ifdef __CUDACC__
// Diagnostics: thread 0 of CTA 0 of a replayed kernel logs (step << 8 | id, %globaltimer).
enum TraceId { TR_ROUTE = 1, TR_DISPATCH, TR_BAR_IN, TR_BAR_OUT, TR_GET, TR_GET_TAIL, TR_CLAIM, TR_MODEL, TR_UNPACK, TR_SCATTER, TR_STEP_END };
__device__ __forceinline__ void trace_mark_q(const QueueDesc* q, int i, int id) {
    if (!q || !q->trace) return;
    unsigned long long* t = q->trace;
    const unsigned long long k = atomicAdd(t, 1ull);
    if (k < t[1]) {
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        t[2 + 2 * k] = ((unsigned long long)i << 8) | (unsigned long long)id;
        t[3 + 2 * k] = ns;
    }
}
__device__ __forceinline__ void trace_mark(const StepRef& r, int id) {
    if (r.q && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) trace_mark_q(r.q, r.i, id);
}

__device__ __forceinline__ StepArgs step_resolve(const StepRef& r) {
    if (!r.q) return r.a;
    const QueueDesc& q = *r.q;
    const long long j = q.j0 + r.i;
    const long long slot = j % q.ring;
    StepArgs a;
    a.kv = q.kv;
    // buffers the descriptor leaves null are fixed ones, frozen in the launch (the exchange's lane buffers)
    a.key_bytes = q.key_bytes ? q.key_bytes + j * q.key_bytes_stride : r.a.key_bytes;
    a.key_off = q.key_off ? q.key_off + j * q.key_off_stride : r.a.key_off;
    a.payload = q.payload ? q.payload + (q.pay_ring ? j % q.pay_ring : j) * q.payload_stride : r.a.payload;
    a.out = q.out ? q.out + slot * q.out_stride : r.a.out;
    a.code = q.hit ? q.hit + slot * q.hit_stride : r.a.code;
    a.now = __dadd_rn(q.now0, __dmul_rn((double)(j + q.jt), q.now_step));  // the host's now0 + j * now_step, no fma
    a.t_get = q.tail0 + 2ull * (unsigned long long)q.batch * (unsigned long long)r.i;
    a.t_set = a.t_get + (unsigned long long)q.batch;
    return a;
}
#endif

constexpr int PT = 256;  // threads per CTA of the pipeline kernels (== ROUTE_THREADS)
constexpr int WCHUNK = 1024;  // log records per victim-window chunk

struct PipeBufs {
    uint8_t* code;
    unsigned long long* value;
    uint8_t* lrank;     // [B] rank of a miss among the misses of its get CTA (PT requests per CTA)
    int32_t* cta_miss;  // [ceil(B / PT)] misses per get CTA
    int32_t* cta_base;  // [ceil(B / PT)] exclusive prefix of cta_miss: rank of a miss = cta_base[q / PT] + lrank[q]
    int32_t* miss_req;  // rank -> request (written by the pack CTAs, read by the GEMM epilogue)
    int32_t* set_slot;
    int32_t* set_row;
    uint32_t* wmask;   // [n_chunks][32]: which of a chunk's 1024 log records are current
    uint32_t* wcount;  // [n_chunks]
    int n_chunks;
    PipeCtrl* ctrl;
};

// Victim window, stage 2 (CTAs of the last kernel of the step): record r of the window is evicted iff it is
// current and fewer than e_final current records precede it.  256 records per CTA.
static __device__ __noinline__ void pipe_evict_cta(const KvView& kv, const PipeBufs& pb, int cta, unsigned long long t_set) {
    __shared__ unsigned s_red[PT / 32];
    __shared__ unsigned s_before;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long e = pb.ctrl->e_final;
    if (e == 0) return;
    const int rec0 = cta * PT;               // first window record of this CTA
    const int chunk = rec0 / WCHUNK;
    if (chunk >= pb.n_chunks) return;
    // current records in all chunks before this one
    unsigned part = 0;
    for (int c = tid; c < chunk; c += PT) part += pb.wcount[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_red[warp] = part;
    __syncthreads();
    if (tid == 0) {
        unsigned t = 0;
        for (int w = 0; w < PT / 32; ++w) t += s_red[w];
        // plus the words of this chunk that precede this CTA's 8 words
        const int w0 = (rec0 % WCHUNK) / 32;
        for (int w = 0; w < w0; ++w) t += __popc(pb.wmask[chunk * 32 + w]);
        s_before = t;
    }
    __syncthreads();
    const int w0 = (rec0 % WCHUNK) / 32;
    unsigned before = s_before;
    for (int w = 0; w < warp; ++w) before += __popc(pb.wmask[chunk * 32 + w0 + w]);
    const unsigned word = pb.wmask[chunk * 32 + w0 + warp];
    const unsigned long long rank = before + __popc(word & ((1u << lane) - 1u));
    const bool mine = (word >> lane) & 1u;
    const unsigned long long pos = pb.ctrl->whead + (unsigned long long)rec0 + tid;
    if (mine && rank < e) {
        Slot* sl = &kv.slots[kv.log[pos & kv.rmask]];
        release_row(kv, sl->value);
        st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
        sl->stamp = 0ull;
        if (rank == e - 1) {  // the newest victim: everything up to it is gone, nobody else touches the counters
            KvState* st = kv.st;
            st->size -= e;
            st->evictions += e;
            st->n_tomb += e;
            st->head = pos + 1;
        }
    }
    // window too short (fewer than e current records in 4 records per request): finish sequentially
    if (cta == 0 && tid == 0) {
        unsigned long long total = 0;
        for (int c = 0; c < pb.n_chunks; ++c) total += pb.wcount[c];
        if (total < e) {
            unsigned long long p = pb.ctrl->whead + (unsigned long long)pb.n_chunks * WCHUNK, left = e - total;
            while (left > 0 && p < t_set) {
                const uint32_t sidx = kv.log[p & kv.rmask];
                if ((unsigned long long)sidx <= kv.mask && ld_u64_cg(&kv.slots[sidx].stamp) == p) {
                    Slot* sl = &kv.slots[sidx];
                    release_row(kv, sl->value);
                    st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
                    sl->stamp = 0ull;
                    --left;
                }
                ++p;
            }
            KvState* st = kv.st;
            if (left > 0) st->status = DIE_ERR_INVALID;  // unreachable while max_batch <= max_size
            st->size -= e;
            st->evictions += e;
            st->n_tomb += e;
            st->head = p;
        }
    }
}

}  // namespace die
