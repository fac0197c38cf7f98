// Device-resident Batcher: per-key accumulation with size- and latency-triggered flush (src/batcher.py:102-200),
// for hosts whose requests already live on the GPU.
//
// The reference keeps one open Batch per "model:version" key.  add_request (batcher.py:102-148) opens it at the key's
// first request (created_at = now; a timer is armed for now + max_latency), appends, and flushes as soon as
// len(requests) >= max_batch_size; the timer (_schedule_flush, batcher.py:150-178) flushes whatever is there when it
// fires; stop() (batcher.py:70-100) flushes the open batches in creation order.  A flush is PACK (batcher.py:211), the
// model callback and UNPACK (batcher.py:225-227).
//
// Here time advances in TICKS: die_batcher_tick(now, arrivals) applies, in this order and without a host round trip,
//   1. every timer that has fired by `now` (now - opened_at >= max_latency), in the order the timers were armed;
//   2. the tick's arrivals in arrival order, with the size flushes they trigger (a key may flush several times);
//   3. with force != 0, stop(): every batch still open, in creation order.
// Nothing arrives between two ticks, so flushing an expired batch at the tick instead of at its deadline changes neither
// the batches' composition nor their order.  Queues are the keys, numbered by the host (at most BKT_MAX_SHARDS).
//
// An open batch never holds more than max_batch requests, so the pending state is one [max_batch] buffer of rows and
// ids per queue (two, used alternately: a tick may flush a queue's buffer and start refilling it).  A tick is four
// launches: stable bucket-by-queue of the arrivals (bucket.cu) -> plan (one CTA: which queues flush, how often, in
// which order) -> move (arrivals and flushed pending rows to their batch: PACK) -> grouped model + UNPACK
// (out[request_id]).
#include <vector>

#include "bucket_device.cuh"
#include "common.cuh"

namespace die {

constexpr int BQ_MAX = BKT_MAX_SHARDS;   // queues
constexpr int BF_MAX = 1024;             // flushes per tick the plan kernel orders in shared memory

struct BatcherState {
    int count[BQ_MAX];                     // requests in the queue's open batch (0: no open batch)
    int cur[BQ_MAX];                       // which of the queue's two pending buffers the open batch uses
    double opened_at[BQ_MAX];              // Batch.created_at (batcher.py:127)
    unsigned long long open_key[BQ_MAX];   // creation order: (tick number << 32) | arrival index of the opening request
    unsigned long long tick;               // ticks so far
    unsigned long long total_requests, total_batches, total_batched_requests;  // batcher.py:58-60
    unsigned long long batch_seq;          // batches flushed before this tick
    unsigned int error;                    // 1: a tick needed more than the flush capacity
};

// What the plan kernel leaves for the move / model kernels of the same tick.
struct BatcherPlan {
    int n_flush;
    unsigned long long batch_seq0;   // number of the tick's first flushed batch
    int c0[BQ_MAX];        // open-batch size when the arrivals start (after step 1)
    int src_buf[BQ_MAX];   // buffer that held the queue's open batch before the tick
    int dst_buf[BQ_MAX];   // buffer the tick's remainder goes to
    int first_flush[BQ_MAX];   // flush index of the batch the old pending rows belong to (-1: they stay)
    int old_count[BQ_MAX];     // pending rows before the tick
    int n_full[BQ_MAX];        // size flushes of the queue this tick
    int size_flush0[BQ_MAX];   // index into size_rank[] of the queue's first size flush
    int final_flush[BQ_MAX];   // force: flush index of the remainder (-1: none)
    int size_rank[BF_MAX];     // unsorted size flush e -> its position in the flush list
};

__global__ void __launch_bounds__(256)
batcher_plan_kernel(BatcherState* st, BatcherPlan* pl, die_flush_desc_t* desc, int flush_cap, const int64_t* __restrict__ offsets,
                    const int32_t* __restrict__ perm, int n, int n_queues, int max_batch, double max_latency, double now,
                    int force) {
    __shared__ int s_dl[BQ_MAX], s_nfull[BQ_MAX], s_sz0[BQ_MAX + 1], s_c0[BQ_MAX], s_m[BQ_MAX], s_rem[BQ_MAX];
    __shared__ unsigned long long s_key[BQ_MAX];
    __shared__ double s_open[BQ_MAX];
    __shared__ int s_fill[BF_MAX];   // arrival index that triggered size flush e
    __shared__ int s_fq[BF_MAX];     // its queue
    __shared__ int s_ndl, s_nsz, s_nfin;
    const int q = threadIdx.x;
    const unsigned long long tick = st->tick;
    if (q < n_queues) {
        const int cnt = st->count[q];
        // 1. timers (batcher.py:150-166): the batch exists, is non-empty and its max_latency has elapsed
        const int dl = cnt > 0 && (now - st->opened_at[q]) >= max_latency ? 1 : 0;
        const int m = n > 0 ? (int)(offsets[q + 1] - offsets[q]) : 0;
        const int c0 = dl ? 0 : cnt;
        s_dl[q] = dl;
        s_c0[q] = c0;
        s_m[q] = m;
        s_nfull[q] = (c0 + m) / max_batch;          // 2. len(batch.requests) >= max_batch_size (batcher.py:143)
        s_rem[q] = (c0 + m) - s_nfull[q] * max_batch;
        s_key[q] = st->open_key[q];
        s_open[q] = st->opened_at[q];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0, ndl = 0;
        for (int k = 0; k < n_queues; ++k) {
            s_sz0[k] = acc;
            acc += s_nfull[k];
            ndl += s_dl[k];
        }
        s_sz0[n_queues] = acc;
        s_ndl = ndl;
        s_nsz = acc;
    }
    __syncthreads();
    const int ndl = s_ndl, nsz = s_nsz;
    const bool overflow = ndl + nsz + n_queues > flush_cap || nsz > BF_MAX;
    if (overflow) {  // refuse loudly: the tick is dropped, die_batcher_tick's caller sees it at the next *_sync
        if (threadIdx.x == 0) {
            st->error = 1u;
            pl->n_flush = 0;
            pl->batch_seq0 = st->batch_seq;
        }
        if (q < n_queues) {
            pl->c0[q] = 0; pl->first_flush[q] = -1; pl->n_full[q] = 0; pl->final_flush[q] = -1; pl->old_count[q] = 0;
            pl->src_buf[q] = pl->dst_buf[q] = st->cur[q];
        }
        return;
    }
    // the arrival that fills batch k of queue q has rank (k + 1) * max_batch - c0 - 1 among the queue's arrivals
    if (q < n_queues)
        for (int k = 0; k < s_nfull[q]; ++k) {
            const int e = s_sz0[q] + k;
            s_fill[e] = perm[offsets[q] + (long long)(k + 1) * max_batch - s_c0[q] - 1];
            s_fq[e] = q;
        }
    __syncthreads();
    // order: fired timers by creation order, then size flushes by the arrival that triggered them
    int my_dl_rank = -1;
    if (q < n_queues && s_dl[q]) {
        int r = 0;
        for (int k = 0; k < n_queues; ++k)
            if (s_dl[k] && (s_open[k] < s_open[q] || (s_open[k] == s_open[q] && s_key[k] < s_key[q]))) ++r;
        my_dl_rank = r;
        desc[r].queue = q;
        desc[r].count = st->count[q];
        desc[r].trigger = DIE_FLUSH_LATENCY;
        desc[r].pad = 0;
        desc[r].opened_at = s_open[q];
    }
    for (int e = threadIdx.x; e < nsz; e += blockDim.x) {
        int r = 0;
        for (int f = 0; f < nsz; ++f) r += s_fill[f] < s_fill[e] ? 1 : 0;
        const int j = ndl + r;
        const int fq = s_fq[e], k = e - s_sz0[fq];
        pl->size_rank[e] = j;
        desc[j].queue = fq;
        desc[j].count = max_batch;
        desc[j].trigger = DIE_FLUSH_SIZE;
        desc[j].pad = 0;
        // the first size flush completes the batch that was already open; later ones were opened by this tick
        desc[j].opened_at = (k == 0 && s_c0[fq] > 0) ? s_open[fq] : now;
    }
    __syncthreads();
    // new open batches: opened by the arrival right after the queue's last flush (or its first arrival)
    unsigned long long new_key = 0ull;
    bool reopened = false;
    if (q < n_queues && s_rem[q] > 0 && s_m[q] > 0 && (s_c0[q] == 0 || s_nfull[q] > 0)) {
        const long long r_open = s_nfull[q] > 0 ? (long long)s_nfull[q] * max_batch - s_c0[q] : 0;
        new_key = (tick << 32) | (unsigned long long)(unsigned)perm[offsets[q] + r_open];
        reopened = true;
    }
    if (q < n_queues) {
        s_key[q] = reopened ? new_key : s_key[q];
        s_open[q] = reopened ? now : s_open[q];
    }
    __syncthreads();
    // 3. stop(): the batches still open, in creation order
    int fin_rank = -1;
    if (q < n_queues && force && s_rem[q] > 0) {
        int r = 0;
        for (int k = 0; k < n_queues; ++k)
            if (s_rem[k] > 0 && (s_open[k] < s_open[q] || (s_open[k] == s_open[q] && s_key[k] < s_key[q]))) ++r;
        fin_rank = ndl + nsz + r;
        desc[fin_rank].queue = q;
        desc[fin_rank].count = s_rem[q];
        desc[fin_rank].trigger = DIE_FLUSH_FORCED;
        desc[fin_rank].pad = 0;
        desc[fin_rank].opened_at = s_open[q];
    }
    if (threadIdx.x == 0) {
        int nfin = 0;
        if (force)
            for (int k = 0; k < n_queues; ++k) nfin += s_rem[k] > 0 ? 1 : 0;
        s_nfin = nfin;
    }
    __syncthreads();
    if (q < n_queues) {
        const int cnt = st->count[q], cur = st->cur[q];
        const bool flushed_old = s_dl[q] || (cnt > 0 && s_nfull[q] > 0) || (cnt > 0 && force && s_m[q] == 0);
        pl->c0[q] = s_c0[q];
        pl->old_count[q] = cnt;
        pl->src_buf[q] = cur;
        // the old rows leave with: the fired timer; else the queue's first size flush; else (force) the final flush
        pl->first_flush[q] = s_dl[q] ? my_dl_rank
                             : (cnt > 0 && s_nfull[q] > 0) ? pl->size_rank[s_sz0[q]]
                             : (cnt > 0 && fin_rank >= 0)  ? fin_rank : -1;
        // the remainder keeps the buffer when the old rows stay with it, else it starts in the other one
        const bool stay = cnt > 0 && !s_dl[q] && s_nfull[q] == 0;
        const int dst = stay ? cur : cur ^ 1;
        pl->dst_buf[q] = dst;
        pl->n_full[q] = s_nfull[q];
        pl->size_flush0[q] = s_sz0[q];
        pl->final_flush[q] = fin_rank;
        st->count[q] = force ? 0 : s_rem[q];
        st->cur[q] = (s_rem[q] > 0 || flushed_old) ? dst : cur;
        st->opened_at[q] = s_open[q];
        st->open_key[q] = s_key[q];
    }
    if (threadIdx.x == 0) {
        const int nf = ndl + nsz + s_nfin;
        pl->n_flush = nf;
        unsigned long long batched = 0ull;
        for (int j = 0; j < nf; ++j) batched += (unsigned long long)desc[j].count;
        st->total_requests += (unsigned long long)n;          // batcher.py:141
        st->total_batches += (unsigned long long)nf;           // batcher.py:206
        st->total_batched_requests += batched;                 // batcher.py:207
        pl->batch_seq0 = st->batch_seq;
        st->batch_seq += (unsigned long long)nf;
        st->tick = tick + 1;
    }
}


// PACK: every request of the tick goes to its row of its batch -- a flushed batch's [max_batch, dim_in] tile or the
// queue's pending buffer -- and the rows an open batch already held follow it when it flushes.  One warp per row.
// Overflowed ticks (n_flush == 0 with the error flag) move nothing.
__global__ void __launch_bounds__(256)
batcher_move_kernel(const BatcherState* __restrict__ st, const BatcherPlan* __restrict__ pl, const int32_t* __restrict__ queue,
                    const int32_t* __restrict__ request_id, const int4* __restrict__ payload, const int64_t* __restrict__ offsets,
                    const int32_t* __restrict__ pos, int n, int n_queues, int max_batch, int vin, int4* __restrict__ pend_rows,
                    int32_t* __restrict__ pend_ids, int4* __restrict__ flush_rows, int32_t* __restrict__ flush_ids) {
    if (st->error) return;
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8, items = (long long)n + (long long)n_queues * max_batch;
    for (long long it = blockIdx.x * 8ll + (threadIdx.x >> 5); it < items; it += warps) {
        const int4* src;
        int id, q, slot, flush;
        if (it < n) {
            q = queue[it];
            if (q < 0 || q >= n_queues) continue;   // not a queue of this batcher: dropped, like an unroutable request
            const int p = pl->c0[q] + (int)(pos[it] - offsets[q]);
            const int k = p / max_batch;
            slot = p - k * max_batch;
            flush = k < pl->n_full[q] ? pl->size_rank[pl->size_flush0[q] + k] : pl->final_flush[q];
            src = payload + it * vin;
            id = request_id[it];
        } else {
            q = (int)((it - n) / max_batch);
            slot = (int)((it - n) - (long long)q * max_batch);
            flush = pl->first_flush[q];
            if (slot >= pl->old_count[q] || flush < 0) continue;   // the open batch stays where it is
            const long long o = ((long long)(pl->src_buf[q] * n_queues + q)) * max_batch + slot;
            src = pend_rows + o * vin;
            id = pend_ids[o];
        }
        int4* dst;
        int32_t* dst_id;
        if (flush >= 0) {
            const long long o = (long long)flush * max_batch + slot;
            dst = flush_rows + o * vin;
            dst_id = flush_ids + o;
        } else {
            const long long o = ((long long)(pl->dst_buf[q] * n_queues + q)) * max_batch + slot;
            dst = pend_rows + o * vin;
            dst_id = pend_ids + o;
        }
        for (int c = lane; c < vin; c += 32) dst[c] = src[c];
        if (lane == 0) *dst_id = id;
    }
}

// The flushed batches through the model (batch_callback, batcher.py:212) and UNPACK (batcher.py:225-227): row r of
// batch j -> out[flush_ids[j][r]].  W == nullptr: FakeModel's echo (fake_model.py:52); else the mock fp32 linear layer
// with the queue's own weights, W[queue][dim_out][dim_in].  One CTA per 32 rows of one batch.
constexpr int BL_ROWS = 32;
__global__ void __launch_bounds__(256)
batcher_model_kernel(const BatcherPlan* __restrict__ pl, const die_flush_desc_t* __restrict__ desc,
                     const float* __restrict__ flush_rows, const int32_t* __restrict__ flush_ids, const float* __restrict__ W,
                     const float* __restrict__ bias, int max_batch, int K, int N, float* __restrict__ out,
                     int32_t* __restrict__ batch_of) {
    extern __shared__ float bl_smem[];
    const int j = blockIdx.y;
    if (j >= pl->n_flush) return;
    const int count = desc[j].count, r0 = blockIdx.x * BL_ROWS;
    if (r0 >= count) return;
    const int rows = min(BL_ROWS, count - r0);
    const int tid = threadIdx.x;
    const float* x = flush_rows + ((long long)j * max_batch + r0) * K;
    const int32_t* ids = flush_ids + (long long)j * max_batch + r0;
    if (batch_of && tid < rows) batch_of[ids[tid]] = (int32_t)(pl->batch_seq0 + (unsigned long long)j);
    if (!W) {
        for (int e = tid; e < rows * N; e += 256) {
            const int r = e / N, c = e - r * N;
            out[(long long)ids[r] * N + c] = x[(long long)r * K + c];
        }
        return;
    }
    float* xs = bl_smem;                 // [BL_ROWS][K]
    float* wt = bl_smem + BL_ROWS * K;   // [32][N + 1]: a k-tile of W, transposed
    for (int e = tid; e < BL_ROWS * K; e += 256) xs[e] = e < rows * K ? x[e] : 0.f;
    const float* Wq = W + (long long)desc[j].queue * N * K;
    const int row = tid >> 3, cg = tid & 7, na = N >> 3;
    float acc[32];
#pragma unroll
    for (int a = 0; a < 32; ++a) acc[a] = 0.f;
    for (int k0 = 0; k0 < K; k0 += 32) {
        __syncthreads();
        for (int e = tid; e < 32 * N; e += 256) {
            const int kk = e & 31, c = e >> 5;
            wt[kk * (N + 1) + c] = k0 + kk < K ? Wq[(long long)c * K + k0 + kk] : 0.f;
        }
        __syncthreads();
        const int kn = min(32, K - k0);
        for (int kk = 0; kk < kn; ++kk) {
            const float xv = xs[row * K + k0 + kk];
#pragma unroll
            for (int a = 0; a < 32; ++a)
                if (a < na) acc[a] = fmaf(xv, wt[kk * (N + 1) + cg + 8 * a], acc[a]);
        }
    }
    if (row < rows) {
        float* o = out + (long long)ids[row] * N;
        const float* bq = bias ? bias + (long long)desc[j].queue * N : nullptr;
#pragma unroll
        for (int a = 0; a < 32; ++a)
            if (a < na) o[cg + 8 * a] = acc[a] + (bq ? bq[cg + 8 * a] : 0.f);
    }
}

}  // namespace die

using namespace die;

struct die_batcher {
    int n_queues, max_batch, dim_in, dim_out, max_arrivals, flush_cap;
    double max_latency;
    BatcherState* st;
    BatcherPlan* pl;
    die_flush_desc_t* desc;
    float* W;
    float* bias;
    float* pend_rows;     // [2][Q][max_batch][dim_in]
    int32_t* pend_ids;    // [2][Q][max_batch]
    float* flush_rows;    // [flush_cap][max_batch][dim_in]
    int32_t* flush_ids;   // [flush_cap][max_batch]
    int64_t* offsets;     // bucket-by-queue scratch
    int32_t* perm;
    int32_t* pos;
    size_t model_smem;
};

extern "C" die_status_t die_batcher_create(int32_t n_queues, int32_t max_batch, double max_latency_s, int32_t dim_in,
                                           int32_t dim_out, int32_t max_arrivals, const float* weights, const float* bias,
                                           die_batcher_t** out) {
    if (!out) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_create: null argument");
    if (n_queues < 1 || n_queues > BQ_MAX) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_create: n_queues must be in [1, %d]", BQ_MAX);
    // the reference's constructor checks (batcher.py:44-47)
    if (max_batch <= 0) DIE_FAIL(DIE_ERR_INVALID, "max_batch_size must be positive");
    if (!(max_latency_s > 0)) DIE_FAIL(DIE_ERR_INVALID, "max_latency_ms must be positive");
    if (dim_in < 4 || dim_in % 4 || dim_out < 4 || dim_out % 4) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_create: dims must be multiples of 4");
    if (!weights && dim_in != dim_out) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_create: the echo model needs dim_in == dim_out");
    if (weights && (dim_out % 8 || dim_out > 256 || dim_in > 1024))
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "die_batcher_create: linear model needs dim_out %% 8 == 0, dim_out <= 256, dim_in <= 1024");
    if (max_arrivals < 1) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_create: max_arrivals must be positive");
    die_batcher* b = new die_batcher();
    memset(b, 0, sizeof(*b));
    b->n_queues = n_queues;
    b->max_batch = max_batch;
    b->dim_in = dim_in;
    b->dim_out = dim_out;
    b->max_arrivals = max_arrivals;
    b->max_latency = max_latency_s;
    b->flush_cap = max_arrivals / max_batch + 2 * n_queues + 1;
    if (max_arrivals / max_batch + n_queues > BF_MAX) {
        delete b;
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_batcher_create: max_arrivals / max_batch exceeds %d flushes per tick", BF_MAX - BQ_MAX);
    }
    const size_t Q = (size_t)n_queues, mb = (size_t)max_batch;
    DIE_CUDA(cudaMalloc(&b->st, sizeof(BatcherState)));
    DIE_CUDA(cudaMemset(b->st, 0, sizeof(BatcherState)));
    DIE_CUDA(cudaMalloc(&b->pl, sizeof(BatcherPlan)));
    DIE_CUDA(cudaMemset(b->pl, 0, sizeof(BatcherPlan)));
    DIE_CUDA(cudaMalloc(&b->desc, (size_t)b->flush_cap * sizeof(die_flush_desc_t)));
    DIE_CUDA(cudaMalloc(&b->pend_rows, 2 * Q * mb * dim_in * sizeof(float)));
    DIE_CUDA(cudaMalloc(&b->pend_ids, 2 * Q * mb * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&b->flush_rows, (size_t)b->flush_cap * mb * dim_in * sizeof(float)));
    DIE_CUDA(cudaMalloc(&b->flush_ids, (size_t)b->flush_cap * mb * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&b->offsets, (BQ_MAX + 2) * sizeof(int64_t)));
    DIE_CUDA(cudaMalloc(&b->perm, (size_t)max_arrivals * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&b->pos, (size_t)max_arrivals * sizeof(int32_t)));
    if (weights) {
        DIE_CUDA(cudaMalloc(&b->W, Q * dim_out * dim_in * sizeof(float)));
        DIE_CUDA(cudaMemcpy(b->W, weights, Q * dim_out * dim_in * sizeof(float), cudaMemcpyHostToDevice));
        if (bias) {
            DIE_CUDA(cudaMalloc(&b->bias, Q * dim_out * sizeof(float)));
            DIE_CUDA(cudaMemcpy(b->bias, bias, Q * dim_out * sizeof(float), cudaMemcpyHostToDevice));
        }
        b->model_smem = ((size_t)BL_ROWS * dim_in + 32 * ((size_t)dim_out + 1)) * sizeof(float);
        DIE_CUDA(cudaFuncSetAttribute(batcher_model_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->model_smem));
    }
    *out = b;
    return DIE_OK;
}

extern "C" die_status_t die_batcher_destroy(die_batcher_t* b) {
    if (!b) return DIE_OK;
    void* bufs[] = {b->st, b->pl, b->desc, b->W, b->bias, b->pend_rows, b->pend_ids, b->flush_rows, b->flush_ids,
                    b->offsets, b->perm, b->pos};
    for (void* p : bufs) cudaFree(p);
    delete b;
    return DIE_OK;
}

extern "C" die_status_t die_batcher_tick(die_batcher_t* b, double now, int32_t force, const int32_t* queue,
                                         const int32_t* request_id, const float* payload, int32_t n, float* out,
                                         int32_t* batch_of, die_stream_t stream) {
    if (!b || n < 0 || !out) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_tick: bad arguments");
    if (n > b->max_arrivals) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_batcher_tick: %d arrivals > max_arrivals %d", n, b->max_arrivals);
    if (n > 0 && (!queue || !request_id || !payload)) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_tick: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n > 0) DIE_TRY(die_bucket_by_shard(queue, n, b->n_queues, b->offsets, b->perm, b->pos, stream));
    DIE_LAUNCH(ST_BUCKET, s, batcher_plan_kernel<<<1, 256, 0, s>>>(b->st, b->pl, b->desc, b->flush_cap, b->offsets, b->perm, n,
                                                                   b->n_queues, b->max_batch, b->max_latency, now, force ? 1 : 0));
    const long long items = (long long)n + (long long)b->n_queues * b->max_batch;
    const int grid = (int)std::min<long long>((items + 7) / 8, (long long)num_sms() * 8);
    DIE_LAUNCH(ST_PACK, s, batcher_move_kernel<<<grid, 256, 0, s>>>(
        b->st, b->pl, queue, request_id, reinterpret_cast<const int4*>(payload), b->offsets, b->pos, n, b->n_queues,
        b->max_batch, b->dim_in / 4, reinterpret_cast<int4*>(b->pend_rows), b->pend_ids,
        reinterpret_cast<int4*>(b->flush_rows), b->flush_ids));
    // at most n / max_batch size flushes, one fired timer and one forced flush per queue
    const int max_flush = std::min(b->flush_cap, n / b->max_batch + 2 * b->n_queues);
    dim3 mg((unsigned)ceil_div(b->max_batch, BL_ROWS), (unsigned)max_flush);
    DIE_LAUNCH(ST_MODEL, s, batcher_model_kernel<<<mg, 256, b->W ? b->model_smem : 0, s>>>(
        b->pl, b->desc, b->flush_rows, b->flush_ids, b->W, b->bias, b->max_batch, b->dim_in, b->dim_out, out, batch_of));
    return DIE_OK;
}

extern "C" die_status_t die_batcher_flushes_sync(die_batcher_t* b, int32_t* n_flush, die_flush_desc_t* desc, int32_t* ids,
                                                 int32_t cap, uint64_t* first_batch, die_stream_t stream) {
    if (!b || !n_flush) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_flushes_sync: null argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    BatcherPlan head;
    unsigned int err = 0;
    DIE_CUDA(cudaMemcpyAsync(&head, b->pl, offsetof(BatcherPlan, c0), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaMemcpyAsync(&err, &b->st->error, sizeof(err), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    if (err) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_batcher: a tick needed more flushes than the batcher was sized for; it was dropped");
    *n_flush = head.n_flush;
    if (first_batch) *first_batch = head.batch_seq0;
    const int m = std::min<int>(head.n_flush, cap);
    if (desc && m > 0) DIE_CUDA(cudaMemcpyAsync(desc, b->desc, (size_t)m * sizeof(die_flush_desc_t), cudaMemcpyDeviceToHost, s));
    if (ids && m > 0)
        DIE_CUDA(cudaMemcpyAsync(ids, b->flush_ids, (size_t)m * b->max_batch * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    return DIE_OK;
}

extern "C" die_status_t die_batcher_stats_sync(die_batcher_t* b, uint64_t* total_requests, uint64_t* total_batches,
                                               uint64_t* total_batched_requests, int32_t* pending, die_stream_t stream) {
    if (!b) DIE_FAIL(DIE_ERR_INVALID, "die_batcher_stats_sync: null argument");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    BatcherState h;
    DIE_CUDA(cudaMemcpyAsync(&h, b->st, sizeof(h), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    if (h.error) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_batcher: a tick needed more flushes than the batcher was sized for; it was dropped");
    if (total_requests) *total_requests = h.total_requests;
    if (total_batches) *total_batches = h.total_batches;
    if (total_batched_requests) *total_batched_requests = h.total_batched_requests;
    if (pending)
        for (int q = 0; q < b->n_queues; ++q) pending[q] = h.count[q];
    return DIE_OK;
}
