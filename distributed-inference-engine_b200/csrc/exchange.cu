// Dispatch / return of requests between the GPUs of one NVSwitch box by direct peer stores.
//
// The reference moves every request to its worker over TCP + JSON, one connection each
// (src/worker.py:84-133).  Here rank r writes the rows it routed to shard d straight into d's receive
// window over NVLink: one kernel packs (gather by bucket order) AND sends, no staging copy, no collective,
// no host round trip.  Peer addresses come from a symmetric allocation the host exchanged once
// (torch.distributed._symmetric_memory); a device-side barrier of that allocation separates the phases.
//
// Receive window of a rank:  rows  [world][cap][4 + dim] fp32   region s is written only by source rank s:
//                                   word 0-3 = MD5 digest of the key, then the payload row
//                            count [world] int32                 how many rows source s wrote this step
// Return window of a rank:   back  [world][cap][dim_out + 4]    region d is written only by owner d:
//                                   result row, then one 16-byte tail whose first word is the hit flag
// cap = the ingest batch size, so no bucket can overflow whatever the key distribution or health state.
// Owners see the sources in rank order, request order kept inside each source: the order the CPU oracle's
// process_exchange applies.
#include "common.cuh"

struct die_pipeline;

namespace die {

constexpr int XCHG_MAX_WORLD = 16;
struct PeerPtrs {
    void* p[XCHG_MAX_WORLD];
};

__device__ __forceinline__ int bucket_of(const long long* off, int world, long long j) {
    int d = 0;
#pragma unroll 1
    while (d < world && j >= off[d + 1]) ++d;
    return d;  // == world for unroutable requests
}

// PACK + SEND: send position j (bucket order) = request perm[j]; row -> peer window of its owner.
__global__ void __launch_bounds__(256)
xchg_dispatch_kernel(const int4* __restrict__ digests, const int4* __restrict__ payload, const int32_t* __restrict__ perm,
                     const int64_t* __restrict__ offsets, int B, int vdim /*payload vectors per row*/, int world, int rank,
                     int cap, PeerPtrs rows, PeerPtrs counts) {
    __shared__ long long s_off[XCHG_MAX_WORLD + 2];
    if (threadIdx.x < world + 2) s_off[threadIdx.x] = offsets[threadIdx.x];
    __syncthreads();
    const int rv = vdim + 1;  // vectors per window row
    const long long n_send = s_off[world];
    const long long total = n_send * rv;
    if (blockIdx.x == 0 && threadIdx.x < world)
        static_cast<int32_t*>(counts.p[threadIdx.x])[rank] = (int32_t)(s_off[threadIdx.x + 1] - s_off[threadIdx.x]);
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long j = v / rv;
        const int c = (int)(v - j * rv);
        const int d = bucket_of(s_off, world, j);
        const long long i = j - s_off[d];
        const int q = perm[j];
        const int4 val = c == 0 ? digests[q] : ldg_stream(payload + (long long)q * vdim + (c - 1));
        int4* dst = static_cast<int4*>(rows.p[d]) + ((long long)rank * cap + i) * rv + c;
        *dst = val;  // NVLink store into the owner's HBM
    }
    __threadfence_system();
}

// Owner: regions -> contiguous digests / payload in (source rank, request) order + the device-side count.
__global__ void __launch_bounds__(256)
xchg_compact_kernel(const int4* __restrict__ rows, const int32_t* __restrict__ counts, int world, int cap, int vdim,
                    long long bmax, int4* __restrict__ digests_out, int4* __restrict__ payload_out,
                    uint32_t* __restrict__ n_out, int32_t* __restrict__ prefix_out /*[world+1]*/,
                    unsigned int* __restrict__ error_flag) {
    __shared__ long long s_pre[XCHG_MAX_WORLD + 1];
    if (threadIdx.x == 0) {
        long long acc = 0;
        for (int s = 0; s < world; ++s) {
            s_pre[s] = acc;
            acc += counts[s];
            if (acc > bmax) {  // more than the owner pipeline was sized for (hot-key skew): refuse loudly
                if (error_flag) *error_flag = 2u;
                acc = bmax;
            }
        }
        s_pre[world] = acc;
        if (blockIdx.x == 0) {
            *n_out = (uint32_t)acc;
            for (int s = 0; s <= world; ++s) prefix_out[s] = (int32_t)s_pre[s];
        }
    }
    __syncthreads();
    const int rv = vdim + 1;
    const long long total = s_pre[world] * rv;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long g = v / rv;
        const int c = (int)(v - g * rv);
        int s = 0;
        while (g >= s_pre[s + 1]) ++s;
        const long long i = g - s_pre[s];
        const int4 val = __ldcg(rows + ((long long)s * cap + i) * rv + c);  // written by a peer this step: read at L2
        if (c == 0)
            digests_out[g] = val;
        else
            stg_stream(payload_out + g * vdim + (c - 1), val);
    }
}

// Owner: result g (source s, position i) -> return window of s.
__global__ void __launch_bounds__(256)
xchg_return_kernel(const int4* __restrict__ out, const uint8_t* __restrict__ hit, const int32_t* __restrict__ prefix,
                   int world, int rank, int cap, int vout, PeerPtrs back) {
    __shared__ int s_pre[XCHG_MAX_WORLD + 1];
    if (threadIdx.x <= world) s_pre[threadIdx.x] = prefix[threadIdx.x];
    __syncthreads();
    const int rv = vout + 1;
    const long long total = (long long)s_pre[world] * rv;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long g = v / rv;
        const int c = (int)(v - g * rv);
        int s = 0;
        while (g >= s_pre[s + 1]) ++s;
        const long long i = g - s_pre[s];
        int4 val;
        if (c < vout)
            val = ldg_stream(out + g * vout + c);
        else
            val = make_int4((int)hit[g], 0, 0, 0);
        *(static_cast<int4*>(back.p[s]) + ((long long)rank * cap + i) * rv + c) = val;
    }
    __threadfence_system();
}

// Source: return window -> results in request order (UNPACK of the exchange).
__global__ void __launch_bounds__(256)
xchg_scatter_kernel(const int4* __restrict__ back, const int32_t* __restrict__ perm, const int64_t* __restrict__ offsets,
                    int B, int world, int cap, int vout, int4* __restrict__ out, uint8_t* __restrict__ hit) {
    __shared__ long long s_off[XCHG_MAX_WORLD + 2];
    if (threadIdx.x < world + 2) s_off[threadIdx.x] = offsets[threadIdx.x];
    __syncthreads();
    const int rv = vout + 1;
    const long long total = (long long)B * rv;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long j = v / rv;
        const int c = (int)(v - j * rv);
        const int d = bucket_of(s_off, world, j);
        const int q = perm[j];
        int4 val = make_int4(0, 0, 0, 0);  // requests the router returned None for have no result
        if (d < world) val = __ldcg(back + ((long long)d * cap + (j - s_off[d])) * rv + c);  // peer-written: read at L2
        if (c < vout)
            stg_stream(out + (long long)q * vout + c, val);
        else
            hit[q] = (uint8_t)val.x;
    }
}

static int grid_for_vectors(long long items) {
    const long long blocks = (items + 255) / 256;
    const long long cap = (long long)num_sms() * 8;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

}  // namespace die

using namespace die;

static die_status_t fill_peers(PeerPtrs* dst, void* const* src, int world, const char* what) {
    if (world < 1 || world > XCHG_MAX_WORLD) DIE_FAIL(DIE_ERR_INVALID, "exchange: world must be in [1, %d]", XCHG_MAX_WORLD);
    for (int i = 0; i < XCHG_MAX_WORLD; ++i) dst->p[i] = i < world ? src[i] : nullptr;
    for (int i = 0; i < world; ++i)
        if (!dst->p[i]) DIE_FAIL(DIE_ERR_INVALID, "exchange: null peer pointer for %s[%d]", what, i);
    return DIE_OK;
}

extern "C" die_status_t die_xchg_dispatch(const uint32_t* digests, const float* payload, const int32_t* perm,
                                          const int64_t* offsets, int32_t B, int32_t dim, int32_t world, int32_t rank,
                                          int32_t cap, void* const* peer_rows, void* const* peer_counts,
                                          die_stream_t stream) {
    if (B < 0 || dim <= 0 || (dim & 3) || cap < B) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_dispatch: bad sizes (cap must be >= B)");
    PeerPtrs rows, counts;
    DIE_TRY(fill_peers(&rows, peer_rows, world, "rows"));
    DIE_TRY(fill_peers(&counts, peer_counts, world, "counts"));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int vdim = dim / 4;
    DIE_LAUNCH(ST_XCHG, s, xchg_dispatch_kernel<<<grid_for_vectors((long long)B * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(digests), reinterpret_cast<const int4*>(payload), perm, offsets, B, vdim, world, rank, cap,
        rows, counts));
    return DIE_OK;
}

extern "C" die_status_t die_xchg_compact(const float* rows, const int32_t* counts, int32_t world, int32_t cap, int32_t dim,
                                         uint32_t* digests_out, float* payload_out, uint32_t* n_out, int32_t* prefix_out,
                                         die_stream_t stream) {
    const long long bmax = (long long)world * cap;
    if (world < 1 || world > XCHG_MAX_WORLD || dim <= 0 || (dim & 3) || cap <= 0)
        DIE_FAIL(DIE_ERR_INVALID, "die_xchg_compact: bad sizes");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int vdim = dim / 4;
    DIE_LAUNCH(ST_XCHG, s, xchg_compact_kernel<<<grid_for_vectors((long long)2 * cap * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(rows), counts, world, cap, vdim, bmax, reinterpret_cast<int4*>(digests_out),
        reinterpret_cast<int4*>(payload_out), n_out, prefix_out, nullptr));
    return DIE_OK;
}

extern "C" die_status_t die_xchg_return(const float* out, const uint8_t* hit, const int32_t* prefix, int32_t world,
                                        int32_t rank, int32_t cap, int32_t dim_out, void* const* peer_back,
                                        die_stream_t stream) {
    if (dim_out <= 0 || (dim_out & 3) || cap <= 0) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_return: bad sizes");
    PeerPtrs back;
    DIE_TRY(fill_peers(&back, peer_back, world, "back"));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int vout = dim_out / 4;
    DIE_LAUNCH(ST_XCHG, s, xchg_return_kernel<<<grid_for_vectors((long long)2 * cap * (vout + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(out), hit, prefix, world, rank, cap, vout, back));
    return DIE_OK;
}

extern "C" die_status_t die_xchg_scatter(const float* back, const int32_t* perm, const int64_t* offsets, int32_t B,
                                         int32_t world, int32_t cap, int32_t dim_out, float* out, uint8_t* hit,
                                         die_stream_t stream) {
    if (B < 0 || world < 1 || world > XCHG_MAX_WORLD || dim_out <= 0 || (dim_out & 3))
        DIE_FAIL(DIE_ERR_INVALID, "die_xchg_scatter: bad sizes");
    if (B == 0) return DIE_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int vout = dim_out / 4;
    DIE_LAUNCH(ST_XCHG, s, xchg_scatter_kernel<<<grid_for_vectors((long long)B * (vout + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(back), perm, offsets, B, world, cap, vout, reinterpret_cast<int4*>(out), hit));
    return DIE_OK;
}

// ================================================================================================
// One call per flushed batch: the whole sharded step enqueued from C (route, bucket, dispatch, barrier,
// compact, owner pipeline, return, barrier, scatter) so that the host cost is a handful of launches.
// ================================================================================================
namespace die {

// Cross-GPU barrier on flags inside the symmetric window: rank r stores `epoch` into flag[r] of every peer
// and waits until all of its own flags reach `epoch`.  One rank per GPU (never launch two ranks' barriers on
// the same device: they would wait on each other inside one GPU).  Gives up after ~4 s and reports.
__global__ void xchg_barrier_kernel(PeerPtrs flags, unsigned int* my_flags, int world, int rank, unsigned int epoch,
                                    unsigned int* error_flag) {
    const int p = threadIdx.x;
    if (p >= world) return;
    __threadfence_system();
    unsigned int* dst = static_cast<unsigned int*>(flags.p[p]) + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(epoch) : "memory");
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        unsigned int v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(my_flags + p) : "memory");
        if ((int)(v - epoch) >= 0) break;
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > 4000000000ull) {
            *error_flag = 1u;
            break;
        }
    }
}

}  // namespace die

struct die_xchg {
    int world, rank, cap, dim;
    int bmax;  // most requests the owner side handles per step (the pipeline's max_batch)
    PeerPtrs rows, back, counts, flags;
    float* my_rows;
    float* my_back;
    int32_t* my_cnt;
    unsigned int* my_flags;
    die_pipeline_t* pipe;
    int32_t* shard;
    uint32_t* digests;
    int64_t* offsets;
    int32_t* perm;
    int32_t* pos;
    uint32_t* recv_dig;
    float* recv_pay;
    float* recv_out;
    uint8_t* recv_hit;
    uint32_t* n_dev;
    int32_t* prefix;
    unsigned int* error_flag;
    unsigned int epoch;
};

static size_t xchg_rows_bytes(int world, int cap, int dim) { return (size_t)world * cap * (dim + 4) * sizeof(float); }

extern "C" uint64_t die_xchg_window_bytes(int32_t world, int32_t cap, int32_t dim) {
    // receive rows + return rows + counts + barrier flags
    return 2 * xchg_rows_bytes(world, cap, dim) + 256 + 256;
}

extern "C" die_status_t die_xchg_create(int32_t world, int32_t rank, int32_t cap, int32_t dim, int32_t owner_max,
                                        die_pipeline_t* pipe, void* const* peer_windows, die_xchg_t** out) {
    if (!out || !pipe || !peer_windows || world < 1 || world > XCHG_MAX_WORLD || rank < 0 || rank >= world || cap <= 0 ||
        dim <= 0 || (dim & 3))
        DIE_FAIL(DIE_ERR_INVALID, "die_xchg_create: bad arguments");
    die_xchg* x = new die_xchg();
    memset(x, 0, sizeof(*x));
    x->world = world;
    x->rank = rank;
    x->cap = cap;
    x->dim = dim;
    x->pipe = pipe;
    x->bmax = owner_max > 0 && owner_max < world * cap ? owner_max : world * cap;
    const size_t rb = xchg_rows_bytes(world, cap, dim);
    for (int i = 0; i < XCHG_MAX_WORLD; ++i) {
        char* base = i < world ? static_cast<char*>(peer_windows[i]) : nullptr;
        if (i < world && !base) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_create: null window for rank %d", i);
        x->rows.p[i] = base;
        x->back.p[i] = base ? base + rb : nullptr;
        x->counts.p[i] = base ? base + 2 * rb : nullptr;
        x->flags.p[i] = base ? base + 2 * rb + 256 : nullptr;
    }
    x->my_rows = static_cast<float*>(x->rows.p[rank]);
    x->my_back = static_cast<float*>(x->back.p[rank]);
    x->my_cnt = static_cast<int32_t*>(x->counts.p[rank]);
    x->my_flags = static_cast<unsigned int*>(x->flags.p[rank]);
    const size_t bmax = (size_t)x->bmax;
    DIE_CUDA(cudaMalloc(&x->shard, cap * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&x->digests, (size_t)cap * 16));
    DIE_CUDA(cudaMalloc(&x->offsets, (world + 2) * sizeof(int64_t)));
    DIE_CUDA(cudaMalloc(&x->perm, cap * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&x->pos, cap * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&x->recv_dig, bmax * 16));
    DIE_CUDA(cudaMalloc(&x->recv_pay, bmax * dim * sizeof(float)));
    DIE_CUDA(cudaMalloc(&x->recv_out, bmax * dim * sizeof(float)));
    DIE_CUDA(cudaMalloc(&x->recv_hit, bmax));
    DIE_CUDA(cudaMalloc(&x->n_dev, sizeof(uint32_t)));
    DIE_CUDA(cudaMalloc(&x->prefix, (world + 1) * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&x->error_flag, sizeof(unsigned int)));
    DIE_CUDA(cudaMemset(x->error_flag, 0, sizeof(unsigned int)));
    *out = x;
    return DIE_OK;
}

extern "C" die_status_t die_xchg_destroy(die_xchg_t* x) {
    if (!x) return DIE_OK;
    void* bufs[] = {x->shard, x->digests, x->offsets, x->perm, x->pos, x->recv_dig, x->recv_pay, x->recv_out,
                    x->recv_hit, x->n_dev, x->prefix, x->error_flag};
    for (void* b : bufs) cudaFree(b);
    delete x;
    return DIE_OK;
}

static die_status_t xchg_barrier(die_xchg* x, cudaStream_t s) {
    x->epoch += 1;
    DIE_LAUNCH(ST_XCHG, s, xchg_barrier_kernel<<<1, 32, 0, s>>>(x->flags, x->my_flags, x->world, x->rank, x->epoch, x->error_flag));
    return DIE_OK;
}

// One flushed batch ingested on this rank; results in request order.  Every rank of the group must make the
// same sequence of calls (the barriers inside are collective).
extern "C" die_status_t die_xchg_step(die_xchg_t* x, const uint8_t* key_bytes, const int64_t* key_off, int32_t B,
                                      const float* payload, double now, float* out, uint8_t* hit_out,
                                      int32_t* shard_out, die_stream_t stream) {
    if (!x || B < 0 || B > x->cap) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step: B must be in [0, cap]");
    if (!key_off || !payload || !out || !hit_out) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int W = x->world, D = x->dim, vdim = D / 4;
    int32_t* shard = shard_out ? shard_out : x->shard;
    DIE_TRY(die_route_md5(key_bytes, key_off, B, W, nullptr, nullptr, 0, 0, shard, x->digests, nullptr, stream));
    DIE_TRY(die_bucket_by_shard(shard, B, W, x->offsets, x->perm, x->pos, stream));
    DIE_LAUNCH(ST_XCHG, s, xchg_dispatch_kernel<<<grid_for_vectors((long long)B * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(x->digests), reinterpret_cast<const int4*>(payload), x->perm, x->offsets, B, vdim, W,
        x->rank, x->cap, x->rows, x->counts));
    DIE_TRY(xchg_barrier(x, s));  // every rank's rows have landed in their owners' windows
    DIE_LAUNCH(ST_XCHG, s, xchg_compact_kernel<<<grid_for_vectors((long long)2 * x->cap * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(x->my_rows), x->my_cnt, W, x->cap, vdim, (long long)x->bmax,
        reinterpret_cast<int4*>(x->recv_dig), reinterpret_cast<int4*>(x->recv_pay), x->n_dev, x->prefix, x->error_flag));
    DIE_TRY(die_pipeline_run_owner_device_n(x->pipe, x->recv_dig, x->bmax, x->n_dev, x->recv_pay, now, x->recv_out,
                                            x->recv_hit, stream));
    DIE_LAUNCH(ST_XCHG, s, xchg_return_kernel<<<grid_for_vectors((long long)2 * x->cap * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(x->recv_out), x->recv_hit, x->prefix, W, x->rank, x->cap, vdim, x->back));
    DIE_TRY(xchg_barrier(x, s));  // every owner's results are back in the sources' windows
    DIE_LAUNCH(ST_XCHG, s, xchg_scatter_kernel<<<grid_for_vectors((long long)B * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(x->my_back), x->perm, x->offsets, B, W, x->cap, vdim, reinterpret_cast<int4*>(out),
        hit_out));
    return DIE_OK;
}

extern "C" die_status_t die_xchg_step_many(die_xchg_t* x, int32_t n_batches, const uint8_t* key_bytes,
                                           int64_t key_bytes_stride, const int64_t* key_off, int64_t key_off_stride,
                                           int32_t B, const float* payload, double now0, double now_step, float* out,
                                           uint8_t* hit_out, int32_t out_ring, die_stream_t stream) {
    if (!x || n_batches < 0 || out_ring < 1) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step_many: bad arguments");
    const size_t row = (size_t)B * x->dim;
    for (int32_t j = 0; j < n_batches; ++j) {
        const int r = j % out_ring;
        DIE_TRY(die_xchg_step(x, key_bytes + (size_t)j * key_bytes_stride, key_off + (size_t)j * key_off_stride, B,
                              payload + (size_t)j * row, now0 + j * now_step, out + (size_t)r * row,
                              hit_out + (size_t)r * B, nullptr, stream));
    }
    return DIE_OK;
}

// Synchronises and reports whether a barrier of this context ever timed out (a peer stopped calling).
extern "C" die_status_t die_xchg_check_sync(die_xchg_t* x, die_stream_t stream) {
    if (!x) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_check_sync: null context");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned int e = 0;
    DIE_CUDA(cudaMemcpyAsync(&e, x->error_flag, sizeof(e), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    if (e == 2) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_xchg: an owner received more requests in one step than owner_max");
    if (e) DIE_FAIL(DIE_ERR_CUDA, "die_xchg: a cross-GPU barrier timed out (a peer rank stopped making calls)");
    return DIE_OK;
}
