// Dispatch / return of requests between the GPUs of one NVSwitch box by direct peer stores.
//
// The reference moves every request to its worker over TCP + JSON, one connection each
// (src/worker.py:84-133).  Here rank r writes the rows it routed to shard d straight into d's receive
// window over NVLink: one kernel packs AND sends (in request order; the slot inside the owner's region is
// computed in the route kernel, there is no bucket pass and no permutation), no staging copy, no collective,
// no host round trip.  Peer addresses come from a symmetric allocation the host exchanged once
// (torch.distributed._symmetric_memory); a device-side barrier of that allocation separates the phases.
//
// Receive window of a rank:  rows  [world][cap][4 + dim] fp32   region s is written only by source rank s:
//                                   word 0-3 = MD5 digest of the key, then the payload row
//                            count [world] int32                 how many rows source s wrote this step
// Return window of a rank:   back  [world][cap][dim_out + 4]    region d is written only by owner d:
//                                   result row, then one 16-byte tail whose first word is the hit flag
// cap = the ingest batch size, so no (source, owner) region can overflow whatever the key distribution.
// Every rank of the group must be up (a rank that is not cannot run its barriers either); which owners the ROUTER
// considers usable is a separate matter: die_xchg_set_health hands the route kernel the health gate and failover list
// of router.py:165-221 (same arithmetic as die_route_md5), requests without an owner come back as zeros with shard -1.
// An owner is sized for owner_max requests per step and a step that exceeds it -- failover can do that -- raises the
// error die_xchg_check_sync reports.
// Owners see the sources in rank order, request order kept inside each source: the order the CPU oracle's
// process_exchange applies.
//
// die_xchg_step / die_xchg_step_many run the step as a three-stage pipeline (ingest | owner pass | way back) over
// XCHG_SETS buffer sets.  A single step overlaps its neighbours through per-stage streams; a queue is replayed in
// groups of 16 steps from captured graphs, software-pipelined across launches so that the owner stage is ONE persistent
// launch per group that never waits for a transfer (xchg_graph_get).  The stand-alone phase entry points
// (die_xchg_dispatch / compact / return / scatter, bucket order from die_bucket_by_shard) stay as a cross-check
// driven from Python.
#include <vector>

#include "pipeline_internal.cuh"
#include "route_device.cuh"

namespace die {

constexpr int XCHG_MAX_WORLD = 16;
// XCHG_SETS (pipeline_device.cuh) buffer sets: steps in flight between ingest and scatter
constexpr int XCHG_GROUP = 16;  // steps per replayed group = owner passes that share one persistent launch
static_assert(XCHG_SETS == 3 * XCHG_GROUP, "three groups in flight: being ingested, with the owner, on the way back");
constexpr int XCHG_MAX_WAYS = 8;  // ingest / return streams (DIE_XCHG_WAYS)
struct PeerPtrs {
    void* p[XCHG_MAX_WORLD];
};

__device__ __forceinline__ int bucket_of(const long long* off, int world, long long j) {
    int d = 0;
#pragma unroll 1
    while (d < world && j >= off[d + 1]) ++d;
    return d;  // == world for unroutable requests
}

// All four data-movement kernels give one warp to one request row (rv = 1 + vdim 16-byte vectors): the row's
// bucket, rank and permutation entry are looked up once per warp, the lanes move the vectors.

// PACK + SEND: send position j (bucket order) = request perm[j]; row -> peer window of its owner.
// No fence here: the stores are complete at system scope when the kernel ends, and the barrier kernel that
// follows on the stream publishes them (fence + st.release.sys).
__global__ void __launch_bounds__(256)
xchg_dispatch_perm_kernel(StepRef sr, const int4* __restrict__ digests, const int32_t* __restrict__ perm,
                     const int64_t* __restrict__ offsets, int B, int vdim /*payload vectors per row*/, int world, int rank,
                     int cap, PeerPtrs rows, PeerPtrs counts) {
    __shared__ long long s_off[XCHG_MAX_WORLD + 2];
    const int4* __restrict__ payload = reinterpret_cast<const int4*>(step_resolve(sr).payload);
    trace_mark(sr, TR_DISPATCH);
    if (threadIdx.x < world + 2) s_off[threadIdx.x] = offsets[threadIdx.x];
    __syncthreads();
    const int rv = vdim + 1;  // vectors per window row
    const long long n_send = s_off[world];
    if (blockIdx.x == 0 && threadIdx.x < world)
        static_cast<int32_t*>(counts.p[threadIdx.x])[rank] = (int32_t)(s_off[threadIdx.x + 1] - s_off[threadIdx.x]);
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long j = blockIdx.x * 8ll + (threadIdx.x >> 5); j < n_send; j += warps) {
        const int d = bucket_of(s_off, world, j);
        const long long i = j - s_off[d];
        const int q = perm[j];
        int4* dst = static_cast<int4*>(rows.p[d]) + ((long long)rank * cap + i) * rv;  // NVLink stores into the owner's HBM
        if (lane == 0) dst[0] = digests[q];
        for (int c = lane; c < vdim; c += 32) dst[1 + c] = ldg_stream(payload + (long long)q * vdim + c);
    }
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
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long g = blockIdx.x * 8ll + (threadIdx.x >> 5); g < s_pre[world]; g += warps) {
        int s = 0;
        while (g >= s_pre[s + 1]) ++s;
        const int4* src = rows + ((long long)s * cap + (g - s_pre[s])) * rv;  // written by a peer this step: read at L2
        if (lane == 0) digests_out[g] = __ldcg(src);
        for (int c = lane; c < vdim; c += 32) stg_stream(payload_out + g * vdim + c, __ldcg(src + 1 + c));
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
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long g = blockIdx.x * 8ll + (threadIdx.x >> 5); g < s_pre[world]; g += warps) {
        int s = 0;
        while (g >= s_pre[s + 1]) ++s;
        int4* dst = static_cast<int4*>(back.p[s]) + ((long long)rank * cap + (g - s_pre[s])) * rv;
        for (int c = lane; c < vout; c += 32) dst[c] = ldg_stream(out + g * vout + c);
        if (lane == 0) dst[vout] = make_int4((int)hit[g], 0, 0, 0);
    }
}

// Source: return window -> results in request order (UNPACK of the exchange).
__global__ void __launch_bounds__(256)
xchg_scatter_perm_kernel(StepRef sr, const int4* __restrict__ back, const int32_t* __restrict__ perm,
                    const int64_t* __restrict__ offsets, int B, int world, int cap, int vout) {
    __shared__ long long s_off[XCHG_MAX_WORLD + 2];
    const StepArgs a = step_resolve(sr);
    trace_mark(sr, TR_SCATTER);
    int4* __restrict__ out = reinterpret_cast<int4*>(a.out);
    uint8_t* __restrict__ hit = a.code;
    if (threadIdx.x < world + 2) s_off[threadIdx.x] = offsets[threadIdx.x];
    __syncthreads();
    const int rv = vout + 1;
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long j = blockIdx.x * 8ll + (threadIdx.x >> 5); j < B; j += warps) {
        const int d = bucket_of(s_off, world, j);
        const int q = perm[j];
        const int4* src = d < world ? back + ((long long)d * cap + (j - s_off[d])) * rv : nullptr;  // peer-written: read at L2
        // requests the router returned None for have no result
        for (int c = lane; c < vout; c += 32)
            stg_stream(out + (long long)q * vout + c, src ? __ldcg(src + c) : make_int4(0, 0, 0, 0));
        if (lane == 0) hit[q] = src ? (uint8_t)__ldcg(src + vout).x : (uint8_t)0;
    }
}

// ---- the step's own ingest kernels: no bucket pass, no permutation ---------------------------------------
// ROUTE: MD5 of every key of the ingested batch -> owner rank (model_registry.py:159), and in the same kernel
// the request's rank among the requests of ITS route CTA that go to the same owner (warp match + a count per
// warp), plus the CTA's count per owner.  A request's slot in its owner's window is then
//     (requests for that owner in earlier CTAs) + (rank inside its CTA)         -- request order, as the
// oracle's process_exchange applies -- and the dispatch and scatter kernels work in request order straight from
// these two small arrays: nothing is sorted, nothing is gathered through a permutation.
// The router's view of the owners (Router.route_request's health gate and _find_alternative_shard, router.py:165-221),
// device-resident so that replayed route kernels see an update without being re-captured: shard p = md5 mod world is
// taken if routable[p]; else, with failover, healthy_idx[md5 mod n_healthy]; else the request has no owner (None).
struct XchgHealth {
    int32_t enabled;   // 0: every rank routable (md5 mod world)
    int32_t failover, n_healthy;
    int32_t healthy_idx[XCHG_MAX_WORLD];
    uint8_t routable[XCHG_MAX_WORLD];
};

__global__ void __launch_bounds__(ROUTE_THREADS)
xchg_route_kernel(StepRef sr, int B, int world, int32_t* __restrict__ shard_out, uint4* __restrict__ digests,
                  int32_t* __restrict__ loc, int32_t* __restrict__ cta_cnt /*[gridDim.x][XCHG_MAX_WORLD]*/,
                  const XchgHealth* __restrict__ health) {
    __shared__ __align__(16) uint8_t stage[ROUTE_STAGE_BYTES + 32];
    __shared__ int wcnt[ROUTE_THREADS / 32][XCHG_MAX_WORLD];
    const StepArgs a = step_resolve(sr);
    trace_mark(sr, TR_ROUTE);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < (ROUTE_THREADS / 32) * XCHG_MAX_WORLD) (&wcnt[0][0])[tid] = 0;
    __syncthreads();
    const Md5State st = cta_hash_keys(a.key_bytes, a.key_off, B, stage);  // contains __syncthreads
    const int i = blockIdx.x * ROUTE_THREADS + tid;
    const bool active = i < B;
    int d = -1;
    if (active) {
        digests[i] = make_uint4(st.a, st.b, st.c, st.d);
        d = (int)md5_mod(st, (uint32_t)world);
        if (health->enabled && !health->routable[d])
            d = health->failover && health->n_healthy > 0 ? health->healthy_idx[md5_mod(st, (uint32_t)health->n_healthy)] : -1;
        shard_out[i] = d;
    }
    const unsigned peers = __match_any_sync(0xffffffffu, d);  // lanes of this warp with the same owner
    const int rank_w = __popc(peers & ((1u << lane) - 1u));
    if (active && d >= 0 && rank_w == 0) wcnt[warp][d] = __popc(peers);
    __syncthreads();
    if (active) {
        int r = rank_w;
        for (int w = 0; w < warp && d >= 0; ++w) r += wcnt[w][d];
        loc[i] = d >= 0 ? (d << 16) | r : -1;   // -1: no owner, the request is answered with zeros (Router returns None)
    }
    if (tid < world) {
        int c = 0;
        for (int w = 0; w < ROUTE_THREADS / 32; ++w) c += wcnt[w][tid];
        cta_cnt[blockIdx.x * XCHG_MAX_WORLD + tid] = c;
    }
}

// Exclusive prefix over the route CTAs of the per-owner counts -> s_off[cta * world + d]; returns nothing,
// fills s_tot[d].  Every CTA of the dispatch / scatter kernels redoes this tiny scan in shared memory.
__device__ __forceinline__ void xchg_cta_offsets(const int32_t* __restrict__ cta_cnt, int n_cta, int world, int* s_off,
                                                 int* s_tot) {
    for (int e = threadIdx.x; e < n_cta * world; e += blockDim.x)
        s_off[e] = cta_cnt[(e / world) * XCHG_MAX_WORLD + (e % world)];
    __syncthreads();
    if ((int)threadIdx.x < world) {
        int acc = 0;
        for (int c = 0; c < n_cta; ++c) {
            const int v = s_off[c * world + threadIdx.x];
            s_off[c * world + threadIdx.x] = acc;
            acc += v;
        }
        s_tot[threadIdx.x] = acc;
    }
    __syncthreads();
}

// PACK + SEND in request order: one warp per request row -> its slot in the owner's window (NVLink stores).
// No fence: the stores are complete at system scope when the kernel ends and the barrier kernel publishes them.
__global__ void __launch_bounds__(256)
xchg_dispatch_kernel(StepRef sr, const int4* __restrict__ digests, const int32_t* __restrict__ loc,
                     const int32_t* __restrict__ cta_cnt, int n_cta, int B, int vdim, int world, int rank, int cap,
                     PeerPtrs rows, PeerPtrs counts) {
    extern __shared__ int s_off[];
    __shared__ int s_tot[XCHG_MAX_WORLD];
    const int4* __restrict__ payload = reinterpret_cast<const int4*>(step_resolve(sr).payload);
    trace_mark(sr, TR_DISPATCH);
    xchg_cta_offsets(cta_cnt, n_cta, world, s_off, s_tot);
    if (blockIdx.x == 0 && (int)threadIdx.x < world) static_cast<int32_t*>(counts.p[threadIdx.x])[rank] = s_tot[threadIdx.x];
    const int rv = vdim + 1;
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long q = blockIdx.x * 8ll + (threadIdx.x >> 5); q < B; q += warps) {
        const int l = loc[q];
        if (l < 0) continue;   // no owner
        const int d = l >> 16;
        const long long i = s_off[(q >> 8) * world + d] + (l & 0xffff);
        int4* dst = static_cast<int4*>(rows.p[d]) + ((long long)rank * cap + i) * rv;
        if (lane == 0) dst[0] = digests[q];
        for (int c = lane; c < vdim; c += 32) dst[1 + c] = ldg_stream(payload + q * vdim + c);
    }
}

// UNPACK of the exchange in request order: the request's row of the return window -> out[q], hit[q].
__global__ void __launch_bounds__(256)
xchg_scatter_kernel(StepRef sr, const int4* __restrict__ back, const int32_t* __restrict__ loc,
                    const int32_t* __restrict__ cta_cnt, int n_cta, int B, int world, int cap, int vout) {
    extern __shared__ int s_off[];
    __shared__ int s_tot[XCHG_MAX_WORLD];
    const StepArgs a = step_resolve(sr);
    trace_mark(sr, TR_SCATTER);
    int4* __restrict__ out = reinterpret_cast<int4*>(a.out);
    uint8_t* __restrict__ hit = a.code;
    xchg_cta_offsets(cta_cnt, n_cta, world, s_off, s_tot);
    const int rv = vout + 1;
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long q = blockIdx.x * 8ll + (threadIdx.x >> 5); q < B; q += warps) {
        const int l = loc[q];
        if (l < 0) {   // no owner: zeros, not a hit
            for (int c = lane; c < vout; c += 32) stg_stream(out + q * vout + c, make_int4(0, 0, 0, 0));
            if (lane == 0) hit[q] = 0;
            continue;
        }
        const int d = l >> 16;
        const long long i = s_off[(q >> 8) * world + d] + (l & 0xffff);
        const int4* src = back + ((long long)d * cap + i) * rv;  // peer-written: read at L2
        for (int c = lane; c < vout; c += 32) stg_stream(out + q * vout + c, __ldcg(src + c));
        if (lane == 0) hit[q] = (uint8_t)__ldcg(src + vout).x;
    }
}

// 8 warps per CTA, one row per warp and pass.  One CTA per SM by default: these kernels are a few dependent loads
// deep, and a grid that fills every SM's thread slots for that long keeps the owner pass of the neighbouring
// step (which runs concurrently, on its own stream) waiting; measured on 2 GPUs, 148 CTAs: 46.7 us per step,
// 1184 CTAs: 52.8 us, 37 CTAs: 53.6 us.
static int grid_for_rows(long long rows) {
    const long long blocks = (rows + 7) / 8;
    static int cap_ctas = 0;
    if (cap_ctas == 0) {
        const char* e = getenv("DIE_XCHG_CTAS");
        cap_ctas = e ? atoi(e) : num_sms();
        if (cap_ctas < 1) cap_ctas = num_sms();
    }
    const long long cap = cap_ctas;
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
    StepRef sr;
    memset(&sr, 0, sizeof(sr));
    sr.a.payload = payload;
    DIE_LAUNCH(ST_XCHG, s, xchg_dispatch_perm_kernel<<<grid_for_rows(B), 256, 0, s>>>(
        sr, reinterpret_cast<const int4*>(digests), perm, offsets, B, vdim, world, rank, cap, rows, counts));
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
    DIE_LAUNCH(ST_XCHG_COMPACT, s, xchg_compact_kernel<<<grid_for_rows(2ll * cap), 256, 0, s>>>(
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
    DIE_LAUNCH(ST_XCHG_RETURN, s, xchg_return_kernel<<<grid_for_rows(2ll * cap), 256, 0, s>>>(
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
    StepRef sr;
    memset(&sr, 0, sizeof(sr));
    sr.a.out = out;
    sr.a.code = hit;
    DIE_LAUNCH(ST_XCHG_SCATTER, s, xchg_scatter_perm_kernel<<<grid_for_rows(B), 256, 0, s>>>(
        sr, reinterpret_cast<const int4*>(back), perm, offsets, B, world, cap, vout));
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
// Replayed from a graph, epoch = the lane's epoch before the replay (descriptor) + the barrier's fixed rank in it.
__global__ void xchg_barrier_kernel(PeerPtrs flags, unsigned int* my_flags, int world, int rank, unsigned int epoch,
                                    const QueueDesc* q, int lane, unsigned int* error_flag) {
    const int p = threadIdx.x;
    if (p >= world) return;
    if (q) epoch += q->epoch0[lane];
    if (p == 0) trace_mark_q(q, (int)epoch, TR_BAR_IN);
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
    __syncwarp();
    if (p == 0) trace_mark_q(q, (int)epoch, TR_BAR_OUT);
}

}  // namespace die

// Everything one in-flight step touches: its part of the symmetric window (rows, back, counts, barrier flags)
// and its scratch.  Step b uses set b % XCHG_SETS, so up to XCHG_SETS steps are in flight between ingest and
// scatter.  Reuse is safe without extra synchronisation across ranks: a peer's dispatch of step b+SETS follows
// its barrier of step b+SETS-1, which the owner only reaches after it compacted step b (and the same one stage
// later for the return windows); on this rank the set's `done` event orders its own reuse.
struct XchgLane {
    PeerPtrs rows, back, counts, flags;
    float* my_rows;
    float* my_back;
    int32_t* my_cnt;
    unsigned int* my_flags;
    int32_t* shard;
    uint32_t* digests;
    uint32_t* recv_dig;
    float* recv_pay;
    float* recv_out;
    uint8_t* recv_hit;
    uint32_t* n_dev;
    int32_t* prefix;
    int32_t* loc;      // [cap] owner << 16 | rank among the route CTA's requests for that owner
    int32_t* cta_cnt;  // [ceil(cap / 256)][XCHG_MAX_WORLD] requests per owner of every route CTA
    unsigned int epoch;
    bool used;             // `done` was recorded since the last fork: the next user of the set waits for it
    cudaEvent_t done;      // the step that used this set last has scattered its results
    cudaEvent_t ingested;  // the step's requests are compacted: the owner pass may start
    cudaEvent_t owned;     // the owner pass is done: the results may go back
};

struct die_xchg {
    int world, rank, cap, dim;
    int bmax;  // most requests the owner side handles per step (the pipeline's max_batch)
    die_pipeline_t* pipe;
    XchgLane lane[XCHG_SETS];
    unsigned int* error_flag;
    unsigned long long steps;  // steps enqueued so far: step b uses set b % XCHG_SETS
    // One in-order stream per stage, every step passes through all three: while the owner pass of step b runs,
    // step b+1 is being ingested and step b-1 is on its way back.
    // Ingest and way back are chains of four / three short dependent kernels (~30 / ~20 us per step end to end, most of
    // it launch-to-launch latency and the cross-GPU barrier) against ~22 us for the owner pass: steps alternate between
    // `ways` streams for those stages so that the owner stream -- the only one whose order matters -- is never starved.
    int ways;
    cudaStream_t s_in[XCHG_MAX_WAYS];   // route, dispatch, barrier, compact of steps with lane % ways == w
    cudaStream_t s_own;                 // the pipeline on the received batch (the only stage that touches the table); high priority
    cudaStream_t s_ret[XCHG_MAX_WAYS];  // return, barrier, scatter
    cudaEvent_t entry, exit_ev;
    // captured replays of K steps
    struct Replay {
        int B, K, set0, parts;
        cudaGraphExec_t exec;
    };
    static constexpr int MAX_GRAPHS = 32;
    Replay graphs[MAX_GRAPHS];
    int n_graphs;
    int graph_steps;
    cudaStream_t cap_stream;
    QueueDesc* qdesc;  // device: [0] the ingested queue, [1] the owner passes
    // the lanes' receive-side buffers are one allocation each (lane l at base + l * stride), so that one persistent
    // owner kernel can walk them as rings over the steps of a replayed group
    uint32_t* ring_dig;
    float* ring_pay;
    float* ring_out;
    uint8_t* ring_hit;
    uint32_t* ring_ndev;
    bool owner_group;     // the owner pipeline has the persistent form: groups are replayed software-pipelined
    XchgHealth* health;   // device: the router's view of the owners (die_xchg_set_health)
};

static size_t xchg_rows_bytes(int world, int cap, int dim) { return (size_t)world * cap * (dim + 4) * sizeof(float); }
static size_t xchg_lane_bytes(int world, int cap, int dim) { return 2 * xchg_rows_bytes(world, cap, dim) + 256 + 256; }

extern "C" uint64_t die_xchg_window_bytes(int32_t world, int32_t cap, int32_t dim) {
    // per buffer set: receive rows + return rows + counts + barrier flags
    return XCHG_SETS * xchg_lane_bytes(world, cap, dim);
}

static std::vector<cudaStream_t> xchg_streams(const die_xchg* x) {  // ingest streams, owner stream, return streams
    std::vector<cudaStream_t> v;
    for (int w = 0; w < x->ways; ++w) v.push_back(x->s_in[w]);
    v.push_back(x->s_own);
    for (int w = 0; w < x->ways; ++w) v.push_back(x->s_ret[w]);
    return v;
}

extern "C" die_status_t die_xchg_create(int32_t world, int32_t rank, int32_t cap, int32_t dim, int32_t owner_max,
                                        die_pipeline_t* pipe, void* const* peer_windows, die_xchg_t** out) {
    if (!out || !pipe || !peer_windows || world < 1 || world > XCHG_MAX_WORLD || rank < 0 || rank >= world || cap <= 0 ||
        dim <= 0 || (dim & 3))
        DIE_FAIL(DIE_ERR_INVALID, "die_xchg_create: bad arguments");
    if (cap > 131072) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_xchg_create: cap %d > 131072 requests per ingested batch", cap);
    for (int i = 0; i < world; ++i)
        if (!peer_windows[i]) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_create: null window for rank %d", i);
    die_xchg* x = new die_xchg();
    memset(x, 0, sizeof(*x));
    x->world = world;
    x->rank = rank;
    x->cap = cap;
    x->dim = dim;
    x->pipe = pipe;
    x->bmax = owner_max > 0 && owner_max < world * cap ? owner_max : world * cap;
    const size_t rb = xchg_rows_bytes(world, cap, dim);
    const size_t lb = xchg_lane_bytes(world, cap, dim);
    const size_t bmax = (size_t)x->bmax;
    int prio_low = 0, prio_high = 0;  // numerically lower = higher priority
    DIE_CUDA(cudaDeviceGetStreamPriorityRange(&prio_low, &prio_high));
    for (int l = 0; l < XCHG_SETS; ++l) {
        XchgLane& L = x->lane[l];
        for (int i = 0; i < XCHG_MAX_WORLD; ++i) {
            char* base = i < world ? static_cast<char*>(peer_windows[i]) + l * lb : nullptr;
            L.rows.p[i] = base;
            L.back.p[i] = base ? base + rb : nullptr;
            L.counts.p[i] = base ? base + 2 * rb : nullptr;
            L.flags.p[i] = base ? base + 2 * rb + 256 : nullptr;
        }
        L.my_rows = static_cast<float*>(L.rows.p[rank]);
        L.my_back = static_cast<float*>(L.back.p[rank]);
        L.my_cnt = static_cast<int32_t*>(L.counts.p[rank]);
        L.my_flags = static_cast<unsigned int*>(L.flags.p[rank]);
        DIE_CUDA(cudaMalloc(&L.shard, cap * sizeof(int32_t)));
        DIE_CUDA(cudaMalloc(&L.digests, (size_t)cap * 16));
        if (l == 0) {
            DIE_CUDA(cudaMalloc(&x->ring_dig, XCHG_SETS * bmax * 16));
            DIE_CUDA(cudaMalloc(&x->ring_pay, XCHG_SETS * bmax * dim * sizeof(float)));
            DIE_CUDA(cudaMalloc(&x->ring_out, XCHG_SETS * bmax * dim * sizeof(float)));
            DIE_CUDA(cudaMalloc(&x->ring_hit, XCHG_SETS * bmax));
            DIE_CUDA(cudaMalloc(&x->ring_ndev, XCHG_SETS * sizeof(uint32_t)));
        }
        L.recv_dig = x->ring_dig + (size_t)l * bmax * 4;
        L.recv_pay = x->ring_pay + (size_t)l * bmax * dim;
        L.recv_out = x->ring_out + (size_t)l * bmax * dim;
        L.recv_hit = x->ring_hit + (size_t)l * bmax;
        L.n_dev = x->ring_ndev + l;
        DIE_CUDA(cudaMalloc(&L.prefix, (world + 1) * sizeof(int32_t)));
        DIE_CUDA(cudaMalloc(&L.loc, cap * sizeof(int32_t)));
        DIE_CUDA(cudaMalloc(&L.cta_cnt, (size_t)ceil_div(cap, ROUTE_THREADS) * XCHG_MAX_WORLD * sizeof(int32_t)));
        DIE_CUDA(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
        DIE_CUDA(cudaEventCreateWithFlags(&L.ingested, cudaEventDisableTiming));
        DIE_CUDA(cudaEventCreateWithFlags(&L.owned, cudaEventDisableTiming));
    }
    x->ways = 4;
    if (const char* e = getenv("DIE_XCHG_WAYS")) x->ways = atoi(e);
    if (x->ways < 1 || x->ways > XCHG_MAX_WAYS) x->ways = 4;
    for (int w = 0; w < x->ways; ++w) {
        DIE_CUDA(cudaStreamCreateWithPriority(&x->s_in[w], cudaStreamNonBlocking, prio_low));
        DIE_CUDA(cudaStreamCreateWithPriority(&x->s_ret[w], cudaStreamNonBlocking, prio_low));
    }
    DIE_CUDA(cudaStreamCreateWithPriority(&x->s_own, cudaStreamNonBlocking, prio_high));
    DIE_CUDA(cudaEventCreateWithFlags(&x->exit_ev, cudaEventDisableTiming));
    DIE_CUDA(cudaEventCreateWithFlags(&x->entry, cudaEventDisableTiming));
    DIE_CUDA(cudaMalloc(&x->error_flag, sizeof(unsigned int)));
    DIE_CUDA(cudaMemset(x->error_flag, 0, sizeof(unsigned int)));
    DIE_CUDA(cudaMalloc(&x->qdesc, 2 * sizeof(QueueDesc)));
    DIE_CUDA(cudaMalloc(&x->health, sizeof(XchgHealth)));
    DIE_CUDA(cudaMemset(x->health, 0, sizeof(XchgHealth)));
    x->owner_group = pipeline_owner_group_supported(pipe, x->bmax);
    DIE_CUDA(cudaStreamCreateWithFlags(&x->cap_stream, cudaStreamNonBlocking));
    x->graph_steps = 16;
    if (const char* e = getenv("DIE_GRAPH_STEPS")) x->graph_steps = atoi(e);
    *out = x;
    return DIE_OK;
}

extern "C" die_status_t die_xchg_destroy(die_xchg_t* x) {
    if (!x) return DIE_OK;
    for (cudaStream_t st : xchg_streams(x))
        if (st) {
            cudaStreamSynchronize(st);
            cudaStreamDestroy(st);
        }
    if (x->exit_ev) cudaEventDestroy(x->exit_ev);
    for (int l = 0; l < XCHG_SETS; ++l) {
        XchgLane& L = x->lane[l];
        void* bufs[] = {L.shard, L.digests, L.prefix, L.loc, L.cta_cnt};
        for (void* b : bufs) cudaFree(b);
        if (L.done) cudaEventDestroy(L.done);
        if (L.ingested) cudaEventDestroy(L.ingested);
        if (L.owned) cudaEventDestroy(L.owned);
    }
    for (int g = 0; g < x->n_graphs; ++g) cudaGraphExecDestroy(x->graphs[g].exec);
    if (x->cap_stream) cudaStreamDestroy(x->cap_stream);
    cudaFree(x->qdesc);
    cudaFree(x->health);
    cudaFree(x->ring_dig); cudaFree(x->ring_pay); cudaFree(x->ring_out); cudaFree(x->ring_hit); cudaFree(x->ring_ndev);
    if (x->entry) cudaEventDestroy(x->entry);
    cudaFree(x->error_flag);
    delete x;
    return DIE_OK;
}

// qd != nullptr: the launch is being captured for replay; `epoch_off` is the barrier's rank on its lane inside
// the replay and the lane's epoch before the replay comes from the descriptor.
static die_status_t xchg_barrier(die_xchg* x, XchgLane& L, int lane, const QueueDesc* qd, unsigned int epoch,
                                 cudaStream_t s) {
    DIE_LAUNCH(ST_XCHG_BARRIER, s, xchg_barrier_kernel<<<1, 32, 0, s>>>(L.flags, L.my_flags, x->world, x->rank, epoch, qd,
                                                                        lane, x->error_flag));
    return DIE_OK;
}

// The launches of one step on lane `lane`.  user: keys / payload / out / hit of the ingested batch;
// owner: table view, clock and stamps of the owner pass (its buffers are the lane's).
// reserve_now != nullptr (stream-launched step): the owner pass's host bookkeeping happens here, after the
// ingest half is enqueued, so that table maintenance -- when due -- is ordered behind the previous owner pass
// without holding back this step's ingest.
enum XchgPart { XCHG_ALL = 0, XCHG_INGEST = 1, XCHG_RETURN = 2 };
static die_status_t xchg_launch_step(die_xchg* x, int lane, const StepRef& user, StepRef owner, int32_t B,
                                     int32_t* shard_out, const QueueDesc* qd, unsigned int epoch_first,
                                     const double* reserve_now, XchgPart part = XCHG_ALL) {
    XchgLane& L = x->lane[lane];
    const int W = x->world, D = x->dim, vdim = D / 4;
    int32_t* shard = shard_out ? shard_out : L.shard;
    const int n_cta = std::max(1, ceil_div(B, ROUTE_THREADS));
    const size_t off_bytes = (size_t)n_cta * W * sizeof(int);
    cudaStream_t s = x->s_in[lane % x->ways], so = x->s_own, sr = x->s_ret[lane % x->ways];
    if (part != XCHG_RETURN) {
        // ---- stage 1, ingest: this set's scratch and windows are free once its previous step has scattered
        if (L.used) DIE_CUDA(cudaStreamWaitEvent(s, L.done, 0));
        DIE_LAUNCH(ST_ROUTE, s, xchg_route_kernel<<<n_cta, ROUTE_THREADS, 0, s>>>(
            user, B, W, shard, reinterpret_cast<uint4*>(L.digests), L.loc, L.cta_cnt, x->health));
        DIE_LAUNCH(ST_XCHG, s, xchg_dispatch_kernel<<<grid_for_rows(B), 256, off_bytes, s>>>(
            user, reinterpret_cast<const int4*>(L.digests), L.loc, L.cta_cnt, n_cta, B, vdim, W, x->rank, x->cap, L.rows,
            L.counts));
        DIE_TRY(xchg_barrier(x, L, lane, qd, epoch_first, s));  // every rank's rows have landed in their owners' windows
        DIE_LAUNCH(ST_XCHG_COMPACT, s, xchg_compact_kernel<<<grid_for_rows(2ll * x->cap), 256, 0, s>>>(
            reinterpret_cast<const int4*>(L.my_rows), L.my_cnt, W, x->cap, vdim, (long long)x->bmax,
            reinterpret_cast<int4*>(L.recv_dig), reinterpret_cast<int4*>(L.recv_pay), L.n_dev, L.prefix, x->error_flag));
        DIE_CUDA(cudaEventRecord(L.ingested, s));
    }
    if (part == XCHG_ALL) {
        // ---- stage 2, owner pass: one in-order stream, so the table sees the batches in step order
        DIE_CUDA(cudaStreamWaitEvent(so, L.ingested, 0));
        if (reserve_now) {
            unsigned long long tail0;
            DIE_TRY(pipeline_owner_reserve(x->pipe, x->bmax, 1, so, &owner.a.kv, &tail0, nullptr));
            owner.a.now = *reserve_now;
            owner.a.t_get = tail0;
            owner.a.t_set = tail0 + (unsigned long long)x->bmax;
        }
        owner.a.payload = L.recv_pay;
        owner.a.out = L.recv_out;
        owner.a.code = L.recv_hit;
        DIE_TRY(pipeline_owner_launch(x->pipe, owner, L.recv_dig, x->bmax, L.n_dev, so));
        DIE_CUDA(cudaEventRecord(L.owned, so));
        DIE_CUDA(cudaStreamWaitEvent(sr, L.owned, 0));
    }
    if (part != XCHG_INGEST) {
        // ---- stage 3, way back (grouped form: the caller has ordered s_ret behind the group's owner launch)
        DIE_LAUNCH(ST_XCHG_RETURN, sr, xchg_return_kernel<<<grid_for_rows(2ll * x->cap), 256, 0, sr>>>(
            reinterpret_cast<const int4*>(L.recv_out), L.recv_hit, L.prefix, W, x->rank, x->cap, vdim, L.back));
        DIE_TRY(xchg_barrier(x, L, lane, qd, epoch_first + 1, sr));  // every owner's results are back in the sources' windows
        DIE_LAUNCH(ST_XCHG_SCATTER, sr, xchg_scatter_kernel<<<grid_for_rows(B), 256, off_bytes, sr>>>(
            user, reinterpret_cast<const int4*>(L.my_back), L.loc, L.cta_cnt, n_cta, B, W, x->cap, vdim));
        DIE_CUDA(cudaEventRecord(L.done, sr));
        L.used = true;
    }
    return DIE_OK;
}

// The stage streams start after the work already on `from`; `to` continues after every stage stream's tail.
static die_status_t xchg_fork(die_xchg* x, cudaStream_t from) {
    DIE_CUDA(cudaEventRecord(x->entry, from));
    for (cudaStream_t st : xchg_streams(x)) DIE_CUDA(cudaStreamWaitEvent(st, x->entry, 0));
    for (int l = 0; l < XCHG_SETS; ++l) x->lane[l].used = false;  // everything earlier is ordered through `from`
    return DIE_OK;
}
static die_status_t xchg_join(die_xchg* x, cudaStream_t to) {
    // a chain through every stage stream: a stream pulled into a capture must be joined back even if it did no work
    const std::vector<cudaStream_t> all = xchg_streams(x);
    for (size_t i = 0; i < all.size(); ++i) {
        DIE_CUDA(cudaEventRecord(x->exit_ev, all[i]));
        DIE_CUDA(cudaStreamWaitEvent(i + 1 < all.size() ? all[i + 1] : to, x->exit_ev, 0));
    }
    return DIE_OK;
}

// One stream-launched step: host bookkeeping of the owner pass, then the launches with their values by value.
static die_status_t xchg_enqueue_step(die_xchg* x, const uint8_t* key_bytes, const int64_t* key_off, int32_t B,
                                      const float* payload, double now, float* out, uint8_t* hit_out,
                                      int32_t* shard_out) {
    const int lane = (int)(x->steps % XCHG_SETS);
    XchgLane& L = x->lane[lane];
    x->steps += 1;
    StepRef user, owner;
    memset(&user, 0, sizeof(user));
    memset(&owner, 0, sizeof(owner));
    user.a.key_bytes = key_bytes;
    user.a.key_off = key_off;
    user.a.payload = payload;
    user.a.out = out;
    user.a.code = hit_out;
    const unsigned int e = L.epoch + 1;
    L.epoch += 2;
    return xchg_launch_step(x, lane, user, owner, B, shard_out, nullptr, e, &now);
}

// Captured replays.  qd[0] describes the ingested queue, qd[1] the owner passes.
//
// parts == 0: one graph holds K whole steps (ingest | one owner launch per step | way back, fork / join inside); used
//   when the owner pipeline has no persistent form.  set0 = the first step's buffer set.
// parts != 0: the software-pipelined form.  A launch is up to three INDEPENDENT branches over three different
//   thirds of the buffer sets -- XP_INGEST: the 16 ingests of group g; XP_OWNER: ONE persistent owner launch over the
//   16 batches of group g-1; XP_RETURN: the 16 returns of group g-2 -- and consecutive launches on the caller's stream
//   provide the order between a group's three stages.  The owner kernel never waits for anything in flight, pays its
//   launch, prologue and drain once per 16 steps, and the short dependent kernels of ingest / way back (whose latency
//   is launch-to-launch gaps and the cross-GPU barrier) hide behind it.  set0 = g % 3.
enum XchgParts { XP_INGEST = 1, XP_OWNER = 2, XP_RETURN = 4 };
static die_status_t xchg_graph_get(die_xchg* x, int32_t B, int K, int parts, int set0, cudaGraphExec_t* out) {
    for (int g = 0; g < x->n_graphs; ++g)
        if (x->graphs[g].B == B && x->graphs[g].K == K && x->graphs[g].set0 == set0 && x->graphs[g].parts == parts) {
            *out = x->graphs[g].exec;
            return DIE_OK;
        }
    if (x->n_graphs == die_xchg::MAX_GRAPHS) {
        cudaGraphExecDestroy(x->graphs[0].exec);
        for (int g = 1; g < x->n_graphs; ++g) x->graphs[g - 1] = x->graphs[g];
        x->n_graphs -= 1;
    }
    cudaGraph_t graph = nullptr;
    DIE_CUDA(cudaStreamBeginCapture(x->cap_stream, cudaStreamCaptureModeThreadLocal));
    die_status_t st = xchg_fork(x, x->cap_stream);
    auto refs = [&](int i, StepRef& user, StepRef& owner) {
        memset(&user, 0, sizeof(user));
        memset(&owner, 0, sizeof(owner));
        user.q = x->qdesc;
        user.i = i;
        owner.q = x->qdesc + 1;
        owner.i = i;
    };
    if (parts) {
        const int third_in = set0, third_own = (set0 + 2) % 3, third_ret = (set0 + 1) % 3;  // groups g, g-1, g-2
        if (parts & XP_INGEST)
            for (int i = 0; i < K && st == DIE_OK; ++i) {
                StepRef user, owner;
                refs(i, user, owner);  // the descriptor's j0 is group g's first batch
                st = xchg_launch_step(x, third_in * XCHG_GROUP + i, user, owner, B, nullptr, x->qdesc, 1u, nullptr, XCHG_INGEST);
            }
        if ((parts & XP_OWNER) && st == DIE_OK) {
            StepLanes ln;
            ln.lanes = XCHG_SETS;
            ln.lane0 = third_own * XCHG_GROUP;
            st = pipeline_owner_launch_group(x->pipe, x->qdesc + 1, 0, K, x->ring_dig, x->bmax, x->ring_ndev, ln, x->s_own);
        }
        if (parts & XP_RETURN)
            for (int i = 0; i < K && st == DIE_OK; ++i) {
                StepRef user, owner;
                refs(i - 2 * K, user, owner);  // group g-2's slots of the caller's out / hit rings
                // the way back's barrier is epoch_first + 1: one past the set's epoch at this launch, like the ingest's
                st = xchg_launch_step(x, third_ret * XCHG_GROUP + i, user, owner, B, nullptr, x->qdesc, 0u, nullptr, XCHG_RETURN);
            }
    } else {
        unsigned int off[XCHG_SETS];
        for (int l = 0; l < XCHG_SETS; ++l) off[l] = 1u;
        for (int i = 0; i < K && st == DIE_OK; ++i) {
            const int lane = (set0 + i) % XCHG_SETS;
            StepRef user, owner;
            refs(i, user, owner);
            st = xchg_launch_step(x, lane, user, owner, B, nullptr, x->qdesc, off[lane], nullptr);
            off[lane] += 2;
        }
    }
    if (st == DIE_OK) st = xchg_join(x, x->cap_stream);
    const cudaError_t e = cudaStreamEndCapture(x->cap_stream, &graph);
    if (st != DIE_OK) {
        if (graph) cudaGraphDestroy(graph);
        return st;
    }
    DIE_CUDA(e);
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    DIE_CUDA(ei);
    die_xchg::Replay& r = x->graphs[x->n_graphs++];
    r.B = B;
    r.K = K;
    r.set0 = set0;
    r.parts = parts;
    r.exec = exec;
    *out = exec;
    return DIE_OK;
}

__global__ void xchg_desc_set_kernel(QueueDesc* dst, QueueDesc user, QueueDesc owner, KvMirrorArgs m) {
    dst[0] = user;
    dst[1] = owner;
    if (m.m) kv_mirror_write(m.st, m.m, m.ops_seen, m.gen, m.seq);  // state before this owner launch -> the host's bounds
}

// One flushed batch ingested on this rank; results in request order.  Every rank of the group must make the
// same sequence of calls (the barriers inside are collective).
extern "C" die_status_t die_xchg_step(die_xchg_t* x, const uint8_t* key_bytes, const int64_t* key_off, int32_t B,
                                      const float* payload, double now, float* out, uint8_t* hit_out,
                                      int32_t* shard_out, die_stream_t stream) {
    if (!x || B < 0 || B > x->cap) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step: B must be in [0, cap]");
    if (!key_off || !payload || !out || !hit_out) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step: null buffer");
    cudaStream_t user = static_cast<cudaStream_t>(stream);
    DIE_TRY(xchg_fork(x, user));
    DIE_TRY(xchg_enqueue_step(x, key_bytes, key_off, B, payload, now, out, hit_out, shard_out));
    return xchg_join(x, user);
}

// A queue of flushed batches.  Consecutive steps use different buffer sets and every stage has its own stream,
// so a step's ingest and return overlap its neighbours' owner passes; groups of K steps are replayed from one captured graph (host
// cost: two launches per K steps).  out / hit_out are rings of out_ring slots (>= 2 when n_batches > 1).
extern "C" die_status_t die_xchg_step_many(die_xchg_t* x, int32_t n_batches, const uint8_t* key_bytes,
                                           int64_t key_bytes_stride, const int64_t* key_off, int64_t key_off_stride,
                                           int32_t B, const float* payload, double now0, double now_step, float* out,
                                           uint8_t* hit_out, int32_t out_ring, die_stream_t stream) {
    if (!x || n_batches < 0 || out_ring < 1) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step_many: bad arguments");
    if (B < 0 || B > x->cap) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step_many: B must be in [0, cap]");
    if (!key_off || !payload || !out || !hit_out) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step_many: null buffer");
    if (n_batches > 1 && out_ring < 2) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step_many: out_ring must be >= 2");
    cudaStream_t user = static_cast<cudaStream_t>(stream);
    const size_t row = (size_t)B * x->dim;
    int32_t j = 0;
    const int K = x->graph_steps;
    if (K >= 2 && B > 0 && n_batches >= K && pipeline_graph_fits(x->pipe, x->bmax, K)) {
        QueueDesc du, dw;
        memset(&du, 0, sizeof(du));
        memset(&dw, 0, sizeof(dw));
        du.key_bytes = key_bytes;
        du.key_off = key_off;
        du.payload = payload;
        du.out = out;
        du.hit = hit_out;
        du.key_bytes_stride = key_bytes_stride;
        du.key_off_stride = key_off_stride;
        du.payload_stride = (long long)row;
        du.out_stride = (long long)row;
        du.hit_stride = B;
        du.ring = out_ring;
        dw.ring = 1;
        dw.batch = x->bmax;
        dw.now0 = now0;
        dw.now_step = now_step;
        // replayed groups start on a group boundary of the buffer sets (stream-launched remainders shift the counter;
        // skipping sets is harmless, every set carries its own barrier epoch)
        x->steps = (x->steps + (unsigned long long)K - 1) / K * K;
        // the way back of a group runs on several streams: two batches of one group must not share an out / hit slot
        if (out_ring < K && n_batches > out_ring)
            DIE_FAIL(DIE_ERR_INVALID, "die_xchg_step_many: out_ring must hold a group of %d batches (or the whole queue)", K);
        const bool pipelined = x->owner_group && K <= XCHG_GROUP;
        const int n_groups = n_batches / K;
        if (pipelined) {
            // launch t: ingest of group t | owner launch of group t-1 | way back of group t-2 (see xchg_graph_get)
            dw.payload = x->ring_pay;  // the owner's persistent launch walks the sets' buffers as rings
            dw.out = x->ring_out;
            dw.hit = x->ring_hit;
            dw.payload_stride = dw.out_stride = (long long)x->bmax * x->dim;
            dw.hit_stride = x->bmax;
            dw.ring = dw.pay_ring = XCHG_SETS;
            for (int t = 0; t < n_groups + 2; ++t) {
                const unsigned long long g = x->steps / K;  // the group launch t would ingest
                const int third_in = (int)(g % 3), third_own = (int)((g + 2) % 3), third_ret = (int)((g + 1) % 3);
                const int parts = (t < n_groups ? XP_INGEST : 0) | (t >= 1 && t - 1 < n_groups ? XP_OWNER : 0) |
                                  (t >= 2 ? XP_RETURN : 0);
                cudaGraphExec_t exec;
                DIE_TRY(xchg_graph_get(x, B, K, parts, third_in, &exec));
                KvMirrorArgs mirror;
                memset(&mirror, 0, sizeof(mirror));
                if (parts & XP_OWNER) {
                    unsigned long long tail0;
                    DIE_TRY(pipeline_owner_reserve(x->pipe, x->bmax, K, user, &dw.kv, &tail0, &mirror));
                    dw.tail0 = tail0;
                    dw.j0 = (long long)third_own * XCHG_GROUP;      // buffers: step i of the group uses set third_own * 16 + i
                    dw.jt = (long long)(t - 1) * K - dw.j0;          // clock: still batch (t-1) * K + i of the queue
                }
                du.j0 = (long long)t * K;  // ingest: batch t*K + i; way back (StepRef.i = i - 2K): batch (t-2)*K + i
                for (int l = 0; l < XCHG_SETS; ++l) du.epoch0[l] = x->lane[l].epoch;
                du.trace = dw.trace = trace_buffer();
                xchg_desc_set_kernel<<<1, 1, 0, user>>>(x->qdesc, du, dw, mirror);
                DIE_LAUNCH_CHECK();
                DIE_CUDA(cudaGraphLaunch(exec, user));
                if (parts & XP_INGEST) count_replay();  // groups replayed (the two draining launches carry no new group)
                count_launch((unsigned)((parts & XP_INGEST ? 4 * K : 0) + (parts & XP_OWNER ? 1 : 0) + (parts & XP_RETURN ? 3 * K : 0)));
                for (int i = 0; i < K; ++i) {  // one barrier per step and stage
                    if (parts & XP_INGEST) x->lane[third_in * XCHG_GROUP + i].epoch += 1u;
                    if (parts & XP_RETURN) x->lane[third_ret * XCHG_GROUP + i].epoch += 1u;
                }
                // the counter advances with every launch, so that the three thirds rotate through the drain as well;
                // the sets the last two launches "ingest" nothing into are simply skipped
                x->steps += (unsigned long long)K;
            }
            j = n_groups * K;
        } else {
            for (; n_batches - j >= K; j += K) {
                const int set0 = (int)(x->steps % XCHG_SETS);
                cudaGraphExec_t exec;
                DIE_TRY(xchg_graph_get(x, B, K, 0, set0, &exec));
                unsigned long long tail0;
                KvMirrorArgs mirror;
                DIE_TRY(pipeline_owner_reserve(x->pipe, x->bmax, K, user, &dw.kv, &tail0, &mirror));
                dw.tail0 = tail0;
                du.j0 = dw.j0 = j;
                for (int l = 0; l < XCHG_SETS; ++l) du.epoch0[l] = x->lane[l].epoch;
                du.trace = dw.trace = trace_buffer();
                xchg_desc_set_kernel<<<1, 1, 0, user>>>(x->qdesc, du, dw, mirror);
                DIE_LAUNCH_CHECK();
                DIE_CUDA(cudaGraphLaunch(exec, user));
                count_replay();
                count_launch((unsigned)(K * (7 + pipeline_launches_per_step(x->pipe))));
                for (int i = 0; i < K; ++i) x->lane[(set0 + i) % XCHG_SETS].epoch += 2u;  // two barriers per step
                x->steps += (unsigned long long)K;
            }
        }
    }
    if (j == n_batches) return DIE_OK;
    DIE_TRY(xchg_fork(x, user));
    for (; j < n_batches; ++j) {
        const int r = j % out_ring;
        DIE_TRY(xchg_enqueue_step(x, key_bytes + (size_t)j * key_bytes_stride, key_off + (size_t)j * key_off_stride, B,
                                  payload + (size_t)j * row, now0 + j * now_step, out + (size_t)r * row,
                                  hit_out + (size_t)r * B, nullptr));
    }
    return xchg_join(x, user);
}

// Builds (captures + instantiates) the replay die_xchg_step_many(B, n_batches >= graph steps) would use next, without
// launching anything: instantiation may allocate and synchronise the device, which a serving loop wants out of its
// first timed call -- and which a group of ranks sharing ONE device (loopback tests) cannot afford while a peer's
// barrier is already spinning on that device.
extern "C" die_status_t die_xchg_prepare(die_xchg_t* x, int32_t B) {
    if (!x || B <= 0 || B > x->cap) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_prepare: B must be in [1, cap]");
    const int K = x->graph_steps;
    if (K < 2 || !pipeline_graph_fits(x->pipe, x->bmax, K)) return DIE_OK;  // steps will be stream-launched
    die_status_t st = DIE_OK;
    cudaGraphExec_t exec;
    if (x->owner_group && K <= XCHG_GROUP) {
        const int all_parts[] = {XP_INGEST, XP_INGEST | XP_OWNER, XP_INGEST | XP_OWNER | XP_RETURN, XP_OWNER | XP_RETURN, XP_RETURN, XP_OWNER};
        for (int parts : all_parts)
            for (int third = 0; third < 3 && st == DIE_OK; ++third) st = xchg_graph_get(x, B, K, parts, third, &exec);
    } else {
        for (int set0 = 0; set0 < XCHG_SETS && st == DIE_OK; set0 += K) st = xchg_graph_get(x, B, K, 0, set0 % XCHG_SETS, &exec);
    }
    return st;
}

// The router's health view for the steps enqueued after this call (every rank of the group must pass the same one).
// routable == nullptr: back to plain md5 mod world.
extern "C" die_status_t die_xchg_set_health(die_xchg_t* x, const uint8_t* routable, const int32_t* healthy_idx,
                                            int32_t n_healthy, int32_t failover, die_stream_t stream) {
    if (!x || n_healthy < 0 || n_healthy > x->world || (n_healthy > 0 && !healthy_idx))
        DIE_FAIL(DIE_ERR_INVALID, "die_xchg_set_health: bad arguments");
    XchgHealth h;
    memset(&h, 0, sizeof(h));
    if (routable) {
        h.enabled = 1;
        h.failover = failover ? 1 : 0;
        h.n_healthy = n_healthy;
        for (int i = 0; i < x->world; ++i) h.routable[i] = routable[i] ? 1 : 0;
        for (int i = 0; i < n_healthy; ++i) {
            if (healthy_idx[i] < 0 || healthy_idx[i] >= x->world) DIE_FAIL(DIE_ERR_INVALID, "die_xchg_set_health: healthy_idx out of range");
            h.healthy_idx[i] = healthy_idx[i];
        }
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_CUDA(cudaMemcpyAsync(x->health, &h, sizeof(h), cudaMemcpyHostToDevice, s));  // pageable source: staged before return
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
