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
                    int4* __restrict__ digests_out, int4* __restrict__ payload_out, uint32_t* __restrict__ n_out,
                    int32_t* __restrict__ prefix_out /*[world+1]*/) {
    __shared__ long long s_pre[XCHG_MAX_WORLD + 1];
    if (threadIdx.x == 0) {
        long long acc = 0;
        for (int s = 0; s < world; ++s) {
            s_pre[s] = acc;
            acc += counts[s];
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
    if (world < 1 || world > XCHG_MAX_WORLD || dim <= 0 || (dim & 3) || cap <= 0)
        DIE_FAIL(DIE_ERR_INVALID, "die_xchg_compact: bad sizes");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int vdim = dim / 4;
    DIE_LAUNCH(ST_XCHG, s, xchg_compact_kernel<<<grid_for_vectors((long long)2 * cap * (vdim + 1)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(rows), counts, world, cap, vdim, reinterpret_cast<int4*>(digests_out),
        reinterpret_cast<int4*>(payload_out), n_out, prefix_out));
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
