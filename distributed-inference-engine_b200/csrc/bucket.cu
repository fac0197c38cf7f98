// Stable bucket-by-destination of a routed batch: the "group requests by owner shard" step that sits
// between Router.route_request (src/router.py:140) and the per-worker hand-off in SURVEY.md call-stack (E).
// The reference has no such code (its coordinator was never written); order inside a bucket is request
// order, which is what fixes the owner-side "request order" the LRU parity is defined against.
//
// One CTA of 1024 threads walks the batch in tiles of 4096 requests; for every destination it ranks the
// tile's requests with a block scan, so positions are deterministic (no atomics).  Requests routed to
// None (-1) form a last bucket.  Sized for dispatch fan-outs (n_shards <= 64), not for general sorting.
#include "bucket_device.cuh"

namespace die {

constexpr int BK_THREADS = 1024;
constexpr int BK_ITEMS = 4;
constexpr int BK_MAX_SHARDS = 64;

__global__ void __launch_bounds__(BK_THREADS)
bucket_by_shard_kernel(const int32_t* __restrict__ shard, int B, int n_shards, int64_t* __restrict__ offsets,
                       int32_t* __restrict__ perm, int32_t* __restrict__ pos) {
    __shared__ unsigned s_warp[32];
    __shared__ unsigned s_total[BK_MAX_SHARDS + 1];
    __shared__ unsigned s_base[BK_MAX_SHARDS + 2];
    const int tid = threadIdx.x;
    const int nb = n_shards + 1;  // last bucket: unroutable
    // pass 1: bucket sizes
    for (int d = tid; d < nb; d += BK_THREADS) s_total[d] = 0;
    __syncthreads();
    for (int i = tid; i < B; i += BK_THREADS) {
        const int s = shard[i];
        atomicAdd(&s_total[s < 0 || s >= n_shards ? n_shards : s], 1u);
    }
    __syncthreads();
    if (tid == 0) {
        unsigned acc = 0;
        for (int d = 0; d < nb; ++d) {
            s_base[d] = acc;
            offsets[d] = acc;
            acc += s_total[d];
        }
        s_base[nb] = acc;
        offsets[nb] = acc;
    }
    __syncthreads();
    // pass 2: stable ranks, tile by tile
    for (int t0 = 0; t0 < B; t0 += BK_THREADS * BK_ITEMS) {
        int my[BK_ITEMS];
#pragma unroll
        for (int k = 0; k < BK_ITEMS; ++k) {
            const int i = t0 + tid * BK_ITEMS + k;
            int s = i < B ? shard[i] : -2;
            if (s != -2 && (s < 0 || s >= n_shards)) s = n_shards;
            my[k] = s;
        }
        for (int d = 0; d < nb; ++d) {
            unsigned cnt = 0;
#pragma unroll
            for (int k = 0; k < BK_ITEMS; ++k) cnt += my[k] == d;
            unsigned excl;
            const unsigned total = block_scan_1024(cnt, excl, s_warp);
            unsigned r = s_base[d] + excl;
#pragma unroll
            for (int k = 0; k < BK_ITEMS; ++k) {
                if (my[k] == d) {
                    const int i = t0 + tid * BK_ITEMS + k;
                    perm[r] = i;
                    pos[i] = (int32_t)r;
                    ++r;
                }
            }
            __syncthreads();
            if (tid == 0) s_base[d] += total;
            __syncthreads();
        }
    }
}

// Fan-outs of one box (<= 19 shards): packed-count ranking, see bucket_device.cuh.
__global__ void __launch_bounds__(BKT_THREADS)
bucket_small_kernel(const int32_t* __restrict__ shard, int B, int n_shards, int64_t* __restrict__ offsets,
                    int32_t* __restrict__ perm, int32_t* __restrict__ pos) {
    __shared__ BktSmem sm;
    bucket_cta(shard, B, n_shards, offsets, perm, pos, sm);
}

}  // namespace die

extern "C" die_status_t die_bucket_by_shard(const int32_t* shard, int32_t B, int32_t n_shards, int64_t* offsets_out,
                                            int32_t* perm_out, int32_t* pos_out, die_stream_t stream) {
    using namespace die;
    if (B < 0 || n_shards < 0 || n_shards > BK_MAX_SHARDS)
        DIE_FAIL(DIE_ERR_INVALID, "die_bucket_by_shard: n_shards must be in [0, %d]", BK_MAX_SHARDS);
    if (!offsets_out || (B > 0 && (!shard || !perm_out || !pos_out)))
        DIE_FAIL(DIE_ERR_INVALID, "die_bucket_by_shard: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n_shards <= BKT_MAX_SHARDS)
        DIE_LAUNCH(ST_BUCKET, s, bucket_small_kernel<<<1, BKT_THREADS, 0, s>>>(shard, B, n_shards, offsets_out, perm_out, pos_out));
    else
        DIE_LAUNCH(ST_BUCKET, s, bucket_by_shard_kernel<<<1, BK_THREADS, 0, s>>>(shard, B, n_shards, offsets_out, perm_out, pos_out));
    return DIE_OK;
}

// ---- rows to / from per-shard regions (the hand-off of the single-GPU Coordinator, die/coordinator.py) -----------
// After die_bucket_by_shard, request q sits at send position pos[q] of bucket shard[q].  bucket_scatter writes row q
// of `src` to region shard[q] of `dst` ([n_shards][cap][row]) at index pos[q] - offsets[shard[q]] -- request order
// inside every shard, which is the order that shard's cache sees -- and the per-shard counts; bucket_gather brings
// result rows back to request order (zeros where the router returned None).  One warp per row, 16-byte vectors.
namespace die {

__global__ void __launch_bounds__(256)
bucket_scatter_kernel(const int4* __restrict__ src, const int32_t* __restrict__ shard, const int32_t* __restrict__ pos,
                      const int64_t* __restrict__ offsets, int B, int n_shards, int cap, int vrow, int4* __restrict__ dst,
                      uint32_t* __restrict__ counts) {
    if (blockIdx.x == 0 && (int)threadIdx.x < n_shards)
        counts[threadIdx.x] = (uint32_t)(offsets[threadIdx.x + 1] - offsets[threadIdx.x]);
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long q = blockIdx.x * 8ll + (threadIdx.x >> 5); q < B; q += warps) {
        const int s = shard[q];
        if (s < 0 || s >= n_shards) continue;
        const long long i = (long long)pos[q] - offsets[s];
        int4* row = dst + ((long long)s * cap + i) * vrow;
        for (int c = lane; c < vrow; c += 32) stg_stream(row + c, ldg_stream(src + q * vrow + c));
    }
}

__global__ void __launch_bounds__(256)
bucket_gather_kernel(const int4* __restrict__ src, const int32_t* __restrict__ shard, const int32_t* __restrict__ pos,
                     const int64_t* __restrict__ offsets, int B, int n_shards, int cap, int vrow, int4* __restrict__ dst,
                     const uint8_t* __restrict__ flag_src, uint8_t* __restrict__ flag_dst) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long q = blockIdx.x * 8ll + (threadIdx.x >> 5); q < B; q += warps) {
        const int s = shard[q];
        const bool routed = s >= 0 && s < n_shards;
        const int4* row = routed ? src + ((long long)s * cap + ((long long)pos[q] - offsets[s])) * vrow : nullptr;
        for (int c = lane; c < vrow; c += 32) stg_stream(dst + q * vrow + c, routed ? ldg_stream(row + c) : make_int4(0, 0, 0, 0));
        if (flag_dst && lane == 0) flag_dst[q] = routed ? flag_src[(long long)s * cap + ((long long)pos[q] - offsets[s])] : (uint8_t)0;
    }
}

}  // namespace die

static int bucket_rows_grid(long long rows) {
    const long long blocks = (rows + 7) / 8, cap = (long long)die::num_sms() * 8;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

extern "C" die_status_t die_bucket_scatter_rows(const void* src, const int32_t* shard, const int32_t* pos,
                                                const int64_t* offsets, int32_t B, int32_t n_shards, int32_t cap,
                                                int32_t row_bytes, void* dst, uint32_t* counts_out, die_stream_t stream) {
    using namespace die;
    if (B < 0 || n_shards < 1 || n_shards > BK_MAX_SHARDS || cap < B || row_bytes <= 0 || (row_bytes & 15))
        DIE_FAIL(DIE_ERR_INVALID, "die_bucket_scatter_rows: need cap >= B and rows of whole 16-byte vectors");
    if (!src || !shard || !pos || !offsets || !dst || !counts_out) DIE_FAIL(DIE_ERR_INVALID, "die_bucket_scatter_rows: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_LAUNCH(ST_BUCKET, s, bucket_scatter_kernel<<<bucket_rows_grid(B), 256, 0, s>>>(
        static_cast<const int4*>(src), shard, pos, offsets, B, n_shards, cap, row_bytes / 16, static_cast<int4*>(dst), counts_out));
    return DIE_OK;
}

extern "C" die_status_t die_bucket_gather_rows(const void* src, const int32_t* shard, const int32_t* pos,
                                               const int64_t* offsets, int32_t B, int32_t n_shards, int32_t cap,
                                               int32_t row_bytes, void* dst, const uint8_t* flag_src, uint8_t* flag_dst,
                                               die_stream_t stream) {
    using namespace die;
    if (B < 0 || n_shards < 1 || n_shards > BK_MAX_SHARDS || cap < B || row_bytes <= 0 || (row_bytes & 15))
        DIE_FAIL(DIE_ERR_INVALID, "die_bucket_gather_rows: need cap >= B and rows of whole 16-byte vectors");
    if (B == 0) return DIE_OK;
    if (!src || !shard || !pos || !offsets || !dst) DIE_FAIL(DIE_ERR_INVALID, "die_bucket_gather_rows: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_LAUNCH(ST_BUCKET, s, bucket_gather_kernel<<<bucket_rows_grid(B), 256, 0, s>>>(
        static_cast<const int4*>(src), shard, pos, offsets, B, n_shards, cap, row_bytes / 16, static_cast<int4*>(dst),
        flag_src, flag_src ? flag_dst : nullptr));
    return DIE_OK;
}
