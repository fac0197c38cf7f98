// K1: batched MD5 key hash -> shard position with the router's health gate and failover.
// Replaces ModelRegistry.get_shard_for_key (src/model_registry.py:149-161), Router.route_request
// (src/router.py:140-184) and Router._find_alternative_shard (src/router.py:186-221).
//
// Layout: keys arrive as CSR (bytes + int64 offsets).  A CTA of 256 threads owns 256 consecutive
// keys; the CTA's contiguous byte span is staged into shared memory with coalesced 128-bit loads
// (warp-cooperative), then every lane hashes its own key out of shared memory.  Spans that do not
// fit (very long keys) are read straight from global memory by the owning lane.
#include "route_device.cuh"

namespace die {

__global__ void __launch_bounds__(ROUTE_THREADS)
route_md5_kernel(const uint8_t* __restrict__ key_bytes, const int64_t* __restrict__ key_off, int B,
                 int n_shards, const uint8_t* __restrict__ routable,
                 const int32_t* __restrict__ healthy_idx, int n_healthy, int failover,
                 int32_t* __restrict__ shard_out, uint4* __restrict__ digest_out,
                 int32_t* __restrict__ dest_counts) {
    __shared__ __align__(16) uint8_t stage[ROUTE_STAGE_BYTES + 32];
    __shared__ int s_counts[ROUTE_SMEM_SHARDS];

    const int tid = threadIdx.x;
    const bool count_smem = dest_counts != nullptr && n_shards <= ROUTE_SMEM_SHARDS;
    if (count_smem)
        for (int s = tid; s < n_shards; s += ROUTE_THREADS) s_counts[s] = 0;
    const Md5State st = cta_hash_keys(key_bytes, key_off, B, stage);
    const int i = blockIdx.x * ROUTE_THREADS + tid;
    int shard = -1;
    if (i < B) {
        if (digest_out) digest_out[i] = make_uint4(st.a, st.b, st.c, st.d);
        if (n_shards > 0) {
            const uint32_t p = md5_mod(st, (uint32_t)n_shards);
            const bool ok = routable ? (routable[p] != 0) : true;
            if (ok)
                shard = (int)p;
            else if (failover && n_healthy > 0)
                shard = healthy_idx[md5_mod(st, (uint32_t)n_healthy)];
        }
        shard_out[i] = shard;
    }
    if (dest_counts) {
        if (count_smem) {
            if (shard >= 0) atomicAdd(&s_counts[shard], 1);
            __syncthreads();
            for (int s = tid; s < n_shards; s += ROUTE_THREADS)
                if (s_counts[s]) atomicAdd(&dest_counts[s], s_counts[s]);
        } else if (shard >= 0) {
            atomicAdd(&dest_counts[shard], 1);
        }
    }
}

}  // namespace die

extern "C" die_status_t die_route_md5(const uint8_t* key_bytes, const int64_t* key_off, int32_t B,
                                      int32_t n_shards, const uint8_t* routable,
                                      const int32_t* healthy_idx, int32_t n_healthy, int32_t failover,
                                      int32_t* shard_out, uint32_t* digest_out, int32_t* dest_counts,
                                      die_stream_t stream) {
    using namespace die;
    if (B < 0 || n_shards < 0 || n_healthy < 0) DIE_FAIL(DIE_ERR_INVALID, "die_route_md5: negative size");
    if (B > 0 && (!key_off || !shard_out)) DIE_FAIL(DIE_ERR_INVALID, "die_route_md5: null key_off/shard_out");
    if (n_healthy > 0 && !healthy_idx) DIE_FAIL(DIE_ERR_INVALID, "die_route_md5: null healthy_idx");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dest_counts && n_shards > 0) DIE_CUDA(cudaMemsetAsync(dest_counts, 0, sizeof(int32_t) * n_shards, s));
    if (B == 0) return DIE_OK;
    DIE_LAUNCH(ST_ROUTE, s, route_md5_kernel<<<ceil_div(B, ROUTE_THREADS), ROUTE_THREADS, 0, s>>>(
        key_bytes, key_off, B, n_shards, routable, healthy_idx, n_healthy, failover, shard_out,
        reinterpret_cast<uint4*>(digest_out), dest_counts));
    return DIE_OK;
}
