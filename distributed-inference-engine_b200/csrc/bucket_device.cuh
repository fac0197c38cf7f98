// Stable bucket-by-destination for dispatch fan-outs of at most 16 shards, run by ONE CTA of 256 threads
// (the stand-alone kernel of bucket.cu, or the last CTA of the exchange's route kernel).
//
// Thread t owns 16 consecutive requests of every 4096-request chunk, so "rank inside the bucket" is
// request order by construction.  Per-thread bucket counts are 16-bit fields packed four to a 64-bit word
// (a chunk holds 4096 requests, so every prefix fits 13 bits): one block scan of the packed words ranks
// all buckets at once instead of one scan per bucket.
#pragma once
#include "common.cuh"

namespace die {

constexpr int BKT_THREADS = 256;
constexpr int BKT_ITEMS = 16;
constexpr int BKT_WORDS = 5;                    // 5 x 4 fields: up to 19 buckets + the "unroutable" one
constexpr int BKT_MAX_SHARDS = BKT_WORDS * 4 - 1;

struct BktWords {
    unsigned long long w[BKT_WORDS];
};

__device__ __forceinline__ void bkt_add(BktWords& a, int d) {
    const unsigned long long inc = 1ull << (16 * (d & 3));
#pragma unroll
    for (int k = 0; k < BKT_WORDS; ++k) a.w[k] += (d >> 2) == k ? inc : 0ull;
}
__device__ __forceinline__ unsigned bkt_field(const BktWords& a, int d) {
    unsigned long long v = 0ull;
#pragma unroll
    for (int k = 0; k < BKT_WORDS; ++k) v = (d >> 2) == k ? a.w[k] : v;
    return (unsigned)(v >> (16 * (d & 3))) & 0xffffu;
}

struct BktSmem {
    unsigned long long warp_tot[BKT_THREADS / 32][BKT_WORDS];
    unsigned base[BKT_MAX_SHARDS + 2];
    unsigned total[BKT_MAX_SHARDS + 1];
};

// shard[] may have been written by other CTAs of the same kernel: read at L2.
__device__ __forceinline__ int bkt_load(const int32_t* shard, int i, int B, int n_shards) {
    if (i >= B) return -1;
    const int s = __ldcg(shard + i);
    return s < 0 || s >= n_shards ? n_shards : s;  // routed to None: the last bucket
}

// All BKT_THREADS threads of the CTA call this.  offsets [n_shards + 2], perm [B], pos [B].
static __device__ void bucket_cta(const int32_t* __restrict__ shard, int B, int n_shards, int64_t* __restrict__ offsets,
                           int32_t* __restrict__ perm, int32_t* __restrict__ pos, BktSmem& sm) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = n_shards + 1;
    constexpr int CHUNK = BKT_THREADS * BKT_ITEMS;
    if (tid <= BKT_MAX_SHARDS) sm.total[tid] = 0u;
    __syncthreads();
    // pass 1: bucket sizes (warp-reduced packed counts, then a handful of shared-memory atomics per warp)
    for (int c0 = 0; c0 < B; c0 += CHUNK) {
        BktWords mine{};
#pragma unroll
        for (int k = 0; k < BKT_ITEMS; ++k) {
            const int d = bkt_load(shard, c0 + tid * BKT_ITEMS + k, B, n_shards);
            if (d >= 0) bkt_add(mine, d);
        }
#pragma unroll
        for (int k = 0; k < BKT_WORDS; ++k)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mine.w[k] += __shfl_xor_sync(0xffffffffu, mine.w[k], o);
        if (lane < nb) {
            const unsigned f = bkt_field(mine, lane);
            if (f) atomicAdd(&sm.total[lane], f);
        }
    }
    __syncthreads();
    if (tid == 0) {
        unsigned acc = 0;
        for (int d = 0; d < nb; ++d) {
            sm.base[d] = acc;
            offsets[d] = acc;
            acc += sm.total[d];
        }
        sm.base[nb] = acc;
        offsets[nb] = acc;
    }
    __syncthreads();
    // pass 2: stable ranks, chunk by chunk
    for (int c0 = 0; c0 < B; c0 += CHUNK) {
        int d[BKT_ITEMS];
        BktWords mine{};
#pragma unroll
        for (int k = 0; k < BKT_ITEMS; ++k) {
            d[k] = bkt_load(shard, c0 + tid * BKT_ITEMS + k, B, n_shards);
            if (d[k] >= 0) bkt_add(mine, d[k]);
        }
        // exclusive block scan of the packed words
        BktWords incl = mine;
#pragma unroll
        for (int k = 0; k < BKT_WORDS; ++k)
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long v = __shfl_up_sync(0xffffffffu, incl.w[k], o);
                if (lane >= o) incl.w[k] += v;
            }
        if (lane == 31)
#pragma unroll
            for (int k = 0; k < BKT_WORDS; ++k) sm.warp_tot[warp][k] = incl.w[k];
        __syncthreads();
        BktWords excl;
#pragma unroll
        for (int k = 0; k < BKT_WORDS; ++k) {
            unsigned long long before = 0ull;
            for (int w = 0; w < warp; ++w) before += sm.warp_tot[w][k];
            excl.w[k] = before + incl.w[k] - mine.w[k];
        }
        BktWords run{};
#pragma unroll
        for (int k = 0; k < BKT_ITEMS; ++k) {
            if (d[k] < 0) continue;
            const int i = c0 + tid * BKT_ITEMS + k;
            const unsigned r = sm.base[d[k]] + bkt_field(excl, d[k]) + bkt_field(run, d[k]);
            perm[r] = i;
            pos[i] = (int32_t)r;
            bkt_add(run, d[k]);
        }
        __syncthreads();
        if (tid < nb) {  // advance the bucket bases by this chunk's totals
            BktWords tot;
#pragma unroll
            for (int k = 0; k < BKT_WORDS; ++k) {
                tot.w[k] = 0ull;
                for (int w = 0; w < BKT_THREADS / 32; ++w) tot.w[k] += sm.warp_tot[w][k];
            }
            sm.base[tid] += bkt_field(tot, tid);
        }
        __syncthreads();
    }
}

}  // namespace die
