// Device side of K1 shared by route.cu and the fused pipeline kernel: CTA-cooperative staging of a
// CTA's contiguous key bytes into shared memory, then one MD5 per lane (md5.cuh).
#pragma once
#include "common.cuh"
#include "md5.cuh"

namespace die {

constexpr int ROUTE_THREADS = 256;
constexpr int ROUTE_STAGE_BYTES = 24 * 1024;
constexpr int ROUTE_SMEM_SHARDS = 1024;

struct SmemKeySrc {
    const uint8_t* base;  // shared-memory address of the key's first byte (any alignment)
    __device__ __forceinline__ uint32_t byte(uint32_t p) const { return base[p]; }
    __device__ __forceinline__ uint32_t word(uint32_t p) const {
        const uintptr_t a = reinterpret_cast<uintptr_t>(base + p);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~uintptr_t(3));
        return __funnelshift_r(w[0], w[1], (uint32_t)(a & 3) * 8);
    }
};

struct GlobalKeySrc {
    const uint8_t* base;
    __device__ __forceinline__ uint32_t byte(uint32_t p) const { return __ldg(base + p); }
    __device__ __forceinline__ uint32_t word(uint32_t p) const {
        return byte(p) | (byte(p + 1) << 8) | (byte(p + 2) << 16) | (byte(p + 3) << 24);
    }
};


// Hash key i (= blockIdx.x * ROUTE_THREADS + threadIdx.x) of a CSR key batch.  All ROUTE_THREADS threads
// of the CTA must call this (it contains __syncthreads); `stage` is ROUTE_STAGE_BYTES + 32 bytes of
// 16-byte aligned shared memory.  Threads with i >= B return an unspecified state.
// cta_hash_keys_at: the same for the ROUTE_THREADS keys starting at k0 (persistent kernels walk the chunks
// themselves); a chunk that starts past the end (k0 >= B) is a no-op for every thread.
__device__ __forceinline__ Md5State cta_hash_keys_at(const uint8_t* __restrict__ key_bytes,
                                                     const int64_t* __restrict__ key_off, int B, int k0, uint8_t* stage) {
    const int tid = threadIdx.x;
    k0 = min(k0, B);
    const int k1 = min(B, k0 + ROUTE_THREADS);
    const int64_t b0 = key_off[k0], b1 = key_off[k1];
    const uintptr_t g0 = reinterpret_cast<uintptr_t>(key_bytes) + (uintptr_t)b0;
    const uintptr_t ga = g0 & ~uintptr_t(15);
    const uint32_t lead = (uint32_t)(g0 - ga);
    const int64_t span = (int64_t)lead + (b1 - b0);
    const bool staged = span <= ROUTE_STAGE_BYTES;
    if (staged && b1 > b0) {
        const int nvec = (int)((span + 15) >> 4);
        const int4* src = reinterpret_cast<const int4*>(ga);
        int4* dst = reinterpret_cast<int4*>(stage);
        for (int v = tid; v < nvec; v += ROUTE_THREADS) dst[v] = ldg_stream(src + v);
    }
    __syncthreads();
    const int i = k0 + tid;
    Md5State st = md5_init();
    if (i < B) {
        const int64_t off = key_off[i];
        const uint32_t len = (uint32_t)(key_off[i + 1] - off);
        if (staged)
            st = md5_key(SmemKeySrc{stage + lead + (uint32_t)(off - b0)}, len);
        else
            st = md5_key(GlobalKeySrc{key_bytes + off}, len);
    }
    return st;
}

__device__ __forceinline__ Md5State cta_hash_keys(const uint8_t* __restrict__ key_bytes,
                                                  const int64_t* __restrict__ key_off, int B, uint8_t* stage) {
    return cta_hash_keys_at(key_bytes, key_off, B, (int)blockIdx.x * ROUTE_THREADS, stage);
}

}  // namespace die
