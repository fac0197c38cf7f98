// K6 on the 5th-generation tensor cores: Y[B,N] = X[B,K] @ W[N,K]^T + b with TF32 inputs and fp32
// accumulation in TMEM (tcgen05.mma.kind::tf32).  BASELINE.json configs[1] "mock linear model"; no
// reference arithmetic exists (fake_model.py:52 is an echo), oracle/models.py defines the result.
//
// One CTA = one 128-row tile of X against the whole W (N <= 256 columns, K <= a few hundred):
//   - all 8 warps copy the X tile and W into shared memory with coalesced 128-bit loads, storing
//     them in the canonical K-major SWIZZLE_128B layout (blocks of [rows x 32 tf32]);
//   - one thread issues K/8 tcgen05.mma (M=128, N, K=8) and commits them to an mbarrier;
//   - warps 0-3 read the fp32 accumulators with tcgen05.ld (32 lanes x 32 columns), add the bias
//     and stage the tile in shared memory; all warps then write whole rows with 512-byte coalesced
//     stores.  The layer is HBM-bound (32 flop/B), so the tensor pipe only has to stay out of the way.
#include <cuda.h>

#include "model_internal.cuh"
#include "tc_common.cuh"

namespace die {

constexpr int TC_M = 128;
constexpr int TC_THREADS = 256;

__device__ __forceinline__ uint32_t pow2_cols(uint32_t n) {
    uint32_t c = 32;
    while (c < n) c <<= 1;
    return c;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(tc::smem_u32(smem_dst)), "l"(gmem_src),
                 "r"(src_bytes)
                 : "memory");
}

// grid = (row tiles, column splits): CTA (x, y) computes rows [128x, 128x+128) x columns [y*Nc, (y+1)*Nc).
// SCATTER (pipeline): row `rank` of X is the rank-th miss of the batch; its result goes straight to
// out[request] and to the cached row of the op that owns the table slot (UNPACK, batcher.py:225-227, fused
// into the GEMM epilogue), and CTAs with blockIdx.x >= tiles evict the batch's LRU victims.
template <bool SCATTER>
__global__ void __launch_bounds__(TC_THREADS, 1)
linear_tf32_kernel(const float* __restrict__ X, const float* __restrict__ W, const float* __restrict__ bias,
                   float* __restrict__ Y, int B, int K, int N, int Nc, const uint32_t* __restrict__ n_rows_dev,
                   TcScatter sc) {
    StepArgs step;
    if (SCATTER) {
        step = step_resolve(sc.sr);
        trace_mark(sc.sr, TR_MODEL);
        sc.pb.code = step.code;
    }
    if (SCATTER && (int)blockIdx.x >= sc.tiles) {  // eviction CTAs (kvstore.py:82-93), 256 window records each
        pdl_wait();
        pipe_evict_cta(step.kv, sc.pb, (int)((blockIdx.x - sc.tiles) * gridDim.y + blockIdx.y), step.t_set);
        return;
    }
    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B atoms must start on 1024-byte boundaries of the shared window
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int m0 = blockIdx.x * TC_M, n0 = blockIdx.y * Nc;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int KB = K / 32;             // 128-byte (32 x tf32) blocks along K
    const int chunks_per_row = K / 4;  // 16-byte chunks per operand row
    uint8_t* a_base = smem;                               // KB blocks of [128 rows x 128 B]
    uint8_t* b_base = smem + (size_t)KB * TC_M * 128;     // KB blocks of [Nc rows x 128 B]
    const uint32_t ncols = pow2_cols((uint32_t)Nc);

    if (warp == 0) tc::tmem_alloc(&tmem_base, ncols);
    if (tid == 32) tc::mbar_init(&bar, 1);

    // operands -> shared memory with cp.async (LDGSTS): a warp owns a row, lanes own its 16-byte chunks, so
    // every request is one contiguous 512-byte row segment and nothing is staged in registers
    // the weights do not depend on the step: their tile is fetched while the previous kernel is still draining
    for (int r = warp; r < Nc; r += TC_THREADS / 32) {
        const float* src = W + (size_t)(n0 + r) * K;
        for (int ck = lane; ck < chunks_per_row; ck += 32)
            cp_async16(b_base + (size_t)(ck >> 3) * ((size_t)Nc * 128) + tc::sw128_offset(r, ck & 7), src + ck * 4, 16u);
    }
    pdl_wait();  // the packed rows and their count come from the previous kernels
    if (n_rows_dev) B = min(B, (int)*n_rows_dev);
    const bool idle = m0 >= B;  // uniform for the CTA: nothing to compute, but the TMEM columns must be returned
    for (int r = warp; r < TC_M && !idle; r += TC_THREADS / 32) {
        const bool live = m0 + r < B;
        const float* src = X + (size_t)(live ? m0 + r : 0) * K;
        for (int ck = lane; ck < chunks_per_row; ck += 32)
            cp_async16(a_base + (size_t)(ck >> 3) * (TC_M * 128) + tc::sw128_offset(r, ck & 7), src + ck * 4,
                       live ? 16u : 0u);  // src-size 0 zero-fills rows past the end of the batch
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t taddr = tmem_base;
    if (idle) {  // tile past the end of the batch (its size is only known on the device)
        if (warp == 0) tc::tmem_dealloc(taddr, ncols);
        return;
    }

    if (tid == 0) {
        const uint32_t idesc = tc::instr_desc(tc::FMT_TF32, TC_M, (uint32_t)Nc);
        const uint32_t a_s = tc::smem_u32(a_base), b_s = tc::smem_u32(b_base);
        for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {  // 4 x (K = 8 tf32 = 32 bytes) per 128-byte block
                const uint64_t ad = tc::smem_desc_k_major_sw128(a_s + kb * (TC_M * 128) + j * 32);
                const uint64_t bd = tc::smem_desc_k_major_sw128(b_s + kb * (Nc * 128) + j * 32);
                tc::mma_tf32(taddr, ad, bd, idesc, (kb | j) != 0);
            }
        }
        tc::mma_commit(&bar);
    }
    tc::mbar_wait(&bar, 0);
    tc::fence_after_thread_sync();

    // epilogue: TMEM -> registers (+bias) -> shared (row pitch Nc+4 floats) -> coalesced row segments
    float* stage = reinterpret_cast<float*>(smem);
    const int pitch = Nc + 4;
    if (warp < 4) {
        const int row = warp * 32 + lane;
        for (int c0 = 0; c0 < Nc; c0 += 32) {
            uint32_t v[32];
            tc::tmem_ld_32x32(taddr + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                if (c0 + j < Nc) {
                    float4 o;
                    o.x = __uint_as_float(v[j + 0]) + (bias ? bias[n0 + c0 + j + 0] : 0.f);
                    o.y = __uint_as_float(v[j + 1]) + (bias ? bias[n0 + c0 + j + 1] : 0.f);
                    o.z = __uint_as_float(v[j + 2]) + (bias ? bias[n0 + c0 + j + 2] : 0.f);
                    o.w = __uint_as_float(v[j + 3]) + (bias ? bias[n0 + c0 + j + 3] : 0.f);
                    *reinterpret_cast<float4*>(stage + (size_t)row * pitch + c0 + j) = o;
                }
            }
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    const int rows = min(TC_M, B - m0);
    const int vec_per_row = Nc / 4;
    const bool any_dup = SCATTER ? (*sc.any_dup != 0u) : false;
    for (int e = tid; e < rows * vec_per_row; e += TC_THREADS) {
        const int r = e / vec_per_row, c = e - r * vec_per_row;
        const int4 o = *reinterpret_cast<const int4*>(stage + (size_t)r * pitch + c * 4);
        if (!SCATTER) {
            stg_stream(reinterpret_cast<int4*>(Y + (size_t)(m0 + r) * N + n0) + c, o);
        } else {
            const int q = sc.miss_req[m0 + r];
            stg_stream(reinterpret_cast<int4*>(step.out + (size_t)q * N + n0) + c, o);
            if (c == 0 && blockIdx.y == 0) step.code[q] = 0;  // callers see 1 = served from the cache, 0 = computed
            long long row = sc.set_row[q];
            if (any_dup) {  // same key twice in the batch: only the last op's result is cached
                const Slot* sl = &step.kv.slots[sc.set_slot[q]];
                row = sl->stamp == step.t_set + (unsigned long long)q ? (long long)sl->value : -1;
            }
            if (row >= 0) stg_stream(reinterpret_cast<int4*>(sc.arena + (size_t)row * N + n0) + c, o);
        }
    }
    if (warp == 0) {
        tc::fence_after_thread_sync();
        tc::tmem_dealloc(taddr, ncols);
    }
}

// Streaming form for large batches (the layer is HBM-bound: 32 flop per byte): persistent CTAs, one per SM, keep W in
// shared memory for their whole life and walk the 128-row tiles of X with a two-deep cp.async pipeline -- tile t+1
// travels while tile t is multiplied (tcgen05, accumulators in TMEM) and written back.  The consumed A buffer doubles
// as the staging area of the epilogue (XOR-swizzled 16-byte chunks, so neither the row-per-lane writes out of TMEM nor
// the row-per-warp reads for the 512-byte coalesced stores conflict on banks); all eight warps drain TMEM (warps w and
// w+4 share a lane quarter and split its columns).
__global__ void __launch_bounds__(TC_THREADS, 1)
linear_tf32_stream_kernel(const float* __restrict__ X, const float* __restrict__ W, const float* __restrict__ bias,
                          float* __restrict__ Y, int B, int K, int N, const uint32_t* __restrict__ n_rows_dev) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base;
    __shared__ __align__(16) float s_bias[256];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int KB = K / 32, chunks_per_row = K / 4;
    uint8_t* b_base = smem;                                    // W: KB blocks of [N rows x 128 B]
    uint8_t* a_base0 = smem + (size_t)N * K * 4;               // two A tiles of KB blocks of [128 rows x 128 B]
    const size_t a_bytes = (size_t)TC_M * K * 4;
    const uint32_t ncols = pow2_cols((uint32_t)N);
    if (n_rows_dev) B = min(B, (int)*n_rows_dev);
    const int tiles = (B + TC_M - 1) / TC_M;
    if (warp == 0) tc::tmem_alloc(&tmem_base, ncols);
    if (tid == 32) tc::mbar_init(&bar, 1);
    if (tid < N) s_bias[tid] = bias ? bias[tid] : 0.f;
    for (int e = tid; e < N * chunks_per_row; e += TC_THREADS) {
        const int r = e / chunks_per_row, ck = e - r * chunks_per_row;
        cp_async16(b_base + (size_t)(ck >> 3) * ((size_t)N * 128) + tc::sw128_offset(r, ck & 7), W + (size_t)r * K + ck * 4, 16u);
    }
    auto load_tile = [&](int t, uint8_t* a_base) {
        const int m0 = t * TC_M;
        for (int e = tid; e < TC_M * chunks_per_row; e += TC_THREADS) {
            const int r = e / chunks_per_row, ck = e - r * chunks_per_row;
            const bool live = m0 + r < B;
            cp_async16(a_base + (size_t)(ck >> 3) * (TC_M * 128) + tc::sw128_offset(r, ck & 7),
                       X + (size_t)(live ? m0 + r : 0) * K + ck * 4, live ? 16u : 0u);
        }
    };
    int t = (int)blockIdx.x;
    if (t < tiles) load_tile(t, a_base0);
    asm volatile("cp.async.commit_group;" ::: "memory");
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t taddr = tmem_base;
    const int row_pitch = N * 4, vec_per_row = N / 4;  // N <= K here, so the staged rows fit the tile
    const int col_half = (N / 2) % 32 == 0 ? N / 2 : N;  // columns per warp of a lane quarter (both warps, or warp < 4 only)
    unsigned phase = 0;
    for (int it = 0; t < tiles; t += (int)gridDim.x, ++it) {
        uint8_t* a_cur = a_base0 + (size_t)(it & 1) * a_bytes;
        const int t_next = t + (int)gridDim.x;
        if (t_next < tiles) load_tile(t_next, a_base0 + (size_t)((it + 1) & 1) * a_bytes);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");  // tile t (and W) have landed
        tc::fence_proxy_async_smem();
        tc::fence_before_thread_sync();
        __syncthreads();
        tc::fence_after_thread_sync();
        if (tid == 0) {
            const uint32_t idesc = tc::instr_desc(tc::FMT_TF32, TC_M, (uint32_t)N);
            const uint32_t a_s = tc::smem_u32(a_cur), b_s = tc::smem_u32(b_base);
            for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    tc::mma_tf32(taddr, tc::smem_desc_k_major_sw128(a_s + kb * (TC_M * 128) + j * 32),
                                 tc::smem_desc_k_major_sw128(b_s + kb * (N * 128) + j * 32), idesc, (kb | j) != 0);
            }
            tc::mma_commit(&bar);
        }
        tc::mbar_wait(&bar, phase & 1u);
        ++phase;
        tc::fence_after_thread_sync();
        // epilogue: TMEM -> registers (+bias) -> the consumed A tile (row r, 16-byte chunk c at c ^ (r & 7)) -> global
        uint8_t* stage = a_cur;
        {
            const int q = warp & 3, row = q * 32 + lane;
            const int c_lo = col_half == N ? 0 : (warp >> 2) * col_half;
            const int c_hi = col_half == N ? (warp < 4 ? N : 0) : c_lo + col_half;
            for (int c0 = c_lo; c0 < c_hi; c0 += 32) {
                uint32_t v[32];
                tc::tmem_ld_32x32(taddr + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    if (c0 + j < N) {
                        const float4 bv = *reinterpret_cast<const float4*>(s_bias + c0 + j);
                        float4 o;
                        o.x = __uint_as_float(v[j + 0]) + bv.x;
                        o.y = __uint_as_float(v[j + 1]) + bv.y;
                        o.z = __uint_as_float(v[j + 2]) + bv.z;
                        o.w = __uint_as_float(v[j + 3]) + bv.w;
                        const int chunk = (c0 + j) >> 2;
                        *reinterpret_cast<float4*>(stage + (size_t)row * row_pitch + (size_t)((chunk ^ (row & 7)) << 4)) = o;
                    }
                }
            }
        }
        tc::fence_before_thread_sync();
        __syncthreads();
        const int m0 = t * TC_M;
        const int rows = min(TC_M, B - m0);
        for (int e = tid; e < rows * vec_per_row; e += TC_THREADS) {
            const int r = e / vec_per_row, c = e - r * vec_per_row;
            const int4 o = *reinterpret_cast<const int4*>(stage + (size_t)r * row_pitch + (size_t)((c ^ (r & 7)) << 4));
            stg_stream(reinterpret_cast<int4*>(Y + (size_t)(m0 + r) * N) + c, o);
        }
        __syncthreads();  // the staging reads are done: the next iteration's prefetch may overwrite this buffer
    }
    if (warp == 0) {
        tc::fence_after_thread_sync();
        tc::tmem_dealloc(taddr, ncols);
    }
}

// The same persistent layer with TMA on both sides: one thread asks for the next A tile as K/32 boxes of
// [128 rows x 32 tf32] (cp.async.bulk.tensor, SWIZZLE_128B: exactly the K-major layout the MMA reads) and, after the
// eight warps have drained the accumulator into the consumed A tile in that same layout, hands the tile back to the
// TMA unit as N/32 boxes of the output.  No thread touches global memory for X or Y; what the 256 threads do per tile
// is tcgen05.ld + st.shared.  Rows past B are zero-filled on the way in and clipped on the way out by the tensor maps.
__global__ void __launch_bounds__(TC_THREADS, 1)
linear_tf32_tma_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmY,
                       const float* __restrict__ W, const float* __restrict__ bias, int B, int K, int N) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t bar, full[2];
    __shared__ uint32_t tmem_base;
    __shared__ __align__(16) float s_bias[256];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int KB = K / 32, chunks_per_row = K / 4;
    uint8_t* b_base = smem;
    uint8_t* a_base0 = smem + (size_t)N * K * 4;
    const size_t a_bytes = (size_t)TC_M * K * 4;
    const uint32_t ncols = pow2_cols((uint32_t)N);
    const int tiles = (B + TC_M - 1) / TC_M;
    if (warp == 0) tc::tmem_alloc(&tmem_base, ncols);
    if (tid == 32) {
        tc::mbar_init(&bar, 1);
        tc::mbar_init(&full[0], 1);
        tc::mbar_init(&full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < N) s_bias[tid] = bias ? bias[tid] : 0.f;
    for (int e = tid; e < N * chunks_per_row; e += TC_THREADS) {
        const int r = e / chunks_per_row, ck = e - r * chunks_per_row;
        cp_async16(b_base + (size_t)(ck >> 3) * ((size_t)N * 128) + tc::sw128_offset(r, ck & 7), W + (size_t)r * K + ck * 4, 16u);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    tc::fence_proxy_async_smem();
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    auto load_tile = [&](int t, int buf) {  // one thread
        tc::mbar_arrive_expect_tx(&full[buf], (uint32_t)a_bytes);
        uint8_t* a = a_base0 + (size_t)buf * a_bytes;
        for (int kb = 0; kb < KB; ++kb) tc::tma_load_2d(a + (size_t)kb * (TC_M * 128), &tmX, kb * 32, t * TC_M, &full[buf]);
    };
    int t = (int)blockIdx.x;
    if (tid == 0 && t < tiles) load_tile(t, 0);
    const uint32_t taddr = tmem_base;
    const int col_half = (N / 2) % 32 == 0 ? N / 2 : N;
    unsigned phase = 0;
    for (int it = 0; t < tiles; t += (int)gridDim.x, ++it) {
        const int buf = it & 1;
        uint8_t* a_cur = a_base0 + (size_t)buf * a_bytes;
        const int t_next = t + (int)gridDim.x;
        if (tid == 0) {
            if (t_next < tiles) {
                tc::bulk_wait_read_all();   // the other buffer's tile (stored one iteration ago) has left shared memory
                load_tile(t_next, buf ^ 1);
            }
            tc::mbar_wait(&full[buf], (unsigned)(it >> 1) & 1u);
            tc::fence_after_thread_sync();
            const uint32_t idesc = tc::instr_desc(tc::FMT_TF32, TC_M, (uint32_t)N);
            const uint32_t a_s = tc::smem_u32(a_cur), b_s = tc::smem_u32(b_base);
            for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    tc::mma_tf32(taddr, tc::smem_desc_k_major_sw128(a_s + kb * (TC_M * 128) + j * 32),
                                 tc::smem_desc_k_major_sw128(b_s + kb * (N * 128) + j * 32), idesc, (kb | j) != 0);
            }
            tc::mma_commit(&bar);
        }
        tc::mbar_wait(&bar, phase & 1u);
        ++phase;
        tc::fence_after_thread_sync();
        {   // accumulator (+bias) -> the consumed A tile, as N/32 blocks of [128 rows x 128 B], SWIZZLE_128B
            const int q = warp & 3, row = q * 32 + lane;
            const int c_lo = col_half == N ? 0 : (warp >> 2) * col_half;
            const int c_hi = col_half == N ? (warp < 4 ? N : 0) : c_lo + col_half;
            for (int c0 = c_lo; c0 < c_hi; c0 += 32) {
                uint32_t v[32];
                tc::tmem_ld_32x32(taddr + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
                uint8_t* blk = a_cur + (size_t)(c0 >> 5) * (TC_M * 128);
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const float4 bv = *reinterpret_cast<const float4*>(s_bias + c0 + j);
                    float4 o;
                    o.x = __uint_as_float(v[j + 0]) + bv.x;
                    o.y = __uint_as_float(v[j + 1]) + bv.y;
                    o.z = __uint_as_float(v[j + 2]) + bv.z;
                    o.w = __uint_as_float(v[j + 3]) + bv.w;
                    *reinterpret_cast<float4*>(blk + tc::sw128_offset((uint32_t)row, (uint32_t)(j >> 2))) = o;
                }
            }
        }
        tc::fence_proxy_async_smem();      // the staged tile is read by the TMA unit
        tc::fence_before_thread_sync();    // ... and the accumulator is free for the next tile's MMA
        __syncthreads();
        if (tid == 0) {
            for (int cg = 0; cg < N / 32; ++cg) tc::tma_store_2d(&tmY, a_cur + (size_t)cg * (TC_M * 128), cg * 32, t * TC_M);
            tc::bulk_commit();
        }
    }
    if (tid == 0) tc::bulk_wait_all();
    __syncthreads();
    if (warp == 0) {
        tc::fence_after_thread_sync();
        tc::tmem_dealloc(taddr, ncols);
    }
}

static die_status_t tf32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                                const uint32_t* n_rows_dev, const TcScatter* scatter, cudaStream_t s) {
    if (B == 0) return DIE_OK;
    if (K % 32 != 0 || N % 16 != 0 || N < 16 || N > 256)
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "linear tf32: needs K %% 32 == 0 and N in [16,256] multiple of 16 (K=%d N=%d)", K, N);
    if ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(Y)) & 15)
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "linear tf32: X, W and Y must be 16-byte aligned");
    const int tiles = ceil_div(B, TC_M);
    {   // large batches: the persistent streaming form (N a multiple of 32 chunks it can swizzle, rows fit the A tile)
        const size_t smem_s = (size_t)N * K * 4 + 2 * (size_t)TC_M * K * 4 + 1024;
        if (!scatter && tiles >= 2 * num_sms() && smem_s <= 200 * 1024 && N <= K && N % 32 == 0) {
            static int use_tma = -1;
            if (use_tma < 0) use_tma = getenv("DIE_TF32_TMA") ? atoi(getenv("DIE_TF32_TMA")) : 1;
            if (use_tma && !n_rows_dev) {  // exact row count on the host: the tensor maps clip the last tile
                CUtensorMap tmX, tmY;
                DIE_TRY(make_tensor_map_2d(&tmX, X, 4, (uint64_t)B, (uint64_t)K, TC_M, 32));
                DIE_TRY(make_tensor_map_2d(&tmY, Y, 4, (uint64_t)B, (uint64_t)N, TC_M, 32));
                static bool configured_t = false;
                if (!configured_t) {
                    DIE_CUDA(cudaFuncSetAttribute(linear_tf32_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                    configured_t = true;
                }
                DIE_LAUNCH(ST_MODEL, s, linear_tf32_tma_kernel<<<num_sms(), TC_THREADS, smem_s, s>>>(tmX, tmY, W, b, B, K, N));
                return DIE_OK;
            }
            static bool configured_s = false;
            if (!configured_s) {
                DIE_CUDA(cudaFuncSetAttribute(linear_tf32_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                configured_s = true;
            }
            DIE_LAUNCH(ST_MODEL, s, linear_tf32_stream_kernel<<<num_sms(), TC_THREADS, smem_s, s>>>(X, W, b, Y, B, K, N, n_rows_dev));
            return DIE_OK;
        }
    }
    // split the columns over CTAs until the grid covers the SMs (the layer is bandwidth/latency bound:
    // more CTAs = more loads in flight; X tiles are re-read from L2, not from HBM)
    int n_split = 1;
    while (tiles * n_split * 2 <= num_sms() && (N / (n_split * 2)) >= 32 && (N / (n_split * 2)) % 16 == 0 &&
           N % (n_split * 2) == 0)
        n_split *= 2;
    const int Nc = N / n_split;
    const size_t operands = (size_t)(TC_M + Nc) * K * sizeof(float);
    const size_t staging = (size_t)TC_M * (Nc + 4) * sizeof(float);
    const size_t smem = std::max(operands, staging) + 1024;  // + slack for the 1024-byte alignment
    if (smem > 200 * 1024)
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "linear tf32: (128+N)*K*4 = %zu bytes of operands exceed shared memory", smem);
    static size_t configured[2] = {0, 0};
    const int which = scatter ? 1 : 0;
    if (smem > configured[which]) {
        if (scatter)
            DIE_CUDA(cudaFuncSetAttribute(linear_tf32_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        else
            DIE_CUDA(cudaFuncSetAttribute(linear_tf32_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[which] = smem;
    }
    if (scatter) {
        TcScatter sc = *scatter;
        sc.tiles = tiles;
        const int evict_ctas = sc.pb.n_chunks * (WCHUNK / PT);
        dim3 grid(tiles + ceil_div(evict_ctas, n_split), n_split);
        DIE_LAUNCH(ST_MODEL, s, launch_pdl(linear_tf32_kernel<true>, grid, dim3(TC_THREADS), smem, s, X, W, b, Y, B, K, N, Nc,
                                           n_rows_dev, sc));
    } else {
        dim3 grid(tiles, n_split);
        DIE_LAUNCH(ST_MODEL, s, linear_tf32_kernel<false><<<grid, TC_THREADS, smem, s>>>(X, W, b, Y, B, K, N, Nc, n_rows_dev, TcScatter{}));
    }
    return DIE_OK;
}

die_status_t linear_tf32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                                const uint32_t* n_rows_dev, cudaStream_t s) {
    return tf32_launch(X, W, b, Y, B, K, N, n_rows_dev, nullptr, s);
}

die_status_t linear_tf32_scatter_launch(const float* X, const float* W, const float* b, int B, int K, int N,
                                        const uint32_t* n_rows_dev, const TcScatter& sc, cudaStream_t s) {
    return tf32_launch(X, W, b, nullptr, B, K, N, n_rows_dev, &sc, s);  // results go to the step's `out`
}

}  // namespace die
