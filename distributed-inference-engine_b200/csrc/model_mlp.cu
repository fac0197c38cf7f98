// K7: the "mock 4096-hidden bf16 MLP" of BASELINE.json configs[4] on the 5th-generation tensor cores.
//   hidden = bf16( relu(X @ W1^T + b1) ),   Y = bf16( hidden @ W2^T + b2 )        (fp32 accumulation)
// No reference arithmetic exists (the reference's model is an echo, fake_model.py:52); oracle/models.py
// defines the result and the tolerance.
//
// Both layers run the same warp-specialised GEMM, C[M,N] = act(A[M,K] . B[N,K]^T + bias):
//   warp 0 (one lane)  TMA producer: cp.async.bulk.tensor 2-D tiles (A 128x64, B 256x64 bf16, SWIZZLE_128B)
//                      into a 4-stage shared-memory ring, completion on mbarriers (complete_tx)
//   warp 1 (one lane)  MMA issuer: 4 x tcgen05.mma.kind::f16 (M128 N256 K16) per stage, accumulator
//                      128 lanes x 256 columns of TMEM; tcgen05.commit releases the stage / signals the epilogue
//   warps 2-5          epilogue: tcgen05.ld -> +bias, relu, bf16 -> global (64 contiguous bytes per row and chunk: whole
//                      sectors; no shared staging, the operand ring belongs to the next tile by then)
// Persistent CTAs (one per SM) walk the 128x256 output tiles; TMEM holds two accumulators so that the epilogue of one
// tile overlaps the MMAs of the next.
#include <cuda.h>
#include <cuda_bf16.h>

#include "model_internal.cuh"
#include "tc_common.cuh"

namespace die {

constexpr int G_BM = 128, G_BN = 256, G_BK = 64, G_STAGES = 4, G_THREADS = 192;
constexpr int G_A_BYTES = G_BM * G_BK * 2, G_B_BYTES = G_BN * G_BK * 2, G_STAGE_BYTES = G_A_BYTES + G_B_BYTES;

using tc::tma_load_2d;

template <bool RELU>
__global__ void __launch_bounds__(G_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const float* __restrict__ bias, __nv_bfloat16* __restrict__ C, int M, int N, int K) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t full[G_STAGES], empty[G_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_k = (K + G_BK - 1) / G_BK;
    const int tiles_n = (N + G_BN - 1) / G_BN, tiles_m = (M + G_BM - 1) / G_BM;
    const int tiles = tiles_m * tiles_n;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
        for (int s = 0; s < G_STAGES; ++s) {
            tc::mbar_init(&full[s], 1);
            tc::mbar_init(&empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&acc_full[a], 1);
            tc::mbar_init(&acc_empty[a], 4);   // one arrival per epilogue warp
        }
    }
    if (warp == 1) tc::tmem_alloc(&tmem_base, 512);   // two 128 x 256 fp32 accumulators: all of TMEM
    tc::fence_before_thread_sync();
    __syncthreads();
    tc::fence_after_thread_sync();
    const uint32_t taddr = tmem_base;

    // Persistent: CTA b works on tiles b, b + grid, ... (n fastest, so the CTAs of a wave share a few A row-blocks and
    // stream the weights from L2).  The operand ring and the accumulator pair run across tile boundaries: while the
    // epilogue warps drain accumulator i & 1, the MMA warp is already filling the other one with tile i + 1.
    if (warp == 0) {
        if (lane == 0) {  // ---- TMA producer
            uint32_t it = 0;
            for (int t = (int)blockIdx.x; t < tiles; t += (int)gridDim.x) {
                const int m0 = (t / tiles_n) * G_BM, n0 = (t % tiles_n) * G_BN;
                for (int kb = 0; kb < num_k; ++kb, ++it) {
                    const uint32_t s = it % G_STAGES, ph = (it / G_STAGES) & 1u;
                    tc::mbar_wait(&empty[s], ph ^ 1u);  // slot free (passes at once on a fresh barrier)
                    tc::mbar_arrive_expect_tx(&full[s], G_STAGE_BYTES);
                    uint8_t* a = smem + (size_t)s * G_STAGE_BYTES;
                    tma_load_2d(a, &tmA, kb * G_BK, m0, &full[s]);
                    tma_load_2d(a + G_A_BYTES, &tmB, kb * G_BK, n0, &full[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ---- MMA issuer
            const uint32_t idesc = tc::instr_desc(tc::FMT_BF16, G_BM, G_BN);
            uint32_t it = 0, i = 0;
            for (int t = (int)blockIdx.x; t < tiles; t += (int)gridDim.x, ++i) {
                const uint32_t acc = i & 1u;
                tc::mbar_wait(&acc_empty[acc], ((i >> 1) & 1u) ^ 1u);   // the epilogue has drained this accumulator
                tc::fence_after_thread_sync();
                const uint32_t d = taddr + acc * (uint32_t)G_BN;
                for (int kb = 0; kb < num_k; ++kb, ++it) {
                    const uint32_t s = it % G_STAGES, ph = (it / G_STAGES) & 1u;
                    tc::mbar_wait(&full[s], ph);  // both tiles of the stage have landed
                    tc::fence_after_thread_sync();
                    const uint32_t a_s = tc::smem_u32(smem + (size_t)s * G_STAGE_BYTES), b_s = a_s + G_A_BYTES;
#pragma unroll
                    for (int j = 0; j < G_BK / 16; ++j)  // K = 16 bf16 = 32 bytes per instruction
                        tc::mma_f16(d, tc::smem_desc_k_major_sw128(a_s + j * 32), tc::smem_desc_k_major_sw128(b_s + j * 32),
                                    idesc, (kb | j) != 0);
                    tc::mma_commit(&empty[s]);  // the stage is free once these MMAs have read it
                }
                tc::mma_commit(&acc_full[acc]);  // accumulator complete
            }
        }
    } else {  // ---- epilogue warps 2..5: TMEM -> (+bias, relu, bf16) -> global, 64 contiguous bytes per row and chunk
        const int q = warp & 3;  // TMEM lane quarter this warp may read
        const int row = q * 32 + lane;
        uint32_t i = 0;
        for (int t = (int)blockIdx.x; t < tiles; t += (int)gridDim.x, ++i) {
            const int m0 = (t / tiles_n) * G_BM, n0 = (t % tiles_n) * G_BN;
            const uint32_t acc = i & 1u;
            tc::mbar_wait(&acc_full[acc], (i >> 1) & 1u);
            tc::fence_after_thread_sync();
            const int gr = m0 + row;
            for (int c0 = 0; c0 < G_BN; c0 += 32) {
                uint32_t v[32];
                tc::tmem_ld_32x32(taddr + ((uint32_t)(q * 32) << 16) + acc * (uint32_t)G_BN + (uint32_t)c0, v);
                uint32_t packed[16];
#pragma unroll
                for (int j = 0; j < 32; j += 2) {
                    float x0 = __uint_as_float(v[j]), x1 = __uint_as_float(v[j + 1]);
                    const int c = n0 + c0 + j;
                    if (bias) {
                        x0 += c < N ? __ldg(bias + c) : 0.f;
                        x1 += c + 1 < N ? __ldg(bias + c + 1) : 0.f;
                    }
                    if (RELU) {
                        x0 = fmaxf(x0, 0.f);
                        x1 = fmaxf(x1, 0.f);
                    }
                    const __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
                    packed[j >> 1] = *reinterpret_cast<const uint32_t*>(&h);
                }
                if (gr < M) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {  // 4 chunks of 8 bf16 (N is a multiple of 8)
                        const int col = n0 + c0 + j * 8;
                        if (col < N)
                            *reinterpret_cast<uint4*>(C + (size_t)gr * N + col) =
                                make_uint4(packed[4 * j], packed[4 * j + 1], packed[4 * j + 2], packed[4 * j + 3]);
                    }
                }
            }
            tc::fence_before_thread_sync();   // this warp's tcgen05.ld of the accumulator are done
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&acc_empty[acc]);
        }
    }
    tc::fence_before_thread_sync();
    __syncthreads();
    if (warp == 1) {
        tc::fence_after_thread_sync();
        tc::tmem_dealloc(taddr, 512);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static die_status_t encode_fn(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        DIE_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (!p || q != cudaDriverEntryPointSuccess) DIE_FAIL(DIE_ERR_CUDA, "cuTensorMapEncodeTiled is not available");
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    *out = fn;
    return DIE_OK;
}

// [rows, cols] row-major, 2- or 4-byte elements -> boxes of [box_rows, box_cols] whose inner extent is 128 bytes = one
// SWIZZLE_128B atom.  Host-side encode only (no CUDA call): safe inside stream capture.
die_status_t make_tensor_map_2d(void* tm, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint32_t box_rows,
                                uint32_t box_cols) {
    EncodeTiledFn fn;
    DIE_TRY(encode_fn(&fn));
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * (uint64_t)elem_bytes};
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(static_cast<CUtensorMap*>(tm), elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                          2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) DIE_FAIL(DIE_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d", (int)r);
    return DIE_OK;
}
static die_status_t make_map(CUtensorMap* tm, const void* base, uint64_t rows, uint64_t K, uint32_t box_rows) {
    return make_tensor_map_2d(tm, base, 2, rows, K, box_rows, (uint32_t)G_BK);
}

static die_status_t gemm_bf16_launch(const void* A, const void* Bm, const float* bias, void* C, int M, int N, int K,
                                     bool relu, cudaStream_t s) {
    if (K % 8 != 0 || N % 8 != 0)
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "bf16 gemm: K and N must be multiples of 8 (K=%d N=%d)", K, N);
    if ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(Bm) | reinterpret_cast<uintptr_t>(C)) & 15)
        DIE_FAIL(DIE_ERR_UNSUPPORTED, "bf16 gemm: operands must be 16-byte aligned");
    CUtensorMap tmA, tmB;
    DIE_TRY(make_map(&tmA, A, (uint64_t)M, (uint64_t)K, G_BM));
    DIE_TRY(make_map(&tmB, Bm, (uint64_t)N, (uint64_t)K, G_BN));
    const size_t smem = (size_t)G_STAGES * G_STAGE_BYTES + 1024;
    static bool configured = false;
    if (!configured) {
        DIE_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        DIE_CUDA(cudaFuncSetAttribute(gemm_bf16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    const int tiles = ceil_div(N, G_BN) * ceil_div(M, G_BM);
    dim3 grid((unsigned)std::min(tiles, num_sms()));   // persistent CTAs, one per SM
    __nv_bfloat16* c = static_cast<__nv_bfloat16*>(C);
    if (relu)
        DIE_LAUNCH(ST_MODEL, s, gemm_bf16_kernel<true><<<grid, G_THREADS, smem, s>>>(tmA, tmB, bias, c, M, N, K));
    else
        DIE_LAUNCH(ST_MODEL, s, gemm_bf16_kernel<false><<<grid, G_THREADS, smem, s>>>(tmA, tmB, bias, c, M, N, K));
    return DIE_OK;
}

die_status_t mlp_bf16_launch(const void* X, const void* W1, const float* b1, const void* W2, const float* b2,
                             void* hidden, void* Y, int B, int D, int H, cudaStream_t s) {
    DIE_TRY(gemm_bf16_launch(X, W1, b1, hidden, B, H, D, true, s));   // hidden = bf16(relu(X W1^T + b1))
    DIE_TRY(gemm_bf16_launch(hidden, W2, b2, Y, B, D, H, false, s));  // Y = bf16(hidden W2^T + b2)
    return DIE_OK;
}

}  // namespace die
