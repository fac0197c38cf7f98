// K6/K8: mock-model forwards on CUDA cores.  The reference's model is an echo
// (src/mock_models/fake_model.py:52); the linear layer is BASELINE.json configs[1] and has no
// reference arithmetic (oracle/models.py defines it).  The tcgen05 (tensor-core) variants live in
// model_tc.cu; this file holds the exact-fp32 path they are checked against.
#include "common.cuh"
#include "model_internal.cuh"

namespace die {

// ------------------------------------------------------------------------------------------------
// echo: Y = X
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
echo_vec_kernel(const int4* __restrict__ in, int4* __restrict__ out, long long nvec) {
    const long long stride = (long long)gridDim.x * 256;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < nvec; v += 4 * stride) {
        int4 val[4];  // four loads in flight per thread before the first store
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (v + j * stride < nvec) val[j] = ldg_stream(in + v + j * stride);
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (v + j * stride < nvec) stg_stream(out + v + j * stride, val[j]);
    }
}
__global__ void __launch_bounds__(256)
echo_byte_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, long long n) {
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < n; v += (long long)gridDim.x * 256) out[v] = in[v];
}

// ------------------------------------------------------------------------------------------------
// linear, fp32 FMA.  64x64 output tile per CTA, 16x16 threads, 4x4 outputs per thread, BK = 16.
// ------------------------------------------------------------------------------------------------
constexpr int LIN_BM = 64, LIN_BN = 64, LIN_BK = 16;

__global__ void __launch_bounds__(256)
linear_f32_kernel(const float* __restrict__ X, const float* __restrict__ W, const float* __restrict__ bias,
                  float* __restrict__ Y, int B, int K, int N, const uint32_t* __restrict__ n_rows_dev) {
    if (n_rows_dev) B = min(B, (int)*n_rows_dev);
    const int m0 = blockIdx.x * LIN_BM, n0 = blockIdx.y * LIN_BN;
    if (m0 >= B) return;
    __shared__ float xs[LIN_BK][LIN_BM + 4];
    __shared__ float ws[LIN_BK][LIN_BN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += LIN_BK) {
        // 64 rows x 16 k = 1024 values per operand, 4 per thread; consecutive threads walk k
        for (int e = tid; e < LIN_BM * LIN_BK; e += 256) {
            const int r = e / LIN_BK, k = e % LIN_BK;
            const int gm = m0 + r, gn = n0 + r, gk = k0 + k;
            xs[k][r] = (gm < B && gk < K) ? X[(long long)gm * K + gk] : 0.f;
            ws[k][r] = (gn < N && gk < K) ? W[(long long)gn * K + gk] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < LIN_BK; ++k) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = xs[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = ws[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= B) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn < N) Y[(long long)gm * N + gn] = acc[i][j] + (bias ? bias[gn] : 0.f);
        }
    }
}

die_status_t echo_launch(const void* in, void* out, uint64_t bytes, cudaStream_t s) {
    if (bytes == 0) return DIE_OK;
    const bool vec = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    const uint64_t nvec = vec ? bytes / 16 : 0;
    if (nvec) {
        const long long blocks = (long long)((nvec + 255) / 256);
        const int grid = (int)std::min<long long>(blocks, (long long)num_sms() * 16);
        DIE_LAUNCH(ST_MODEL, s, echo_vec_kernel<<<grid, 256, 0, s>>>(static_cast<const int4*>(in),
                                                                     static_cast<int4*>(out), (long long)nvec));
    }
    const uint64_t tail = bytes - nvec * 16;
    if (tail) {
        const long long blocks = (long long)((tail + 255) / 256);
        const int grid = (int)std::min<long long>(blocks, (long long)num_sms() * 16);
        DIE_LAUNCH(ST_MODEL, s, echo_byte_kernel<<<grid, 256, 0, s>>>(static_cast<const uint8_t*>(in) + nvec * 16,
                                                                      static_cast<uint8_t*>(out) + nvec * 16,
                                                                      (long long)tail));
    }
    return DIE_OK;
}

die_status_t linear_f32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                               const uint32_t* n_rows_dev, cudaStream_t s) {
    if (B == 0 || N == 0) return DIE_OK;
    dim3 grid(ceil_div(B, LIN_BM), ceil_div(N, LIN_BN));
    DIE_LAUNCH(ST_MODEL, s, linear_f32_kernel<<<grid, 256, 0, s>>>(X, W, b, Y, B, K, N, n_rows_dev));
    return DIE_OK;
}

}  // namespace die

using namespace die;

extern "C" die_status_t die_echo_fwd(const void* in, void* out, uint64_t bytes, die_stream_t stream) {
    if (bytes && (!in || !out)) DIE_FAIL(DIE_ERR_INVALID, "die_echo_fwd: null buffer");
    return echo_launch(in, out, bytes, static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_linear_fwd_f32(const float* X, const float* W, const float* b, float* Y, int32_t B,
                                           int32_t K, int32_t N, int32_t precision, die_stream_t stream) {
    if (B < 0 || K <= 0 || N <= 0) DIE_FAIL(DIE_ERR_INVALID, "die_linear_fwd_f32: bad sizes");
    if (B == 0) return DIE_OK;
    if (!X || !W || !Y) DIE_FAIL(DIE_ERR_INVALID, "die_linear_fwd_f32: null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (precision == DIE_LINEAR_FP32) return linear_f32_launch(X, W, b, Y, B, K, N, nullptr, s);
    if (precision == DIE_LINEAR_TF32) return linear_tf32_launch(X, W, b, Y, B, K, N, nullptr, s);
    DIE_FAIL(DIE_ERR_INVALID, "die_linear_fwd_f32: unknown precision %d", precision);
}

extern "C" die_status_t die_mlp_fwd_bf16(const void* X, const void* W1, const float* b1, const void* W2,
                                         const float* b2, void* hidden, void* Y, int32_t B, int32_t D, int32_t H,
                                         die_stream_t stream) {
    if (B < 0 || D <= 0 || H <= 0) DIE_FAIL(DIE_ERR_INVALID, "die_mlp_fwd_bf16: bad sizes");
    if (B == 0) return DIE_OK;
    if (!X || !W1 || !W2 || !hidden || !Y) DIE_FAIL(DIE_ERR_INVALID, "die_mlp_fwd_bf16: null buffer");
    return mlp_bf16_launch(X, W1, b1, W2, b2, hidden, Y, B, D, H, static_cast<cudaStream_t>(stream));
}
