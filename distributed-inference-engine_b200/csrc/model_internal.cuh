// Internal launchers of the mock-model kernels (shared by model.cu, model_tc.cu and pipeline.cu).
#pragma once
#include <algorithm>

#include "common.cuh"

namespace die {

die_status_t echo_launch(const void* in, void* out, uint64_t bytes, cudaStream_t s);
// n_rows_dev (optional): device-side row count, rows >= *n_rows_dev are not computed.
die_status_t linear_f32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                               const uint32_t* n_rows_dev, cudaStream_t s);
die_status_t linear_tf32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                                const uint32_t* n_rows_dev, cudaStream_t s);
die_status_t mlp_bf16_launch(const void* X, const void* W1, const float* b1, const void* W2, const float* b2,
                             void* hidden, void* Y, int B, int D, int H, cudaStream_t s);

}  // namespace die
