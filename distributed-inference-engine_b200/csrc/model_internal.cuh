// Internal launchers of the mock-model kernels (shared by model.cu, model_tc.cu and pipeline.cu).
#pragma once
#include <algorithm>

#include "pipeline_device.cuh"

namespace die {

die_status_t echo_launch(const void* in, void* out, uint64_t bytes, cudaStream_t s);
// n_rows_dev (optional): device-side row count, rows >= *n_rows_dev are not computed.
die_status_t linear_f32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                               const uint32_t* n_rows_dev, cudaStream_t s);
die_status_t linear_tf32_launch(const float* X, const float* W, const float* b, float* Y, int B, int K, int N,
                                const uint32_t* n_rows_dev, cudaStream_t s);
// Fused pipeline epilogue of the TF32 linear kernel (see model_tc.cu): scatter results to request order and to
// the cached rows, and evict the batch's LRU victims in the same launch.
struct TcScatter {
    StepRef sr;                // table view, stamps, out and hit codes of the step
    const int32_t* miss_req;   // rank -> request index
    const int32_t* set_row;    // request -> cached-row handle (-1: ask the slot)
    const int32_t* set_slot;   // request -> table slot
    PipeBufs pb;               // victim window + control block of the batch
    const unsigned int* any_dup;
    float* arena;              // [n_rows, N]
    int tiles;
};
die_status_t linear_tf32_scatter_launch(const float* X, const float* W, const float* b, int B, int K, int N,
                                        const uint32_t* n_rows_dev, const TcScatter& sc, cudaStream_t s);
die_status_t make_tensor_map_2d(void* tm /*CUtensorMap*/, const void* base, int elem_bytes, uint64_t rows, uint64_t cols,
                                uint32_t box_rows, uint32_t box_cols);
die_status_t mlp_bf16_launch(const void* X, const void* W1, const float* b1, const void* W2, const float* b2,
                             void* hidden, void* Y, int B, int D, int H, cudaStream_t s);

}  // namespace die
