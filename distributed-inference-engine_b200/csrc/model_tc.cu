// K6/K7 on the 5th-generation tensor cores (tcgen05 + TMEM).  Placeholder launchers until the
// kernels land: they fail loudly instead of falling back to another path.
#include "model_internal.cuh"

namespace die {

die_status_t linear_tf32_launch(const float*, const float*, const float*, float*, int, int, int,
                                const uint32_t*, cudaStream_t) {
    DIE_FAIL(DIE_ERR_UNSUPPORTED, "linear tf32 (tcgen05) kernel is not built yet");
}

die_status_t mlp_bf16_launch(const void*, const void*, const float*, const void*, const float*, void*, void*, int,
                             int, int, cudaStream_t) {
    DIE_FAIL(DIE_ERR_UNSUPPORTED, "bf16 MLP (tcgen05) kernel is not built yet");
}

}  // namespace die
