// What the exchange (exchange.cu) needs from the owner pipeline (pipeline.cu) beyond the C-ABI: the launches
// of one step without the host bookkeeping, and the bookkeeping for K steps at once, so that K sharded steps
// can be captured into one CUDA graph.
#pragma once
#include "pipeline_device.cuh"

struct die_pipeline;

namespace die {

// True when K steps of up to B_max requests may be reserved at once on this pipeline's table.
bool pipeline_graph_fits(const die_pipeline* p, int B_max, int K);
// Host side of K owner steps: table maintenance if a bound is near, LRU stamps handed out.
// mirror (optional): arguments of kv_mirror_write for the kernel the caller enqueues before the K steps.
die_status_t pipeline_owner_reserve(die_pipeline* p, int B_max, int K, cudaStream_t s, KvView* view,
                                    unsigned long long* tail0, KvMirrorArgs* mirror);
// Device side of one owner step (no host state touched): digests in, results to the step's out / code.
die_status_t pipeline_owner_launch(die_pipeline* p, const StepRef& sr, uint32_t* digests, int B_max,
                                   const uint32_t* n_req_dev, cudaStream_t s);
int pipeline_launches_per_step(const die_pipeline* p);

// ---- the persistent step kernel (pipe_step.cu) ------------------------------------------------------------
die_status_t pipe_step_grid_alloc(void** out);
die_status_t pipe_step_marks(void* grid, unsigned long long out[8]);
die_status_t pipe_step_cta_marks(void* grid, int n_cta, unsigned long long* out);
// True (and the launch shape) when batches of up to B requests of a K -> N TF32 linear layer with an n_chunks
// victim window fit one co-resident grid.
bool pipe_step_shape(int B, int K, int N, int n_chunks, int* R_out, int* n_req_out);
// n_steps consecutive batches starting at `sr` in one launch.  qv (optional): the queue descriptor, passed by value
// in the launch (stream launches); otherwise sr.q is device-resident or null (values in sr.a).
// lanes (optional, the exchange's owner pass over several steps in one launch): `digests` and `n_req_dev` are rings of
// `lanes` per-step buffers, step s of the launch uses lane (lane0 + s) % lanes.
struct StepLanes {
    int lanes, lane0;
};
die_status_t pipe_step_launch(const StepRef& sr, const QueueDesc* qv, int n_steps, int B, int n_chunks, int K, int N,
                              bool from_keys, const uint32_t* n_req_dev, uint32_t* digests, const float* W,
                              const float* bias, float* arena, void* grid, const KvMirrorArgs* mirror, cudaStream_t s,
                              const StepLanes* lanes = nullptr);
// K owner steps of the exchange (steps first .. first+K-1 of the device-resident descriptor) in ONE persistent launch;
// pipeline_owner_group_supported: false when this pipeline / shape has no persistent kernel (or DIE_XCHG_OWNER_GROUP=0)
// and the caller launches the steps one by one.
bool pipeline_owner_group_supported(const die_pipeline* p, int B_max);
die_status_t pipeline_owner_launch_group(die_pipeline* p, const QueueDesc* qd_dev, int first, int K, uint32_t* digests_ring,
                                         int B_max, const uint32_t* n_dev_ring, const StepLanes& lanes, cudaStream_t s);

}  // namespace die
