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

}  // namespace die
