// Composed hot path for one owner shard: route -> get -> pack misses -> model -> unpack -> set.
// This is SURVEY.md call-stack (E) -- the coordinator the reference's README.md:56-60 describes
// but never wrote -- built from the same kernels the stand-alone C-ABI entry points launch.
//
// Per flushed batch of B fixed-width requests (all device-side, no host sync):
//   1 route_md5_kernel        keys -> 128-bit digests                       (model_registry.py:159)
//   2 kv_get_kernel           hit / miss / expired, LRU stamps, row handles (kvstore.py:136-164)
//   3 miss_scan_kernel        rank of every miss in request order           (batcher.py:211 order)
//   4 pack_hits_kernel        misses: payload -> packed[rank];  hits: result row -> out
//   5 model forward           packed -> packed_out (echo / linear)          (fake_model.py:52)
//   6 kv set (3 kernels)      misses inserted in request order, LRU evictions (kvstore.py:104-134)
//   7 unpack_store_kernel     packed_out[rank] -> out[q] and -> the cached result row of the
//                             op that owns the slot (last duplicate wins)   (batcher.py:225-227)
// Cached values are result rows in `arena`; the table hands out row handles (kv_internal.cuh).
#include "kv_internal.cuh"
#include "model_internal.cuh"

struct die_pipeline {
    die_pipeline_config_t cfg;
    die_kv* kv;
    uint64_t n_rows;
    float* arena;     // [n_rows, dim_out]
    float* W;         // [dim_out, dim_in]
    float* bias;      // [dim_out]
    uint32_t* digests;  // [B,4]
    int32_t* shard;     // [B]
    uint8_t* code;      // [B]
    uint64_t* value;    // [B]
    int32_t* miss_rank; // [B]
    uint32_t* n_miss;   // [1]
    int32_t* set_slot;  // [B]
    float* packed_in;   // [B, dim_in]
    float* packed_out;  // [B, dim_out]
    // staging for the host-buffer entry point
    uint8_t* d_key_bytes;
    int64_t* d_key_off;
    float* d_payload;
    float* d_out;
    uint8_t* d_hit;
};

namespace die {

// rank[q] = number of misses before q (request order) for misses, -1 for hits; *n_miss = total.
__global__ void __launch_bounds__(1024)
miss_scan_kernel(const uint8_t* __restrict__ code, int B, int32_t* __restrict__ rank, uint32_t* __restrict__ n_miss) {
    __shared__ unsigned s_warp[32];
    unsigned base = 0;
    for (int pos = 0; pos < B; pos += 1024 * 4) {
        const int q0 = pos + threadIdx.x * 4;
        bool miss[4];
        unsigned cnt = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            miss[k] = q0 + k < B && code[q0 + k] != 1;
            cnt += miss[k];
        }
        unsigned excl;
        const unsigned total = block_scan_1024(cnt, excl, s_warp);
        unsigned r = base + excl;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (q0 + k < B) rank[q0 + k] = miss[k] ? (int32_t)r++ : -1;
        base += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_miss = base;
}

// misses: packed_in[rank[q]] = payload[q]   (PACK, batcher.py:211)
// hits  : out[q] = arena[value[q]]          (the cached result)
__global__ void __launch_bounds__(256)
pack_hits_kernel(const int4* __restrict__ payload, const uint8_t* __restrict__ code,
                 const int32_t* __restrict__ rank, const unsigned long long* __restrict__ value,
                 const int4* __restrict__ arena, long long B, int vin, int vout, int4* __restrict__ packed_in,
                 int4* __restrict__ out) {
    const int vmax = vin > vout ? vin : vout;
    const long long total = B * vmax;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long q = v / vmax;
        const int c = (int)(v - q * vmax);
        if (code[q] == 1) {
            if (c < vout) stg_stream(out + q * vout + c, ldg_stream(arena + (long long)value[q] * vout + c));
        } else if (c < vin) {
            stg_stream(packed_in + (long long)rank[q] * vin + c, ldg_stream(payload + q * vin + c));
        }
    }
}

__global__ void __launch_bounds__(256)
echo_rows_kernel(const int4* __restrict__ in, int4* __restrict__ out, int vrow, const uint32_t* __restrict__ n_rows) {
    const long long total = (long long)*n_rows * vrow;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256)
        stg_stream(out + v, ldg_stream(in + v));
}

// UNPACK (batcher.py:225-227): result row of miss q goes to out[q]; the op that owns the table
// slot after the batch (stamp == t_set + q, i.e. the last duplicate) also stores it in its row.
__global__ void __launch_bounds__(256)
unpack_store_kernel(KvView kv, const int4* __restrict__ packed_out, uint8_t* __restrict__ code,
                    const int32_t* __restrict__ rank, const int32_t* __restrict__ set_slot, unsigned long long t_set,
                    long long B, int vout, int4* __restrict__ out, int4* __restrict__ arena) {
    const long long total = B * vout;
    for (long long v = blockIdx.x * 256ll + threadIdx.x; v < total; v += (long long)gridDim.x * 256) {
        const long long q = v / vout;
        const int c = (int)(v - q * vout);
        if (code[q] == 1) continue;
        const int4 y = ldg_stream(packed_out + (long long)rank[q] * vout + c);
        stg_stream(out + v, y);
        if (c == 0) code[q] = 0;  // callers see 1 = served from the cache, 0 = computed
        const int32_t s = set_slot[q];
        if (s >= 0) {
            const Slot* sl = &kv.slots[s];
            if (sl->stamp == t_set + (unsigned long long)q) stg_stream(arena + (long long)sl->value * vout + c, y);
        }
    }
}

static int grid_256(long long items) {
    const long long blocks = (items + 255) / 256;
    const long long cap = (long long)num_sms() * 16;
    return (int)(blocks < 1 ? 1 : (blocks < cap ? blocks : cap));
}

static die_status_t run_owner(die_pipeline* p, const uint32_t* digests, int B, const float* payload, double now,
                              float* out, uint8_t* hit_out, cudaStream_t s) {
    const die_pipeline_config_t& c = p->cfg;
    const int vin = c.dim_in / 4, vout = c.dim_out / 4;
    uint8_t* code = hit_out ? hit_out : p->code;
    DIE_TRY(kv_get_batch_impl(p->kv, digests, B, now, code, p->value, nullptr, s));
    DIE_LAUNCH(ST_MISS_SCAN, s, miss_scan_kernel<<<1, 1024, 0, s>>>(code, B, p->miss_rank, p->n_miss));
    DIE_LAUNCH(ST_PACK_HITS, s, pack_hits_kernel<<<grid_256((long long)B * std::max(vin, vout)), 256, 0, s>>>(
        reinterpret_cast<const int4*>(payload), code, p->miss_rank,
        reinterpret_cast<const unsigned long long*>(p->value), reinterpret_cast<const int4*>(p->arena), B, vin, vout,
        reinterpret_cast<int4*>(p->packed_in), reinterpret_cast<int4*>(out)));
    if (c.model == DIE_MODEL_ECHO) {
        DIE_LAUNCH(ST_MODEL, s, echo_rows_kernel<<<grid_256((long long)B * vin), 256, 0, s>>>(reinterpret_cast<const int4*>(p->packed_in),
                                                                       reinterpret_cast<int4*>(p->packed_out), vin,
                                                                       p->n_miss));
    } else if (c.model == DIE_MODEL_LINEAR_FP32) {
        DIE_TRY(linear_f32_launch(p->packed_in, p->W, p->bias, p->packed_out, B, c.dim_in, c.dim_out, p->n_miss, s));
    } else {
        DIE_TRY(linear_tf32_launch(p->packed_in, p->W, p->bias, p->packed_out, B, c.dim_in, c.dim_out, p->n_miss, s));
    }
    const unsigned long long t_set = p->kv->tail;
    DIE_TRY(kv_set_batch_impl(p->kv, digests, nullptr, nullptr, B, now, 0, code, 1, p->n_miss, p->set_slot, nullptr, s));
    DIE_LAUNCH(ST_UNPACK_STORE, s, unpack_store_kernel<<<grid_256((long long)B * vout), 256, 0, s>>>(
        view_of(p->kv), reinterpret_cast<const int4*>(p->packed_out), code, p->miss_rank, p->set_slot, t_set, B, vout,
        reinterpret_cast<int4*>(out), reinterpret_cast<int4*>(p->arena)));
    return DIE_OK;
}

}  // namespace die

using namespace die;

extern "C" die_status_t die_pipeline_create(const die_pipeline_config_t* cfg, die_pipeline_t** out) {
    if (!cfg || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: null argument");
    if (cfg->max_batch <= 0 || cfg->dim_in <= 0 || cfg->dim_out <= 0 || (cfg->dim_in & 3) || (cfg->dim_out & 3))
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: max_batch > 0 and dims multiple of 4 fp32 required");
    if (cfg->model == DIE_MODEL_ECHO && cfg->dim_in != cfg->dim_out)
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: echo needs dim_in == dim_out");
    if (cfg->model != DIE_MODEL_ECHO && (!cfg->weight || !cfg->bias))
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: linear model needs weight and bias");
    if (cfg->model < DIE_MODEL_ECHO || cfg->model > DIE_MODEL_LINEAR_TF32)
        DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_create: unknown model %d", cfg->model);
    die_pipeline* p = new die_pipeline();
    memset(p, 0, sizeof(*p));
    p->cfg = *cfg;
    p->cfg.weight = nullptr;
    p->cfg.bias = nullptr;
    const uint64_t B = (uint64_t)cfg->max_batch;
    p->n_rows = cfg->kv_max_size + B + 64;
    DIE_TRY(kv_create_impl(cfg->kv_max_size, cfg->default_ttl, p->n_rows, B, &p->kv));
    if ((uint64_t)cfg->max_batch > p->kv->max_chunk)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_pipeline_create: max_batch %d exceeds the table's batch limit %llu",
                 cfg->max_batch, (unsigned long long)p->kv->max_chunk);
    DIE_CUDA(cudaMalloc(&p->arena, p->n_rows * cfg->dim_out * sizeof(float)));
    if (cfg->model != DIE_MODEL_ECHO) {
        DIE_CUDA(cudaMalloc(&p->W, (size_t)cfg->dim_out * cfg->dim_in * sizeof(float)));
        DIE_CUDA(cudaMalloc(&p->bias, (size_t)cfg->dim_out * sizeof(float)));
        DIE_CUDA(cudaMemcpy(p->W, cfg->weight, (size_t)cfg->dim_out * cfg->dim_in * sizeof(float),
                            cudaMemcpyHostToDevice));
        DIE_CUDA(cudaMemcpy(p->bias, cfg->bias, (size_t)cfg->dim_out * sizeof(float), cudaMemcpyHostToDevice));
    }
    DIE_CUDA(cudaMalloc(&p->digests, B * 16));
    DIE_CUDA(cudaMalloc(&p->shard, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->code, B));
    DIE_CUDA(cudaMalloc(&p->value, B * sizeof(uint64_t)));
    DIE_CUDA(cudaMalloc(&p->miss_rank, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->n_miss, sizeof(uint32_t)));
    DIE_CUDA(cudaMalloc(&p->set_slot, B * sizeof(int32_t)));
    DIE_CUDA(cudaMalloc(&p->packed_in, B * cfg->dim_in * sizeof(float)));
    DIE_CUDA(cudaMalloc(&p->packed_out, B * cfg->dim_out * sizeof(float)));
    const size_t key_bytes = cfg->max_key_bytes > 0 ? (size_t)cfg->max_key_bytes : B * 64;
    p->cfg.max_key_bytes = (int32_t)key_bytes;
    DIE_CUDA(cudaMalloc(&p->d_key_bytes, key_bytes + 16));
    DIE_CUDA(cudaMalloc(&p->d_key_off, (B + 1) * sizeof(int64_t)));
    DIE_CUDA(cudaMalloc(&p->d_payload, B * cfg->dim_in * sizeof(float)));
    DIE_CUDA(cudaMalloc(&p->d_out, B * cfg->dim_out * sizeof(float)));
    DIE_CUDA(cudaMalloc(&p->d_hit, B));
    *out = p;
    return DIE_OK;
}

extern "C" die_status_t die_pipeline_destroy(die_pipeline_t* p) {
    if (!p) return DIE_OK;
    die_kv_destroy(p->kv);
    void* bufs[] = {p->arena, p->W, p->bias, p->digests, p->shard, p->code, p->value, p->miss_rank, p->n_miss,
                    p->set_slot, p->packed_in, p->packed_out, p->d_key_bytes, p->d_key_off, p->d_payload, p->d_out,
                    p->d_hit};
    for (void* b : bufs) cudaFree(b);
    delete p;
    return DIE_OK;
}

extern "C" die_kv_t* die_pipeline_kv(die_pipeline_t* p) { return p ? p->kv : nullptr; }

extern "C" die_status_t die_pipeline_run_owner_device(die_pipeline_t* p, const uint32_t* digests, int32_t B,
                                                      const float* payload, double now, float* out,
                                                      uint8_t* hit_out, die_stream_t stream) {
    if (!p || B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_owner_device: bad B");
    if (B == 0) return DIE_OK;
    if (!digests || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_owner_device: null buffer");
    return run_owner(p, digests, B, payload, now, out, hit_out, static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_pipeline_run_device(die_pipeline_t* p, const uint8_t* key_bytes, const int64_t* key_off,
                                                int32_t B, const float* payload, double now, float* out,
                                                uint8_t* hit_out, uint32_t* digest_out, die_stream_t stream) {
    if (!p || B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device: bad B");
    if (B == 0) return DIE_OK;
    if (!key_off || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_device: null buffer");
    uint32_t* dg = digest_out ? digest_out : p->digests;
    DIE_TRY(die_route_md5(key_bytes, key_off, B, 1, nullptr, nullptr, 0, 0, p->shard, dg, nullptr, stream));
    return run_owner(p, dg, B, payload, now, out, hit_out, static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_pipeline_run_host(die_pipeline_t* p, const uint8_t* key_bytes, const int64_t* key_off,
                                              int32_t B, const float* payload, double now, float* out,
                                              uint8_t* hit_out, die_stream_t stream) {
    if (!p || B < 0 || B > p->cfg.max_batch) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_host: bad B");
    if (B == 0) return DIE_OK;
    if (!key_off || !payload || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pipeline_run_host: null buffer");
    const int64_t nbytes = key_off[B] - key_off[0];
    if (key_off[0] != 0 || nbytes < 0 || nbytes > p->cfg.max_key_bytes)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_pipeline_run_host: key bytes %lld outside [0, %d] (key_off[0] must be 0)",
                 (long long)nbytes, p->cfg.max_key_bytes);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const die_pipeline_config_t& c = p->cfg;
    if (nbytes) DIE_CUDA(cudaMemcpyAsync(p->d_key_bytes, key_bytes, (size_t)nbytes, cudaMemcpyHostToDevice, s));
    DIE_CUDA(cudaMemcpyAsync(p->d_key_off, key_off, (size_t)(B + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    DIE_CUDA(cudaMemcpyAsync(p->d_payload, payload, (size_t)B * c.dim_in * sizeof(float), cudaMemcpyHostToDevice, s));
    DIE_TRY(die_pipeline_run_device(p, p->d_key_bytes, p->d_key_off, B, p->d_payload, now, p->d_out, p->d_hit, nullptr,
                                    stream));
    DIE_CUDA(cudaMemcpyAsync(out, p->d_out, (size_t)B * c.dim_out * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (hit_out) DIE_CUDA(cudaMemcpyAsync(hit_out, p->d_hit, (size_t)B, cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    return DIE_OK;
}
