/*
 * die.h -- C-ABI of libdie.so: the B200 (sm_100a) implementation of the reference's one
 * data-parallel hot path, route -> probe -> pack -> infer -> unpack.
 *
 * The reference (a pure-Python prototype) has no FFI; the interface each entry point
 * replaces is a Python method, cited as file:line relative to the reference checkout.
 * INTEGRATION.md shows the ctypes stub a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns a die_status_t (0 = DIE_OK) and never throws or aborts;
 *     die_last_error() gives the message of the calling thread's last failure;
 *   - unless a parameter is marked [host], pointers are DEVICE pointers owned by the caller
 *     (e.g. torch tensors' data_ptr()); only die_kv_t / die_pipeline_t handles own memory;
 *   - all device work is enqueued on the cudaStream_t passed as `stream` (a void* here so the
 *     header needs no CUDA include); functions whose name ends in _sync or that return
 *     values through [host] pointers synchronise that stream;
 *   - one host thread per handle; handles are bound to the device current at creation.
 */
#ifndef DIE_H_
#define DIE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DIE_ABI_VERSION 1

typedef int32_t die_status_t;
enum {
    DIE_OK = 0,
    DIE_ERR_INVALID = 1,        /* bad argument */
    DIE_ERR_CUDA = 2,           /* a CUDA runtime call failed */
    DIE_ERR_NOMEM = 3,
    DIE_ERR_STOP_ITERATION = 4, /* the reference would have raised StopIteration (kvstore.py:89) */
    DIE_ERR_TOO_LARGE = 5,
    DIE_ERR_UNSUPPORTED = 6,
    DIE_ERR_VALUE_ERROR = 7     /* the reference would have raised ValueError: LFU min() of an empty dict (kvstore.py:95) */
};

typedef void* die_stream_t; /* cudaStream_t */

int32_t die_abi_version(void);
const char* die_last_error(void);
/* Makes `device` current for this thread inside the library's CUDA runtime instance. */
die_status_t die_set_device(int32_t device);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
uint64_t die_launch_count(void);
/* Number of CUDA-graph replays (groups of steps of a queue, see die_pipeline_run_device_many) launched so far. */
uint64_t die_replay_count(void);

/* Diagnostics with no counterpart in the reference: while a profile is open every kernel this
 * library launches is bracketed by CUDA events on the kernel's own stream; die_profile_end()
 * synchronises the device and returns, per stage, the summed device time (ms) and launch count.
 * Used by bench.py for the roofline of the dominant kernel; adds event overhead, so never left
 * on during a throughput measurement. */
die_status_t die_profile_begin(void);
die_status_t die_profile_end(int32_t cap, double* ms_out /*[host]*/, uint64_t* count_out /*[host]*/,
                             int32_t* n_stages /*[host]*/);
const char* die_profile_stage_name(int32_t stage);
/* The open profile as a timeline, one record per launch (call before die_profile_end): stage, a small
 * integer per distinct stream, start / end in microseconds after the first recorded launch began. */
die_status_t die_profile_timeline(int32_t cap, int32_t* stage_out /*[host]*/, int32_t* stream_out /*[host]*/,
                                  double* start_us /*[host]*/, double* end_us /*[host]*/, int32_t* n_out /*[host]*/);

/* Diagnostics (no reference counterpart): while enabled, the kernels of steps replayed from a CUDA graph log
 * (step << 8 | kernel id, %globaltimer ns) when they start; read_sync returns and clears the log. */
die_status_t die_debug_trace_enable(uint64_t n_records);
die_status_t die_debug_trace_read_sync(uint64_t cap, uint64_t* tag_out /*[host]*/, uint64_t* ns_out /*[host]*/,
                                       uint64_t* n_out /*[host]*/);

/* Calibration (no reference counterpart): n_threads threads each issue per_thread independent random reads of
 * bytes_each (16/32/64) bytes over buf[0..bytes).  bench.py times it to obtain the random-sector ceiling the
 * table kernels are held against, next to the streaming-copy peak. */
die_status_t die_debug_random_read(const void* buf, uint64_t bytes, int32_t n_threads, int32_t per_thread,
                                   int32_t bytes_each, uint64_t seed, uint64_t* sink, die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * K1  route: shard = int(md5(key).hexdigest(), 16) % n_shards, then the health gate + failover.
 * Replaces ModelRegistry.get_shard_for_key (src/model_registry.py:149-161),
 * Router.route_request (src/router.py:140-184) and Router._find_alternative_shard
 * (src/router.py:186-221) for a whole batch of keys.
 *
 *   key_bytes/key_off : CSR of the UTF-8 key bytes; key i = key_bytes[key_off[i] .. key_off[i+1])
 *   routable[s]       : 1 if shard position s may be returned directly, i.e. its worker is
 *                       registered and (healthy or !prefer_healthy); NULL = all routable
 *   healthy_idx[h]    : shard positions whose worker is HEALTHY, in shard-list order
 *                       (router.py:197-201); used when the primary is not routable
 *   failover          : Router.failover_enabled
 *   shard_out[i]      : shard LIST POSITION (not shard_id), or -1 where the reference returns None
 *   digest_out        : optional [B,4] uint32, the MD5 state words A,B,C,D (digest bytes, LE words)
 *   dest_counts       : optional [n_shards] int32, number of requests routed to each position
 * ------------------------------------------------------------------------------------------ */
die_status_t die_route_md5(const uint8_t* key_bytes, const int64_t* key_off, int32_t B,
                           int32_t n_shards, const uint8_t* routable,
                           const int32_t* healthy_idx, int32_t n_healthy, int32_t failover,
                           int32_t* shard_out, uint32_t* digest_out, int32_t* dest_counts,
                           die_stream_t stream);

/* Stable grouping of a routed batch by destination shard: the hand-off between Router.route_request
 * (src/router.py:140) and the owner workers in the composed path (the reference's coordinator, README.md:56-60,
 * was never written).  Requests keep request order inside their bucket; shard -1 (None) forms a last bucket.
 *   offsets_out [n_shards+2] : bucket starts (int64), offsets_out[n_shards+1] == B
 *   perm_out    [B]          : request index placed at send position j
 *   pos_out     [B]          : send position of request i (inverse of perm)                                  */
die_status_t die_bucket_by_shard(const int32_t* shard, int32_t B, int32_t n_shards, int64_t* offsets_out,
                                 int32_t* perm_out, int32_t* pos_out, die_stream_t stream);
/* Rows to / from per-shard regions after die_bucket_by_shard (same hand-off; used by the single-GPU coordinator):
 * scatter writes row q of src [B, row_bytes] to dst [n_shards][cap][row_bytes] at (shard[q], pos[q] - offsets[shard[q]])
 * and counts_out[s] = requests of shard s (uint32, device: the owner pass takes its batch size from there);
 * gather is the way back to request order (zero rows where shard[q] is -1), with one optional byte per row (the hit
 * flags) taken along.  row_bytes a multiple of 16, cap >= B. */
die_status_t die_bucket_scatter_rows(const void* src, const int32_t* shard, const int32_t* pos, const int64_t* offsets,
                                     int32_t B, int32_t n_shards, int32_t cap, int32_t row_bytes, void* dst,
                                     uint32_t* counts_out, die_stream_t stream);
die_status_t die_bucket_gather_rows(const void* src, const int32_t* shard, const int32_t* pos, const int64_t* offsets,
                                    int32_t B, int32_t n_shards, int32_t cap, int32_t row_bytes, void* dst,
                                    const uint8_t* flag_src /*optional [n_shards][cap]*/, uint8_t* flag_dst /*[B]*/,
                                    die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * K2/K3  KV store: GPU-resident open-addressing table keyed by the 128-bit MD5 of the key,
 * with the reference's LRU + TTL semantics applied "as if" the batch ran in request order.
 * Replaces KVCache (src/kvstore.py:26-236, eviction_policy="lru").
 *   default_ttl : NaN = None.   ttl arrays: NaN = "ttl=None" (use default_ttl); NULL = all None.
 *   now         : the value time.time() would return for every op of this batch.
 * ------------------------------------------------------------------------------------------ */
typedef struct die_kv die_kv_t;

#define DIE_KV_POLICY_LRU 0
#define DIE_KV_POLICY_LFU 1  /* victim = fewest uses, ties by the order the keys were first counted (kvstore.py:94-98) */
#define DIE_KV_POLICY_FIFO 2 /* victim = oldest insertion; accesses do not matter (kvstore.py:99-102) */

/* KVCache.__init__ (kvstore.py:38-61).  LFU evicts by a scan of the table per victim (the reference's min() over its
 * dict is O(n) too): meant for policy caches of modest size, not for the 10^7-entry tables of the LRU path. */
die_status_t die_kv_create(uint64_t max_size, int32_t policy, double default_ttl, die_kv_t** out);
die_status_t die_kv_destroy(die_kv_t* kv);

/* KVCache.get / batch_get (kvstore.py:136-168) for B keys in request order.
 *   hit_out[i]   : 1 hit; 0 miss (absent); 2 miss because expired -- this op deleted the entry
 *                  and value_out[i] holds the dropped value
 *   value_out[i] : stored value on hit
 *   slot_out     : optional, table slot index of a hit (-1 otherwise)                        */
die_status_t die_kv_get_batch(die_kv_t* kv, const uint32_t* digests, int32_t B, double now,
                              uint8_t* hit_out, uint64_t* value_out, int32_t* slot_out,
                              die_stream_t stream);

/* KVCache.set applied to M (key, value, ttl) triples in request order (kvstore.py:104-134);
 * deferred_stamp == 1 gives KVCache.batch_set (kvstore.py:170-176; keys must be distinct);
 * deferred_stamp == 2 gives set(..., batch=True) on its own: inserted but never put in the LRU order.
 *   values NULL        : the table picks the value itself (slot index, or a result-row handle
 *                        when it belongs to a die_pipeline_t)
 *   slot_out           : optional [M], slot holding op i's key after the batch
 *   dropped_out        : optional [2*M] values of entries evicted or overwritten by this batch;
 *   the count and the StopIteration condition are read back with die_kv_sync_status().        */
die_status_t die_kv_set_batch(die_kv_t* kv, const uint32_t* digests, const uint64_t* values,
                              const double* ttl, int32_t M, double now, int32_t deferred_stamp,
                              int32_t* slot_out, uint64_t* dropped_out, die_stream_t stream);

/* No counterpart in the reference.  die_kv_set_batch for callers that are not allowed to allocate device
 * memory once they are up: a serving loop that reserved all HBM at start, a latency-critical section that
 * cannot take cudaMalloc's device-wide synchronisation, a runtime that forbids allocation after
 * initialisation.  Same arguments and same results as die_kv_set_batch, but
 *   - the per-op scratch (one int32 per op) is a fixed 256 KiB array inside the library's device image
 *     (every GPU that loads libdie.so holds it until the process exits), so M must be
 *     <= DIE_KV_NOALLOC_MAX_OPS and <= the table's per-launch limit, else DIE_ERR_TOO_LARGE;
 *   - it never calls cudaMalloc/cudaFree and never starts table housekeeping (which allocates): when the
 *     host-side bounds say housekeeping is due it returns DIE_ERR_NOMEM and the caller runs
 *     die_kv_maintain_sync where allocation is allowed, then retries;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress); make
 *     that first call before starting any stream capture on that device;
 *   - all calls share the one array and it keeps the previous call's contents (nothing in it needs
 *     clearing: each call overwrites what it reads).  Two calls whose kernels overlap on the GPU
 *     (different streams, different tables) race on it: serialise them. */
#define DIE_KV_NOALLOC_MAX_OPS 65536
die_status_t die_kv_set_batch_noalloc(die_kv_t* kv, const uint32_t* digests, const uint64_t* values,
                                      const double* ttl, int32_t M, double now, int32_t deferred_stamp,
                                      int32_t* slot_out, uint64_t* dropped_out, die_stream_t stream);

/* KVCache.delete (kvstore.py:178-187): deleted_out[i] = 1 if the key was present (value_out[i] = its value). */
die_status_t die_kv_delete_batch(die_kv_t* kv, const uint32_t* digests, int32_t B,
                                 uint8_t* deleted_out, uint64_t* value_out, die_stream_t stream);
/* KVCache.__contains__ (kvstore.py:221-228): 1 present; 0 absent; 2 expired (deleted, value_out set). */
die_status_t die_kv_contains_batch(die_kv_t* kv, const uint32_t* digests, int32_t B, double now,
                                   uint8_t* out, uint64_t* value_out, die_stream_t stream);
/* KVCache.__len__ (kvstore.py:230-236): purge every expired entry, then count.  purged_out (optional,
 * device, capacity purged_cap) receives the dropped values. Synchronises. */
die_status_t die_kv_len_sync(die_kv_t* kv, double now, uint64_t* purged_out, uint64_t purged_cap,
                             uint64_t* n_purged /*[host]*/, uint64_t* len_out /*[host]*/, die_stream_t stream);
/* KVCache.clear (kvstore.py:189-198). */
die_status_t die_kv_clear(die_kv_t* kv, die_stream_t stream);
/* KVCache.get_stats (kvstore.py:206-219): out = {hits, misses, evictions, size (unpurged), max_size,
 * capacity, tombstones, lru_log_span}. Synchronises. */
die_status_t die_kv_stats_sync(die_kv_t* kv, uint64_t out[8] /*[host]*/, die_stream_t stream);
/* Synchronises, returns and clears the sticky batch status: DIE_OK or DIE_ERR_STOP_ITERATION
 * (in *status), and how many values the last set batch wrote to dropped_out. */
die_status_t die_kv_sync_status(die_kv_t* kv, int32_t* status /*[host]*/, uint64_t* n_dropped /*[host]*/,
                                uint32_t* slow_path_batches /*[host], optional*/, die_stream_t stream);
/* Parity dump: the live entries in LRU order (head = next victim), unstamped entries last in slot
 * order; arrays are [host] with room for `cap` entries.  Mirrors list(_access_order) + self.cache. */
die_status_t die_kv_dump_sync(die_kv_t* kv, uint64_t cap, uint32_t* digests /*[cap,4]*/, uint64_t* values,
                              double* created_at, double* last_accessed, uint32_t* access_count,
                              double* ttl, uint8_t* stamped, uint64_t* n_out, die_stream_t stream,
                              uint32_t* lfu_count /*optional: _access_count of an LFU table*/);
/* Persistence (README.md:14 "optional persistence"; the reference never wrote it).  The whole table -- entries,
 * eviction order, counters -- as one flat [host] image of die_kv_snapshot_bytes(kv) bytes, and back into a table
 * created with the same max_size and policy.  Values are the caller's 64-bit handles, stored as they are.
 * Not for tables that belong to a die_pipeline_t (their values are result-row handles).  Both synchronise. */
uint64_t die_kv_snapshot_bytes(die_kv_t* kv);
die_status_t die_kv_snapshot_sync(die_kv_t* kv, void* image /*[host]*/, uint64_t image_bytes, die_stream_t stream);
die_status_t die_kv_restore_sync(die_kv_t* kv, const void* image /*[host]*/, uint64_t image_bytes, die_stream_t stream);
/* Housekeeping that has no counterpart in the reference: renumber the LRU log and/or rebuild the
 * table when tombstones pile up.  Called automatically by the batch entry points when their
 * conservative host-side bounds require it; force != 0 does both now.  Synchronises. */
die_status_t die_kv_maintain_sync(die_kv_t* kv, int32_t force, die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * K4/K5  pack / unpack.  Replace Batcher._process_batch's gather
 * `inputs_list = [req.inputs for req in requests_to_process]` (src/batcher.py:211) and scatter
 * `for req, result in zip(...): req.future.set_result(result)` (src/batcher.py:225-227).
 *   values/off : ragged CSR of request payloads, off in ELEMENTS (elem_bytes each)
 *   src_index  : optional [R]; packed row r holds request src_index[r] (NULL: request r)
 *   out        : [R, Lpad] row-major, rows zero-padded; lengths_out[r] = element count of row r
 * Rows longer than Lpad are truncated to Lpad (lengths_out still holds the full length).
 * ------------------------------------------------------------------------------------------ */
die_status_t die_pack_ragged(const void* values, const int64_t* off, const int32_t* src_index,
                             int32_t R, int32_t elem_bytes, int32_t Lpad, void* out,
                             int32_t* lengths_out, die_stream_t stream);
/* Row r of packed [R, Lpad] -> request q = dst_index ? dst_index[r] : r, written to
 * out_values[off_out[q] .. off_out[q+1]). */
die_status_t die_unpack_ragged(const void* packed, const int64_t* off_out, const int32_t* dst_index,
                               int32_t R, int32_t elem_bytes, int32_t Lpad, void* out_values,
                               die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * K6-K8  mock-model forward.  The reference's model is an echo (src/mock_models/fake_model.py:52);
 * the linear layer and the MLP are the configs BASELINE.json names and have no reference arithmetic.
 * ------------------------------------------------------------------------------------------ */
die_status_t die_echo_fwd(const void* in, void* out, uint64_t bytes, die_stream_t stream);
#define DIE_LINEAR_FP32 0 /* CUDA-core FMA, fp32 accumulate */
#define DIE_LINEAR_TF32 1 /* tcgen05 kind::tf32, fp32 accumulate in TMEM */
/* Y[B,N] = X[B,K] @ W[N,K]^T + b[N], fp32 storage. */
die_status_t die_linear_fwd_f32(const float* X, const float* W, const float* b, float* Y,
                                int32_t B, int32_t K, int32_t N, int32_t precision, die_stream_t stream);
/* Y[B,D] = bf16( bf16(relu(X[B,D] @ W1[H,D]^T + b1)) @ W2[D,H]^T + b2 ); X,W1,W2,Y bf16, b1,b2 fp32,
 * hidden [B,H] bf16 scratch supplied by the caller. */
die_status_t die_mlp_fwd_bf16(const void* X, const void* W1, const float* b1, const void* W2,
                              const float* b2, void* hidden, void* Y, int32_t B, int32_t D, int32_t H,
                              die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Composed pipeline for one owner shard (SURVEY.md call-stack (E); the coordinator the
 * reference's README.md:56-60 describes but never wrote).  One call = one flushed batch:
 * route -> get -> pack misses -> model -> unpack -> set misses, results in request order.
 * Fixed-width payloads (dim_in elements in, dim_out elements out, fp32).
 * ------------------------------------------------------------------------------------------ */
typedef struct die_pipeline die_pipeline_t;
#define DIE_MODEL_ECHO 0
#define DIE_MODEL_LINEAR_FP32 1
#define DIE_MODEL_LINEAR_TF32 2
#define DIE_MODEL_MLP_BF16 3   /* Y = relu(X W1^T + b1) W2^T + b2, bf16 storage, fp32 accumulation (die_mlp_fwd_bf16); dim_in ==
                                * dim_out = D, hidden = H; weight = W1 [H,D] then W2 [D,H], bias = b1 [H] then b2 [D], fp32 on
                                * the host, converted once; payload and result rows stay fp32 at the interface */

typedef struct {
    uint64_t kv_max_size;
    double default_ttl;     /* NaN = None */
    int32_t max_batch;      /* largest B a call may pass */
    int32_t dim_in;         /* elements per request payload */
    int32_t dim_out;        /* elements per result (== dim_in for echo) */
    int32_t model;          /* DIE_MODEL_* */
    int32_t max_key_bytes;  /* upper bound of key bytes per batch (host-buffer entry point) */
    int32_t hidden;         /* DIE_MODEL_MLP_BF16: H (multiple of 8); other models: 0 */
    const float* weight;    /* [host] [dim_out, dim_in] (linear), W1 then W2 (MLP), or NULL */
    const float* bias;      /* [host] [dim_out] (linear), b1 then b2 (MLP), or NULL */
} die_pipeline_config_t;

die_status_t die_pipeline_create(const die_pipeline_config_t* cfg, die_pipeline_t** out);
die_status_t die_pipeline_destroy(die_pipeline_t* p);
die_kv_t* die_pipeline_kv(die_pipeline_t* p);
/* Diagnostics (no reference counterpart): %globaltimer marks (ns) written by the last fused get kernel:
 * [0] start, [1] keys hashed, [2] all probes done, [3] misses ranked, [4] victim window done, [5] victims found. */
die_status_t die_pipeline_debug_marks_sync(die_pipeline_t* p, uint64_t out[8] /*[host]*/);
/* Diagnostics of the persistent step kernel (no reference counterpart): the same marks of the last batch for the first
 * n_cta (<= 160) CTAs of its grid, out[n_cta][8]: [0] batch start, [4] probes issued, [1] barrier 1 passed, [5] inserts
 * done, [7] accumulators ready (request CTAs) / victim bitmap done (window CTAs), [2] barrier 2 passed, [6] epilogue
 * done, [3] barrier 3 passed. */
die_status_t die_pipeline_debug_cta_marks_sync(die_pipeline_t* p, int32_t n_cta, uint64_t* out /*[host]*/);
/* Inputs already resident in HBM.  payload [B, dim_in] fp32; out [B, dim_out] fp32;
 * hit_out [B] uint8 (1 = served from the cache); digest_out optional [B,4]. */
die_status_t die_pipeline_run_device(die_pipeline_t* p, const uint8_t* key_bytes, const int64_t* key_off,
                                     int32_t B, const float* payload, double now, float* out,
                                     uint8_t* hit_out, uint32_t* digest_out, die_stream_t stream);
/* The owner-side half of the path for requests whose digests are already known (after the
 * all-to-all dispatch): get -> pack -> model -> unpack -> set. */
die_status_t die_pipeline_run_owner_device(die_pipeline_t* p, const uint32_t* digests, int32_t B,
                                           const float* payload, double now, float* out, uint8_t* hit_out,
                                           die_stream_t stream);
/* Same through HOST buffers (pinned recommended): copies keys + payload in, results + hit flags
 * out, and synchronises.  This is the call the reference-facing Python plugin makes. */
die_status_t die_pipeline_run_host(die_pipeline_t* p, const uint8_t* key_bytes /*[host]*/,
                                   const int64_t* key_off /*[host]*/, int32_t B,
                                   const float* payload /*[host]*/, double now, float* out /*[host]*/,
                                   uint8_t* hit_out /*[host]*/, die_stream_t stream);

/* The owner-side pass when the request count is only known on the device (*n_req_dev <= B_max), i.e. right
 * after peers have written their requests into this rank's receive window. */
die_status_t die_pipeline_run_owner_device_n(die_pipeline_t* p, const uint32_t* digests, int32_t B_max,
                                             const uint32_t* n_req_dev, const float* payload, double now, float* out,
                                             uint8_t* hit_out, die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Ingress data format.  The reference's worker takes one JSON document per TCP connection,
 * {"model", "version", "inputs", "request_id"} (src/worker.py:92-100: reader.read(4096) + json.loads, so a request
 * cannot exceed 4096 bytes).  The binary replacement is a length-prefixed frame, little-endian, a multiple of 4 bytes:
 *     die_frame_header_t | key bytes (UTF-8), zero-padded to 4 | fp32 inputs[n_floats]
 * The host keeps the transport: it hands over a buffer of frames back to back and frame_off[n+1], where each frame
 * starts (the receive loop knows that from frame_bytes).  die_frame_decode turns them into what the path consumes:
 * CSR keys (key_bytes_out, key_off_out[n+1]) for die_route_md5 / the pipeline, fixed-width rows payload_out[n, dim]
 * zero-padded (PACK, src/batcher.py:211, done at the door), request ids, queue = index of "model:version".
 * status_out[i]: DIE_FRAME_OK, DIE_FRAME_BAD (magic / lengths inconsistent with the frame's extent: empty key, zero row)
 * or DIE_FRAME_TOO_WIDE (n_floats > dim); *n_bad_out (device) counts the frames that are not OK.
 * ------------------------------------------------------------------------------------------ */
typedef struct die_frame_header {
    uint32_t magic;        /* 0x31454944, "DIE1" */
    uint32_t frame_bytes;  /* header + padded key + inputs */
    uint64_t request_id;
    uint16_t key_len;
    uint16_t queue;
    uint32_t n_floats;
} die_frame_header_t;
enum { DIE_FRAME_OK = 0, DIE_FRAME_BAD = 1, DIE_FRAME_TOO_WIDE = 2 };
die_status_t die_frame_decode(const uint8_t* frames, const int64_t* frame_off /*[n+1]*/, int32_t n, int32_t dim,
                              uint8_t* key_bytes_out, int64_t* key_off_out /*[n+1]*/, float* payload_out /*[n, dim]*/,
                              uint64_t* request_id_out /*[n]*/, int32_t* queue_out /*[n]*/, uint8_t* status_out /*[n]*/,
                              uint32_t* n_bad_out, die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Device-resident Batcher: per-key accumulation with size- and latency-triggered flush, for requests that already
 * live on the GPU.  Replaces Batcher.add_request / _schedule_flush / _flush_batch / _process_batch / stop
 * (src/batcher.py:70-230) for fixed-width payloads: one open batch per key ("model:version", numbered 0..n_queues-1 by
 * the host), flushed when it holds max_batch requests (batcher.py:143) or when max_latency has passed since it was
 * opened (batcher.py:150-166).  Time advances in ticks:
 *   die_batcher_tick(now, force, arrivals): 1. every open batch with now - opened_at >= max_latency flushes, in the
 *     order the batches were opened; 2. the arrivals are appended in order, flushing every batch they fill; 3. force != 0
 *     is stop() (batcher.py:70-100): the batches still open flush in creation order.
 *   A flushed batch is packed (batcher.py:211), run through the model -- echo (weights NULL, fake_model.py:52) or the
 *   mock fp32 linear layer with the queue's own weights -- and unpacked (batcher.py:225-227): the result row of request r
 *   lands in out[request_id[r]], and batch_of[request_id[r]] = the number of the batch it was served in (optional).
 *   request_id is the caller's handle for the future (batcher.py:118): an index into out.
 * All on the stream, no host round trip; die_batcher_flushes_sync reads back what the LAST tick flushed (descriptors
 * in flush order, optionally the request ids of every batch, [cap][max_batch]) and die_batcher_stats_sync the
 * reference's counters (batcher.py:58-60) plus the open batches' sizes.  A tick that would flush more batches than
 * max_arrivals / max_batch + 2 * n_queues is dropped and the next *_sync call returns DIE_ERR_TOO_LARGE.
 * ------------------------------------------------------------------------------------------ */
typedef struct die_batcher die_batcher_t;
enum { DIE_FLUSH_LATENCY = 0, DIE_FLUSH_SIZE = 1, DIE_FLUSH_FORCED = 2 };
typedef struct die_flush_desc {
    int32_t queue;     /* key of the batch */
    int32_t count;     /* requests in it */
    int32_t trigger;   /* DIE_FLUSH_* */
    int32_t pad;
    double opened_at;  /* Batch.created_at (batcher.py:127) */
} die_flush_desc_t;
die_status_t die_batcher_create(int32_t n_queues, int32_t max_batch, double max_latency_s, int32_t dim_in, int32_t dim_out,
                                int32_t max_arrivals, const float* weights /*[host] [n_queues][dim_out][dim_in] or NULL*/,
                                const float* bias /*[host] [n_queues][dim_out] or NULL*/, die_batcher_t** out);
die_status_t die_batcher_destroy(die_batcher_t* b);
die_status_t die_batcher_tick(die_batcher_t* b, double now, int32_t force, const int32_t* queue /*[n]*/,
                              const int32_t* request_id /*[n]*/, const float* payload /*[n, dim_in]*/, int32_t n,
                              float* out /*[*, dim_out], indexed by request id*/, int32_t* batch_of /*or NULL*/,
                              die_stream_t stream);
die_status_t die_batcher_flushes_sync(die_batcher_t* b, int32_t* n_flush /*[host]*/, die_flush_desc_t* desc /*[host] [cap] or NULL*/,
                                      int32_t* ids /*[host] [cap][max_batch] or NULL*/, int32_t cap,
                                      uint64_t* first_batch /*[host] or NULL*/, die_stream_t stream);
die_status_t die_batcher_stats_sync(die_batcher_t* b, uint64_t* total_requests /*[host]*/, uint64_t* total_batches /*[host]*/,
                                    uint64_t* total_batched_requests /*[host]*/, int32_t* pending /*[host] [n_queues] or NULL*/,
                                    die_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Dispatch / return between the GPUs of one NVSwitch box by direct peer stores.  Replaces the reference's
 * TCP + JSON hop, one request per connection (src/worker.py:84-133).  peer_* are [host] arrays of `world`
 * device pointers into a symmetric allocation (one window per rank, see csrc/exchange.cu for the layout);
 * the caller separates the phases with a device-side barrier of that allocation.
 *   die_xchg_dispatch : PACK (bucket order = perm/offsets of die_bucket_by_shard) + SEND rows [digest | payload]
 *                       into each owner's receive window, and the per-owner counts
 *   die_xchg_compact  : owner: window regions -> contiguous digests / payload in (source rank, request) order,
 *                       *n_out = number of requests received, prefix_out[world+1] = per-source starts
 *   die_xchg_return   : owner: results + hit flags -> each source's return window
 *   die_xchg_scatter  : source: return window -> out / hit in request order (requests routed to None get zeros)
 * ------------------------------------------------------------------------------------------ */
die_status_t die_xchg_dispatch(const uint32_t* digests, const float* payload, const int32_t* perm,
                               const int64_t* offsets, int32_t B, int32_t dim, int32_t world, int32_t rank, int32_t cap,
                               void* const* peer_rows /*[host]*/, void* const* peer_counts /*[host]*/,
                               die_stream_t stream);
die_status_t die_xchg_compact(const float* rows, const int32_t* counts, int32_t world, int32_t cap, int32_t dim,
                              uint32_t* digests_out, float* payload_out, uint32_t* n_out, int32_t* prefix_out,
                              die_stream_t stream);
die_status_t die_xchg_return(const float* out, const uint8_t* hit, const int32_t* prefix, int32_t world, int32_t rank,
                             int32_t cap, int32_t dim_out, void* const* peer_back /*[host]*/, die_stream_t stream);
die_status_t die_xchg_scatter(const float* back, const int32_t* perm, const int64_t* offsets, int32_t B, int32_t world,
                              int32_t cap, int32_t dim_out, float* out, uint8_t* hit, die_stream_t stream);

/* The whole sharded step in one call: route (+ slot ranking), dispatch, flag barrier inside the window, compact,
 * owner pipeline, return, barrier, scatter -- three in-order stage streams over four buffer sets, so consecutive
 * steps overlap; die_xchg_step_many replays groups of steps from one CUDA graph (host: two launches per group).
 * The caller allocates one symmetric window of die_xchg_window_bytes() per rank (zero-filled; it holds four
 * buffer sets), exchanges the addresses and passes all of them as
 * peer_windows [host, world entries].  Every rank must make the same sequence of calls: the barriers are collective.
 * Requires one rank per GPU.  die_xchg_check_sync reports a barrier that timed out. */
typedef struct die_xchg die_xchg_t;
uint64_t die_xchg_window_bytes(int32_t world, int32_t cap, int32_t dim);
/* owner_max: the most requests one owner handles per step (= max_batch of `pipe`); 0 = world*cap.  Hash sharding
 * spreads a batch evenly, so 2*cap is ample; a step that exceeds it sets an error die_xchg_check_sync reports. */
die_status_t die_xchg_create(int32_t world, int32_t rank, int32_t cap, int32_t dim, int32_t owner_max,
                             die_pipeline_t* pipe, void* const* peer_windows /*[host]*/, die_xchg_t** out);
die_status_t die_xchg_destroy(die_xchg_t* x);
die_status_t die_xchg_step(die_xchg_t* x, const uint8_t* key_bytes, const int64_t* key_off, int32_t B,
                           const float* payload, double now, float* out, uint8_t* hit_out, int32_t* shard_out,
                           die_stream_t stream);
die_status_t die_xchg_step_many(die_xchg_t* x, int32_t n_batches, const uint8_t* key_bytes, int64_t key_bytes_stride,
                                const int64_t* key_off, int64_t key_off_stride, int32_t B, const float* payload,
                                double now0, double now_step, float* out, uint8_t* hit_out, int32_t out_ring,
                                die_stream_t stream);
die_status_t die_xchg_check_sync(die_xchg_t* x, die_stream_t stream);
/* No reference counterpart.  Captures and instantiates, without launching, the replay the next
 * die_xchg_step_many(B, ...) of this context would use: keeps graph instantiation (which may allocate and
 * synchronise the device) out of the first serving call. */
die_status_t die_xchg_prepare(die_xchg_t* x, int32_t B);
/* The Router's view of the owners for the steps enqueued afterwards (Router.route_request's health gate and
 * _find_alternative_shard, src/router.py:165-221; same arithmetic as die_route_md5): routable[world] [host] and the
 * ordered positions of the HEALTHY shards.  A request whose shard is not routable goes to
 * healthy_idx[md5 mod n_healthy] when failover is on and to nobody otherwise (shard -1, zero row, not a hit).
 * routable == NULL: every rank routable.  Every rank of the group passes the same view; a rank marked unhealthy still
 * takes part in the step's barriers (it is the router's opinion, not a dead process).  With failover an owner can
 * receive more than its share: size owner_max for it, an overflow is reported by die_xchg_check_sync. */
die_status_t die_xchg_set_health(die_xchg_t* x, const uint8_t* routable /*[host]*/, const int32_t* healthy_idx /*[host]*/,
                                 int32_t n_healthy, int32_t failover, die_stream_t stream);

/* A queue of n flushed batches in one call (each batch is still one pass of the path, applied strictly in
 * order): batch j reads key_bytes + j*key_bytes_stride, key_off + j*key_off_stride (0 = shared offsets),
 * payload + j*B*dim_in; device form writes out + (j % out_ring)*B*dim_out and hit_out + (j % out_ring)*B.
 * Groups of 16 batches (environment DIE_GRAPH_STEPS; 0 = off) are replayed from one captured CUDA graph whose
 * kernels read their per-batch arguments from a device-resident descriptor: same results as n single calls
 * (tests/test_gpu_pipeline.py), host cost two launches per group. */
die_status_t die_pipeline_run_device_many(die_pipeline_t* p, int32_t n_batches, const uint8_t* key_bytes,
                                          int64_t key_bytes_stride, const int64_t* key_off, int64_t key_off_stride,
                                          int32_t B, const float* payload, double now0, double now_step, float* out,
                                          uint8_t* hit_out, int32_t out_ring, die_stream_t stream);
/* Host-buffer form ([host] pinned pointers; out [n,B,dim_out], hit_out [n,B]): one copy per array and one
 * replay per group of batches; copy-in, kernels and copy-out of consecutive groups overlap on three streams
 * over three staging sets.  Synchronises before returning. */
die_status_t die_pipeline_run_host_many(die_pipeline_t* p, int32_t n_batches, const uint8_t* key_bytes /*[host]*/,
                                        int64_t key_bytes_stride, const int64_t* key_off /*[host]*/,
                                        int64_t key_off_stride, int32_t B, const float* payload /*[host]*/,
                                        double now0, double now_step, float* out /*[host]*/,
                                        uint8_t* hit_out /*[host]*/, die_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DIE_H_ */
