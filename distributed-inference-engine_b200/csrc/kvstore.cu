// K2/K3 kernels + host entry points of the KV store (src/kvstore.py:26-236).
// Layout and batch semantics are documented at the top of kv_internal.cuh.
#include <algorithm>
#include <vector>

#include <atomic>

#include "kv_internal.cuh"

namespace die {

constexpr int KV_THREADS = 256;

// ------------------------------------------------------------------------------------------------
// get / contains  (kvstore.py:136-164, 221-228)
// ------------------------------------------------------------------------------------------------
template <bool CONTAINS>
__global__ void __launch_bounds__(KV_THREADS)
kv_get_kernel(KvView kv, const uint4* __restrict__ digests, int B, double now, unsigned long long t,
              uint8_t* __restrict__ code_out, unsigned long long* __restrict__ value_out,
              int32_t* __restrict__ slot_out) {
    const int i = blockIdx.x * KV_THREADS + threadIdx.x;
    const bool active = i < B;
    int code = 0;
    if (active) {
        unsigned long long value;
        long long s;
        code = kv_get_one<CONTAINS>(kv, fp_of(digests[i]), now, t + (unsigned long long)i, value, s);
        code_out[i] = (uint8_t)code;
        if (value_out) value_out[i] = value;
        if (slot_out) slot_out[i] = code == 1 ? (int32_t)s : -1;
    }
    if (!CONTAINS) {
        block_add(&kv.st->hit_part[blockIdx.x & 15][0], active && code == 1);
        block_add(&kv.st->miss_part[blockIdx.x & 15][0], active && code != 1);
    }
    const bool deleted = active && code == 2;
    block_sub(&kv.st->size, deleted);
    block_add(&kv.st->n_tomb, deleted);
}

// delete (kvstore.py:178-187)
__global__ void __launch_bounds__(KV_THREADS)
kv_delete_kernel(KvView kv, const uint4* __restrict__ digests, int B, uint8_t* __restrict__ deleted_out,
                 unsigned long long* __restrict__ value_out) {
    const int i = blockIdx.x * KV_THREADS + threadIdx.x;
    const bool active = i < B;
    bool deleted = false;
    if (active) {
        const u128 key = fp_of(digests[i]);
        const long long s = kv_find(kv, key);
        unsigned long long value = 0;
        if (s >= 0) {
            Slot* sl = &kv.slots[s];
            const u128 old = cas128(sl, key, u128{FP_TOMB_WORD, FP_TOMB_WORD});
            if (fp_eq(old, key)) {
                value = ld_u64_cg(&sl->value);
                sl->stamp = 0ull;
                release_row(kv, value);
                deleted = true;
            }
        }
        deleted_out[i] = deleted ? 1 : 0;
        if (value_out) value_out[i] = value;
    }
    block_sub(&kv.st->size, deleted);
    block_add(&kv.st->n_tomb, deleted);
}

// __len__ purge (kvstore.py:230-236): every expired entry is deleted.
__global__ void __launch_bounds__(KV_THREADS)
kv_purge_kernel(KvView kv, double now, unsigned long long* __restrict__ purged_out,
                unsigned long long purged_cap, unsigned long long* __restrict__ n_purged) {
    const unsigned long long cap = kv.mask + 1;
    for (unsigned long long s = blockIdx.x * (unsigned long long)KV_THREADS + threadIdx.x; s < cap;
         s += (unsigned long long)gridDim.x * KV_THREADS) {
        Slot* sl = &kv.slots[s];
        const u128 fp = ld_fp(sl);
        if (fp_empty(fp) || fp_tomb(fp)) continue;
        if (ttl_expired(now, sl->created_at, sl->ttl)) {
            const unsigned long long value = sl->value;
            st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
            sl->stamp = 0ull;
            release_row(kv, value);
            atomicAdd(&kv.st->size, (unsigned long long)(-1ll));
            atomicAdd(&kv.st->n_tomb, 1ull);
            const unsigned long long k = atomicAdd(n_purged, 1ull);
            if (purged_out && k < purged_cap) purged_out[k] = value;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// set, phase A: read-only precheck.  Finds the batch keys that already live in the table.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(KV_THREADS)
kv_set_precheck_kernel(KvView kv, const uint4* __restrict__ digests, int M, const uint8_t* __restrict__ skip,
                       int32_t* __restrict__ op_slot) {
    const int i = blockIdx.x * KV_THREADS + threadIdx.x;
    const bool active = i < M && !(skip && skip[i] == 1);
    bool pre = false;
    if (active) {
        const long long s = kv_find(kv, fp_of(digests[i]));
        op_slot[i] = (int32_t)s;
        if (s >= 0) {
            pre = true;
            const unsigned long long p = ld_u64_cg(&kv.slots[s].stamp);
            if (p > 0) atomicMin(&kv.st->min_prev, p);
        }
    }
    block_add(&kv.st->n_pre, active && pre);
    block_add(&kv.st->n_absent, active && !pre);
}

// ------------------------------------------------------------------------------------------------
// LRU window walk, one CTA of 1024 threads.  Visits log records from `head` in stamp order and
// selects the first E current ones (the E least-recently-used entries).  apply=true evicts them.
// Returns (to all threads) one past the E-th current record, or ~0 when fewer than E exist
// before `limit`.
// ------------------------------------------------------------------------------------------------
constexpr int WIN_THREADS = 1024;
constexpr int WIN_ITEMS = 4;

__device__ __forceinline__ bool record_current(const KvView& kv, unsigned long long p, uint32_t& slot) {
    slot = kv.log[p & kv.rmask];
    return (unsigned long long)slot <= kv.mask && ld_u64_cg(&kv.slots[slot].stamp) == p;
}

__device__ __forceinline__ void evict_slot(const KvView& kv, uint32_t s, unsigned long long* dropped_out) {
    Slot* sl = &kv.slots[s];
    const unsigned long long value = sl->value;
    if (dropped_out) dropped_out[atomicAdd(&kv.st->n_dropped, 1ull)] = value;
    release_row(kv, value);
    st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
    sl->stamp = 0ull;
}

__device__ unsigned long long lru_window_select(const KvView& kv, unsigned long long head,
                                                unsigned long long limit, unsigned long long E, bool apply,
                                                unsigned long long* dropped_out,
                                                unsigned long long* first_current /*optional*/) {
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned long long s_hnew;
    __shared__ unsigned long long s_first;
    const int tid = threadIdx.x;
    if (tid == 0) {
        s_hnew = ~0ull;
        s_first = ~0ull;
    }
    __syncthreads();
    unsigned long long found = 0, pos = head;
    while (found < E && pos < limit) {
        bool valid[WIN_ITEMS];
        uint32_t slot[WIN_ITEMS];
        unsigned cnt = 0;
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; ++k) {
            const unsigned long long p = pos + (unsigned long long)tid * WIN_ITEMS + k;
            valid[k] = p < limit && record_current(kv, p, slot[k]);
            cnt += valid[k];
        }
        unsigned excl;
        const unsigned total = block_scan_1024(cnt, excl, s_warp);
        unsigned long long rank = found + excl;
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; ++k) {
            if (!valid[k]) continue;
            const unsigned long long p = pos + (unsigned long long)tid * WIN_ITEMS + k;
            if (rank == 0) s_first = p;
            if (rank < E) {
                if (apply) evict_slot(kv, slot[k], dropped_out);
                if (rank == E - 1) s_hnew = p + 1;
            }
            ++rank;
        }
        found += total;
        pos += (unsigned long long)WIN_THREADS * WIN_ITEMS;
        __syncthreads();
    }
    __syncthreads();
    if (first_current) *first_current = s_first;
    return s_hnew;
}

// ------------------------------------------------------------------------------------------------
// set, phase B: can the batch's evictions reach a key of this batch, or run out of victims?
// The LRU log is read from `head` in tiles of WIN_TILE positions by independent CTAs (B1: current records per
// tile); one CTA then scans the tile counts, finds the tile that holds the e_ub-th current record and walks only
// that tile (B2).  The evictions after the claims reuse the same tiles and prefixes (phase D).  A window of
// WIN_FACTOR positions per op that turns out too short (sparse log) falls back to the one-CTA walk, exactly.
// ------------------------------------------------------------------------------------------------
constexpr unsigned long long WIN_TILE = (unsigned long long)WIN_THREADS * WIN_ITEMS;
constexpr int WIN_FACTOR = 8;

__device__ __forceinline__ unsigned win_tile_flags(const KvView& kv, unsigned long long pos, unsigned long long limit,
                                                   bool (&valid)[WIN_ITEMS], uint32_t (&slot)[WIN_ITEMS]) {
    unsigned cnt = 0;
#pragma unroll
    for (int k = 0; k < WIN_ITEMS; ++k) {
        const unsigned long long p = pos + (unsigned long long)threadIdx.x * WIN_ITEMS + k;
        valid[k] = p < limit && record_current(kv, p, slot[k]);
        cnt += valid[k];
    }
    return cnt;
}

__device__ __forceinline__ unsigned long long set_e_ub(const KvView& kv, const uint32_t* n_active_dev) {
    const KvState* st = kv.st;
    const unsigned long long size = st->size;
    const unsigned long long n_absent = n_active_dev ? (unsigned long long)*n_active_dev : st->n_absent;
    return size + n_absent > kv.max_size ? size + n_absent - kv.max_size : 0ull;
}

// B1: tile_cnt[tile] = current records among the tile's positions (before `t`, the batch's first stamp)
__global__ void __launch_bounds__(WIN_THREADS)
kv_set_window_count_kernel(KvView kv, unsigned long long t, const uint32_t* __restrict__ n_active_dev, int deferred,
                           unsigned int* __restrict__ tile_cnt) {
    __shared__ unsigned int s_sum;
    if (deferred == 2 || set_e_ub(kv, n_active_dev) == 0) return;  // uniform for the whole grid
    if (threadIdx.x == 0) s_sum = 0u;
    __syncthreads();
    bool valid[WIN_ITEMS];
    uint32_t slot[WIN_ITEMS];
    unsigned w = win_tile_flags(kv, kv.st->head + blockIdx.x * WIN_TILE, t, valid, slot);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
    if ((threadIdx.x & 31) == 0 && w) atomicAdd(&s_sum, w);
    __syncthreads();
    if (threadIdx.x == 0) tile_cnt[blockIdx.x] = s_sum;
}

// Position of the record with rank k (0-based) among the current records of one tile; ~0 if there is none.
__device__ unsigned long long win_tile_select(const KvView& kv, unsigned long long pos, unsigned long long limit,
                                              unsigned k, unsigned int* s_warp, unsigned long long* s_res) {
    if (threadIdx.x == 0) *s_res = ~0ull;
    __syncthreads();
    bool valid[WIN_ITEMS];
    uint32_t slot[WIN_ITEMS];
    const unsigned cnt = win_tile_flags(kv, pos, limit, valid, slot);
    unsigned excl;
    block_scan_1024(cnt, excl, s_warp);
    unsigned rank = excl;
#pragma unroll
    for (int j = 0; j < WIN_ITEMS; ++j) {
        if (!valid[j]) continue;
        if (rank == k) *s_res = pos + (unsigned long long)threadIdx.x * WIN_ITEMS + j;
        ++rank;
    }
    __syncthreads();
    return *s_res;
}

// B2: one CTA.  tile_cnt becomes its exclusive prefix.
__global__ void __launch_bounds__(WIN_THREADS)
kv_set_window_decide_kernel(KvView kv, unsigned long long t, const uint32_t* __restrict__ n_active_dev, int deferred,
                            int n_tiles, unsigned int* __restrict__ tile_cnt) {
    KvState* st = kv.st;
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned long long s_res, s_first;
    __shared__ int s_tile_e, s_tile_f;
    __shared__ unsigned s_rank_e;
    const int tid = threadIdx.x;
    if (deferred == 2) {  // set(..., batch=True) on its own: entries stay outside the LRU order
        if (tid == 0) {
            st->conflict = 1;
            st->win_mode = 0;
        }
        return;
    }
    const unsigned long long e_ub = set_e_ub(kv, n_active_dev);
    const unsigned long long n_pre = n_active_dev ? 0ull : st->n_pre;
    const unsigned long long min_prev = st->min_prev;
    const unsigned long long head = st->head;
    if (e_ub == 0) {  // conflict stays 0, nothing to evict
        if (tid == 0) st->win_mode = 0;
        return;
    }
    if (tid == 0) {
        s_tile_e = -1;
        s_tile_f = 0x7fffffff;
        s_first = ~0ull;
    }
    __syncthreads();
    unsigned long long base = 0;
    for (int t0 = 0; t0 < n_tiles; t0 += WIN_THREADS) {
        const int tile = t0 + tid;
        const unsigned cnt = tile < n_tiles ? tile_cnt[tile] : 0u;
        unsigned excl;
        const unsigned total = block_scan_1024(cnt, excl, s_warp);
        const unsigned long long pre = base + excl;
        if (tile < n_tiles) {
            tile_cnt[tile] = (unsigned)pre;  // live entries < 2^32
            if (cnt) atomicMin(&s_tile_f, tile);
            if (pre < e_ub && e_ub <= pre + cnt) {
                s_tile_e = tile;
                s_rank_e = (unsigned)(e_ub - 1 - pre);
            }
        }
        base += total;
        __syncthreads();
    }
    const bool covered_all = head + (unsigned long long)n_tiles * WIN_TILE >= t;
    unsigned long long h_ub = ~0ull;
    unsigned long long mode = 1;
    if (base >= e_ub) {
        h_ub = win_tile_select(kv, head + (unsigned long long)s_tile_e * WIN_TILE, t, s_rank_e, s_warp, &s_res) + 1ull;
        __syncthreads();
        if (s_tile_f != 0x7fffffff) {
            const unsigned long long f = win_tile_select(kv, head + (unsigned long long)s_tile_f * WIN_TILE, t, 0u, s_warp, &s_res);
            if (tid == 0) s_first = f;
        }
    } else if (covered_all) {
        mode = 0;  // fewer stamped pre-batch entries than evictions that may be needed: h_ub stays ~0
        if (s_tile_f != 0x7fffffff) {
            const unsigned long long f = win_tile_select(kv, head + (unsigned long long)s_tile_f * WIN_TILE, t, 0u, s_warp, &s_res);
            if (tid == 0) s_first = f;
        }
    } else {
        mode = 2;  // the window was too short for this log: exact one-CTA walk, now and in phase D
        h_ub = lru_window_select(kv, head, t, e_ub, false, nullptr, &s_first);
    }
    __syncthreads();
    if (tid == 0) {
        if (h_ub == ~0ull)
            st->conflict = 1;  // fewer stamped pre-batch entries than evictions that may be needed
        else if (n_pre > 0 && min_prev < h_ub)
            st->conflict = 1;  // an entry this batch re-sets is within reach of its evictions
        st->win_mode = mode;
        st->win_origin = head;
        st->win_hnew = ~0ull;
        if (s_first != ~0ull && s_first > head) st->head = s_first;  // drop leading stale records
    }
}

// set, phase C: parallel find-or-claim + stamp.  No-op when the batch goes to the sequential path.
// have_precheck == 0: every active key is known to be absent (no precheck was run).
__global__ void __launch_bounds__(KV_THREADS)
kv_set_claim_kernel(KvView kv, const uint4* __restrict__ digests, int M, double now, unsigned long long t,
                    int own_values, const uint8_t* __restrict__ skip, int have_precheck,
                    int32_t* __restrict__ op_slot,
                    int32_t* __restrict__ slot_out, unsigned long long* __restrict__ dropped_out) {
    if (kv.st->conflict) return;
    const int i = blockIdx.x * KV_THREADS + threadIdx.x;
    const bool active = i < M && !(skip && skip[i] == 1);
    bool claimed = false, from_empty = false;
    if (active) {
        const u128 key = fp_of(digests[i]);
        long long s = have_precheck ? op_slot[i] : -1;
        const bool pre = s >= 0;
        if (!pre) s = kv_claim_absent(kv, key, claimed, from_empty);
        Slot* sl = &kv.slots[s];
        const unsigned long long stamp = t + (unsigned long long)i;
        const unsigned long long old_stamp = atomicMax(&sl->stamp, stamp);
        const bool first_touch = pre && old_stamp < t;  // set of an existing key = delete + insert
        if (first_touch && dropped_out && !own_values)
            dropped_out[atomicAdd(&kv.st->n_dropped, 1ull)] = ld_u64_cg(&sl->value);
        if (claimed || first_touch) {
            sl->created_at = now;
            sl->last_accessed = now;
            sl->access_count = 0u;
            if (own_values) {  // the table owns the value: a result-row handle, or the slot index
                if (claimed) sl->value = kv.free_rows ? acquire_row(kv) : (unsigned long long)s;
                sl->ttl = kv.default_ttl;
            }
        }
        kv.log[stamp & kv.rmask] = (uint32_t)s;
        op_slot[i] = (int32_t)s;
        if (slot_out) slot_out[i] = (int32_t)s;
    } else if (i < M && slot_out) {
        slot_out[i] = -1;
    }
    block_add(&kv.st->n_new, active && claimed);
    block_add(&kv.st->size, active && claimed);
    block_add(&kv.st->n_empty_claimed, active && claimed && from_empty);
    block_sub(&kv.st->n_tomb, active && claimed && !from_empty);
}

// set, phase C2: the last op that touched a slot (largest stamp) owns its value / ttl.
__global__ void __launch_bounds__(KV_THREADS)
kv_set_finalize_kernel(KvView kv, int M, unsigned long long t, const unsigned long long* __restrict__ values,
                       const double* __restrict__ ttl, const uint8_t* __restrict__ skip,
                       const int32_t* __restrict__ op_slot) {
    if (kv.st->conflict) return;
    const int i = blockIdx.x * KV_THREADS + threadIdx.x;
    if (i >= M || (skip && skip[i] == 1)) return;
    const int32_t s = op_slot[i];
    Slot* sl = &kv.slots[s];
    if (ld_u64_cg(&sl->stamp) != t + (unsigned long long)i) return;
    if (values) sl->value = values[i];
    const double x = ttl ? ttl[i] : NAN;
    sl->ttl = isnan(x) ? kv.default_ttl : x;
}

// Exact sequential replay of a set batch (kvstore.py:104-134, 170-176), one thread.
__device__ void kv_set_sequential(const KvView& kv, const uint4* digests, const unsigned long long* values,
                                  const double* ttl, int M, double now, unsigned long long t, int deferred,
                                  const uint8_t* skip, int32_t* slot_out, unsigned long long* dropped_out) {
    KvState* st = kv.st;
    unsigned long long size = st->size, head = st->head, evictions = st->evictions;
    long long n_tomb = (long long)st->n_tomb;
    unsigned long long n_empty_claimed = st->n_empty_claimed;
    const bool own_values = values == nullptr;
    bool aborted = false;
    int i = 0;
    for (; i < M; ++i) {
        if (skip && skip[i] == 1) {
            if (slot_out) slot_out[i] = -1;
            continue;
        }
        const u128 key = fp_of(digests[i]);
        long long s = kv_find(kv, key);
        if (s >= 0) {  // if key in self.cache: self.delete(key)
            Slot* sl = &kv.slots[s];
            const unsigned long long v = sl->value;
            if (dropped_out && !own_values) dropped_out[st->n_dropped++] = v;
            release_row(kv, v);
            st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
            sl->stamp = 0ull;
            --size;
            ++n_tomb;
        }
        if (size >= kv.max_size) {  // _evict_if_needed
            ++evictions;            // counted before the pop (kvstore.py:87)
            const unsigned long long bound = deferred ? t : t + (unsigned long long)i;
            long long victim = -1;
            while (head < bound) {
                const unsigned long long p = head++;
                uint32_t r;
                if (record_current(kv, p, r)) {
                    victim = (long long)r;
                    break;
                }
            }
            if (victim < 0) {  // next(iter(OrderedDict())) -> StopIteration
                st->status = DIE_ERR_STOP_ITERATION;
                aborted = true;
                break;
            }
            Slot* v = &kv.slots[victim];
            const unsigned long long vv = v->value;
            if (dropped_out) dropped_out[st->n_dropped++] = vv;
            release_row(kv, vv);
            st128(v, u128{FP_TOMB_WORD, FP_TOMB_WORD});
            v->stamp = 0ull;
            --size;
            ++n_tomb;
        }
        unsigned long long cur = key.lo & kv.mask;
        for (;;) {
            const u128 fp = ld_fp(&kv.slots[cur]);
            if (fp_empty(fp)) {
                ++n_empty_claimed;
                break;
            }
            if (fp_tomb(fp)) {
                --n_tomb;
                break;
            }
            cur = (cur + 1) & kv.mask;
        }
        Slot* sl = &kv.slots[cur];
        const double x = ttl ? ttl[i] : NAN;
        sl->created_at = now;
        sl->ttl = isnan(x) ? kv.default_ttl : x;
        sl->value = values ? values[i] : (kv.free_rows ? acquire_row(kv) : cur);
        sl->last_accessed = now;
        sl->access_count = 0u;
        if (deferred) {
            sl->stamp = 0ull;
        } else {
            const unsigned long long stamp = t + (unsigned long long)i;
            sl->stamp = stamp;
            kv.log[stamp & kv.rmask] = (uint32_t)cur;
        }
        st128(sl, key);
        ++size;
        if (slot_out) slot_out[i] = (int32_t)cur;
    }
    if (aborted && slot_out)
        for (int j = i; j < M; ++j) slot_out[j] = -1;
    if (deferred == 1 && !aborted) {  // for key in items: self._update_access(key)
        for (int j = 0; j < M; ++j) {
            if (skip && skip[j] == 1) continue;
            const long long s = kv_find(kv, fp_of(digests[j]));
            if (s < 0) continue;
            const unsigned long long stamp = t + (unsigned long long)j;
            kv.slots[s].stamp = stamp;
            kv.log[stamp & kv.rmask] = (uint32_t)s;
        }
    }
    st->size = size;
    st->head = head;
    st->evictions = evictions;
    st->n_tomb = (unsigned long long)n_tomb;
    st->n_empty_claimed = n_empty_claimed;
    st->slow_batches += 1;
}

// set, phase D1: evict the E least-recently-used entries, tile by tile (tiles and prefixes of phase B; the claims
// only touched records beyond the window's reach, or the batch went to the sequential path).
__global__ void __launch_bounds__(WIN_THREADS)
kv_set_evict_tiles_kernel(KvView kv, unsigned long long t, const unsigned int* __restrict__ tile_pre,
                          unsigned long long* __restrict__ dropped_out) {
    __shared__ unsigned int s_warp[32];
    KvState* st = kv.st;
    if (st->conflict || st->win_mode != 1ull) return;
    const unsigned long long size = st->size;  // not modified by this kernel (phase D2 applies the counters)
    const unsigned long long E = size > kv.max_size ? size - kv.max_size : 0ull;
    const unsigned long long base = tile_pre[blockIdx.x];
    if (E == 0 || base >= E) return;
    const unsigned long long pos = st->win_origin + blockIdx.x * WIN_TILE;
    bool valid[WIN_ITEMS];
    uint32_t slot[WIN_ITEMS];
    const unsigned cnt = win_tile_flags(kv, pos, t, valid, slot);
    unsigned excl;
    block_scan_1024(cnt, excl, s_warp);
    unsigned long long rank = base + excl;
#pragma unroll
    for (int k = 0; k < WIN_ITEMS; ++k) {
        if (!valid[k]) continue;
        if (rank < E) {
            evict_slot(kv, slot[k], dropped_out);
            if (rank == E - 1) st->win_hnew = pos + (unsigned long long)threadIdx.x * WIN_ITEMS + k + 1ull;
        }
        ++rank;
    }
}

// set, phase D2 (one CTA): counters of the evictions, or the one-CTA walk, or the sequential replay.
__global__ void __launch_bounds__(WIN_THREADS)
kv_set_evict_kernel(KvView kv, const uint4* __restrict__ digests, const unsigned long long* __restrict__ values,
                    const double* __restrict__ ttl, int M, double now, unsigned long long t, int deferred,
                    const uint8_t* __restrict__ skip, int32_t* __restrict__ slot_out,
                    unsigned long long* __restrict__ dropped_out) {
    KvState* st = kv.st;
    const unsigned long long conflict = st->conflict, size = st->size, head = st->head, mode = st->win_mode;
    __syncthreads();
    if (conflict) {
        if (threadIdx.x == 0)
            kv_set_sequential(kv, digests, values, ttl, M, now, t, deferred, skip, slot_out, dropped_out);
    } else {
        const unsigned long long E = size > kv.max_size ? size - kv.max_size : 0ull;
        if (E > 0) {
            unsigned long long hnew;
            if (mode == 1ull)
                hnew = st->win_hnew;  // the tiles of phase D1 did the evictions
            else
                hnew = lru_window_select(kv, head, t, E, true, dropped_out, nullptr);
            if (threadIdx.x == 0) {
                st->size = size - E;
                st->evictions += E;
                st->n_tomb += E;
                if (hnew != ~0ull) st->head = hnew;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {  // hand the scratch back clean for the next batch
        st->n_dropped_last = st->n_dropped;
        st->n_pre = st->n_absent = st->n_new = st->conflict = st->n_dropped = 0ull;
        st->min_prev = ~0ull;
        st->win_mode = 0ull;
        st->win_hnew = ~0ull;
    }
}

// ------------------------------------------------------------------------------------------------
// LFU policy (kvstore.py:67-69,79-80,94-98): a set batch is applied op by op by one CTA.  The reference picks its
// victim with min() over the whole _access_count dict (O(n), first key with the smallest count in the order the keys
// were first counted); here the CTA's 1024 threads scan the table for the smallest (count, stamp) pair per eviction.
// New entries start with count 1, i.e. below every entry that was ever read, so a burst of inserts evicts its own
// earlier inserts: the result depends on the order of the ops, which is why nothing here runs in parallel across ops.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
kv_set_lfu_kernel(KvView kv, const uint4* __restrict__ digests, const unsigned long long* __restrict__ values,
                  const double* __restrict__ ttl, int M, double now, unsigned long long t, int deferred,
                  int32_t* __restrict__ slot_out, unsigned long long* __restrict__ dropped_out) {
    __shared__ unsigned long long s_cnt_stamp[32][2];
    __shared__ unsigned int s_slot[32];
    __shared__ int s_need, s_abort;
    __shared__ unsigned long long s_size;
    KvState* st = kv.st;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s_size = st->size;
        s_abort = 0;
    }
    __syncthreads();
    const unsigned long long cap = kv.mask + 1;
    int i = 0;
    for (; i < M; ++i) {
        if (tid == 0) {
            const u128 key = fp_of(digests[i]);
            const long long s = kv_find(kv, key);
            if (s >= 0) {  // if key in self.cache: self.delete(key)
                Slot* sl = &kv.slots[s];
                if (dropped_out && values) dropped_out[st->n_dropped++] = sl->value;
                st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
                sl->stamp = 0ull;
                --s_size;
                st->n_tomb += 1;
            }
            s_need = s_size >= kv.max_size;
            if (s_need) st->evictions += 1;  // counted before min() can raise (kvstore.py:87)
            __threadfence_block();
        }
        __syncthreads();
        if (s_need) {
            unsigned long long best_c = ~0ull, best_s = ~0ull;
            unsigned int best_slot = 0xffffffffu;
            for (unsigned long long o = tid; o < cap; o += 1024) {
                const Slot* sl = &kv.slots[o];
                const u128 fp = ld_fp(sl);
                if (fp_empty(fp) || fp_tomb(fp)) continue;
                const unsigned long long stamp = ld_u64_cg(&sl->stamp);
                if (stamp == 0ull) continue;  // not in _access_count: never a candidate
                const unsigned long long c = (unsigned long long)__ldcg(&sl->lfu_count);
                if (c < best_c || (c == best_c && stamp < best_s)) {
                    best_c = c;
                    best_s = stamp;
                    best_slot = (unsigned int)o;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long c2 = __shfl_xor_sync(0xffffffffu, best_c, o);
                const unsigned long long s2 = __shfl_xor_sync(0xffffffffu, best_s, o);
                const unsigned int sl2 = __shfl_xor_sync(0xffffffffu, best_slot, o);
                if (c2 < best_c || (c2 == best_c && s2 < best_s)) {
                    best_c = c2;
                    best_s = s2;
                    best_slot = sl2;
                }
            }
            if (lane == 0) {
                s_cnt_stamp[warp][0] = best_c;
                s_cnt_stamp[warp][1] = best_s;
                s_slot[warp] = best_slot;
            }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < 32; ++w)
                    if (s_cnt_stamp[w][0] < best_c || (s_cnt_stamp[w][0] == best_c && s_cnt_stamp[w][1] < best_s)) {
                        best_c = s_cnt_stamp[w][0];
                        best_s = s_cnt_stamp[w][1];
                        best_slot = s_slot[w];
                    }
                if (best_slot == 0xffffffffu) {  // min() of an empty _access_count -> ValueError
                    st->status = DIE_ERR_VALUE_ERROR;
                    s_abort = 1;
                } else {
                    Slot* v = &kv.slots[best_slot];
                    if (dropped_out) dropped_out[st->n_dropped++] = v->value;
                    st128(v, u128{FP_TOMB_WORD, FP_TOMB_WORD});
                    v->stamp = 0ull;
                    --s_size;
                    st->n_tomb += 1;
                }
                __threadfence_block();
            }
            __syncthreads();
            if (s_abort) break;
        }
        if (tid == 0) {
            const u128 key = fp_of(digests[i]);
            unsigned long long cur = key.lo & kv.mask;
            for (;;) {
                const u128 fp = ld_fp(&kv.slots[cur]);
                if (fp_empty(fp)) {
                    st->n_empty_claimed += 1;
                    break;
                }
                if (fp_tomb(fp)) {
                    st->n_tomb -= 1;
                    break;
                }
                cur = (cur + 1) & kv.mask;
            }
            Slot* sl = &kv.slots[cur];
            const double x = ttl ? ttl[i] : NAN;
            sl->created_at = now;
            sl->ttl = isnan(x) ? kv.default_ttl : x;
            sl->value = values ? values[i] : cur;
            sl->last_accessed = now;
            sl->access_count = 0u;
            if (deferred) {
                sl->stamp = 0ull;
                sl->lfu_count = 0u;
            } else {  // self._update_access(key): count 0 + 1, appended to _access_count
                const unsigned long long stamp = t + (unsigned long long)i;
                sl->stamp = stamp;
                sl->lfu_count = 1u;
                kv.log[stamp & kv.rmask] = (uint32_t)cur;
            }
            st128(sl, key);
            ++s_size;
            if (slot_out) slot_out[i] = (int32_t)cur;
            __threadfence_block();
        }
        __syncthreads();
    }
    if (tid == 0) {
        if (s_abort && slot_out)
            for (int j = i; j < M; ++j) slot_out[j] = -1;
        if (deferred == 1 && !s_abort) {  // for key in items: self._update_access(key)
            for (int j = 0; j < M; ++j) {
                const long long s = kv_find(kv, fp_of(digests[j]));
                if (s < 0) continue;
                Slot* sl = &kv.slots[s];
                const unsigned long long stamp = t + (unsigned long long)j;
                sl->lfu_count += 1u;
                if (sl->stamp == 0ull) {
                    sl->stamp = stamp;
                    kv.log[stamp & kv.rmask] = (uint32_t)s;
                }
            }
        }
        st->size = s_size;
        st->slow_batches += 1;
        st->n_dropped_last = st->n_dropped;
        st->n_dropped = 0ull;
    }
}

// ------------------------------------------------------------------------------------------------
// maintenance: re-append the current LRU records at the tail of a fresh ring (order kept)
// ------------------------------------------------------------------------------------------------
// Three launches over tiles of WIN_THREADS * WIN_ITEMS log positions, any number of CTAs:
//   count   current records per tile
//   scan    exclusive prefix of the tile counts (one CTA; one value per 4096 positions)
//   move    record r of the old log -> position tail + (current records before it), its slot re-stamped
// Whether position p is current depends only on the stamp of the slot log[p] names, and a slot that is moved
// was current at some other position p' != p, so moving stamps never changes the verdict of another
// position: the passes need no ordering beyond the kernel boundaries, and `move` may recompute the verdicts.
__device__ __forceinline__ unsigned compact_tile_flags(const KvView& kv, unsigned long long pos, unsigned long long tail,
                                                       bool (&valid)[WIN_ITEMS], uint32_t (&slot)[WIN_ITEMS]) {
    unsigned cnt = 0;
#pragma unroll
    for (int k = 0; k < WIN_ITEMS; ++k) {
        const unsigned long long p = pos + (unsigned long long)threadIdx.x * WIN_ITEMS + k;
        valid[k] = p < tail && record_current(kv, p, slot[k]);
        cnt += valid[k];
    }
    return cnt;
}

__global__ void __launch_bounds__(WIN_THREADS)
kv_compact_count_kernel(KvView kv, unsigned long long head, unsigned long long tail, unsigned long long n_tiles,
                        unsigned int* __restrict__ tile_cnt) {
    constexpr unsigned long long TILE = (unsigned long long)WIN_THREADS * WIN_ITEMS;
    for (unsigned long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        bool valid[WIN_ITEMS];
        uint32_t slot[WIN_ITEMS];
        const unsigned cnt = compact_tile_flags(kv, head + t * TILE, tail, valid, slot);
        __shared__ unsigned int s_sum;
        if (threadIdx.x == 0) s_sum = 0u;
        __syncthreads();
        unsigned w = cnt;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if ((threadIdx.x & 31) == 0 && w) atomicAdd(&s_sum, w);
        __syncthreads();
        if (threadIdx.x == 0) tile_cnt[t] = s_sum;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(1024)
kv_compact_scan_kernel(KvView kv, unsigned int* __restrict__ tile_cnt, unsigned long long n_tiles,
                       unsigned long long tail, unsigned long long* __restrict__ n_kept) {
    __shared__ unsigned int s_warp[32];
    unsigned long long base = 0;
    for (unsigned long long t0 = 0; t0 < n_tiles; t0 += 1024) {
        const unsigned long long t = t0 + threadIdx.x;
        const unsigned cnt = t < n_tiles ? tile_cnt[t] : 0u;
        unsigned excl;
        const unsigned total = block_scan_1024(cnt, excl, s_warp);
        // live records never exceed the table's capacity (< 2^32), so 32-bit prefixes are exact
        if (t < n_tiles) tile_cnt[t] = (unsigned)(base + excl);
        base += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *n_kept = base;
        kv.st->head = tail;
    }
}

__global__ void __launch_bounds__(WIN_THREADS)
kv_compact_move_kernel(KvView kv, uint32_t* __restrict__ new_log, unsigned long long head, unsigned long long tail,
                       unsigned long long n_tiles, const unsigned int* __restrict__ tile_base) {
    __shared__ unsigned int s_warp[32];
    constexpr unsigned long long TILE = (unsigned long long)WIN_THREADS * WIN_ITEMS;
    for (unsigned long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        bool valid[WIN_ITEMS];
        uint32_t slot[WIN_ITEMS];
        const unsigned cnt = compact_tile_flags(kv, head + t * TILE, tail, valid, slot);
        unsigned excl;
        block_scan_1024(cnt, excl, s_warp);
        unsigned long long rank = (unsigned long long)tile_base[t] + excl;
#pragma unroll
        for (int k = 0; k < WIN_ITEMS; ++k) {
            if (!valid[k]) continue;
            const unsigned long long np = tail + rank;  // beyond every old position: no aliasing
            new_log[np & kv.rmask] = slot[k];
            kv.slots[slot[k]].stamp = np;
            ++rank;
        }
        __syncthreads();
    }
}

// maintenance: rebuild the table without tombstones
__global__ void __launch_bounds__(KV_THREADS)
kv_rehash_kernel(KvView kv, const Slot* __restrict__ old_slots, unsigned long long old_cap) {
    for (unsigned long long o = blockIdx.x * (unsigned long long)KV_THREADS + threadIdx.x; o < old_cap;
         o += (unsigned long long)gridDim.x * KV_THREADS) {
        const Slot* src = &old_slots[o];
        const u128 key{src->fp_lo, src->fp_hi};
        if (fp_empty(key) || fp_tomb(key)) continue;
        unsigned long long cur = key.lo & kv.mask;
        for (;;) {
            const u128 old = cas128(&kv.slots[cur], u128{0ull, 0ull}, key);
            if (fp_empty(old)) break;
            cur = (cur + 1) & kv.mask;
        }
        Slot* dst = &kv.slots[cur];
        const unsigned long long stamp = src->stamp;
        dst->created_at = src->created_at;
        dst->ttl = src->ttl;
        dst->stamp = stamp;
        dst->value = src->value;
        dst->last_accessed = src->last_accessed;
        dst->access_count = src->access_count;
        dst->lfu_count = src->lfu_count;
        if (stamp) kv.log[stamp & kv.rmask] = (uint32_t)cur;
    }
}

__global__ void kv_state_init_kernel(KvView kv, unsigned long long head, unsigned long long n_rows) {
    KvState* st = kv.st;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->hits = st->misses = st->evictions = 0ull;
        for (int i = 0; i < 16; ++i) st->hit_part[i][0] = st->miss_part[i][0] = 0ull;
        st->size = 0ull;
        st->head = head;
        st->n_tomb = st->n_empty_claimed = 0ull;
        st->free_head = 0ull;
        st->free_tail = n_rows;
        st->n_pre = st->n_absent = st->n_new = st->conflict = st->n_dropped = 0ull;
        st->min_prev = ~0ull;
        st->win_mode = st->win_origin = 0ull;
        st->win_hnew = ~0ull;
    }
    for (unsigned long long r = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; r < n_rows;
         r += (unsigned long long)gridDim.x * blockDim.x)
        kv.free_rows[r] = (uint32_t)r;  // row 0 is popped first
}

static uint64_t next_pow2(uint64_t x) {
    uint64_t p = 1;
    while (p < x) p <<= 1;
    return p;
}

die_status_t kv_create_impl(uint64_t max_size, double default_ttl, uint64_t n_rows, uint64_t min_batch,
                            die_kv** out, int policy) {
    if (max_size > (1ull << 28)) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_kv_create: max_size > 2^28 (slot indices are 32-bit)");
    die_kv* kv = new die_kv();
    memset(kv, 0, sizeof(*kv));
    DIE_CUDA(cudaGetDevice(&kv->device));
    kv->max_size = max_size;
    kv->policy = policy;
    kv->default_ttl = default_ttl;
    // slots per entry (power-of-two capacity >= factor * max_size): 8 keeps the live load at 1/8, so probe chains stay at
    // one or two slots and most evicted slots have an EMPTY successor (they become EMPTY, not tombstones)
    static int factor = 0;
    if (factor == 0) factor = getenv("DIE_KV_SLOTS_PER_ENTRY") ? std::max(2, atoi(getenv("DIE_KV_SLOTS_PER_ENTRY"))) : 8;
    kv->capacity = next_pow2(std::max<uint64_t>(std::max<uint64_t>((uint64_t)factor * max_size, 16 * min_batch), 4096));
    // LRU log: 4 bytes per position.  Positions are handed out per batch *slot* (an owner step of the exchange
    // stamps its whole capacity, not only the requests that arrived), so the log is sparse; 16 positions per
    // entry keeps compaction (kv_maintain) out of the steady state of a table that evicts.
    kv->ring_cap = next_pow2(std::max<uint64_t>(std::max<uint64_t>(16 * max_size, 8 * min_batch), 1ull << 20));
    kv->max_chunk = std::min<uint64_t>(kv->capacity / 16, std::min<uint64_t>(kv->ring_cap / 4, 1ull << 22));
    kv->tail = 1;
    kv->head_lb = 1;
    kv->n_rows = n_rows;
    DIE_CUDA(cudaMalloc(&kv->slots, kv->capacity * sizeof(Slot)));
    DIE_CUDA(cudaMalloc(&kv->log, kv->ring_cap * sizeof(uint32_t)));
    DIE_CUDA(cudaMalloc(&kv->st, sizeof(KvState)));
    kv->row_ring = n_rows ? next_pow2(n_rows) : 0;
    if (n_rows) DIE_CUDA(cudaMalloc(&kv->free_rows, kv->row_ring * sizeof(uint32_t)));
    DIE_CUDA(cudaMemset(kv->slots, 0, kv->capacity * sizeof(Slot)));
    DIE_CUDA(cudaMemset(kv->log, 0xff, kv->ring_cap * sizeof(uint32_t)));
    DIE_CUDA(cudaMemset(kv->st, 0, sizeof(KvState)));
    DIE_CUDA(cudaMalloc(&kv->n_kept_dev, sizeof(unsigned long long)));
    DIE_CUDA(cudaMalloc(&kv->tile_cnt, (kv->ring_cap / ((uint64_t)WIN_THREADS * WIN_ITEMS) + 2) * sizeof(unsigned int)));
    DIE_CUDA(cudaHostAlloc(&kv->mirror, sizeof(KvMirror), cudaHostAllocMapped | cudaHostAllocPortable));
    memset(kv->mirror, 0, sizeof(KvMirror));
    kv->mirror_gen = 1;
    DIE_LAUNCH(ST_KV_MAINT, (cudaStream_t)0, kv_state_init_kernel<<<n_rows ? 256 : 1, 256>>>(view_of(kv), 1ull, n_rows));
    DIE_CUDA(cudaDeviceSynchronize());
    *out = kv;
    return DIE_OK;
}

// Tiles of the victim window of a set batch of n ops (WIN_FACTOR log positions per op, at least one tile).
static int set_window_tiles(const die_kv* kv, int64_t n) {
    const uint64_t cap = kv->ring_cap / WIN_TILE + 1;
    uint64_t tiles = ((uint64_t)WIN_FACTOR * (uint64_t)std::max<int64_t>(n, 1) + WIN_TILE - 1) / WIN_TILE;
    return (int)std::max<uint64_t>(1, std::min<uint64_t>(tiles, cap));
}

static die_status_t ensure_scratch(die_kv* kv, uint64_t n) {
    if (n <= kv->op_cap) return DIE_OK;
    const uint64_t cap = next_pow2(std::max<uint64_t>(n, 1024));
    if (kv->op_slot) DIE_CUDA(cudaFree(kv->op_slot));
    kv->op_slot = nullptr;
    kv->op_cap = 0;
    DIE_CUDA(cudaMalloc(&kv->op_slot, cap * sizeof(int32_t)));
    kv->op_cap = cap;
    return DIE_OK;
}

static die_status_t kv_maintain(die_kv* kv, int force, cudaStream_t s) {
    KvState h;
    DIE_CUDA(cudaMemcpyAsync(&h, kv->st, sizeof(KvState), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    kv->head_lb = h.head;
    const uint64_t span = kv->tail - h.head;
    // compaction walks the whole span in one CTA (milliseconds for millions of records): only when the ring is
    // really filling up.  A table that evicts consumes its log as fast as it appends and never gets here.
    if (force || span > kv->ring_cap / 4 * 3) {
        uint32_t* new_log = kv->log_spare;  // kept between compactions: no cudaMalloc / cudaFree on the serving path
        unsigned long long* n_kept = kv->n_kept_dev;
        if (!new_log) DIE_CUDA(cudaMalloc(&new_log, kv->ring_cap * sizeof(uint32_t)));
        DIE_CUDA(cudaMemsetAsync(new_log, 0xff, kv->ring_cap * sizeof(uint32_t), s));
        const unsigned long long n_tiles = (span + (unsigned long long)WIN_THREADS * WIN_ITEMS - 1) / ((unsigned long long)WIN_THREADS * WIN_ITEMS);
        if (n_tiles) {
            const int grid = (int)std::min<unsigned long long>(n_tiles, (unsigned long long)num_sms() * 2);
            DIE_LAUNCH(ST_KV_MAINT, s, kv_compact_count_kernel<<<grid, WIN_THREADS, 0, s>>>(view_of(kv), h.head, kv->tail, n_tiles, kv->tile_cnt));
            DIE_LAUNCH(ST_KV_MAINT, s, kv_compact_scan_kernel<<<1, 1024, 0, s>>>(view_of(kv), kv->tile_cnt, n_tiles, kv->tail, n_kept));
            DIE_LAUNCH(ST_KV_MAINT, s, kv_compact_move_kernel<<<grid, WIN_THREADS, 0, s>>>(view_of(kv), new_log, h.head, kv->tail, n_tiles, kv->tile_cnt));
        } else {
            DIE_LAUNCH(ST_KV_MAINT, s, kv_compact_scan_kernel<<<1, 1024, 0, s>>>(view_of(kv), kv->tile_cnt, 0ull, kv->tail, n_kept));
        }
        unsigned long long kept = 0;
        DIE_CUDA(cudaMemcpyAsync(&kept, n_kept, sizeof(kept), cudaMemcpyDeviceToHost, s));
        DIE_CUDA(cudaStreamSynchronize(s));
        kv->log_spare = kv->log;
        kv->log = new_log;
        kv->head_lb = kv->tail;
        kv->tail += kept;
    }
    // probe chains stay short while used slots (live + tombstones) stay under half of the capacity
    if (force || h.size + h.n_tomb + kv->max_chunk > kv->capacity / 2) {
        Slot* fresh = kv->slots_spare;
        if (!fresh) DIE_CUDA(cudaMalloc(&fresh, kv->capacity * sizeof(Slot)));
        DIE_CUDA(cudaMemsetAsync(fresh, 0, kv->capacity * sizeof(Slot), s));
        Slot* old = kv->slots;
        kv->slots = fresh;
        DIE_LAUNCH(ST_KV_MAINT, s, kv_rehash_kernel<<<num_sms() * 8, KV_THREADS, 0, s>>>(view_of(kv), old, kv->capacity));
        h.n_tomb = 0;
        h.n_empty_claimed = h.size;
        DIE_CUDA(cudaMemcpyAsync(&kv->st->n_tomb, &h.n_tomb, sizeof(h.n_tomb), cudaMemcpyHostToDevice, s));
        DIE_CUDA(cudaMemcpyAsync(&kv->st->n_empty_claimed, &h.n_empty_claimed, sizeof(h.n_empty_claimed),
                                 cudaMemcpyHostToDevice, s));
        DIE_CUDA(cudaStreamSynchronize(s));
        kv->slots_spare = old;  // kept for the next rebuild: no cudaMalloc/cudaFree on the serving path
        kv->live_at_rebuild = h.size;
    } else {
        kv->live_at_rebuild = h.size + h.n_tomb;  // tighten the conservative bound with the real count
    }
    kv->claims_ub = 0;
    kv->mirror_gen += 1;
    return DIE_OK;
}

KvMirrorArgs kv_mirror_args(die_kv* kv) {
    KvMirrorArgs a;
    a.st = kv->st;
    a.m = kv->mirror;
    a.ops_seen = kv->ops_enq;
    a.gen = kv->mirror_gen;
    a.seq = ++kv->mirror_seq;
    return a;
}

// Tighten the host bounds from the device's latest snapshot, if one of this table generation is there.
static void kv_mirror_refresh(die_kv* kv) {
    const volatile KvMirror* m = kv->mirror;
    if (!m) return;
    for (int attempt = 0; attempt < 4; ++attempt) {
        const unsigned long long s1 = m->seq1;
        std::atomic_thread_fence(std::memory_order_acquire);
        const unsigned long long head = m->head, size = m->size, tomb = m->n_tomb, seen = m->ops_seen, gen = m->gen;
        std::atomic_thread_fence(std::memory_order_acquire);
        if (m->seq0 != s1) continue;  // torn: the device is writing
        if (s1 == 0 || gen != kv->mirror_gen) return;
        if (head > kv->head_lb) kv->head_lb = head;  // the head only moves forward between compactions
        const uint64_t used = size + tomb, since = kv->ops_enq - seen;
        if (used + since < kv->live_at_rebuild + kv->claims_ub) {
            kv->live_at_rebuild = used;
            kv->claims_ub = since;
        }
        return;
    }
}

// Called before a batch of n ops: keeps the ring and the EMPTY-slot reserve safe using host bounds only.
die_status_t kv_reserve(die_kv* kv, uint64_t n, bool inserts, cudaStream_t s) {
    kv_mirror_refresh(kv);
    bool need = (kv->tail + n - kv->head_lb) > kv->ring_cap - 16;
    if (inserts && kv->live_at_rebuild + kv->claims_ub + n > kv->capacity / 2) need = true;
    if (need) DIE_TRY(kv_maintain(kv, 0, s));
    if ((kv->tail + n - kv->head_lb) > kv->ring_cap - 16) DIE_TRY(kv_maintain(kv, 1, s));
    if (inserts) {
        kv->claims_ub += n;
        kv->ops_enq += n;
    }
    return DIE_OK;
}

die_status_t kv_get_batch_impl(die_kv* kv, const uint32_t* digests, int32_t B, double now, uint8_t* hit_out,
                               uint64_t* value_out, int32_t* slot_out, cudaStream_t s) {
    for (int64_t done = 0; done < B;) {
        const int n = (int)std::min<int64_t>(B - done, (int64_t)kv->max_chunk);
        DIE_TRY(kv_reserve(kv, n, false, s));
        DIE_LAUNCH(ST_KV_GET, s, kv_get_kernel<false><<<ceil_div(n, KV_THREADS), KV_THREADS, 0, s>>>(
            view_of(kv), reinterpret_cast<const uint4*>(digests) + done, n, now, kv->tail, hit_out + done,
            reinterpret_cast<unsigned long long*>(value_out ? value_out + done : nullptr),
            slot_out ? slot_out + done : nullptr));
        kv->tail += n;
        done += n;
    }
    return DIE_OK;
}

// LFU: the whole batch in one launch of the op-by-op kernel (no scratch, no housekeeping beyond kv_reserve's).
static die_status_t kv_set_lfu(die_kv* kv, const uint32_t* digests, const uint64_t* values, const double* ttl, int32_t M,
                               double now, int32_t deferred, int32_t* slot_out, uint64_t* dropped_out, cudaStream_t s) {
    if ((int64_t)M > (int64_t)kv->max_chunk)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "kv_set_batch (lfu): %d ops exceed this table's per-call limit of %llu", M,
                 (unsigned long long)kv->max_chunk);
    kv->dropped_prefix = 0;
    DIE_LAUNCH(ST_KV_EVICT, s, kv_set_lfu_kernel<<<1, 1024, 0, s>>>(
        view_of(kv), reinterpret_cast<const uint4*>(digests), reinterpret_cast<const unsigned long long*>(values), ttl, M, now,
        kv->tail, deferred, slot_out, reinterpret_cast<unsigned long long*>(dropped_out)));
    kv->tail += (uint64_t)M;
    return DIE_OK;
}

die_status_t kv_set_batch_impl(die_kv* kv, const uint32_t* digests, const uint64_t* values, const double* ttl,
                               int32_t M, double now, int32_t deferred, const uint8_t* skip, int known_absent,
                               const uint32_t* n_active_dev, int32_t* slot_out, uint64_t* dropped_out,
                               cudaStream_t s) {
    if (kv->policy == DIE_KV_POLICY_FIFO) deferred = 0;  // the eviction order is the insertion order: stamped at insert
    if (kv->policy == DIE_KV_POLICY_LFU) {
        if (skip || known_absent) DIE_FAIL(DIE_ERR_UNSUPPORTED, "kv_set_batch: the composed pipeline runs on LRU tables");
        DIE_TRY(kv_reserve(kv, (uint64_t)M, true, s));
        return kv_set_lfu(kv, digests, values, ttl, M, now, deferred, slot_out, dropped_out, s);
    }
    if (known_absent && ((int64_t)M > (int64_t)kv->max_chunk || !n_active_dev))
        DIE_FAIL(DIE_ERR_TOO_LARGE, "kv_set_batch: known_absent batches must fit one chunk (%d > %llu)", M,
                 (unsigned long long)kv->max_chunk);
    if (deferred && (int64_t)M > (int64_t)kv->max_chunk)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "kv_set_batch: a batch_set of %d items exceeds this table's limit of %llu", M,
                 (unsigned long long)kv->max_chunk);
    uint64_t dropped_base = 0;
    kv->dropped_prefix = 0;
    for (int64_t done = 0; done < M;) {
        const int n = (int)std::min<int64_t>(M - done, (int64_t)kv->max_chunk);
        DIE_TRY(ensure_scratch(kv, n));
        DIE_TRY(kv_reserve(kv, n, true, s));
        const KvView v = view_of(kv);
        const uint4* dg = reinterpret_cast<const uint4*>(digests) + done;
        const unsigned long long* vals =
            reinterpret_cast<const unsigned long long*>(values ? values + done : nullptr);
        const double* tt = ttl ? ttl + done : nullptr;
        const uint8_t* sk = skip ? skip + done : nullptr;
        int32_t* so = slot_out ? slot_out + done : nullptr;
        unsigned long long* dr =
            reinterpret_cast<unsigned long long*>(dropped_out ? dropped_out + dropped_base : nullptr);
        const int grid = ceil_div(n, KV_THREADS);
        const int own_values = values ? 0 : 1;
        if (!known_absent) {
            DIE_LAUNCH(ST_KV_PRECHECK, s, kv_set_precheck_kernel<<<grid, KV_THREADS, 0, s>>>(v, dg, n, sk, kv->op_slot));
        }
        const int n_tiles = set_window_tiles(kv, n);
        const uint32_t* n_act = known_absent ? n_active_dev : nullptr;
        DIE_LAUNCH(ST_KV_WINDOW, s, kv_set_window_count_kernel<<<n_tiles, WIN_THREADS, 0, s>>>(v, kv->tail, n_act, deferred, kv->tile_cnt));
        DIE_LAUNCH(ST_KV_WINDOW, s, kv_set_window_decide_kernel<<<1, WIN_THREADS, 0, s>>>(v, kv->tail, n_act, deferred, n_tiles, kv->tile_cnt));
        DIE_LAUNCH(ST_KV_CLAIM, s, kv_set_claim_kernel<<<grid, KV_THREADS, 0, s>>>(v, dg, n, now, kv->tail, own_values, sk,
                                                        known_absent ? 0 : 1, kv->op_slot, so, dr));
        if (values || ttl) {
            DIE_LAUNCH(ST_KV_FINALIZE, s, kv_set_finalize_kernel<<<grid, KV_THREADS, 0, s>>>(v, n, kv->tail, vals, tt, sk, kv->op_slot));
        }
        DIE_LAUNCH(ST_KV_EVICT, s, kv_set_evict_tiles_kernel<<<n_tiles, WIN_THREADS, 0, s>>>(v, kv->tail, kv->tile_cnt, dr));
        DIE_LAUNCH(ST_KV_EVICT, s, kv_set_evict_kernel<<<1, WIN_THREADS, 0, s>>>(v, dg, vals, tt, n, now, kv->tail, deferred, sk, so, dr));
        kv->tail += n;
        done += n;
        if (done < M) {
            // chunked batches are rare (tiny tables / huge batches): keep the dropped list contiguous
            // and stop where the reference would have raised
            KvState h;
            DIE_CUDA(cudaMemcpyAsync(&h, kv->st, sizeof(KvState), cudaMemcpyDeviceToHost, s));
            DIE_CUDA(cudaStreamSynchronize(s));
            dropped_base += h.n_dropped_last;
            kv->dropped_prefix = dropped_base;
            if (h.status) {
                if (slot_out) DIE_CUDA(cudaMemsetAsync(slot_out + done, 0xff, (size_t)(M - done) * sizeof(int32_t), s));
                break;
            }
        }
    }
    return DIE_OK;
}

}  // namespace die

// ================================================================================================
// C-ABI
// ================================================================================================
using namespace die;

extern "C" die_status_t die_kv_create(uint64_t max_size, int32_t policy, double default_ttl, die_kv_t** out) {
    if (!out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_create: null out");
    if (policy != DIE_KV_POLICY_LRU && policy != DIE_KV_POLICY_LFU && policy != DIE_KV_POLICY_FIFO)
        DIE_FAIL(DIE_ERR_INVALID, "die_kv_create: unknown eviction policy %d", policy);
    return kv_create_impl(max_size, default_ttl, 0, 0, out, policy);
}

extern "C" die_status_t die_kv_destroy(die_kv_t* kv) {
    if (!kv) return DIE_OK;
    cudaFree(kv->slots);
    cudaFree(kv->slots_spare);
    cudaFree(kv->log);
    cudaFree(kv->log_spare);
    cudaFree(kv->tile_cnt);
    cudaFree(kv->n_kept_dev);
    if (kv->mirror) cudaFreeHost(kv->mirror);
    cudaFree(kv->st);
    cudaFree(kv->op_slot);
    cudaFree(kv->free_rows);
    delete kv;
    return DIE_OK;
}

extern "C" die_status_t die_kv_maintain_sync(die_kv_t* kv, int32_t force, die_stream_t stream) {
    if (!kv) DIE_FAIL(DIE_ERR_INVALID, "die_kv_maintain_sync: null kv");
    return kv_maintain(kv, force, static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_kv_get_batch(die_kv_t* kv, const uint32_t* digests, int32_t B, double now,
                                         uint8_t* hit_out, uint64_t* value_out, int32_t* slot_out,
                                         die_stream_t stream) {
    if (!kv || B < 0) DIE_FAIL(DIE_ERR_INVALID, "die_kv_get_batch: bad arguments");
    if (B == 0) return DIE_OK;
    if (!digests || !hit_out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_get_batch: null digests/hit_out");
    return kv_get_batch_impl(kv, digests, B, now, hit_out, value_out, slot_out, static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_kv_contains_batch(die_kv_t* kv, const uint32_t* digests, int32_t B, double now,
                                              uint8_t* out, uint64_t* value_out, die_stream_t stream) {
    if (!kv || B < 0) DIE_FAIL(DIE_ERR_INVALID, "die_kv_contains_batch: bad arguments");
    if (B == 0) return DIE_OK;
    if (!digests || !out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_contains_batch: null digests/out");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_LAUNCH(ST_KV_CONTAINS, s, kv_get_kernel<true><<<ceil_div(B, KV_THREADS), KV_THREADS, 0, s>>>(
        view_of(kv), reinterpret_cast<const uint4*>(digests), B, now, 0ull, out,
        reinterpret_cast<unsigned long long*>(value_out), nullptr));
    return DIE_OK;
}

extern "C" die_status_t die_kv_delete_batch(die_kv_t* kv, const uint32_t* digests, int32_t B,
                                            uint8_t* deleted_out, uint64_t* value_out, die_stream_t stream) {
    if (!kv || B < 0) DIE_FAIL(DIE_ERR_INVALID, "die_kv_delete_batch: bad arguments");
    if (B == 0) return DIE_OK;
    if (!digests || !deleted_out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_delete_batch: null digests/deleted_out");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_LAUNCH(ST_KV_DELETE, s, kv_delete_kernel<<<ceil_div(B, KV_THREADS), KV_THREADS, 0, s>>>(
        view_of(kv), reinterpret_cast<const uint4*>(digests), B, deleted_out,
        reinterpret_cast<unsigned long long*>(value_out)));
    return DIE_OK;
}

extern "C" die_status_t die_kv_set_batch(die_kv_t* kv, const uint32_t* digests, const uint64_t* values,
                                         const double* ttl, int32_t M, double now, int32_t deferred_stamp,
                                         int32_t* slot_out, uint64_t* dropped_out, die_stream_t stream) {
    if (!kv || M < 0) DIE_FAIL(DIE_ERR_INVALID, "die_kv_set_batch: bad arguments");
    if (M == 0) return DIE_OK;
    if (!digests) DIE_FAIL(DIE_ERR_INVALID, "die_kv_set_batch: null digests");
    return kv_set_batch_impl(kv, digests, values, ttl, M, now, deferred_stamp, nullptr, 0, nullptr, slot_out,
                             dropped_out, static_cast<cudaStream_t>(stream));
}

extern "C" die_status_t die_kv_len_sync(die_kv_t* kv, double now, uint64_t* purged_out, uint64_t purged_cap,
                                        uint64_t* n_purged, uint64_t* len_out, die_stream_t stream) {
    if (!kv || !len_out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_len_sync: bad arguments");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned long long* d_n = kv->n_kept_dev;  // the handle's one-word device counter (also used by compaction)
    DIE_CUDA(cudaMemsetAsync(d_n, 0, sizeof(unsigned long long), s));
    DIE_LAUNCH(ST_KV_PURGE, s, kv_purge_kernel<<<num_sms() * 8, KV_THREADS, 0, s>>>(view_of(kv), now,
                                                         reinterpret_cast<unsigned long long*>(purged_out),
                                                         purged_cap, d_n));
    unsigned long long n = 0, size = 0;
    DIE_CUDA(cudaMemcpyAsync(&n, d_n, sizeof(n), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaMemcpyAsync(&size, &kv->st->size, sizeof(size), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    if (n_purged) *n_purged = n;
    *len_out = size;
    return DIE_OK;
}

extern "C" die_status_t die_kv_clear(die_kv_t* kv, die_stream_t stream) {
    if (!kv) DIE_FAIL(DIE_ERR_INVALID, "die_kv_clear: null kv");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_CUDA(cudaMemsetAsync(kv->slots, 0, kv->capacity * sizeof(Slot), s));
    DIE_LAUNCH(ST_KV_MAINT, s, kv_state_init_kernel<<<kv->n_rows ? 256 : 1, 256, 0, s>>>(view_of(kv), kv->tail, kv->n_rows));
    kv->head_lb = kv->tail;
    kv->claims_ub = 0;
    kv->live_at_rebuild = 0;
    return DIE_OK;
}

extern "C" die_status_t die_kv_stats_sync(die_kv_t* kv, uint64_t out[8], die_stream_t stream) {
    if (!kv || !out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_stats_sync: bad arguments");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    KvState h;
    DIE_CUDA(cudaMemcpyAsync(&h, kv->st, sizeof(KvState), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    kv->head_lb = h.head;
    out[0] = h.hits;
    out[1] = h.misses;
    for (int i = 0; i < 16; ++i) {
        out[0] += h.hit_part[i][0];
        out[1] += h.miss_part[i][0];
    }
    out[2] = h.evictions;
    out[3] = h.size;
    out[4] = kv->max_size;
    out[5] = kv->capacity;
    out[6] = h.n_tomb;
    out[7] = kv->tail - h.head;
    return DIE_OK;
}

extern "C" die_status_t die_kv_sync_status(die_kv_t* kv, int32_t* status, uint64_t* n_dropped,
                                           uint32_t* slow_path_batches, die_stream_t stream) {
    if (!kv) DIE_FAIL(DIE_ERR_INVALID, "die_kv_sync_status: null kv");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    KvState h;
    DIE_CUDA(cudaMemcpyAsync(&h, kv->st, sizeof(KvState), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    kv->head_lb = h.head;
    if (status) *status = (int32_t)h.status;
    if (n_dropped) *n_dropped = kv->dropped_prefix + h.n_dropped_last;
    if (slow_path_batches) *slow_path_batches = (uint32_t)h.slow_batches;
    if (h.status) {
        const unsigned long long zero = 0;
        DIE_CUDA(cudaMemcpyAsync(&kv->st->status, &zero, sizeof(zero), cudaMemcpyHostToDevice, s));
        DIE_CUDA(cudaStreamSynchronize(s));
    }
    return DIE_OK;
}

extern "C" die_status_t die_kv_dump_sync(die_kv_t* kv, uint64_t cap, uint32_t* digests, uint64_t* values,
                                         double* created_at, double* last_accessed, uint32_t* access_count,
                                         double* ttl, uint8_t* stamped, uint64_t* n_out, die_stream_t stream,
                                         uint32_t* lfu_count) {
    if (!kv || !n_out) DIE_FAIL(DIE_ERR_INVALID, "die_kv_dump_sync: bad arguments");
    if (kv->capacity > (1ull << 24)) DIE_FAIL(DIE_ERR_TOO_LARGE, "die_kv_dump_sync: parity dump is for small tables");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    std::vector<Slot> host(kv->capacity);
    DIE_CUDA(cudaMemcpyAsync(host.data(), kv->slots, kv->capacity * sizeof(Slot), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    std::vector<const Slot*> live;
    for (const Slot& sl : host) {
        const bool empty = sl.fp_lo == 0 && sl.fp_hi == 0;
        const bool tomb = sl.fp_lo == ~0ull && sl.fp_hi == ~0ull;
        if (!empty && !tomb) live.push_back(&sl);
    }
    std::stable_sort(live.begin(), live.end(), [](const Slot* a, const Slot* b) {
        const unsigned long long sa = a->stamp ? a->stamp : ~0ull, sb = b->stamp ? b->stamp : ~0ull;
        return sa < sb;
    });
    *n_out = live.size();
    for (uint64_t k = 0; k < live.size() && k < cap; ++k) {
        const Slot* sl = live[k];
        if (digests) {
            digests[4 * k + 0] = (uint32_t)sl->fp_lo;
            digests[4 * k + 1] = (uint32_t)(sl->fp_lo >> 32);
            digests[4 * k + 2] = (uint32_t)sl->fp_hi;
            digests[4 * k + 3] = (uint32_t)(sl->fp_hi >> 32);
        }
        if (values) values[k] = sl->value;
        if (created_at) created_at[k] = sl->created_at;
        if (last_accessed) last_accessed[k] = sl->last_accessed;
        if (access_count) access_count[k] = sl->access_count;
        if (ttl) ttl[k] = sl->ttl;
        if (stamped) stamped[k] = sl->stamp ? 1 : 0;
        if (lfu_count) lfu_count[k] = sl->lfu_count;
    }
    return DIE_OK;
}

// ------------------------------------------------------------------------------------------------
// Snapshot / restore (the reference's README.md:14 promises "optional persistence" and never wrote it): the table as
// one flat image -- header, counters, slots.  The LRU log is not stored: it is rebuilt from the slots' stamps.
// ------------------------------------------------------------------------------------------------
namespace die {
struct KvImageHeader {
    uint64_t magic, version, max_size, capacity, ring_cap, policy, tail;
    double default_ttl;
    KvState st;
};
constexpr uint64_t KV_IMAGE_MAGIC = 0x314d494b45494444ull;  // "DDIEKIM1"

__global__ void __launch_bounds__(KV_THREADS)
kv_relog_kernel(KvView kv) {  // log[stamp] = slot for every entry that is in the eviction order
    const unsigned long long cap = kv.mask + 1;
    for (unsigned long long o = blockIdx.x * (unsigned long long)KV_THREADS + threadIdx.x; o < cap;
         o += (unsigned long long)gridDim.x * KV_THREADS) {
        const Slot* sl = &kv.slots[o];
        const u128 fp{sl->fp_lo, sl->fp_hi};
        if (fp_empty(fp) || fp_tomb(fp)) continue;
        if (sl->stamp) kv.log[sl->stamp & kv.rmask] = (uint32_t)o;
    }
}
}  // namespace die

extern "C" uint64_t die_kv_snapshot_bytes(die_kv_t* kv) {
    return kv ? sizeof(KvImageHeader) + kv->capacity * sizeof(Slot) : 0;
}

extern "C" die_status_t die_kv_snapshot_sync(die_kv_t* kv, void* image, uint64_t image_bytes, die_stream_t stream) {
    if (!kv || !image) DIE_FAIL(DIE_ERR_INVALID, "die_kv_snapshot_sync: null argument");
    if (kv->free_rows) DIE_FAIL(DIE_ERR_UNSUPPORTED, "die_kv_snapshot_sync: a pipeline-owned table holds result-row handles, not values");
    if (image_bytes < die_kv_snapshot_bytes(kv))
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_kv_snapshot_sync: the image needs %llu bytes", (unsigned long long)die_kv_snapshot_bytes(kv));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    KvImageHeader* h = static_cast<KvImageHeader*>(image);
    h->magic = KV_IMAGE_MAGIC;
    h->version = 1;
    h->max_size = kv->max_size;
    h->capacity = kv->capacity;
    h->ring_cap = kv->ring_cap;
    h->policy = (uint64_t)kv->policy;
    h->tail = kv->tail;
    h->default_ttl = kv->default_ttl;
    DIE_CUDA(cudaMemcpyAsync(&h->st, kv->st, sizeof(KvState), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaMemcpyAsync(h + 1, kv->slots, kv->capacity * sizeof(Slot), cudaMemcpyDeviceToHost, s));
    DIE_CUDA(cudaStreamSynchronize(s));
    return DIE_OK;
}

extern "C" die_status_t die_kv_restore_sync(die_kv_t* kv, const void* image, uint64_t image_bytes, die_stream_t stream) {
    if (!kv || !image) DIE_FAIL(DIE_ERR_INVALID, "die_kv_restore_sync: null argument");
    if (kv->free_rows) DIE_FAIL(DIE_ERR_UNSUPPORTED, "die_kv_restore_sync: not for pipeline-owned tables");
    const KvImageHeader* h = static_cast<const KvImageHeader*>(image);
    if (image_bytes < sizeof(KvImageHeader) || h->magic != KV_IMAGE_MAGIC || h->version != 1)
        DIE_FAIL(DIE_ERR_INVALID, "die_kv_restore_sync: not a table image");
    if (h->max_size != kv->max_size || h->capacity != kv->capacity || h->ring_cap != kv->ring_cap ||
        h->policy != (uint64_t)kv->policy || image_bytes < sizeof(KvImageHeader) + kv->capacity * sizeof(Slot))
        DIE_FAIL(DIE_ERR_INVALID, "die_kv_restore_sync: the image was taken from a table of another shape "
                                  "(max_size %llu, policy %llu)", (unsigned long long)h->max_size, (unsigned long long)h->policy);
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_CUDA(cudaMemcpyAsync(kv->slots, h + 1, kv->capacity * sizeof(Slot), cudaMemcpyHostToDevice, s));
    DIE_CUDA(cudaMemcpyAsync(kv->st, &h->st, sizeof(KvState), cudaMemcpyHostToDevice, s));
    DIE_CUDA(cudaMemsetAsync(kv->log, 0xff, kv->ring_cap * sizeof(uint32_t), s));
    DIE_LAUNCH(ST_KV_MAINT, s, kv_relog_kernel<<<num_sms() * 8, KV_THREADS, 0, s>>>(view_of(kv)));
    DIE_CUDA(cudaStreamSynchronize(s));
    kv->tail = h->tail;
    kv->default_ttl = h->default_ttl;
    kv->head_lb = h->st.head;
    kv->live_at_rebuild = h->st.size + h->st.n_tomb;
    kv->claims_ub = 0;
    kv->dropped_prefix = 0;
    kv->mirror_gen += 1;  // snapshots of the table that was here describe something else
    return DIE_OK;
}

// ================================================================================================
// die_kv_set_batch_noalloc: the same batched set, for hosts that must not touch the CUDA allocator
// ================================================================================================
// A serving loop that has reserved all of its device memory up front, code running inside a
// latency-critical section where cudaMalloc's implicit device-wide synchronisation is not acceptable,
// or an embedding runtime that simply forbids allocation after start-up cannot use die_kv_set_batch:
// that entry point grows its per-op scratch with cudaMalloc and may rebuild the table.  This variant
// takes its per-op scratch (one int32 slot index per op) from a fixed array that lives in the
// library's device image, so every GPU that loads libdie.so carries 256 KiB for it until the
// process exits.  Inputs and outputs are still the caller's buffers.
namespace die {
constexpr int32_t KV_NOALLOC_MAX_OPS = 65536;  // DIE_KV_NOALLOC_MAX_OPS in include/die.h
static __device__ __align__(256) int32_t g_noalloc_op_slot[KV_NOALLOC_MAX_OPS];
constexpr int KV_NOALLOC_MAX_DEVICES = 64;
// Device address of g_noalloc_op_slot on each device.  The symbol's name is only a host-side stand-in,
// never a device pointer: the address comes from cudaGetSymbolAddress, once per device.
static int32_t* g_noalloc_scratch[KV_NOALLOC_MAX_DEVICES] = {nullptr};
}  // namespace die

extern "C" die_status_t die_kv_set_batch_noalloc(die_kv_t* kv, const uint32_t* digests, const uint64_t* values,
                                                 const double* ttl, int32_t M, double now, int32_t deferred_stamp,
                                                 int32_t* slot_out, uint64_t* dropped_out, die_stream_t stream) {
    if (!kv || M < 0) DIE_FAIL(DIE_ERR_INVALID, "die_kv_set_batch_noalloc: bad arguments");
    if (M == 0) return DIE_OK;
    if (!digests) DIE_FAIL(DIE_ERR_INVALID, "die_kv_set_batch_noalloc: null digests");
    if (M > KV_NOALLOC_MAX_OPS)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_kv_set_batch_noalloc: %d ops exceed the fixed scratch of %d ops; split the batch",
                 M, KV_NOALLOC_MAX_OPS);
    if ((uint64_t)M > kv->max_chunk)
        DIE_FAIL(DIE_ERR_TOO_LARGE, "die_kv_set_batch_noalloc: %d ops exceed this table's per-launch limit of %llu", M,
                 (unsigned long long)kv->max_chunk);
    if (kv->policy == DIE_KV_POLICY_FIFO) deferred_stamp = 0;
    int dev = -1;
    DIE_CUDA(cudaGetDevice(&dev));
    if (dev != kv->device || dev < 0 || dev >= KV_NOALLOC_MAX_DEVICES)
        DIE_FAIL(DIE_ERR_INVALID, "die_kv_set_batch_noalloc: table lives on device %d, current device is %d", kv->device, dev);
    if (!g_noalloc_scratch[dev]) {  // one-time set-up for this device (do it before any stream capture)
        void* p = nullptr;
        DIE_CUDA(cudaGetSymbolAddress(&p, g_noalloc_op_slot));
        g_noalloc_scratch[dev] = static_cast<int32_t*>(p);
    }
    int32_t* scratch = g_noalloc_scratch[dev];
    // Housekeeping (LRU-ring compaction, table rebuild) allocates, so it is never started from here: when the
    // conservative host-side bounds say it is due, the caller is told to run it where allocation is allowed.
    const uint64_t n = (uint64_t)M;
    if ((kv->tail + n - kv->head_lb) > kv->ring_cap - 16 ||
        kv->live_at_rebuild + kv->claims_ub + n > kv->capacity / 2)
        DIE_FAIL(DIE_ERR_NOMEM,
                 "die_kv_set_batch_noalloc: table housekeeping is due and would allocate; call die_kv_maintain_sync "
                 "at a point where allocation is allowed, then retry");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (kv->policy == DIE_KV_POLICY_LFU) {  // the op-by-op LFU kernel uses no scratch at all
        DIE_TRY(kv_set_lfu(kv, digests, values, ttl, M, now, deferred_stamp, slot_out, dropped_out, s));
        kv->claims_ub += n;
        kv->ops_enq += n;
        return DIE_OK;
    }
    const KvView v = view_of(kv);
    const uint4* dg = reinterpret_cast<const uint4*>(digests);
    const unsigned long long* vals = reinterpret_cast<const unsigned long long*>(values);
    unsigned long long* dr = reinterpret_cast<unsigned long long*>(dropped_out);
    const int grid = ceil_div(M, KV_THREADS);
    // Nothing in the scratch has to start at zero: the precheck writes op_slot[i] for every op before any
    // later kernel of this call reads it.  What the previous call left there is simply overwritten.
    DIE_LAUNCH(ST_KV_PRECHECK, s, kv_set_precheck_kernel<<<grid, KV_THREADS, 0, s>>>(v, dg, M, nullptr, scratch));
    const int n_tiles = set_window_tiles(kv, M);
    DIE_LAUNCH(ST_KV_WINDOW, s, kv_set_window_count_kernel<<<n_tiles, WIN_THREADS, 0, s>>>(v, kv->tail, nullptr, deferred_stamp, kv->tile_cnt));
    DIE_LAUNCH(ST_KV_WINDOW, s, kv_set_window_decide_kernel<<<1, WIN_THREADS, 0, s>>>(v, kv->tail, nullptr, deferred_stamp, n_tiles, kv->tile_cnt));
    DIE_LAUNCH(ST_KV_CLAIM, s, kv_set_claim_kernel<<<grid, KV_THREADS, 0, s>>>(v, dg, M, now, kv->tail, values ? 0 : 1,
                                                                              nullptr, 1, scratch, slot_out, dr));
    if (values || ttl)
        DIE_LAUNCH(ST_KV_FINALIZE, s, kv_set_finalize_kernel<<<grid, KV_THREADS, 0, s>>>(v, M, kv->tail, vals, ttl,
                                                                                         nullptr, scratch));
    DIE_LAUNCH(ST_KV_EVICT, s, kv_set_evict_tiles_kernel<<<n_tiles, WIN_THREADS, 0, s>>>(v, kv->tail, kv->tile_cnt, dr));
    DIE_LAUNCH(ST_KV_EVICT, s, kv_set_evict_kernel<<<1, WIN_THREADS, 0, s>>>(v, dg, vals, ttl, M, now, kv->tail,
                                                                             deferred_stamp, nullptr, slot_out, dr));
    // Host bookkeeping only once every launch of the batch is enqueued (a failed launch returned above and left
    // the bounds untouched): the same three fields kv_reserve() advances for die_kv_set_batch.  ops_enq feeds
    // kv_mirror_refresh(): without it a pipeline-owned table would under-count the inserts in flight.
    kv->claims_ub += n;
    kv->ops_enq += n;
    kv->dropped_prefix = 0;
    kv->tail += n;
    return DIE_OK;
}
