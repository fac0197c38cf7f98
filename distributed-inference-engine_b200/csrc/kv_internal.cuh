// K2/K3: GPU-resident KV store with the reference's LRU + TTL semantics (src/kvstore.py:26-236).
//
// HBM layout
//   slots : capacity x 64 B (power-of-two capacity >= 2*max_size), open addressing, linear probing.
//           sector 0 = {fingerprint (128-bit MD5 of the key), created_at f64, ttl f64}   (read-mostly)
//           sector 1 = {lru stamp u64, value u64, last_accessed f64, access_count u32}  (hot, mutable)
//           fingerprint (0,0) = EMPTY, (~0,~0) = TOMBSTONE.
//   log   : ring of uint32 slot indices indexed by LRU stamp.  Every touch (get hit, insert) of
//           batch op i gets stamp t+i and writes log[t+i] = slot.  The reference's OrderedDict
//           (_access_order, kvstore.py:71-78) is the subsequence of log records that are still
//           current (slot.stamp == position); its head (next victim, kvstore.py:89-91) is the first
//           current record at or after `head`.  Stale records are skipped lazily.
//   rows  : optional free-list of result-row handles (pipeline mode): an insert pops a row for its
//           value, every delete/eviction pushes it back.
//
// Batch semantics ("as if applied in request order")
//   get   : order-independent once stamps are op indices (atomicMax) -> fully parallel.
//   set   : parallel fast path when provably identical to the sequential result, i.e. no key of
//           the batch that already lives in the table sits inside the window of log records the
//           batch's evictions can reach, and that window does not run out of pre-batch entries.
//           Otherwise an exact sequential kernel replays the ops one by one (kvstore.py:104-134),
//           including the reference's StopIteration when nothing stamped is left to evict.
#pragma once
#include <math.h>

#include "common.cuh"

namespace die {

struct __align__(64) Slot {
    unsigned long long fp_lo, fp_hi;
    double created_at;
    double ttl;                // NaN = None
    unsigned long long stamp;  // 0 = not in the LRU order (batch_set insert not yet stamped)
    unsigned long long value;
    double last_accessed;
    unsigned int access_count;
    unsigned int lfu_count;    // LFU policy: the reference's _access_count[key] (kvstore.py:79-80); unused otherwise
};
static_assert(sizeof(Slot) == 64, "slot must be one 64-byte record");

struct KvState {
    // persistent counters
    unsigned long long size, hits, misses, evictions, head, n_tomb, n_empty_claimed;
    unsigned long long free_head, free_tail;  // result-row queue: pops at free_head, pushes at free_tail (monotone)
    unsigned long long status, slow_batches, n_dropped_last;
    // per set-batch scratch (reset by the last kernel of each batch)
    unsigned long long n_pre, n_absent, min_prev, n_new, conflict, n_dropped;
    // victim window of the stand-alone set path: 0 none / sequential, 1 tiles (parallel), 2 one-CTA walk
    unsigned long long win_mode, win_origin, win_hnew;
    // hit / miss counters of the batched get, spread over 16 x 64-byte lines so that thousands of CTAs do not
    // serialise on one L2 atomic unit; readers add them to hits / misses
    unsigned long long hit_part[16][8], miss_part[16][8];
};

// Snapshot of the counters the host's maintenance decisions need, written by the device into pinned host
// memory at every replay boundary (seq0 first, seq1 last: the host retries on a torn read).  With it the host
// refreshes its conservative bounds without synchronising the stream.
struct KvMirror {
    unsigned long long seq0, head, size, n_tomb, ops_seen, gen, seq1;
};

struct KvView {
    Slot* slots;
    uint32_t* log;
    KvState* st;
    uint32_t* free_rows;       // nullptr unless the table manages result rows
    unsigned long long row_mask; // free_rows ring size - 1 (a power of two >= the number of rows)
    unsigned long long mask;   // capacity - 1
    unsigned long long rmask;  // ring_cap - 1
    unsigned long long max_size;
    double default_ttl;
    int policy;                // DIE_KV_POLICY_*
};

#ifdef __CUDACC__
constexpr unsigned long long FP_TOMB_WORD = ~0ull;
__device__ __forceinline__ bool fp_eq(u128 a, u128 b) { return a.lo == b.lo && a.hi == b.hi; }
__device__ __forceinline__ bool fp_empty(u128 a) { return a.lo == 0ull && a.hi == 0ull; }
__device__ __forceinline__ bool fp_tomb(u128 a) { return a.lo == FP_TOMB_WORD && a.hi == FP_TOMB_WORD; }
__device__ __forceinline__ u128 fp_of(uint4 d) {
    return u128{(unsigned long long)d.x | ((unsigned long long)d.y << 32),
                (unsigned long long)d.z | ((unsigned long long)d.w << 32)};
}

__device__ __forceinline__ u128 ld_fp(const Slot* s) {
    u128 r;  // .cg: slots are mutated by other SMs inside the same kernel, never trust L1
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(r.lo), "=l"(r.hi) : "l"(s) : "memory");
    return r;
}
__device__ __forceinline__ unsigned long long ld_u64_cg(const unsigned long long* p) {
    unsigned long long r;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void ld_created_ttl(const Slot* s, double& created, double& ttl) {
    asm volatile("ld.global.cg.v2.f64 {%0, %1}, [%2];"
                 : "=d"(created), "=d"(ttl)
                 : "l"(&s->created_at)
                 : "memory");
}

// Linear probe for `key`; returns slot index or -1 (stops at the first EMPTY slot).
__device__ __forceinline__ long long kv_find(const KvView& kv, u128 key) {
    unsigned long long s = key.lo & kv.mask;
    for (unsigned long long n = 0; n <= kv.mask; ++n) {
        const u128 fp = ld_fp(&kv.slots[s]);
        if (fp_eq(fp, key)) return (long long)s;
        if (fp_empty(fp)) return -1;
        s = (s + 1) & kv.mask;
    }
    return -1;
}

__device__ __forceinline__ bool ttl_expired(double now, double created, double ttl) {
    return !isnan(ttl) && (now - created) > ttl;  // strict '>' (kvstore.py:204)
}

__device__ __forceinline__ void warp_add(unsigned long long* ctr, bool pred) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m && lane_id() == 0) atomicAdd(ctr, (unsigned long long)__popc(m));
}
__device__ __forceinline__ void warp_sub(unsigned long long* ctr, bool pred) {
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m && lane_id() == 0) atomicAdd(ctr, (unsigned long long)(-(long long)__popc(m)));
}

// One atomic per CTA instead of one per warp: counters shared by a whole launch live in one L2 sector, and
// same-address atomics from every warp of a large batch would serialise there.  All threads must call.
__device__ __forceinline__ void block_add(unsigned long long* ctr, bool pred) {
    const int n = __syncthreads_count(pred);
    if (threadIdx.x == 0 && n) atomicAdd(ctr, (unsigned long long)n);
}
__device__ __forceinline__ void block_sub(unsigned long long* ctr, bool pred) {
    const int n = __syncthreads_count(pred);
    if (threadIdx.x == 0 && n) atomicAdd(ctr, (unsigned long long)(-(long long)n));
}

// result-row handles (pipeline mode): a power-of-two ring queue, popped at free_head, pushed at free_tail.
// Pushes and pops may run side by side as long as the pops of a phase never outnumber the handles that were free
// when the phase began (a pop then never reaches an entry whose push is still in flight): the table owns
// max_batch + 64 more rows than entries, and a phase pops at most one row per request of a batch.
__device__ __forceinline__ void release_row(const KvView& kv, unsigned long long value) {
    if (kv.free_rows) kv.free_rows[atomicAdd(&kv.st->free_tail, 1ull) & kv.row_mask] = (uint32_t)value;
}
__device__ __forceinline__ unsigned long long acquire_row(const KvView& kv) {
    return kv.free_rows[atomicAdd(&kv.st->free_head, 1ull) & kv.row_mask];
}
// The same for a whole warp with one atomic (all 32 lanes call; lanes with !want get 0 / do nothing).
__device__ __forceinline__ unsigned long long warp_acquire_row(const KvView& kv, bool want) {
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (!m) return 0ull;
    const int leader = __ffs(m) - 1;
    unsigned long long base = 0;
    if ((int)lane_id() == leader) base = atomicAdd(&kv.st->free_head, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (!want) return 0ull;
    return kv.free_rows[(base + __popc(m & ((1u << lane_id()) - 1u))) & kv.row_mask];
}
__device__ __forceinline__ void warp_release_row(const KvView& kv, bool have, unsigned long long value) {
    const unsigned m = __ballot_sync(0xffffffffu, have);
    if (!m || !kv.free_rows) return;
    const int leader = __ffs(m) - 1;
    unsigned long long base = 0;
    if ((int)lane_id() == leader) base = atomicAdd(&kv.st->free_tail, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (have) kv.free_rows[(base + __popc(m & ((1u << lane_id()) - 1u))) & kv.row_mask] = (uint32_t)value;
}
// One KVCache.get (kvstore.py:136-164) / __contains__ (:221-228) probe.  Returns the code
// (1 hit, 0 absent, 2 expired-and-deleted-by-this-op); `deleted` tells the caller to account for it.
// Both 32-byte sectors of a slot, fetched with three independent 16-byte loads so that a hit costs one
// DRAM round trip instead of three dependent ones.
struct SlotLine {
    u128 fp;
    double created, ttl;
    unsigned long long stamp, value;
};
__device__ __forceinline__ SlotLine ld_slot_line(const Slot* s) {
    SlotLine l;
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(l.fp.lo), "=l"(l.fp.hi) : "l"(s));
    asm volatile("ld.global.cg.v2.f64 {%0, %1}, [%2];" : "=d"(l.created), "=d"(l.ttl) : "l"(&s->created_at));
    asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(l.stamp), "=l"(l.value) : "l"(&s->stamp));
    return l;
}

// One KVCache.get (kvstore.py:136-164) / __contains__ (:221-228) probe.  Returns the code
// (1 hit, 0 absent, 2 expired-and-deleted-by-this-op).
constexpr int PROBE_AHEAD = 4;
// kv_get_one_from: the same with the key's home slot (key.lo & mask) already fetched by the caller as `l`, so that
// the caller can overlap that round trip with other work.
template <bool CONTAINS>
__device__ __forceinline__ int kv_get_one_from(const KvView& kv, u128 key, SlotLine l, double now, unsigned long long stamp,
                                               unsigned long long& value, long long& slot) {
    value = 0;
    slot = -1;
    unsigned long long cur = key.lo & kv.mask;
    bool found = fp_eq(l.fp, key);
    if (!found && !fp_empty(l.fp)) {
        // The home slot is somebody else's: walk the chain PROBE_AHEAD slots per round trip (their fingerprints
        // are fetched together), so that the longest chain of a batch costs a few DRAM latencies, not one each.
        cur = (cur + 1) & kv.mask;
        for (unsigned long long n = 1; n <= kv.mask && !found; n += PROBE_AHEAD) {
            u128 f[PROBE_AHEAD];
#pragma unroll
            for (int k = 0; k < PROBE_AHEAD; ++k) f[k] = ld_fp(&kv.slots[(cur + k) & kv.mask]);
            int at = -1;
            bool end = false;
#pragma unroll
            for (int k = 0; k < PROBE_AHEAD; ++k) {
                if (at < 0 && !end) {
                    if (fp_eq(f[k], key)) at = k;
                    else if (fp_empty(f[k])) end = true;
                }
            }
            if (at >= 0) {
                cur = (cur + at) & kv.mask;
                l = ld_slot_line(&kv.slots[cur]);
                found = fp_eq(l.fp, key);  // false: an op of this launch deleted it (expired) in the meantime
                break;
            }
            if (end) break;
            cur = (cur + PROBE_AHEAD) & kv.mask;
        }
    }
    if (!found) return 0;
    Slot* sl = &kv.slots[cur];
    if (ttl_expired(now, l.created, l.ttl)) {
        // delete(key) (kvstore.py:153-156); exactly one of several duplicate ops wins
        const u128 old = cas128(sl, key, u128{FP_TOMB_WORD, FP_TOMB_WORD});
        if (!fp_eq(old, key)) return 0;
        value = l.value;
        sl->stamp = 0ull;
        release_row(kv, value);
        return 2;
    }
    slot = (long long)cur;
    value = l.value;
    if (!CONTAINS) {
        if (kv.policy == DIE_KV_POLICY_LRU) {
            atomicMax(&sl->stamp, stamp);  // move_to_end: the last toucher of the batch wins
            kv.log[stamp & kv.rmask] = (uint32_t)slot;
        } else if (kv.policy == DIE_KV_POLICY_LFU) {
            // _access_count[key] = get(key, 0) + 1 (kvstore.py:79-80).  An entry that is not counted yet (inserted with
            // batch=True) enters the dict now: its place in the tie-break order is this access -- the earliest one
            // when several ops of the batch touch it.
            atomicAdd(&sl->lfu_count, 1u);
            if (l.stamp == 0ull) {
                unsigned long long cur = 0ull;
                for (;;) {
                    const unsigned long long old = atomicCAS(&sl->stamp, cur, stamp);
                    if (old == cur || (old != 0ull && old <= stamp)) break;
                    cur = old;
                }
                kv.log[stamp & kv.rmask] = (uint32_t)slot;
            }
        }  // FIFO: an access changes nothing in the eviction order (kvstore.py:99-102)
        atomicAdd(&sl->access_count, 1u);
        sl->last_accessed = now;
    }
    return 1;
}

template <bool CONTAINS>
__device__ __forceinline__ int kv_get_one(const KvView& kv, u128 key, double now, unsigned long long stamp,
                                          unsigned long long& value, long long& slot) {
    return kv_get_one_from<CONTAINS>(kv, key, ld_slot_line(&kv.slots[key.lo & kv.mask]), now, stamp, value, slot);
}

// Insert of a key that is known not to live in the table before this batch: claim the first free
// slot of the probe chain with a 128-bit CAS on the fingerprint.  Duplicate ops of the same batch
// converge on one slot (claimed == true for exactly one of them).
__device__ __forceinline__ long long kv_claim_absent(const KvView& kv, u128 key, bool& claimed, bool& from_empty) {
    claimed = false;
    from_empty = false;
    for (;;) {
        unsigned long long cur = key.lo & kv.mask;
        long long first_free = -1;
        bool first_free_empty = false;
        // first the home slot alone (most chains end there), then PROBE_AHEAD fingerprints per round trip
        bool done = false;
        for (unsigned long long n = 0; n <= kv.mask && !done;) {
            const int width = n == 0 ? 1 : PROBE_AHEAD;
            u128 f[PROBE_AHEAD];
#pragma unroll
            for (int k = 0; k < PROBE_AHEAD; ++k)
                if (k < width) f[k] = ld_fp(&kv.slots[(cur + k) & kv.mask]);
#pragma unroll
            for (int k = 0; k < PROBE_AHEAD; ++k) {
                if (k < width && !done) {
                    const long long at = (long long)((cur + k) & kv.mask);
                    if (fp_eq(f[k], key)) return at;  // a duplicate op of this batch already inserted it
                    if (fp_tomb(f[k]) && first_free < 0) first_free = at;
                    if (fp_empty(f[k])) {
                        if (first_free < 0) {
                            first_free = at;
                            first_free_empty = true;
                        }
                        done = true;
                    }
                }
            }
            cur = (cur + width) & kv.mask;
            n += width;
        }
        if (first_free < 0) continue;  // cannot happen: the host keeps >= capacity/4 EMPTY slots
        const u128 expect = first_free_empty ? u128{0ull, 0ull} : u128{FP_TOMB_WORD, FP_TOMB_WORD};
        const u128 old = cas128(&kv.slots[first_free], expect, key);
        if (fp_eq(old, expect)) {
            claimed = true;
            from_empty = first_free_empty;
            return first_free;
        }
        if (fp_eq(old, key)) return first_free;
        // somebody else's key took the slot: walk the chain again
    }
}
#endif  // __CUDACC__

}  // namespace die

struct die_kv {
    int device;
    int policy;  // DIE_KV_POLICY_*
    uint64_t max_size, capacity, ring_cap;
    double default_ttl;
    die::Slot* slots;
    die::Slot* slots_spare;  // second table buffer, allocated at the first rebuild and reused
    uint32_t* log;
    die::KvState* st;
    uint32_t* free_rows;  // pipeline mode only
    uint64_t n_rows;
    uint64_t row_ring;    // size of the free_rows ring (power of two >= n_rows)
    uint64_t tail;        // next LRU stamp to hand out (host-tracked: every batch op consumes one)
    uint64_t head_lb;     // host lower bound of the device-side head
    uint64_t claims_ub;   // upper bound of EMPTY slots consumed since the last rebuild
    uint64_t live_at_rebuild;
    uint64_t max_chunk;   // largest batch one launch sequence may take
    int32_t* op_slot;     // scratch [op_cap]
    uint64_t op_cap;
    uint64_t dropped_prefix;  // dropped values written by earlier chunks of the last set batch
    uint32_t* log_spare;      // second LRU-log buffer, allocated at the first compaction and reused
    unsigned int* tile_cnt;   // compaction scratch: current records per tile of 4096 log positions
    unsigned long long* n_kept_dev;
    die::KvMirror* mirror;    // pinned + mapped host memory (device-visible at the same address)
    uint64_t ops_enq;         // insert ops reserved so far (monotone)
    uint64_t mirror_gen;      // bumped by every maintenance: older snapshots describe a table that is gone
    uint64_t mirror_seq;
};

namespace die {

inline KvView view_of(const die_kv* kv) {
    KvView v;
    v.slots = kv->slots;
    v.log = kv->log;
    v.st = kv->st;
    v.free_rows = kv->free_rows;
    v.row_mask = kv->row_ring ? kv->row_ring - 1 : 0;
    v.mask = kv->capacity - 1;
    v.rmask = kv->ring_cap - 1;
    v.max_size = kv->max_size;
    v.default_ttl = kv->default_ttl;
    v.policy = kv->policy;
    return v;
}

// n_rows > 0: the table hands out result-row handles in [0, n_rows) as values (pipeline mode).
// min_batch: the table is sized so that one launch sequence can take a batch of this many ops.
die_status_t kv_create_impl(uint64_t max_size, double default_ttl, uint64_t n_rows, uint64_t min_batch,
                            die_kv** out, int policy = DIE_KV_POLICY_LRU);

#ifdef __CUDACC__
// One thread: snapshot -> pinned host mirror.  ops_seen = insert ops the host had reserved when it enqueued
// the kernel this runs in (all of them are stream-ordered before it).
__device__ __forceinline__ void kv_mirror_write(const KvState* st, KvMirror* m, unsigned long long ops_seen,
                                                unsigned long long gen, unsigned long long seq) {
    volatile KvMirror* vm = m;
    vm->seq0 = seq;
    __threadfence_system();
    vm->head = st->head;
    vm->size = st->size;
    vm->n_tomb = st->n_tomb;
    vm->ops_seen = ops_seen;
    vm->gen = gen;
    __threadfence_system();
    vm->seq1 = seq;
}
#endif
// Arguments of kv_mirror_write for a kernel enqueued now.
struct KvMirrorArgs {
    const KvState* st;
    KvMirror* m;
    unsigned long long ops_seen, gen, seq;
};
KvMirrorArgs kv_mirror_args(die_kv* kv);

// Host-side guard run before a batch of n ops: keeps the LRU ring and the EMPTY-slot reserve safe
// using conservative host bounds (synchronises only when a bound is about to be crossed).
die_status_t kv_reserve(die_kv* kv, uint64_t n, bool inserts, cudaStream_t s);

// Batched KVCache.get; stamps are kv->tail + i.  Advances kv->tail by B.
die_status_t kv_get_batch_impl(die_kv* kv, const uint32_t* digests, int32_t B, double now, uint8_t* hit_out,
                               uint64_t* value_out, int32_t* slot_out, cudaStream_t s);

// Batched KVCache.set.  skip (optional, [M]): ops with skip[i] == 1 are not applied (their stamp
// is still reserved).  known_absent != 0: the caller guarantees that no active key lives in the
// table (true for the misses of the get batch just before) and n_active_dev holds their count.
die_status_t kv_set_batch_impl(die_kv* kv, const uint32_t* digests, const uint64_t* values, const double* ttl,
                               int32_t M, double now, int32_t deferred, const uint8_t* skip, int known_absent,
                               const uint32_t* n_active_dev, int32_t* slot_out, uint64_t* dropped_out,
                               cudaStream_t s);

}  // namespace die
