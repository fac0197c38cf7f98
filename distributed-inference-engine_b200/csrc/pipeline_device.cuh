// Device-side pieces of the fused pipeline that more than one translation unit needs (pipeline.cu and the
// tensor-core GEMM whose extra CTAs evict): per-batch control block, scratch pointers, eviction stage.
#pragma once
#include "kv_internal.cuh"

struct PipeCtrl {
    unsigned int n_miss, ticket_get, ticket_claim, any_dup;
    unsigned long long e_ub, e_final;
    unsigned long long whead;  // LRU-log position the victim window of this batch starts at
    unsigned long long ts[8];  // %globaltimer marks of the last get kernel (ns): diagnostics only
};

namespace die {

constexpr int PT = 256;  // threads per CTA of the pipeline kernels (== ROUTE_THREADS)
constexpr int WCHUNK = 1024;  // log records per victim-window chunk

struct PipeBufs {
    uint8_t* code;
    unsigned long long* value;
    int32_t* miss_rank;
    int32_t* miss_req;
    int32_t* set_slot;
    int32_t* set_row;
    uint32_t* wmask;   // [n_chunks][32]: which of a chunk's 1024 log records are current
    uint32_t* wcount;  // [n_chunks]
    int n_chunks;
    PipeCtrl* ctrl;
};

// Victim window, stage 2 (CTAs of the last kernel of the step): record r of the window is evicted iff it is
// current and fewer than e_final current records precede it.  256 records per CTA.
static __device__ __noinline__ void pipe_evict_cta(const KvView& kv, const PipeBufs& pb, int cta, unsigned long long t_set) {
    __shared__ unsigned s_red[PT / 32];
    __shared__ unsigned s_before;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long e = pb.ctrl->e_final;
    if (e == 0) return;
    const int rec0 = cta * PT;               // first window record of this CTA
    const int chunk = rec0 / WCHUNK;
    if (chunk >= pb.n_chunks) return;
    // current records in all chunks before this one
    unsigned part = 0;
    for (int c = tid; c < chunk; c += PT) part += pb.wcount[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_red[warp] = part;
    __syncthreads();
    if (tid == 0) {
        unsigned t = 0;
        for (int w = 0; w < PT / 32; ++w) t += s_red[w];
        // plus the words of this chunk that precede this CTA's 8 words
        const int w0 = (rec0 % WCHUNK) / 32;
        for (int w = 0; w < w0; ++w) t += __popc(pb.wmask[chunk * 32 + w]);
        s_before = t;
    }
    __syncthreads();
    const int w0 = (rec0 % WCHUNK) / 32;
    unsigned before = s_before;
    for (int w = 0; w < warp; ++w) before += __popc(pb.wmask[chunk * 32 + w0 + w]);
    const unsigned word = pb.wmask[chunk * 32 + w0 + warp];
    const unsigned long long rank = before + __popc(word & ((1u << lane) - 1u));
    const bool mine = (word >> lane) & 1u;
    const unsigned long long pos = pb.ctrl->whead + (unsigned long long)rec0 + tid;
    if (mine && rank < e) {
        Slot* sl = &kv.slots[kv.log[pos & kv.rmask]];
        release_row(kv, sl->value);
        st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
        sl->stamp = 0ull;
        if (rank == e - 1) {  // the newest victim: everything up to it is gone, nobody else touches the counters
            KvState* st = kv.st;
            st->size -= e;
            st->evictions += e;
            st->n_tomb += e;
            st->head = pos + 1;
        }
    }
    // window too short (fewer than e current records in 4 records per request): finish sequentially
    if (cta == 0 && tid == 0) {
        unsigned long long total = 0;
        for (int c = 0; c < pb.n_chunks; ++c) total += pb.wcount[c];
        if (total < e) {
            unsigned long long p = pb.ctrl->whead + (unsigned long long)pb.n_chunks * WCHUNK, left = e - total;
            while (left > 0 && p < t_set) {
                const uint32_t sidx = kv.log[p & kv.rmask];
                if ((unsigned long long)sidx <= kv.mask && ld_u64_cg(&kv.slots[sidx].stamp) == p) {
                    Slot* sl = &kv.slots[sidx];
                    release_row(kv, sl->value);
                    st128(sl, u128{FP_TOMB_WORD, FP_TOMB_WORD});
                    sl->stamp = 0ull;
                    --left;
                }
                ++p;
            }
            KvState* st = kv.st;
            if (left > 0) st->status = DIE_ERR_INVALID;  // unreachable while max_batch <= max_size
            st->size -= e;
            st->evictions += e;
            st->n_tomb += e;
            st->head = p;
        }
    }
}

}  // namespace die
