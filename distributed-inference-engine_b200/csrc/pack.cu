// K4/K5: ragged pack / unpack.  Replace the Batcher's gather and scatter one-liners
// (src/batcher.py:211 and :225-227) for device-resident CSR payloads.
//
// pack: one thread per 16-byte vector of the padded batch (grid-stride); unpack: one warp per row, lanes
// stream the row's vectors.  The padded side is always 16-byte aligned; the ragged side
// is read/written at its natural alignment: reads use two aligned 128-bit loads + a funnel shift
// (pack_device.cuh), so DRAM traffic stays at the algorithmic bytes; stores to ragged rows are full 128-bit
// where the vector lies inside the row and element-wise at the two ragged ends.
// Grid = up to 16 CTAs per SM, grid-stride over rows.
#include "common.cuh"
#include "pack_device.cuh"
#include "tc_common.cuh"

namespace die {

constexpr int PACK_THREADS = 256;

// out[r, :] = values[off[q] : off[q+1]] zero-padded, q = src_index ? src_index[r] : r
__global__ void __launch_bounds__(PACK_THREADS)
pack_vec_kernel(const uint8_t* __restrict__ values, const int64_t* __restrict__ off,
                const int32_t* __restrict__ src_index, long long R, int elem_bytes, int vpr /*vectors per row*/,
                int4* __restrict__ out, int32_t* __restrict__ lengths_out) {
    const long long total = R * vpr;
    const int row_bytes = vpr * 16;
    for (long long v = blockIdx.x * (long long)PACK_THREADS + threadIdx.x; v < total;
         v += (long long)gridDim.x * PACK_THREADS) {
        const long long r = v / vpr;
        const int c = (int)(v - r * vpr);
        const long long q = src_index ? src_index[r] : r;
        const long long o0 = off[q], o1 = off[q + 1];
        const long long len_bytes_full = (o1 - o0) * elem_bytes;
        const int len_bytes = (int)(len_bytes_full < row_bytes ? len_bytes_full : row_bytes);
        if (c == 0 && lengths_out) lengths_out[r] = (int32_t)(o1 - o0);
        const int keep = len_bytes - c * 16;  // bytes of this vector that carry payload
        int4 val = make_int4(0, 0, 0, 0);
        if (keep > 0) {
            const uint8_t* src = values + o0 * elem_bytes + (long long)c * 16;
            val = load16_shifted(src, 0, keep < 16 ? keep : 16);
            if (keep < 16) val = mask_tail16(val, keep);
        }
        stg_stream(out + v, val);
    }
}

// generic fallback (row bytes not a multiple of 16 or unaligned buffers): one element per thread
__global__ void __launch_bounds__(PACK_THREADS)
pack_elem_kernel(const uint8_t* __restrict__ values, const int64_t* __restrict__ off,
                 const int32_t* __restrict__ src_index, long long R, int elem_bytes, int Lpad,
                 uint8_t* __restrict__ out, int32_t* __restrict__ lengths_out) {
    const long long total = R * Lpad;
    for (long long e = blockIdx.x * (long long)PACK_THREADS + threadIdx.x; e < total;
         e += (long long)gridDim.x * PACK_THREADS) {
        const long long r = e / Lpad;
        const int c = (int)(e - r * Lpad);
        const long long q = src_index ? src_index[r] : r;
        const long long o0 = off[q], len = off[q + 1] - o0;
        if (c == 0 && lengths_out) lengths_out[r] = (int32_t)len;
        for (int b = 0; b < elem_bytes; ++b)
            out[e * elem_bytes + b] = c < len ? values[(o0 + c) * elem_bytes + b] : (uint8_t)0;
    }
}

// out_values[off_out[q] : off_out[q+1]] = packed[r, :len], q = dst_index ? dst_index[r] : r.
// One warp per row; the work unit of a lane is one 16-byte ALIGNED vector of the ragged destination, so
// stores are full 128-bit except at the two ragged ends of the row (element-wise there: neighbouring rows
// share those vectors).  Only the live part of the padded row is read.
__global__ void __launch_bounds__(PACK_THREADS)
unpack_vec_kernel(const uint8_t* __restrict__ packed, const int64_t* __restrict__ off_out,
                  const int32_t* __restrict__ dst_index, long long R, int elem_bytes, int row_bytes, int vpr1,
                  uint8_t* __restrict__ out_values) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * (PACK_THREADS / 32);
    for (long long r = (long long)blockIdx.x * (PACK_THREADS / 32) + (threadIdx.x >> 5); r < R; r += warps) {
        const long long q = dst_index ? dst_index[r] : r;
        const long long o0 = off_out[q];
        long long len_bytes = (off_out[q + 1] - o0) * elem_bytes;
        if (len_bytes > row_bytes) len_bytes = row_bytes;
        if (len_bytes <= 0) continue;
        const uintptr_t d0 = reinterpret_cast<uintptr_t>(out_values) + (uintptr_t)(o0 * elem_bytes);
        const uintptr_t d1 = d0 + (uintptr_t)len_bytes;
        const uintptr_t a0 = d0 & ~uintptr_t(15);
        const int nvec = (int)((d1 - a0 + 15) >> 4);  // aligned destination vectors the row touches
        const uint8_t* row = packed + r * (long long)row_bytes;
        for (int c0 = lane; c0 < nvec; c0 += 128) {
          int4 vals[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {  // issue the loads of four vectors before any store
            const int c = c0 + 32 * j;
            vals[j] = make_int4(0, 0, 0, 0);
            if (c < nvec) {
                const uintptr_t va = a0 + (uintptr_t)c * 16;
                const int first = va < d0 ? (int)(d0 - va) : 0;
                const int last = va + 16 > d1 ? (int)(d1 - va) : 16;
                vals[j] = load16_shifted(row + ((long long)va - (long long)d0), first, last);
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = c0 + 32 * j;
            if (c >= nvec) continue;
            const uintptr_t va = a0 + (uintptr_t)c * 16;
            const int first = va < d0 ? (int)(d0 - va) : 0;       // first byte of the vector inside the row
            const int last = va + 16 > d1 ? (int)(d1 - va) : 16;  // one past the last byte inside the row
            // source byte for destination byte va+k is row + (va + k - d0)
            const int4 val = vals[j];
            if (first == 0 && last == 16) {
                stg_stream(reinterpret_cast<int4*>(va), val);
            } else {
                const unsigned w[4] = {(unsigned)val.x, (unsigned)val.y, (unsigned)val.z, (unsigned)val.w};
                uint8_t* dst = reinterpret_cast<uint8_t*>(va);
                if ((elem_bytes & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (4 * j >= first && 4 * j + 4 <= last) reinterpret_cast<unsigned*>(dst)[j] = w[j];
                } else {
                    for (int k = first; k < last; ++k) dst[k] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
                }
            }
          }
        }
    }
}

__global__ void __launch_bounds__(PACK_THREADS)
unpack_elem_kernel(const uint8_t* __restrict__ packed, const int64_t* __restrict__ off_out,
                   const int32_t* __restrict__ dst_index, long long R, int elem_bytes, int Lpad,
                   uint8_t* __restrict__ out_values) {
    const long long total = R * Lpad;
    for (long long e = blockIdx.x * (long long)PACK_THREADS + threadIdx.x; e < total;
         e += (long long)gridDim.x * PACK_THREADS) {
        const long long r = e / Lpad;
        const int c = (int)(e - r * Lpad);
        const long long q = dst_index ? dst_index[r] : r;
        const long long o0 = off_out[q], len = off_out[q + 1] - o0;
        if (c >= len) continue;
        for (int b = 0; b < elem_bytes; ++b) out_values[(o0 + c) * elem_bytes + b] = packed[e * elem_bytes + b];
    }
}

// ---- bulk-copy forms for rows of a kilobyte and more ------------------------------------------------------------
// The vector kernels above move every 16 bytes with a load and a store of their own thread (ncu: long-scoreboard
// stalls, DRAM 43 % / 31 % busy on the C3 shape).  Here the copy engine does the global traffic: a row's bytes arrive in shared memory as ONE bulk copy (cp.async.bulk, completion on an mbarrier), the
// threads only realign them there (ragged rows start at any element, bulk copies want 16-byte addresses), and the
// result leaves as ONE bulk copy.  Each CTA keeps BK_STAGES rows in flight; several CTAs share an SM.
constexpr int BK_THREADS = 128;
constexpr int BK_STAGES = 6;
constexpr int BK_LAG = 3;   // rows between a buffer's store and its refill
constexpr int BK_VEC_PER_THREAD = 8;  // BK_MAX_ROW / 16 / BK_THREADS
constexpr int BK_MAX_ROW = 16384;  // bytes per padded row the staging buffers are sized for
constexpr int BK_MIN_ROW = 1024;

struct BulkSeg {      // what one row moves
    const uint8_t* src;   // first payload byte (pack: ragged source; unpack: padded row)
    uint8_t* dst;         // first destination byte (pack: padded row; unpack: ragged destination)
    int len;              // payload bytes (clamped to the row)
};

__device__ __forceinline__ BulkSeg pack_seg(const uint8_t* values, const int64_t* off, const int32_t* src_index, long long r,
                                            int elem_bytes, int row_bytes, uint8_t* out, int32_t* lengths_out, bool note) {
    const long long q = src_index ? src_index[r] : r;
    const long long o0 = off[q], o1 = off[q + 1];
    if (note && lengths_out) lengths_out[r] = (int32_t)(o1 - o0);
    const long long full = (o1 - o0) * elem_bytes;
    BulkSeg g;
    g.src = values + o0 * elem_bytes;
    g.dst = out + r * (long long)row_bytes;
    g.len = (int)(full < row_bytes ? full : row_bytes);
    return g;
}

// out[r, :] = values[off[q] : off[q+1]] zero-padded (pack_vec_kernel's contract), row_bytes a multiple of 16.
__global__ void __launch_bounds__(BK_THREADS)
pack_bulk_kernel(const uint8_t* __restrict__ values, const int64_t* __restrict__ off, const int32_t* __restrict__ src_index,
                 long long R, int elem_bytes, int row_bytes, uint8_t* __restrict__ out, int32_t* __restrict__ lengths_out) {
    extern __shared__ __align__(128) uint8_t bk_smem[];
    __shared__ __align__(8) uint64_t full[BK_STAGES];
    const int tid = threadIdx.x;
    const int in_bytes = row_bytes + 32;   // aligned superset of an unaligned segment of up to row_bytes
    uint8_t* in_buf = bk_smem;             // [BK_STAGES][in_bytes]: a row is realigned in place
    if (tid == 0)
        for (int s = 0; s < BK_STAGES; ++s) tc::mbar_init(&full[s], 1);
    __syncthreads();
    auto fetch = [&](long long r, int s) {  // thread 0: ask for row r's bytes (their 16-byte aligned superset)
        const BulkSeg g = pack_seg(values, off, src_index, r, elem_bytes, row_bytes, out, lengths_out, true);
        const uintptr_t a = reinterpret_cast<uintptr_t>(g.src), a0 = a & ~uintptr_t(15);
        const uint32_t bytes = g.len > 0 ? (uint32_t)(((a + (uintptr_t)g.len + 15) & ~uintptr_t(15)) - a0) : 0u;
        tc::mbar_arrive_expect_tx(&full[s], bytes);
        if (bytes) tc::bulk_load_1d(in_buf + (size_t)s * in_bytes, reinterpret_cast<const void*>(a0), bytes, &full[s]);
    };
    const long long r0 = blockIdx.x, stride = gridDim.x;
    if (tid == 0)
        for (int s = 0; s < BK_STAGES; ++s)
            if (r0 + s * stride < R) fetch(r0 + s * stride, s);
    const int vpr = row_bytes / 16;
    long long i = 0;
    for (long long r = r0; r < R; r += stride, ++i) {
        const int s = (int)(i % BK_STAGES);
        const BulkSeg g = pack_seg(values, off, src_index, r, elem_bytes, row_bytes, out, nullptr, false);
        tc::mbar_wait(&full[s], (uint32_t)((i / BK_STAGES) & 1));
        uint8_t* buf = in_buf + (size_t)s * in_bytes;
        const uint8_t* in = buf + (reinterpret_cast<uintptr_t>(g.src) & 15);
        int4 v[BK_VEC_PER_THREAD];
#pragma unroll
        for (int k = 0; k < BK_VEC_PER_THREAD; ++k) {
            const int c = tid + k * BK_THREADS;
            const int keep = g.len - c * 16;
            v[k] = make_int4(0, 0, 0, 0);
            if (c < vpr && keep > 0) {
                v[k] = lds16_shifted(in + c * 16);
                if (keep < 16) v[k] = mask_tail16(v[k], keep);
            }
        }
        __syncthreads();   // every vector has been read: the row can be written back at its aligned position
#pragma unroll
        for (int k = 0; k < BK_VEC_PER_THREAD; ++k) {
            const int c = tid + k * BK_THREADS;
            if (c < vpr) reinterpret_cast<int4*>(buf)[c] = v[k];
        }
        tc::fence_proxy_async_smem();
        __syncthreads();
        if (tid == 0) {
            tc::bulk_store_1d(g.dst, buf, (uint32_t)row_bytes);
            tc::bulk_commit();
            // refill the buffer whose row was stored BK_LAG rows ago (its store has had time to read shared memory;
            // waiting for the store just issued would put the copy engine's drain on every row's critical path)
            const long long rn = r + (BK_STAGES - BK_LAG) * stride;
            if (i >= BK_LAG && rn < R) {
                tc::bulk_wait_read<BK_LAG>();
                fetch(rn, (int)((i + BK_STAGES - BK_LAG) % BK_STAGES));
            }
        }
    }
    if (tid == 0) tc::bulk_wait_all();
}

// out_values[off_out[q] : off_out[q+1]] = packed[r, :len] (unpack_vec_kernel's contract).  The live part of the padded
// row arrives as one bulk copy; the 16-byte aligned interior of the ragged destination leaves as one bulk copy from a
// realigned staging buffer; the two ragged ends (< 16 bytes each, shared with the neighbouring rows' vectors) are
// written element-wise.
__global__ void __launch_bounds__(BK_THREADS)
unpack_bulk_kernel(const uint8_t* __restrict__ packed, const int64_t* __restrict__ off_out, const int32_t* __restrict__ dst_index,
                   long long R, int elem_bytes, int row_bytes, uint8_t* __restrict__ out_values) {
    extern __shared__ __align__(128) uint8_t bk_smem[];
    __shared__ __align__(8) uint64_t full[BK_STAGES];
    const int tid = threadIdx.x;
    const int in_bytes = row_bytes + 32;
    uint8_t* in_buf = bk_smem;   // [BK_STAGES][in_bytes]: a row is realigned in place
    if (tid == 0)
        for (int s = 0; s < BK_STAGES; ++s) tc::mbar_init(&full[s], 1);
    __syncthreads();
    auto seg = [&](long long r) {
        const long long q = dst_index ? dst_index[r] : r;
        const long long o0 = off_out[q];
        const long long full_len = (off_out[q + 1] - o0) * elem_bytes;
        BulkSeg g;
        g.src = packed + r * (long long)row_bytes;
        g.dst = out_values + o0 * elem_bytes;
        g.len = (int)(full_len < row_bytes ? full_len : row_bytes);
        return g;
    };
    auto fetch = [&](long long r, int s) {
        const BulkSeg g = seg(r);
        const uint32_t bytes = g.len > 0 ? (uint32_t)((g.len + 15) & ~15) : 0u;   // padded rows are 16-byte aligned
        tc::mbar_arrive_expect_tx(&full[s], bytes);
        if (bytes) tc::bulk_load_1d(in_buf + (size_t)s * in_bytes, g.src, bytes, &full[s]);
    };
    const long long r0 = blockIdx.x, stride = gridDim.x;
    if (tid == 0)
        for (int s = 0; s < BK_STAGES; ++s)
            if (r0 + s * stride < R) fetch(r0 + s * stride, s);
    long long i = 0;
    for (long long r = r0; r < R; r += stride, ++i) {
        const int s = (int)(i % BK_STAGES);
        const BulkSeg g = seg(r);
        tc::mbar_wait(&full[s], (uint32_t)((i / BK_STAGES) & 1));
        uint8_t* buf = in_buf + (size_t)s * in_bytes;
        const uintptr_t d0 = reinterpret_cast<uintptr_t>(g.dst), d1 = d0 + (uintptr_t)g.len;
        uintptr_t A0 = (d0 + 15) & ~uintptr_t(15), A1 = d1 & ~uintptr_t(15);
        if (A1 < A0) A1 = A0 = d1;          // the row lies inside one vector: everything is "ragged end"
        const int head = (int)((A0 < d1 ? A0 : d1) - d0);   // bytes before the aligned interior
        const int nvec = (int)((A1 - A0) >> 4);
        // ragged ends, element-wise from the staged row: [d0, A0) and [A1, d1)
        const int tail0 = head + nvec * 16, ntail = g.len - tail0;
        if ((elem_bytes & 3) == 0) {
            const int nh = head >> 2, nt = ntail >> 2;
            if (tid < nh) reinterpret_cast<unsigned*>(g.dst)[tid] = reinterpret_cast<const unsigned*>(buf)[tid];
            if (tid >= 32 && tid - 32 < nt)
                reinterpret_cast<unsigned*>(g.dst + tail0)[tid - 32] = reinterpret_cast<const unsigned*>(buf + tail0)[tid - 32];
        } else {
            if (tid < head) g.dst[tid] = buf[tid];
            if (tid >= 32 && tid - 32 < ntail) g.dst[tail0 + tid - 32] = buf[tail0 + tid - 32];
        }
        int4 v[BK_VEC_PER_THREAD];   // interior: source byte of A0 + 16 c is row byte head + 16 c
#pragma unroll
        for (int k = 0; k < BK_VEC_PER_THREAD; ++k) {
            const int c = tid + k * BK_THREADS;
            if (c < nvec) v[k] = lds16_shifted(buf + head + c * 16);
        }
        __syncthreads();   // ends and interior have been read: the interior can be written back at aligned positions
#pragma unroll
        for (int k = 0; k < BK_VEC_PER_THREAD; ++k) {
            const int c = tid + k * BK_THREADS;
            if (c < nvec) reinterpret_cast<int4*>(buf)[c] = v[k];
        }
        tc::fence_proxy_async_smem();
        __syncthreads();
        if (tid == 0) {
            if (nvec) tc::bulk_store_1d(reinterpret_cast<void*>(A0), buf, (uint32_t)nvec * 16u);
            tc::bulk_commit();
            const long long rn = r + (BK_STAGES - BK_LAG) * stride;
            if (i >= BK_LAG && rn < R) {
                tc::bulk_wait_read<BK_LAG>();
                fetch(rn, (int)((i + BK_STAGES - BK_LAG) % BK_STAGES));
            }
        }
    }
    if (tid == 0) tc::bulk_wait_all();
}

static int grid_for(long long work_items) {
    const long long blocks = (work_items + PACK_THREADS - 1) / PACK_THREADS;
    const long long cap = (long long)num_sms() * 16;  // 16 resident CTAs of 256 threads cover 2 full waves
    if (blocks <= cap) return (int)(blocks > 0 ? blocks : 1);
    return (int)cap;
}

// The bulk kernels: padded rows of 1-16 KB, 4-byte aligned ragged buffer (the ragged ends are written as 32-bit words).
// Opt-in (DIE_PACK_BULK=1): measured on B200 at the C3 shape (ncu, 16384 rows of 16..2048 fp32 -> 8 KB rows) they are no
// faster than the vector kernels -- pack 44.6 vs 42.0 us, unpack 40.4 vs 35.0 us, with 3 or 6 rows in flight per CTA
// and with the refill lagged behind the store -- so per-thread load latency is not what holds these copies at 0.7 /
// 0.6 of the copy peak; the vector kernels stay the default.
static bool bulk_rows_ok(long long row_bytes, const void* ragged) {
    const char* e = getenv("DIE_PACK_BULK");
    return e && atoi(e) && row_bytes >= BK_MIN_ROW && row_bytes <= BK_MAX_ROW && (reinterpret_cast<uintptr_t>(ragged) & 3) == 0;
}
static size_t bulk_smem(long long row_bytes) { return (size_t)BK_STAGES * ((size_t)row_bytes + 32); }
static die_status_t bulk_configure() {
    static bool done = false;
    if (!done) {
        const int max_smem = (int)bulk_smem(BK_MAX_ROW);
        DIE_CUDA(cudaFuncSetAttribute(pack_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
        DIE_CUDA(cudaFuncSetAttribute(unpack_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
        done = true;
    }
    return DIE_OK;
}
static int bulk_grid(long long R, size_t smem) {
    const long long per_sm = std::max<long long>(1, std::min<long long>(8, (220 * 1024) / (long long)(smem + 1024)));
    return (int)std::max<long long>(1, std::min<long long>(R, per_sm * num_sms()));
}

}  // namespace die

using namespace die;

extern "C" die_status_t die_pack_ragged(const void* values, const int64_t* off, const int32_t* src_index,
                                        int32_t R, int32_t elem_bytes, int32_t Lpad, void* out,
                                        int32_t* lengths_out, die_stream_t stream) {
    if (R < 0 || elem_bytes <= 0 || Lpad < 0) DIE_FAIL(DIE_ERR_INVALID, "die_pack_ragged: bad sizes");
    if (R == 0 || Lpad == 0) return DIE_OK;
    if (!off || !out) DIE_FAIL(DIE_ERR_INVALID, "die_pack_ragged: null off/out");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const long long row_bytes = (long long)Lpad * elem_bytes;
    const bool vec = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    if (vec && bulk_rows_ok(row_bytes, values)) {
        const size_t smem = bulk_smem(row_bytes);
        DIE_TRY(bulk_configure());
        DIE_LAUNCH(ST_PACK, s, pack_bulk_kernel<<<bulk_grid(R, smem), BK_THREADS, smem, s>>>(
            static_cast<const uint8_t*>(values), off, src_index, R, elem_bytes, (int)row_bytes, static_cast<uint8_t*>(out),
            lengths_out));
    } else if (vec) {
        const int vpr = (int)(row_bytes / 16);
        DIE_LAUNCH(ST_PACK, s, pack_vec_kernel<<<grid_for((long long)R * vpr), PACK_THREADS, 0, s>>>(
            static_cast<const uint8_t*>(values), off, src_index, R, elem_bytes, vpr, static_cast<int4*>(out),
            lengths_out));
    } else {
        DIE_LAUNCH(ST_PACK, s, pack_elem_kernel<<<grid_for((long long)R * Lpad), PACK_THREADS, 0, s>>>(
            static_cast<const uint8_t*>(values), off, src_index, R, elem_bytes, Lpad, static_cast<uint8_t*>(out),
            lengths_out));
    }
    return DIE_OK;
}

extern "C" die_status_t die_unpack_ragged(const void* packed, const int64_t* off_out, const int32_t* dst_index,
                                          int32_t R, int32_t elem_bytes, int32_t Lpad, void* out_values,
                                          die_stream_t stream) {
    if (R < 0 || elem_bytes <= 0 || Lpad < 0) DIE_FAIL(DIE_ERR_INVALID, "die_unpack_ragged: bad sizes");
    if (R == 0 || Lpad == 0) return DIE_OK;
    if (!off_out || !packed) DIE_FAIL(DIE_ERR_INVALID, "die_unpack_ragged: null off_out/packed");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const long long row_bytes = (long long)Lpad * elem_bytes;
    const bool vec = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(packed) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(out_values) % elem_bytes) == 0 &&
                     (elem_bytes == 1 || elem_bytes == 2 || (elem_bytes & 3) == 0);
    if (vec && bulk_rows_ok(row_bytes, out_values)) {
        const size_t smem = bulk_smem(row_bytes);
        DIE_TRY(bulk_configure());
        DIE_LAUNCH(ST_UNPACK, s, unpack_bulk_kernel<<<bulk_grid(R, smem), BK_THREADS, smem, s>>>(
            static_cast<const uint8_t*>(packed), off_out, dst_index, R, elem_bytes, (int)row_bytes,
            static_cast<uint8_t*>(out_values)));
    } else if (vec) {
        const int vpr1 = (int)(row_bytes / 16) + 1;
        DIE_LAUNCH(ST_UNPACK, s, unpack_vec_kernel<<<grid_for((long long)R * 32), PACK_THREADS, 0, s>>>(
            static_cast<const uint8_t*>(packed), off_out, dst_index, R, elem_bytes, (int)row_bytes, vpr1,
            static_cast<uint8_t*>(out_values)));
    } else {
        DIE_LAUNCH(ST_UNPACK, s, unpack_elem_kernel<<<grid_for((long long)R * Lpad), PACK_THREADS, 0, s>>>(
            static_cast<const uint8_t*>(packed), off_out, dst_index, R, elem_bytes, Lpad,
            static_cast<uint8_t*>(out_values)));
    }
    return DIE_OK;
}
