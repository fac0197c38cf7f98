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

static int grid_for(long long work_items) {
    const long long blocks = (work_items + PACK_THREADS - 1) / PACK_THREADS;
    const long long cap = (long long)num_sms() * 16;  // 16 resident CTAs of 256 threads cover 2 full waves
    if (blocks <= cap) return (int)(blocks > 0 ? blocks : 1);
    return (int)cap;
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
    if (vec) {
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
    if (vec) {
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
