// Device helpers shared by the pack / unpack kernels and the fused pipeline kernels.
#pragma once
#include "common.cuh"

namespace die {

// Load the 16 bytes at `src` (any alignment) with at most two aligned 128-bit loads.
// Only bytes [first, last) of the 16 are guaranteed meaningful; no aligned 16-byte chunk that
// holds none of those bytes is touched, so ragged row ends never read outside their buffer.
__device__ __forceinline__ int4 load16_shifted(const uint8_t* src, int first, int last) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(src);
    const unsigned m = (unsigned)(a & 15);
    const int4* lo_p = reinterpret_cast<const int4*>(a - m);
    if (m == 0) return ldg_stream(lo_p);
    int4 lo = make_int4(0, 0, 0, 0), hi = make_int4(0, 0, 0, 0);
    const int split = 16 - (int)m;  // bytes [0, split) live in lo, [split, 16) in hi
    if (first < split && first < last) lo = ldg_stream(lo_p);
    if (last > split) hi = ldg_stream(lo_p + 1);
    const unsigned w[8] = {(unsigned)lo.x, (unsigned)lo.y, (unsigned)lo.z, (unsigned)lo.w,
                           (unsigned)hi.x, (unsigned)hi.y, (unsigned)hi.z, (unsigned)hi.w};
    const unsigned sb = (m & 3) * 8;
    unsigned o[4];
    switch (m >> 2) {
        case 0:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j], w[j + 1], sb);
            break;
        case 1:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j + 1], w[j + 2], sb);
            break;
        case 2:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j + 2], w[j + 3], sb);
            break;
        default:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j + 3], w[j + 4], sb);
            break;
    }
    return make_int4((int)o[0], (int)o[1], (int)o[2], (int)o[3]);
}

// The 16 bytes at shared-memory address `src` (any alignment) from the two aligned vectors around them.  The caller's
// buffer extends to the next 16-byte boundary past the last byte it wants (staging buffers are padded by 16).
__device__ __forceinline__ int4 lds16_shifted(const uint8_t* src) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(src);
    const unsigned m = (unsigned)(a & 15);
    const int4* lo_p = reinterpret_cast<const int4*>(a - m);
    const int4 lo = lo_p[0];
    if (m == 0) return lo;
    const int4 hi = lo_p[1];
    const unsigned w[8] = {(unsigned)lo.x, (unsigned)lo.y, (unsigned)lo.z, (unsigned)lo.w,
                           (unsigned)hi.x, (unsigned)hi.y, (unsigned)hi.z, (unsigned)hi.w};
    const unsigned sb = (m & 3) * 8;
    unsigned o[4];
    switch (m >> 2) {
        case 0:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j], w[j + 1], sb);
            break;
        case 1:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j + 1], w[j + 2], sb);
            break;
        case 2:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j + 2], w[j + 3], sb);
            break;
        default:
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = __funnelshift_r(w[j + 3], w[j + 4], sb);
            break;
    }
    return make_int4((int)o[0], (int)o[1], (int)o[2], (int)o[3]);
}

// Zero the bytes of v at positions >= keep (0..16).
__device__ __forceinline__ int4 mask_tail16(int4 v, int keep) {
    unsigned w[4] = {(unsigned)v.x, (unsigned)v.y, (unsigned)v.z, (unsigned)v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k = keep - 4 * j;  // bytes of word j to keep
        if (k <= 0)
            w[j] = 0u;
        else if (k < 4)
            w[j] &= (1u << (8 * k)) - 1u;
    }
    return make_int4((int)w[0], (int)w[1], (int)w[2], (int)w[3]);
}

}  // namespace die
