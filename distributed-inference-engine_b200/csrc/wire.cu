// Ingress data format: length-prefixed binary request frames -> the device CSR buffers the path consumes
// (SURVEY.md section 8(f4)).  The reference's worker reads one JSON document per TCP connection
// (src/worker.py:92-100: reader.read(4096), json.loads) whose fields are model, version, inputs and request_id; a
// frame carries the same fields in binary (include/die.h, die_frame_header_t).  The transport stays with the host; what
// it received -- any number of frames back to back in one buffer, plus where each starts, which the receive loop
// knows from the length prefix -- is decoded here in two launches:
//   frame_scan_kernel  one CTA: validate every header, exclusive scan of the key lengths -> key_off[n + 1]
//   frame_copy_kernel  one warp per frame: key bytes -> key_bytes[key_off[i] ..), payload -> row i of [n, dim]
//                      (zero-padded: this is PACK, batcher.py:211, done at the door), request id, queue, status
// A malformed frame (bad magic, lengths that do not fit the frame, more floats than the row holds) gets status != 0, an
// empty key and a zero row; *n_bad counts them.  Nothing is trusted: every length is checked against the frame's extent.
#include "common.cuh"

namespace die {

constexpr uint32_t FRAME_MAGIC = 0x31454944u;  // "DIE1"

struct FrameInfo {
    uint32_t key_len, n_floats;
    uint8_t status;
};

__device__ __forceinline__ FrameInfo frame_check(const uint8_t* __restrict__ frames, const int64_t* __restrict__ off, int i,
                                                 int dim, die_frame_header_t& h) {
    FrameInfo f{0u, 0u, DIE_FRAME_OK};
    const int64_t o = off[i], len = off[i + 1] - o;
    if (len < (int64_t)sizeof(die_frame_header_t) || (o & 3)) {
        f.status = DIE_FRAME_BAD;
        return f;
    }
    const uint32_t* w = reinterpret_cast<const uint32_t*>(frames + o);
    h.magic = w[0];
    h.frame_bytes = w[1];
    h.request_id = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
    h.key_len = (uint16_t)(w[4] & 0xffffu);
    h.queue = (uint16_t)(w[4] >> 16);
    h.n_floats = w[5];
    const uint64_t key_pad = ((uint64_t)h.key_len + 3u) & ~3ull;
    const uint64_t need = sizeof(die_frame_header_t) + key_pad + 4ull * h.n_floats;
    if (h.magic != FRAME_MAGIC || (int64_t)h.frame_bytes != len || need != (uint64_t)len) {
        f.status = DIE_FRAME_BAD;
        return f;
    }
    if (h.n_floats > (uint32_t)dim) {
        f.status = DIE_FRAME_TOO_WIDE;
        return f;
    }
    f.key_len = h.key_len;
    f.n_floats = h.n_floats;
    return f;
}

__global__ void __launch_bounds__(1024)
frame_scan_kernel(const uint8_t* __restrict__ frames, const int64_t* __restrict__ off, int n, int dim,
                  int64_t* __restrict__ key_off, uint8_t* __restrict__ status, uint32_t* __restrict__ n_bad) {
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned long long s_carry;
    __shared__ unsigned s_bad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s_carry = 0ull;
        s_bad = 0u;
    }
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        unsigned long long v = 0ull;
        if (i < n) {
            die_frame_header_t h;
            const FrameInfo f = frame_check(frames, off, i, dim, h);
            v = f.key_len;
            status[i] = f.status;
            if (f.status != DIE_FRAME_OK) atomicAdd(&s_bad, 1u);
        }
        unsigned long long x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned long long t = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            s_warp[lane] = t;
        }
        __syncthreads();
        const unsigned long long before = s_carry + (warp ? s_warp[warp - 1] : 0ull) + x - v;
        if (i < n) key_off[i] = (int64_t)before;
        __syncthreads();
        if (tid == 1023) s_carry = before + v;
        __syncthreads();
    }
    if (tid == 0) {
        key_off[n] = (int64_t)s_carry;
        *n_bad = s_bad;
    }
}

__global__ void __launch_bounds__(256)
frame_copy_kernel(const uint8_t* __restrict__ frames, const int64_t* __restrict__ off, int n, int dim,
                  const int64_t* __restrict__ key_off, const uint8_t* __restrict__ status, uint8_t* __restrict__ key_bytes,
                  float* __restrict__ payload, uint64_t* __restrict__ request_id, int32_t* __restrict__ queue) {
    const int lane = threadIdx.x & 31;
    const long long warps = (long long)gridDim.x * 8;
    for (long long i = blockIdx.x * 8ll + (threadIdx.x >> 5); i < n; i += warps) {
        float* row = payload + i * dim;
        if (status[i] != DIE_FRAME_OK) {
            for (int c = lane; c < dim; c += 32) row[c] = 0.f;
            if (lane == 0) {
                request_id[i] = 0ull;
                queue[i] = -1;
            }
            continue;
        }
        const uint8_t* f = frames + off[i];
        const uint32_t* w = reinterpret_cast<const uint32_t*>(f);
        const uint32_t key_len = w[4] & 0xffffu, n_floats = w[5];
        if (lane == 0) {
            request_id[i] = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
            queue[i] = (int32_t)(w[4] >> 16);
        }
        const uint8_t* kb = f + sizeof(die_frame_header_t);
        uint8_t* kd = key_bytes + key_off[i];
        for (uint32_t c = lane; c < key_len; c += 32) kd[c] = kb[c];
        const float* x = reinterpret_cast<const float*>(kb + ((key_len + 3u) & ~3u));
        for (int c = lane; c < dim; c += 32) row[c] = c < (int)n_floats ? x[c] : 0.f;
    }
}

}  // namespace die

using namespace die;

extern "C" die_status_t die_frame_decode(const uint8_t* frames, const int64_t* frame_off, int32_t n, int32_t dim,
                                         uint8_t* key_bytes_out, int64_t* key_off_out, float* payload_out,
                                         uint64_t* request_id_out, int32_t* queue_out, uint8_t* status_out,
                                         uint32_t* n_bad_out, die_stream_t stream) {
    if (n < 0 || dim <= 0) DIE_FAIL(DIE_ERR_INVALID, "die_frame_decode: n >= 0 and dim > 0 required");
    if (!frame_off || !key_off_out || !n_bad_out ||
        (n > 0 && (!frames || !key_bytes_out || !payload_out || !request_id_out || !queue_out || !status_out)))
        DIE_FAIL(DIE_ERR_INVALID, "die_frame_decode: null buffer");
    if (reinterpret_cast<uintptr_t>(frames) & 3) DIE_FAIL(DIE_ERR_INVALID, "die_frame_decode: the frame buffer must be 4-byte aligned");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DIE_LAUNCH(ST_PACK, s, frame_scan_kernel<<<1, 1024, 0, s>>>(frames, frame_off, n, dim, key_off_out, status_out, n_bad_out));
    if (n > 0) {
        const int grid = (int)std::min<long long>(((long long)n + 7) / 8, (long long)num_sms() * 8);
        DIE_LAUNCH(ST_PACK, s, frame_copy_kernel<<<grid, 256, 0, s>>>(frames, frame_off, n, dim, key_off_out, status_out,
                                                                     key_bytes_out, payload_out, request_id_out, queue_out));
    }
    return DIE_OK;
}
