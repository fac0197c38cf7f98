// Host build of the PRODUCT's MD5 device code (csrc/md5.cuh is __host__ __device__) so the exact
// arithmetic the kernel runs can be compared with hashlib on a machine without a GPU.
// TEST INFRASTRUCTURE: built into oracle/_build/, loaded only by tests/.
#include <stdint.h>

#include "../distributed-inference-engine_b200/csrc/md5.cuh"

namespace {
struct HostKeySrc {
    const uint8_t* base;
    uint32_t byte(uint32_t p) const { return base[p]; }
    uint32_t word(uint32_t p) const {
        return (uint32_t)base[p] | ((uint32_t)base[p + 1] << 8) | ((uint32_t)base[p + 2] << 16) |
               ((uint32_t)base[p + 3] << 24);
    }
};
}  // namespace

extern "C" void kernel_md5_host(const uint8_t* key_bytes, const int64_t* key_off, int64_t B, int32_t n_shards,
                                uint32_t* digest_out /*[B,4]*/, int32_t* shard_out) {
    for (int64_t i = 0; i < B; ++i) {
        const die::Md5State st = die::md5_key(HostKeySrc{key_bytes + key_off[i]}, (uint32_t)(key_off[i + 1] - key_off[i]));
        digest_out[4 * i + 0] = st.a;
        digest_out[4 * i + 1] = st.b;
        digest_out[4 * i + 2] = st.c;
        digest_out[4 * i + 3] = st.d;
        shard_out[i] = n_shards > 0 ? (int32_t)die::md5_mod(st, (uint32_t)n_shards) : -1;
    }
}
