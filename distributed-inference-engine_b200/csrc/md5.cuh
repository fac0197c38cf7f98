// MD5 (RFC 1321) for short keys, one key per thread.  Replaces hashlib.md5(key.encode())
// at src/model_registry.py:159 and src/router.py:214.  Host+device so the host build can be
// checked against hashlib on a CPU-only machine (tests/test_cabi_cpu.py).
#pragma once
#include <stdint.h>

#if defined(__CUDA_ARCH__)
#define DIE_ROTL(x, s) __funnelshift_l((x), (x), (s))
#else
#define DIE_ROTL(x, s) (((x) << (s)) | ((x) >> (32 - (s))))
#endif

#ifdef __CUDACC__
#define DIE_HD __host__ __device__ __forceinline__
#else
#define DIE_HD inline
#endif

namespace die {

struct Md5State {
    uint32_t a, b, c, d;
};

DIE_HD Md5State md5_init() { return Md5State{0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u}; }

#define DIE_MD5_F(x, y, z) ((z) ^ ((x) & ((y) ^ (z))))
#define DIE_MD5_G(x, y, z) ((y) ^ ((z) & ((x) ^ (y))))
#define DIE_MD5_H(x, y, z) ((x) ^ (y) ^ (z))
#define DIE_MD5_I(x, y, z) ((y) ^ ((x) | ~(z)))
#define DIE_MD5_STEP(f, a, b, c, d, x, t, s)        \
    (a) += f((b), (c), (d)) + (x) + (uint32_t)(t);  \
    (a) = DIE_ROTL((a), (s));                       \
    (a) += (b);

// One 64-byte block; m[] are the 16 little-endian message words.
DIE_HD void md5_block(Md5State& st, const uint32_t (&m)[16]) {
    uint32_t a = st.a, b = st.b, c = st.c, d = st.d;
    DIE_MD5_STEP(DIE_MD5_F, a, b, c, d, m[0], 0xd76aa478, 7)
    DIE_MD5_STEP(DIE_MD5_F, d, a, b, c, m[1], 0xe8c7b756, 12)
    DIE_MD5_STEP(DIE_MD5_F, c, d, a, b, m[2], 0x242070db, 17)
    DIE_MD5_STEP(DIE_MD5_F, b, c, d, a, m[3], 0xc1bdceee, 22)
    DIE_MD5_STEP(DIE_MD5_F, a, b, c, d, m[4], 0xf57c0faf, 7)
    DIE_MD5_STEP(DIE_MD5_F, d, a, b, c, m[5], 0x4787c62a, 12)
    DIE_MD5_STEP(DIE_MD5_F, c, d, a, b, m[6], 0xa8304613, 17)
    DIE_MD5_STEP(DIE_MD5_F, b, c, d, a, m[7], 0xfd469501, 22)
    DIE_MD5_STEP(DIE_MD5_F, a, b, c, d, m[8], 0x698098d8, 7)
    DIE_MD5_STEP(DIE_MD5_F, d, a, b, c, m[9], 0x8b44f7af, 12)
    DIE_MD5_STEP(DIE_MD5_F, c, d, a, b, m[10], 0xffff5bb1, 17)
    DIE_MD5_STEP(DIE_MD5_F, b, c, d, a, m[11], 0x895cd7be, 22)
    DIE_MD5_STEP(DIE_MD5_F, a, b, c, d, m[12], 0x6b901122, 7)
    DIE_MD5_STEP(DIE_MD5_F, d, a, b, c, m[13], 0xfd987193, 12)
    DIE_MD5_STEP(DIE_MD5_F, c, d, a, b, m[14], 0xa679438e, 17)
    DIE_MD5_STEP(DIE_MD5_F, b, c, d, a, m[15], 0x49b40821, 22)

    DIE_MD5_STEP(DIE_MD5_G, a, b, c, d, m[1], 0xf61e2562, 5)
    DIE_MD5_STEP(DIE_MD5_G, d, a, b, c, m[6], 0xc040b340, 9)
    DIE_MD5_STEP(DIE_MD5_G, c, d, a, b, m[11], 0x265e5a51, 14)
    DIE_MD5_STEP(DIE_MD5_G, b, c, d, a, m[0], 0xe9b6c7aa, 20)
    DIE_MD5_STEP(DIE_MD5_G, a, b, c, d, m[5], 0xd62f105d, 5)
    DIE_MD5_STEP(DIE_MD5_G, d, a, b, c, m[10], 0x02441453, 9)
    DIE_MD5_STEP(DIE_MD5_G, c, d, a, b, m[15], 0xd8a1e681, 14)
    DIE_MD5_STEP(DIE_MD5_G, b, c, d, a, m[4], 0xe7d3fbc8, 20)
    DIE_MD5_STEP(DIE_MD5_G, a, b, c, d, m[9], 0x21e1cde6, 5)
    DIE_MD5_STEP(DIE_MD5_G, d, a, b, c, m[14], 0xc33707d6, 9)
    DIE_MD5_STEP(DIE_MD5_G, c, d, a, b, m[3], 0xf4d50d87, 14)
    DIE_MD5_STEP(DIE_MD5_G, b, c, d, a, m[8], 0x455a14ed, 20)
    DIE_MD5_STEP(DIE_MD5_G, a, b, c, d, m[13], 0xa9e3e905, 5)
    DIE_MD5_STEP(DIE_MD5_G, d, a, b, c, m[2], 0xfcefa3f8, 9)
    DIE_MD5_STEP(DIE_MD5_G, c, d, a, b, m[7], 0x676f02d9, 14)
    DIE_MD5_STEP(DIE_MD5_G, b, c, d, a, m[12], 0x8d2a4c8a, 20)

    DIE_MD5_STEP(DIE_MD5_H, a, b, c, d, m[5], 0xfffa3942, 4)
    DIE_MD5_STEP(DIE_MD5_H, d, a, b, c, m[8], 0x8771f681, 11)
    DIE_MD5_STEP(DIE_MD5_H, c, d, a, b, m[11], 0x6d9d6122, 16)
    DIE_MD5_STEP(DIE_MD5_H, b, c, d, a, m[14], 0xfde5380c, 23)
    DIE_MD5_STEP(DIE_MD5_H, a, b, c, d, m[1], 0xa4beea44, 4)
    DIE_MD5_STEP(DIE_MD5_H, d, a, b, c, m[4], 0x4bdecfa9, 11)
    DIE_MD5_STEP(DIE_MD5_H, c, d, a, b, m[7], 0xf6bb4b60, 16)
    DIE_MD5_STEP(DIE_MD5_H, b, c, d, a, m[10], 0xbebfbc70, 23)
    DIE_MD5_STEP(DIE_MD5_H, a, b, c, d, m[13], 0x289b7ec6, 4)
    DIE_MD5_STEP(DIE_MD5_H, d, a, b, c, m[0], 0xeaa127fa, 11)
    DIE_MD5_STEP(DIE_MD5_H, c, d, a, b, m[3], 0xd4ef3085, 16)
    DIE_MD5_STEP(DIE_MD5_H, b, c, d, a, m[6], 0x04881d05, 23)
    DIE_MD5_STEP(DIE_MD5_H, a, b, c, d, m[9], 0xd9d4d039, 4)
    DIE_MD5_STEP(DIE_MD5_H, d, a, b, c, m[12], 0xe6db99e5, 11)
    DIE_MD5_STEP(DIE_MD5_H, c, d, a, b, m[15], 0x1fa27cf8, 16)
    DIE_MD5_STEP(DIE_MD5_H, b, c, d, a, m[2], 0xc4ac5665, 23)

    DIE_MD5_STEP(DIE_MD5_I, a, b, c, d, m[0], 0xf4292244, 6)
    DIE_MD5_STEP(DIE_MD5_I, d, a, b, c, m[7], 0x432aff97, 10)
    DIE_MD5_STEP(DIE_MD5_I, c, d, a, b, m[14], 0xab9423a7, 15)
    DIE_MD5_STEP(DIE_MD5_I, b, c, d, a, m[5], 0xfc93a039, 21)
    DIE_MD5_STEP(DIE_MD5_I, a, b, c, d, m[12], 0x655b59c3, 6)
    DIE_MD5_STEP(DIE_MD5_I, d, a, b, c, m[3], 0x8f0ccc92, 10)
    DIE_MD5_STEP(DIE_MD5_I, c, d, a, b, m[10], 0xffeff47d, 15)
    DIE_MD5_STEP(DIE_MD5_I, b, c, d, a, m[1], 0x85845dd1, 21)
    DIE_MD5_STEP(DIE_MD5_I, a, b, c, d, m[8], 0x6fa87e4f, 6)
    DIE_MD5_STEP(DIE_MD5_I, d, a, b, c, m[15], 0xfe2ce6e0, 10)
    DIE_MD5_STEP(DIE_MD5_I, c, d, a, b, m[6], 0xa3014314, 15)
    DIE_MD5_STEP(DIE_MD5_I, b, c, d, a, m[13], 0x4e0811a1, 21)
    DIE_MD5_STEP(DIE_MD5_I, a, b, c, d, m[4], 0xf7537e82, 6)
    DIE_MD5_STEP(DIE_MD5_I, d, a, b, c, m[11], 0xbd3af235, 10)
    DIE_MD5_STEP(DIE_MD5_I, c, d, a, b, m[2], 0x2ad7d2bb, 15)
    DIE_MD5_STEP(DIE_MD5_I, b, c, d, a, m[9], 0xeb86d391, 21)
    st.a += a;
    st.b += b;
    st.c += c;
    st.d += d;
}

// Little-endian message word at byte position `pos` of the PADDED message of a `len`-byte key.
// ByteSrc::byte(p) returns key byte p (p < len); ByteSrc::word(p) returns the 4 bytes at p (p+4 <= len).
template <class ByteSrc>
DIE_HD uint32_t md5_msg_word(const ByteSrc& src, uint32_t len, uint32_t pos) {
    if (pos + 4 <= len) return src.word(pos);
    if (pos > len) return 0u;
    uint32_t w = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t p = pos + k;
        const uint32_t byte = p < len ? (uint32_t)src.byte(p) : (p == len ? 0x80u : 0u);
        w |= byte << (8 * k);
    }
    return w;
}

// Full digest of one key. Keys of any length (multi-block) are supported.
template <class ByteSrc>
DIE_HD Md5State md5_key(const ByteSrc& src, uint32_t len) {
    Md5State st = md5_init();
    const uint32_t nblk = (len + 8) / 64 + 1;
    for (uint32_t blk = 0; blk < nblk; ++blk) {
        uint32_t m[16];
#pragma unroll
        for (int w = 0; w < 16; ++w) m[w] = md5_msg_word(src, len, blk * 64 + 4 * w);
        if (blk == nblk - 1) {
            m[14] = len << 3;   // bit length, low word
            m[15] = len >> 29;  // high word
        }
        md5_block(st, m);
    }
    return st;
}

DIE_HD uint32_t bswap32(uint32_t x) {
    return (x >> 24) | ((x >> 8) & 0xff00u) | ((x << 8) & 0xff0000u) | (x << 24);
}

// int(hexdigest, 16) % n : the 16 digest bytes read big-endian (model_registry.py:159-160).
DIE_HD uint32_t md5_mod(const Md5State& st, uint32_t n) {
    if ((n & (n - 1)) == 0) return bswap32(st.d) & (n - 1);  // power of two: low bits of last bytes
    unsigned long long r = bswap32(st.a) % n;
    r = ((r << 32) | bswap32(st.b)) % n;
    r = ((r << 32) | bswap32(st.c)) % n;
    r = ((r << 32) | bswap32(st.d)) % n;
    return (uint32_t)r;
}

}  // namespace die
