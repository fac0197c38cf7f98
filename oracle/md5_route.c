/* CPU oracle (TEST INFRASTRUCTURE, never linked into the product): plain-C restatement of the
 * reference's routing arithmetic.
 *
 *   key_hash = int(hashlib.md5(key.encode()).hexdigest(), 16)     src/model_registry.py:159
 *   shard    = shards[key_hash % len(shards)]                     src/model_registry.py:160-161
 *   gate + failover: healthy[key_hash % len(healthy)]             src/router.py:166-184, 197-215
 *
 * hashlib's MD5 is not part of the reference tree (CPython stdlib / OpenSSL); it is restated here
 * from RFC 1321 with the sine-derived constants computed at start-up, and pinned by
 * tests/test_oracle_golden.py against digests produced by the reference's own call site.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

static uint32_t K[64];
static const int S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22,
                          5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20, 5, 9,  14, 20,
                          4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23,
                          6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
static int k_ready = 0;

static void init_k(void) {
    for (int i = 0; i < 64; ++i) K[i] = (uint32_t)floor(fabs(sin((double)(i + 1))) * 4294967296.0);
    k_ready = 1;
}

static uint32_t rol(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }

static void block(uint32_t h[4], const uint8_t* p) {
    uint32_t m[16];
    for (int i = 0; i < 16; ++i)
        m[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) |
               ((uint32_t)p[4 * i + 3] << 24);
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    for (int i = 0; i < 64; ++i) {
        uint32_t f;
        int g;
        if (i < 16) {
            f = (b & c) | (~b & d);
            g = i;
        } else if (i < 32) {
            f = (d & b) | (~d & c);
            g = (5 * i + 1) & 15;
        } else if (i < 48) {
            f = b ^ c ^ d;
            g = (3 * i + 5) & 15;
        } else {
            f = c ^ (b | ~d);
            g = (7 * i) & 15;
        }
        const uint32_t t = d;
        d = c;
        c = b;
        b = b + rol(a + f + K[i] + m[g], S[i]);
        a = t;
    }
    h[0] += a;
    h[1] += b;
    h[2] += c;
    h[3] += d;
}

/* digest[16] of msg[0..len) */
void oracle_md5(const uint8_t* msg, uint64_t len, uint8_t digest[16]) {
    if (!k_ready) init_k();
    uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    uint64_t done = 0;
    for (; done + 64 <= len; done += 64) block(h, msg + done);
    uint8_t tail[128];
    const uint64_t rem = len - done;
    memset(tail, 0, sizeof tail);
    memcpy(tail, msg + done, rem);
    tail[rem] = 0x80;
    const uint64_t padded = rem + 9 <= 64 ? 64 : 128;
    const uint64_t bits = len * 8;
    for (int i = 0; i < 8; ++i) tail[padded - 8 + i] = (uint8_t)(bits >> (8 * i));
    block(h, tail);
    if (padded == 128) block(h, tail + 64);
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) digest[4 * i + j] = (uint8_t)(h[i] >> (8 * j));
}

/* big-endian 128-bit integer of the digest, modulo n (Horner over the bytes) */
static uint32_t digest_mod(const uint8_t d[16], uint32_t n) {
    uint64_t r = 0;
    for (int i = 0; i < 16; ++i) r = (r * 256 + d[i]) % n;
    return (uint32_t)r;
}

/* Batched route_request over CSR keys; shard_out[i] = list position or -1 (None). */
void oracle_route(const uint8_t* key_bytes, const int64_t* key_off, int64_t B, int32_t n_shards,
                  const uint8_t* routable, const int32_t* healthy_idx, int32_t n_healthy, int32_t failover,
                  int32_t* shard_out, uint8_t* digest_out /* [B,16] or NULL */) {
    for (int64_t i = 0; i < B; ++i) {
        uint8_t d[16];
        oracle_md5(key_bytes + key_off[i], (uint64_t)(key_off[i + 1] - key_off[i]), d);
        if (digest_out) memcpy(digest_out + 16 * i, d, 16);
        int32_t s = -1;
        if (n_shards > 0) {
            const uint32_t p = digest_mod(d, (uint32_t)n_shards);
            if (!routable || routable[p])
                s = (int32_t)p;
            else if (failover && n_healthy > 0)
                s = healthy_idx[digest_mod(d, (uint32_t)n_healthy)];
        }
        shard_out[i] = s;
    }
}
