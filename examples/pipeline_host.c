/* Plain-C caller of libdie.so: one flushed batch through the composed path (route -> probe -> pack -> infer ->
 * unpack) with host buffers in and host buffers out.  This is what a binding in any language does; the Python
 * package (die/_cabi.py) is the same calls through ctypes.
 *
 *   gcc -std=c99 -Iinclude examples/pipeline_host.c -o pipeline_host \
 *       -Ldistributed-inference-engine_b200/die/_lib -ldie -Wl,-rpath,$PWD/distributed-inference-engine_b200/die/_lib
 *
 * Needs a B200 to run; it only needs gcc to build (tests/test_cpu_host.py builds it).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "die.h"

#define CHECK(call)                                                              \
    do {                                                                         \
        die_status_t st__ = (call);                                              \
        if (st__ != DIE_OK) {                                                    \
            fprintf(stderr, "%s -> %d: %s\n", #call, (int)st__, die_last_error()); \
            return 1;                                                            \
        }                                                                        \
    } while (0)

int main(void) {
    enum { B = 8, DIM = 4 };
    const char* keys[B] = {"user-1", "user-2", "user-3", "user-1", "user-4", "user-2", "user-5", "user-1"};

    /* keys as one byte string + offsets (CSR), 16 bytes of slack after the last key */
    uint8_t key_bytes[256];
    int64_t key_off[B + 1];
    int64_t n = 0;
    memset(key_bytes, 0, sizeof(key_bytes));
    for (int i = 0; i < B; ++i) {
        key_off[i] = n;
        memcpy(key_bytes + n, keys[i], strlen(keys[i]));
        n += (int64_t)strlen(keys[i]);
    }
    key_off[B] = n;

    float payload[B * DIM], out[B * DIM];
    uint8_t hit[B];
    for (int i = 0; i < B * DIM; ++i) payload[i] = (float)i;

    die_pipeline_config_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.kv_max_size = 1024;   /* KVCache(max_size=1024, "lru", default_ttl=60) */
    cfg.default_ttl = 60.0;
    cfg.max_batch = B;
    cfg.dim_in = DIM;
    cfg.dim_out = DIM;
    cfg.model = 0;            /* echo: the reference's FakeModel (fake_model.py:52) */
    cfg.max_key_bytes = 256;

    die_pipeline_t* pipe = NULL;
    printf("libdie ABI %d\n", (int)die_abi_version());
    CHECK(die_set_device(0));
    CHECK(die_pipeline_create(&cfg, &pipe));
    for (int pass = 0; pass < 2; ++pass) { /* second pass: every key is served from the cache */
        CHECK(die_pipeline_run_host(pipe, key_bytes, key_off, B, payload, 100.0 + pass, out, hit, NULL));
        printf("pass %d hits:", pass);
        for (int i = 0; i < B; ++i) printf(" %d", (int)hit[i]);
        printf("\n");
    }
    CHECK(die_pipeline_destroy(pipe));
    return 0;
}
