/* oracle/philox_ref.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Plain-C restatement of the Philox4x32-10 counter-based generator
 * (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11).
 * The product's device implementation lives in covid-abm_b200/csrc/cabm_device.cuh; the two are
 * written independently. tests/test_philox.py checks this one against the published known-answer
 * vectors; the GPU parity tests then require identical draws from the device one.
 *
 * Stream convention shared by oracle, GSL shim tape mode and the CUDA kernels:
 *   key     = (seed_lo, seed_hi)
 *   counter = (block, a, b, (day << 8) | purpose)
 * where a/b are the entity ids the draw is keyed on (person, location / aux), `purpose` is one of
 * the CABM_P_* slots of include/cabm.h and `block` counts groups of four 32-bit draws consumed
 * in order by one event.
 */
#ifndef ORACLE_PHILOX_REF_H
#define ORACLE_PHILOX_REF_H
#include <stdint.h>

static inline void philox4x32_10_ref(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
    uint32_t k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Sequential draw stream for one keyed event. */
typedef struct {
    uint32_t key[2];
    uint32_t a, b, dp;      /* counter words 1..3 */
    uint32_t idx;           /* number of 32-bit draws consumed so far */
    uint32_t buf[4];
} philox_stream_ref;

static inline void philox_stream_begin(philox_stream_ref* s, uint64_t seed, uint32_t day, uint32_t purpose,
                                       uint32_t a, uint32_t b) {
    s->key[0] = (uint32_t)seed; s->key[1] = (uint32_t)(seed >> 32);
    s->a = a; s->b = b; s->dp = (day << 8) | (purpose & 0xFFu);
    s->idx = 0;
}

static inline uint32_t philox_stream_next(philox_stream_ref* s) {
    const uint32_t lane = s->idx & 3u;
    if (lane == 0) {
        uint32_t ctr[4] = { s->idx >> 2, s->a, s->b, s->dp };
        philox4x32_10_ref(ctr, s->key, s->buf);
    }
    s->idx++;
    return s->buf[lane];
}

/* uniform on [0,1) with 32-bit resolution, exactly representable in double */
static inline double philox_u01(uint32_t x) { return (double)x * (1.0 / 4294967296.0); }

#endif
