/* oracle/gsl_shim/gsl/gsl_rng.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Header-only stand-in for the subset of GSL's gsl_rng API that the reference uses
 * (call sites listed in SURVEY.md 8c: simulator.h:53-55, Community.cpp:976, Person.cpp:326...,
 * behavior_tuner.h:61-63,500-502). GSL is not installed in this image and is not vendored by the
 * reference, so this shim exists only so that the UNMODIFIED reference sources compile here.
 *
 * The generators are restated from their published recurrences (L'Ecuyer 1996 "taus88" with an
 * LCG-69069 seeding; Matsumoto & Nishimura 1998 MT19937). Equality with GSL's exact streams is
 * NOT claimed and nothing in the reference pins it.
 *
 * TAPE MODE (parity tooling): when a generator has a tape attached (gsl_shim_tape_attach), every
 * draw is served from a Philox stream keyed by gsl_shim_tape_begin(day, purpose, a, b) instead of
 * the generator's own recurrence, and every distribution in gsl_randist.h becomes a single-draw
 * table lookup. That turns the unmodified reference functions into deterministic functions of
 * (state, key), which the tests compare bit-for-bit with the oracle restatement.
 */
#ifndef GSL_SHIM_RNG_H
#define GSL_SHIM_RNG_H
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cstdio>
#include "../../philox_ref.h"

struct gsl_rng_type { const char* name; int kind; };
inline const gsl_rng_type gsl_shim_taus2_type   = { "taus2", 0 };
inline const gsl_rng_type gsl_shim_mt19937_type = { "mt19937", 1 };
inline const gsl_rng_type* const gsl_rng_taus2   = &gsl_shim_taus2_type;
inline const gsl_rng_type* const gsl_rng_mt19937 = &gsl_shim_mt19937_type;

struct gsl_shim_tape {
    uint64_t seed = 0;
    philox_stream_ref stream;
    uint64_t draws_total = 0;
};

/* VISIT SCRIPT (FoI parity tooling, `ref_driver_probe foi`): Community::public_activity draws one Poisson per patron in
 * id order and then that patron's location indices (Community.cpp:1006-1022). While a script is set, the k-th Poisson call
 * on a taped generator first re-keys the tape to (day, purpose, pids[k], 0): the visit stream of the restatement. */
struct gsl_shim_script { const uint32_t* pids = nullptr; size_t n = 0, next = 0; uint32_t day = 0, purpose = 0; bool overrun = false; };
inline gsl_shim_script gsl_shim_visits;

struct gsl_rng {
    const gsl_rng_type* type;
    uint32_t s1, s2, s3;        /* taus2 */
    uint32_t mt[624]; int mti;  /* mt19937 */
    gsl_shim_tape* tape;        /* non-null => tape mode */
};

/* ---- taus88 ---- */
static inline uint32_t gsl_shim_taus_step(gsl_rng* r) {
    r->s1 = ((r->s1 & 4294967294u) << 12) ^ (((r->s1 << 13) ^ r->s1) >> 19);
    r->s2 = ((r->s2 & 4294967288u) << 4)  ^ (((r->s2 << 2)  ^ r->s2) >> 25);
    r->s3 = ((r->s3 & 4294967280u) << 17) ^ (((r->s3 << 3)  ^ r->s3) >> 11);
    return r->s1 ^ r->s2 ^ r->s3;
}

/* ---- mt19937 ---- */
static inline void gsl_shim_mt_seed(gsl_rng* r, uint32_t s) {
    r->mt[0] = s;
    for (int i = 1; i < 624; ++i) r->mt[i] = 1812433253u * (r->mt[i-1] ^ (r->mt[i-1] >> 30)) + (uint32_t)i;
    r->mti = 624;
}
static inline uint32_t gsl_shim_mt_step(gsl_rng* r) {
    if (r->mti >= 624) {
        for (int k = 0; k < 624; ++k) {
            const uint32_t y = (r->mt[k] & 0x80000000u) | (r->mt[(k + 1) % 624] & 0x7fffffffu);
            r->mt[k] = r->mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        r->mti = 0;
    }
    uint32_t y = r->mt[r->mti++];
    y ^= (y >> 11); y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= (y >> 18);
    return y;
}

static inline void gsl_rng_set(gsl_rng* r, unsigned long int seed) {
    if (r->type->kind == 0) {
        uint32_t s = (uint32_t)seed; if (s == 0) s = 1;
        #define GSL_SHIM_LCG(n) (69069u * (n))
        r->s1 = GSL_SHIM_LCG(s);      if (r->s1 < 2u)  r->s1 += 2u;
        r->s2 = GSL_SHIM_LCG(r->s1);  if (r->s2 < 8u)  r->s2 += 8u;
        r->s3 = GSL_SHIM_LCG(r->s2);  if (r->s3 < 16u) r->s3 += 16u;
        #undef GSL_SHIM_LCG
        for (int i = 0; i < 6; ++i) gsl_shim_taus_step(r);
    } else {
        gsl_shim_mt_seed(r, seed == 0 ? 4357u : (uint32_t)seed);
    }
}

static inline gsl_rng* gsl_rng_alloc(const gsl_rng_type* T) {
    gsl_rng* r = (gsl_rng*)std::calloc(1, sizeof(gsl_rng));
    r->type = T; r->tape = nullptr;
    gsl_rng_set(r, 0);
    return r;
}
static inline void gsl_rng_free(gsl_rng* r) { std::free(r); }
static inline gsl_rng* gsl_rng_clone(const gsl_rng* r) {
    gsl_rng* c = (gsl_rng*)std::malloc(sizeof(gsl_rng)); std::memcpy(c, r, sizeof(gsl_rng)); return c;
}
static inline int gsl_rng_memcpy(gsl_rng* dest, const gsl_rng* src) { std::memcpy(dest, src, sizeof(gsl_rng)); return 0; }
static inline const char* gsl_rng_name(const gsl_rng* r) { return r->type->name; }

/* raw 32-bit draw; both generators span [0, 2^32-1] */
static inline unsigned long int gsl_rng_get(const gsl_rng* cr) {
    gsl_rng* r = const_cast<gsl_rng*>(cr);
    if (r->tape) { r->tape->draws_total++; return philox_stream_next(&r->tape->stream); }
    return r->type->kind == 0 ? gsl_shim_taus_step(r) : gsl_shim_mt_step(r);
}
static inline unsigned long int gsl_rng_max(const gsl_rng*) { return 0xffffffffUL; }
static inline unsigned long int gsl_rng_min(const gsl_rng*) { return 0; }

static inline double gsl_rng_uniform(const gsl_rng* r) { return (double)gsl_rng_get(r) / 4294967296.0; }
static inline double gsl_rng_uniform_pos(const gsl_rng* r) {
    double x; do { x = gsl_rng_uniform(r); } while (x == 0.0); return x;
}
static inline unsigned long int gsl_rng_uniform_int(const gsl_rng* r, unsigned long int n) {
    if (r->tape) { return (unsigned long int)(((uint64_t)gsl_rng_get(r) * (uint64_t)n) >> 32); }
    const unsigned long int scale = 0xffffffffUL / n;
    unsigned long int k;
    do { k = gsl_rng_get(r) / scale; } while (k >= n);
    return k;
}

/* ---- tape control ---- */
static inline void gsl_shim_tape_attach(gsl_rng* r, gsl_shim_tape* t) { r->tape = t; }
static inline void gsl_shim_tape_begin(gsl_shim_tape* t, uint32_t day, uint32_t purpose, uint32_t a, uint32_t b) {
    philox_stream_begin(&t->stream, t->seed, day, purpose, a, b);
}
static inline void gsl_shim_script_step(const gsl_rng* r) {
    gsl_shim_script& s = gsl_shim_visits;
    if (!r->tape || !s.pids) return;
    if (s.next < s.n) gsl_shim_tape_begin(r->tape, s.day, s.purpose, s.pids[s.next++], 0); else s.overrun = true;
}
#endif
