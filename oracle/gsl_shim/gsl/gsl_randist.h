/* oracle/gsl_shim/gsl/gsl_randist.h -- TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Stand-in for the GSL random-variate functions the reference calls:
 *   gsl_ran_poisson, gsl_ran_gamma, gsl_ran_binomial, gsl_ran_weibull, gsl_ran_negative_binomial,
 *   gsl_ran_geometric, gsl_ran_gaussian, gsl_ran_flat, gsl_ran_choose, gsl_ran_shuffle,
 *   gsl_ran_multinomial.
 * Each is written from the textbook algorithm named beside it; distributional correctness is
 * checked against scipy.stats in tests/test_gsl_shim.py. In TAPE MODE (see gsl_rng.h) every
 * variate is one table lookup on one 32-bit draw; the tables are the same integer-support CDF
 * tables that the product uploads to the GPU (registered with gsl_shim_register_table).
 */
#ifndef GSL_SHIM_RANDIST_H
#define GSL_SHIM_RANDIST_H
#include <cmath>
#include <vector>
#include <cstddef>
#include "gsl_rng.h"

enum gsl_shim_table_kind {
    GSL_SHIM_T_GAMMA_FLOOR = 0, /* p1=a, p2=b : cdf[k] = P(floor(G) <= k)                       */
    GSL_SHIM_T_BINOMIAL,        /* p1=p, p2=n : cdf[k] = P(X <= k)                              */
    GSL_SHIM_T_NEGBIN,          /* p1=p, p2=n                                                   */
    GSL_SHIM_T_GEOMETRIC,       /* p1=p       : cdf[k] = P(X <= k+1), support starts at 1       */
    GSL_SHIM_T_POISSON,         /* p1=mu                                                        */
    GSL_SHIM_T_WEIBULL_GT       /* p1=a, p2=b : cdf[0] = P(X > threshold the caller compares to) */
};
struct gsl_shim_table { int kind; double p1, p2; std::vector<double> cdf; };
inline std::vector<gsl_shim_table> gsl_shim_tables;

static inline void gsl_shim_register_table(int kind, double p1, double p2, const double* cdf, int n) {
    gsl_shim_table t; t.kind = kind; t.p1 = p1; t.p2 = p2; t.cdf.assign(cdf, cdf + n);
    gsl_shim_tables.push_back(t);
}
static inline void gsl_shim_clear_tables() { gsl_shim_tables.clear(); }

static inline bool gsl_shim_close(double x, double y) { return std::fabs(x - y) <= 1e-9 * (std::fabs(x) + std::fabs(y) + 1e-300); }
static inline const gsl_shim_table& gsl_shim_find_table(int kind, double p1, double p2) {
    for (const auto& t : gsl_shim_tables) {
        if (t.kind == kind && gsl_shim_close(t.p1, p1) && gsl_shim_close(t.p2, p2)) return t;
    }
    std::fprintf(stderr, "gsl_shim: tape mode needs a registered table kind=%d p1=%.17g p2=%.17g\n", kind, p1, p2);
    std::abort();
}
/* first k with u < cdf[k]; the last entry is treated as 1.0 */
static inline unsigned gsl_shim_table_draw(const gsl_shim_table& t, const gsl_rng* r) {
    const double u = gsl_rng_uniform(r);
    unsigned k = 0; const unsigned last = (unsigned)t.cdf.size() - 1;
    while (k < last && !(u < t.cdf[k])) ++k;
    return k;
}

/* ---- continuous helpers (normal mode) ---- */
static inline double gsl_ran_gaussian(const gsl_rng* r, double sigma) {
    /* Marsaglia polar method */
    double x, y, r2;
    do {
        x = -1.0 + 2.0 * gsl_rng_uniform_pos(r);
        y = -1.0 + 2.0 * gsl_rng_uniform_pos(r);
        r2 = x * x + y * y;
    } while (r2 > 1.0 || r2 == 0.0);
    return sigma * y * std::sqrt(-2.0 * std::log(r2) / r2);
}

static inline double gsl_ran_gamma(const gsl_rng* r, double a, double b) {
    if (r->tape) { return (double)gsl_shim_table_draw(gsl_shim_find_table(GSL_SHIM_T_GAMMA_FLOOR, a, b), r) + 0.5; }
    /* Marsaglia & Tsang (2000), with the a<1 boost G(a) = G(a+1) * U^(1/a) */
    if (a < 1.0) {
        const double u = gsl_rng_uniform_pos(r);
        return gsl_ran_gamma(r, 1.0 + a, b) * std::pow(u, 1.0 / a);
    }
    const double d = a - 1.0 / 3.0, c = (1.0 / 3.0) / std::sqrt(d);
    double x, v, u;
    for (;;) {
        do { x = gsl_ran_gaussian(r, 1.0); v = 1.0 + c * x; } while (v <= 0.0);
        v = v * v * v;
        u = gsl_rng_uniform_pos(r);
        if (u < 1.0 - 0.0331 * x * x * x * x) break;
        if (std::log(u) < 0.5 * x * x + d * (1.0 - v + std::log(v))) break;
    }
    return b * d * v;
}

static inline double gsl_shim_ran_beta(const gsl_rng* r, double a, double b) {
    const double x1 = gsl_ran_gamma(r, a, 1.0), x2 = gsl_ran_gamma(r, b, 1.0);
    return x1 / (x1 + x2);
}

static inline unsigned int gsl_ran_binomial(const gsl_rng* r, double p, unsigned int n) {
    if (r->tape) { return gsl_shim_table_draw(gsl_shim_find_table(GSL_SHIM_T_BINOMIAL, p, (double)n), r); }
    /* Knuth TAOCP 3.4.1 recursive beta-splitting, then direct Bernoulli sum */
    unsigned int k = 0;
    while (n > 10) {
        const unsigned int a = 1 + (n / 2), b = 1 + n - a;
        const double X = gsl_shim_ran_beta(r, (double)a, (double)b);
        if (X >= p) { n = a - 1; p /= X; }
        else        { k += a; n = b - 1; p = (p - X) / (1.0 - X); }
    }
    for (unsigned int i = 0; i < n; ++i) { if (gsl_rng_uniform(r) < p) ++k; }
    return k;
}

static inline unsigned int gsl_ran_poisson(const gsl_rng* r, double mu) {
    if (r->tape) { gsl_shim_script_step(r); return gsl_shim_table_draw(gsl_shim_find_table(GSL_SHIM_T_POISSON, mu, 0.0), r); }
    /* Knuth TAOCP 3.4.1: gamma/binomial reduction for large mu, product-of-uniforms below 10 */
    unsigned int k = 0;
    while (mu > 10.0) {
        const unsigned int m = (unsigned int)(mu * (7.0 / 8.0));
        const double X = gsl_ran_gamma(r, (double)m, 1.0);
        if (X >= mu) { return k + gsl_ran_binomial(r, mu / X, m - 1); }
        k += m; mu -= X;
    }
    const double emu = std::exp(-mu);
    double prod = 1.0;
    do { prod *= gsl_rng_uniform(r); ++k; } while (prod > emu);
    return k - 1;
}

static inline unsigned int gsl_ran_negative_binomial(const gsl_rng* r, double p, double n) {
    if (r->tape) { return gsl_shim_table_draw(gsl_shim_find_table(GSL_SHIM_T_NEGBIN, p, n), r); }
    /* gamma-Poisson mixture */
    const double X = gsl_ran_gamma(r, n, 1.0);
    return gsl_ran_poisson(r, X * (1.0 - p) / p);
}

static inline unsigned int gsl_ran_geometric(const gsl_rng* r, double p) {
    if (r->tape) { return 1 + gsl_shim_table_draw(gsl_shim_find_table(GSL_SHIM_T_GEOMETRIC, p, 0.0), r); }
    if (p == 1.0) return 1;
    const double u = gsl_rng_uniform_pos(r);
    return (unsigned int)(std::log(u) / std::log(1.0 - p) + 1.0);   /* inversion, support {1,2,...} */
}

static inline double gsl_ran_weibull(const gsl_rng* r, double a, double b) {
    if (r->tape) {
        /* the reference only ever compares the deviate with one threshold (Person.cpp:352,355) */
        const gsl_shim_table& t = gsl_shim_find_table(GSL_SHIM_T_WEIBULL_GT, a, b);
        return gsl_rng_uniform(r) < t.cdf[0] ? 1e6 : 0.0;
    }
    const double u = gsl_rng_uniform_pos(r);
    return a * std::pow(-std::log(u), 1.0 / b);                     /* inversion */
}

static inline double gsl_ran_flat(const gsl_rng* r, double a, double b) {
    const double u = gsl_rng_uniform(r);
    return a * (1.0 - u) + b * u;
}

/* Knuth Algorithm S (selection sampling): k of n, order preserved */
static inline int gsl_ran_choose(const gsl_rng* r, void* dest, size_t k, void* src, size_t n, size_t size) {
    if (k > n) return 1;
    size_t j = 0;
    for (size_t i = 0; i < n && j < k; ++i) {
        if ((double)(n - i) * gsl_rng_uniform(r) < (double)(k - j)) {
            std::memcpy((char*)dest + size * j, (char*)src + size * i, size);
            ++j;
        }
    }
    return 0;
}

/* Fisher-Yates */
static inline void gsl_ran_shuffle(const gsl_rng* r, void* base, size_t n, size_t size) {
    char* b = (char*)base; std::vector<char> tmp(size);
    for (size_t i = n; i > 1; --i) {
        const size_t j = gsl_rng_uniform_int(r, i);
        if (j != i - 1) {
            std::memcpy(tmp.data(), b + (i - 1) * size, size);
            std::memcpy(b + (i - 1) * size, b + j * size, size);
            std::memcpy(b + j * size, tmp.data(), size);
        }
    }
}

/* conditional-binomial construction */
static inline void gsl_ran_multinomial(const gsl_rng* r, size_t K, unsigned int N, const double p[], unsigned int n[]) {
    double norm = 0.0, sum_p = 0.0; unsigned int sum_n = 0;
    for (size_t k = 0; k < K; ++k) norm += p[k];
    for (size_t k = 0; k < K; ++k) {
        if (p[k] > 0.0 && sum_n < N && norm - sum_p > 0.0) {
            double q = p[k] / (norm - sum_p); if (q > 1.0) q = 1.0;
            n[k] = gsl_ran_binomial(r, q, N - sum_n);
        } else { n[k] = 0; }
        sum_p += p[k]; sum_n += n[k];
    }
}
#endif
