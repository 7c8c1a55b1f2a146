/* oracle/stream.h -- TEST INFRASTRUCTURE (shared by the reference shim and the C restatement).
 *
 * One deterministic stream of uniforms in [0,1), 53 bits each (u = m * 2^-53), replacing the
 * reference's clock-seeded engines (main.cpp:323-326, BRDF.cpp:38-39, Mylight.cpp:107-108,
 * 268-269, 432-433).  Two modes:
 *   list mode : the next value of a caller-supplied array (function-level parity tests,
 *               "fixed uniform inputs");
 *   hash mode : u_k = mix64(key + k*GOLDEN) >> 11, key = f(seed, pixel, sample) -- a
 *               counter-based stream restarted at every (pixel, sample) so that images are
 *               reproducible and independent of how rows are split over processes.
 * Draw ORDER is whatever the consuming code does; the restatement consumes in the same order
 * as the reference (SURVEY.md 8a "Uniform-consumption order"), which is what makes
 * restatement-vs-reference images comparable value by value.
 */
#ifndef MCPT_ORACLE_STREAM_H
#define MCPT_ORACLE_STREAM_H
#include <stddef.h>
#include <stdint.h>

typedef struct oracle_stream {
    uint64_t key;
    uint64_t ctr;
    const double* list;
    size_t list_n;
    size_t list_i;
    uint64_t draws;
} oracle_stream;

static inline uint64_t oracle_mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static inline void oracle_stream_seed(oracle_stream* s, uint64_t seed, uint64_t pixel, uint64_t sample)
{
    s->key = oracle_mix64(oracle_mix64(oracle_mix64(seed) + pixel) + sample);
    s->ctr = 0;
    s->list = 0;
    s->list_n = s->list_i = 0;
}

static inline void oracle_stream_list(oracle_stream* s, const double* list, size_t n)
{
    s->list = list;
    s->list_n = n;
    s->list_i = 0;
}

/* 53-bit integer m; u = m * 2^-53 */
static inline uint64_t oracle_stream_next_m53(oracle_stream* s)
{
    s->draws++;
    if (s->list) {
        double u = (s->list_i < s->list_n) ? s->list[s->list_i] : 0.5;
        s->list_i++;
        return (uint64_t)(u * 9007199254740992.0);
    }
    return oracle_mix64(s->key + (s->ctr++) * 0x9E3779B97F4A7C15ull) >> 11;
}

static inline double oracle_stream_next(oracle_stream* s)
{
    return (double)oracle_stream_next_m53(s) * (1.0 / 9007199254740992.0);
}
#endif
