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
    int keyed;   /* list entries are POSITIONS of a keyed table (oracle_keyed_draw4): a consumer that skips a draw
                    (std::discrete_distribution below two weights) still steps over its position */
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
    s->keyed = 0;
}

/* ---- keyed mode: the product renderer's own uniforms ------------------------------------------------
 * Philox4x32-10 (Salmon et al., SC'11; public algorithm, constants from the paper) with counter
 * (pixel, global sample, path-tree node lo, node hi * 4 + call) and key = seed: 4 uniforms of 24 bits each,
 * u = (x >> 8) * 2^-24 -- the table libmcpt's renderer draws from (csrc/rng.cuh restated in C).  With
 * ORACLE_RNG_KEYED the restatement feeds every sampling call from this table instead of the sequential
 * stream, call 0 = (roulette, u_lobe, xi1, xi2), call 1 = (u_tri, xi1, xi2) / (u_class, u_tri, xi1, xi2),
 * so that a PARITY render of libmcpt and the oracle walk the SAME paths and their films can be compared
 * pixel by pixel. */
static inline void oracle_keyed_draw4(uint64_t seed, uint32_t pixel, uint32_t sample, uint64_t node, uint32_t call, double out[4])
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    uint32_t c0 = pixel, c1 = sample, c2 = (uint32_t)node, c3 = (uint32_t)(node >> 32) * 4u + call;
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = (double)(c0 >> 8) * (1.0 / 16777216.0); out[1] = (double)(c1 >> 8) * (1.0 / 16777216.0);
    out[2] = (double)(c2 >> 8) * (1.0 / 16777216.0); out[3] = (double)(c3 >> 8) * (1.0 / 16777216.0);
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
