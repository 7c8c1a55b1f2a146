// Deterministic RNG injected into the UNMODIFIED reference sources.  <random> is included
// first so its own typedef of default_random_engine is parsed untouched; afterwards every
// `std::default_random_engine generator(seed1)` in the reference constructs one of the
// engines below.  libstdc++'s discrete_distribution / uniform_real_distribution each draw ONE
// generate_canonical<double,53>; with a 64-bit engine that is one call: u = x * 2^-64
// (bits/random.tcc generate_canonical), so returning m53 << 11 delivers u = m53 * 2^-53 exactly.
#pragma once
#include <random>
#include <cstdint>
#include "../stream.h"

extern oracle_stream g_oracle_stream;
extern int g_oracle_depth;          // live RR engines == depth of the vertex being shaded
extern int g_oracle_max_depth;      // <=0: unlimited (the reference has no depth cap, F5)
extern uint64_t g_oracle_vertices;  // RR engines constructed == non-emissive vertices shaded

namespace std {
// BRDF.cpp / Mylight.cpp: plain stream reader.
struct oracle_engine {
    typedef uint64_t result_type;
    explicit oracle_engine(unsigned long long = 0) {}
    static constexpr result_type min() { return 0; }
    static constexpr result_type max() { return ~0ull; }
    result_type operator()() { return oracle_stream_next_m53(&g_oracle_stream) << 11; }
};
// main.cpp: the only engines that are *used* there are the Russian-roulette ones
// (main.cpp:323-326,377-380,431-434).  Each lives until its shade*() frame returns, i.e. across
// the recursive call, so the number of live engines is the depth of the current vertex.  That
// is how a max_depth is imposed without touching the reference: at depth >= max_depth the RR
// draw is forced to fail (u -> 1 > P_RR) and consumes nothing from the stream.
struct oracle_rr_engine {
    typedef uint64_t result_type;
    explicit oracle_rr_engine(unsigned long long = 0) { ++g_oracle_depth; ++g_oracle_vertices; }
    oracle_rr_engine(const oracle_rr_engine&) = delete;
    ~oracle_rr_engine() { --g_oracle_depth; }
    static constexpr result_type min() { return 0; }
    static constexpr result_type max() { return ~0ull; }
    result_type operator()()
    {
        if (g_oracle_max_depth > 0 && g_oracle_depth >= g_oracle_max_depth) return max();
        return oracle_stream_next_m53(&g_oracle_stream) << 11;
    }
};
}
#ifdef ORACLE_RR_TU
#define default_random_engine oracle_rr_engine
#else
#define default_random_engine oracle_engine
#endif
