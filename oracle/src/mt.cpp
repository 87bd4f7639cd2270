// ORACLE (test infrastructure).  MT19937 restated from the published mt19937ar
// algorithm; follows the reference's abb_manipulator/src/mtrand.cpp:12-50 (gen_state, seed,
// seed(array)) and include/mtrand.h:77-104 (tempering, MTRand [0,1)).
// Pinned against the reference's own mtrand.cpp through oracle/_ref (tests/test_oracle.py: test_mt_against_live_reference_build, test_mt_seeds_match_reference_fixture).
#include "../oracle.h"
#include <cmath>
#include <cstdlib>

struct orc_mt {
    uint32_t s[624];
    int p;
};

static void mt_seed(orc_mt* m, uint32_t seed) {           // mtrand.cpp:21-31
    m->s[0] = seed;
    for (int i = 1; i < 624; ++i)
        m->s[i] = 1812433253u * (m->s[i - 1] ^ (m->s[i - 1] >> 30)) + (uint32_t)i;
    m->p = 624;
}

static void mt_seed_array(orc_mt* m, const uint32_t* key, int len) {   // mtrand.cpp:33-50
    mt_seed(m, 19650218u);
    int i = 1, j = 0;
    for (int k = (624 > len ? 624 : len); k; --k) {
        m->s[i] = (m->s[i] ^ ((m->s[i - 1] ^ (m->s[i - 1] >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
        if (++i == 624) { m->s[0] = m->s[623]; i = 1; }
        if (++j == len) j = 0;
    }
    for (int k = 623; k; --k) {
        m->s[i] = (m->s[i] ^ ((m->s[i - 1] ^ (m->s[i - 1] >> 30)) * 1566083941u)) - (uint32_t)i;
        if (++i == 624) { m->s[0] = m->s[623]; i = 1; }
    }
    m->s[0] = 0x80000000u;
    m->p = 624;
}

static inline uint32_t twist(uint32_t u, uint32_t v) {     // mtrand.h:77-82
    return (((u & 0x80000000u) | (v & 0x7FFFFFFFu)) >> 1) ^ ((v & 1u) ? 0x9908B0DFu : 0u);
}

static void mt_gen(orc_mt* m) {                            // mtrand.cpp:12-19
    uint32_t* s = m->s;
    for (int i = 0; i < 624 - 397; ++i) s[i] = s[i + 397] ^ twist(s[i], s[i + 1]);
    for (int i = 624 - 397; i < 623; ++i) s[i] = s[i + 397 - 624] ^ twist(s[i], s[i + 1]);
    s[623] = s[396] ^ twist(s[623], s[0]);
    m->p = 0;
}

extern "C" {

orc_mt* orc_mt_new_seed(uint32_t seed) {
    orc_mt* m = (orc_mt*)malloc(sizeof(orc_mt));
    mt_seed(m, seed);
    return m;
}
orc_mt* orc_mt_new_array(const uint32_t* key, int len) {
    orc_mt* m = (orc_mt*)malloc(sizeof(orc_mt));
    mt_seed_array(m, key, len);
    return m;
}
void orc_mt_free(orc_mt* m) { free(m); }

uint32_t orc_mt_u32(orc_mt* m) {                           // mtrand.h:84-93
    if (m->p == 624) mt_gen(m);
    uint32_t x = m->s[m->p++];
    x ^= (x >> 11);
    x ^= (x << 7) & 0x9D2C5680u;
    x ^= (x << 15) & 0xEFC60000u;
    return x ^ (x >> 18);
}

double orc_mt_real(orc_mt* m) {                            // mtrand.h:103-104
    return (double)orc_mt_u32(m) * (1. / 4294967296.);
}

// std::uniform_real_distribution<double>(a,b) over std::mt19937 as libstdc++ does it
// (MS:1171-1184): generate_canonical<double,53> draws TWO 32-bit words,
// sum = w0 + w1*2^32 accumulated in double, divided by 2^64, clamped below 1.
double orc_mt_std_uniform(orc_mt* m, double a, double b) {
    double sum = 0.0, tmp = 1.0;
    for (int k = 0; k < 2; ++k) {
        sum += (double)orc_mt_u32(m) * tmp;
        tmp *= 4294967296.0;
    }
    double r = sum / tmp;
    if (r >= 1.0) r = std::nextafter(1.0, 0.0);
    return r * (b - a) + a;
}

void orc_mt_fill_u32(orc_mt* m, uint32_t* out, int64_t n) {
    for (int64_t i = 0; i < n; ++i) out[i] = orc_mt_u32(m);
}
void orc_mt_fill_real(orc_mt* m, double* out, int64_t n) {
    for (int64_t i = 0; i < n; ++i) out[i] = orc_mt_real(m);
}

}  // extern "C"
