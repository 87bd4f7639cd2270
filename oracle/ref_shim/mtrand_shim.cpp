// ORACLE support (test infrastructure).  C shim over the REFERENCE's own MT19937
// (abb_manipulator/include/mtrand.h + src/mtrand.cpp, compiled from where they lie under
// /root/reference into oracle/_ref/libmtrand_ref.so — no reference source is copied).
// Used only to pin oracle/src/mt.cpp and to regenerate tests/golden/mt19937_*.json.
// The reference class keeps its state in static members (mtrand.h:68-70), so there is
// one global stream per process.
#include "mtrand.h"
#include <stdint.h>

extern "C" {
void ref_mt_seed(uint32_t s) { MTRand_int32 g((unsigned long)s); (void)g; }
void ref_mt_seed_array(const uint32_t* key, int len) {
    unsigned long k[64];
    for (int i = 0; i < len && i < 64; ++i) k[i] = key[i];
    MTRand_int32 g(k, len);
    (void)g;
}
uint32_t ref_mt_u32(void) { MTRand_int32 g; return (uint32_t)g(); }
double ref_mt_real(void) { MTRand g; return g(); }
}
