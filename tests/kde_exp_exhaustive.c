/* tests/kde_exp_exhaustive.c — every float z in [0, 15]: kde_term(z) (charon_b200/csrc/kde_fast_exp.inc, the code the CUDA
 * kernel compiles) against the reference expression float(exp(-z*z/2) / 2.5066282746310002) evaluated with this machine's
 * libm.  Prints "mismatches N"; the pytest wrapper (tests/test_kde_exp.py) expects 0.  ~20 s of CPU time over all cores. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../charon_b200/csrc/kde_exp_table.inc"
static const double T64[64] = CG_EXP_T64_INIT;
static inline uint64_t bits_of(double d) { uint64_t b; memcpy(&b, &d, 8); return b; }
static inline double from_bits(uint64_t b) { double d; memcpy(&d, &b, 8); return d; }
static inline double mul_rn(double a, double b) { volatile double x = a * b; return x; }
static inline double add_rn(double a, double b) { volatile double x = a + b; return x; }
#define CG_KDE_FN static inline
#define CG_KDE_TPARAM
#define CG_KDE_TABLE(j) T64[j]
#define CG_KDE_MUL(a, b) mul_rn(a, b)
#define CG_KDE_ADD(a, b) add_rn(a, b)
#define CG_KDE_BITS(d) bits_of(d)
#define CG_KDE_FROM_BITS(b) from_bits(b)
#include "../charon_b200/csrc/kde_fast_exp.inc"
int main(void) {
    const double inv_sqrt = 2.5066282746310002;
    const float top = 15.0f;
    uint32_t ntop; memcpy(&ntop, &top, 4);
    unsigned long long mism = 0;
#pragma omp parallel for schedule(dynamic, 1 << 20) reduction(+ : mism)
    for (uint32_t u = 0; u <= ntop; ++u) {
        float z; memcpy(&z, &u, 4);
        const double zz = (double)z * (double)z;
        const float ref = (float)(exp(-zz / 2.0) / inv_sqrt);
        const float neg = kde_term(-z), got = kde_term(z);
        if (memcmp(&got, &ref, 4) != 0 || memcmp(&neg, &ref, 4) != 0) ++mism;
    }
    printf("tested %u mismatches %llu\n", ntop + 1, mism);
    return mism != 0;
}
