/* TEST INFRASTRUCTURE.  Exhaustive sweep of the glibc-sincosf restatement used by the rBRIEF kernel
 * (orb_slam2-chinese-annotation_b200/csrc/sincosf_glibc.h) against the host libm's sincosf, which is
 * what the reference's cos(float)/sin(float) (src/ORBextractor.cc:124-125) compile to.
 *
 *   sincosf_sweep <first_bits> <last_bits>
 * prints: "<count> <mismatches> <checksum_libm> <checksum_restatement>"
 * The checksum (sum over inputs of a 64-bit mix of the sin and cos bit patterns) is the same one
 * orb_sincosf_sweep_device() computes on the GPU (include/orb_b200.h). */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../orb_slam2-chinese-annotation_b200/csrc/sincosf_glibc.h"

static inline uint64_t mix(uint32_t xb, uint32_t sb, uint32_t cb) {
    uint64_t v = ((uint64_t)sb << 32) | cb;
    v ^= (uint64_t)xb * 0x9E3779B97F4A7C15ull;
    v *= 0xD6E8FEB86659FD93ull;
    v ^= v >> 32;
    return v;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const uint32_t first = (uint32_t)strtoul(argv[1], 0, 0), last = (uint32_t)strtoul(argv[2], 0, 0);
    uint64_t mism = 0, ck_m = 0, ck_r = 0;
#pragma omp parallel for reduction(+ : mism, ck_m, ck_r) schedule(static)
    for (int64_t i = first; i <= (int64_t)last; ++i) {
        const uint32_t u = (uint32_t)i;
        float x, s0, c0, s1, c1;
        memcpy(&x, &u, 4);
        sincosf(x, &s0, &c0);
        orb_sincosf(x, &s1, &c1);
        uint32_t a, b, c, d;
        memcpy(&a, &s0, 4); memcpy(&b, &c0, 4); memcpy(&c, &s1, 4); memcpy(&d, &c1, 4);
        mism += (a != c) || (b != d);
        ck_m += mix(u, a, b);
        ck_r += mix(u, c, d);
    }
    printf("%llu %llu %llu %llu\n", (unsigned long long)((int64_t)last - first + 1), (unsigned long long)mism,
           (unsigned long long)ck_m, (unsigned long long)ck_r);
    return 0;
}
