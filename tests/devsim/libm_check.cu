// CPU-side check of csrc/rtb_libm.cuh against the host libm (glibc).  Test tooling: host-compiled by nvcc.
// usage: libm_check [stride]   -> prints "<name> tested mismatches" per function; exit 1 on any mismatch.
#include "../../raytracing-hw_b200/csrc/rtb_libm.cuh"
#include <cmath>
#include <cstdio>
#include <cstdlib>

int main(int argc, char **argv) {
    const uint32_t stride = argc > 1 ? (uint32_t) std::atoi(argv[1]) : 97u;
    unsigned long long n = 0, bad = 0;
    // logf over positive normal + subnormal floats up to 4.0 (r2 of the polar method lies in (0,1])
    for (uint32_t u = 1; u < 0x40800000u; u += stride) {
        const float x = rtb::u2f(u);
        const float a = rtb::glibc_logf(x), b = logf(x);
        ++n;
        if (rtb::f2u(a) != rtb::f2u(b)) { if (bad < 5) printf("logf(%a): %a vs %a\n", x, a, b); ++bad; }
    }
    printf("logf %llu %llu\n", n, bad);
    unsigned long long total_bad = bad;
    n = bad = 0;
    const float g = 1.f / 2.2f;
    for (uint32_t u = 0; u <= 0x3f800000u; u += stride) {
        const float x = rtb::u2f(u);
        const float a = rtb::glibc_powf_pos(x, g), b = powf(x, g);
        ++n;
        if (rtb::f2u(a) != rtb::f2u(b)) { if (bad < 5) printf("powf(%a): %a vs %a\n", x, a, b); ++bad; }
    }
    for (uint32_t u = 0; u <= 0x3f800000u; u += stride * 3) {   // the sRGB decode exponent too
        const float x = rtb::u2f(u);
        const float a = rtb::glibc_powf_pos(x, 2.2f), b = powf(x, 2.2f);
        ++n;
        if (rtb::f2u(a) != rtb::f2u(b)) { if (bad < 5) printf("powf2.2(%a): %a vs %a\n", x, a, b); ++bad; }
    }
    printf("powf %llu %llu\n", n, bad);
    total_bad += bad;
    n = bad = 0;
    for (uint32_t u = 0; u <= 0x3f800000u; u += stride) {
        const float x = rtb::u2f(u);
        const float a = rtb::pow5_via_double(x), b = (float) std::pow((double) x, 5);
        ++n;
        if (rtb::f2u(a) != rtb::f2u(b)) { if (bad < 5) printf("pow5(%a): %a vs %a\n", x, a, b); ++bad; }
    }
    printf("pow5 %llu %llu\n", n, bad);
    total_bad += bad;
    return total_bad ? 1 : 0;
}
