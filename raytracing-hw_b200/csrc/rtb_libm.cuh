// glibc-compatible logf / powf / pow(x,5) for device code.
//
// The reference's sample stream runs through libstdc++'s normal_distribution, i.e. through glibc's logf
// (random.tcc:1836: mult = sqrt(-2 * log(r2) / r2)), and its resolve pass through powf(x, 1/2.2f)
// (src/core/scene.cpp:61).  CUDA's own logf/powf differ from glibc in the last ulp, which would make every path
// diverge from the reference after its first Gaussian draw.  glibc >= 2.28 implements both with the table-driven
// double-precision algorithms of ARM's optimized-routines (sysdeps/ieee754/flt-32/e_logf.c, e_powf.c, LGPL; the
// algorithm and its public coefficient tables are restated here, no code copied).  x86-64 hosts with FMA select
// the *_fma ifunc variants, i.e. the same source compiled with contraction; the contractions are written out
// explicitly with fma() below.  tests/test_libm.py checks these restatements against the host libm over
// millions of arguments (0 mismatches expected); the one-in-2^29 double-rounding differences a different
// contraction choice could cause would only perturb single samples.
#pragma once
#include "rtb_device.cuh"

namespace rtb {

RTB_HD double dfma(double a, double b, double c) {
#ifdef __CUDA_ARCH__
    return __fma_rn(a, b, c);
#else
    return std::fma(a, b, c);
#endif
}
RTB_HD uint64_t d2u(double d) {
#ifdef __CUDA_ARCH__
    return (uint64_t) __double_as_longlong(d);
#else
    uint64_t u;
    std::memcpy(&u, &d, 8);
    return u;
#endif
}
RTB_HD double u2d(uint64_t u) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long) u);
#else
    double d;
    std::memcpy(&d, &u, 8);
    return d;
#endif
}

// tables: one copy in constant memory for the kernels, one for host-compiled debugging
#ifdef __CUDA_ARCH__
#define RTB_TBL(name) name##_dev
#else
#define RTB_TBL(name) name##_host
#endif
#define RTB_DEFINE_TABLE(type, name, ...)                    \
    static __constant__ type name##_dev[] = __VA_ARGS__;     \
    static const type name##_host[] = __VA_ARGS__;

// __logf_data: {invc, logc} x 16, c near the centre of each sub-interval of [0x1.66p-1, 0x1.66p0)
RTB_DEFINE_TABLE(double, kLogfTab, {
    0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2,
    0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2, 0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3,
    0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3,
    0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4,
    0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5, 0x1p+0, 0x0p+0,
    0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5, 0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4,
    0x1.b2036576afce6p-1, 0x1.526e57720db08p-3, 0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3,
    0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2, 0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2})

// glibc logf (e_logf.c), FMA-contracted variant
RTB_HD float glibc_logf(float x) {
    uint32_t ix = f2u(x);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return NAN;
        ix = f2u(x * 0x1p23f);   // subnormal: normalise
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int) ((tmp >> 19) & 15u);
    const int k = (int32_t) tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = RTB_TBL(kLogfTab)[2 * i], logc = RTB_TBL(kLogfTab)[2 * i + 1];
    const double z = (double) u2f(iz);
    const double r = dfma(z, invc, -1.0);
    const double y0 = dfma((double) k, 0x1.62e42fefa39efp-1, logc);
    const double r2 = r * r;
    double y = dfma(0x1.5575b0be00b6ap-2, r, -0x1.ffffef20a4123p-2);
    y = dfma(-0x1.00ea348b88334p-2, r2, y);
    y = dfma(y, r2, y0 + r);
    return (float) y;
}

// __powf_log2_data (POWF_SCALE = 1 on x86-64: TOINT_INTRINSICS is 0): {invc, log2(c)} x 16, then the polynomial
RTB_DEFINE_TABLE(double, kPowfLog2Tab, {
    0x1.661ec79f8f3bep+0, -0x1.efec65b963019p-2, 0x1.571ed4aaf883dp+0, -0x1.b0b6832d4fca4p-2,
    0x1.49539f0f010bp+0, -0x1.7418b0a1fb77bp-2, 0x1.3c995b0b80385p+0, -0x1.39de91a6dcf7bp-2,
    0x1.30d190c8864a5p+0, -0x1.01d9bf3f2b631p-2, 0x1.25e227b0b8eap+0, -0x1.97c1d1b3b7afp-3,
    0x1.1bb4a4a1a343fp+0, -0x1.2f9e393af3c9fp-3, 0x1.12358f08ae5bap+0, -0x1.960cbbf788d5cp-4,
    0x1.0953f419900a7p+0, -0x1.a6f9db6475fcep-5, 0x1p+0, 0x0p+0,
    0x1.e608cfd9a47acp-1, 0x1.338ca9f24f53dp-4, 0x1.ca4b31f026aap-1, 0x1.476a9543891bap-3,
    0x1.b2036576afce6p-1, 0x1.e840b4ac4e4d2p-3, 0x1.9c2d163a1aa2dp-1, 0x1.40645f0c6651cp-2,
    0x1.886e6037841edp-1, 0x1.88e9c2c1b9ff8p-2, 0x1.767dcf5534862p-1, 0x1.ce0a44eb17bccp-2})

// __exp2f_data.tab: 2^(i/32) with the exponent bits pre-subtracted
RTB_DEFINE_TABLE(uint64_t, kExp2fTab, {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull,
    0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull,
    0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull,
    0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull,
    0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull,
    0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull,
    0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull})

// glibc powf (e_powf.c) for x >= 0 (or NaN) and a finite, positive, non-integer y -- the only way the hot path
// calls it (resolve: pow(saturate(aces(c)), 1/2.2f), src/core/scene.cpp:60-61; vector.h:361-363).
RTB_HD float glibc_powf_pos(float x, float y) {
    uint32_t ix = f2u(x);
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return 0.f;            // +-0 ^ y = +0
        if (ix * 2 > 0xff000000u) return x + y;  // NaN
        if (ix == 0x7f800000u) return x;         // +inf
        if (ix & 0x80000000u) return NAN;        // negative base, non-integer exponent
        ix = f2u(x * 0x1p23f);                   // subnormal
        ix &= 0x7fffffffu;
        ix -= 23u << 23;
    }
    // log2_inline
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int) ((tmp >> 19) & 15u);
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz = ix - top;
    const int k = (int32_t) top >> 23;
    const double invc = RTB_TBL(kPowfLog2Tab)[2 * i], logc = RTB_TBL(kPowfLog2Tab)[2 * i + 1];
    const double z = (double) u2f(iz);
    const double r = dfma(z, invc, -1.0);
    const double y0 = logc + (double) k;
    const double r2 = r * r;
    double yy = dfma(0x1.27616c9496e0bp-2, r, -0x1.71969a075c67ap-2);
    const double p = dfma(0x1.ec70a6ca7baddp-2, r, -0x1.7154748bef6c8p-1);
    const double r4 = r2 * r2;
    double q = dfma(0x1.71547652ab82bp+0, r, y0);
    q = dfma(p, r2, q);
    yy = dfma(yy, r4, q);
    const double ylogx = (double) y * yy;
    if (((d2u(ylogx) >> 47) & 0xffffu) >= (d2u(126.0) >> 47)) {
        if (ylogx > 0x1.fffffffd1d571p+6) return INFINITY;
        if (ylogx <= -150.0) return 0.f;
    }
    // exp2_inline, N = 32
    const double shift = 0x1.8p+47;   // 0x1.8p+52 / N
    double kd = ylogx + shift;
    const uint64_t ki = d2u(kd);
    kd -= shift;
    const double rr = ylogx - kd;
    uint64_t t = RTB_TBL(kExp2fTab)[ki & 31u];
    t += ki << (52 - 5);
    const double s = u2d(t);
    const double zz = dfma(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3);
    const double rr2 = rr * rr;
    double e = dfma(0x1.62e42ff0c52d6p-1, rr, 1.0);
    e = dfma(zz, rr2, e);
    e = e * s;
    return (float) e;
}

// (float) std::pow(double(x), 5)  (vector.h:464,470): glibc's pow is correctly rounded for all practical purposes,
// so x^5 is formed in double-double and rounded once to double, then to float.
RTB_HD float pow5_via_double(float x) {
    const double xd = (double) x;
    const double x2 = xd * xd;                 // exact: 48 significant bits
    const double hi = x2 * x2;
    const double lo = dfma(x2, x2, -hi);       // exact error of hi
    const double p = hi * xd;
    const double e = dfma(hi, xd, -p);
    return (float) (p + (e + lo * xd));
}

}  // namespace rtb
