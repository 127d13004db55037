/*
 * TEST INFRASTRUCTURE ONLY (oracle/): restatement of the libm expf/logf the compiled reference calls.
 *
 * The reference tree calls exp()/log() on floats (ctree_muzero/lib/cnode.cpp:129, :673), which bind
 * to expf@GLIBC_2.27 / logf@GLIBC_2.27 (nm -D on the built reference).  libm is a third-party
 * dependency absent from /root/reference: glibc 2.39 (Ubuntu 2.39-0ubuntu8.5), whose float exp/log
 * are the ARM "optimized-routines" algorithms (glibc sysdeps/ieee754/flt-32/e_expf.c, e_logf.c,
 * e_exp2f_data.c, e_logf_data.c): a 32-entry 2^(i/32) table + cubic for expf, a 16-entry
 * (1/c, log c) table + cubic for logf, both evaluated in double and rounded once to float.
 * On x86-64 hosts with FMA, glibc's ifunc selects the -mfma build (__expf_fma/__logf_fma); the
 * fused operations below are the ones that build performs (read from its disassembly), so results
 * are bit-identical to the host libm on such machines.  tests/test_libm_restatement.py pins this
 * against the live libm over hundreds of millions of inputs.
 *
 * Provenance and licences (Arm Optimized Routines / glibc expf, logf: MIT / LGPL-2.1+; fdlibm tanhf, expm1f: Copyright (C)
 * 1993 by Sun Microsystems, Inc., "Permission to use, copy, modify, and distribute this software is freely granted,
 * provided that this notice is preserved."): THIRD_PARTY_NOTICES.md at the repo root.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "oracle.h"

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint64_t d2u(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
static inline double u2d(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }

/* 2^(i/32) bit patterns minus (i << 47): glibc __exp2f_data.tab */
static const uint64_t EXP_T[32] = {
    0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,
    0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,
    0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,
    0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,
    0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,
    0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,
    0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,
    0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL,
};
#define EXP_SHIFT 0x1.8p+52
#define EXP_INVLN2N 0x1.71547652b82fep+5 /* 32/ln2 */
#define EXP_C0 0x1.c6af84b912394p-20
#define EXP_C1 0x1.ebfce50fac4f3p-13
#define EXP_C2 0x1.62e42ff0c52d6p-6

float orc_expf(float x)
{
    uint32_t ux = f2u(x);
    uint32_t abstop = (ux >> 20) & 0x7ff;
    double xd = (double)x;
    if (abstop >= 0x42b) { /* |x| >= 88 or nan/inf */
        if (ux == 0xff800000u) return 0.0f;
        if (abstop >= 0x7f8) return x + x;
        if (x > 0x1.62e42ep6f) return INFINITY;                    /* overflow */
        if (x < -0x1.9fe368p6f) return 0.0f;                       /* underflow */
        if (x < -0x1.9d1d9ep6f) return 0x1p-149f;                  /* may-underflow: 0x1.4p-75f^2 rounds to the least subnormal */
    }
    double kd = fma(EXP_INVLN2N, xd, EXP_SHIFT);
    uint64_t ki = d2u(kd);
    kd -= EXP_SHIFT;
    double r = fma(EXP_INVLN2N, xd, -kd);
    uint64_t t = EXP_T[ki & 31] + (ki << 47);
    double s = u2d(t);
    double z = fma(EXP_C0, r, EXP_C1);
    double r2 = r * r;
    double y = fma(EXP_C2, r, 1.0);
    y = fma(z, r2, y);
    y = y * s;
    return (float)y;
}

/* glibc __logf_data: {1/c, log(c)} for 16 subintervals of [0x1.66p-1, 0x1.66p0) */
static const double LOG_T[16][2] = {
    {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
    {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
    {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
    {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
    {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
    {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
    {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
    {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2},
};
#define LOG_LN2 0x1.62e42fefa39efp-1
#define LOG_A0 -0x1.00ea348b88334p-2
#define LOG_A1 0x1.5575b0be00b6ap-2
#define LOG_A2 -0x1.ffffef20a4123p-2

float orc_logf(float x)
{
    uint32_t ix = f2u(x);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2u == 0) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2u >= 0xff000000u) return NAN;
        ix = f2u(x * 0x1p23f);
        ix -= 23u << 23;
    }
    uint32_t tmp = ix - 0x3f330000u;
    int i = (tmp >> 19) & 15;
    int k = (int32_t)tmp >> 23;
    uint32_t iz = ix - (tmp & 0xff800000u);
    double invc = LOG_T[i][0], logc = LOG_T[i][1];
    double z = (double)u2f(iz);
    double r = fma(z, invc, -1.0);
    double y0 = fma((double)k, LOG_LN2, logc);
    double y = fma(LOG_A1, r, LOG_A2);
    double r2 = r * r;
    double hi = r + y0;
    y = fma(LOG_A0, r2, y);
    y = fma(r2, y, hi);
    return (float)y;
}

/* bulk drivers used by the pinning test */
void orc_expf_array(const float *x, float *out, long n) { for (long i = 0; i < n; ++i) out[i] = orc_expf(x[i]); }
void orc_logf_array(const float *x, float *out, long n) { for (long i = 0; i < n; ++i) out[i] = orc_logf(x[i]); }

/* count bit mismatches against the live libm over the raw float bit range [lo, hi) */
long orc_expf_mismatch_range(uint32_t lo, uint32_t hi)
{
    long bad = 0;
    for (uint64_t u = lo; u < hi; ++u) {
        float x = u2f((uint32_t)u);
        float a = orc_expf(x), b = expf(x);
        if (f2u(a) != f2u(b) && !(a != a && b != b)) ++bad;
    }
    return bad;
}

long orc_logf_mismatch_range(uint32_t lo, uint32_t hi)
{
    long bad = 0;
    for (uint64_t u = lo; u < hi; ++u) {
        float x = u2f((uint32_t)u);
        float a = orc_logf(x), b = logf(x);
        if (f2u(a) != f2u(b) && !(a != a && b != b)) ++bad;
    }
    return bad;
}

/* ---------------------------------------------------------------------------------------------------------------
 * tanhf, as the Sampled-EfficientZero tree calls it on every sampled action (ctree_sampled_efficientzero/lib/
 * cnode.cpp:268: tanh() on a float resolves to the float overload, i.e. tanhf@GLIBC_2.2.5).  glibc 2.39 still ships
 * the Sun fdlibm routines for it (sysdeps/ieee754/flt-32/s_tanhf.c on top of s_expm1f.c; no FMA ifunc variant):
 * plain float arithmetic, restated here operation for operation.  tests/test_sampler_oracle.py pins it against the
 * live libm (orc_tanhf_mismatch_range).  Constants are the published fdlibm ones.
 */
static float orc_expm1f(float x)
{
    static const float one = 1.0f, huge = 1.0e+30f, tiny = 1.0e-30f, o_threshold = 8.8721679688e+01f,
                       ln2_hi = 6.9313812256e-01f, ln2_lo = 9.0580006145e-06f, invln2 = 1.4426950216e+00f,
                       Q1 = -3.3333335072e-02f, Q2 = 1.5873016091e-03f, Q3 = -7.9365076090e-05f,
                       Q4 = 4.0082177293e-06f, Q5 = -2.0109921195e-07f;
    float y, hi, lo, c = 0.0f, t, e, hxs, hfx, r1;
    int32_t k;
    uint32_t hx = f2u(x);
    const uint32_t xsb = hx & 0x80000000u;
    hx &= 0x7fffffffu;
    if (hx >= 0x4195b844u) { /* |x| >= 27 ln2 */
        if (hx >= 0x42b17218u) {
            if (hx > 0x7f800000u) return x + x;
            if (hx == 0x7f800000u) return xsb == 0 ? x : -1.0f;
            if (x > o_threshold) return huge * huge;
        }
        if (xsb != 0) return tiny - one;
    }
    if (hx > 0x3eb17218u) { /* |x| > 0.5 ln2 */
        if (hx < 0x3F851592u) { /* |x| < 1.5 ln2 */
            if (xsb == 0) { hi = x - ln2_hi; lo = ln2_lo; k = 1; }
            else { hi = x + ln2_hi; lo = -ln2_lo; k = -1; }
        } else {
            k = (int32_t)(invln2 * x + (xsb == 0 ? 0.5f : -0.5f));
            t = (float)k;
            hi = x - t * ln2_hi;
            lo = t * ln2_lo;
        }
        x = hi - lo;
        c = (hi - x) - lo;
    } else if (hx < 0x33000000u) { /* |x| < 2^-25 */
        t = huge + x;
        return x - (t - (huge + x));
    } else k = 0;
    hfx = 0.5f * x;
    hxs = x * hfx;
    r1 = one + hxs * (Q1 + hxs * (Q2 + hxs * (Q3 + hxs * (Q4 + hxs * Q5))));
    t = 3.0f - r1 * hfx;
    e = hxs * ((r1 - t) / (6.0f - x * t));
    if (k == 0) return x - (x * e - hxs);
    e = (x * (e - c) - c);
    e -= hxs;
    if (k == -1) return 0.5f * (x - e) - 0.5f;
    if (k == 1) return x < -0.25f ? -2.0f * (e - (x + 0.5f)) : one + 2.0f * (x - e);
    if (k <= -2 || k > 56) {
        y = one - (e - x);
        y = u2f(f2u(y) + ((uint32_t)k << 23));
        return y - one;
    }
    if (k < 23) {
        t = u2f(0x3f800000u - (0x1000000u >> k));
        y = t - (e - x);
        y = u2f(f2u(y) + ((uint32_t)k << 23));
    } else {
        t = u2f((uint32_t)(0x7f - k) << 23);
        y = x - (e + t);
        y += one;
        y = u2f(f2u(y) + ((uint32_t)k << 23));
    }
    return y;
}

float orc_tanhf(float x)
{
    const float one = 1.0f, two = 2.0f, tiny = 1.0e-30f;
    float t, z;
    const int32_t jx = (int32_t)f2u(x), ix = jx & 0x7fffffff;
    if (ix >= 0x7f800000) return jx >= 0 ? one / x + one : one / x - one;
    if (ix < 0x41b00000) { /* |x| < 22 */
        if (ix == 0) return x;
        if (ix < 0x24000000) return x * (one + x);
        if (ix >= 0x3f800000) {
            t = orc_expm1f(two * fabsf(x));
            z = one - two / (t + two);
        } else {
            t = orc_expm1f(-two * fabsf(x));
            z = -t / (t + two);
        }
    } else z = one - tiny;
    return jx >= 0 ? z : -z;
}

void orc_tanhf_array(const float *x, float *out, long n) { for (long i = 0; i < n; ++i) out[i] = orc_tanhf(x[i]); }

/* bit mismatches against the live libm over the raw float bit range [lo, hi), every `stride`-th pattern */
long orc_tanhf_mismatch_range(uint32_t lo, uint32_t hi, uint32_t stride)
{
    long bad = 0;
    for (uint64_t u = lo; u < hi; u += stride ? stride : 1) {
        float x = u2f((uint32_t)u);
        float a = orc_tanhf(x), b = tanhf(x);
        if (f2u(a) != f2u(b) && !(a != a && b != b)) ++bad;
    }
    return bad;
}
