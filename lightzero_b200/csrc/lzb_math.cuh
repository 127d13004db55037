// Device math that has to agree bit-for-bit with what the reference tree computes on the host.
//
// The reference (lzero/mcts/ctree/ctree_muzero/lib/cnode.cpp:129,673 and siblings) calls exp()/log()
// on floats, i.e. glibc's expf/logf.  glibc >= 2.27 implements them with the ARM optimized-routines
// algorithms (sysdeps/ieee754/flt-32/e_expf.c, e_logf.c): table + cubic in double, one final rounding
// to float.  CUDA's expf/logf differ in the last ulp, which would change softmax priors and, through
// them, pUCT arg-maxes.  So the same algorithm is evaluated here in fp64 with the fused operations
// the x86-64 FMA build of glibc performs.  This translation unit is compiled with -fmad=false: the
// only fused multiply-adds are the explicit fma() calls below.
// Provenance and licences of the restated routines (Arm Optimized Routines / glibc expf, logf: MIT / LGPL-2.1+;
// fdlibm tanhf, expm1f: Copyright (C) 1993 by Sun Microsystems, Inc., "Permission to use, copy, modify, and distribute
// this software is freely granted, provided that this notice is preserved."): THIRD_PARTY_NOTICES.md at the repo root.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lzb {

// The two lookup tables are indexed by a DIFFERENT entry in every lane.  In __constant__ memory such a load (LDC with a
// register index) is replayed once per distinct address in the warp - up to 18 replays per expf of a softmax over 18
// logits; as ordinary global memory read through the read-only path (LDG.CONSTANT) the 256 bytes are two cache lines
// and a warp's lookups are one or two L1 hits.  -DLZB_TABLES_IN_CONSTANT restores the round-1 placement for A/B runs.
#ifdef LZB_TABLES_IN_CONSTANT
#define LZB_TABLE __device__ __constant__
#define LZB_TABLE_AT(tab, i) (tab[i])
#else
#define LZB_TABLE __device__
#define LZB_TABLE_AT(tab, i) __ldg(&tab[i])
#endif

LZB_TABLE uint64_t kExpTab[32] = {
    0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,
    0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,
    0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,
    0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,
    0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,
    0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,
    0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,
    0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL,
};

// {1/c, log c} pairs of glibc's __logf_data
LZB_TABLE double kLogTab[32] = {
    0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2,
    0x1.49539f0f010bp+0,  -0x1.01eae7f513a67p-2, 0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3,
    0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0,  -0x1.1aa2bc79c81p-3,
    0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4,
    0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5, 0x1p+0,               0x0p+0,
    0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5,  0x1.ca4b31f026aap-1,  0x1.c5e53aa362eb4p-4,
    0x1.b2036576afce6p-1, 0x1.526e57720db08p-3,  0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3,
    0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2,  0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2,
};

// Tables in shared memory (round 2; -DLZB_X_NO_SHARED_TABLES = before: mz 17.32 -> 17.23 us per step, ez 29.45 -> 29.03,
// gmz 22.96 -> 22.40): the lane-group kernels copy the two libm tables (2 x 256 B) into dynamic shared memory
// before they wait on the previous launch, and their expf / logf read them with LDS - the per-lane-indexed lookup is a
// dependent load in the middle of every expf, and as a global load it depends on L1 residency next to the row traffic
extern __shared__ unsigned long long lzb_tab_smem[]; // [0..32): kExpTab, [32..64): kLogTab bit patterns

// expf as glibc 2.39 x86-64 (FMA ifunc variant) computes it.
template <bool SH>
__device__ __forceinline__ float glibc_expf_impl(float x)
{
    const uint32_t ux = __float_as_uint(x);
    const uint32_t abstop = (ux >> 20) & 0x7ffu;
    if (abstop >= 0x42bu) { // |x| >= 88, inf or nan
        if (ux == 0xff800000u) return 0.0f;
        if (abstop >= 0x7f8u) return x + x;
        if (x > 0x1.62e42ep6f) return __int_as_float(0x7f800000);
        if (x < -0x1.9fe368p6f) return 0.0f;
        if (x < -0x1.9d1d9ep6f) return __uint_as_float(1u); // least subnormal
    }
    const double xd = (double)x;
    double kd = fma(0x1.71547652b82fep+5, xd, 0x1.8p+52);
    const uint64_t ki = (uint64_t)__double_as_longlong(kd);
    kd = __dsub_rn(kd, 0x1.8p+52);
    const double r = fma(0x1.71547652b82fep+5, xd, -kd);
    const uint64_t tab = SH ? (uint64_t)lzb_tab_smem[ki & 31u] : LZB_TABLE_AT(kExpTab, ki & 31u);
    const double s = __longlong_as_double((long long)(tab + (ki << 47)));
    const double z = fma(0x1.c6af84b912394p-20, r, 0x1.ebfce50fac4f3p-13);
    const double r2 = __dmul_rn(r, r);
    double y = fma(0x1.62e42ff0c52d6p-6, r, 1.0);
    y = fma(z, r2, y);
    y = __dmul_rn(y, s);
    return __double2float_rn(y);
}
__device__ __forceinline__ float glibc_expf(float x) { return glibc_expf_impl<false>(x); }
__device__ __forceinline__ float glibc_expf_s(float x) // lane-group kernels: table in shared memory when the experiment is on
{
#ifndef LZB_X_NO_SHARED_TABLES
    return glibc_expf_impl<true>(x);
#else
    return glibc_expf_impl<false>(x);
#endif
}

// logf as glibc 2.39 x86-64 (FMA ifunc variant) computes it.
template <bool SH>
__device__ __forceinline__ float glibc_logf_impl(float x)
{
    uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2u == 0u) return __int_as_float(0xff800000);
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2u >= 0xff000000u) return __int_as_float(0x7fc00000);
        ix = __float_as_uint(__fmul_rn(x, 0x1p23f));
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = SH ? __longlong_as_double((long long)lzb_tab_smem[32 + 2 * i]) : LZB_TABLE_AT(kLogTab, 2 * i);
    const double logc = SH ? __longlong_as_double((long long)lzb_tab_smem[32 + 2 * i + 1]) : LZB_TABLE_AT(kLogTab, 2 * i + 1);
    const double z = (double)__uint_as_float(iz);
    const double r = fma(z, invc, -1.0);
    const double y0 = fma((double)k, 0x1.62e42fefa39efp-1, logc);
    double y = fma(0x1.5575b0be00b6ap-2, r, -0x1.ffffef20a4123p-2);
    const double r2 = __dmul_rn(r, r);
    const double hi = __dadd_rn(r, y0);
    y = fma(-0x1.00ea348b88334p-2, r2, y);
    y = fma(r2, y, hi);
    return __double2float_rn(y);
}
__device__ __forceinline__ float glibc_logf(float x) { return glibc_logf_impl<false>(x); }
__device__ __forceinline__ float glibc_logf_s(float x)
{
#ifndef LZB_X_NO_SHARED_TABLES
    return glibc_logf_impl<true>(x);
#else
    return glibc_logf_impl<false>(x);
#endif
}

// the lane-group kernels call this first (before griddepcontrol.wait: the tables are constants)
__device__ __forceinline__ void fill_shared_tables()
{
#ifndef LZB_X_NO_SHARED_TABLES
    if (threadIdx.x < 32) lzb_tab_smem[threadIdx.x] = kExpTab[threadIdx.x];
    else if (threadIdx.x < 64) lzb_tab_smem[threadIdx.x] = (unsigned long long)__double_as_longlong(kLogTab[threadIdx.x - 32]);
    __syncthreads();
#endif
}

// tanhf as glibc 2.39 computes it (sysdeps/ieee754/flt-32/s_tanhf.c over s_expm1f.c: the Sun fdlibm routines, plain
// float arithmetic, no FMA variant): what the Sampled-EfficientZero tree applies to every sampled action (sez
// cnode.cpp:268).  Same operations in the same order (this TU never contracts a * b + c); pinned against the host
// restatement oracle/glibc_flt32.c:orc_tanhf, itself pinned against the live libm.  Used by the parity sampler only.
__device__ __forceinline__ float glibc_expm1f(float x)
{
    const float one = 1.0f, huge = 1.0e+30f, tiny = 1.0e-30f, o_threshold = 8.8721679688e+01f, ln2_hi = 6.9313812256e-01f,
                ln2_lo = 9.0580006145e-06f, invln2 = 1.4426950216e+00f, Q1 = -3.3333335072e-02f, Q2 = 1.5873016091e-03f,
                Q3 = -7.9365076090e-05f, Q4 = 4.0082177293e-06f, Q5 = -2.0109921195e-07f;
    float y, hi, lo, c = 0.0f, t, e, hxs, hfx, r1;
    int k;
    uint32_t hx = __float_as_uint(x);
    const uint32_t xsb = hx & 0x80000000u;
    hx &= 0x7fffffffu;
    if (hx >= 0x4195b844u) {
        if (hx >= 0x42b17218u) {
            if (hx > 0x7f800000u) return x + x;
            if (hx == 0x7f800000u) return xsb == 0 ? x : -1.0f;
            if (x > o_threshold) return huge * huge;
        }
        if (xsb != 0) return tiny - one;
    }
    if (hx > 0x3eb17218u) {
        if (hx < 0x3F851592u) {
            if (xsb == 0) { hi = x - ln2_hi; lo = ln2_lo; k = 1; }
            else { hi = x + ln2_hi; lo = -ln2_lo; k = -1; }
        } else {
            k = __float2int_rz(invln2 * x + (xsb == 0 ? 0.5f : -0.5f));
            t = (float)k;
            hi = x - t * ln2_hi;
            lo = t * ln2_lo;
        }
        x = hi - lo;
        c = (hi - x) - lo;
    } else if (hx < 0x33000000u) {
        t = huge + x;
        return x - (t - (huge + x));
    } else k = 0;
    hfx = 0.5f * x;
    hxs = x * hfx;
    r1 = one + hxs * (Q1 + hxs * (Q2 + hxs * (Q3 + hxs * (Q4 + hxs * Q5))));
    t = 3.0f - r1 * hfx;
    e = hxs * __fdiv_rn(r1 - t, 6.0f - x * t);
    if (k == 0) return x - (x * e - hxs);
    e = (x * (e - c) - c);
    e -= hxs;
    if (k == -1) return 0.5f * (x - e) - 0.5f;
    if (k == 1) return x < -0.25f ? -2.0f * (e - (x + 0.5f)) : one + 2.0f * (x - e);
    if (k <= -2 || k > 56) {
        y = one - (e - x);
        y = __uint_as_float(__float_as_uint(y) + ((uint32_t)k << 23));
        return y - one;
    }
    if (k < 23) {
        t = __uint_as_float(0x3f800000u - (0x1000000u >> k));
        y = t - (e - x);
        y = __uint_as_float(__float_as_uint(y) + ((uint32_t)k << 23));
    } else {
        t = __uint_as_float((uint32_t)(0x7f - k) << 23);
        y = x - (e + t);
        y += one;
        y = __uint_as_float(__float_as_uint(y) + ((uint32_t)k << 23));
    }
    return y;
}

__device__ __forceinline__ float glibc_tanhf(float x)
{
    const float one = 1.0f, two = 2.0f, tiny = 1.0e-30f;
    float t, z;
    const int jx = __float_as_int(x), ix = jx & 0x7fffffff;
    if (ix >= 0x7f800000) return jx >= 0 ? __fdiv_rn(one, x) + one : __fdiv_rn(one, x) - one;
    if (ix < 0x41b00000) {
        if (ix == 0) return x;
        if (ix < 0x24000000) return x * (one + x);
        if (ix >= 0x3f800000) {
            t = glibc_expm1f(two * fabsf(x));
            z = one - __fdiv_rn(two, t + two);
        } else {
            t = glibc_expm1f(-two * fabsf(x));
            z = __fdiv_rn(-t, t + two);
        }
    } else z = one - tiny;
    return jx >= 0 ? z : -z;
}

// ---------------------------------------------------------------- tie-break stream (include/lzb200.h)
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__host__ __device__ __forceinline__ uint32_t tiebreak_hash(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth)
{
    uint64_t x = mix64(seed + 0x9E3779B97F4A7C15ULL * (uint64_t)(sim + 1u));
    x = mix64(x ^ (((uint64_t)root << 32) | (uint64_t)depth));
    return (uint32_t)(x >> 33);
}

// chance-draw bits of the stochastic tree (include/lzb200.h: lzb_chance_bits)
__host__ __device__ __forceinline__ uint64_t chance_bits(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth)
{
    const uint64_t x = mix64(seed + 0x9E3779B97F4A7C15ULL * (uint64_t)(sim + 1u));
    return mix64(x ^ (((uint64_t)root << 32) | (uint64_t)depth) ^ 0xC2B2AE3D27D4EB4FULL);
}

// ---------------------------------------------------------------- warp helpers
constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ float warp_min(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ int warp_max_i(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ int warp_min_i(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

} // namespace lzb
