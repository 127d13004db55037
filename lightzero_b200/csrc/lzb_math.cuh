// Device math that has to agree bit-for-bit with what the reference tree computes on the host.
//
// The reference (lzero/mcts/ctree/ctree_muzero/lib/cnode.cpp:129,673 and siblings) calls exp()/log()
// on floats, i.e. glibc's expf/logf.  glibc >= 2.27 implements them with the ARM optimized-routines
// algorithms (sysdeps/ieee754/flt-32/e_expf.c, e_logf.c): table + cubic in double, one final rounding
// to float.  CUDA's expf/logf differ in the last ulp, which would change softmax priors and, through
// them, pUCT arg-maxes.  So the same algorithm is evaluated here in fp64 with the fused operations
// the x86-64 FMA build of glibc performs.  This translation unit is compiled with -fmad=false: the
// only fused multiply-adds are the explicit fma() calls below.
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

// expf as glibc 2.39 x86-64 (FMA ifunc variant) computes it.
__device__ __forceinline__ float glibc_expf(float x)
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
    const double s = __longlong_as_double((long long)(LZB_TABLE_AT(kExpTab, ki & 31u) + (ki << 47)));
    const double z = fma(0x1.c6af84b912394p-20, r, 0x1.ebfce50fac4f3p-13);
    const double r2 = __dmul_rn(r, r);
    double y = fma(0x1.62e42ff0c52d6p-6, r, 1.0);
    y = fma(z, r2, y);
    y = __dmul_rn(y, s);
    return __double2float_rn(y);
}

// logf as glibc 2.39 x86-64 (FMA ifunc variant) computes it.
__device__ __forceinline__ float glibc_logf(float x)
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
    const double invc = LZB_TABLE_AT(kLogTab, 2 * i), logc = LZB_TABLE_AT(kLogTab, 2 * i + 1);
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
