// Action sampler of the Sampled-EfficientZero tree (K6): the K actions CNode::expand draws for a new node
// (ctree_sampled_efficientzero/lib/cnode.cpp:236-283 continuous, :326-406 discrete).  The reference seeds a
// std::default_random_engine from the wall clock inside expand(), so its draws cannot be reproduced; the
// DISTRIBUTIONS are restated here over a counter-based generator (Philox4x32-10 keyed by seed, offset by
// (call counter, global root, k)), which makes a search reproducible and independent of how roots are sharded.
//   continuous: a[k][d] = tanhf(mu[d] + sigma[d] * z),  z ~ N(0,1)            logits = {mu[0..D), sigma[0..D)}
//   discrete:   K actions without replacement: keys u^(1/p) sorted descending, p = exp(l) / (sum exp(l) + 1e-6),
//               compared as log(u) / p (same order); a[k][0] = (float)action id
// `uniforms` (optional, tests): pre-drawn u in (0,1), [N, A] discrete / [N, K, D, 2] continuous, used instead of Philox.
#pragma once
#include "lzb_math.cuh"

namespace lzb {

struct Philox {
    uint32_t key[2];
    __device__ __forceinline__ Philox(uint64_t seed) { key[0] = (uint32_t)seed; key[1] = (uint32_t)(seed >> 32); }
    __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const
    {
        uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
            c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};

// (0, 1) from 32 random bits: never 0 or 1, so log(u) is finite
__device__ __forceinline__ float unit_open(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

struct SampleArgs {
    const float *logits;   // [N, 2D] continuous | [N, A] discrete
    const float *uniforms; // optional
    float *out;            // [N, K, D]
    int N, K, D, A, continuous, root_offset;
    unsigned long long seed;
    unsigned counter;      // which expansion of the search (0 = roots, s + 1 = simulation s)
    const unsigned long long *seed_dev; // optional: the seed is read from device memory (a captured CUDA graph replays
                                        // with fresh draws when the caller rewrites this word between replays)
};

// one warp per root
__global__ void k_sample_actions(SampleArgs a)
{
    const int lane = threadIdx.x & 31, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= a.N) return;
    const Philox rng(a.seed_dev ? *a.seed_dev : a.seed);
    const uint32_t groot = (uint32_t)(a.root_offset + i);
    if (a.continuous) {
        const float *mu = a.logits + (size_t)i * 2 * a.D, *sg = mu + a.D;
        const int KD = a.K * a.D;
#ifndef LZB_X_SAMPLER_ONE_PER_PAIR
        // Philox path: one block (four words) serves FOUR samples - each (u1, u2) pair gives both Box-Muller variates
        // (radius * cos and radius * sin), halving the Philox / log / sqrt work per sample (round 2: the sampler kernel was
        // 26 % of a Sampled search).  Injected uniforms (tests) keep one pair per sample and the cosine variate, below.
        if (!a.uniforms) {
            for (int q = lane; 4 * q < KD; q += 32) {
                const uint4 r = rng(a.counter, groot, (uint32_t)q, 0x5a3c0001u);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float u1 = unit_open(h ? r.z : r.x), u2 = unit_open(h ? r.w : r.y);
                    const float rad = sqrtf(-2.0f * logf(u1));
                    float sn, cs;
                    sincospif(2.0f * u2, &sn, &cs);
#pragma unroll
                    for (int w = 0; w < 2; ++w) {
                        const int t = 4 * q + 2 * h + w;
                        if (t < KD) {
                            const int d = t - (t / a.D) * a.D;
                            a.out[(size_t)i * KD + t] = tanhf(mu[d] + sg[d] * (rad * (w ? sn : cs)));
                        }
                    }
                }
            }
            return;
        }
#endif
        // one Philox block (four words) serves the two samples 2q and 2q + 1; `uniforms` keeps one (u1, u2) pair per sample
        for (int q = lane; 2 * q < KD; q += 32) {
            uint4 r = make_uint4(0, 0, 0, 0);
            if (!a.uniforms) r = rng(a.counter, groot, (uint32_t)q, 0x5a3c0001u);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int t = 2 * q + h;
                if (t < KD) {
                    const int d = t - (t / a.D) * a.D;
                    float u1, u2;
                    if (a.uniforms) {
                        u1 = a.uniforms[((size_t)i * KD + t) * 2];
                        u2 = a.uniforms[((size_t)i * KD + t) * 2 + 1];
                    } else {
                        u1 = unit_open(h ? r.z : r.x);
                        u2 = unit_open(h ? r.w : r.y);
                    }
                    const float z = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2); // Box-Muller
                    a.out[(size_t)i * KD + t] = tanhf(mu[d] + sg[d] * z);
                }
            }
        }
        return;
    }
    // discrete: Gumbel-top-K style keys; each lane owns actions lane, lane+32, ...; K rounds of warp arg-max
    const float *l = a.logits + (size_t)i * a.A;
    float sum = 0.0f; // left-to-right like the reference loop (cnode.cpp:340-343); every lane runs it
    for (int j = 0; j < a.A; ++j) sum += expf(l[j]);
    constexpr int kMaxPerLane = 8; // A <= 256
    float key[kMaxPerLane];
#pragma unroll
    for (int r = 0; r < kMaxPerLane; ++r) {
        const int j = lane + 32 * r;
        key[r] = -INFINITY;
        if (j < a.A) {
            const float p = expf(l[j]) / (sum + 1e-6f);
            const float u = a.uniforms ? a.uniforms[(size_t)i * a.A + j] : unit_open(rng(a.counter, groot, (uint32_t)j, 0x5a3c0002u).x);
            key[r] = (p > 0.0f) ? logf(u) / p : -INFINITY; // log of u^(1/p); larger = earlier in the descending sort
        }
    }
    unsigned taken = 0; // bit r: this lane's action lane + 32 r has been drawn
    for (int k = 0; k < a.K; ++k) {
        float best = -INFINITY;
        int bj = 0x7fffffff;
#pragma unroll
        for (int r = 0; r < kMaxPerLane; ++r) {
            const int j = lane + 32 * r;
            // actions whose key is -inf (p underflowed to 0, or log(u) / p overflowed) stay candidates until drawn: they
            // fill the table, lowest id first, once the finite keys are used up - K distinct actions like the reference's sort
            if (j < a.A && !((taken >> r) & 1u) && (key[r] > best || (key[r] == best && j < bj))) { best = key[r]; bj = j; }
        }
        // warp arg-max, lowest action id on equal keys (std::sort on equal keys is unspecified in the reference)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(kFull, best, o);
            const int oj = __shfl_xor_sync(kFull, bj, o);
            if (ob > best || (ob == best && oj < bj)) { best = ob; bj = oj; }
        }
        if (bj == 0x7fffffff) bj = 0; // unreachable for K <= A (checked by the C ABI)
        if (lane == 0) a.out[(size_t)i * a.K + k] = (float)bj;
#pragma unroll
        for (int r = 0; r < kMaxPerLane; ++r)
            if (lane + 32 * r == bj) taken |= 1u << r; // without replacement
    }
}

// ------------------------------------------------------------------------------------------------ parity sampler
// The reference's OWN draws (sez cnode.cpp:251-283, :336-406): std::default_random_engine (minstd_rand0) seeded with the
// truncated wall clock, std::normal_distribution<float> (Marsaglia polar, a fresh object per draw) or
// std::uniform_real_distribution<double>, as GCC 13's libstdc++ computes them - restated in oracle/libstdcxx_rng.c (where
// the algorithm is described and pinned against libstdc++ itself) and here for the device.  Root i is seeded with
// (unsigned)(seed_base + i): what the reference's i-th expand() of a batched call sees when its clock reads seed_base + i.
// One thread per root (the engine is a serial chain with data-dependent rejections); production sampling is
// k_sample_actions above - this mode exists so that the sampler itself is testable bit for bit against the reference.
struct ParitySampleArgs {
    const float *logits; // [N, 2D] continuous | [N, A] discrete
    float *out;          // [N, K, D]
    int N, K, D, A, continuous;
    unsigned long long seed_base;
};

struct MinStd { // linear_congruential_engine<uint_fast32_t, 16807, 0, 2147483647>
    unsigned long long x;
    __device__ __forceinline__ explicit MinStd(unsigned seed)
    {
        x = (unsigned long long)seed % 2147483647ull;
        if (x == 0) x = 1;
    }
    __device__ __forceinline__ unsigned long long next() { return x = (16807ull * x) % 2147483647ull; }
    // std::generate_canonical<float, 24>: one draw, float(x - 1) / float(R), R = 2147483646 rounds to 2^31
    __device__ __forceinline__ float canonical_f32()
    {
        const float r = __fdiv_rn(__uint2float_rn((unsigned)(next() - 1ull)), 2147483648.0f);
        return r >= 1.0f ? __uint_as_float(0x3f7fffffu) : r;
    }
    // std::generate_canonical<double, 53>: two draws, (x1 - 1 + (x2 - 1) R) / double(R * R)
    __device__ __forceinline__ double canonical_f64()
    {
        double sum = (double)(next() - 1ull);
        sum = __dadd_rn(sum, __dmul_rn((double)(next() - 1ull), 2147483646.0));
        const double r = __ddiv_rn(sum, 0x1.fffffff000000p+61); // double(2147483646^2)
        return r >= 1.0 ? __longlong_as_double(0x3fefffffffffffffLL) : r;
    }
    __device__ __forceinline__ float normal(float mean, float stddev)
    {
        float a, b, r2;
        do {
            a = 2.0f * canonical_f32() - 1.0f;
            b = 2.0f * canonical_f32() - 1.0f;
            r2 = a * a + b * b;
        } while (r2 > 1.0f || r2 == 0.0f);
        const float mult = __fsqrt_rn(__fdiv_rn(-2.0f * glibc_logf(r2), r2));
        return (b * mult) * stddev + mean;
    }
};

__global__ void k_sample_actions_libstdcxx(ParitySampleArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.N) return;
    MinStd gen((unsigned)(a.seed_base + (unsigned long long)i));
    if (a.continuous) {
        const float *mu = a.logits + (size_t)i * 2 * a.D, *sg = mu + a.D;
        for (int k = 0; k < a.K; ++k)
            for (int d = 0; d < a.D; ++d) a.out[((size_t)i * a.K + k) * a.D + d] = glibc_tanhf(gen.normal(mu[d], sg[d]));
        return;
    }
    const float *l = a.logits + (size_t)i * a.A;
    float sum = 0.0f;
    for (int j = 0; j < a.A; ++j) sum += glibc_expf(l[j]);
    double key[256];
    const double den = __dadd_rn((double)sum, 1e-6);
    for (int j = 0; j < a.A; ++j) {
        const float p = __double2float_rn(__ddiv_rn((double)glibc_expf(l[j]), den));
        key[j] = pow(gen.canonical_f64(), __ddiv_rn(1.0, (double)p)); // std::pow(u, 1. / prob), sez cnode.cpp:377
    }
    for (int k = 0; k < a.K; ++k) { // the K largest keys, descending (:391-397); equal keys: the lower index first
        int best = -1;
        for (int j = 0; j < a.A; ++j)
            if (key[j] >= 0.0 && (best < 0 || key[j] > key[best])) best = j;
        if (best < 0) best = 0;
        key[best] = -1.0; // taken (keys are >= 0)
        a.out[(size_t)i * a.K + k] = (float)best;
    }
}

} // namespace lzb
