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
        // one Philox block (four words) serves the two samples 2q and 2q + 1; `uniforms` keeps one (u1, u2) pair per sample
        const int KD = a.K * a.D;
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
    for (int k = 0; k < a.K; ++k) {
        float best = -INFINITY;
        int bj = 0x7fffffff;
#pragma unroll
        for (int r = 0; r < kMaxPerLane; ++r) {
            const int j = lane + 32 * r;
            if (j < a.A && (key[r] > best || (key[r] == best && j < bj))) { best = key[r]; bj = j; }
        }
        // warp arg-max, lowest action id on equal keys (std::sort on equal keys is unspecified in the reference)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(kFull, best, o);
            const int oj = __shfl_xor_sync(kFull, bj, o);
            if (ob > best || (ob == best && oj < bj)) { best = ob; bj = oj; }
        }
        if (bj == 0x7fffffff) bj = k; // fewer than K actions with p > 0: fill deterministically
        if (lane == 0) a.out[(size_t)i * a.K + k] = (float)bj;
#pragma unroll
        for (int r = 0; r < kMaxPerLane; ++r)
            if (lane + 32 * r == bj) key[r] = -INFINITY; // without replacement
    }
}

} // namespace lzb
