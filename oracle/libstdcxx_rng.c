/*
 * TEST INFRASTRUCTURE ONLY (oracle/): restatement of the libstdc++ random-number arithmetic the Sampled-EfficientZero
 * reference tree samples its K actions with, so that the sampler - not only the tree below it - can be pinned.
 *
 * The algorithm lives in a third-party dependency that is absent from /root/reference: GCC's libstdc++ (this image:
 * GCC 13.3, libstdc++.so.6.0.33; headers bits/random.h, bits/random.tcc).  Restated here from the published algorithm:
 *   std::default_random_engine = minstd_rand0 = linear_congruential_engine<uint_fast32_t, 16807, 0, 2147483647>
 *       seed(s): x = s mod m, 0 -> 1 (random.h "linear_congruential_engine::seed");  next: x = 16807 x mod m
 *   std::generate_canonical<T, digits>(g) (random.tcc): k = max(1, ceil(digits / log2(R))) draws with R = max - min + 1
 *       = 2147483646;  sum = sum_j T(x_j - 1) R^j;  ret = sum / R^k;  ret >= 1 -> nextafter(1, 0)
 *       float: one draw, R^1 rounded to float = 2^31;  double: two draws, R^2 rounded to double
 *   std::normal_distribution<float>::operator() (random.tcc): Marsaglia polar method,
 *       do { x = 2 u1 - 1; y = 2 u2 - 1; r2 = x x + y y; } while (r2 > 1 || r2 == 0);
 *       mult = sqrt(-2 log(r2) / r2);  ret = y mult (x mult is saved - the reference builds a fresh distribution per
 *       draw, cnode.cpp:262, so the saved half is never used);  ret = ret * stddev + mean
 *   std::uniform_real_distribution<double>(0, 1): generate_canonical<double, 53>.
 * Anchored on the reference's own call sites: ctree_sampled_efficientzero/lib/cnode.cpp:251-283 (continuous) and
 * :326-406 (discrete).  Parity: PINNED - tests/test_sampler_oracle.py compares this file with libstdc++ itself
 * (oracle/ref_harness_sampled.cpp: ref_libstdcxx_normal / ref_libstdcxx_uniform) and with the action tables the compiled
 * reference tree samples under a pinned clock, bit for bit; tests/golden/sez_sampler_*.npz carry those tables to machines
 * without /root/reference.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

float orc_expf(float x);
float orc_logf(float x);
float orc_tanhf(float x);

#define LCG_M 2147483647ull

static uint64_t lcg_seed(uint32_t s)
{
    uint64_t x = (uint64_t)s % LCG_M;
    return x == 0 ? 1 : x;
}
static uint64_t lcg_next(uint64_t *x)
{
    *x = (16807ull * *x) % LCG_M;
    return *x;
}

/* generate_canonical<float, 24>: one draw; float(x - 1) rounds to nearest, R rounds to 2^31 */
static float canonical_f32(uint64_t *x)
{
    const float sum = (float)(uint32_t)(lcg_next(x) - 1);
    float ret = sum / 2147483648.0f;
    if (ret >= 1.0f) ret = 0x1.fffffep-1f; /* nextafter(1, 0) */
    return ret;
}

/* generate_canonical<double, 53>: two draws; tmp = R then R * R (the product is taken in long double, then rounded) */
static double canonical_f64(uint64_t *x)
{
    const double r = 2147483646.0;
    double sum = (double)(lcg_next(x) - 1);
    sum += (double)(lcg_next(x) - 1) * r;
    const double tmp = (double)(2147483646.0L * 2147483646.0L);
    double ret = sum / tmp;
    if (ret >= 1.0) ret = 0x1.fffffffffffffp-1;
    return ret;
}

static float normal_f32(uint64_t *x, float mean, float stddev)
{
    float a, b, r2;
    do {
        a = 2.0f * canonical_f32(x) - 1.0f;
        b = 2.0f * canonical_f32(x) - 1.0f;
        r2 = a * a + b * b;
    } while (r2 > 1.0f || r2 == 0.0f);
    const float mult = sqrtf(-2.0f * orc_logf(r2) / r2);
    float ret = b * mult;
    ret = ret * stddev + mean;
    return ret;
}

void orc_libstdcxx_normal(uint32_t seed, float mu, float sigma, int count, float *out)
{
    uint64_t x = lcg_seed(seed);
    for (int i = 0; i < count; ++i) out[i] = normal_f32(&x, mu, sigma);
}

void orc_libstdcxx_uniform(uint32_t seed, int count, double *out)
{
    uint64_t x = lcg_seed(seed);
    for (int i = 0; i < count; ++i) out[i] = canonical_f64(&x);
}

/* The K actions CNode::expand samples for root i of a batched call whose k-th expand() sees the clock value
 * seed_base + k (cnode.cpp:236-283 continuous, :326-406 discrete).  out: [n, K, D] (D = 1 and the action id as a float
 * for discrete).  logits: [n, 2 D] (mu | sigma) continuous, [n, A] discrete.
 * continuous: a = tanhf(normal(mu_d, sigma_d)) for k in K, d in D, one engine per root.
 * discrete: p_j = float(double(expf(l_j)) / (double(sum_j expf(l_j)) + 1e-6)), the sum taken in float in index order
 *   (:340-347); key_j = pow(u_j, 1. / p_j) with u_j = uniform_real_distribution<double>(0, 1) in index order (:377);
 *   the K largest keys in descending order (:391-397).  Equal keys: the lower index first (std::sort leaves that order
 *   unspecified; keys tie only when they underflow to 0). */
void orc_sample_libstdcxx(const float *logits, int n, int K, int D, int A, int continuous, uint64_t seed_base, float *out)
{
    for (int i = 0; i < n; ++i) {
        uint64_t x = lcg_seed((uint32_t)(seed_base + (uint64_t)i));
        if (continuous) {
            const float *mu = logits + (size_t)i * 2 * D, *sg = mu + D;
            for (int k = 0; k < K; ++k)
                for (int d = 0; d < D; ++d) out[((size_t)i * K + k) * D + d] = orc_tanhf(normal_f32(&x, mu[d], sg[d]));
            continue;
        }
        const float *l = logits + (size_t)i * A;
        float sum = 0;
        for (int j = 0; j < A; ++j) sum += orc_expf(l[j]);
        double key[512];
        unsigned char taken[512];
        const int An = A < 512 ? A : 512;
        for (int j = 0; j < An; ++j) {
            const float p = (float)((double)orc_expf(l[j]) / ((double)sum + 1e-6));
            key[j] = pow(canonical_f64(&x), 1. / p);
            taken[j] = 0;
        }
        for (int k = 0; k < K; ++k) {
            int best = -1;
            for (int j = 0; j < An; ++j)
                if (!taken[j] && (best < 0 || key[j] > key[best])) best = j;
            if (best < 0) best = 0;
            taken[best] = 1;
            out[(size_t)i * K + k] = (float)best;
        }
    }
}
