// Single-node Gumbel MuZero helpers: the free functions the reference exposes to its unit tests through gmz_tree.pyx:97-149
// (csoftmax, compute_mixed_value, rescale_qvalues, qtransform_completed_by_mix_value, score_considered, and the two
// single-node selections cselect_root_child / cselect_interior_child; gmz cnode.cpp:701-790,904-1040,1097-1132) over
// vectors in device memory.  They are serial by definition (one node, left-to-right fp32 sums), so ONE thread evaluates
// them with the tree's own libm (glibc_expf / glibc_logf, no FMA contraction) - a parity surface for a maintainer's unit
// tests, not a hot path: the batched search does the same arithmetic in select_gumbel_g / k_root_outputs.
#pragma once
#include <cuda_runtime.h>

#include "lzb_math.cuh"

namespace lzb {

struct GumbelHelperArgs {
    int op, n;
    const float *a, *b, *c; // op-specific float vectors [n]
    const int *v;           // visit counts [n]
    float s0, s1, s2, s3;   // op-specific scalars
    int i0, i1;
    float *out;             // [n] (or [1])
    int *out_i;             // [1]
    float *tmp;             // [3 n] scratch
};

// csoftmax (gmz cnode.cpp:904-933): exp(x - m - log(sum exp(x - m))), in place
__device__ inline void gh_softmax(float *x, int n)
{
    float m = x[0];
    for (int i = 1; i < n; ++i)
        if (x[i] > m) m = x[i];
    float sum = 0.0f;
    for (int i = 0; i < n; ++i) sum += glibc_expf(x[i] - m);
    const float ls = glibc_logf(sum);
    for (int i = 0; i < n; ++i) x[i] = glibc_expf(x[i] - m - ls);
}

// compute_mixed_value (gmz cnode.cpp:935-970); the priors are clamped in place like the reference's by-reference vector
__device__ inline float gh_mixed_value(float raw_value, const float *q, const int *visit, float *prior, int n)
{
    float visit_count_sum = 0.0f, probs_sum = 0.0f, weighted_q_sum = 0.0f;
    const float min_num = -10e7f;
    for (int i = 0; i < n; ++i) visit_count_sum += (float)visit[i];
    for (int i = 0; i < n; ++i) prior[i] = (prior[i] < min_num) ? min_num : prior[i];
    for (int i = 0; i < n; ++i)
        if (visit[i] > 0) probs_sum += prior[i];
    for (int i = 0; i < n; ++i)
        if (visit[i] > 0) weighted_q_sum += prior[i] * q[i] / probs_sum;
    return (raw_value + visit_count_sum * weighted_q_sum) / (visit_count_sum + 1);
}

// rescale_qvalues (gmz cnode.cpp:972-987), in place
__device__ inline void gh_rescale(float *x, int n, float epsilon)
{
    float mx = x[0], mn = x[0];
    for (int i = 1; i < n; ++i) {
        if (x[i] > mx) mx = x[i];
        if (x[i] < mn) mn = x[i];
    }
    float gap = mx - mn;
    gap = (gap < epsilon) ? epsilon : gap;
    for (int i = 0; i < n; ++i) x[i] = (x[i] - mn) / gap;
}

// qtransform_completed_by_mix_value (gmz cnode.cpp:989-1040): q = the node's get_q(discount), prior = the caller's child
// priors (soft-maxed again inside, as the reference does); out[n]
__device__ inline void gh_qtransform(const float *q, float raw_value, const int *visit, const float *prior, int n, float maxvisit_init,
                                     float value_scale, int rescale, float epsilon, float *out, float *ptmp)
{
    for (int i = 0; i < n; ++i) ptmp[i] = prior[i];
    gh_softmax(ptmp, n);
    const float value = gh_mixed_value(raw_value, q, visit, ptmp, n);
    for (int i = 0; i < n; ++i) out[i] = visit[i] > 0 ? q[i] : value;
    if (rescale) gh_rescale(out, n, epsilon);
    int maxv = visit[0];
    for (int i = 1; i < n; ++i)
        if (visit[i] > maxv) maxv = visit[i];
    const float visit_scale = maxvisit_init + (float)maxv;
    for (int i = 0; i < n; ++i) out[i] = out[i] * visit_scale * value_scale;
}

// score_considered (gmz cnode.cpp:1097-1132)
__device__ inline void gh_score_considered(int considered_visit, const float *gumbel, const float *logits, const float *nq, const int *visit,
                                           int n, float *out)
{
    float max_logit = logits[0];
    for (int i = 1; i < n; ++i)
        if (logits[i] > max_logit) max_logit = logits[i];
    for (int i = 0; i < n; ++i) {
        const float penalty = (visit[i] == considered_visit) ? 0.0f : -INFINITY;
        const float s = gumbel[i] + (logits[i] - max_logit) + nq[i];
        out[i] = ((-1e9f < s) ? s : -1e9f) + penalty;
    }
}

__device__ inline int gh_first_argmax(const float *s, int n) // the reference's strict '>' scan from -inf
{
    float best = -INFINITY;
    int pos = 0;
    for (int i = 0; i < n; ++i)
        if (s[i] > best) { best = s[i]; pos = i; }
    return pos;
}

enum { GH_SOFTMAX = 0, GH_MIXED_VALUE = 1, GH_RESCALE = 2, GH_QTRANSFORM = 3, GH_SCORE_CONSIDERED = 4, GH_SELECT_ROOT = 5, GH_SELECT_INTERIOR = 6 };

__global__ void k_gumbel_helper(GumbelHelperArgs a)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    const int n = a.n;
    float *t0 = a.tmp, *t1 = a.tmp + n, *t2 = a.tmp + 2 * n;
    switch (a.op) {
    case GH_SOFTMAX:
        for (int i = 0; i < n; ++i) a.out[i] = a.a[i];
        gh_softmax(a.out, n);
        break;
    case GH_MIXED_VALUE: // s0 raw value, a = q, v = visits, b = priors
        for (int i = 0; i < n; ++i) t0[i] = a.b[i];
        a.out[0] = gh_mixed_value(a.s0, a.a, a.v, t0, n);
        break;
    case GH_RESCALE: // a = values, s0 = epsilon
        for (int i = 0; i < n; ++i) a.out[i] = a.a[i];
        gh_rescale(a.out, n, a.s0);
        break;
    case GH_QTRANSFORM: // a = q, s0 raw value, v, b = priors, s1 maxvisit_init, s2 value_scale, i0 rescale, s3 epsilon
        gh_qtransform(a.a, a.s0, a.v, a.b, n, a.s1, a.s2, a.i0, a.s3, a.out, t0);
        break;
    case GH_SCORE_CONSIDERED: // i0 considered visit, a = gumbel, b = logits, c = normalised q, v
        gh_score_considered(a.i0, a.a, a.b, a.c, a.v, n, a.out);
        break;
    case GH_SELECT_ROOT: // cselect_root_child (:701-745): a = q, s0 raw value, v, b = priors, c = gumbel, i0 = considered visit
        gh_qtransform(a.a, a.s0, a.v, a.b, n, 50.0f, 0.1f, 1, 1e-8f, t1, t0);
        gh_score_considered(a.i0, a.c, a.b, t1, a.v, n, t2);
        a.out_i[0] = gh_first_argmax(t2, n);
        break;
    case GH_SELECT_INTERIOR: { // cselect_interior_child (:747-790): a = q, s0 raw value, v, b = priors
        gh_qtransform(a.a, a.s0, a.v, a.b, n, 50.0f, 0.1f, 1, 1e-8f, t1, t0);
        int vsum = 0;
        for (int i = 0; i < n; ++i) { t1[i] = a.b[i] + t1[i]; vsum += a.v[i]; }
        gh_softmax(t1, n);
        for (int i = 0; i < n; ++i) t2[i] = t1[i] - (float)a.v[i] / (float)(1 + vsum);
        a.out_i[0] = gh_first_argmax(t2, n);
        break;
    }
    default: break;
    }
}

} // namespace lzb
