/*
 * TEST INFRASTRUCTURE ONLY - see oracle/oracle.h for the role and parity status of this file.
 *
 * Plain-C restatement of the reference's batched MCTS tree, kept in the reference's own shape
 * (heap nodes with per-node legal-action lists and child arrays, serial loops over roots) so that
 * it reads against the four lib/cnode.cpp files under lzero/mcts/ctree.  All tree arithmetic is float, evaluated
 * operation by operation (compiled with -ffp-contract=off), as the reference's -O2 x86-64 build
 * does.  expf/logf are the restated glibc routines of oracle/glibc_flt32.c.
 *
 * File:line citations are relative to /root/reference/lzero/mcts/ctree/ ; "mz" = ctree_muzero/lib/cnode.cpp,
 * "ez" = ctree_efficientzero/lib/cnode.cpp, "gmz" = ctree_gumbel_muzero/lib/cnode.cpp,
 * "sez" = ctree_sampled_efficientzero/lib/cnode.cpp, "smz" = ctree_stochastic_muzero/lib/cnode.cpp,
 * "mm" = common_lib/cminimax.cpp.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

#define FLOAT_MAX 1000000.0f /* common_lib/cminimax.h:9 */
#define FLOAT_MIN (-FLOAT_MAX)

typedef struct Node {
    int visit_count, to_play, latent_index, batch_index, best_action, is_reset;
    int is_chance; /* smz cnode.h:25: outcome node (children drawn from the priors), else decision node */
    float reward; /* `reward` (mz, gmz) or `value_prefix` (ez, sez) */
    float prior, value_sum, raw_value;
    int n_legal;   /* legal_actions.size() */
    int *legal;    /* discrete trees: action ids in list order */
    float *actions;/* sampled tree: [n_legal][D] action vectors */
    int n_children;/* children.size(); 0 == not expanded (mz:210-217) */
    struct Node *children; /* child j <-> legal[j] */
} Node;

typedef struct { float maximum, minimum, value_delta_max; } MinMax; /* cminimax.h:14-24 */

struct OrcForest {
    int variant, n, A, K, D, continuous;
    int C, W; /* smz: chance_space_size and the row stride max(A, C) of every logits array */
    Node *roots;
    MinMax *mm;
    float *gumbel; /* gmz: shared N(0)-seeded table, see orc_gumbel_noise */
    /* CSearchResults (mz cnode.h:67-79), recreated by every traverse */
    Node ***paths;
    int *path_len;
    /* pinned tie-break stream */
    int rng_mode;
    uint64_t rng_seed;
    uint32_t root_offset, sim;
};

/* ------------------------------------------------------------------ tie-break stream */
static inline uint64_t mix64(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

/* Same contract as oracle/rng_pin.c and include/lzb200.h. */
uint32_t orc_tiebreak_hash(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth)
{
    uint64_t x = mix64(seed + 0x9E3779B97F4A7C15ULL * (uint64_t)(sim + 1u));
    x = mix64(x ^ (((uint64_t)root << 32) | (uint64_t)depth));
    return (uint32_t)(x >> 33);
}

static int pinned_rand(const OrcForest *f, uint32_t root, uint32_t depth)
{
    if (f->rng_mode == 2) return (int)orc_tiebreak_hash(f->rng_seed, f->sim, f->root_offset + root, depth);
    return (int)orc_tiebreak_hash(f->rng_seed, f->sim, 0u, 0u);
}

/* ------------------------------------------------------------------ MinMaxStats (mm:7-45) */
static void mm_init(MinMax *m) { m->maximum = FLOAT_MIN; m->minimum = FLOAT_MAX; m->value_delta_max = 0.f; }
static void mm_update(MinMax *m, float v)
{
    if (v > m->maximum) m->maximum = v;
    if (v < m->minimum) m->minimum = v;
}
static float mm_normalize(const MinMax *m, float value)
{
    float norm = value;
    float delta = m->maximum - m->minimum;
    if (delta > 0) {
        if (delta < m->value_delta_max) norm = (norm - m->minimum) / m->value_delta_max;
        else norm = (norm - m->minimum) / delta;
    }
    return norm;
}

/* ------------------------------------------------------------------ nodes */
static void node_init(Node *nd, float prior)
{ /* mz:61-79 */
    memset(nd, 0, sizeof(*nd));
    nd->prior = prior;
    nd->best_action = -1;
    nd->latent_index = -1;
    nd->batch_index = -1;
}

static void node_free(Node *nd)
{
    for (int j = 0; j < nd->n_children; ++j) node_free(&nd->children[j]);
    free(nd->children);
    free(nd->legal);
    free(nd->actions);
}

static float node_value(const Node *nd)
{ /* mz:219-235 */
    if (nd->visit_count == 0) return 0.0f;
    return nd->value_sum / (float)nd->visit_count;
}

/* CNode::expand for the discrete trees (mz:83-147, ez:88-151, gmz:94-156). */
static void node_expand(Node *nd, int to_play, int latent_index, int batch_index, float reward, float raw_value,
                        const float *logits, int A)
{
    nd->to_play = to_play;
    nd->latent_index = latent_index;
    nd->batch_index = batch_index;
    nd->reward = reward;
    nd->raw_value = raw_value;
    if (nd->n_legal == 0) { /* mz:101-107: only roots carry masks */
        nd->legal = (int *)malloc(sizeof(int) * (size_t)A);
        for (int a = 0; a < A; ++a) nd->legal[a] = a;
        nd->n_legal = A;
    }
    float policy_max = FLOAT_MIN;
    for (int j = 0; j < nd->n_legal; ++j)
        if (policy_max < logits[nd->legal[j]]) policy_max = logits[nd->legal[j]];
    float *tmp = (float *)malloc(sizeof(float) * (size_t)nd->n_legal);
    float policy_sum = 0.0f;
    for (int j = 0; j < nd->n_legal; ++j) { /* mz:127-132: serial fp32 sum in legal order */
        float t = orc_expf(logits[nd->legal[j]] - policy_max);
        policy_sum += t;
        tmp[j] = t;
    }
    nd->children = (Node *)malloc(sizeof(Node) * (size_t)nd->n_legal);
    nd->n_children = nd->n_legal;
    for (int j = 0; j < nd->n_legal; ++j) node_init(&nd->children[j], tmp[j] / policy_sum);
    free(tmp);
}

/* Sampled CNode::expand (sez:194-452) with the K sampled actions INJECTED (the sampler itself is
 * wall-clock seeded, sez:251,336).  Priors never reach selection (sez:1054-1079 takes the uniform
 * branch because std::string::compare returns 0 on equality), so they are not restated. */
static void node_expand_sampled(Node *nd, int to_play, int latent_index, int batch_index, float value_prefix,
                                const float *actions, int K, int D)
{
    nd->to_play = to_play;
    nd->latent_index = latent_index;
    nd->batch_index = batch_index;
    nd->reward = value_prefix;
    nd->actions = (float *)malloc(sizeof(float) * (size_t)K * (size_t)D);
    memcpy(nd->actions, actions, sizeof(float) * (size_t)K * (size_t)D);
    nd->n_legal = K;
    nd->children = (Node *)malloc(sizeof(Node) * (size_t)K);
    nd->n_children = K; /* children.size(); assumes no 6-decimal hash collisions (sez:79-111) */
    for (int j = 0; j < K; ++j) node_init(&nd->children[j], 0.0f);
}

/* Stochastic MuZero CNode::expand (smz:96-166): a chance node takes C chance logits, has its reward zeroed and
 * decision children; a decision node takes A policy logits and has chance children.  The `child_is_chance`
 * argument of the reference is overwritten before use (smz:115-122), so it is not restated. */
static void node_expand_smz(Node *nd, int to_play, int latent_index, int batch_index, float reward, const float *logits,
                            int A, int C)
{
    node_expand(nd, to_play, latent_index, batch_index, reward, 0.0f, logits, nd->is_chance ? C : A);
    if (nd->is_chance) nd->reward = 0.0f;
    for (int j = 0; j < nd->n_children; ++j) nd->children[j].is_chance = !nd->is_chance;
}

/* Counter-based source of the chance draws: 64 bits per (seed, simulation, global root, depth).  The reference
 * seeds an mt19937 from std::random_device for every draw (smz:622-623), which nothing can reproduce; the
 * contract here is explicit, like the tie-break stream's (include/lzb200.h). */
uint64_t orc_chance_bits(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth)
{
    uint64_t x = mix64(seed + 0x9E3779B97F4A7C15ULL * (uint64_t)(sim + 1u));
    return mix64(x ^ (((uint64_t)root << 32) | (uint64_t)depth) ^ 0xC2B2AE3D27D4EB4FULL);
}

/* std::discrete_distribution<>(probs)(gen) of libstdc++ 13 (bits/random.tcc:2657-2714) for a uniform u in [0, 1):
 * probabilities normalised by their double sum, partial sums with the last one forced to 1.0, lower_bound.
 * probs are the children's float priors widened to double (smz:617-620). */
int orc_discrete_draw(const double *probs, int n, double u)
{
    if (n < 2) return 0;
    double sum = 0.0;
    for (int j = 0; j < n; ++j) sum += probs[j];
    double cp = 0.0;
    int pos = 0;
    for (int j = 0; j < n; ++j) {
        cp += probs[j] / sum;
        double c = (j == n - 1) ? 1.0 : cp;
        if (c < u) pos = j + 1; /* lower_bound over a sequence partitioned by (c < u) */
    }
    return pos < n ? pos : n - 1;
}

static int chance_outcome(const Node *nd, double u)
{
    double *p = (double *)malloc(sizeof(double) * (size_t)(nd->n_children > 0 ? nd->n_children : 1));
    for (int j = 0; j < nd->n_children; ++j) p[j] = (double)nd->children[j].prior;
    int pos = orc_discrete_draw(p, nd->n_children, u);
    free(p);
    return pos;
}

static void node_add_noise(Node *nd, float frac, const float *noises)
{ /* mz:149-167: noises indexed by position in the legal list */
    for (int j = 0; j < nd->n_legal; ++j) {
        float prior = nd->children[j].prior;
        nd->children[j].prior = prior * (1 - frac) + noises[j] * frac;
    }
}

static int is_vp(const OrcForest *f) { return f->variant == ORC_EZ || f->variant == ORC_SEZ; }

/* true reward of `child` seen from `parent` (mz:186; ez:191-195) */
static float true_reward(const OrcForest *f, const Node *parent, const Node *child)
{
    if (!is_vp(f)) return child->reward;
    float r = child->reward - parent->reward;
    if (parent->is_reset == 1) r = child->reward;
    return r;
}

static float compute_mean_q(const OrcForest *f, const Node *nd, int is_root, float parent_q, float discount)
{ /* mz:169-203, ez:173-212 */
    float total_unsigned_q = 0.0f;
    int total_visits = 0;
    for (int j = 0; j < nd->n_legal; ++j) {
        const Node *c = &nd->children[j];
        if (c->visit_count > 0) {
            float qsa = true_reward(f, nd, c) + discount * node_value(c);
            total_unsigned_q += qsa;
            total_visits += 1;
        }
    }
    if (is_root && total_visits > 0) return total_unsigned_q / (float)total_visits;
    return (parent_q + total_unsigned_q) / (float)(total_visits + 1);
}

static float ucb_score(const OrcForest *f, const Node *parent, const Node *child, const MinMax *mm, float parent_mean_q,
                       float total_children_visit_counts, float pb_c_base, float pb_c_init, float discount, int players)
{ /* mz:655-699, ez:756-814, sez:1026-1108 */
    float pb_c = orc_logf((total_children_visit_counts + pb_c_base + 1) / pb_c_base) + pb_c_init;
    pb_c *= (sqrtf(total_children_visit_counts) / (float)(child->visit_count + 1));
    float prior_score;
    if (f->variant == ORC_SEZ) prior_score = pb_c * 1 / (float)parent->n_children; /* sez:1078 */
    else prior_score = pb_c * child->prior;
    float value_score = 0.0f;
    if (child->visit_count == 0) {
        value_score = parent_mean_q;
    } else {
        float r = true_reward(f, parent, child);
        if (players == 1) value_score = r + discount * node_value(child);
        else if (players == 2) value_score = r + discount * (-node_value(child));
    }
    value_score = mm_normalize(mm, value_score);
    if (value_score < 0) value_score = 0;
    if (value_score > 1) value_score = 1;
    return prior_score + value_score;
}

/* cselect_child (mz:551-596): order-dependent tie list, then rand() % size. Returns child position. */
static int select_child(const OrcForest *f, const Node *nd, const MinMax *mm, int pb_c_base, float pb_c_init,
                        float discount, float mean_q, int players, uint32_t root, uint32_t depth)
{
    float max_score = FLOAT_MIN;
    const float epsilon = 0.000001f;
    int *ties = (int *)malloc(sizeof(int) * (size_t)(nd->n_legal > 0 ? nd->n_legal : 1));
    int n_ties = 0;
    for (int j = 0; j < nd->n_legal; ++j) {
        float s = ucb_score(f, nd, &nd->children[j], mm, mean_q, (float)(nd->visit_count - 1), (float)pb_c_base,
                            pb_c_init, discount, players);
        if (max_score < s) {
            max_score = s;
            n_ties = 0;
            ties[n_ties++] = j;
        } else if (s >= max_score - epsilon) {
            ties[n_ties++] = j;
        }
    }
    int pos = -1; /* reference: action = 0 when the list is empty (mz:589) */
    if (n_ties > 0) pos = ties[pinned_rand(f, root, depth) % n_ties];
    free(ties);
    return pos;
}

/* carm_score (mz:700-753; ez:816-878): the root child on the trajectory's true action is scored with the value the
 * NEXT state's search produced (ReZero, arXiv 2404.16364) and, once visited, loses its exploration term. */
static float arm_score(const OrcForest *f, const Node *parent, const Node *child, const MinMax *mm, float parent_mean_q,
                       float reuse_value, float total_children_visit_counts, float pb_c_base, float pb_c_init,
                       float discount, int players)
{
    float pb_c = orc_logf((total_children_visit_counts + pb_c_base + 1) / pb_c_base) + pb_c_init;
    pb_c *= (sqrtf(total_children_visit_counts) / (float)(child->visit_count + 1));
    float prior_score = pb_c * child->prior;
    float value_score = 0.0f;
    if (child->visit_count == 0) {
        value_score = parent_mean_q;
    } else {
        float r = true_reward(f, parent, child); /* ez:833-837 */
        if (players == 1) value_score = r + discount * reuse_value;
        else if (players == 2) value_score = r + discount * (-reuse_value);
    }
    value_score = mm_normalize(mm, value_score);
    if (value_score < 0) value_score = 0;
    if (value_score > 1) value_score = 1;
    return (child->visit_count == 0) ? prior_score + value_score : value_score; /* mz:743-751 */
}

/* cselect_root_child (mz:598-653; ez:697-752). Returns child position. */
static int select_root_child(const OrcForest *f, const Node *nd, const MinMax *mm, int pb_c_base, float pb_c_init,
                             float discount, float mean_q, int players, int true_action, float reuse_value,
                             uint32_t root, uint32_t depth)
{
    float max_score = FLOAT_MIN;
    const float epsilon = 0.000001f;
    int *ties = (int *)malloc(sizeof(int) * (size_t)(nd->n_legal > 0 ? nd->n_legal : 1));
    int n_ties = 0;
    for (int j = 0; j < nd->n_legal; ++j) {
        float s;
        if (nd->legal[j] == true_action)
            s = arm_score(f, nd, &nd->children[j], mm, mean_q, reuse_value, (float)(nd->visit_count - 1),
                          (float)pb_c_base, pb_c_init, discount, players);
        else
            s = ucb_score(f, nd, &nd->children[j], mm, mean_q, (float)(nd->visit_count - 1), (float)pb_c_base,
                          pb_c_init, discount, players);
        if (max_score < s) {
            max_score = s;
            n_ties = 0;
            ties[n_ties++] = j;
        } else if (s >= max_score - epsilon) {
            ties[n_ties++] = j;
        }
    }
    int pos = -1;
    if (n_ties > 0) pos = ties[pinned_rand(f, root, depth) % n_ties];
    free(ties);
    return pos;
}

/* ------------------------------------------------------------------ Gumbel MuZero (gmz:701-790, 904-1152) */

/* csoftmax (gmz:904-933): exp(x - m - log(sum)) form, float throughout. */
static void csoftmax(float *x, int n)
{
    float m = x[0];
    for (int i = 1; i < n; ++i)
        if (x[i] > m) m = x[i];
    float sum = 0;
    for (int i = 0; i < n; ++i) sum += orc_expf(x[i] - m);
    float ls = orc_logf(sum);
    for (int i = 0; i < n; ++i) x[i] = orc_expf(x[i] - m - ls);
}
void orc_softmax(float *x, int n) { csoftmax(x, n); }

static float compute_mixed_value(float raw_value, const float *q, const int *visit, float *prior, int n)
{ /* gmz:935-970 */
    float visit_count_sum = 0.0f, probs_sum = 0.0f, weighted_q_sum = 0.0f;
    float min_num = -10e7f;
    for (int i = 0; i < n; ++i) visit_count_sum += (float)visit[i];
    for (int i = 0; i < n; ++i) prior[i] = (prior[i] < min_num) ? min_num : prior[i]; /* std::max(prior, min_num) */
    for (int i = 0; i < n; ++i)
        if (visit[i] > 0) probs_sum += prior[i];
    for (int i = 0; i < n; ++i)
        if (visit[i] > 0) weighted_q_sum += prior[i] * q[i] / probs_sum;
    return (raw_value + visit_count_sum * weighted_q_sum) / (visit_count_sum + 1);
}

/* qtransform_completed_by_mix_value (gmz:989-1040) with the header defaults
 * maxvisit_init=50, value_scale=0.1, rescale_values=true, epsilon=1e-8 (gmz cnode.h:101-102).
 * NB: `prior` are probabilities (post-noise) and are soft-maxed again - reproduced as is. */
static void completed_q(const Node *nd, float discount, int *visit, float *prior, float *out)
{
    int n = nd->n_legal;
    float *q = (float *)malloc(sizeof(float) * (size_t)n);
    float *p = (float *)malloc(sizeof(float) * (size_t)n);
    for (int j = 0; j < n; ++j) {
        const Node *c = &nd->children[j];
        visit[j] = c->visit_count;
        prior[j] = c->prior;
        p[j] = c->prior;
        q[j] = c->reward + discount * node_value(c); /* get_q, gmz:181-197 */
    }
    csoftmax(p, n);
    float value = compute_mixed_value(nd->raw_value, q, visit, p, n);
    for (int j = 0; j < n; ++j) out[j] = visit[j] > 0 ? q[j] : value;
    /* rescale_qvalues (gmz:972-987) */
    float mx = out[0], mn = out[0];
    for (int j = 1; j < n; ++j) {
        if (out[j] > mx) mx = out[j];
        if (out[j] < mn) mn = out[j];
    }
    float gap = mx - mn;
    gap = (gap < 1e-8f) ? 1e-8f : gap; /* std::max(gap, epsilon) */
    for (int j = 0; j < n; ++j) out[j] = (out[j] - mn) / gap;
    int maxv = visit[0];
    for (int j = 1; j < n; ++j)
        if (visit[j] > maxv) maxv = visit[j];
    float visit_scale = 50.0f + (float)maxv;
    for (int j = 0; j < n; ++j) out[j] = out[j] * visit_scale * 0.1f;
    free(q);
    free(p);
}

/* ---- the single-node helpers of gmz_tree.pyx:97-149 over plain arrays (tests/test_gumbel_helpers*.py) */
float orc_mixed_value(float raw_value, const float *q, const int *visit, const float *prior, int n)
{
    float *p = (float *)malloc(sizeof(float) * (size_t)n);
    memcpy(p, prior, sizeof(float) * (size_t)n);
    float r = compute_mixed_value(raw_value, q, visit, p, n);
    free(p);
    return r;
}

void orc_rescale(float *x, int n, float epsilon)
{ /* rescale_qvalues, gmz:972-987 */
    float mx = x[0], mn = x[0];
    for (int j = 1; j < n; ++j) {
        if (x[j] > mx) mx = x[j];
        if (x[j] < mn) mn = x[j];
    }
    float gap = mx - mn;
    gap = (gap < epsilon) ? epsilon : gap;
    for (int j = 0; j < n; ++j) x[j] = (x[j] - mn) / gap;
}

void orc_qtransform(const float *q, float raw_value, const int *visit, const float *prior, int n, float maxvisit_init,
                    float value_scale, int rescale_values, float epsilon, float *out)
{ /* qtransform_completed_by_mix_value, gmz:989-1040, with its arguments instead of the header defaults */
    float *p = (float *)malloc(sizeof(float) * (size_t)n);
    memcpy(p, prior, sizeof(float) * (size_t)n);
    csoftmax(p, n);
    float value = compute_mixed_value(raw_value, q, visit, p, n);
    for (int j = 0; j < n; ++j) out[j] = visit[j] > 0 ? q[j] : value;
    if (rescale_values) orc_rescale(out, n, epsilon);
    int maxv = visit[0];
    for (int j = 1; j < n; ++j)
        if (visit[j] > maxv) maxv = visit[j];
    float visit_scale = maxvisit_init + (float)maxv;
    for (int j = 0; j < n; ++j) out[j] = out[j] * visit_scale * value_scale;
    free(p);
}

void orc_score_considered(int considered_visit, const float *gumbel, const float *logits, const float *nq, const int *visit, int n,
                          float *out)
{ /* score_considered, gmz:1097-1132 */
    float max_logit = logits[0];
    for (int j = 1; j < n; ++j)
        if (logits[j] > max_logit) max_logit = logits[j];
    for (int j = 0; j < n; ++j) {
        float penalty = (visit[j] == considered_visit) ? 0.0f : -INFINITY;
        float s = gumbel[j] + (logits[j] - max_logit) + nq[j];
        out[j] = ((-1e9f < s) ? s : -1e9f) + penalty;
    }
}

static int first_argmax(const float *s, int n)
{
    float best = -INFINITY;
    int pos = 0;
    for (int j = 0; j < n; ++j)
        if (s[j] > best) { best = s[j]; pos = j; }
    return pos;
}

int orc_select_root_arrays(const float *q, float raw_value, const int *visit, const float *prior, const float *gumbel, int n,
                           int considered_visit)
{ /* cselect_root_child, gmz:701-745 */
    float *cq = (float *)malloc(sizeof(float) * (size_t)n), *sc = (float *)malloc(sizeof(float) * (size_t)n);
    orc_qtransform(q, raw_value, visit, prior, n, 50.0f, 0.1f, 1, 1e-8f, cq);
    orc_score_considered(considered_visit, gumbel, prior, cq, visit, n, sc);
    int pos = first_argmax(sc, n);
    free(cq); free(sc);
    return pos;
}

int orc_select_interior_arrays(const float *q, float raw_value, const int *visit, const float *prior, int n)
{ /* cselect_interior_child, gmz:747-790 */
    float *cq = (float *)malloc(sizeof(float) * (size_t)n);
    orc_qtransform(q, raw_value, visit, prior, n, 50.0f, 0.1f, 1, 1e-8f, cq);
    int vsum = 0;
    for (int j = 0; j < n; ++j) { cq[j] = prior[j] + cq[j]; vsum += visit[j]; }
    csoftmax(cq, n);
    for (int j = 0; j < n; ++j) cq[j] = cq[j] - (float)visit[j] / (float)(1 + vsum);
    int pos = first_argmax(cq, n);
    free(cq);
    return pos;
}

static int considered_seq(int m, int S, int *out)
{ /* get_sequence_of_considered_visits (gmz:1042-1077) */
    if (m <= 1) {
        for (int i = 0; i < S; ++i) out[i] = i;
        return S;
    }
    int log2max = 0; /* ceil(log2(m)) */
    while ((1 << log2max) < m) ++log2max;
    int *visits = (int *)calloc((size_t)m, sizeof(int));
    int len = 0, num_considered = m;
    while (len < S) {
        int extra = S / (log2max * num_considered);
        if (extra < 1) extra = 1;
        for (int i = 0; i < extra; ++i) {
            for (int j = 0; j < num_considered; ++j)
                if (len < S) out[len++] = visits[j];
                else ++len; /* the reference grows past S and slices (gmz:1074-1076) */
            for (int j = 0; j < num_considered; ++j) visits[j] += 1;
        }
        num_considered = num_considered / 2 > 2 ? num_considered / 2 : 2;
    }
    free(visits);
    return S;
}

void orc_gumbel_table(int m, int S, int *out)
{ /* get_table_of_considered_visits (gmz:1079-1095): rows 0..m */
    for (int r = 0; r <= m; ++r) considered_seq(r, S, out + (size_t)r * S);
}

/* generate_gumbel (gmz:1134-1152): std::mt19937(0) + std::extreme_value_distribution<float>(0,1),
 * libstdc++ 13: generate_canonical<float,24> draws one 32-bit word, u = float(w) / 2^32 (clamped
 * below 1), value = a - b * log(-log(1 - u)). Every node seeds with 0, so all tables are prefixes
 * of one sequence. */
void orc_gumbel_noise(float scale, int shape, float *out)
{
    uint32_t mt[624];
    mt[0] = 0u;
    for (int i = 1; i < 624; ++i) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
    int idx = 624;
    for (int k = 0; k < shape; ++k) {
        if (idx >= 624) {
            for (int i = 0; i < 624; ++i) {
                uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
                mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        float u = (float)y / 4294967296.0f;
        if (u >= 1.0f) u = nextafterf(1.0f, 0.0f);
        float g = 0.0f - 1.0f * orc_logf(-orc_logf(1.0f - u));
        out[k] = scale * g;
    }
}

static int gumbel_select_root(const OrcForest *f, const Node *nd, float discount, int S, int m)
{ /* cselect_root_child (gmz:701-745) + score_considered (gmz:1097-1132) */
    int n = nd->n_legal;
    int *visit = (int *)malloc(sizeof(int) * (size_t)n);
    float *prior = (float *)malloc(sizeof(float) * (size_t)n);
    float *cq = (float *)malloc(sizeof(float) * (size_t)n);
    completed_q(nd, discount, visit, prior, cq);
    int num_considered = m < S ? m : S;
    int *seq = (int *)malloc(sizeof(int) * (size_t)S);
    considered_seq(num_considered, S, seq);
    int sim_idx = 0;
    for (int j = 0; j < n; ++j) sim_idx += visit[j];
    int considered_visit = seq[sim_idx];
    float max_logit = prior[0];
    for (int j = 1; j < n; ++j)
        if (prior[j] > max_logit) max_logit = prior[j];
    float best = -INFINITY;
    int pos = 0;
    for (int j = 0; j < n; ++j) {
        float logit = prior[j] - max_logit;
        float penalty = (visit[j] == considered_visit) ? 0.0f : -INFINITY;
        float s = f->gumbel[j] + logit + cq[j];
        s = ((-1e9f < s) ? s : -1e9f) + penalty; /* std::max(low_logit, ...) */
        if (s > best) { best = s; pos = j; }
    }
    free(visit); free(prior); free(cq); free(seq);
    return pos;
}

static int gumbel_select_interior(const Node *nd, float discount)
{ /* cselect_interior_child (gmz:747-790) */
    int n = nd->n_legal;
    int *visit = (int *)malloc(sizeof(int) * (size_t)n);
    float *prior = (float *)malloc(sizeof(float) * (size_t)n);
    float *cq = (float *)malloc(sizeof(float) * (size_t)n);
    completed_q(nd, discount, visit, prior, cq);
    for (int j = 0; j < n; ++j) cq[j] = prior[j] + cq[j];
    csoftmax(cq, n);
    int vsum = 0;
    for (int j = 0; j < n; ++j) vsum += visit[j];
    float best = -INFINITY;
    int pos = 0;
    for (int j = 0; j < n; ++j) {
        float s = cq[j] - (float)visit[j] / (float)(1 + vsum);
        if (s > best) { best = s; pos = j; }
    }
    free(visit); free(prior); free(cq);
    return pos;
}

/* ------------------------------------------------------------------ forest API */
OrcForest *orc_new(int variant, int n, int A, const int *legal_flat, const int *legal_off, int K, int D, int continuous)
{ /* CRoots::CRoots (mz:301-317; sez:625-673) */
    OrcForest *f = (OrcForest *)calloc(1, sizeof(OrcForest));
    f->variant = variant; f->n = n; f->A = A; f->K = K; f->D = D; f->continuous = continuous;
    f->C = (variant == ORC_SMZ) ? K : 0; /* smz: K carries chance_space_size (smz:327-347) */
    f->W = (f->C > A) ? f->C : A;
    f->roots = (Node *)malloc(sizeof(Node) * (size_t)n);
    f->mm = (MinMax *)malloc(sizeof(MinMax) * (size_t)n);
    f->paths = (Node ***)calloc((size_t)n, sizeof(Node **));
    f->path_len = (int *)calloc((size_t)n, sizeof(int));
    f->rng_mode = 1;
    for (int i = 0; i < n; ++i) {
        node_init(&f->roots[i], 0.0f);
        mm_init(&f->mm[i]);
        if (variant != ORC_SEZ && legal_off) {
            int m = legal_off[i + 1] - legal_off[i];
            if (m > 0) {
                f->roots[i].legal = (int *)malloc(sizeof(int) * (size_t)m);
                memcpy(f->roots[i].legal, legal_flat + legal_off[i], sizeof(int) * (size_t)m);
                f->roots[i].n_legal = m;
            }
        }
    }
    if (variant == ORC_GMZ) {
        f->gumbel = (float *)malloc(sizeof(float) * (size_t)(A > 0 ? A : 1));
        orc_gumbel_noise(10.0f, A, f->gumbel); /* gmz:87-89: gumbel_scale=10, gumbel_rng=0 */
    }
    return f;
}

void orc_free(OrcForest *f)
{
    for (int i = 0; i < f->n; ++i) { node_free(&f->roots[i]); free(f->paths[i]); }
    free(f->roots); free(f->mm); free(f->paths); free(f->path_len); free(f->gumbel); free(f);
}

void orc_set_delta(OrcForest *f, float d) { for (int i = 0; i < f->n; ++i) f->mm[i].value_delta_max = d; }

void orc_set_rng(OrcForest *f, int mode, uint64_t seed, uint32_t root_offset)
{
    f->rng_mode = mode; f->rng_seed = seed; f->root_offset = root_offset;
}

void orc_prepare(OrcForest *f, int with_noise, float noise_w, const float *noise_flat, const int *noise_off,
                 const float *rewards, const float *values, const float *logits, const int *to_play,
                 const float *sampled_actions)
{ /* CRoots::prepare / prepare_no_noise (mz:321-358; gmz:418-455; sez:677-715) */
    for (int i = 0; i < f->n; ++i) {
        Node *r = &f->roots[i];
        if (f->variant == ORC_SEZ)
            node_expand_sampled(r, to_play[i], 0, i, rewards[i], sampled_actions + (size_t)i * f->K * f->D, f->K, f->D);
        else if (f->variant == ORC_SMZ) /* smz:349-386; logits rows have stride W */
            node_expand_smz(r, to_play[i], 0, i, rewards[i], logits + (size_t)i * f->W, f->A, f->C);
        else
            node_expand(r, to_play[i], 0, i, rewards[i], values ? values[i] : 0.0f, logits + (size_t)i * f->A, f->A);
        if (with_noise && f->variant != ORC_SEZ) node_add_noise(r, noise_w, noise_flat + noise_off[i]);
        r->visit_count += 1;
    }
}

void orc_traverse(OrcForest *f, int pb_c_base, float pb_c_init, float discount, int num_simulations,
                  int max_num_considered_actions, const int *to_play, int *ix, int *iy, int *last_action,
                  float *last_action_f, int *search_len, int *vtp, int *path_actions, int path_stride)
{ /* cbatch_traverse (mz:755-824; ez:886-963; gmz:834-898; sez:1110-1187) */
    float parent_q = 0.0f; /* function-scope in the reference: carried across roots (mz:773,794) */
    int largest = to_play[0];
    for (int i = 1; i < f->n; ++i)
        if (to_play[i] > largest) largest = to_play[i];
    int players = (largest == -1) ? 1 : 2; /* batch-global (mz:776-781) */
    int last = -1;
    for (int i = 0; i < f->n; ++i) {
        Node *node = &f->roots[i];
        int is_root = 1, len = 0, cap = 8, v = to_play[i];
        const float *last_f = NULL;
        Node **path = (Node **)malloc(sizeof(Node *) * (size_t)cap);
        int plen = 0;
        path[plen++] = node;
        if (f->rng_mode == 2) parent_q = 0.0f; /* per-root driving == one reference call per root */
        while (node->n_children > 0) {
            int pos;
            if (f->variant == ORC_GMZ) {
                pos = is_root ? gumbel_select_root(f, node, discount, num_simulations, max_num_considered_actions)
                              : gumbel_select_interior(node, discount);
                is_root = 0;
            } else {
                float mean_q = compute_mean_q(f, node, is_root, parent_q, discount);
                is_root = 0;
                parent_q = mean_q;
                pos = select_child(f, node, &f->mm[i], pb_c_base, pb_c_init, discount, mean_q, players, (uint32_t)i,
                                   (uint32_t)len);
                if (pos < 0) pos = 0;
                if (players > 1) v = (v == 1) ? 2 : 1; /* mz:797-804 */
            }
            int action = (f->variant == ORC_SEZ) ? pos : node->legal[pos];
            node->best_action = action;
            if (f->variant == ORC_SEZ) last_f = node->actions + (size_t)pos * f->D;
            if (path_actions && len < path_stride) path_actions[(size_t)i * path_stride + len] = action;
            node = &node->children[pos];
            last = action;
            if (plen == cap) { cap *= 2; path = (Node **)realloc(path, sizeof(Node *) * (size_t)cap); }
            path[plen++] = node;
            len += 1;
        }
        Node *parent = path[plen - 2];
        ix[i] = parent->latent_index;
        iy[i] = parent->batch_index;
        if (last_action) last_action[i] = last;
        if (last_action_f && last_f) memcpy(last_action_f + (size_t)i * f->D, last_f, sizeof(float) * (size_t)f->D);
        search_len[i] = len;
        vtp[i] = v;
        free(f->paths[i]);
        f->paths[i] = path;
        f->path_len[i] = plen;
    }
    f->sim += 1;
}

/* Stochastic MuZero cbatch_traverse (smz:713-787) + cselect_child (smz:596-660): decision levels are pUCT with the
 * pinned tie-break stream (its position counts rand() calls, i.e. decision levels); chance levels take the outcome in
 * forced[i * forced_stride + depth] when that is >= 0 (outcomes read back from the compiled reference), else draw
 * it from the children's priors with orc_chance_bits / orc_discrete_draw.  compute_mean_q runs on both kinds. */
void orc_traverse_chance(OrcForest *f, int pb_c_base, float pb_c_init, float discount, const int *to_play,
                         const int *forced, int forced_stride, int *ix, int *iy, int *last_action, int *search_len,
                         int *vtp, int *leaf_is_chance, int *path_actions, int path_stride)
{
    float parent_q = 0.0f;
    int largest = to_play[0];
    for (int i = 1; i < f->n; ++i)
        if (to_play[i] > largest) largest = to_play[i];
    int players = (largest == -1) ? 1 : 2;
    int last = -1;
    for (int i = 0; i < f->n; ++i) {
        Node *node = &f->roots[i];
        int is_root = 1, len = 0, cap = 8, v = to_play[i];
        Node **path = (Node **)malloc(sizeof(Node *) * (size_t)cap);
        int plen = 0;
        path[plen++] = node;
        if (f->rng_mode == 2) parent_q = 0.0f;
        while (node->n_children > 0) {
            float mean_q = compute_mean_q(f, node, is_root, parent_q, discount);
            is_root = 0;
            parent_q = mean_q;
            int pos;
            if (node->is_chance) {
                int fo = (forced && len < forced_stride) ? forced[(size_t)i * forced_stride + len] : -1;
                if (fo >= 0) {
                    pos = fo;
                } else {
                    uint64_t bits = orc_chance_bits(f->rng_seed, f->sim, f->root_offset + (uint32_t)i, (uint32_t)len);
                    pos = chance_outcome(node, (double)(bits >> 11) * (1.0 / 9007199254740992.0));
                }
            } else {
                pos = select_child(f, node, &f->mm[i], pb_c_base, pb_c_init, discount, mean_q, players, (uint32_t)i,
                                   (uint32_t)(len >> 1));
                if (pos < 0) pos = 0;
            }
            if (players > 1) v = (v == 1) ? 2 : 1; /* smz:756-763: flips on chance levels too */
            int action = node->legal[pos];
            node->best_action = action;
            if (path_actions && len < path_stride) path_actions[(size_t)i * path_stride + len] = action;
            node = &node->children[pos];
            last = action;
            if (plen == cap) { cap *= 2; path = (Node **)realloc(path, sizeof(Node *) * (size_t)cap); }
            path[plen++] = node;
            len += 1;
        }
        Node *parent = path[plen - 2];
        ix[i] = parent->latent_index;
        iy[i] = parent->batch_index;
        last_action[i] = last;
        search_len[i] = len;
        vtp[i] = v;
        if (leaf_is_chance) leaf_is_chance[i] = node->is_chance; /* smz:782 */
        free(f->paths[i]);
        f->paths[i] = path;
        f->path_len[i] = plen;
    }
    f->sim += 1;
}

static void backpropagate(const OrcForest *f, Node **path, int plen, MinMax *mm, int to_play, float value, float discount)
{ /* cbackpropagate (mz:419-478; ez:482-575; gmz:605-631; sez:860-945) */
    float bootstrap = value;
    int two = (to_play != -1) && f->variant != ORC_GMZ; /* gmz asserts to_play == -1, compiled out */
    for (int i = plen - 1; i >= 0; --i) {
        Node *node = path[i];
        if (!two || node->to_play == to_play) node->value_sum += bootstrap;
        else node->value_sum += -bootstrap;
        node->visit_count += 1;
        if (is_vp(f)) {
            float parent_vp = 0.0f;
            int is_reset = 0;
            if (i >= 1) { parent_vp = path[i - 1]->reward; is_reset = path[i - 1]->is_reset; }
            float r = node->reward - parent_vp;
            mm_update(mm, r + discount * node_value(node)); /* ez:516,558: pre-reset reward, value not negated */
            if (is_reset == 1) r = node->reward;
            if (two && node->to_play == to_play) bootstrap = -r + discount * bootstrap;
            else bootstrap = r + discount * bootstrap;
        } else {
            float r = node->reward;
            if (two) mm_update(mm, r + discount * -node_value(node)); /* mz:470 */
            else mm_update(mm, r + discount * node_value(node));
            if (two && node->to_play == to_play) bootstrap = -r + discount * bootstrap;
            else bootstrap = r + discount * bootstrap;
        }
    }
}

void orc_backprop(OrcForest *f, int latent_index, float discount, const float *rewards, const float *values,
                  const float *logits, const int *is_reset, const int *to_play, const float *sampled_actions)
{ /* cbatch_backpropagate (mz:480-500; ez:577-601; gmz:633-652; sez:947-971) */
    for (int i = 0; i < f->n; ++i) {
        Node *leaf = f->paths[i][f->path_len[i] - 1];
        if (f->variant == ORC_SEZ)
            node_expand_sampled(leaf, to_play[i], latent_index, i, rewards[i], sampled_actions + (size_t)i * f->K * f->D,
                                f->K, f->D);
        else if (f->variant == ORC_SMZ) /* smz:559-593: the driver's two calls cover every tree exactly once */
            node_expand_smz(leaf, to_play[i], latent_index, i, rewards[i], logits + (size_t)i * f->W, f->A, f->C);
        else
            node_expand(leaf, to_play[i], latent_index, i, rewards[i], values[i], logits + (size_t)i * f->A, f->A);
        if (is_vp(f)) leaf->is_reset = is_reset ? is_reset[i] : 0;
        backpropagate(f, f->paths[i], f->path_len[i], &f->mm[i], to_play[i], values[i], discount);
    }
}

/* cbatch_traverse_with_reuse (mz:826-928; ez:965-1073), MuZero / EfficientZero trees only. */
void orc_traverse_reuse(OrcForest *f, int pb_c_base, float pb_c_init, float discount, const int *to_play,
                        const int *true_action, const float *reuse_value, int *ix, int *iy, int *last_action,
                        int *search_len, int *vtp)
{
    float parent_q = 0.0f;
    int largest = to_play[0];
    for (int i = 1; i < f->n; ++i)
        if (to_play[i] > largest) largest = to_play[i];
    int players = (largest == -1) ? 1 : 2;
    int last = -1;
    for (int i = 0; i < f->n; ++i) {
        Node *node = &f->roots[i];
        int is_root = 1, len = 0, cap = 8, v = to_play[i];
        Node **path = (Node **)malloc(sizeof(Node *) * (size_t)cap);
        int plen = 0;
        path[plen++] = node;
        if (f->rng_mode == 2) parent_q = 0.0f;
        while (node->n_children > 0) {
            float mean_q = compute_mean_q(f, node, is_root, parent_q, discount);
            parent_q = mean_q;
            int pos;
            if (is_root)
                pos = select_root_child(f, node, &f->mm[i], pb_c_base, pb_c_init, discount, mean_q, players,
                                        true_action[i], reuse_value[i], (uint32_t)i, (uint32_t)len);
            else
                pos = select_child(f, node, &f->mm[i], pb_c_base, pb_c_init, discount, mean_q, players, (uint32_t)i,
                                   (uint32_t)len);
            if (pos < 0) pos = 0;
            if (players > 1) v = (v == 1) ? 2 : 1;
            int action = node->legal[pos];
            node->best_action = action;
            node = &node->children[pos];
            last = action;
            if (plen == cap) { cap *= 2; path = (Node **)realloc(path, sizeof(Node *) * (size_t)cap); }
            path[plen++] = node;
            len += 1;
            if (is_root && action == true_action[i]) break; /* mz:895-898: stop on the trajectory's own action */
            is_root = 0;
        }
        if (node->n_children > 0) { /* already expanded: nothing to infer (mz:904-913) */
            ix[i] = -1;
            iy[i] = i;
        } else {
            Node *parent = path[plen - 2];
            ix[i] = parent->latent_index;
            iy[i] = parent->batch_index;
        }
        last_action[i] = last;
        search_len[i] = len;
        vtp[i] = v;
        free(f->paths[i]);
        f->paths[i] = path;
        f->path_len[i] = plen;
    }
    f->sim += 1;
}

/* cbatch_backpropagate_with_reuse (mz:502-549; ez:603-644).  rewards / values / logits are COMPACTED over the trees
 * that needed inference (count_b); no_inference_lst and reuse_lst are ascending tree indices terminated by -1;
 * reuse_value and is_reset are indexed by tree. */
void orc_backprop_reuse(OrcForest *f, int latent_index, float discount, const float *rewards, const float *values,
                        const float *logits, const int *is_reset, const int *to_play, const int *no_inference_lst,
                        const int *reuse_lst, const float *reuse_value)
{
    int count_a = 0, count_b = 0, count_c = 0;
    float value_propagate = 0.0f;
    for (int i = 0; i < f->n; ++i) {
        Node *leaf = f->paths[i][f->path_len[i] - 1];
        if (i == no_inference_lst[count_a]) {
            count_a += 1;
            value_propagate = reuse_value[i];
        } else {
            node_expand(leaf, to_play[i], latent_index, count_b, rewards[count_b], values[count_b],
                        logits + (size_t)count_b * f->A, f->A);
            if (i == reuse_lst[count_c]) {
                value_propagate = reuse_value[i];
                count_c += 1;
            } else {
                value_propagate = values[count_b];
            }
            count_b += 1;
        }
        if (is_vp(f)) leaf->is_reset = is_reset ? is_reset[i] : 0; /* ez:640 */
        backpropagate(f, f->paths[i], f->path_len[i], &f->mm[i], to_play[i], value_propagate, discount);
    }
}

void orc_distributions(OrcForest *f, int *dist, int *len)
{ /* get_distributions (mz:259-277,387-403): legal order, length n_legal; [] when not expanded */
    int W = (f->variant == ORC_SEZ) ? f->K : f->W;
    for (int i = 0; i < f->n; ++i) {
        Node *r = &f->roots[i];
        int m = r->n_children > 0 ? r->n_legal : 0;
        len[i] = m;
        for (int j = 0; j < W; ++j) dist[(size_t)i * W + j] = j < m ? r->children[j].visit_count : -1;
    }
}

void orc_values(OrcForest *f, float *out) { for (int i = 0; i < f->n; ++i) out[i] = node_value(&f->roots[i]); }

void orc_minmax(OrcForest *f, float *mn, float *mx)
{
    for (int i = 0; i < f->n; ++i) { mn[i] = f->mm[i].minimum; mx[i] = f->mm[i].maximum; }
}

void orc_gumbel_outputs(OrcForest *f, float discount, float *policies, float *children_values)
{ /* get_policy / get_children_value (gmz:309-385) */
    int A = f->A;
    for (int i = 0; i < f->n; ++i) {
        Node *r = &f->roots[i];
        int n = r->n_legal;
        int *visit = (int *)malloc(sizeof(int) * (size_t)n);
        float *prior = (float *)malloc(sizeof(float) * (size_t)n);
        float *cq = (float *)malloc(sizeof(float) * (size_t)n);
        completed_q(r, discount, visit, prior, cq);
        float *p = policies + (size_t)i * A, *v = children_values + (size_t)i * A;
        for (int a = 0; a < A; ++a) { p[a] = -INFINITY; v[a] = -INFINITY; }
        for (int j = 0; j < n; ++j) { v[r->legal[j]] = cq[j]; p[r->legal[j]] = prior[j] + cq[j]; }
        csoftmax(p, A);
        free(visit); free(prior); free(cq);
    }
}

void orc_dump_tree(OrcForest *f, int i, int E, int *visit, float *value_sum, float *prior, float *reward,
                   int *child_e, int *hdr, float *raw_value)
{ /* same layout as ref_dump_tree in oracle/ref_harness.cpp */
    int W = (f->variant == ORC_SEZ) ? f->K : f->W;
    int cap = 64, top = 0;
    Node **stack = (Node **)malloc(sizeof(Node *) * (size_t)cap);
    int *pes = (int *)malloc(sizeof(int) * (size_t)cap);
    stack[top] = &f->roots[i]; pes[top++] = -1;
    while (top > 0) {
        Node *nd = stack[--top];
        int pe = pes[top];
        int e = nd->latent_index;
        if (e < 0 || e >= E) continue;
        hdr[e * 4 + 0] = nd->n_legal;
        hdr[e * 4 + 1] = nd->to_play;
        hdr[e * 4 + 2] = is_vp(f) ? nd->is_reset : (f->variant == ORC_SMZ ? nd->is_chance : 0);
        hdr[e * 4 + 3] = pe;
        if (raw_value && f->variant == ORC_GMZ) raw_value[e] = nd->raw_value;
        for (int j = 0; j < nd->n_legal; ++j) {
            Node *c = &nd->children[j];
            size_t o = (size_t)e * W + j;
            visit[o] = c->visit_count;
            value_sum[o] = c->value_sum;
            prior[o] = c->prior;
            reward[o] = c->n_children > 0 ? c->reward : 0.0f;
            child_e[o] = c->n_children > 0 ? c->latent_index : -1;
            if (c->n_children > 0) {
                if (top == cap) {
                    cap *= 2;
                    stack = (Node **)realloc(stack, sizeof(Node *) * (size_t)cap);
                    pes = (int *)realloc(pes, sizeof(int) * (size_t)cap);
                }
                stack[top] = c; pes[top++] = e;
            }
        }
    }
    free(stack); free(pes);
}

/* CRoots::get_trajectories -> CNode::get_trajectory (mz:237-257, 369-385): follow best_action (an action id, set by
 * every traverse that selected from the node, mz:806) from the root until a node that was never selected from.
 * out [n, stride] (entries beyond stride dropped), len[i] = the true length. */
void orc_trajectories(OrcForest *f, int *out, int stride, int *len)
{
    for (int i = 0; i < f->n; ++i) {
        Node *nd = &f->roots[i];
        int d = 0;
        while (nd->best_action >= 0) {
            if (d < stride) out[(size_t)i * stride + d] = nd->best_action;
            d += 1;
            Node *next = NULL; /* get_child(action): the child stored under that action id (mz:279-288) */
            for (int j = 0; j < nd->n_children; ++j)
                if (nd->legal[j] == nd->best_action) { next = &nd->children[j]; break; }
            if (!next) break;
            nd = next;
        }
        len[i] = d;
    }
}

void orc_root_stats(OrcForest *f, int *visit, float *value_sum, float *reward)
{
    for (int i = 0; i < f->n; ++i) {
        visit[i] = f->roots[i].visit_count;
        value_sum[i] = f->roots[i].value_sum;
        reward[i] = f->roots[i].reward;
    }
}

void orc_sampled_actions(OrcForest *f, float *out)
{ /* get_sampled_actions (sez:759-787) */
    for (int i = 0; i < f->n; ++i)
        memcpy(out + (size_t)i * f->K * f->D, f->roots[i].actions, sizeof(float) * (size_t)f->K * f->D);
}
