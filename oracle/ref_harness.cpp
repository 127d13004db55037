// TEST INFRASTRUCTURE ONLY (oracle/): C-callable harness around the UNMODIFIED reference ctree.
//
// Built by oracle/build.py into oracle/_ref/libref_{mz,ez,gmz}.so with
//     g++ -O2 -DNDEBUG -DREF_{MZ,EZ,GMZ} -I/root/reference/lzero/mcts/ctree/ctree_<variant> ...
// (same optimisation flags as the reference's own setup.py:53-68 Cython build: -O2 -DNDEBUG, no
// -march, no -ffast-math).  The reference sources are #included from where they lie under
// /root/reference; nothing is copied.  This file only marshals flat arrays <-> std::vector (the job
// the reference's .pyx files do, e.g. ctree_muzero/mz_tree.pyx:74-100) and walks public members
// of tree::CNode (cnode.h:21-46) for introspection.
//
// Two driving modes:
//   per_root = 0 : one CRoots(N); cbatch_* called exactly as mz_tree.pyx does.
//   per_root = 1 : N CRoots(1) objects, cbatch_* called root by root, so that the pinned rand()
//                  (oracle/rng_pin.c) can key the tie-break stream by (sim, root, depth).
// Headers the reference TU relies on its includer (the Cython-generated .cpp) to have pulled in.
#include <algorithm>
#include <cassert>
#include <chrono>
#include <limits>
#include <numeric>
#include <random>
#include <string>

#include "lib/cnode.cpp"
#include "../common_lib/cminimax.cpp"

#include <cstdint>
#include <cstring>

extern "C" void ref_rng_set(uint32_t sim, uint32_t root);

namespace {

struct Ref {
    int n = 0, A = 0, per_root = 0, sim = 0;
    std::vector<tree::CRoots *> roots;
    std::vector<tools::CMinMaxStatsList *> mm;
    std::vector<tree::CSearchResults *> res;
    int groups() const { return per_root ? n : 1; }
    int gsize() const { return per_root ? 1 : n; }
    tree::CNode *root(int i) { return per_root ? &roots[i]->roots[0] : &roots[0]->roots[i]; }
};

#if defined(REF_EZ)
inline float node_reward(tree::CNode *c) { return c->value_prefix; }
#else
inline float node_reward(tree::CNode *c) { return c->reward; }
#endif

} // namespace

extern "C" {

// legal_flat/legal_off: CSR of per-root legal action lists (mz_tree.pyx:30-32 -> cnode.cpp:301-317).
void *ref_new(int n, int A, const int *legal_flat, const int *legal_off, int per_root)
{
    Ref *r = new Ref();
    r->n = n; r->A = A; r->per_root = per_root;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<int> > legal;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            legal.push_back(std::vector<int>(legal_flat + legal_off[i], legal_flat + legal_off[i + 1]));
        }
        r->roots.push_back(new tree::CRoots(gs, legal));
        r->mm.push_back(new tools::CMinMaxStatsList(gs));
        r->res.push_back(nullptr);
    }
    return r;
}

void ref_free(void *h)
{
    Ref *r = (Ref *)h;
    for (auto p : r->roots) delete p;
    for (auto p : r->mm) delete p;
    for (auto p : r->res) delete p;
    delete r;
}

void ref_set_delta(void *h, float d)
{
    Ref *r = (Ref *)h;
    for (auto p : r->mm) p->set_delta(d);
}

// noises: CSR by legal position (cnode.cpp:159-166); values only used by the Gumbel tree
// (ctree_gumbel_muzero/lib/cnode.cpp:418-437).  with_noise=0 -> prepare_no_noise.
void ref_prepare(void *h, int with_noise, float noise_w, const float *noise_flat, const int *noise_off,
                 const float *rewards, const float *values, const float *logits, const int *to_play)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<float> > nz, pol;
        std::vector<float> rw, vals;
        std::vector<int> tp;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            if (with_noise) nz.push_back(std::vector<float>(noise_flat + noise_off[i], noise_flat + noise_off[i + 1]));
            pol.push_back(std::vector<float>(logits + (size_t)i * r->A, logits + (size_t)(i + 1) * r->A));
            rw.push_back(rewards[i]);
            if (values) vals.push_back(values[i]);
            tp.push_back(to_play[i]);
        }
#if defined(REF_GMZ)
        if (with_noise) r->roots[g]->prepare(noise_w, nz, rw, vals, pol, tp);
        else r->roots[g]->prepare_no_noise(rw, vals, pol, tp);
#else
        (void)vals;
        if (with_noise) r->roots[g]->prepare(noise_w, nz, rw, pol, tp);
        else r->roots[g]->prepare_no_noise(rw, pol, tp);
#endif
    }
}

// One cbatch_traverse (cnode.cpp:755-824; gumbel :834-898).  A fresh CSearchResults per call, as
// mcts_ctree.py:261 does.  to_play is copied (Cython copies the list), outputs are flat [n].
// path_actions: [n, path_stride] actions taken from the root, length search_len[i].
void ref_traverse(void *h, int pb_c_base, float pb_c_init, float discount, int num_simulations,
                  int max_num_considered_actions, const int *to_play, int *ix, int *iy, int *last_action,
                  int *search_len, int *vtp, int *path_actions, int path_stride)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        delete r->res[g];
        r->res[g] = new tree::CSearchResults(gs);
        std::vector<int> tp(to_play + g * gs, to_play + (g + 1) * gs);
        ref_rng_set((uint32_t)r->sim, (uint32_t)g);
#if defined(REF_GMZ)
        (void)pb_c_base; (void)pb_c_init;
        tree::cbatch_traverse(r->roots[g], num_simulations, max_num_considered_actions, discount, *r->res[g], tp);
#else
        (void)num_simulations; (void)max_num_considered_actions;
        tree::cbatch_traverse(r->roots[g], pb_c_base, pb_c_init, discount, r->mm[g], *r->res[g], tp);
#endif
        tree::CSearchResults &s = *r->res[g];
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            ix[i] = s.latent_state_index_in_search_path[k];
            iy[i] = r->per_root ? i : s.latent_state_index_in_batch[k];
            last_action[i] = s.last_actions[k];
            search_len[i] = s.search_lens[k];
            vtp[i] = s.virtual_to_play_batchs[k];
            if (path_actions) {
                std::vector<tree::CNode *> &p = s.search_paths[k];
                for (size_t d = 0; d + 1 < p.size() && (int)d < path_stride; ++d) {
                    int found = -1;
                    for (int a : p[d]->legal_actions)
                        if (p[d]->get_child(a) == p[d + 1]) { found = a; break; }
                    path_actions[(size_t)i * path_stride + d] = found;
                }
            }
        }
    }
    r->sim += 1;
}

// One cbatch_backpropagate (cnode.cpp:480-500; EZ :577-601 with is_reset; gumbel :633-652).
void ref_backprop(void *h, int latent_index, float discount, const float *rewards, const float *values,
                  const float *logits, const int *is_reset, const int *to_play)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<float> > pol;
        std::vector<float> rw, vals;
        std::vector<int> tp, rst;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            pol.push_back(std::vector<float>(logits + (size_t)i * r->A, logits + (size_t)(i + 1) * r->A));
            rw.push_back(rewards[i]);
            vals.push_back(values[i]);
            tp.push_back(to_play[i]);
            rst.push_back(is_reset ? is_reset[i] : 0);
        }
#if defined(REF_GMZ)
        tree::cbatch_back_propagate(latent_index, discount, rw, vals, pol, r->mm[g], *r->res[g], tp);
#elif defined(REF_EZ)
        tree::cbatch_backpropagate(latent_index, discount, rw, vals, pol, r->mm[g], *r->res[g], rst, tp);
#else
        tree::cbatch_backpropagate(latent_index, discount, rw, vals, pol, r->mm[g], *r->res[g], tp);
#endif
        if (r->per_root) {
            // the reference stamps batch_index = position in the call (cnode.cpp:497); restore the global one
            r->res[g]->nodes[0]->batch_index = g;
        }
    }
}

#if !defined(REF_GMZ)
// One cbatch_traverse_with_reuse (ctree_muzero cnode.cpp:826-928; ctree_efficientzero :965-1073).
void ref_traverse_reuse(void *h, int pb_c_base, float pb_c_init, float discount, const int *to_play,
                        const int *true_action, const float *reuse_value, int *ix, int *iy, int *last_action,
                        int *search_len, int *vtp)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        delete r->res[g];
        r->res[g] = new tree::CSearchResults(gs);
        std::vector<int> tp(to_play + g * gs, to_play + (g + 1) * gs);
        std::vector<int> ta(true_action + g * gs, true_action + (g + 1) * gs);
        std::vector<float> rv(reuse_value + g * gs, reuse_value + (g + 1) * gs);
        ref_rng_set((uint32_t)r->sim, (uint32_t)g);
        tree::cbatch_traverse_with_reuse(r->roots[g], pb_c_base, pb_c_init, discount, r->mm[g], *r->res[g], tp, ta, rv);
        tree::CSearchResults &s = *r->res[g];
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            ix[i] = s.latent_state_index_in_search_path[k];
            iy[i] = s.latent_state_index_in_batch[k];
            if (r->per_root && ix[i] <= 0) iy[i] = i; // roots and "no inference" entries report the tree's own index
            last_action[i] = s.last_actions[k];
            search_len[i] = s.search_lens[k];
            vtp[i] = s.virtual_to_play_batchs[k];
        }
    }
    r->sim += 1;
}

// One cbatch_backpropagate_with_reuse (ctree_muzero cnode.cpp:502-549; ctree_efficientzero :603-644).
// rewards / values / logits are compacted over the trees that infer; the two lists end with -1.
void ref_backprop_reuse(void *h, int latent_index, float discount, const float *rewards, const float *values,
                        const float *logits, const int *is_reset, const int *to_play, const int *no_inference_lst,
                        const int *reuse_lst, const float *reuse_value, int n_infer)
{
    Ref *r = (Ref *)h;
    if (!r->per_root) {
        std::vector<std::vector<float> > pol;
        std::vector<float> rw(rewards, rewards + n_infer), vals(values, values + n_infer);
        for (int k = 0; k < n_infer; ++k)
            pol.push_back(std::vector<float>(logits + (size_t)k * r->A, logits + (size_t)(k + 1) * r->A));
        std::vector<int> tp(to_play, to_play + r->n), rst, noinf, reuse;
        for (int i = 0; i < r->n; ++i) rst.push_back(is_reset ? is_reset[i] : 0);
        for (int k = 0;; ++k) { noinf.push_back(no_inference_lst[k]); if (no_inference_lst[k] == -1) break; }
        for (int k = 0;; ++k) { reuse.push_back(reuse_lst[k]); if (reuse_lst[k] == -1) break; }
        std::vector<float> rv(reuse_value, reuse_value + r->n);
#if defined(REF_EZ)
        tree::cbatch_backpropagate_with_reuse(latent_index, discount, rw, vals, pol, r->mm[0], *r->res[0], rst, tp, noinf, reuse, rv);
#else
        (void)rst;
        tree::cbatch_backpropagate_with_reuse(latent_index, discount, rw, vals, pol, r->mm[0], *r->res[0], tp, noinf, reuse, rv);
#endif
        return;
    }
    int count_a = 0, count_b = 0, count_c = 0;
    for (int i = 0; i < r->n; ++i) {
        std::vector<std::vector<float> > pol;
        std::vector<float> rw, vals, rv(1, reuse_value[i]);
        std::vector<int> tp(1, to_play[i]), rst(1, is_reset ? is_reset[i] : 0), noinf, reuse;
        bool infer = true;
        if (i == no_inference_lst[count_a]) {
            count_a += 1;
            noinf.push_back(0);
            infer = false;
        } else {
            rw.push_back(rewards[count_b]);
            vals.push_back(values[count_b]);
            pol.push_back(std::vector<float>(logits + (size_t)count_b * r->A, logits + (size_t)(count_b + 1) * r->A));
            if (i == reuse_lst[count_c]) { count_c += 1; reuse.push_back(0); }
        }
        noinf.push_back(-1);
        reuse.push_back(-1);
#if defined(REF_EZ)
        tree::cbatch_backpropagate_with_reuse(latent_index, discount, rw, vals, pol, r->mm[i], *r->res[i], rst, tp, noinf, reuse, rv);
#else
        tree::cbatch_backpropagate_with_reuse(latent_index, discount, rw, vals, pol, r->mm[i], *r->res[i], tp, noinf, reuse, rv);
#endif
        if (infer) {
            r->res[i]->nodes[0]->batch_index = count_b; // the compacted index a whole-batch call would have stamped
            count_b += 1;
        }
    }
}
#endif

// get_distributions (cnode.cpp:387-403): child visit counts in legal order; dist is [n, A] padded with -1.
void ref_distributions(void *h, int *dist, int *len)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) {
        std::vector<int> d = r->root(i)->get_children_distribution();
        len[i] = (int)d.size();
        for (int j = 0; j < r->A; ++j) dist[(size_t)i * r->A + j] = j < (int)d.size() ? d[j] : -1;
    }
}

void ref_values(void *h, float *out)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) out[i] = r->root(i)->value();
}

void ref_minmax(void *h, float *mn, float *mx)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) {
        tools::CMinMaxStats &s = r->per_root ? r->mm[i]->stats_lst[0] : r->mm[0]->stats_lst[i];
        mn[i] = s.minimum; mx[i] = s.maximum;
    }
}

#if defined(REF_GMZ)
// get_policies / get_children_values (ctree_gumbel_muzero/lib/cnode.cpp:506-541): [n, A].
void ref_gumbel_outputs(void *h, float discount, float *policies, float *children_values)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) {
        std::vector<float> p = r->root(i)->get_policy(discount, r->A);
        std::vector<float> v = r->root(i)->get_children_value(discount, r->A);
        memcpy(policies + (size_t)i * r->A, p.data(), sizeof(float) * r->A);
        memcpy(children_values + (size_t)i * r->A, v.data(), sizeof(float) * r->A);
    }
}

void ref_gumbel_table(int max_num_considered_actions, int num_simulations, int *out /*[(m+1), S]*/)
{
    std::vector<std::vector<int> > t = tree::get_table_of_considered_visits(max_num_considered_actions, num_simulations);
    for (int m = 0; m <= max_num_considered_actions; ++m)
        for (int s = 0; s < num_simulations; ++s) out[m * num_simulations + s] = t[m][s];
}

void ref_gumbel_noise(float scale, float rng, int shape, float *out)
{
    std::vector<float> g = tree::generate_gumbel(scale, rng, shape);
    for (int i = 0; i < shape; ++i) out[i] = g[i];
}

void ref_softmax(float *x, int n)
{
    std::vector<float> v(x, x + n);
    tree::csoftmax(v, n);
    for (int i = 0; i < n; ++i) x[i] = v[i];
}

// ---- the single-node helpers gmz_tree.pyx:97-149 hands to the reference's unit tests, called as they are
float ref_mixed_value(float raw_value, const float *q, const int *visit, const float *prior, int n)
{
    std::vector<float> vq(q, q + n), vp(prior, prior + n);
    std::vector<int> vv(visit, visit + n);
    return tree::compute_mixed_value(raw_value, vq, vv, vp);
}

void ref_rescale(float *x, int n, float epsilon)
{
    std::vector<float> v(x, x + n);
    tree::rescale_qvalues(v, epsilon);
    for (int i = 0; i < n; ++i) x[i] = v[i];
}

void ref_score_considered(int considered_visit, const float *gumbel, const float *logits, const float *nq, const int *visit, int n,
                          float *out)
{
    std::vector<float> s = tree::score_considered(considered_visit, std::vector<float>(gumbel, gumbel + n), std::vector<float>(logits, logits + n),
                                                  std::vector<float>(nq, nq + n), std::vector<int>(visit, visit + n));
    for (int i = 0; i < n; ++i) out[i] = s[i];
}

// One node built like CRoots builds a root (CNode(prior, legal_actions): the constructor that draws the node's gumbel
// noise - the .pyx Node wraps a default-constructed CNode whose gumbel vector is EMPTY, so select_root_child on it reads
// past the end), expanded, then handed to the three node-taking helpers.  child_visit / child_prior are the caller's
// vectors for qtransform (pqtransform_completed_by_mix_value's arguments); the selections read the node's own children.
void ref_node_helpers(int to_play, float value_prefix, float value, const float *logits, int A, float discount, int num_simulations,
                      int max_num_considered_actions, const int *child_visit, const float *child_prior, float maxvisit_init,
                      float value_scale, int rescale_values, float epsilon, float *priors_out, float *qtransform_out, int *root_action,
                      int *interior_action)
{
    std::vector<int> legal(A);
    for (int i = 0; i < A; ++i) legal[i] = i;
    tree::CNode node(0.0f, legal);
    node.expand(to_play, 0, 0, value_prefix, value, std::vector<float>(logits, logits + A));
    for (int i = 0; i < A; ++i) priors_out[i] = node.get_child(i)->prior;
    std::vector<int> cv(child_visit, child_visit + A);
    std::vector<float> cp(child_prior, child_prior + A);
    std::vector<float> t = tree::qtransform_completed_by_mix_value(&node, cv, cp, discount, maxvisit_init, value_scale, rescale_values != 0, epsilon);
    for (int i = 0; i < A; ++i) qtransform_out[i] = t[i];
    *root_action = tree::cselect_root_child(&node, discount, num_simulations, max_num_considered_actions);
    *interior_action = tree::cselect_interior_child(&node, discount);
}
#endif

// Whole-tree dump keyed by current_latent_state_index e (unique per tree, 0 = root): children of the
// node expanded as e are written in legal-action order at [e*A + j].  Unwritten entries keep the
// caller's fill.  hdr[e*4 + {0,1,2,3}] = {n_children, to_play, is_reset, parent's e}.
void ref_dump_tree(void *h, int i, int E, int *visit, float *value_sum, float *prior, float *reward,
                   int *child_e, int *hdr, float *raw_value)
{
    Ref *r = (Ref *)h;
    int A = r->A;
    std::vector<std::pair<tree::CNode *, int> > stack;
    stack.push_back(std::make_pair(r->root(i), -1));
    while (!stack.empty()) {
        tree::CNode *nd = stack.back().first;
        int pe = stack.back().second;
        stack.pop_back();
        int e = nd->current_latent_state_index;
        if (e < 0 || e >= E) continue;
        hdr[e * 4 + 0] = (int)nd->legal_actions.size();
        hdr[e * 4 + 1] = nd->to_play;
#if defined(REF_EZ)
        hdr[e * 4 + 2] = nd->is_reset;
#else
        hdr[e * 4 + 2] = 0;
#endif
        hdr[e * 4 + 3] = pe;
#if defined(REF_GMZ)
        if (raw_value) raw_value[e] = nd->raw_value;
#else
        (void)raw_value;
#endif
        int j = 0;
        for (int a : nd->legal_actions) {
            tree::CNode *c = nd->get_child(a);
            size_t o = (size_t)e * A + j;
            visit[o] = c->visit_count;
            value_sum[o] = c->value_sum;
            prior[o] = c->prior;
            reward[o] = c->expanded() ? node_reward(c) : 0.0f; // mz:61-79 leaves `reward` uninitialised until expand
            child_e[o] = c->expanded() ? c->current_latent_state_index : -1;
            if (c->expanded()) stack.push_back(std::make_pair(c, e));
            ++j;
        }
    }
}

// CRoots::get_trajectories (cnode.cpp:369-385; Gumbel :563-579): out [n, stride], len[i] = true length
void ref_trajectories(void *h, int *out, int stride, int *len)
{
    Ref *r = (Ref *)h;
    for (int g = 0; g < r->groups(); ++g) {
        std::vector<std::vector<int> > t = r->roots[g]->get_trajectories();
        for (int k = 0; k < r->gsize(); ++k) {
            int i = g * r->gsize() + k;
            len[i] = (int)t[k].size();
            for (int d = 0; d < (int)t[k].size() && d < stride; ++d) out[(size_t)i * stride + d] = t[k][d];
        }
    }
}

void ref_root_stats(void *h, int *visit, float *value_sum, float *reward)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) {
        visit[i] = r->root(i)->visit_count;
        value_sum[i] = r->root(i)->value_sum;
        reward[i] = node_reward(r->root(i));
    }
}

} // extern "C"
