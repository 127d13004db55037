// TEST INFRASTRUCTURE ONLY (oracle/): C-callable harness around the UNMODIFIED Stochastic-MuZero reference tree
// (/root/reference/lzero/mcts/ctree/ctree_stochastic_muzero/lib/cnode.cpp), built by oracle/build.py into
// oracle/_ref/libref_smz.so with the flags of the reference's own Cython build (-O2 -DNDEBUG).  The reference source
// is #included from where it lies; nothing is copied or edited.
//
// The tree alternates decision nodes (pUCT over the A actions) and chance nodes (an outcome drawn from the node's
// priors by std::discrete_distribution over an mt19937 seeded from std::random_device, cnode.cpp:610-630), so the
// chance draws cannot be dictated from outside.  They do not have to be: the harness lets the reference draw and
// reports the action taken at every level of every search path (ref_traverse's path_actions).  The oracle and the
// CUDA path are then handed the outcomes of the chance levels, and everything else - pUCT on the decision levels,
// tie breaking, the alternating expansion widths, reward zeroing on chance nodes, back-propagation, MinMax - is
// compared bit for bit.  The arithmetic of the draw itself (normalise, partial sums, lower_bound) is pinned
// separately by ref_discrete below, which runs libstdc++'s discrete_distribution under a scripted engine.
//
// Same two driving modes as ref_harness.cpp: per_root = 0 (one CRoots(N)) and per_root = 1 (N CRoots(1) objects, so
// that oracle/rng_pin.c can key the tie-break stream by (sim, root, k-th rand() call)).
#include <algorithm>
#include <cassert>
#include <chrono>
#include <limits>
#include <numeric>
#include <random>
#include <stack>
#include <string>

#include "lib/cnode.cpp"
#include "../common_lib/cminimax.cpp"

#include <cstdint>
#include <cstring>

extern "C" void ref_rng_set(uint32_t sim, uint32_t root);

namespace {

struct Ref {
    int n = 0, A = 0, C = 0, per_root = 0, sim = 0;
    std::vector<tree::CRoots *> roots;
    std::vector<tools::CMinMaxStatsList *> mm;
    std::vector<tree::CSearchResults *> res;
    int groups() const { return per_root ? n : 1; }
    int gsize() const { return per_root ? 1 : n; }
    int W() const { return A > C ? A : C; }
    tree::CNode *root(int i) { return per_root ? &roots[i]->roots[0] : &roots[0]->roots[i]; }
};

// a uniform random bit generator that hands out two scripted 32-bit words (what generate_canonical<double, 53>
// consumes from an mt19937) and then repeats the last one
struct Scripted {
    typedef uint32_t result_type;
    uint32_t w[2];
    int k = 0;
    static constexpr result_type min() { return 0u; }
    static constexpr result_type max() { return 0xffffffffu; }
    result_type operator()() { uint32_t v = w[k < 2 ? k : 1]; ++k; return v; }
};

} // namespace

extern "C" {

// Roots(root_num, legal_actions_list, chance_space_size) (stochastic_mz_tree.pyx:31-33 -> cnode.cpp:327-347)
void *ref_new(int n, int A, int C, const int *legal_flat, const int *legal_off, int per_root)
{
    Ref *r = new Ref();
    r->n = n; r->A = A; r->C = C; r->per_root = per_root;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<int> > legal;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            legal.push_back(std::vector<int>(legal_flat + legal_off[i], legal_flat + legal_off[i + 1]));
        }
        r->roots.push_back(new tree::CRoots(gs, legal, C));
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

// prepare / prepare_no_noise (cnode.cpp:349-386).  logits [n, A]; noises: CSR by legal position.
void ref_prepare(void *h, int with_noise, float noise_w, const float *noise_flat, const int *noise_off,
                 const float *rewards, const float *logits, const int *to_play)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<float> > nz, pol;
        std::vector<float> rw;
        std::vector<int> tp;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            if (with_noise) nz.push_back(std::vector<float>(noise_flat + noise_off[i], noise_flat + noise_off[i + 1]));
            pol.push_back(std::vector<float>(logits + (size_t)i * r->A, logits + (size_t)(i + 1) * r->A));
            rw.push_back(rewards[i]);
            tp.push_back(to_play[i]);
        }
        if (with_noise) r->roots[g]->prepare(noise_w, nz, rw, pol, tp);
        else r->roots[g]->prepare_no_noise(rw, pol, tp);
    }
}

// One cbatch_traverse (cnode.cpp:713-787).  path_actions [n, path_stride]: the action (decision levels) or outcome
// (chance levels) taken at each depth of the path, length search_len[i].
void ref_traverse(void *h, int pb_c_base, float pb_c_init, float discount, const int *to_play, int *ix, int *iy,
                  int *last_action, int *search_len, int *vtp, int *leaf_is_chance, int *path_actions, int path_stride)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        delete r->res[g];
        r->res[g] = new tree::CSearchResults(gs);
        std::vector<int> tp(to_play + g * gs, to_play + (g + 1) * gs);
        ref_rng_set((uint32_t)r->sim, (uint32_t)g);
        tree::cbatch_traverse(r->roots[g], pb_c_base, pb_c_init, discount, r->mm[g], *r->res[g], tp);
        tree::CSearchResults &s = *r->res[g];
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            ix[i] = s.latent_state_index_in_search_path[k];
            iy[i] = r->per_root ? i : s.latent_state_index_in_batch[k];
            last_action[i] = s.last_actions[k];
            search_len[i] = s.search_lens[k];
            vtp[i] = s.virtual_to_play_batchs[k];
            leaf_is_chance[i] = s.leaf_node_is_chance[k] ? 1 : 0;
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

// The two cbatch_backpropagate calls of one simulation as the driver issues them (mcts_ctree_stochastic.py:232-249 ->
// cnode.cpp:559-593): first the leaves that are chance nodes, then the decision leaves, each with its model outputs
// compacted and the tree indices in leaf_idx_list.  logits [n, W]: row i holds C chance logits when leaf i is a chance
// node, else A policy logits.
void ref_backprop(void *h, int latent_index, float discount, const float *rewards, const float *values,
                  const float *logits, int W, const int *to_play)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        tree::CSearchResults &s = *r->res[g];
        std::vector<int> tp;
        std::vector<bool> child_is_chance;
        for (int k = 0; k < gs; ++k) {
            tp.push_back(to_play[g * gs + k]);
            child_is_chance.push_back(s.leaf_node_is_chance[k]); // child_is_chance_batch (:196); ignored by expand (:115-122)
        }
        for (int pass = 1; pass >= 0; --pass) { // chance leaves first (:232), then decision leaves (:241)
            std::vector<std::vector<float> > pol;
            std::vector<float> rw, vals;
            std::vector<int> idx;
            for (int k = 0; k < gs; ++k) {
                if ((s.leaf_node_is_chance[k] ? 1 : 0) != pass) continue;
                int i = g * gs + k;
                int width = pass ? r->C : r->A;
                pol.push_back(std::vector<float>(logits + (size_t)i * W, logits + (size_t)i * W + width));
                rw.push_back(rewards[i]);
                vals.push_back(values[i]);
                idx.push_back(k);
            }
            if (idx.empty()) continue;
            tree::cbatch_backpropagate(latent_index, discount, rw, vals, pol, r->mm[g], s, tp, child_is_chance, idx);
        }
        if (r->per_root) s.nodes[0]->batch_index = g; // cnode.cpp:588 stamps the position in the call
    }
}

void ref_distributions(void *h, int *dist, int *len)
{
    Ref *r = (Ref *)h;
    int W = r->W();
    for (int i = 0; i < r->n; ++i) {
        std::vector<int> d = r->root(i)->get_children_distribution();
        len[i] = (int)d.size();
        for (int j = 0; j < W; ++j) dist[(size_t)i * W + j] = j < (int)d.size() ? d[j] : -1;
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

// Whole-tree dump keyed by current_latent_state_index e, children in list order at [e*W + j], W = max(A, C);
// hdr[e*4 + {0,1,2,3}] = {n_children, to_play, is_chance, parent's e} (see ref_harness.cpp).
void ref_dump_tree(void *h, int i, int E, int *visit, float *value_sum, float *prior, float *reward, int *child_e,
                   int *hdr, float *raw_value)
{
    (void)raw_value;
    Ref *r = (Ref *)h;
    int W = r->W();
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
        hdr[e * 4 + 2] = nd->is_chance ? 1 : 0;
        hdr[e * 4 + 3] = pe;
        int j = 0;
        for (int a : nd->legal_actions) {
            tree::CNode *c = nd->get_child(a);
            size_t o = (size_t)e * W + j;
            visit[o] = c->visit_count;
            value_sum[o] = c->value_sum;
            prior[o] = c->prior;
            reward[o] = c->expanded() ? c->reward : 0.0f;
            child_e[o] = c->expanded() ? c->current_latent_state_index : -1;
            if (c->expanded()) stack.push_back(std::make_pair(c, e));
            ++j;
        }
    }
}

void ref_root_stats(void *h, int *visit, float *value_sum, float *reward)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) {
        visit[i] = r->root(i)->visit_count;
        value_sum[i] = r->root(i)->value_sum;
        reward[i] = r->root(i)->reward;
    }
}

// What cselect_child does at a chance node (cnode.cpp:617-628) with the engine's two 32-bit outputs scripted:
// std::discrete_distribution<>(probs)(engine).  Pins the draw's arithmetic for a given uniform.
int ref_discrete(const double *probs, int n, uint32_t w0, uint32_t w1)
{
    std::vector<double> p(probs, probs + n);
    std::discrete_distribution<> dist(p.begin(), p.end());
    Scripted eng;
    eng.w[0] = w0; eng.w[1] = w1;
    return dist(eng);
}

} // extern "C"
