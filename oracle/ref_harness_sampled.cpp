// TEST INFRASTRUCTURE ONLY (oracle/): C-callable harness around the UNMODIFIED Sampled-EfficientZero reference tree
// (/root/reference/lzero/mcts/ctree/ctree_sampled_efficientzero/lib/cnode.cpp), built by oracle/build.py into
// oracle/_ref/libref_sez.so with the flags of the reference's own Cython build (-O2 -DNDEBUG).  The reference source is
// #included from where it lies; nothing is copied or edited.
//
// The reference draws each node's K actions inside CNode::expand from a std::default_random_engine seeded with the
// wall clock (cnode.cpp:251,336), so the draws cannot be dictated from outside.  They do not have to be: this harness
// lets the reference sample as it likes and READS the K actions of every node back out of the tree
// (CNode::legal_actions, public member, cnode.h:47) right after it was expanded.  The oracle and the CUDA path are then
// handed exactly those action tables, so everything downstream of the sampler - uniform-prior pUCT over the K
// children, value-prefix differencing, is_reset, tie breaking, back-propagation, [N][D] float last_actions - is
// compared bit for bit against the real reference.  Only the sampler's distribution stays unpinned (checked
// statistically in tests/test_gpu_sampler.py).
//
// Same two driving modes as ref_harness.cpp: per_root = 0 (one CRoots(N)) and per_root = 1 (N CRoots(1) objects, so
// that oracle/rng_pin.c can key the tie-break stream by (sim, root, depth)).
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

// ---- pinned wall clock.  CNode::expand seeds its std::default_random_engine with
//   unsigned seed = std::chrono::system_clock::now().time_since_epoch().count();          (cnode.cpp:251,336)
// system_clock::now() is an out-of-line function of libstdc++; this translation unit (which #includes the reference
// source) defines it, and the library is linked with -Wl,-Bsymbolic, so the reference's calls bind HERE.  Unarmed it
// reads the real clock; armed (ref_clock_pin) the k-th call returns base + k nanoseconds, which makes the seed of the
// k-th expand() of a batched call (roots are expanded in index order, cnode.cpp:690-700, 960-967) equal to
// (unsigned)(base + k).  No reference source is edited.
#include <time.h>
namespace {
bool g_clock_pinned = false;
uint64_t g_clock_next = 0;
}
namespace std { namespace chrono { inline namespace _V2 {
system_clock::time_point system_clock::now() noexcept
{
    if (g_clock_pinned) return time_point(duration(static_cast<rep>(g_clock_next++)));
    timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    return time_point(duration(static_cast<rep>(ts.tv_sec) * 1000000000LL + ts.tv_nsec));
}
}}}
extern "C" void ref_clock_pin(int on, uint64_t next_ns) { g_clock_pinned = on != 0; g_clock_next = next_ns; }

// libstdc++'s own engine + distributions, exposed so that tests can pin the oracle's restatement of their arithmetic
// against the library itself (independent of the tree): count draws of normal_distribution<float>(mu, sigma), a fresh
// distribution object per draw exactly as cnode.cpp:262-263 constructs it, from one default_random_engine(seed);
// and count draws of uniform_real_distribution<double>(0, 1) as cnode.cpp:369,377 uses it.
extern "C" void ref_libstdcxx_normal(unsigned seed, float mu, float sigma, int count, float *out)
{
    std::default_random_engine gen(seed);
    for (int i = 0; i < count; ++i) {
        std::normal_distribution<float> d(mu, sigma);
        out[i] = d(gen);
    }
}
extern "C" void ref_libstdcxx_uniform(unsigned seed, int count, double *out)
{
    std::default_random_engine gen(seed);
    std::uniform_real_distribution<double> d(0.0, 1.0);
    for (int i = 0; i < count; ++i) out[i] = d(gen);
}

namespace {

struct Ref {
    int n = 0, A = 0, K = 0, D = 0, continuous = 0, per_root = 0, sim = 0;
    std::vector<tree::CRoots *> roots;
    std::vector<tools::CMinMaxStatsList *> mm;
    std::vector<tree::CSearchResults *> res;
    int groups() const { return per_root ? n : 1; }
    int gsize() const { return per_root ? 1 : n; }
    tree::CNode *root(int i) { return per_root ? &roots[i]->roots[0] : &roots[0]->roots[i]; }
};

// the K actions of an expanded node in child order, [K][D] (zero-filled if the node holds fewer, e.g. hash collisions)
void read_actions(tree::CNode *nd, int K, int D, float *out)
{
    memset(out, 0, sizeof(float) * (size_t)K * D);
    int j = 0;
    for (auto &a : nd->legal_actions) {
        if (j >= K) break;
        for (int d = 0; d < D && d < (int)a.value.size(); ++d) out[(size_t)j * D + d] = a.value[d];
        ++j;
    }
}

} // namespace

extern "C" {

// A = action_space_size of the env (continuous: action dimension D; discrete: number of actions), K = sampled actions.
// Roots are created the way the policies create them for "all actions legal": legal_actions_list[i] = {-1, ...}
// (sampled_efficientzero.py:845-851 -> cnode.cpp:645-655).
void *ref_new(int n, int A, int K, int continuous, int per_root)
{
    Ref *r = new Ref();
    r->n = n; r->A = A; r->K = K; r->continuous = continuous; r->per_root = per_root;
    r->D = continuous ? A : 1;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<float> > legal(gs, std::vector<float>(A, -1.0f));
        r->roots.push_back(new tree::CRoots(gs, legal, A, K, continuous != 0));
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

// prepare / prepare_no_noise (cnode.cpp:675-757).  logits: [n, W] with W = 2A (continuous: mu, sigma) or A.
// noises: [n, K] by child position.  actions_out [n, K, D]: the actions the reference sampled for each root.
void ref_prepare(void *h, int with_noise, float noise_w, const float *noises, const float *value_prefixs,
                 const float *logits, int W, const int *to_play, float *actions_out)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<float> > nz, pol;
        std::vector<float> vp;
        std::vector<int> tp;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            if (with_noise) nz.push_back(std::vector<float>(noises + (size_t)i * r->K, noises + (size_t)(i + 1) * r->K));
            pol.push_back(std::vector<float>(logits + (size_t)i * W, logits + (size_t)(i + 1) * W));
            vp.push_back(value_prefixs[i]);
            tp.push_back(to_play[i]);
        }
        if (with_noise) r->roots[g]->prepare(noise_w, nz, vp, pol, tp);
        else r->roots[g]->prepare_no_noise(vp, pol, tp);
    }
    for (int i = 0; i < r->n; ++i) read_actions(r->root(i), r->K, r->D, actions_out + (size_t)i * r->K * r->D);
}

// One cbatch_traverse (cnode.cpp:1110-1187).  last_action_f: [n, D].
void ref_traverse(void *h, int pb_c_base, float pb_c_init, float discount, const int *to_play, int *ix, int *iy,
                  float *last_action_f, int *search_len, int *vtp)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        delete r->res[g];
        r->res[g] = new tree::CSearchResults(gs);
        std::vector<int> tp(to_play + g * gs, to_play + (g + 1) * gs);
        ref_rng_set((uint32_t)r->sim, (uint32_t)g);
        tree::cbatch_traverse(r->roots[g], pb_c_base, pb_c_init, discount, r->mm[g], *r->res[g], tp, r->continuous != 0);
        tree::CSearchResults &s = *r->res[g];
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            ix[i] = s.latent_state_index_in_search_path[k];
            iy[i] = r->per_root ? i : s.latent_state_index_in_batch[k];
            for (int d = 0; d < r->D; ++d)
                last_action_f[(size_t)i * r->D + d] = d < (int)s.last_actions[k].size() ? s.last_actions[k][d] : 0.0f;
            search_len[i] = s.search_lens[k];
            vtp[i] = s.virtual_to_play_batchs[k];
        }
    }
    r->sim += 1;
}

// One cbatch_backpropagate (cnode.cpp:947-971).  actions_out [n, K, D]: what the reference sampled for the new leaves.
void ref_backprop(void *h, int latent_index, float discount, const float *value_prefixs, const float *values,
                  const float *logits, int W, const int *is_reset, const int *to_play, float *actions_out)
{
    Ref *r = (Ref *)h;
    int G = r->groups(), gs = r->gsize();
    for (int g = 0; g < G; ++g) {
        std::vector<std::vector<float> > pol;
        std::vector<float> vp, vals;
        std::vector<int> tp, rst;
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            pol.push_back(std::vector<float>(logits + (size_t)i * W, logits + (size_t)(i + 1) * W));
            vp.push_back(value_prefixs[i]);
            vals.push_back(values[i]);
            tp.push_back(to_play[i]);
            rst.push_back(is_reset ? is_reset[i] : 0);
        }
        tree::cbatch_backpropagate(latent_index, discount, vp, vals, pol, r->mm[g], *r->res[g], rst, tp);
        for (int k = 0; k < gs; ++k) {
            int i = g * gs + k;
            if (r->per_root) r->res[g]->nodes[k]->batch_index = i; // cnode.cpp:965 stamps the position in the call
            read_actions(r->res[g]->nodes[k], r->K, r->D, actions_out + (size_t)i * r->K * r->D);
        }
    }
}

// get_distributions: child visit counts in child order, [n, K] padded with -1.
void ref_distributions(void *h, int *dist, int *len)
{
    Ref *r = (Ref *)h;
    for (int i = 0; i < r->n; ++i) {
        std::vector<int> d = r->root(i)->get_children_distribution();
        len[i] = (int)d.size();
        for (int j = 0; j < r->K; ++j) dist[(size_t)i * r->K + j] = j < (int)d.size() ? d[j] : -1;
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

// Whole-tree dump keyed by current_latent_state_index e, children in child order at [e*K + j] (see ref_harness.cpp).
void ref_dump_tree(void *h, int i, int E, int *visit, float *value_sum, float *prior, float *reward, int *child_e,
                   int *hdr, float *raw_value)
{
    (void)raw_value;
    Ref *r = (Ref *)h;
    int K = r->K;
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
        hdr[e * 4 + 2] = nd->is_reset;
        hdr[e * 4 + 3] = pe;
        int j = 0;
        for (auto &a : nd->legal_actions) {
            if (j >= K) break;
            tree::CNode *c = nd->get_child(a);
            size_t o = (size_t)e * K + j;
            visit[o] = c->visit_count;
            value_sum[o] = c->value_sum;
            prior[o] = c->prior;
            reward[o] = c->expanded() ? c->value_prefix : 0.0f;
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
        reward[i] = r->root(i)->value_prefix;
    }
}

} // extern "C"
