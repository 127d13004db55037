// Selection / traverse kernels (K2, K7) with the leaf gather (K3) fused as an epilogue.
// One warp walks one tree from the root to the first unexpanded child.
#pragma once
#include "lzb_math.cuh"
#include "lzb_pool.cuh"

namespace lzb {

struct Gather {
    const char *src; // [E][N][row_bytes] pool, or nullptr: rows come from the slot table (Pool::slots)
    char *dst;       // [N][row_bytes]
    long long row_bytes;
    int zero_if_reset; // LSTM state: rows of a parent that was expanded with is_reset = 1 read as zeros (mcts_ctree.py:812-816)
};

struct SelectArgs {
    int c_base, players, sim, rng_mode, num_simulations, ng;
    float c_init, discount, delta;
    unsigned long long seed;
    const unsigned long long *seed_dev; // device word replacing `seed` (lzb_search_cfg_t::rng_seed_dev), or nullptr
    const int *to_play;
    int *ix, *iy, *search_len, *vtp;
    long long *last_action;
    float *last_action_f;
    Gather g[kMaxGathers];
    int new_e;                        // fused step: the expansion index whose slot pointers are not in the table yet, else -1
    const char *new_slot[kMaxGathers];
    // ReZero reuse traverse (cbatch_traverse_with_reuse, cnode.cpp:826-928), lane-group kernels only
    const int *true_action;    // [N] the trajectory's action at each root
    const float *reuse_value;  // [N] value found by the next state's search
    int *infer_count;          // += trees that stopped on an unexpanded leaf (need inference), or nullptr
    // Stochastic MuZero (ctree_stochastic_muzero/lib/cnode.cpp:596-660,713-787), lane-group kernels only
    const int *forced_outcome; // [N][forced_stride] outcome to take at a chance level when >= 0, or nullptr
    int forced_stride;
    int *leaf_is_chance;       // [N] or nullptr
    const unsigned long long *chance_seed; // device word replacing `seed` for the chance draws, or nullptr
    int slot_by_kind;          // stochastic tree with two heads: a node's rows live in slot row 1 if it is a chance node, else row 0
};

// dst[i,:] = rows of the leaf's parent (expansion ix) with 16-byte accesses when rows allow it.  The source is
// streamed (ld.global.cs): each row is read once per simulation and must not push the node pool out of L2.
__device__ __forceinline__ void gather_rows(const Pool &p, const SelectArgs &a, int i, int ix, int parent_reset, int lane)
{
    for (int k = 0; k < a.ng; ++k) {
        const long long rb = a.g[k].row_bytes;
        const char *s;
        if (a.g[k].src) s = a.g[k].src + ((long long)ix * p.N + i) * rb;
        else s = ((ix == a.new_e) ? pick3(a.new_slot, k) : p.slots[(size_t)k * p.E + ix]) + (long long)i * rb;
        char *d = a.g[k].dst + (long long)i * rb;
        const bool zero = a.g[k].zero_if_reset && parent_reset == 1;
        if ((rb & 15) == 0 && ((((unsigned long long)s) | ((unsigned long long)d)) & 15ull) == 0) {
            const int4 *s4 = reinterpret_cast<const int4 *>(s);
            int4 *d4 = reinterpret_cast<int4 *>(d);
            const int n4 = (int)(rb >> 4);
            if (zero) {
                for (int t = lane; t < n4; t += 32) d4[t] = make_int4(0, 0, 0, 0);
            } else {
#pragma unroll 4
                for (int t = lane; t < n4; t += 32) d4[t] = __ldcs(s4 + t);
            }
        } else {
            const int *s1 = reinterpret_cast<const int *>(s);
            int *d1 = reinterpret_cast<int *>(d);
            const int n1 = (int)(rb >> 2);
            for (int t = lane; t < n1; t += 32) d1[t] = zero ? 0 : __ldcs(s1 + t);
        }
    }
}

// Leaf bookkeeping common to every flavour (cnode.cpp:814-822).
__device__ __forceinline__ void write_leaf(const Pool &p, const SelectArgs &a, int i, int e, int pos, int action,
                                           int depth, int vtp, int parent_reset, int lane)
{
    if (lane == 0) {
        p.res[i] = make_int4(depth, e, pos, vtp);
        if (a.ix) a.ix[i] = e;
        if (a.iy) a.iy[i] = i;
        if (a.search_len) a.search_len[i] = depth;
        if (a.vtp) a.vtp[i] = vtp;
        if (a.last_action) a.last_action[i] = (long long)action;
    }
    if (p.variant == LZB_SEZ) {
        const float *act = p.actions + (((size_t)i * p.E + e) * p.A + pos) * p.D;
        if (a.last_action_f)
            for (int d = lane; d < p.D; d += 32) a.last_action_f[(size_t)i * p.D + d] = act[d];
        if (a.last_action && lane == 0) a.last_action[i] = (long long)act[0]; // discrete sampled: float(index)
    }
    gather_rows(p, a, i, e, parent_reset, lane);
}

__device__ __forceinline__ int4 shfl4(int4 v, int src)
{
    return make_int4(__shfl_sync(kFull, v.x, src), __shfl_sync(kFull, v.y, src), __shfl_sync(kFull, v.z, src),
                     __shfl_sync(kFull, v.w, src));
}

// pUCT traverse of tree i by one warp: compute_mean_q + cucb_score + cselect_child per level
// (mz cnode.cpp:169-203,551-596,655-699,755-824; ez :173-212,651-695,756-814,886-963; sez :483-522,973-1187).
// s_child / s_score: warp-private shared memory, A entries each.
template <int VAR>
__device__ __forceinline__ void select_puct_tree(const Pool &p, const SelectArgs &a, int i, int lane, int4 *s_child,
                                                 float *s_score)
{
    constexpr bool VP = (VAR == LZB_EZ || VAR == LZB_SEZ);
    const int B = p.B;
    const int4 *tree = p.tree(i);
    int4 *pa = p.path_a + (size_t)i * (p.S + 1);
    int2 *pb = p.path_b + (size_t)i * (p.S + 1);
    // independent loads, issued together with the root block below
    const float2 mm = p.minmax[i];
    const int4 rt = p.root[i];
    int vtp = a.to_play[i];
    const int players = a.players ? a.players : ((p.scratch[0] == -1) ? 1 : 2); // cnode.cpp:776-781
    const float g = a.discount;
    const float c_base = (float)a.c_base;
    const float mm_delta = mm.y - mm.x;
    const unsigned long long tie_seed = a.seed_dev ? *a.seed_dev : a.seed;

    int parent_visits = rt.x;
    float parent_reward = __int_as_float(rt.z);
    int e = 0, depth = 0, pos = 0, action = 0, parent_reset = 0;
    float parent_q = 0.0f;
    bool is_root = true;

    while (true) {
        // ---- ONE dependent read per level: record `lane` of block e (header in lane 0, children after it)
        const int4 *blk = tree + (size_t)e * B;
        const int4 first = (lane < B) ? blk[lane] : make_int4(0, 0, 0, 0);
        const int4 h = shfl4(first, 0);
        const int n = h.x;
        parent_reset = h.z;
        // ---- stage children, mean-Q over visited children in legal order (compute_mean_q)
        float total = 0.0f;
        int nvis = 0;
        for (int r0 = 0; r0 <= n; r0 += 32) {
            const int r = r0 + lane, j = r - 1;
            const bool valid = (r >= 1) && (r <= n);
            int4 c = make_int4(0, 0, 0, 0);
            if (valid) { c = (r0 == 0) ? first : blk[r]; s_child[j] = c; }
            const int vc = child_visits(c.x);
            float qsa = 0.0f;
            if (vc > 0) {
                const float rew = __int_as_float(c.w);
                float tr = rew;
                if (VP) tr = (parent_reset == 1) ? rew : rew - parent_reward;
                qsa = tr + g * (__int_as_float(c.y) / (float)vc);
            }
            unsigned vm = __ballot_sync(kFull, valid && vc > 0);
            while (vm) {
                const int src = __ffs(vm) - 1;
                vm &= vm - 1;
                total += __shfl_sync(kFull, qsa, src);
                nvis += 1;
            }
        }
        __syncwarp();
        const float mean_q = (is_root && nvis > 0) ? total / (float)nvis : (parent_q + total) / (float)(nvis + 1);

        // ---- pUCT scores (cucb_score)
        const float T = (float)(parent_visits - 1);
        const float pbc0 = glibc_logf((T + c_base + 1) / c_base) + a.c_init;
        const float sqT = sqrtf(T);
        float lmax = -INFINITY;
        for (int j = lane; j < n; j += 32) {
            const int4 c = s_child[j];
            const int vc = child_visits(c.x);
            const float pb_c = pbc0 * (sqT / (float)(vc + 1));
            float prior_score;
            if (VAR == LZB_SEZ) prior_score = pb_c / (float)n; // uniform branch, sez cnode.cpp:1078
            else prior_score = pb_c * __int_as_float(c.z);
            float vs;
            if (vc == 0) {
                vs = mean_q;
            } else {
                const float rew = __int_as_float(c.w);
                float tr = rew;
                if (VP) tr = (parent_reset == 1) ? rew : rew - parent_reward;
                const float val = __int_as_float(c.y) / (float)vc;
                vs = (players == 1) ? tr + g * val : tr + g * (-val);
            }
            if (mm_delta > 0) { // CMinMaxStats::normalize (cminimax.cpp:33-45)
                if (mm_delta < a.delta) vs = (vs - mm.x) / a.delta;
                else vs = (vs - mm.x) / mm_delta;
            }
            if (vs < 0) vs = 0;
            if (vs > 1) vs = 1;
            const float sc = prior_score + vs;
            s_score[j] = sc;
            lmax = fmaxf(lmax, sc);
        }
        __syncwarp();
        const float gmax = warp_max(lmax);

        // ---- order-dependent tie set + pinned draw (cselect_child, cnode.cpp:567-596)
        pos = 0;
        if (gmax > kFloatMin) {
            int jstar = 0x7fffffff;
            for (int j = lane; j < n; j += 32)
                if (s_score[j] == gmax) { jstar = j; break; }
            __syncwarp();
            jstar = warp_min_i(jstar);
            const float thr = gmax - 0.000001f;
            int later = 0;
            for (int j0 = 0; j0 < n; j0 += 32) {
                const int j = j0 + lane;
                later += __popc(__ballot_sync(kFull, j < n && j > jstar && s_score[j] >= thr));
            }
            pos = jstar;
            if (later > 0) {
                const unsigned r = (a.rng_mode == LZB_RNG_PER_ROOT)
                                       ? tiebreak_hash(tie_seed, (unsigned)a.sim, (unsigned)(p.root_offset + i), (unsigned)depth)
                                       : tiebreak_hash(tie_seed, (unsigned)a.sim, 0u, 0u);
                int k = (int)(r % (unsigned)(later + 1));
                if (k > 0) {
                    for (int j0 = 0; j0 < n; j0 += 32) {
                        const int j = j0 + lane;
                        const unsigned m = __ballot_sync(kFull, j < n && j > jstar && s_score[j] >= thr);
                        const int c = __popc(m);
                        if (k <= c) { pos = j0 + (int)__fns(m, 0, k); break; }
                        k -= c;
                    }
                }
            }
        }
        const int4 ch = s_child[pos];
        action = (VAR != LZB_SEZ && e == 0) ? p.legal[(size_t)i * p.A + pos] : pos;
        if (players > 1) vtp = (vtp == 1) ? 2 : 1; // cnode.cpp:797-804
        if (lane == 0) {
            pa[depth] = make_int4(e * B + 1 + pos, ch.x, ch.y, ch.w);
            pb[depth] = make_int2(h.y, h.z);
            p.best[(size_t)i * p.E + e] = pos;
        }
        depth += 1;
        const int ce = child_expansion(ch.x);
        if (ce < 0) break;
        __syncwarp(); // s_child / s_score are rewritten by the next level
        parent_visits = child_visits(ch.x);
        parent_reward = __int_as_float(ch.w);
        parent_q = mean_q;
        is_root = false;
        e = ce;
    }
    write_leaf(p, a, i, e, pos, action, depth, vtp, parent_reset, lane);
}

template <int VAR>
__global__ void k_select_puct(Pool p, SelectArgs a)
{
    extern __shared__ int4 smem4[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int i = blockIdx.x * wpb + wib;
    if (i >= p.N) return;
    int4 *s_child = smem4 + (size_t)wib * p.A;
    float *s_score = reinterpret_cast<float *>(smem4 + (size_t)wpb * p.A) + (size_t)wib * p.A;
    select_puct_tree<VAR>(p, a, i, lane, s_child, s_score);
}

// ---------------------------------------------------------------------------------------------
// Gumbel MuZero (ctree_gumbel_muzero/lib/cnode.cpp).  Warp-private smem: s_child[A], s_q[A], s_p[A], s_c[A].

// csoftmax (gmz cnode.cpp:904-933) over s[0..n): max, serial fp32 sum of expf(x-m) in index order,
// then expf(x - m - logf(sum)).  In place.
__device__ __forceinline__ void warp_csoftmax(float *s, int n, int lane)
{
    float m = -INFINITY;
    for (int j = lane; j < n; j += 32) m = fmaxf(m, s[j]);
    m = warp_max(m);
    float sum = 0.0f;
    for (int j0 = 0; j0 < n; j0 += 32) {
        const int j = j0 + lane;
        const float ex = (j < n) ? glibc_expf(s[j] - m) : 0.0f;
        const int cnt = min(32, n - j0);
        for (int l = 0; l < cnt; ++l) sum += __shfl_sync(kFull, ex, l);
    }
    const float ls = glibc_logf(sum);
    for (int j = lane; j < n; j += 32) s[j] = glibc_expf(s[j] - m - ls);
    __syncwarp();
}

// qtransform_completed_by_mix_value (gmz cnode.cpp:989-1040) with the header defaults
// (maxvisit_init 50, value_scale 0.1, rescale on, epsilon 1e-8; gmz cnode.h:101-102) for the node whose
// children are staged in s_child[0..n).  Leaves completed Q in s_c; s_q, s_p are scratch.  Returns sum of visits.
__device__ __forceinline__ int gumbel_completed_q(const int4 *s_child, float *s_q, float *s_p, float *s_c, int n,
                                                  float raw_value, float g, int lane)
{
    int vsum_i = 0, vmax_i = 0;
    for (int j = lane; j < n; j += 32) {
        const int4 c = s_child[j];
        const int vc = child_visits(c.x);
        const float val = vc ? __int_as_float(c.y) / (float)vc : 0.0f;
        s_q[j] = __int_as_float(c.w) + g * val; // get_q (:181-197)
        s_p[j] = __int_as_float(c.z);
        vsum_i += vc;
        vmax_i = max(vmax_i, vc);
    }
    vsum_i = warp_sum_i(vsum_i);
    vmax_i = warp_max_i(vmax_i);
    __syncwarp();
    warp_csoftmax(s_p, n, lane); // priors (probabilities) are soft-maxed again (:1014) - reproduced as is
    // compute_mixed_value (:935-970): serial sums over visited children in index order
    const float vsum_f = (float)vsum_i; // sum of small ints: exact in any order
    float probs_sum = 0.0f;
    for (int j = 0; j < n; ++j) {
        const float pj = s_p[j];
        const float pc = (pj < -10e7f) ? -10e7f : pj; // std::max(prior, min_num)
        if (child_visits(s_child[j].x) > 0) probs_sum += pc;
    }
    float wq = 0.0f;
    for (int j = 0; j < n; ++j) {
        const float pj = s_p[j];
        const float pc = (pj < -10e7f) ? -10e7f : pj;
        if (child_visits(s_child[j].x) > 0) wq += pc * s_q[j] / probs_sum;
    }
    const float mixed = (raw_value + vsum_f * wq) / (vsum_f + 1);
    float mx = -INFINITY, mn = INFINITY;
    for (int j = lane; j < n; j += 32) {
        const float c = child_visits(s_child[j].x) > 0 ? s_q[j] : mixed;
        s_c[j] = c;
        mx = fmaxf(mx, c);
        mn = fminf(mn, c);
    }
    mx = warp_max(mx);
    mn = warp_min(mn);
    float gap = mx - mn; // rescale_qvalues (:972-987)
    gap = (gap < 1e-8f) ? 1e-8f : gap;
    const float visit_scale = 50.0f + (float)vmax_i;
    for (int j = lane; j < n; j += 32) s_c[j] = (s_c[j] - mn) / gap * visit_scale * 0.1f;
    __syncwarp();
    return vsum_i;
}

// first index attaining the maximum of s[0..n) under the reference's strict '>' scan from -inf
// (gmz cnode.cpp:733-742,778-787); 0 when nothing beats -inf.
__device__ __forceinline__ int warp_first_argmax(const float *s, int n, int lane)
{
    float m = -INFINITY;
    for (int j = lane; j < n; j += 32) m = fmaxf(m, s[j]);
    m = warp_max(m);
    int best = 0x7fffffff;
    if (m > -INFINITY)
        for (int j = lane; j < n; j += 32)
            if (s[j] == m) { best = j; break; }
    __syncwarp();
    best = warp_min_i(best);
    return best == 0x7fffffff ? 0 : best;
}

__device__ __forceinline__ void select_gumbel_tree(const Pool &p, const SelectArgs &a, int i, int lane, int4 *s_child,
                                                   float *s_q, float *s_p, float *s_c)
{
    const int B = p.B;
    const int4 *tree = p.tree(i);
    int4 *pa = p.path_a + (size_t)i * (p.S + 1);
    int2 *pb = p.path_b + (size_t)i * (p.S + 1);
    const float g = a.discount;
    const int vtp = a.to_play[i]; // never flipped (gmz cnode.cpp:870-885)
    int e = 0, depth = 0, pos = 0, action = 0;

    while (true) {
        const int4 *blk = tree + (size_t)e * B;
        const int4 first = (lane < B) ? blk[lane] : make_int4(0, 0, 0, 0);
        const int4 h = shfl4(first, 0);
        const int n = h.x;
        for (int r = lane; r <= n; r += 32)
            if (r >= 1) s_child[r - 1] = (r < 32) ? first : blk[r];
        __syncwarp();
        const int vsum_i = gumbel_completed_q(s_child, s_q, s_p, s_c, n, __int_as_float(h.w), g, lane);
        if (e == 0) {
            // cselect_root_child (:701-745) + score_considered (:1097-1132)
            const int considered = p.vtable[vsum_i];
            float ml = -INFINITY;
            for (int j = lane; j < n; j += 32) ml = fmaxf(ml, __int_as_float(s_child[j].z));
            ml = warp_max(ml);
            for (int j = lane; j < n; j += 32) {
                const int4 c = s_child[j];
                const float logit = __int_as_float(c.z) - ml;
                float sc = p.gumbel[j] + logit + s_c[j];
                sc = (-1e9f < sc) ? sc : -1e9f;
                sc += (child_visits(c.x) == considered) ? 0.0f : -INFINITY;
                s_q[j] = sc;
            }
        } else {
            // cselect_interior_child (:747-790)
            for (int j = lane; j < n; j += 32) s_q[j] = __int_as_float(s_child[j].z) + s_c[j];
            __syncwarp();
            warp_csoftmax(s_q, n, lane);
            const float den = (float)(1 + vsum_i);
            for (int j = lane; j < n; j += 32) s_q[j] = s_q[j] - (float)child_visits(s_child[j].x) / den;
        }
        __syncwarp();
        pos = warp_first_argmax(s_q, n, lane);
        const int4 ch = s_child[pos];
        action = (e == 0) ? p.legal[(size_t)i * p.A + pos] : pos;
        if (lane == 0) {
            pa[depth] = make_int4(e * B + 1 + pos, ch.x, ch.y, ch.w);
            pb[depth] = make_int2(h.y, h.z);
            p.best[(size_t)i * p.E + e] = pos;
        }
        depth += 1;
        const int ce = child_expansion(ch.x);
        if (ce < 0) break;
        __syncwarp();
        e = ce;
    }
    write_leaf(p, a, i, e, pos, action, depth, vtp, 0, lane);
}

__global__ void k_select_gumbel(Pool p, SelectArgs a)
{
    extern __shared__ int4 smem4[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int i = blockIdx.x * wpb + wib;
    if (i >= p.N) return;
    const int A = p.A;
    int4 *s_child = smem4 + (size_t)wib * A;
    float *fbase = reinterpret_cast<float *>(smem4 + (size_t)wpb * A) + (size_t)wib * 3 * A;
    select_gumbel_tree(p, a, i, lane, s_child, fbase, fbase + A, fbase + 2 * A);
}

} // namespace lzb
