// Root prepare (K1) and fused expand + backpropagate (K4).  One warp per tree.
#pragma once
#include "lzb_math.cuh"
#include "lzb_pool.cuh"

namespace lzb {

// CNode::expand's masked softmax (cnode.cpp:118-140) for the n children of expansion `e`:
// max over legal logits, t = expf(l - max), fp32 sum strictly left-to-right in legal order,
// prior = t / sum; optional root noise mix (cnode.cpp:159-166).  s_tmp: >= n floats of warp-private smem.
__device__ __forceinline__ void expand_children(const Pool &p, int i, int e, int n, const int *legal,
                                                const float *logits, const float *noise, float frac,
                                                float *s_tmp, int lane)
{
    float mx = kFloatMin; // cnode.cpp:118
    for (int j = lane; j < n; j += 32) {
        const float l = logits[legal ? legal[j] : j];
        mx = fmaxf(mx, l);
    }
    mx = warp_max(mx);
    for (int j = lane; j < n; j += 32) s_tmp[j] = glibc_expf(logits[legal ? legal[j] : j] - mx);
    __syncwarp();
    float sum = 0.0f;
    for (int j = 0; j < n; ++j) sum += s_tmp[j]; // every lane, same order: bit-identical to the serial host loop
    int4 *dst = p.block(i, e) + 1;
    for (int j = lane; j < n; j += 32) {
        float prior = s_tmp[j] / sum;
        if (noise) prior = prior * (1 - frac) + noise[j] * frac;
        dst[j] = make_int4(0, 0, __float_as_int(prior), 0);
    }
    __syncwarp();
}

struct PrepArgs {
    const float *logits, *noise, *reward, *value, *sampled_actions;
    const int *to_play;
    float frac;
};

// CRoots::prepare / prepare_no_noise (cnode.cpp:321-358; gmz :418-455; sez :677-715).
__global__ void k_prepare(Pool p, PrepArgs a)
{
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int i = blockIdx.x * (blockDim.x >> 5) + wib;
    if (i >= p.N) return;
    float *s_tmp = smem + (size_t)wib * p.A;
    int n;
    if (p.variant == LZB_SEZ) {
        n = p.A; // K sampled children, priors unused by selection (sez cnode.cpp:1054-1079)
        int4 *dst = p.block(i, 0) + 1;
        for (int j = lane; j < n; j += 32) dst[j] = make_int4(0, 0, 0, 0);
        const int KD = p.A * p.D;
        float *tab = p.actions + (size_t)i * p.E * KD;
        for (int t = lane; t < KD; t += 32) tab[t] = a.sampled_actions[(size_t)i * KD + t];
    } else {
        const int *legal = p.legal + (size_t)i * p.A;
        n = p.block(i, 0)[0].x; // legal count was stored by forest_create
        expand_children(p, i, 0, n, legal, a.logits + (size_t)i * p.A, a.noise ? a.noise + (size_t)i * p.A : nullptr,
                        a.frac, s_tmp, lane);
    }
    if (lane == 0) {
        const float rv = a.value ? a.value[i] : 0.0f;
        p.block(i, 0)[0] = make_int4(n, a.to_play[i], 0, __float_as_int(rv));
        p.root[i] = make_int4(1, 0, __float_as_int(a.reward[i]), 0); // visit_count += 1 (cnode.cpp:338)
    }
}

struct BackupArgs {
    int latent_index, horizon;
    float discount;
    const float *reward, *value, *logits, *sampled_actions;
    const int *to_play, *is_reset;
    int *is_reset_out;
    int n_slots;                       // slot pointers of the new node (rows its children's leaves will gather)
    const char *new_slot[kMaxGathers];
    // ReZero reuse backup (cbatch_backpropagate_with_reuse, cnode.cpp:502-549), lane-group kernels only
    const int *true_action;    // [N]
    const float *reuse_value;  // [N]
    const int *reuse_flags;    // [N] bit 0: tree is in no_inference_lst, bit 1: in reuse_lst; nullptr: derived from the traverse
    const int *batch_index;    // [N] count_b of tree i in the caller's compacted batch; nullptr: i
    const int *active;         // [N] or nullptr: trees with active[i] == 0 are left untouched (lane-group backup kernel only)
    // Stochastic tree, both heads on the whole batch (lzb_chance_head_t): a chance leaf's outputs; nullptr = one batch
    const float *reward_c, *value_c, *logits_c;
    int logits_c_stride, logits_stride; // floats per logits row (logits_stride 0 = the pool width)
};

// cbatch_backpropagate: CNode::expand of the leaf + cbackpropagate along the recorded path
// (cnode.cpp:419-500; ez :482-601; gmz :605-652; sez :860-971).
template <int VAR>
__device__ __forceinline__ void expand_backup_tree(const Pool &p, const BackupArgs &a, int i, int lane, float *s_tmp)
{
    constexpr bool VP = (VAR == LZB_EZ || VAR == LZB_SEZ);
    const int4 res = p.res[i];
    const int depth = res.x; // search_len >= 1
    const int e_new = a.latent_index;
    const int to_play = a.to_play[i];
    const float value = a.value[i];
    const float leaf_reward = a.reward[i];
    int leaf_reset = 0;
    if (VP) {
        if (a.is_reset) leaf_reset = a.is_reset[i];
        else if (a.horizon > 0) leaf_reset = (depth % a.horizon == 0) ? 1 : 0; // mcts_ctree.py:810
        if (a.is_reset_out && lane == 0) a.is_reset_out[i] = leaf_reset;
    }

    // ---- expand the leaf as expansion e_new
    if (VAR == LZB_SEZ) {
        int4 *dst = p.block(i, e_new) + 1;
        for (int j = lane; j < p.A; j += 32) dst[j] = make_int4(0, 0, 0, 0);
        const int KD = p.A * p.D;
        float *tab = p.actions + ((size_t)i * p.E + e_new) * KD;
        for (int t = lane; t < KD; t += 32) tab[t] = a.sampled_actions[(size_t)i * KD + t];
    } else {
        expand_children(p, i, e_new, p.A, nullptr, a.logits + (size_t)i * p.A, nullptr, 0.0f, s_tmp, lane);
    }
    if (lane == 0) {
        p.block(i, e_new)[0] = make_int4(p.A, to_play, leaf_reset, __float_as_int(value));
        p.best[(size_t)i * p.E + e_new] = -1;
    }

    // ---- backup: node q = depth (leaf) ... 1 are path records q-1, node 0 is the root
    const int4 *pa = p.path_a + (size_t)i * (p.S + 1);
    const int2 *pb = p.path_b + (size_t)i * (p.S + 1);
    const bool two = (VAR != LZB_GMZ) && (to_play != -1); // cnode.cpp:432; gmz asserts -1 (compiled out)
    const float g = a.discount;
    float b = value; // bootstrap_value entering the current chunk
    float cmin = kFloatMax, cmax = kFloatMin;
    for (int qhi = depth; qhi >= 0; qhi -= 32) {
        const int q = qhi - lane;
        const bool valid = q >= 0;
        int x = 0, slot = -1, node_tp = 0, parent_reset = 0;
        float vsum = 0.f, rew = 0.f, parent_vp = 0.f;
        if (valid) {
            if (q == 0) {
                const int4 r = p.root[i];
                x = r.x; vsum = __int_as_float(r.y); rew = __int_as_float(r.z);
            } else {
                const int4 r = pa[q - 1];
                slot = r.x; x = r.y; vsum = __int_as_float(r.z); rew = __int_as_float(r.w);
                if (q == depth) rew = leaf_reward; // expand() stamps the leaf's reward / value_prefix first
            }
            if (two) node_tp = (q == depth) ? to_play : pb[q].x;
            if (VP && q >= 1) {
                parent_vp = (q == 1) ? __int_as_float(p.root[i].z) : __int_as_float(pa[q - 2].w);
                parent_reset = pb[q - 1].y;
            }
        }
        float r_pre = rew, r_boot = rew;
        if (VP) {
            r_pre = rew - parent_vp;              // ez cnode.cpp:515
            r_boot = (parent_reset == 1) ? rew : r_pre; // :518-522
        }
        const bool same = two && (node_tp == to_play);
        const float r_signed = same ? -r_boot : r_boot;
        // serial leaf->root recurrence (cnode.cpp:447,473-475): lane l receives b_q, then b = r + g*b
        float my_b = 0.f;
        const int cnt = min(32, qhi + 1);
        for (int l = 0; l < cnt; ++l) {
            const float rl = __shfl_sync(kFull, r_signed, l);
            if (lane == l) my_b = b;
            b = rl + g * b;
        }
        if (valid) {
            const float add = (two && !same) ? -my_b : my_b;
            vsum += add;
            const int visits = (q == 0 ? x : child_visits(x)) + 1;
            const float nv = vsum / (float)visits;
            float cand;
            if (VP) cand = r_pre + g * nv;             // ez :516,558 (value not negated)
            else if (two) cand = rew + g * -nv;        // mz :470
            else cand = rew + g * nv;                  // mz :445, gmz :627
            cmin = fminf(cmin, cand);
            cmax = fmaxf(cmax, cand);
            if (q == 0) {
                p.root[i] = make_int4(visits, __float_as_int(vsum), __float_as_int(rew), 0);
            } else {
                int4 *rec = p.tree(i) + slot;
                const int exp_idx = (q == depth) ? e_new : child_expansion(x);
                *reinterpret_cast<int2 *>(rec) = make_int2(child_pack(visits, exp_idx), __float_as_int(vsum));
                if (q == depth) reinterpret_cast<int *>(rec)[3] = __float_as_int(rew);
            }
        }
    }
    cmin = warp_min(cmin);
    cmax = warp_max(cmax);
    if (lane == 0) {
        float2 mm = p.minmax[i];
        mm.x = fminf(mm.x, cmin); // CMinMaxStats::update (cminimax.cpp:23-31)
        mm.y = fmaxf(mm.y, cmax);
        p.minmax[i] = mm;
    }
}

template <int VAR>
__global__ void k_expand_backup(Pool p, BackupArgs a)
{
    extern __shared__ float smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int i = blockIdx.x * (blockDim.x >> 5) + wib;
    if (blockIdx.x == 0 && threadIdx.x < a.n_slots) p.slots[(size_t)threadIdx.x * p.E + a.latent_index] = pick3(a.new_slot, threadIdx.x);
    if (i >= p.N) return;
    expand_backup_tree<VAR>(p, a, i, lane, smem + (size_t)wib * p.A);
}

} // namespace lzb
