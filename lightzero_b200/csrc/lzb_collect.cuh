// Root-batch prologue / epilogue around a search: what LightZero's policies and reanalyze
// buffers do per root in Python before `Roots.prepare` and after `get_distributions`, as batched kernels so that
// masks, noise, visit counts, targets and chosen actions never have to leave the device.
//   k_legal_from_mask   legal_actions = [a for a, x in enumerate(action_mask[j]) if x == 1]   (policy/muzero.py:671;
//                       buffer/game_buffer_muzero.py:552)
//   k_dirichlet_noise   np.random.dirichlet([alpha] * n_legal)                                (policy/muzero.py:674-677)
//   k_select_action     select_action(visit_counts, temperature, deterministic)               (policy/utils.py:515-539)
//                       + action = np.where(action_mask == 1)[0][action_pos]                  (policy/muzero.py:709)
//   k_target_policies   visit_count / sum_visits scattered to action ids                      (game_buffer_muzero.py:617-656)
// Integer / index results are exact; the float64 arithmetic follows the reference's evaluation order (Python's
// left-to-right sum, numpy's sequential cumsum); the Dirichlet sampler is a distributional restatement (the
// reference draws from numpy's global Mersenne Twister), keyed by a counter-based generator like lzb_sample.cuh.
#pragma once
#include "lzb_sample.cuh"

namespace lzb {

// ------------------------------------------------------------------------------------------------ masks -> lists
// one warp per root: stable compaction of the legal action ids, rest of the row zero (as Forest expects it)
__global__ void k_legal_from_mask(const signed char *mask, int N, int A, int *count, int *ids)
{
    const int lane = threadIdx.x & 31, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= N) return;
    int base = 0;
    for (int j0 = 0; j0 < A; j0 += 32) {
        const int j = j0 + lane;
        const bool legal = j < A && mask[(size_t)i * A + j] == 1;
        const unsigned m = __ballot_sync(kFull, legal);
        if (legal) ids[(size_t)i * A + base + __popc(m & ((1u << lane) - 1u))] = j;
        base += __popc(m);
    }
    for (int j = base + lane; j < A; j += 32) ids[(size_t)i * A + j] = 0;
    if (lane == 0) count[i] = base;
}

// ------------------------------------------------------------------------------------------------ Dirichlet noise
// 53-bit uniform in [0, 1) from two 32-bit words (the construction of numpy's random_sample)
__device__ __forceinline__ double unit53(uint32_t a, uint32_t b)
{
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}

// Gamma(alpha, 1), Marsaglia-Tsang squeeze with the alpha < 1 boost; draws come from Philox blocks (c0 = attempt)
__device__ double gamma_draw(const Philox &rng, double alpha, uint32_t c1, uint32_t c2, uint32_t c3)
{
    const double a = (alpha < 1.0) ? alpha + 1.0 : alpha;
    const double d = a - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
    double g = 0.0;
    uint32_t attempt = 0;
    for (;; ++attempt) {
        const uint4 r = rng(attempt, c1, c2, c3);
        const double u1 = ((double)r.x + 0.5) * (1.0 / 4294967296.0), u2 = ((double)r.y + 0.5) * (1.0 / 4294967296.0);
        const double z = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2); // Box-Muller
        const double v0 = 1.0 + c * z;
        if (v0 <= 0.0) continue;
        const double v = v0 * v0 * v0;
        const double u = ((double)r.z + 0.5) * (1.0 / 4294967296.0);
        if (log(u) < 0.5 * z * z + d - d * v + d * log(v) || attempt > 64) {
            g = d * v;
            if (alpha < 1.0) {
                const double ub = ((double)r.w + 0.5) * (1.0 / 4294967296.0);
                g *= pow(ub, 1.0 / alpha);
            }
            break;
        }
    }
    return g;
}

// one warp per root; noise[i, j] for j < n_legal_i (position in the legal list, what Roots.prepare indexes), 0 beyond
__global__ void k_dirichlet_noise(float alpha, const int *count, float *noise, int N, int A, int root_offset,
                                  unsigned long long seed, unsigned counter)
{
    const int lane = threadIdx.x & 31, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= N) return;
    const Philox rng(seed);
    const int n = count ? min(count[i], A) : A;
    double total = 0.0;
    for (int j0 = 0; j0 < A; j0 += 32) { // pass 1: draws are recomputed in pass 2 instead of being staged
        const int j = j0 + lane;
        double g = (j < n) ? gamma_draw(rng, (double)alpha, counter, (uint32_t)(root_offset + i), (uint32_t)j) : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) g += __shfl_xor_sync(kFull, g, o);
        total += g;
    }
    for (int j0 = 0; j0 < A; j0 += 32) {
        const int j = j0 + lane;
        if (j >= A) continue;
        float out = 0.0f;
        if (j < n) {
            const double g = gamma_draw(rng, (double)alpha, counter, (uint32_t)(root_offset + i), (uint32_t)j);
            out = (float)(total > 0.0 ? g / total : 1.0 / n); // .astype(np.float32)
        }
        noise[(size_t)i * A + j] = out;
    }
}

// ------------------------------------------------------------------------------------------------ select_action
struct SelectActionArgs {
    const int *visits;        // [N, A] child visit counts by position in the legal list (get_distributions)
    const int *count;         // [N] n_legal, or nullptr = A
    const int *ids;           // [N, A] legal action ids by position, or nullptr = identity
    const double *uniforms;   // [N] optional injected draws in [0, 1)
    int *action_pos, *action; // [N] index within the legal set / action id
    double *entropy;          // [N] base-2 entropy of the visit-count distribution
    double temperature;
    int N, A, deterministic, root_offset;
    unsigned long long seed;
    unsigned counter;
};

// visit_count ** (1 / temperature) (policy/utils.py:530): exact for the integral exponents LightZero's temperature
// schedule produces (1, 2, 4), pow() otherwise
__device__ __forceinline__ double visit_power(int v, double inv_t, int int_exp)
{
    const double x = (double)v;
    if (int_exp > 0) {
        double r = x;
        for (int k = 1; k < int_exp; ++k) r *= x;
        return r;
    }
    return pow(x, inv_t);
}

// one thread per root: every sum is sequential in list order, as the Python / numpy code evaluates it
__global__ void k_select_action(SelectActionArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.N) return;
    const int n = a.count ? min(a.count[i], a.A) : a.A;
    const int *v = a.visits + (size_t)i * a.A;
    const double inv_t = 1.0 / a.temperature;
    int int_exp = 0;
    if (inv_t == floor(inv_t) && inv_t >= 1.0 && inv_t <= 8.0) int_exp = (int)inv_t;
    double total = 0.0; // sum(action_probs), left to right
    int best = 0, best_v = INT_MIN;
    for (int j = 0; j < n; ++j) {
        total += visit_power(v[j], inv_t, int_exp);
        if (v[j] > best_v) { best_v = v[j]; best = j; } // np.argmax: first maximum
    }
    int pos = best;
    double pk_total = 0.0, cdf_last = 0.0;
    for (int j = 0; j < n; ++j) {
        const double p = visit_power(v[j], inv_t, int_exp) / total;
        pk_total += p;
        cdf_last += p; // p.cumsum()[-1]
    }
    if (!a.deterministic) { // np.random.choice(n, p=action_probs): cdf /= cdf[-1]; searchsorted(u, side='right')
        double u;
        if (a.uniforms) u = a.uniforms[i];
        else {
            const uint4 r = Philox(a.seed)(0u, a.counter, (uint32_t)(a.root_offset + i), 0x53454c41u);
            u = unit53(r.x, r.y);
        }
        double c = 0.0;
        pos = 0;
        for (int j = 0; j < n; ++j) {
            c += visit_power(v[j], inv_t, int_exp) / total;
            if (c / cdf_last <= u) pos = j + 1;
        }
        if (pos >= n) pos = n - 1;
    }
    double s = 0.0; // scipy.stats.entropy(action_probs, base=2): pk / sum(pk), sum(-pk log pk) / log(2)
    for (int j = 0; j < n; ++j) {
        const double pk = (visit_power(v[j], inv_t, int_exp) / total) / pk_total;
        if (pk > 0.0) s += -pk * log(pk);
    }
    if (a.action_pos) a.action_pos[i] = pos;
    if (a.action) a.action[i] = a.ids ? a.ids[(size_t)i * a.A + pos] : pos;
    if (a.entropy) a.entropy[i] = s / log(2.0);
}

// ------------------------------------------------------------------------------------------------ target policies
// policy[i, :] = 0 where policy_mask[i] == 0; else visits / sum(visits) in float64, cast to float32, placed at the
// action ids when `ids` is given (board games) or at the positions themselves (fixed action space)
__global__ void k_target_policies(const int *visits, const int *count, const int *ids, const int *policy_mask, int N,
                                  int A, float *policy)
{
    const int lane = threadIdx.x & 31, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= N) return;
    const int n = count ? min(count[i], A) : A;
    for (int j = lane; j < A; j += 32) policy[(size_t)i * A + j] = 0.0f;
    if (policy_mask && policy_mask[i] == 0) return;
    long long sum = 0;
    for (int j0 = 0; j0 < n; j0 += 32) {
        long long s = (j0 + lane < n) ? visits[(size_t)i * A + j0 + lane] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(kFull, s, o);
        sum += s;
    }
    if (sum == 0) return;
    __syncwarp(); // the zero fill above is ordered before the scattered writes below
    for (int j = lane; j < n; j += 32) {
        const int at = ids ? ids[(size_t)i * A + j] : j;
        policy[(size_t)i * A + at] = (float)((double)visits[(size_t)i * A + j] / (double)sum);
    }
}

// ------------------------------------------------------------------------------------------------ value targets
// The n-step value targets of the reanalyze buffer (lzero/mcts/buffer/game_buffer_muzero.py:463-513), one thread per
// transition (segment b, unroll step u).  Everything is float64 like the reference's Python floats / numpy arrays
// (-fmad=false: mul and add stay separate roundings); discount ** k comes from a host table so that it is the host
// libm's pow().  board_games reproduces the reference's to_play indexing as written (:496: the reward's OFFSET i, not
// its position, indexes to_play_list).
struct ValueTargetArgs {
    const float *root_values;   // [B * (U + 1)] searched root values (use_root_value) or predicted values
    const int *td_steps;        // [B * (U + 1)]
    const float *value_mask;    // [B * (U + 1)]
    const float *rewards;       // [B, reward_stride] reward_list of each game segment, padded
    const int *reward_len;      // [B] len(reward_list)
    const int *segment_len;     // [B] game_segment_lens
    const int *pos;             // [B] pos_in_game_segment_list
    const int *to_play;         // [B, to_play_stride] to_play_segment (board games), or nullptr
    const double *pow_table;    // [pow_len] discount ** k
    double *target_values;      // [B, U + 1]
    double *target_rewards;     // [B, U + 1]
    int B, U, reward_stride, to_play_stride, pow_len, board_games;
    int horizon;                // EfficientZero: lstm_horizon_len (value-prefix targets), 0 = MuZero's reward targets
};

__global__ void k_value_targets(ValueTargetArgs a)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int per = a.U + 1;
    if (t >= a.B * per) return;
    const int b = t / per, u = t - b * per;
    const int state_index = a.pos[b], current_index = state_index + u;
    const int td = a.td_steps[t];
    const double dpow = a.pow_table[min(max(td, 0), a.pow_len - 1)];
    double v = (double)a.root_values[t];
    if (a.board_games) v = __dmul_rn(v, (td % 2 == 0) ? dpow : -dpow); // :466-473
    else v = __dmul_rn(v, dpow);                                       // :474-477
    v = __dmul_rn(v, (double)a.value_mask[t]);                          // :479
    const int len = a.reward_len[b];
    const float *rew = a.rewards + (size_t)b * a.reward_stride;
    // EfficientZero (game_buffer_efficientzero.py:278,298-301): horizon_id == t runs over the whole batch; the reset of step u'
    // takes effect AFTER that step's value sum, so this step's base index is the last reset strictly before u, its value
    // prefix restarts at the last reset at or before u
    int base_index = state_index, prefix_from = 0;
    if (a.horizon > 0) {
        for (int w = 0; w <= u; ++w)
            if ((b * per + w) % a.horizon == 0) {
                if (w < u) base_index = state_index + w;
                prefix_from = w;
            }
    }
    // reward_list[current_index:bootstrap_index]: Python slice clipping (:490-491)
    const int lo = min(max(current_index, 0), len), hi = min(max(current_index + td, 0), len);
    for (int k = lo; k < hi; ++k) {
        const int i = k - lo;
        const double term = __dmul_rn((double)rew[k], a.pow_table[min(i, a.pow_len - 1)]);
        bool same = true;
        if (a.board_games) {
            const int *tp = a.to_play + (size_t)b * a.to_play_stride;
            same = tp[min(base_index, a.to_play_stride - 1)] == tp[min(i, a.to_play_stride - 1)]; // to_play_list[base_index] == to_play_list[i] (:496)
        }
        v = same ? __dadd_rn(v, term) : __dadd_rn(v, -term);
    }
    const bool in_segment = current_index < a.segment_len[b]; // :505-510
    a.target_values[t] = in_segment ? v : 0.0;
    if (a.horizon > 0) { // value_prefix += reward_list[current_index] for the steps inside the segment (:303-311)
        double prefix = 0.0;
        for (int w = prefix_from; w <= u; ++w)
            if (state_index + w < a.segment_len[b]) prefix = __dadd_rn(prefix, (double)rew[min(state_index + w, a.reward_stride - 1)]);
        a.target_rewards[t] = prefix;
        return;
    }
    a.target_rewards[t] = in_segment ? (double)rew[min(current_index, a.reward_stride - 1)] : 0.0;
}

} // namespace lzb
