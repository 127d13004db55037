// liblzb200.so: C ABI of include/lzb200.h over the kernels in lzb_expand.cuh / lzb_select.cuh.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo -O3 -std=c++17 --shared ...
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <random>
#include <vector>

#include "../../include/lzb200.h"
#include "lzb_group.cuh"
#include "lzb_sample.cuh"
#include "lzb_collect.cuh"
#include "lzb_peer.cuh"
#include "lzb_scaling.cuh"
#include "lzb_step.cuh"
#include "lzb_gumbel_helpers.cuh"

using namespace lzb;

struct lzb_forest {
    Pool pool;
    size_t bytes;
    float delta;
    int sim;      // traverses issued since the last prepare (keys the tie-break stream)
    int pbc_base; float pbc_init; // pUCT constants the pbc table in the workspace was computed for
    const void *players_for; // to_play pointer whose batch maximum is in pool.scratch[0] for this search
    int expanded; // highest latent_index expanded so far
    bool slot_by_kind; // stochastic tree: new nodes' rows were registered per kind (lzb_chance_head_t::latent_dev)
    int vt_m, vt_s;
    std::vector<int> vtable_host;
    std::vector<float> gumbel_host;
};

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define LZB_CUDA(expr)                                                                                  \
    do {                                                                                                \
        cudaError_t err__ = (expr);                                                                     \
        if (err__ != cudaSuccess) return fail(LZB_E_CUDA, "%s: %s", #expr, cudaGetErrorString(err__)); \
    } while (0)

constexpr int kWarpsPerBlock = 4;
constexpr size_t kAlign = 256;

inline size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

struct Layout {
    size_t node, root, legal, minmax, path_a, path_b, res, best, bidx, actions, gumbel, vtable, scratch, slots, pbc, pbc2, cval, gstat, total;
};

constexpr size_t kPbc2MaxW = 1024;

Layout make_layout(int variant, int N, int A, int S, int D)
{
    const size_t E = (size_t)S + 1;
    Layout l;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes); return at; };
    l.node = take((size_t)N * E * (A + 1) * sizeof(int4));
    l.root = take((size_t)N * sizeof(int4));
    l.legal = take((size_t)N * A * sizeof(int));
    l.minmax = take((size_t)N * sizeof(float2));
    l.path_a = take((size_t)N * E * sizeof(int4));
    l.path_b = take((size_t)N * E * sizeof(int2));
    l.res = take((size_t)N * sizeof(int4));
    l.best = take((size_t)N * E * sizeof(int));
    l.bidx = take((size_t)N * E * sizeof(int));
    l.actions = take(variant == LZB_SEZ ? (size_t)N * E * A * D * sizeof(float) : 0);
    l.gumbel = take((size_t)A * sizeof(float));
    l.vtable = take(E * sizeof(int));
    l.scratch = take(4 * sizeof(int));
    l.slots = take((size_t)kMaxGathers * E * sizeof(char *));
    l.pbc = take((E + 1) * sizeof(float2));
    l.pbc2 = l.cval = 0;
#ifdef LZB_X_PBC2 // experiment: the whole pb_c factor from a table (searches of up to kPbc2MaxSims simulations)
    l.pbc2 = take((variant != LZB_GMZ && E + 1 <= kPbc2MaxW) ? (E + 1) * (E + 1) * sizeof(float) : 0);
#endif
#ifdef LZB_X_CVAL // experiment: child values cached by the backup
    l.cval = take(variant != LZB_GMZ ? (size_t)N * E * (A + 1) * sizeof(float) : 0);
#endif
    l.gstat = 0;
#ifndef LZB_X_NO_GMZ_STATS // per-node constants of the Gumbel tree's softmax over priors, computed once (round 2: 24.1 -> 23.0 us per step)
    l.gstat = take(variant == LZB_GMZ ? (size_t)N * E * sizeof(float2) : 0);
#endif
    l.total = o;
    return l;
}

int check_dims(int variant, int N, int A, int S, int D)
{
    if (variant < LZB_MZ || variant > LZB_SMZ) return fail(LZB_E_ARG, "unknown variant %d", variant);
    if (N <= 0 || A <= 0 || S <= 0) return fail(LZB_E_ARG, "num_roots, action_space and max_sims must be positive");
    if (S > 65000) return fail(LZB_E_UNSUPPORTED, "max_sims %d exceeds the 16-bit visit/expansion packing", S);
    if (variant == LZB_SEZ && D <= 0) return fail(LZB_E_ARG, "sampled tree needs action_dim > 0");
    if (variant == LZB_SMZ && D <= 0) return fail(LZB_E_ARG, "stochastic tree needs chance_space_size > 0 (passed as action_dim)");
    if (variant == LZB_SMZ && std::max(A, D) > 128) return fail(LZB_E_UNSUPPORTED, "stochastic tree supports max(action_space, chance_space) <= 128");
    if ((size_t)A * 28 > 200 * 1024) return fail(LZB_E_UNSUPPORTED, "action_space %d too large for shared-memory staging", A);
    return LZB_OK;
}

// child records per node: the stochastic tree's nodes hold policy OR chance children (action_dim carries C)
int record_width(int variant, int A, int D) { return variant == LZB_SMZ ? std::max(A, D) : A; }

// warps per block so that `bytes_per_warp` of dynamic shared memory fits the default 48 KB window
int warps_for(size_t bytes_per_warp) { return (bytes_per_warp * kWarpsPerBlock <= 48 * 1024) ? kWarpsPerBlock : 1; }

template <typename K>
int launch_cfg(K kernel, size_t bytes_per_warp, int N, dim3 &grid, dim3 &block, size_t &smem)
{
    const int wpb = warps_for(bytes_per_warp);
    smem = bytes_per_warp * wpb;
    if (smem > 48 * 1024) LZB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    block = dim3(32 * wpb);
    grid = dim3((N + wpb - 1) / wpb);
    return LZB_OK;
}

__global__ void k_init(Pool p, const int *legal_count, const int *legal_actions)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.N) return;
    if (i == 0) p.scratch[1] = 0; // the checked build's violation word (LZB_CHK)
    int n = legal_count ? legal_count[i] : 0;
    const bool all = (n <= 0); // empty list == every action legal (cnode.cpp:101-107)
    if (all) n = p.Ad;
    for (int j = 0; j < p.A; ++j) p.legal[(size_t)i * p.A + j] = (all || !legal_actions) ? j : legal_actions[(size_t)i * p.A + j];
    p.block(i, 0)[0] = make_int4(n, 0, 0, 0);
    p.root[i] = make_int4(0, 0, 0, 0);
    p.minmax[i] = make_float2(kFloatMax, kFloatMin); // CMinMaxStats ctor (cminimax.cpp:7-11)
    p.res[i] = make_int4(0, 0, 0, 0);
    p.best[(size_t)i * p.E] = -1; // CNode::best_action = -1 (cnode.cpp:55-66)
    p.bidx[(size_t)i * p.E] = i;  // roots expand with batch_index = i (cnode.cpp:335)
}

__global__ void k_reset_minmax(Pool p)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.N) p.minmax[i] = make_float2(kFloatMax, kFloatMin);
}

// players = (max(to_play_batch) == -1) ? 1 : 2 is batch-global (cnode.cpp:776-781): out[0] = max over the batch
__global__ void k_max_to_play_init(int *out) { out[0] = INT_MIN; }
__global__ void k_max_to_play(const int *to_play, int n, int *out)
{
    int m = INT_MIN;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = max(m, to_play[i]);
    m = warp_max_i(m);
    if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

// pb_c factors of cucb_score (cnode.cpp:673-674) for every possible parent visit count, with the tree's own logf
__global__ void k_pbc_table(Pool p, int c_base, float c_init)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > p.E) return;
    const float T = (float)t, c2 = (float)c_base;
    p.pbc[t] = make_float2(glibc_logf((T + c2 + 1) / c2) + c_init, sqrtf(T));
}

// experiment LZB_X_PBC2: pb_c = pbc0 * (sqrt(T) / (child visits + 1)) for every (T, child visits) pair, evaluated with the
// very expression the selection kernels use (same fp32 operations, no contraction), so a lookup is bit-identical
__global__ void k_pbc2_table(Pool p, int c_base, float c_init)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int W = p.pbc2_w;
    if (t >= W * W) return;
    const int Ti = t / W, vc = t - Ti * W;
    const float T = (float)Ti, c2 = (float)c_base;
    const float pbc0 = glibc_logf((T + c2 + 1) / c2) + c_init, sqT = sqrtf(T);
    p.pbc2[t] = pbc0 * (sqT / (float)(vc + 1));
}

__global__ void k_set_slots(Pool p, int e, int n, const char *s0, const char *s1, const char *s2)
{
    if (threadIdx.x < n) p.slots[(size_t)threadIdx.x * p.E + e] = threadIdx.x == 0 ? s0 : (threadIdx.x == 1 ? s1 : s2);
}

// dst[i,:] = is_reset[i] ? 0 : src[i,:]; 16-byte accesses when VEC (row bytes and both bases 16-B aligned)
template <bool VEC>
__global__ void k_store_hidden(const float *src, float *dst, const int *is_reset, int n, long long row)
{
    if (VEC) {
        const long long row4 = row >> 2, total = (long long)n * row4;
        const int4 *s4 = reinterpret_cast<const int4 *>(src);
        int4 *d4 = reinterpret_cast<int4 *>(dst);
        for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
            int4 v = __ldg(s4 + t);
            if (is_reset && is_reset[t / row4]) v = make_int4(0, 0, 0, 0);
            d4[t] = v;
        }
    } else {
        const long long total = (long long)n * row;
        for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x)
            dst[t] = (is_reset && is_reset[t / row]) ? 0.0f : src[t];
    }
}

// get_distributions / get_values (cnode.cpp:387-417) and the Gumbel extras (gmz :309-385,506-541).
__global__ void k_root_outputs(Pool p, int *visits, int *legal_count, float *values, float discount, float *policies,
                               float *child_values)
{
    extern __shared__ int4 smem4[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int i = blockIdx.x * wpb + wib;
    if (i >= p.N) return;
    const int A = p.A;
    const int4 h = p.block(i, 0)[0];
    const int n = h.x;
    const int4 *kids = p.block(i, 0) + 1;
    if (visits)
        for (int j = lane; j < A; j += 32) visits[(size_t)i * A + j] = j < n ? child_visits(kids[j].x) : 0;
    if (lane == 0) {
        if (legal_count) legal_count[i] = n;
        if (values) {
            const int4 r = p.root[i];
            values[i] = r.x ? __int_as_float(r.y) / (float)r.x : 0.0f; // CNode::value (cnode.cpp:219-235)
        }
    }
    if (p.variant == LZB_GMZ && (policies || child_values)) {
        int4 *s_child = smem4 + (size_t)wib * A;
        float *fbase = reinterpret_cast<float *>(smem4 + (size_t)wpb * A) + (size_t)wib * 3 * A;
        float *s_q = fbase, *s_p = fbase + A, *s_c = fbase + 2 * A;
        for (int j = lane; j < n; j += 32) s_child[j] = kids[j];
        __syncwarp();
        gumbel_completed_q(s_child, s_q, s_p, s_c, n, __int_as_float(h.w), discount, lane);
        const int *legal = p.legal + (size_t)i * A;
        if (child_values) {
            for (int a = lane; a < A; a += 32) child_values[(size_t)i * A + a] = -INFINITY;
            __syncwarp();
            for (int j = lane; j < n; j += 32) child_values[(size_t)i * A + legal[j]] = s_c[j];
        }
        if (policies) {
            for (int a = lane; a < A; a += 32) s_q[a] = -INFINITY;
            __syncwarp();
            for (int j = lane; j < n; j += 32) s_q[legal[j]] = __int_as_float(s_child[j].z) + s_c[j];
            __syncwarp();
            warp_csoftmax(s_q, A, lane); // over the whole action space, -inf for illegal (gmz :374-382)
            for (int a = lane; a < A; a += 32) policies[(size_t)i * A + a] = s_q[a];
        }
    }
}

__global__ void k_dump_tree(Pool p, int i, int upto, int *visit, float *vsum, float *prior, float *reward, int *child_e,
                            int *hdr, float *raw)
{
    const int A = p.A;
    for (int e = 0; e <= upto; ++e) {
        const int4 h = p.block(i, e)[0];
        if (h.x == 0) continue; // a ReZero simulation that needed no inference expanded nothing in this tree
        if (threadIdx.x == 0) {
            hdr[e * 4 + 0] = h.x;
            hdr[e * 4 + 1] = h.y;
            hdr[e * 4 + 2] = h.z;
            if (e == 0) hdr[3] = -1;
            if (raw && p.variant == LZB_GMZ) raw[e] = __int_as_float(h.w);
        }
        for (int j = threadIdx.x; j < h.x; j += blockDim.x) {
            const int4 c = p.block(i, e)[1 + j];
            const size_t o = (size_t)e * A + j;
            const int ce = child_expansion(c.x);
            visit[o] = child_visits(c.x);
            vsum[o] = __int_as_float(c.y);
            prior[o] = __int_as_float(c.z);
            reward[o] = ce >= 0 ? __int_as_float(c.w) : 0.0f;
            child_e[o] = ce;
            if (ce >= 0) hdr[ce * 4 + 3] = e;
        }
    }
}

__global__ void k_dump_paths(Pool p, int *out, int stride)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.N) return;
    const int depth = p.res[i].x;
    for (int d = 0; d < depth && d < stride; ++d) {
        const int slot = p.path_a[(size_t)i * (p.S + 1) + d].x;
        const int e = slot / p.B, pos = slot % p.B - 1;
        out[(size_t)i * stride + d] = (e == 0 && p.variant != LZB_SEZ) ? p.legal[(size_t)i * p.A + pos] : pos;
    }
}

// CRoots::get_trajectories (cnode.cpp:237-257,369-385): follow best_action from the root while it is >= 0
__global__ void k_trajectories(Pool p, int *out, int stride, int *len)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.N) return;
    int e = 0, d = 0;
    while (e >= 0) {
        const int pos = p.best[(size_t)i * p.E + e];
        if (pos < 0) break;
        if (d < stride) out[(size_t)i * stride + d] = (e == 0 && p.variant != LZB_SEZ) ? p.legal[(size_t)i * p.A + pos] : pos;
        d += 1;
        e = child_expansion(p.block(i, e)[1 + pos].x);
    }
    len[i] = d;
}

__global__ void k_minmax(Pool p, float *mn, float *mx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.N) { mn[i] = p.minmax[i].x; mx[i] = p.minmax[i].y; }
}

__global__ void k_root_stats(Pool p, int *visit, float *vsum, float *reward)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p.N) {
        const int4 r = p.root[i];
        visit[i] = r.x; vsum[i] = __int_as_float(r.y); reward[i] = __int_as_float(r.z);
    }
}

__global__ void k_test_expf(const float *x, float *out, long long n)
{
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) out[t] = glibc_expf(x[t]);
}
__global__ void k_test_logf(const float *x, float *out, long long n)
{
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) out[t] = glibc_logf(x[t]);
}

// get_sequence_of_considered_visits (gmz cnode.cpp:1042-1077)
std::vector<int> considered_sequence(int m, int S)
{
    std::vector<int> seq;
    if (m <= 1) {
        for (int i = 0; i < S; ++i) seq.push_back(i);
        return seq;
    }
    int log2max = 0;
    while ((1 << log2max) < m) ++log2max;
    std::vector<int> visits(m, 0);
    int considered = m;
    while ((int)seq.size() < S) {
        const int extra = std::max(1, S / (log2max * considered));
        for (int i = 0; i < extra; ++i) {
            seq.insert(seq.end(), visits.begin(), visits.begin() + considered);
            for (int j = 0; j < considered; ++j) visits[j] += 1;
        }
        considered = std::max(2, considered / 2);
    }
    seq.resize(S);
    return seq;
}

// generate_gumbel (gmz cnode.cpp:1134-1152): the host C++ library's own mt19937 / extreme_value_distribution
void gumbel_noise(float scale, int shape, float *out)
{
    std::mt19937 gen(0u);
    std::extreme_value_distribution<float> d(0, 1);
    for (int i = 0; i < shape; ++i) out[i] = scale * d(gen);
}

} // namespace

extern "C" {

const char *lzb_last_error(void) { return g_err; }
int lzb_abi_version(void) { return LZB_ABI_VERSION; }
#ifdef LZB_TIMELINE
// tuning builds only (scripts/gpu_timeline.py): the buffer [launch][warp][24] of u64 the fused step kernel stamps
int lzb_debug_timeline(void *buf_dev) { return cudaMemcpyToSymbol(g_timeline, &buf_dev, sizeof(buf_dev)) == cudaSuccess ? 0 : -1; }
#endif
uint32_t lzb_tiebreak_hash(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth) { return tiebreak_hash(seed, sim, root, depth); }
uint64_t lzb_chance_bits(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth) { return chance_bits(seed, sim, root, depth); }

int lzb_forest_bytes(int variant, int num_roots, int action_space, int max_sims, int action_dim, size_t *bytes)
{
    if (!bytes) return fail(LZB_E_ARG, "bytes is NULL");
    if (int rc = check_dims(variant, num_roots, action_space, max_sims, action_dim)) return rc;
    *bytes = make_layout(variant, num_roots, record_width(variant, action_space, action_dim), max_sims, action_dim).total;
    return LZB_OK;
}

int lzb_forest_create(lzb_forest_t **out, void *workspace_dev, size_t workspace_bytes, int variant, int num_roots,
                      int action_space, int max_sims, int action_dim, int root_offset, const int32_t *legal_count_dev,
                      const int32_t *legal_actions_dev, void *stream)
{
    if (!out || !workspace_dev) return fail(LZB_E_ARG, "out / workspace is NULL");
    if (int rc = check_dims(variant, num_roots, action_space, max_sims, action_dim)) return rc;
    if (((uintptr_t)workspace_dev & (kAlign - 1)) != 0) return fail(LZB_E_ARG, "workspace must be %zu-byte aligned", kAlign);
    const int decision_actions = action_space, chance_space = (variant == LZB_SMZ) ? action_dim : 0;
    action_space = record_width(variant, action_space, action_dim);
    if (variant == LZB_SMZ) action_dim = 0;
    const Layout l = make_layout(variant, num_roots, action_space, max_sims, action_dim);
    if (workspace_bytes < l.total) return fail(LZB_E_ARG, "workspace too small: %zu < %zu", workspace_bytes, l.total);
    if ((legal_count_dev == nullptr) != (legal_actions_dev == nullptr)) return fail(LZB_E_ARG, "legal_count / legal_actions must both be given or both NULL");
    lzb_forest *f = new lzb_forest();
    char *w = (char *)workspace_dev;
    Pool &p = f->pool;
    p.variant = variant; p.N = num_roots; p.A = action_space; p.S = max_sims; p.E = max_sims + 1; p.D = action_dim;
    p.root_offset = root_offset; p.B = action_space + 1; p.Ad = decision_actions; p.C = chance_space;
    p.node = (int4 *)(w + l.node); p.root = (int4 *)(w + l.root); p.slots = (const char **)(w + l.slots); p.pbc = (float2 *)(w + l.pbc);
    p.legal = (int *)(w + l.legal); p.minmax = (float2 *)(w + l.minmax);
    p.path_a = (int4 *)(w + l.path_a); p.path_b = (int2 *)(w + l.path_b); p.res = (int4 *)(w + l.res);
    p.best = (int *)(w + l.best); p.bidx = (int *)(w + l.bidx); p.actions = (float *)(w + l.actions); p.gumbel = (float *)(w + l.gumbel);
    p.vtable = (int *)(w + l.vtable); p.scratch = (int *)(w + l.scratch);
    p.pbc2 = nullptr; p.pbc2_w = 0; p.cval = nullptr; p.gstat = nullptr;
#ifndef LZB_X_NO_GMZ_STATS
    if (variant == LZB_GMZ) p.gstat = (float2 *)(w + l.gstat);
#endif
#ifdef LZB_X_PBC2
    if (variant != LZB_GMZ && (size_t)p.E + 1 <= kPbc2MaxW) { p.pbc2 = (float *)(w + l.pbc2); p.pbc2_w = p.E + 1; }
#endif
#ifdef LZB_X_CVAL
    if (variant != LZB_GMZ) p.cval = (float *)(w + l.cval);
#endif
    f->bytes = l.total; f->delta = 0.0f; f->sim = 0; f->expanded = -1; f->slot_by_kind = false; f->vt_m = -1; f->vt_s = -1; f->players_for = nullptr; f->pbc_base = -1; f->pbc_init = 0.0f;
    cudaStream_t st = (cudaStream_t)stream;
#ifdef LZB_CHECKED // a slot that was never registered reads as NULL (check 11) instead of whatever the workspace held
    LZB_CUDA(cudaMemsetAsync((void *)p.slots, 0, (size_t)kMaxGathers * p.E * sizeof(char *), st));
#endif
    k_init<<<(num_roots + 127) / 128, 128, 0, st>>>(p, legal_count_dev, legal_actions_dev);
    if (variant == LZB_GMZ) {
        f->gumbel_host.resize(action_space);
        gumbel_noise(10.0f, action_space, f->gumbel_host.data()); // gumbel_scale = 10, gumbel_rng = 0 (gmz :87-89)
        cudaError_t e = cudaMemcpyAsync(p.gumbel, f->gumbel_host.data(), sizeof(float) * action_space, cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) { delete f; return fail(LZB_E_CUDA, "gumbel upload: %s", cudaGetErrorString(e)); }
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { delete f; return fail(LZB_E_CUDA, "forest init: %s", cudaGetErrorString(e)); }
    *out = f;
    return LZB_OK;
}

void lzb_forest_destroy(lzb_forest_t *f) { delete f; }

int lzb_forest_reset(lzb_forest_t *f, const int32_t *legal_count_dev, const int32_t *legal_actions_dev, void *stream)
{
    if (!f) return fail(LZB_E_ARG, "forest is NULL");
    if ((legal_count_dev == nullptr) != (legal_actions_dev == nullptr)) return fail(LZB_E_ARG, "legal_count / legal_actions must both be given or both NULL");
    k_init<<<(f->pool.N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(f->pool, legal_count_dev, legal_actions_dev);
    LZB_CUDA(cudaGetLastError());
    f->sim = 0;
    f->expanded = -1;
    f->slot_by_kind = false;
    f->players_for = nullptr;
    return LZB_OK;
}

int lzb_set_delta(lzb_forest_t *f, float value_delta_max, void *stream)
{
    if (!f) return fail(LZB_E_ARG, "forest is NULL");
    f->delta = value_delta_max;
    k_reset_minmax<<<(f->pool.N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(f->pool);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_prepare(lzb_forest_t *f, const float *logits_dev, const float *noise_dev, float noise_weight,
                const float *reward_dev, const float *value_dev, const int32_t *to_play_dev,
                const float *sampled_actions_dev, void *stream)
{
    if (!f) return fail(LZB_E_ARG, "forest is NULL");
    const Pool &p = f->pool;
    if (!reward_dev || !to_play_dev) return fail(LZB_E_ARG, "reward / to_play is NULL");
    if (p.variant == LZB_SEZ ? !sampled_actions_dev : !logits_dev) return fail(LZB_E_ARG, "logits (or sampled_actions for the sampled tree) is NULL");
    if (p.variant == LZB_GMZ && !value_dev) return fail(LZB_E_ARG, "the Gumbel tree needs root values");
    PrepArgs a{logits_dev, noise_dev, reward_dev, value_dev, sampled_actions_dev, to_play_dev, noise_weight};
    if (p.gstat) LZB_CUDA(cudaMemsetAsync(p.gstat, 0xff, (size_t)p.N * p.E * sizeof(float2), (cudaStream_t)stream)); // NaN: nothing cached for the new roots' trees
    dim3 grid, block; size_t smem;
    if (int rc = launch_cfg(k_prepare, (size_t)p.A * sizeof(float), p.N, grid, block, smem)) return rc;
    k_prepare<<<grid, block, smem, (cudaStream_t)stream>>>(p, a);
    LZB_CUDA(cudaGetLastError());
    f->sim = 0;
    f->expanded = 0;
    f->slot_by_kind = false;
    f->players_for = nullptr;
    return LZB_OK;
}

} // extern "C"

namespace {

// Validates the traverse arguments and fills the kernel-side struct; launches the batch max(to_play) reduction
// the first time a search sees this to_play buffer (players is batch-global, cnode.cpp:776-781).
// ReZero reuse lives in the lane-group kernels of the two trees that have it in the reference (mz, ez)
int check_reuse(const Pool &p)
{
    if (p.variant != LZB_MZ && p.variant != LZB_EZ) return fail(LZB_E_UNSUPPORTED, "reuse search exists for the MuZero and EfficientZero trees only");
    if (p.A > 128 || getenv("LZB_FORCE_WARP_PER_TREE")) return fail(LZB_E_UNSUPPORTED, "reuse search supports action_space <= 128");
    return LZB_OK;
}

int fill_select(lzb_forest_t *f, const lzb_search_cfg_t *cfg, const lzb_traverse_io_t *io, SelectArgs &a, cudaStream_t st)
{
    if (!f || !cfg || !io || !io->to_play_dev) return fail(LZB_E_ARG, "forest / cfg / io / to_play is NULL");
    if (f->expanded < 0) return fail(LZB_E_STATE, "traverse before prepare");
    if (io->num_gathers < 0 || io->num_gathers > kMaxGathers || (io->num_gathers && !io->gathers)) return fail(LZB_E_ARG, "0..3 gathers supported");
    if (cfg->rng_mode != LZB_RNG_BATCH && cfg->rng_mode != LZB_RNG_PER_ROOT) return fail(LZB_E_ARG, "rng_mode must be LZB_RNG_BATCH or LZB_RNG_PER_ROOT");
    if (cfg->players < 0 || cfg->players > 2) return fail(LZB_E_ARG, "players must be 0, 1 or 2");
    const Pool &p = f->pool;
    a = SelectArgs{};
    a.c_base = cfg->pb_c_base; a.c_init = cfg->pb_c_init; a.discount = cfg->discount; a.delta = f->delta;
    a.players = (p.variant == LZB_GMZ) ? 1 : cfg->players; a.sim = f->sim; a.rng_mode = cfg->rng_mode; a.seed = cfg->rng_seed;
    a.seed_dev = (const unsigned long long *)cfg->rng_seed_dev;
    a.num_simulations = cfg->num_simulations; a.ng = io->num_gathers; a.new_e = -1;
    a.slot_by_kind = f->slot_by_kind ? 1 : 0;
    a.to_play = io->to_play_dev; a.ix = io->ix_dev; a.iy = io->iy_dev; a.search_len = io->search_len_dev; a.vtp = io->vtp_dev;
    a.last_action = (long long *)io->last_action_dev; a.last_action_f = io->last_action_f_dev;
    for (int k = 0; k < io->num_gathers; ++k) {
        const lzb_gather_t &g = io->gathers[k];
        if (!g.dst_dev || g.row_bytes <= 0 || (g.row_bytes & 3))
            return fail(LZB_E_ARG, "gather %d: NULL destination or row_bytes not a positive multiple of 4", k);
        a.g[k] = Gather{(const char *)g.src_dev, (char *)g.dst_dev, (long long)g.row_bytes, g.zero_if_reset};
    }
    if (io->reuse) {
        if (int rc = check_reuse(p)) return rc;
        if (!io->reuse->true_action_dev || !io->reuse->reuse_value_dev) return fail(LZB_E_ARG, "reuse: true_action / reuse_value is NULL");
        a.true_action = io->reuse->true_action_dev; a.reuse_value = io->reuse->reuse_value_dev; a.infer_count = io->reuse->infer_count_dev;
    }
    if (io->chance) {
        if (p.variant != LZB_SMZ) return fail(LZB_E_ARG, "chance arguments exist for the stochastic tree only");
        if (io->chance->forced_outcome_dev && io->chance->forced_stride <= 0) return fail(LZB_E_ARG, "chance: forced_stride must be positive");
        a.forced_outcome = io->chance->forced_outcome_dev; a.forced_stride = io->chance->forced_stride;
        a.leaf_is_chance = io->chance->leaf_is_chance_dev;
        a.chance_seed = (const unsigned long long *)io->chance->seed_dev;
    }
    if (p.variant == LZB_SMZ && getenv("LZB_FORCE_WARP_PER_TREE")) return fail(LZB_E_UNSUPPORTED, "the stochastic tree has lane-group kernels only");
    if (p.variant != LZB_GMZ && (f->pbc_base != cfg->pb_c_base || f->pbc_init != cfg->pb_c_init)) {
        k_pbc_table<<<(p.E + 128) / 128, 128, 0, st>>>(p, cfg->pb_c_base, cfg->pb_c_init);
        if (p.pbc2) k_pbc2_table<<<(p.pbc2_w * p.pbc2_w + 255) / 256, 256, 0, st>>>(p, cfg->pb_c_base, cfg->pb_c_init);
        f->pbc_base = cfg->pb_c_base; f->pbc_init = cfg->pb_c_init;
    }
    if (a.players == 0 && f->players_for != (const void *)io->to_play_dev) {
        k_max_to_play_init<<<1, 1, 0, st>>>(p.scratch);
        k_max_to_play<<<std::min((p.N + 255) / 256, 148), 256, 0, st>>>(io->to_play_dev, p.N, p.scratch);
        f->players_for = io->to_play_dev;
    }
    if (p.variant == LZB_GMZ) {
        const int S = cfg->num_simulations, m = cfg->max_num_considered_actions;
        if (S <= 0 || m < 0) return fail(LZB_E_ARG, "Gumbel needs num_simulations > 0 and max_num_considered_actions >= 0");
        if (f->sim >= S) return fail(LZB_E_STATE, "more traverses (%d) than num_simulations (%d)", f->sim + 1, S);
        if (f->vt_m != m || f->vt_s != S) {
            f->vtable_host = considered_sequence(std::min(m, S), S); // row min(m, S) of the table (gmz :724-729)
            const int cnt = std::min(S, p.E);
            LZB_CUDA(cudaMemcpyAsync(p.vtable, f->vtable_host.data(), sizeof(int) * cnt, cudaMemcpyHostToDevice, st));
            f->vt_m = m; f->vt_s = S;
        }
    }
    return LZB_OK;
}

int fill_backup(lzb_forest_t *f, const lzb_backprop_io_t *io, BackupArgs &b)
{
    if (!f || !io) return fail(LZB_E_ARG, "forest / io is NULL");
    const Pool &p = f->pool;
    if (!io->reward_dev || !io->value_dev || !io->to_play_dev) return fail(LZB_E_ARG, "reward / value / to_play is NULL");
    if (p.variant == LZB_SEZ ? !io->sampled_actions_dev : !io->logits_dev) return fail(LZB_E_ARG, "logits (or sampled_actions for the sampled tree) is NULL");
    if (f->sim <= 0) return fail(LZB_E_STATE, "backprop before traverse");
    if (io->latent_index <= 0 || io->latent_index > p.S) return fail(LZB_E_STATE, "latent_index %d outside the pool (max_sims %d)", io->latent_index, p.S);
    if (io->new_slots.n < 0 || io->new_slots.n > kMaxGathers) return fail(LZB_E_ARG, "0..3 slot pointers supported");
    b = BackupArgs{io->latent_index, io->lstm_horizon_len, io->discount, io->reward_dev, io->value_dev, io->logits_dev,
                   io->sampled_actions_dev, io->to_play_dev, io->is_reset_dev, io->is_reset_out_dev, io->new_slots.n, {}};
    for (int k = 0; k < io->new_slots.n; ++k) b.new_slot[k] = (const char *)io->new_slots.ptr[k];
    b.active = io->active_dev;
    if (io->chance_head) {
        const lzb_chance_head_t &h = *io->chance_head;
        if (p.variant != LZB_SMZ) return fail(LZB_E_ARG, "chance_head exists for the stochastic tree only");
        if (!h.reward_dev || !h.value_dev || !h.logits_dev || h.logits_stride < p.C || (h.decision_logits_stride && h.decision_logits_stride < p.Ad))
            return fail(LZB_E_ARG, "chance_head: NULL outputs or a logits stride below the node width");
        b.reward_c = h.reward_dev; b.value_c = h.value_dev; b.logits_c = h.logits_dev;
        b.logits_c_stride = h.logits_stride; b.logits_stride = h.decision_logits_stride;
        if (h.latent_dev) { // slot row 1 of the new node = its rows if it is a chance node
            if (io->new_slots.n != 1) return fail(LZB_E_ARG, "chance_head.latent_dev goes with exactly one new_slots pointer (the decision head's rows)");
            b.new_slot[1] = (const char *)h.latent_dev;
            b.n_slots = 2;
            f->slot_by_kind = true;
        }
    }
    if (io->reuse) {
        if (int rc = check_reuse(p)) return rc;
        if (!io->reuse->reuse_value_dev || (!io->reuse->flags_dev && !io->reuse->true_action_dev))
            return fail(LZB_E_ARG, "reuse: reuse_value is NULL, or neither flags nor true_action given");
        b.true_action = io->reuse->true_action_dev; b.reuse_value = io->reuse->reuse_value_dev;
        b.reuse_flags = io->reuse->flags_dev; b.batch_index = io->reuse->batch_index_dev;
    }
    return LZB_OK;
}

// dynamic shared memory per warp: selection stages A child records + A scores (Gumbel: 3 A-vectors), backup A floats
size_t select_bytes_per_warp(const Pool &p) { return (size_t)p.A * (p.variant == LZB_GMZ ? 28 : 20); }

// Lane-group kernels (lzb_group.cuh): G lanes per tree, R child records per lane, chosen from the action space size.
enum { kSelect = 0, kBackup = 1, kStep = 2 };

template <int VAR, int G, int R, bool REUSE>
void launch_group(int which, const Pool &p, const BackupArgs *b, const SelectArgs *a, cudaStream_t st)
{
    const int per_block = 4 * (32 / G); // four warps of 32 / G trees (lzb_group.cuh: group_tree)
    // programmatic dependent launch: the kernel may be scheduled while its predecessor drains; it reads nothing
    // before its griddepcontrol.wait (LZB_NO_PDL=1 launches it the ordinary way)
    static const bool pdl = getenv("LZB_NO_PDL") == nullptr;
    cudaLaunchAttribute attr[2];
    int n_attr = 0;
    if (pdl) {
        attr[n_attr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n_attr].val.programmaticStreamSerializationAllowed = 1;
        n_attr += 1;
    }
    // experiment (LZB_L2_PERSIST_MB=<MB>, default off): the node pool as a persisting L2 window, so that the 2 x 4 KB of
    // leaf rows per tree streamed through L2 every step do not evict the blocks the next step walks
    static const long long persist_mb = getenv("LZB_L2_PERSIST_MB") ? atoll(getenv("LZB_L2_PERSIST_MB")) : 0;
    if (persist_mb > 0) {
        static bool limit_set = false;
        if (!limit_set) { cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)persist_mb << 20); limit_set = true; }
        int max_win = 0, dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
        const size_t pool_bytes = (size_t)p.N * p.E * p.B * sizeof(int4);
        const size_t win = std::min(pool_bytes, (size_t)max_win);
        attr[n_attr].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[n_attr].val.accessPolicyWindow.base_ptr = (void *)p.node;
        attr[n_attr].val.accessPolicyWindow.num_bytes = win;
        attr[n_attr].val.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)((size_t)persist_mb << 20) / (double)win);
        attr[n_attr].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[n_attr].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        n_attr += 1;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((p.N + per_block - 1) / per_block);
#ifndef LZB_X_NO_SHARED_TABLES
    cfg.dynamicSmemBytes = 64 * sizeof(unsigned long long); // the two libm tables (fill_shared_tables)
#ifndef LZB_X_NO_LONE_STAGE
    // lone walker: one mbarrier and one staged row per warp (lzb_group.cuh, lone_stage_rows), only where the kernel can use them
    if (which != kBackup && a && LoneWalker<VAR, G, R, REUSE>::value && VAR != LZB_SMZ && VAR != LZB_GMZ && lone_stage_rows(*a))
        cfg.dynamicSmemBytes = kStageRowWord * 8 + 4 * (size_t)a->g[0].row_bytes;
#endif
#endif
    cfg.blockDim = dim3(128);
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = n_attr;
    if (which == kSelect) cudaLaunchKernelEx(&cfg, k_select_g<VAR, G, R, REUSE>, p, *a);
    else if (which == kBackup) cudaLaunchKernelEx(&cfg, k_expand_backup_g<VAR, G, R, REUSE>, p, *b);
    else cudaLaunchKernelEx(&cfg, k_step_g<VAR, G, R, REUSE>, p, *b, *a);
}

// Lanes per tree G and child records per lane R (G * R >= A).  Fewer records per lane = a shorter dependent chain
// per tree level, more lanes per tree = more warps; past one resident wave (148 SMs x 32 warps of these kernels)
// extra warps only queue, so the widest group that still fits one wave is taken, else the narrowest.
template <int VAR, bool REUSE>
bool launch_group_for(int which, const Pool &p, const BackupArgs *b, const SelectArgs *a, cudaStream_t st)
{
    const int A = p.A;
    // tuning hook: 8, 10, 16 or 32 lanes per tree (LZB_GROUP_LANES=0 / unset: chosen below)
    static const int forced = getenv("LZB_GROUP_LANES") ? atoi(getenv("LZB_GROUP_LANES")) : 0;
    const long long one_wave = 148LL * 28; // warps: __launch_bounds__(128, 7) keeps 7 blocks of 4 warps resident per SM
    auto fits = [&](int G) { return forced ? forced == G : ((long long)p.N + (32 / G) - 1) / (32 / G) <= one_wave; };
    if (A <= 8) {
        launch_group<VAR, 8, 1, REUSE>(which, p, b, a, st);
    } else if (A <= 10 && forced == 10) {
        launch_group<VAR, 10, 1, REUSE>(which, p, b, a, st); // experiment: three trees per warp
    } else if (A <= 16) {
        if (fits(16)) launch_group<VAR, 16, 1, REUSE>(which, p, b, a, st);
        else launch_group<VAR, 8, 2, REUSE>(which, p, b, a, st);
    } else if (A <= 20 && forced == 10) {
        // experiment (round 2, measured and rejected as a default: 22.4 vs 17.4 us per step at 8192 x A=18):
        // 10 lanes x 2 records = the chain of 16 x 2 with two thirds of the warps, three trees per warp
        launch_group<VAR, 10, 2, REUSE>(which, p, b, a, st);
    } else if (A <= 24) {
        if (fits(32)) launch_group<VAR, 32, 1, REUSE>(which, p, b, a, st);
        else if (fits(16)) launch_group<VAR, 16, 2, REUSE>(which, p, b, a, st);
        else launch_group<VAR, 8, 3, REUSE>(which, p, b, a, st);
    } else if (A <= 32) {
        if (fits(32)) launch_group<VAR, 32, 1, REUSE>(which, p, b, a, st);
        else launch_group<VAR, 16, 2, REUSE>(which, p, b, a, st);
    } else if (A <= 64) {
        launch_group<VAR, 32, 2, REUSE>(which, p, b, a, st);
    } else if (A <= 128) {
        launch_group<VAR, 32, 4, REUSE>(which, p, b, a, st);
    } else {
        return false; // wider action spaces: warp-per-tree kernels with shared-memory staging
    }
    return true;
}

// true when a lane-group kernel was launched for this forest's flavour / width
bool launch_grouped(int which, const Pool &p, const BackupArgs *b, const SelectArgs *a, cudaStream_t st)
{
    if (getenv("LZB_FORCE_WARP_PER_TREE")) return false; // test hook: exercise the generic kernels on small A
    const bool reuse = (a && a->true_action) || (b && b->reuse_value);
    switch (p.variant) {
    case LZB_MZ: return reuse ? launch_group_for<LZB_MZ, true>(which, p, b, a, st) : launch_group_for<LZB_MZ, false>(which, p, b, a, st);
    case LZB_EZ: return reuse ? launch_group_for<LZB_EZ, true>(which, p, b, a, st) : launch_group_for<LZB_EZ, false>(which, p, b, a, st);
    case LZB_SEZ: return launch_group_for<LZB_SEZ, false>(which, p, b, a, st);
    case LZB_SMZ: return launch_group_for<LZB_SMZ, false>(which, p, b, a, st);
    default: return launch_group_for<LZB_GMZ, false>(which, p, b, a, st);
    }
}

} // namespace

extern "C" {

int lzb_set_slots(lzb_forest_t *f, int latent_index, const lzb_slots_t *slots, void *stream)
{
    if (!f || !slots) return fail(LZB_E_ARG, "forest / slots is NULL");
    if (latent_index < 0 || latent_index > f->pool.S) return fail(LZB_E_ARG, "latent_index outside the pool");
    if (slots->n < 0 || slots->n > kMaxGathers) return fail(LZB_E_ARG, "0..3 slot pointers supported");
    k_set_slots<<<1, 32, 0, (cudaStream_t)stream>>>(f->pool, latent_index, slots->n, (const char *)slots->ptr[0],
                                                   (const char *)slots->ptr[1], (const char *)slots->ptr[2]);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_traverse(lzb_forest_t *f, const lzb_search_cfg_t *cfg, const lzb_traverse_io_t *io, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    SelectArgs a;
    if (int rc = fill_select(f, cfg, io, a, st)) return rc;
    const Pool &p = f->pool;
    dim3 grid, block; size_t smem;
    const size_t bpw = select_bytes_per_warp(p);
    if (launch_grouped(kSelect, p, nullptr, &a, st)) {
        LZB_CUDA(cudaGetLastError());
        f->sim += 1;
        return LZB_OK;
    }
    if (p.variant == LZB_SMZ) return fail(LZB_E_UNSUPPORTED, "the stochastic tree has lane-group kernels only");
    switch (p.variant) {
    case LZB_GMZ:
        if (int rc = launch_cfg(k_select_gumbel, bpw, p.N, grid, block, smem)) return rc;
        k_select_gumbel<<<grid, block, smem, st>>>(p, a);
        break;
    case LZB_MZ:
        if (int rc = launch_cfg(k_select_puct<LZB_MZ>, bpw, p.N, grid, block, smem)) return rc;
        k_select_puct<LZB_MZ><<<grid, block, smem, st>>>(p, a);
        break;
    case LZB_EZ:
        if (int rc = launch_cfg(k_select_puct<LZB_EZ>, bpw, p.N, grid, block, smem)) return rc;
        k_select_puct<LZB_EZ><<<grid, block, smem, st>>>(p, a);
        break;
    default:
        if (int rc = launch_cfg(k_select_puct<LZB_SEZ>, bpw, p.N, grid, block, smem)) return rc;
        k_select_puct<LZB_SEZ><<<grid, block, smem, st>>>(p, a);
        break;
    }
    LZB_CUDA(cudaGetLastError());
    f->sim += 1;
    return LZB_OK;
}

int lzb_backprop(lzb_forest_t *f, const lzb_backprop_io_t *io, void *stream)
{
    BackupArgs a;
    if (int rc = fill_backup(f, io, a)) return rc;
    const Pool &p = f->pool;
    dim3 grid, block; size_t smem;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bpw = (size_t)p.A * sizeof(float);
    if (launch_grouped(kBackup, p, &a, nullptr, st)) {
        LZB_CUDA(cudaGetLastError());
        if (io->latent_index > f->expanded) f->expanded = io->latent_index;
        return LZB_OK;
    }
    if (p.variant == LZB_SMZ) return fail(LZB_E_UNSUPPORTED, "the stochastic tree has lane-group kernels only");
    if (io->active_dev) return fail(LZB_E_UNSUPPORTED, "backprop: the active mask needs the lane-group kernels (action space <= 128)");
    switch (p.variant) {
    case LZB_MZ:
        if (int rc = launch_cfg(k_expand_backup<LZB_MZ>, bpw, p.N, grid, block, smem)) return rc;
        k_expand_backup<LZB_MZ><<<grid, block, smem, st>>>(p, a);
        break;
    case LZB_EZ:
        if (int rc = launch_cfg(k_expand_backup<LZB_EZ>, bpw, p.N, grid, block, smem)) return rc;
        k_expand_backup<LZB_EZ><<<grid, block, smem, st>>>(p, a);
        break;
    case LZB_GMZ:
        if (int rc = launch_cfg(k_expand_backup<LZB_GMZ>, bpw, p.N, grid, block, smem)) return rc;
        k_expand_backup<LZB_GMZ><<<grid, block, smem, st>>>(p, a);
        break;
    default:
        if (int rc = launch_cfg(k_expand_backup<LZB_SEZ>, bpw, p.N, grid, block, smem)) return rc;
        k_expand_backup<LZB_SEZ><<<grid, block, smem, st>>>(p, a);
        break;
    }
    LZB_CUDA(cudaGetLastError());
    if (io->latent_index > f->expanded) f->expanded = io->latent_index;
    return LZB_OK;
}

int lzb_step(lzb_forest_t *f, const lzb_search_cfg_t *cfg, const lzb_backprop_io_t *bio, const lzb_traverse_io_t *tio,
             void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    BackupArgs b;
    SelectArgs a;
    if (int rc = fill_backup(f, bio, b)) return rc;
    if (int rc = fill_select(f, cfg, tio, a, st)) return rc;
    if ((bio->reuse != nullptr) != (tio->reuse != nullptr)) return fail(LZB_E_ARG, "step: reuse must be given for both halves or neither");
    if (bio->active_dev) return fail(LZB_E_ARG, "step: an active mask is accepted by lzb_backprop only");
    a.new_e = b.latent_index;
    for (int k = 0; k < kMaxGathers; ++k) a.new_slot[k] = b.new_slot[k];
    for (int k = 0; k < a.ng; ++k)
        if (!a.g[k].src && k >= b.n_slots) return fail(LZB_E_ARG, "gather %d reads the slot table but the step carries no slot pointer for it", k);
    const Pool &p = f->pool;
    dim3 grid, block; size_t smem;
    const size_t bpw = select_bytes_per_warp(p);
    if (launch_grouped(kStep, p, &b, &a, st)) {
        LZB_CUDA(cudaGetLastError());
        if (bio->latent_index > f->expanded) f->expanded = bio->latent_index;
        f->sim += 1;
        return LZB_OK;
    }
    if (p.variant == LZB_SMZ) return fail(LZB_E_UNSUPPORTED, "the stochastic tree has lane-group kernels only");
    switch (p.variant) {
    case LZB_MZ:
        if (int rc = launch_cfg(k_step<LZB_MZ>, bpw, p.N, grid, block, smem)) return rc;
        k_step<LZB_MZ><<<grid, block, smem, st>>>(p, b, a);
        break;
    case LZB_EZ:
        if (int rc = launch_cfg(k_step<LZB_EZ>, bpw, p.N, grid, block, smem)) return rc;
        k_step<LZB_EZ><<<grid, block, smem, st>>>(p, b, a);
        break;
    case LZB_GMZ:
        if (int rc = launch_cfg(k_step<LZB_GMZ>, bpw, p.N, grid, block, smem)) return rc;
        k_step<LZB_GMZ><<<grid, block, smem, st>>>(p, b, a);
        break;
    default:
        if (int rc = launch_cfg(k_step<LZB_SEZ>, bpw, p.N, grid, block, smem)) return rc;
        k_step<LZB_SEZ><<<grid, block, smem, st>>>(p, b, a);
        break;
    }
    LZB_CUDA(cudaGetLastError());
    if (bio->latent_index > f->expanded) f->expanded = bio->latent_index;
    f->sim += 1;
    return LZB_OK;
}

int lzb_store_hidden(const float *src_dev, float *dst_dev, const int32_t *is_reset_dev, int num_roots,
                     int64_t row_floats, void *stream)
{
    if (!src_dev || !dst_dev || num_roots <= 0 || row_floats <= 0) return fail(LZB_E_ARG, "store_hidden: bad argument");
    const bool vec = (row_floats % 4 == 0) && ((((uintptr_t)src_dev) | ((uintptr_t)dst_dev)) & 15) == 0;
    const long long total = (long long)num_roots * (vec ? row_floats / 4 : row_floats);
    const int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 8);
    if (vec) k_store_hidden<true><<<blocks, 256, 0, (cudaStream_t)stream>>>(src_dev, dst_dev, is_reset_dev, num_roots, row_floats);
    else k_store_hidden<false><<<blocks, 256, 0, (cudaStream_t)stream>>>(src_dev, dst_dev, is_reset_dev, num_roots, row_floats);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_root_outputs(lzb_forest_t *f, int32_t *visits_dev, int32_t *legal_count_dev, float *values_dev,
                     float discount, float *gumbel_policies_dev, float *gumbel_children_values_dev, void *stream)
{
    if (!f) return fail(LZB_E_ARG, "forest is NULL");
    const Pool &p = f->pool;
    if ((gumbel_policies_dev || gumbel_children_values_dev) && p.variant != LZB_GMZ)
        return fail(LZB_E_ARG, "policies / children values exist only for the Gumbel tree");
    dim3 grid, block; size_t smem;
    if (int rc = launch_cfg(k_root_outputs, (size_t)p.A * 28, p.N, grid, block, smem)) return rc;
    k_root_outputs<<<grid, block, smem, (cudaStream_t)stream>>>(p, visits_dev, legal_count_dev, values_dev, discount,
                                                               gumbel_policies_dev, gumbel_children_values_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_dump_tree(lzb_forest_t *f, int tree, int32_t *visit_dev, float *value_sum_dev, float *prior_dev,
                  float *reward_dev, int32_t *child_e_dev, int32_t *hdr_dev, float *raw_value_dev, void *stream)
{
    if (!f || tree < 0 || tree >= f->pool.N) return fail(LZB_E_ARG, "dump_tree: bad tree index");
    if (f->expanded < 0) return fail(LZB_E_STATE, "dump_tree before prepare");
    k_dump_tree<<<1, 32, 0, (cudaStream_t)stream>>>(f->pool, tree, f->expanded, visit_dev, value_sum_dev, prior_dev,
                                                   reward_dev, child_e_dev, hdr_dev, raw_value_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_dump_paths(lzb_forest_t *f, int32_t *path_actions_dev, int path_stride, void *stream)
{
    if (!f || !path_actions_dev || path_stride <= 0) return fail(LZB_E_ARG, "dump_paths: bad argument");
    k_dump_paths<<<(f->pool.N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(f->pool, path_actions_dev, path_stride);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_trajectories(lzb_forest_t *f, int32_t *actions_dev, int stride, int32_t *len_dev, void *stream)
{
    if (!f || !actions_dev || !len_dev || stride <= 0) return fail(LZB_E_ARG, "trajectories: bad argument");
    if (f->expanded < 0) return fail(LZB_E_STATE, "trajectories before prepare");
    k_trajectories<<<(f->pool.N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(f->pool, actions_dev, stride, len_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_minmax(lzb_forest_t *f, float *min_dev, float *max_dev, void *stream)
{
    if (!f || !min_dev || !max_dev) return fail(LZB_E_ARG, "minmax: bad argument");
    k_minmax<<<(f->pool.N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(f->pool, min_dev, max_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_check_word(lzb_forest_t *f, int32_t *word_host, int32_t *checked_build, void *stream)
{
    if (!f || !word_host) return fail(LZB_E_ARG, "check_word: bad argument");
    LZB_CUDA(cudaMemcpyAsync(word_host, f->pool.scratch + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    LZB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (checked_build) {
#ifdef LZB_CHECKED
        *checked_build = 1;
#else
        *checked_build = 0;
#endif
    }
    return LZB_OK;
}

int lzb_gumbel_helper(const lzb_gumbel_helper_t *h, void *stream)
{
    if (!h || h->n <= 0 || h->op < GH_SOFTMAX || h->op > GH_SELECT_INTERIOR || !h->tmp_dev) return fail(LZB_E_ARG, "gumbel_helper: bad argument");
    const bool needs_out_i = h->op == GH_SELECT_ROOT || h->op == GH_SELECT_INTERIOR;
    if (needs_out_i ? !h->out_i_dev : !h->out_dev) return fail(LZB_E_ARG, "gumbel_helper: the op's output pointer is NULL");
    if (!h->a_dev) return fail(LZB_E_ARG, "gumbel_helper: a_dev is NULL");
    const bool needs_vb = h->op == GH_MIXED_VALUE || h->op == GH_QTRANSFORM || h->op == GH_SCORE_CONSIDERED || needs_out_i;
    if (needs_vb && (!h->v_dev || !h->b_dev)) return fail(LZB_E_ARG, "gumbel_helper: the op needs visit counts and a second vector");
    if ((h->op == GH_SCORE_CONSIDERED || h->op == GH_SELECT_ROOT) && !h->c_dev) return fail(LZB_E_ARG, "gumbel_helper: the op needs a third vector");
    GumbelHelperArgs a{h->op, h->n, h->a_dev, h->b_dev, h->c_dev, h->v_dev, h->s0, h->s1, h->s2, h->s3, h->i0, h->i1, h->out_dev, h->out_i_dev, h->tmp_dev};
    k_gumbel_helper<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_root_stats(lzb_forest_t *f, int32_t *visit_dev, float *value_sum_dev, float *reward_dev, void *stream)
{
    if (!f || !visit_dev || !value_sum_dev || !reward_dev) return fail(LZB_E_ARG, "root_stats: bad argument");
    k_root_stats<<<(f->pool.N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(f->pool, visit_dev, value_sum_dev, reward_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_sample_actions_seeded(const float *logits_dev, const float *uniforms_dev, float *out_dev, int num_roots, int num_sampled,
                              int action_dim, int action_space, int continuous, int root_offset, uint64_t seed,
                              const uint64_t *seed_dev, uint32_t counter, void *stream)
{
    if (!logits_dev || !out_dev || num_roots <= 0 || num_sampled <= 0) return fail(LZB_E_ARG, "sample_actions: bad argument");
    if (continuous ? action_dim <= 0 : (action_space <= 0 || action_space > 256 || action_dim != 1))
        return fail(LZB_E_ARG, "sample_actions: continuous needs action_dim > 0; discrete needs 0 < action_space <= 256 and action_dim == 1");
    if (!continuous && num_sampled > action_space) return fail(LZB_E_ARG, "sample_actions: cannot draw %d of %d actions without replacement", num_sampled, action_space);
    SampleArgs a{logits_dev, uniforms_dev, out_dev, num_roots, num_sampled, action_dim, action_space, continuous, root_offset, seed, counter, (const unsigned long long *)seed_dev};
    k_sample_actions<<<(num_roots + 3) / 4, 128, 0, (cudaStream_t)stream>>>(a);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_sample_actions_libstdcxx(const float *logits_dev, float *out_dev, int num_roots, int num_sampled, int action_dim,
                                 int action_space, int continuous, uint64_t seed_base, void *stream)
{
    if (!logits_dev || !out_dev || num_roots <= 0 || num_sampled <= 0) return fail(LZB_E_ARG, "sample_actions_libstdcxx: bad argument");
    if (continuous ? action_dim <= 0 : (action_space <= 0 || action_space > 256 || action_dim != 1))
        return fail(LZB_E_ARG, "sample_actions_libstdcxx: continuous needs action_dim > 0; discrete needs 0 < action_space <= 256 and action_dim == 1");
    if (!continuous && num_sampled > action_space) return fail(LZB_E_ARG, "sample_actions_libstdcxx: cannot draw %d of %d actions without replacement", num_sampled, action_space);
    ParitySampleArgs a{logits_dev, out_dev, num_roots, num_sampled, action_dim, action_space, continuous, seed_base};
    k_sample_actions_libstdcxx<<<(num_roots + 63) / 64, 64, 0, (cudaStream_t)stream>>>(a);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_sample_actions(const float *logits_dev, const float *uniforms_dev, float *out_dev, int num_roots, int num_sampled,
                       int action_dim, int action_space, int continuous, int root_offset, uint64_t seed, uint32_t counter,
                       void *stream)
{
    return lzb_sample_actions_seeded(logits_dev, uniforms_dev, out_dev, num_roots, num_sampled, action_dim, action_space,
                                     continuous, root_offset, seed, nullptr, counter, stream);
}

int lzb_legal_from_mask(const int8_t *mask_dev, int num_roots, int action_space, int32_t *legal_count_dev,
                        int32_t *legal_actions_dev, void *stream)
{
    if (!mask_dev || !legal_count_dev || !legal_actions_dev || num_roots <= 0 || action_space <= 0) return fail(LZB_E_ARG, "legal_from_mask: bad argument");
    k_legal_from_mask<<<(num_roots + 3) / 4, 128, 0, (cudaStream_t)stream>>>((const signed char *)mask_dev, num_roots, action_space,
                                                                             legal_count_dev, legal_actions_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_dirichlet_noise(float alpha, const int32_t *legal_count_dev, float *noise_dev, int num_roots, int action_space,
                        int root_offset, uint64_t seed, uint32_t counter, void *stream)
{
    if (!noise_dev || num_roots <= 0 || action_space <= 0 || !(alpha > 0.0f)) return fail(LZB_E_ARG, "dirichlet_noise: bad argument");
    k_dirichlet_noise<<<(num_roots + 3) / 4, 128, 0, (cudaStream_t)stream>>>(alpha, legal_count_dev, noise_dev, num_roots, action_space,
                                                                             root_offset, seed, counter);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_select_action(const lzb_select_action_t *args, void *stream)
{
    if (!args || !args->visits_dev || args->num_roots <= 0 || args->action_space <= 0) return fail(LZB_E_ARG, "select_action: bad argument");
    if (!(args->temperature > 0.0)) return fail(LZB_E_ARG, "select_action: temperature must be positive");
    SelectActionArgs a{args->visits_dev, args->legal_count_dev, args->legal_actions_dev, args->uniforms_dev, args->action_pos_dev,
                       args->action_dev, args->entropy_dev, args->temperature, args->num_roots, args->action_space,
                       args->deterministic, args->root_offset, args->seed, args->counter};
    k_select_action<<<(args->num_roots + 127) / 128, 128, 0, (cudaStream_t)stream>>>(a);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_target_policies(const int32_t *visits_dev, const int32_t *legal_count_dev, const int32_t *legal_actions_dev,
                        const int32_t *policy_mask_dev, int num_roots, int action_space, float *policy_dev, void *stream)
{
    if (!visits_dev || !policy_dev || num_roots <= 0 || action_space <= 0) return fail(LZB_E_ARG, "target_policies: bad argument");
    k_target_policies<<<(num_roots + 3) / 4, 128, 0, (cudaStream_t)stream>>>(visits_dev, legal_count_dev, legal_actions_dev, policy_mask_dev,
                                                                             num_roots, action_space, policy_dev);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_value_targets(const lzb_value_targets_t *x, void *stream)
{
    if (!x || !x->root_values_dev || !x->value_mask_dev || !x->rewards_dev || !x->td_steps_dev || !x->reward_len_dev || !x->segment_len_dev ||
        !x->pos_dev || !x->pow_table_dev || !x->target_values_dev || !x->target_rewards_dev)
        return fail(LZB_E_ARG, "value_targets: NULL argument");
    if (x->num_segments <= 0 || x->num_unroll_steps < 0 || x->reward_stride <= 0 || x->pow_len <= 0)
        return fail(LZB_E_ARG, "value_targets: num_segments, reward_stride and pow_len must be positive");
    if (x->board_games && (!x->to_play_dev || x->to_play_stride <= 0)) return fail(LZB_E_ARG, "value_targets: board_games needs to_play [B, to_play_stride]");
    ValueTargetArgs a{x->root_values_dev, x->td_steps_dev, x->value_mask_dev, x->rewards_dev, x->reward_len_dev, x->segment_len_dev, x->pos_dev,
                      x->to_play_dev, x->pow_table_dev, x->target_values_dev, x->target_rewards_dev, x->num_segments, x->num_unroll_steps,
                      x->reward_stride, x->to_play_stride, x->pow_len, x->board_games, x->lstm_horizon_len};
    if (x->lstm_horizon_len < 0) return fail(LZB_E_ARG, "value_targets: lstm_horizon_len must be >= 0");
    const long long total = (long long)x->num_segments * (x->num_unroll_steps + 1);
    k_value_targets<<<(int)((total + 127) / 128), 128, 0, (cudaStream_t)stream>>>(a);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

namespace {
// Grid of the scalar-transform kernels: one row (or root) per warp, in as many waves of blocks as that takes.  16,384 rows are
// 2.31 waves of 148 SMs x 48 resident warps and ncu shows the SMs active for 72 % of the launch, so a balanced grid was tried
// (-DLZB_X_SCALAR_BALANCED: P = ceil(units / resident warps) passes over ceil(units / P) warps, every warp walks P rows and
// all end together): 14.6-14.9 us against 14.5 us - no gain, because at 39 MB the launch is in the small-kernel regime where
// even torch's plain row sum over the same bytes takes 13.6 us (profiles/r02_scalar_transform.md).
void scalar_grid(long long units, int blocks_per_sm, long long wpb, int &blocks, int &iters)
{
#ifndef LZB_X_SCALAR_BALANCED
    (void)blocks_per_sm;
    blocks = (int)std::min<long long>((units + wpb - 1) / wpb, 1LL << 22);
    iters = (int)((units + wpb * blocks - 1) / (wpb * blocks));
#else
    static int sms = 0;
    if (sms == 0) {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        sms = n;
    }
    const long long resident = (long long)sms * blocks_per_sm * wpb;
    const long long passes = std::max<long long>(1, (units + resident - 1) / resident);
    const long long warps = (units + passes - 1) / passes;
    blocks = (int)((warps + wpb - 1) / wpb);
    iters = (int)((units + wpb * blocks - 1) / (wpb * blocks));
#endif
}

// LZB_SCALAR_THREADS / 32 warps per block (lzb_scaling.cuh; -DLZB_X_SCALAR_PIPELINED: 4 persistent blocks per SM)
int launch_scalar(const ScalarHeads &a, int W, long long rows, cudaStream_t st)
{
    constexpr long long wpb = kScalarThreads / 32;
    int blocks = 0, iters = 0;
#if defined(LZB_X_SCALAR_PAIR) && !defined(LZB_X_SCALAR_PIPELINED) && !defined(LZB_X_SCALAR_ROWLOADS)
    if (a.n_heads == 2 && W <= 32 * 19) { // experiment: both heads of a root by one warp, packed fp32 arithmetic (lzb_scaling.cuh)
        scalar_grid(rows / 2, LZB_SCALAR_PAIR_MIN_BLOCKS, wpb, blocks, iters);
        if (W <= 32 * 4) k_inverse_scalar_transform_pair<4><<<blocks, kScalarThreads, 0, st>>>(a, iters);
        else if (W <= 32 * 8) k_inverse_scalar_transform_pair<8><<<blocks, kScalarThreads, 0, st>>>(a, iters);
        else k_inverse_scalar_transform_pair<19><<<blocks, kScalarThreads, 0, st>>>(a, iters);
        LZB_CUDA(cudaGetLastError());
        return LZB_OK;
    }
#endif
#ifdef LZB_X_SCALAR_PIPELINED
    blocks = (int)std::min<long long>((rows + wpb - 1) / wpb, 148LL * 4);
    iters = (int)((rows + wpb * blocks - 1) / (wpb * blocks));
#else
    scalar_grid(rows, LZB_SCALAR_MIN_BLOCKS, wpb, blocks, iters);
#endif
    if (W <= 32 * 4) k_inverse_scalar_transform<4><<<blocks, kScalarThreads, 0, st>>>(a, iters);
    else if (W <= 32 * 8) k_inverse_scalar_transform<8><<<blocks, kScalarThreads, 0, st>>>(a, iters);
    else if (W <= 32 * 19) k_inverse_scalar_transform<19><<<blocks, kScalarThreads, 0, st>>>(a, iters); // support 300: 601 logits per row
    else k_inverse_scalar_transform<0><<<blocks, kScalarThreads, 0, st>>>(a, iters);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}
} // namespace

int lzb_scalar_workspace_bytes(int n_heads, int num_rows, size_t *bytes)
{
    if (!bytes || n_heads <= 0 || n_heads > kMaxScalarHeads || num_rows <= 0) return fail(LZB_E_ARG, "scalar_workspace_bytes: 1..%d heads, num_rows > 0", kMaxScalarHeads);
    *bytes = 16 + (size_t)n_heads * (size_t)num_rows * sizeof(float);
    return LZB_OK;
}

int lzb_inverse_scalar_transform(const lzb_support_head_t *heads, int n_heads, int num_rows, int support_size, float epsilon,
                                 void *workspace_dev, void *stream)
{
    if (!heads || n_heads <= 0 || n_heads > kMaxScalarHeads || num_rows <= 0 || support_size <= 0 || !workspace_dev)
        return fail(LZB_E_ARG, "inverse_scalar_transform: 1..%d heads, num_rows > 0, support_size > 0, workspace required", kMaxScalarHeads);
    if (((uintptr_t)workspace_dev & 15) != 0) return fail(LZB_E_ARG, "inverse_scalar_transform: workspace must be 16-byte aligned");
    if (!(epsilon > 0.0f)) return fail(LZB_E_ARG, "inverse_scalar_transform: epsilon must be positive");
    const int W = 2 * support_size + 1;
    ScalarHeads a = {};
    for (int h = 0; h < n_heads; ++h) {
        if (!heads[h].logits_dev || !heads[h].scalar_dev || heads[h].row_stride < W)
            return fail(LZB_E_ARG, "inverse_scalar_transform: head %d has a NULL pointer or a row stride below %d", h, W);
        a.logits[h] = heads[h].logits_dev; a.out[h] = heads[h].scalar_dev; a.stride[h] = heads[h].row_stride;
    }
    a.n_heads = n_heads; a.N = num_rows; a.support = support_size; a.eps = epsilon;
    a.ctl = (unsigned *)workspace_dev; a.raw = (float *)((char *)workspace_dev + 16);
    return launch_scalar(a, W, (long long)n_heads * num_rows, (cudaStream_t)stream);
}

// ---- lzb_inverse_scalar_transform_static: the same kernel with its scratch in a file-scope device array, for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
namespace {
constexpr int kStaticScalarRows = LZB_STATIC_SCALAR_MAX_ROWS;
constexpr size_t kStaticScalarCtlBytes = 256; // ticket + per-head counters, padded so that the scalars behind stay aligned
constexpr size_t kStaticScalarBytes = kStaticScalarCtlBytes + (size_t)kMaxScalarHeads * kStaticScalarRows * sizeof(float);
}
// 524,544 bytes in every GPU context that loads the library, until the process exits.  In HOST code this name is only a
// stand-in for the symbol, not a device address: the address comes from cudaGetSymbolAddress (below), per device.
static __device__ __align__(256) unsigned char g_static_scalar_scratch[kStaticScalarBytes];

namespace {
constexpr int kMaxDevices = 64;
void *g_static_scalar_ptr[kMaxDevices]; // device address of the array on each device, resolved on first use
std::mutex g_static_scalar_mutex;

// the array's address on the CURRENT device (resolved once per device, with that device current)
int static_scalar_scratch(void **out)
{
    int dev = -1;
    LZB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(LZB_E_UNSUPPORTED, "inverse_scalar_transform_static: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_static_scalar_mutex);
    if (!g_static_scalar_ptr[dev]) {
        void *ptr = nullptr;
        LZB_CUDA(cudaGetSymbolAddress(&ptr, g_static_scalar_scratch));
        if (!ptr) return fail(LZB_E_CUDA, "inverse_scalar_transform_static: cudaGetSymbolAddress returned a NULL address");
        g_static_scalar_ptr[dev] = ptr;
    }
    *out = g_static_scalar_ptr[dev];
    return LZB_OK;
}
} // namespace

int lzb_inverse_scalar_transform_static(const lzb_support_head_t *heads, int n_heads, int num_rows, int support_size, float epsilon,
                                        void *stream)
{
    if (!heads || n_heads <= 0 || n_heads > kMaxScalarHeads || num_rows <= 0 || support_size <= 0)
        return fail(LZB_E_ARG, "inverse_scalar_transform_static: 1..%d heads, num_rows > 0, support_size > 0", kMaxScalarHeads);
    if (num_rows > kStaticScalarRows)
        return fail(LZB_E_ARG, "inverse_scalar_transform_static: num_rows %d exceeds the static scratch (LZB_STATIC_SCALAR_MAX_ROWS = %d rows "
                               "per head); use lzb_inverse_scalar_transform with a caller-owned workspace", num_rows, kStaticScalarRows);
    if (!(epsilon > 0.0f)) return fail(LZB_E_ARG, "inverse_scalar_transform_static: epsilon must be positive");
    const int W = 2 * support_size + 1;
    ScalarHeads a = {};
    for (int h = 0; h < n_heads; ++h) {
        if (!heads[h].logits_dev || !heads[h].scalar_dev || heads[h].row_stride < W)
            return fail(LZB_E_ARG, "inverse_scalar_transform_static: head %d has a NULL pointer or a row stride below %d", h, W);
        a.logits[h] = heads[h].logits_dev; a.out[h] = heads[h].scalar_dev; a.stride[h] = heads[h].row_stride;
    }
    void *scratch = nullptr;
    if (int rc = static_scalar_scratch(&scratch)) return rc;
    a.n_heads = n_heads; a.N = num_rows; a.support = support_size; a.eps = epsilon;
    a.ctl = (unsigned *)scratch; a.raw = (float *)((char *)scratch + kStaticScalarCtlBytes);
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps whatever the last call (of any caller, possibly one that was aborted) left in it: the ticket and the
    // per-head counters must start at zero; the scalars behind them are written before they are read
    LZB_CUDA(cudaMemsetAsync(scratch, 0, 16, st));
    return launch_scalar(a, W, (long long)n_heads * num_rows, st);
}

// ---- peer exchange (lzb_peer.cuh).  Layout of one rank's buffer: [2][world*rows][row_words] data, [world] flags, [2] ctl
namespace {
struct PeerLayout { size_t epoch_words, flags_at, ctl_at, total; };
PeerLayout peer_layout(int world, int64_t rows, int row_words)
{
    PeerLayout l;
    l.epoch_words = (size_t)world * (size_t)rows * (size_t)row_words;
    l.flags_at = 2 * l.epoch_words;
    l.ctl_at = l.flags_at + (size_t)world;
    l.total = (l.ctl_at + 2) * sizeof(uint32_t);
    return l;
}
} // namespace

int lzb_peer_bytes(int world, int64_t rows_per_rank, int row_words, size_t *bytes)
{
    if (!bytes || world <= 0 || world > kMaxPeers || rows_per_rank <= 0 || row_words <= 0) return fail(LZB_E_ARG, "peer_bytes: bad argument (1..%d ranks)", kMaxPeers);
    *bytes = peer_layout(world, rows_per_rank, row_words).total;
    return LZB_OK;
}

int lzb_peer_alloc(size_t bytes, void **buf_dev)
{
    if (!buf_dev || bytes == 0) return fail(LZB_E_ARG, "peer_alloc: bad argument");
    LZB_CUDA(cudaMalloc(buf_dev, bytes)); // plain cudaMalloc: shareable through cudaIpcGetMemHandle
    LZB_CUDA(cudaMemset(*buf_dev, 0, bytes));
    LZB_CUDA(cudaDeviceSynchronize());
    return LZB_OK;
}

int lzb_peer_free(void *buf_dev)
{
    if (buf_dev) LZB_CUDA(cudaFree(buf_dev));
    return LZB_OK;
}

int lzb_peer_export(void *buf_dev, void *handle_64_bytes)
{
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!buf_dev || !handle_64_bytes) return fail(LZB_E_ARG, "peer_export: bad argument");
    cudaIpcMemHandle_t h;
    LZB_CUDA(cudaIpcGetMemHandle(&h, buf_dev));
    memcpy(handle_64_bytes, &h, sizeof(h));
    return LZB_OK;
}

int lzb_peer_import(const void *handle_64_bytes, void **buf_dev)
{
    if (!buf_dev || !handle_64_bytes) return fail(LZB_E_ARG, "peer_import: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_64_bytes, sizeof(h));
    LZB_CUDA(cudaIpcOpenMemHandle(buf_dev, h, cudaIpcMemLazyEnablePeerAccess));
    return LZB_OK;
}

int lzb_peer_unimport(void *buf_dev)
{
    if (buf_dev) LZB_CUDA(cudaIpcCloseMemHandle(buf_dev));
    return LZB_OK;
}

int lzb_peer_allgather(void *const *bufs_host, int rank, int world, int64_t rows_per_rank, const void *const *pieces_dev,
                       const int32_t *piece_words, int n_pieces, uint32_t epoch, int64_t timeout_ms, void **result_dev,
                       void *stream)
{
    if (!bufs_host || !pieces_dev || !piece_words || !result_dev) return fail(LZB_E_ARG, "peer_allgather: NULL argument");
    if (world <= 0 || world > kMaxPeers || rank < 0 || rank >= world || rows_per_rank <= 0) return fail(LZB_E_ARG, "peer_allgather: bad rank / world / rows");
    if (n_pieces <= 0 || n_pieces > kMaxPieces) return fail(LZB_E_ARG, "peer_allgather: 1..%d pieces", kMaxPieces);
    if (epoch == 0) return fail(LZB_E_ARG, "peer_allgather: epochs start at 1");
    PeerArgs a = {};
    for (int k = 0; k < n_pieces; ++k) {
        if (!pieces_dev[k] || piece_words[k] <= 0) return fail(LZB_E_ARG, "peer_allgather: piece %d is empty", k);
        a.piece[k] = (const uint32_t *)pieces_dev[k];
        a.words[k] = piece_words[k];
        a.row_words += piece_words[k];
    }
    const PeerLayout l = peer_layout(world, rows_per_rank, a.row_words);
    for (int w = 0; w < world; ++w) {
        if (!bufs_host[w]) return fail(LZB_E_ARG, "peer_allgather: buffer of rank %d is NULL", w);
        uint32_t *base = (uint32_t *)bufs_host[w];
        a.data[w] = base + (size_t)(epoch & 1u) * l.epoch_words;
        a.flags[w] = base + l.flags_at;
    }
    a.ctl = (uint32_t *)bufs_host[rank] + l.ctl_at;
    a.n_pieces = n_pieces; a.rows = (int)rows_per_rank; a.rank = rank; a.world = world; a.epoch = epoch;
    a.timeout_ns = (timeout_ms > 0 ? timeout_ms : 30000) * 1000000LL;
    const long long total = (long long)rows_per_rank * a.row_words;
    const int blocks = (int)std::min<long long>((total + 255) / 256, 128);
    k_peer_allgather<<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
    LZB_CUDA(cudaGetLastError());
    *result_dev = a.data[rank];
    return LZB_OK;
}

int lzb_peer_error(void *own_buf_dev, int world, int64_t rows_per_rank, int row_words, int *error_host, void *stream)
{
    if (!own_buf_dev || !error_host) return fail(LZB_E_ARG, "peer_error: bad argument");
    const PeerLayout l = peer_layout(world, rows_per_rank, row_words);
    uint32_t e = 0;
    LZB_CUDA(cudaMemcpyAsync(&e, (uint32_t *)own_buf_dev + l.ctl_at + 1, sizeof(e), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    LZB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    *error_host = (int)e;
    if (e != 0) // reported once: the word is cleared so that a later failure can be told from this one
        LZB_CUDA(cudaMemsetAsync((uint32_t *)own_buf_dev + l.ctl_at + 1, 0, sizeof(uint32_t), (cudaStream_t)stream));
    return LZB_OK;
}

int lzb_gumbel_table(int m, int S, int32_t *out_host)
{
    if (m < 0 || S <= 0 || !out_host) return fail(LZB_E_ARG, "gumbel_table: bad argument");
    for (int r = 0; r <= m; ++r) {
        std::vector<int> seq = considered_sequence(r, S);
        memcpy(out_host + (size_t)r * S, seq.data(), sizeof(int) * S);
    }
    return LZB_OK;
}

int lzb_gumbel_noise(float scale, int shape, float *out_host)
{
    if (shape < 0 || !out_host) return fail(LZB_E_ARG, "gumbel_noise: bad argument");
    gumbel_noise(scale, shape, out_host);
    return LZB_OK;
}

int lzb_test_expf(const float *x_dev, float *out_dev, int64_t n, void *stream)
{
    k_test_expf<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(x_dev, out_dev, (long long)n);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

int lzb_test_logf(const float *x_dev, float *out_dev, int64_t n, void *stream)
{
    k_test_logf<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(x_dev, out_dev, (long long)n);
    LZB_CUDA(cudaGetLastError());
    return LZB_OK;
}

} // extern "C"
