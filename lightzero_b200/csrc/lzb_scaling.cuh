// InverseScalarTransform on the device: what the reference driver applies to the model's value and reward /
// value_prefix heads every simulation (lzero/mcts/tree_search/mcts_ctree.py:302-303 -> lzero/policy/
// scaling_transform.py:118-128, with ensure_softmax :36-62): categorical logits over the support
// {-s .. s} -> softmax -> expectation -> h^-1.  One launch serves both heads: it reads every logit ONCE (the path is
// HBM-bound: 2 x N x (2s+1) floats, 39 MB at 8192 roots and the Atari support of 300) and writes N scalars per head
// that lzb_backprop / lzb_step consume as they are.
//
// ensure_softmax is a batch-global decision (torch.allclose over ALL row sums): a head whose rows all sum to 1 is
// taken as probabilities, any other head is soft-maxed.  Rows of real networks are raw logits, so the soft-max branch
// is written straight away; the probability-branch value of every row goes to a scratch array, and the LAST block to
// finish overwrites a head's scalars from it if no row of that head was found un-normalised (one atomic ticket per
// block; the counters are reset by that block, so the workspace stays ready for the next call).
//
// This is not bit-pinned by the tree (torch's own reductions fix no summation order); tests/ hold it to 1e-5 relative
// against the reference class loaded from its file.  exp is ex2.approx (2 ulp), the row sums are taken per lane in
// fp32 over <= ceil(W / 32) terms and across lanes in fp64.
#pragma once
#include <cuda_runtime.h>

#include "lzb_math.cuh"

namespace lzb {

constexpr int kMaxScalarHeads = 2;

struct ScalarHeads {
    const float *logits[kMaxScalarHeads]; // [N, W] rows, row stride in floats below
    float *out[kMaxScalarHeads];          // [N]
    long long stride[kMaxScalarHeads];
    int n_heads, N, support;              // W = 2 * support + 1
    float eps;
    float *raw;                           // workspace: [n_heads][N] probability-branch scalars
    unsigned *ctl;                        // workspace: [0] ticket, [1 + h] un-normalised rows seen for head h
};

__device__ __forceinline__ double warp_sum_d(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

// h^-1 (scaling_transform.py:124-126): sign(x) * (((sqrt(1 + 4 eps (|x| + 1 + eps)) - 1) / (2 eps))^2 - 1), in fp32
// like the reference's tensor arithmetic (this translation unit never contracts a * b + c)
__device__ __forceinline__ float inverse_h(float x, float eps)
{
    const float t = (sqrtf(1.0f + 4.0f * eps * (fabsf(x) + 1.0f + eps)) - 1.0f) / (2.0f * eps);
    const float sgn = (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f);
    return sgn * (t * t - 1.0f);
}

template <int T> // T = loads per lane held in registers: W <= 32 * T
__device__ __forceinline__ void scalar_row(const ScalarHeads &a, int h, int i, int lane, bool &not_normalised)
{
    const int W = 2 * a.support + 1;
    const float *row = a.logits[h] + (long long)i * a.stride[h];
    float x[T];
#pragma unroll
    for (int t = 0; t < T; ++t) { // all loads of the row in flight together; streamed: read once
        const int k = lane + 32 * t;
        x[t] = (k < W) ? __ldcs(row + k) : -INFINITY;
    }
    float m = -INFINITY;
#pragma unroll
    for (int t = 0; t < T; ++t) m = fmaxf(m, x[t]);
    m = warp_max(m);
    float se = 0.0f, te = 0.0f, sx = 0.0f, tx = 0.0f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const int k = lane + 32 * t;
        if (k < W) {
            const float sup = (float)(k - a.support);
            const float e = __expf(x[t] - m);
            se += e;
            te += e * sup;
            sx += x[t];
            tx += x[t] * sup;
        }
    }
    const double dse = warp_sum_d((double)se), dte = warp_sum_d((double)te);
    const double dsx = warp_sum_d((double)sx), dtx = warp_sum_d((double)tx);
    if (lane == 0) {
        a.out[h][i] = inverse_h((float)(dte / dse), a.eps);
        a.raw[(size_t)h * a.N + i] = inverse_h((float)dtx, a.eps);
        // torch.allclose(sum, 1, atol=1e-5) with the default rtol=1e-5: |sum - 1| <= 1e-5 + 1e-5 * 1; NaN is not close
        not_normalised = !(fabs(dsx - 1.0) <= 2e-5);
    }
}

// wide supports: the row is walked twice (max, then sums) instead of held in registers
__device__ __forceinline__ void scalar_row_wide(const ScalarHeads &a, int h, int i, int lane, bool &not_normalised)
{
    const int W = 2 * a.support + 1;
    const float *row = a.logits[h] + (long long)i * a.stride[h];
    float m = -INFINITY;
    for (int k = lane; k < W; k += 32) m = fmaxf(m, row[k]);
    m = warp_max(m);
    double se = 0.0, te = 0.0, sx = 0.0, tx = 0.0;
    for (int k = lane; k < W; k += 32) {
        const float v = row[k], sup = (float)(k - a.support), e = __expf(v - m);
        se += e; te += (double)(e * sup); sx += v; tx += (double)(v * sup);
    }
    se = warp_sum_d(se); te = warp_sum_d(te); sx = warp_sum_d(sx); tx = warp_sum_d(tx);
    if (lane == 0) {
        a.out[h][i] = inverse_h((float)(te / se), a.eps);
        a.raw[(size_t)h * a.N + i] = inverse_h((float)tx, a.eps);
        not_normalised = !(fabs(sx - 1.0) <= 2e-5);
    }
}

// one warp per row; rows of head 0 first, then head 1 (the grid strides over n_heads * N rows)
template <int T>
__global__ void __launch_bounds__(256) k_inverse_scalar_transform(ScalarHeads a)
{
    __shared__ unsigned s_bad[kMaxScalarHeads];
    __shared__ unsigned s_last;
    if (threadIdx.x < kMaxScalarHeads) s_bad[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int warps = gridDim.x * (blockDim.x >> 5);
    const int total = a.n_heads * a.N;
    for (int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < total; r += warps) {
        const int h = r / a.N, i = r - h * a.N;
        bool bad = false;
        if constexpr (T > 0) scalar_row<T>(a, h, i, lane, bad);
        else scalar_row_wide(a, h, i, lane, bad);
        if (lane == 0 && bad) s_bad[h] = 1; // benign race: every writer stores 1
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int h = 0; h < a.n_heads; ++h)
            if (s_bad[h]) atomicAdd(a.ctl + 1 + h, 1u);
        __threadfence(); // this block's scalars and counters are visible before its ticket
        s_last = (atomicAdd(a.ctl, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int h = 0; h < a.n_heads; ++h) {
        const unsigned bad = *((volatile unsigned *)(a.ctl + 1 + h));
        if (bad == 0) // every row of this head sums to 1: the head emitted probabilities (ensure_softmax's else branch)
            for (int i = threadIdx.x; i < a.N; i += blockDim.x) a.out[h][i] = __ldcg(a.raw + (size_t)h * a.N + i);
    }
    __syncthreads();
    if (threadIdx.x <= a.n_heads) a.ctl[threadIdx.x] = 0; // ready for the next call
}

} // namespace lzb
