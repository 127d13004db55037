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
// against the reference class loaded from its file.  exp is ex2.approx (2 ulp); the row sums are fp32 like the
// reference's own tensor arithmetic (per lane over <= 4 * ceil(W / 128) terms with explicit FMAs, then a butterfly).
//
// Round 2 profile (profiles/r02_scalar_*): the first version (one 4-byte load and ~35 instructions per logit, fp64
// butterflies) executed 12.4 M warp-instructions per launch of 8192 x 2 rows x 601 bins and took 21-22 us for 39 MB -
// issue-bound, not bandwidth-bound.  This version reads 16-byte vectors (the row's unaligned head and tail through
// predicated scalar loads, so nothing outside a row is ever read), folds the support value into compile-time
// constants and defers the probability-branch h^-1 to the epilogue.  -DLZB_X_SCALAR_ROWLOADS restores the first version.
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
    float *raw;                           // workspace: [n_heads][N] probability-branch expectations (h^-1 not applied yet)
    unsigned *ctl;                        // workspace: [0] ticket, [1 + h] un-normalised rows seen for head h
    // head h of the per-head arrays without indexing kernel parameters by a run-time value (which would copy the
    // whole struct to local memory)
    __device__ __forceinline__ const float *row_of(int h, int i) const
    {
        return (h == 0 ? logits[0] : logits[1]) + (long long)i * (h == 0 ? stride[0] : stride[1]);
    }
    __device__ __forceinline__ float *out_of(int h) const { return h == 0 ? out[0] : out[1]; }
};
static_assert(kMaxScalarHeads == 2, "row_of / out_of select between two heads");

__device__ __forceinline__ double warp_sum_d(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

__device__ __forceinline__ float exp2f_approx(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// h^-1 (scaling_transform.py:124-126): sign(x) * (((sqrt(1 + 4 eps (|x| + 1 + eps)) - 1) / (2 eps))^2 - 1), in fp32
// like the reference's tensor arithmetic (this translation unit never contracts a * b + c)
__device__ __forceinline__ float inverse_h(float x, float eps)
{
    const float t = (sqrtf(1.0f + 4.0f * eps * (fabsf(x) + 1.0f + eps)) - 1.0f) / (2.0f * eps);
    const float sgn = (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f);
    return sgn * (t * t - 1.0f);
}

// the same with the division by 2 eps as a multiplication by its reciprocal (one rounding apart: 1.2e-7 relative, inside
// the cancellation noise sqrt(1 + d) - 1 already carries; the square root stays IEEE).  The hot path only: every lane of
// a row's warp evaluates h^-1, so the ~20 instructions of two full-precision divisions are paid once per row.
__device__ __forceinline__ float inverse_h_fast(float x, float eps, float inv_2eps)
{
    const float t = (sqrtf(1.0f + 4.0f * eps * (fabsf(x) + 1.0f + eps)) - 1.0f) * inv_2eps;
    const float sgn = (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f);
    return sgn * (t * t - 1.0f);
}

template <int T> // T = loads per lane held in registers: W <= 32 * T
__device__ __forceinline__ void scalar_row(const ScalarHeads &a, int h, int i, int lane, bool &not_normalised)
{
    const int W = 2 * a.support + 1;
    const float *row = a.row_of(h, i);
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
        a.out_of(h)[i] = inverse_h((float)(dte / dse), a.eps);
        a.raw[(size_t)h * a.N + i] = (float)dtx;
        // torch.allclose(sum, 1, atol=1e-5) with the default rtol=1e-5: |sum - 1| <= 1e-5 + 1e-5 * 1; NaN is not close
        not_normalised = !(fabs(dsx - 1.0) <= 2e-5);
    }
}

__device__ __forceinline__ float warp_sum_f(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

// One row by one warp with 16-byte loads.  The row starts `off` floats past a 16-byte boundary (off = 0..3, any row
// stride); vector v of the row covers positions 4v - off .. 4v - off + 3 and lane l takes vectors l, l + 32, ...
// Vectors that lie wholly inside the row are one LDG.128; the (at most two) vectors that straddle the row's ends are
// read element by element under predicates, so no byte outside [row, row + W) is touched.  V = vectors per lane:
// 4 * 32 * V >= W + 3.  Positions outside the row hold -inf (their exp is 0) and a clear bit in the lane's mask.
template <int V>
struct RowRegs {
    float x[V][4];
    unsigned vm[V]; // which of a vector's four positions belong to the row
    int off;
};

template <int V>
__device__ __forceinline__ void load_row_v4(const float *row, int W, int lane, RowRegs<V> &r)
{
    r.off = (int)((reinterpret_cast<unsigned long long>(row) >> 2) & 3ull);
    const float *base = row - r.off; // 16-byte aligned
#pragma unroll
    for (int j = 0; j < V; ++j) {
        const int lo = 4 * (lane + 32 * j) - r.off;
        const bool inside = lo >= 0 && lo + 3 < W;
        r.vm[j] = inside ? 0xfu : 0u;
        float4 v = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        if (inside) v = __ldcs(reinterpret_cast<const float4 *>(base) + lane + 32 * j);
        r.x[j][0] = v.x; r.x[j][1] = v.y; r.x[j][2] = v.z; r.x[j][3] = v.w;
    }
    // the straddling vectors: lane 0's first one when the row is not 16-byte aligned, and the row's last one - both
    // conditions are warp-uniform, so rows and slices without a straddling vector skip the element-wise code
    const int jend = ((r.off + W - 1) >> 2) >> 5;
#pragma unroll
    for (int j = 0; j < V; ++j) {
        if ((j == 0 && r.off != 0) || j == jend) {
            const int lo = 4 * (lane + 32 * j) - r.off;
            if (r.vm[j] == 0u && lo + 3 >= 0 && lo < W) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (lo + c >= 0 && lo + c < W) { r.x[j][c] = __ldcs(row + lo + c); r.vm[j] |= 1u << c; }
            }
        }
    }
}

// softmax expectation and raw (probability-branch) expectation of the row held in r; the support value of position p is
// p - support = (4l - off - support) + (128 j + c): the second term is a compile-time constant, so the weighted sums are
// sum e * (128 j + c)  +  (4l - off - support) * sum e.
template <int V>
__device__ __forceinline__ void reduce_row_v4(const ScalarHeads &a, const RowRegs<V> &r, int h, int i, int lane, bool live, float inv_2eps,
                                              bool &not_normalised)
{
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < V; ++j) m = fmaxf(fmaxf(fmaxf(r.x[j][0], r.x[j][1]), fmaxf(r.x[j][2], r.x[j][3])), m);
    m = warp_max(m);
    const float kLog2e = 1.4426950408889634f;
    const float mneg = -m * kLog2e;
    float se = 0.0f, te = 0.0f, sx = 0.0f, tx = 0.0f;
    // vectors 1 .. jend - 1 lie inside the row in EVERY lane (jend: the slice that holds the row's last vector, warp-
    // uniform), so only slice 0 and slice jend need the position masks, and slices beyond jend hold nothing
    const int jend = ((r.off + 2 * a.support) >> 2) >> 5;
#pragma unroll
    for (int j = 0; j < V; ++j) {
        if (j > jend) continue;
        if (j > 0 && j < jend) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float rel = (float)(128 * j + c); // compile-time constant
                const float e = exp2f_approx(__fmaf_rn(r.x[j][c], kLog2e, mneg));
                se += e;
                te = __fmaf_rn(e, rel, te);
                sx += r.x[j][c];
                tx = __fmaf_rn(r.x[j][c], rel, tx);
            }
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float rel = (float)(128 * j + c);
                const float e = exp2f_approx(__fmaf_rn(r.x[j][c], kLog2e, mneg)); // -inf outside the row: e = 0
                const float xr = ((r.vm[j] >> c) & 1u) ? r.x[j][c] : 0.0f;         // the raw sums skip positions outside the row
                se += e;
                te = __fmaf_rn(e, rel, te);
                sx += xr;
                tx = __fmaf_rn(xr, rel, tx);
            }
        }
    }
    const float lane_sup = (float)(4 * lane - r.off - a.support);
    te = __fmaf_rn(lane_sup, se, te);
    tx = __fmaf_rn(lane_sup, sx, tx);
    se = warp_sum_f(se); te = warp_sum_f(te); sx = warp_sum_f(sx); tx = warp_sum_f(tx);
#ifdef LZB_X_SCALAR_PRECISE_DIV // round-2 first form: two IEEE divisions per row on every lane
    const float scalar = inverse_h(te / se, a.eps);
#else
    const float scalar = inverse_h_fast(__fdividef(te, se), a.eps, inv_2eps);
#endif
    if (lane == 0 && live) {
        a.out_of(h)[i] = scalar;
        a.raw[(size_t)h * a.N + i] = tx;
        not_normalised = !(fabsf(sx - 1.0f) <= 2e-5f); // torch.allclose(sum, 1, atol=1e-5) + default rtol=1e-5; NaN is not close
    }
}

// ---------------------------------------------------------------------------------------------- two rows per warp
// EXPERIMENT (-DLZB_X_SCALAR_PAIR; measured and not adopted: 14.3-16.3 us against 14.4 us, profiles/r02_scalar_transform.md
// - at 39 MB per launch the instruction count is not what bounds the kernel).  The one-row kernel executes 586
// warp-instructions per row, 7.75 per logit in the element loop.  Here a warp takes ROOT i's value row and
// reward row together - ten 16-byte loads in flight per lane - and does the element arithmetic with sm_100's packed fp32
// instructions (FFMA2 / FADD2 on register pairs): per 16-byte vector two FFMA2 (x * log2e - max * log2e), four MUFU.EX2 and
// four FADD2 per accumulated quantity, with SUFFIX sums instead of per-position weights -
//     sum_j j * e_j  =  sum_{k >= 1} R_k,   R_k = sum_{j >= k} e_j       (slices walked from the last to the first)
// so the support values need no constants at all: sum_p e_p * (p - s) = 128 * sum_c Q_c + sum_c c * R_c(0) + (4 lane - off
// - s) * sum_c R_c(0).  3.5 instead of 7.75 instructions per logit; the eight warp sums of the two rows are reduced as four
// packed pairs, and h^-1 is evaluated once per warp (lane 0 for the value head, lane 1 for the reward head).
__device__ __forceinline__ unsigned long long pk2(float lo, float hi)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void upk2(unsigned long long v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b)
{
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ unsigned long long warp_sum_2(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = add2(v, __shfl_xor_sync(kFull, v, o));
    return v;
}

template <int V>
__device__ __forceinline__ float row_max_v4(const RowRegs<V> &r)
{
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < V; ++j) m = fmaxf(fmaxf(fmaxf(r.x[j][0], r.x[j][1]), fmaxf(r.x[j][2], r.x[j][3])), m);
    return m;
}

// per-lane partial sums of one row: {sum e, sum e * (p - s), sum x, sum x * (p - s)} over the lane's positions p
template <int V>
__device__ __forceinline__ float4 row_sums_packed(const RowRegs<V> &r, int support, int lane, float m)
{
    const float kLog2e = 1.4426950408889634f;
    const unsigned long long L2E = pk2(kLog2e, kLog2e), MN = pk2(-m * kLog2e, -m * kLog2e);
    const int jend = ((r.off + 2 * support) >> 2) >> 5; // slice of the row's last vector (warp-uniform)
    unsigned long long Re01 = 0ull, Re23 = 0ull, Qe01 = 0ull, Qe23 = 0ull; // 0ull = {+0.0f, +0.0f}
    unsigned long long Rx01 = 0ull, Rx23 = 0ull, Qx01 = 0ull, Qx23 = 0ull;
#pragma unroll
    for (int j = V - 1; j >= 0; --j) {
        if (j > jend) continue; // nothing of the row in this slice, in any lane
        float t0, t1, t2, t3;
        upk2(fma2(pk2(r.x[j][0], r.x[j][1]), L2E, MN), t0, t1);
        upk2(fma2(pk2(r.x[j][2], r.x[j][3]), L2E, MN), t2, t3);
        Re01 = add2(Re01, pk2(exp2f_approx(t0), exp2f_approx(t1))); // -inf outside the row: e = 0
        Re23 = add2(Re23, pk2(exp2f_approx(t2), exp2f_approx(t3)));
        if (j > 0 && j < jend) { // every lane's vector lies inside the row
            Rx01 = add2(Rx01, pk2(r.x[j][0], r.x[j][1]));
            Rx23 = add2(Rx23, pk2(r.x[j][2], r.x[j][3]));
        } else { // the raw sums skip positions outside the row
            const unsigned vm = r.vm[j];
            Rx01 = add2(Rx01, pk2((vm & 1u) ? r.x[j][0] : 0.0f, (vm & 2u) ? r.x[j][1] : 0.0f));
            Rx23 = add2(Rx23, pk2((vm & 4u) ? r.x[j][2] : 0.0f, (vm & 8u) ? r.x[j][3] : 0.0f));
        }
        if (j >= 1) { // Q = sum_{k >= 1} R_k
            Qe01 = add2(Qe01, Re01); Qe23 = add2(Qe23, Re23);
            Qx01 = add2(Qx01, Rx01); Qx23 = add2(Qx23, Rx23);
        }
    }
    float e0, e1, e2, e3, q0, q1, q2, q3;
    upk2(Re01, e0, e1); upk2(Re23, e2, e3);
    upk2(add2(Qe01, Qe23), q0, q1);
    const float lane_sup = (float)(4 * lane - r.off - support);
    const float se = (e0 + e1) + (e2 + e3);
    float te = __fmaf_rn(128.0f, q0 + q1, __fmaf_rn(3.0f, e3, __fmaf_rn(2.0f, e2, e1)));
    te = __fmaf_rn(lane_sup, se, te);
    upk2(Rx01, e0, e1); upk2(Rx23, e2, e3);
    upk2(add2(Qx01, Qx23), q2, q3);
    const float sx = (e0 + e1) + (e2 + e3);
    float tx = __fmaf_rn(128.0f, q2 + q3, __fmaf_rn(3.0f, e3, __fmaf_rn(2.0f, e2, e1)));
    tx = __fmaf_rn(lane_sup, sx, tx);
    return make_float4(se, te, sx, tx);
}

// wide supports: the row is walked twice (max, then sums) instead of held in registers
__device__ __forceinline__ void scalar_row_wide(const ScalarHeads &a, int h, int i, int lane, bool live, bool &not_normalised)
{
    const int W = 2 * a.support + 1;
    const float *row = a.row_of(h, i);
    float m = -INFINITY;
    for (int k = lane; k < W; k += 32) m = fmaxf(m, row[k]);
    m = warp_max(m);
    double se = 0.0, te = 0.0, sx = 0.0, tx = 0.0;
    for (int k = lane; k < W; k += 32) {
        const float v = row[k], sup = (float)(k - a.support), e = __expf(v - m);
        se += e; te += (double)(e * sup); sx += v; tx += (double)(v * sup);
    }
    se = warp_sum_d(se); te = warp_sum_d(te); sx = warp_sum_d(sx); tx = warp_sum_d(tx);
    if (lane == 0 && live) {
        a.out_of(h)[i] = inverse_h((float)(te / se), a.eps);
        a.raw[(size_t)h * a.N + i] = (float)tx;
        not_normalised = !(fabs(sx - 1.0) <= 2e-5);
    }
}

// Default: ONE row per warp, the grid holds a warp for every row and `iters` = 1.  There is no branch around the row
// pass: a warp without a row (tail of the last block) recomputes the last row with its stores predicated off, and the
// trip count `iters` is a kernel argument - warp-uniform BY CONSTRUCTION - so the compiler keeps the 25 shuffles of a
// row as plain SHFL (inside a loop bounded by a per-thread value each one is wrapped in a WARPSYNC.COLLECTIVE protocol).
// Measured at 8192 x 2 rows x 601 bins (profiles/r02_scalar_transform.md): 4-byte loads + fp64 butterflies in a
// grid-stride loop 12.4 M warp-instructions / 21.5 us; this form 7.7 M / 16.0 us; persistent warps (-DLZB_X_SCALAR_PIPELINED:
// 4 blocks per SM, several rows per warp, the vectors of row it + 1 requested before row it is reduced) 9.7 M / 17.7 us -
// the dead iterations of the last pass and the register copies cost more than the prefetch hides.
//
// Epilogue: the block's warps meet at a barrier, thread 0 forwards the block's "un-normalised" flags and takes the global
// ticket, a second barrier tells the block whether it is the grid's last.  ncu attributes 28 % of all warp-resident time
// to those two barriers (profiles/r02_scalar_transform.md), so a form without them was tried (-DLZB_X_SCALAR_WARP_TICKET:
// a warp that has stored its scalar fences, counts itself in a shared word and leaves; the block's last warp takes the
// ticket; the grid's last warp does the rare rewrite alone) and REJECTED: 17.3 us against 14.4 us back to back in a CUDA
// graph - one MEMBAR.GPU per warp (16,384 per launch instead of 2,048) costs more than the waiting it removes; blocks of
// 128 threads 14.8 us, of 64 threads 20.0 us (-DLZB_SCALAR_THREADS=...).
#ifndef LZB_SCALAR_THREADS
#define LZB_SCALAR_THREADS 256
#endif
constexpr int kScalarThreads = LZB_SCALAR_THREADS;
#ifdef LZB_X_SCALAR_PIPELINED
#define LZB_SCALAR_MIN_BLOCKS 4
#else
#define LZB_SCALAR_MIN_BLOCKS (1536 / LZB_SCALAR_THREADS) // 40 registers: 48 resident warps per SM
#endif
template <int T>
__global__ void __launch_bounds__(kScalarThreads, LZB_SCALAR_MIN_BLOCKS) k_inverse_scalar_transform(ScalarHeads a, int iters)
{
    __shared__ unsigned s_bad[kMaxScalarHeads];
    __shared__ unsigned s_last;
    __shared__ unsigned s_arrived;
    if (threadIdx.x < kMaxScalarHeads) s_bad[threadIdx.x] = 0;
    if (threadIdx.x == kMaxScalarHeads) s_arrived = 0;
    __syncthreads(); // the block's warps start together: nobody waits here
    const float inv_2eps = 1.0f / (2.0f * a.eps);
    const int lane = threadIdx.x & 31;
    const int total = a.n_heads * a.N;
    const int warps = gridDim.x * (blockDim.x >> 5);
    const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    bool bad0 = false, bad1 = false;
#ifndef LZB_X_SCALAR_ROWLOADS
    if constexpr (T > 0) {
        constexpr int V = (32 * T + 3 + 127) / 128; // W <= 32 T; the vectors cover W + 3 positions
        const int W = 2 * a.support + 1;
#ifdef LZB_X_SCALAR_PIPELINED
        RowRegs<V> cur, nxt;
        {
            const int r = min(gw, total - 1);
            const int h = (r >= a.N) ? 1 : 0;
            load_row_v4<V>(a.row_of(h, r - h * a.N), W, lane, nxt);
        }
#endif
#pragma unroll 1
        for (int it = 0; it < iters; ++it) {
            const int r0 = it * warps + gw;
            const bool live = r0 < total;
            const int r = live ? r0 : total - 1;
            const int h = (r >= a.N) ? 1 : 0;
#ifdef LZB_X_SCALAR_PIPELINED
            cur = nxt;
            if (it + 1 < iters) { // warp-uniform
                const int rn = min((it + 1) * warps + gw, total - 1);
                const int hn = (rn >= a.N) ? 1 : 0;
                load_row_v4<V>(a.row_of(hn, rn - hn * a.N), W, lane, nxt);
            }
#else
            RowRegs<V> cur;
            load_row_v4<V>(a.row_of(h, r - h * a.N), W, lane, cur);
#endif
            bool bad = false;
            reduce_row_v4<V>(a, cur, h, r - h * a.N, lane, live, inv_2eps, bad);
            bad0 = bad0 || (bad && h == 0);
            bad1 = bad1 || (bad && h == 1);
        }
    } else
#endif
    {
#pragma unroll 1
        for (int it = 0; it < iters; ++it) {
            const int r0 = it * warps + gw;
            const bool live = r0 < total;
            const int r = live ? r0 : total - 1;
            const int h = (r >= a.N) ? 1 : 0;
            bool bad = false;
#ifdef LZB_X_SCALAR_ROWLOADS
            if constexpr (T > 0) scalar_row<T>(a, h, r - h * a.N, lane, bad);
            else
#endif
            scalar_row_wide(a, h, r - h * a.N, lane, live, bad);
            bad0 = bad0 || (bad && h == 0);
            bad1 = bad1 || (bad && h == 1);
        }
    }
#ifndef LZB_X_SCALAR_WARP_TICKET
    if (lane == 0 && bad0) s_bad[0] = 1; // benign race: every writer stores 1
    if (lane == 0 && bad1) s_bad[1] = 1;
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
            for (int i = threadIdx.x; i < a.N; i += blockDim.x) a.out_of(h)[i] = inverse_h(__ldcg(a.raw + (size_t)h * a.N + i), a.eps);
    }
    __syncthreads();
    if (threadIdx.x <= a.n_heads) a.ctl[threadIdx.x] = 0; // ready for the next call
#else
    (void)s_last;
    unsigned last = 0;
    if (lane == 0) {
        if (bad0) atomicOr(&s_bad[0], 1u);
        if (bad1) atomicOr(&s_bad[1], 1u);
        __threadfence(); // this warp's scalars (and its flags) are visible before its arrival is counted
        if (atomicAdd(&s_arrived, 1u) == (blockDim.x >> 5) - 1) { // the block's last warp: every other warp's flags are in
            for (int h = 0; h < a.n_heads; ++h)
                if (*((volatile unsigned *)&s_bad[h])) atomicAdd(a.ctl + 1 + h, 1u);
            __threadfence(); // the block's counters are visible before its ticket
            last = (atomicAdd(a.ctl, 1u) == gridDim.x - 1) ? 1u : 0u;
        }
    }
    last = __shfl_sync(kFull, last, 0);
    if (!last) return;
    __threadfence();
    for (int h = 0; h < a.n_heads; ++h) {
        const unsigned bad = *((volatile unsigned *)(a.ctl + 1 + h));
        if (bad == 0) // every row of this head sums to 1: the head emitted probabilities (ensure_softmax's else branch)
            for (int i = lane; i < a.N; i += 32) a.out_of(h)[i] = inverse_h(__ldcg(a.raw + (size_t)h * a.N + i), a.eps);
    }
    __syncwarp();
    if (lane <= a.n_heads) a.ctl[lane] = 0; // ready for the next call
#endif
}

// the two-rows-per-warp kernel: n_heads == 2, one warp per root (see above).  Same epilogue as the one-row kernel.
#ifndef LZB_SCALAR_PAIR_MIN_BLOCKS
#define LZB_SCALAR_PAIR_MIN_BLOCKS 3
#endif
template <int T>
__global__ void __launch_bounds__(kScalarThreads, LZB_SCALAR_PAIR_MIN_BLOCKS) k_inverse_scalar_transform_pair(ScalarHeads a, int iters)
{
    static_assert(T > 0, "the pair kernel holds both rows in registers");
    __shared__ unsigned s_bad[kMaxScalarHeads];
    __shared__ unsigned s_last;
    if (threadIdx.x < kMaxScalarHeads) s_bad[threadIdx.x] = 0;
    __syncthreads();
    constexpr int V = (32 * T + 3 + 127) / 128;
    const float inv_2eps = 1.0f / (2.0f * a.eps);
    const int lane = threadIdx.x & 31;
    const int W = 2 * a.support + 1;
    const int warps = gridDim.x * (blockDim.x >> 5);
    const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    bool bad0 = false, bad1 = false;
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        const int r0 = it * warps + gw;
        const bool live = r0 < a.N;
        const int i = live ? r0 : a.N - 1;
        RowRegs<V> rv, rr;
        load_row_v4<V>(a.logits[0] + (long long)i * a.stride[0], W, lane, rv);
        load_row_v4<V>(a.logits[1] + (long long)i * a.stride[1], W, lane, rr);
        float mv = row_max_v4<V>(rv), mr = row_max_v4<V>(rr);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mv = fmaxf(mv, __shfl_xor_sync(kFull, mv, o));
            mr = fmaxf(mr, __shfl_xor_sync(kFull, mr, o));
        }
        const float4 pv = row_sums_packed<V>(rv, a.support, lane, mv);
        const float4 pr = row_sums_packed<V>(rr, a.support, lane, mr);
        float se0, se1, te0, te1, sx0, sx1, tx0, tx1; // head 0 = value row, head 1 = reward row
        upk2(warp_sum_2(pk2(pv.x, pr.x)), se0, se1);
        upk2(warp_sum_2(pk2(pv.y, pr.y)), te0, te1);
        upk2(warp_sum_2(pk2(pv.z, pr.z)), sx0, sx1);
        upk2(warp_sum_2(pk2(pv.w, pr.w)), tx0, tx1);
        const bool second = lane & 1; // lane 0 finishes the value head, lane 1 the reward head
        const float scalar = inverse_h_fast(__fdividef(second ? te1 : te0, second ? se1 : se0), a.eps, inv_2eps);
        if (live && lane < 2) {
            (second ? a.out[1] : a.out[0])[i] = scalar;
            a.raw[(size_t)(second ? a.N : 0) + i] = second ? tx1 : tx0;
        }
        // torch.allclose(sum, 1, atol=1e-5) with the default rtol=1e-5; NaN is not close
        bad0 = bad0 || (live && !(fabsf(sx0 - 1.0f) <= 2e-5f));
        bad1 = bad1 || (live && !(fabsf(sx1 - 1.0f) <= 2e-5f));
    }
    if (lane == 0 && bad0) s_bad[0] = 1; // benign race: every writer stores 1
    if (lane == 0 && bad1) s_bad[1] = 1;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int h = 0; h < 2; ++h)
            if (s_bad[h]) atomicAdd(a.ctl + 1 + h, 1u);
        __threadfence(); // this block's scalars and counters are visible before its ticket
        s_last = (atomicAdd(a.ctl, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int h = 0; h < 2; ++h) {
        const unsigned bad = *((volatile unsigned *)(a.ctl + 1 + h));
        if (bad == 0) // every row of this head sums to 1: the head emitted probabilities (ensure_softmax's else branch)
            for (int i = threadIdx.x; i < a.N; i += blockDim.x) a.out_of(h)[i] = inverse_h(__ldcg(a.raw + (size_t)h * a.N + i), a.eps);
    }
    __syncthreads();
    if (threadIdx.x <= 2) a.ctl[threadIdx.x] = 0; // ready for the next call
}

} // namespace lzb
