// Lane-group kernels: G lanes (8, 16 or 32) cooperate on one tree, so a warp walks 32/G trees at once.
//
// The tree work is a few hundred dependent scalar operations per level on at most A child records; with one
// whole warp per tree (lzb_select.cuh) most issue slots carry 18 of 32 lanes through shuffles and reductions
// and the kernel is instruction-issue bound (ncu: issue active 42 %, 1.8 k warp-instructions per tree-step).
// Here lane k of a group keeps children j = k + G*r, r < R, in registers (no shared memory), every loop over
// children is unrolled at compile time, and group-wide exchanges use shuffles of width G under the group's
// member mask, so the 32/G trees of a warp share every instruction while staying free to diverge in depth.
//
// Arithmetic is identical to the warp-per-tree kernels (and hence to the reference): fp32, no FMA contraction,
// serial left-to-right sums in legal-action order (adding +0.0f for skipped entries is exact because a running
// sum that starts at +0.0f can never become -0.0f), glibc-exact expf / logf.
#pragma once
#if defined(LZB_X_CARRY) && !defined(LZB_X_EARLY_PATH)
#define LZB_X_EARLY_PATH
#endif
#include <type_traits>

#include "lzb_expand.cuh"
#include "lzb_select.cuh"

#ifndef LZB_MIN_BLOCKS
#define LZB_MIN_BLOCKS 7 // resident blocks of 128 threads per SM the lane-group kernels are compiled for (72 registers)
#endif

namespace lzb {

// All 32 lanes of a warp execute every instruction of these kernels together (warp-synchronous style): exchanges
// are full-mask shuffles of width G and full-mask ballots shifted to the group's bits, and a group that has
// finished (reached its leaf, or has no tree) keeps executing with its writes predicated off.  Sub-warp member
// masks would let groups diverge, but the compiler then wraps every shuffle in a convergence protocol
// (WARPSYNC.COLLECTIVE / REDUX / BSSY...BSYNC) that more than doubled the instruction count when measured.
// G need not be a power of two: 10-lane groups put THREE trees into a warp (lanes 30, 31 idle) and serve action
// spaces up to 10 (R = 1) / 20 (R = 2) with the same chain length as 16-lane groups but two thirds of the warps.
// Power-of-two groups exchange with shuffles of width G and butterfly reductions; the others address the source lane
// absolutely (first lane of the group + index) and reduce over a ring (offsets 1, 2, 4, 8: windows double, and for the
// idempotent max / min an overlapping last window is harmless).
template <int G>
struct Grp {
    static constexpr bool kPow2 = (G & (G - 1)) == 0;
    static constexpr int kLanes = G;
    int lane;  // 0 .. G-1 within the group
    int shift; // first warp lane of the group
    __device__ __forceinline__ Grp()
    {
        const int wl = threadIdx.x & 31;
        lane = kPow2 ? (wl & (G - 1)) : (wl % G);
        shift = wl - lane;
    }
    __device__ __forceinline__ unsigned ballot(bool pred) const
    {
        const unsigned m = __ballot_sync(kFull, pred) >> shift;
        return (G == 32) ? m : (m & ((1u << (G & 31)) - 1u));
    }
    // lane of the warp that holds group member (lane + o) mod G
    __device__ __forceinline__ int ring(int o) const
    {
        const int t = lane + o;
        return shift + (t >= G ? t - G : t);
    }
    template <typename T>
    __device__ __forceinline__ T shfl_(T v, int src) const
    {
        if constexpr (kPow2) return __shfl_sync(kFull, v, src, G);
        else return __shfl_sync(kFull, v, shift + src);
    }
    __device__ __forceinline__ float shfl(float v, int src) const { return shfl_(v, src); }
    __device__ __forceinline__ int shfl(int v, int src) const { return shfl_(v, src); }
    __device__ __forceinline__ double shfl(double v, int src) const { return shfl_(v, src); }
    __device__ __forceinline__ int4 shfl(int4 v, int src) const
    {
        return make_int4(shfl(v.x, src), shfl(v.y, src), shfl(v.z, src), shfl(v.w, src));
    }
    // sum += v of lanes 0 .. cnt-1, strictly in lane order (cnt is warp-uniform; cnt == G: fully unrolled)
    __device__ __forceinline__ float ordered_add(float sum, float v, int cnt) const
    {
        if (cnt >= G) {
#pragma unroll
            for (int l = 0; l < G; ++l) sum += shfl(v, l);
        } else {
#pragma unroll 1
            for (int l = 0; l < cnt; ++l) sum += shfl(v, l);
        }
        return sum;
    }
    // sum += v of lanes 0 .. cnt-1 in lane order, in unrolled blocks of eight: lanes cnt .. (next multiple of 8) are added too,
    // so the caller guarantees they hold +0.0f (exact to add).  For a group wider than the action space (the lone walker).
    __device__ __forceinline__ float ordered_add_upto(float sum, float v, int cnt) const
    {
#pragma unroll
        for (int b = 0; b < G; b += 8) {
            if (b < cnt) {
#pragma unroll
                for (int l = b; l < b + 8 && l < G; ++l) sum += shfl(v, l);
            }
        }
        return sum;
    }
    // slice r (of RR) of an ordered sum over the A values a group holds as v[r] for j = lane + G * r, lanes beyond the action
    // space holding +0.0f: the last slice stops at the action space; the lone walker's single slice (32 lanes, A <= 32) adds
    // blocks of eight lanes up to A
    template <int RR>
    __device__ __forceinline__ float ordered_slice(float sum, float v, int r, int A) const
    {
        if constexpr (G == 32 && RR == 1) return ordered_add_upto(sum, v, A);
        else return (r == RR - 1 && RR > 1) ? ordered_add(sum, v, A - G * (RR - 1)) : ordered_add(sum, v, G);
    }
    // the same ordered sum when only the lanes in `u` (group-relative bits, the UNION over the warp's groups, warp-uniform)
    // can hold a non-zero term: adding +0.0f for the others is exact, so they are skipped
    __device__ __forceinline__ float ordered_add_sparse(float sum, float v, unsigned u) const
    {
#pragma unroll 1
        while (u) {
            const int l = __ffs(u) - 1;
            u &= u - 1;
            sum += shfl(v, l);
        }
        return sum;
    }
    __device__ __forceinline__ unsigned union_mask(bool pred) const // group-relative lanes where `pred` holds in ANY group of the warp
    {
        const unsigned w = __ballot_sync(kFull, pred);
        unsigned u = 0;
#pragma unroll
        for (int s = 0; s + G <= 32; s += G) u |= w >> s;
        return (G == 32) ? u : (u & ((1u << (G & 31)) - 1u));
    }
#ifndef LZB_X_NO_REDUX
    // Group reductions by the warp-wide reduction unit (sm_100a: redux.sync.{max,min}.f32 -> CREDUX, integers REDUX) instead of
    // log2(G) dependent shuffle + op steps (round 2; -DLZB_X_NO_REDUX = before: mz 16.66 -> 16.35 us per step, tree-only 13.20 ->
    // 12.75, gmz 22.57 -> 22.09, ez 28.94 -> 28.77).  Two 16-lane groups = two independent full-mask reductions with
    // the other group's lanes holding the operation's identity.  NaN handling as fmaxf / fminf (the number wins).
    static __device__ __forceinline__ float rx_fmax(float v) { float r; asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r; }
    static __device__ __forceinline__ float rx_fmin(float v) { float r; asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r; }
    template <typename T, typename F>
    __device__ __forceinline__ T rx_group(T v, T ident, F f) const
    {
        if constexpr (G == 32) return f(v);
        else {
            const bool hi = shift != 0;
            const T lo_r = f(hi ? ident : v), hi_r = f(hi ? v : ident);
            return hi ? hi_r : lo_r;
        }
    }
#endif
    __device__ __forceinline__ float fmax(float v) const
    {
#ifndef LZB_X_NO_REDUX
        if constexpr (G == 32 || G == 16) return rx_group(v, -INFINITY, [](float x) { return rx_fmax(x); });
#endif
        if constexpr (kPow2) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o, G));
        } else {
#pragma unroll
            for (int o = 1; o < G; o <<= 1) v = fmaxf(v, __shfl_sync(kFull, v, ring(o)));
        }
        return v;
    }
    __device__ __forceinline__ float fmin(float v) const
    {
#ifndef LZB_X_NO_REDUX
        if constexpr (G == 32 || G == 16) return rx_group(v, INFINITY, [](float x) { return rx_fmin(x); });
#endif
        if constexpr (kPow2) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(kFull, v, o, G));
        } else {
#pragma unroll
            for (int o = 1; o < G; o <<= 1) v = fminf(v, __shfl_sync(kFull, v, ring(o)));
        }
        return v;
    }
    __device__ __forceinline__ int isum(int v) const
    {
#ifndef LZB_X_NO_REDUX
        if constexpr (G == 32 || G == 16) return rx_group(v, 0, [](int x) { return __reduce_add_sync(kFull, x); });
#endif
        if constexpr (kPow2) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o, G);
            return v;
        } else { // exact ring windows: w[k] = sum over 2^k members from this lane on; G = sum of its binary digits' windows
            int total = 0, at = 0, w = v;
#pragma unroll
            for (int o = 1; o <= G; o <<= 1) {
                if (G & o) { // take the window of o members that starts `at` members further on
                    total += (at == 0) ? w : __shfl_sync(kFull, w, ring(at));
                    at += o;
                }
                if ((o << 1) <= G) w += __shfl_sync(kFull, w, ring(o));
            }
            return total;
        }
    }
    __device__ __forceinline__ int imax(int v) const
    {
#ifndef LZB_X_NO_REDUX
        if constexpr (G == 32 || G == 16) return rx_group(v, (int)0x80000000, [](int x) { return __reduce_max_sync(kFull, x); });
#endif
        if constexpr (kPow2) {
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(kFull, v, o, G));
        } else {
#pragma unroll
            for (int o = 1; o < G; o <<= 1) v = max(v, __shfl_sync(kFull, v, ring(o)));
        }
        return v;
    }
};

// tree handled by this thread's lane group: (32 / G) trees per warp, four warps per block; lanes of a warp beyond the
// last whole group (G = 10: lanes 30, 31) form a group without a tree
template <int G>
__device__ __forceinline__ int group_tree(bool &has_tree)
{
    constexpr int kPerWarp = 32 / G;
    const int g = (threadIdx.x & 31) / G;
    has_tree = g < kPerWarp;
    return (blockIdx.x * 4 + (threadIdx.x >> 5)) * kPerWarp + g;
}

// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-serialization attribute may start
// while its predecessor in the stream drains; pdl_wait() is the point past which the predecessor's writes are visible
// (a no-op for a normally launched kernel), pdl_release() lets the successor's blocks be scheduled as ours exit.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Tuning aid, compiled only with -DLZB_TIMELINE (never in the shipped library): every warp of the fused step kernel
// stamps clock64() at its phase boundaries into a caller-provided buffer [launch][warp][24] (scripts/gpu_timeline.py).
// LZB_TLD takes the stamp only once `dep` has arrived (the clock read is predicated on it).
#ifdef LZB_TIMELINE
__device__ unsigned long long *g_timeline;
__device__ __forceinline__ void tl_store(int row, int slot, unsigned long long v)
{
    if (row >= 0 && g_timeline && (threadIdx.x & 31) == 0)
        g_timeline[(((size_t)row * gridDim.x + blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 24 + slot] = v;
}
__device__ __forceinline__ void tl_mark(int row, int slot)
{
    unsigned long long v;
    if (slot == 1) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(v));
    else v = clock64();
    tl_store(row, slot, v);
}
__device__ __forceinline__ void tl_mark_dep(int row, int slot, int dep)
{
    unsigned long long v = 0;
    asm volatile("{ .reg .pred p; setp.ne.s32 p, %1, 0x7ffffff1; @p mov.u64 %0, %clock64; }" : "+l"(v) : "r"(dep));
    tl_store(row, slot, v);
}
#define LZB_TL(row, slot) tl_mark(row, slot)
#define LZB_TLD(row, slot, dep) tl_mark_dep(row, slot, dep)
#else
#define LZB_TL(row, slot) ((void)0)
#define LZB_TLD(row, slot, dep) ((void)0)
#endif

// Leaf rows of one tree moved by the G lanes of its group (same semantics as gather_rows in lzb_select.cuh).
// slot_row >= 0: the slot-table row to read instead of the gather's own index (stochastic tree: the row of the parent's kind)
// experiment LZB_X_ROW_PREFETCH: a walk that ENTERS a node with visit count 1 ends there for certain (none of its children has
// been visited, so whichever is chosen is unexpanded), i.e. the rows this tree will gather are that node's - known one whole
// level (block read + scores, ~3.6 k cycles) before the gather starts.  The group asks L2 for them right away
// (prefetch.global.L2: no registers, no wasted traffic since the prediction is exact); the deepest trees - the ones that set
// the launch time - almost always end on such a node.
template <int G>
__device__ __forceinline__ void prefetch_rows_g(const Pool &p, const SelectArgs &a, int i, int ix, int lane)
{
    for (int k = 0; k < a.ng; ++k) {
        const long long rb = a.g[k].row_bytes;
        const char *s;
        if (a.g[k].src) s = a.g[k].src + ((long long)ix * p.N + i) * rb;
        else {
            const char *slot = (ix == a.new_e) ? pick3(a.new_slot, k) : p.slots[(size_t)k * p.E + ix];
            if (slot == nullptr) continue;
            s = slot + (long long)i * rb;
        }
        for (long long off = (long long)lane * 128; off < rb; off += (long long)G * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(s + off));
    }
}

// experiment LZB_X_GATHER_STCS: the gathered rows written with the streaming (evict-first) policy, so that 33 MB of rows per
// step do not push the node pool out of L2
// experiments LZB_X_GATHER_LDG / LZB_X_GATHER_LDNC: the row loads with the default policy / through the read-only path instead
// of the streaming (evict-first) one
#if defined(LZB_X_GATHER_LDG)
#define LZB_GATHER_LD(ptr) (*(ptr))
#elif defined(LZB_X_GATHER_LDNC)
#define LZB_GATHER_LD(ptr) __ldg(ptr)
#elif defined(LZB_X_GATHER_LDLU) // "last use"
#define LZB_GATHER_LD(ptr) __ldlu(ptr)
#elif defined(LZB_X_GATHER_LDCG) // L2 only
#define LZB_GATHER_LD(ptr) __ldcg(ptr)
#else
#define LZB_GATHER_LD(ptr) __ldcs(ptr)
#endif
#ifdef LZB_X_GATHER_STCS
#define LZB_GATHER_ST(ptr, val) __stcs((ptr), (val))
#else
#define LZB_GATHER_ST(ptr, val) (*(ptr) = (val))
#endif

// Lone walker, rows (round 2; -DLZB_X_NO_LONE_STAGE = before, -DLZB_X_LONE_STAGE_LSU = the cp.async form): when a warp's first tree
// ends its walk, its row is requested from the bulk-copy engine into shared memory (one cp.async.bulk by one lane, completion on an
// mbarrier: no registers, no load / store unit traffic) while the other tree walks on with all 32 lanes; afterwards the engine
// writes it to the model batch and the walker's own row is gathered by all 32 lanes in ONE round trip (8 x 16 bytes per lane)
// instead of the two (12 + 4 vectors per lane) of the 16-lane form.  One row of at most LZB_LONE_STAGE_BYTES per tree (the
// MuZero case); anything else takes the ordinary path.  Measured: mz 16.66 -> 16.34 us per step; the same staging through
// cp.async (LDGSTS, 16 requests per lane) was SLOWER, 18.1 us - it is the load / store path the early rows must stay out of.
#if defined(LZB_X_NO_SHARED_TABLES) && !defined(LZB_X_NO_LONE_STAGE)
#define LZB_X_NO_LONE_STAGE // the staging area sits behind the tables in dynamic shared memory
#endif
#ifndef LZB_LONE_STAGE_BYTES
#define LZB_LONE_STAGE_BYTES 4096
#endif
// dynamic shared memory of the lane-group kernels, in 8-byte words: [0, 64) libm tables, [64, 68) one mbarrier per warp,
// [68, ...) one staged row per warp
constexpr int kStageBarWord = 64, kStageRowWord = 68;
__host__ __device__ __forceinline__ bool lone_stage_rows(const SelectArgs &a)
{
    return a.ng == 1 && a.g[0].row_bytes <= LZB_LONE_STAGE_BYTES && (a.g[0].row_bytes & 15) == 0 && (((unsigned long long)a.g[0].dst) & 15ull) == 0;
}
template <int VAR, int G, int R, bool REUSE>
struct LoneWalker {
#if defined(LZB_X_LONE_WALKER_ALL)
    static constexpr bool value = G == 16 && R == 2;
#elif !defined(LZB_X_NO_LONE_WALKER)
    static constexpr bool value = !REUSE && VAR != LZB_SMZ && G == 16 && R == 2;
#else
    static constexpr bool value = false;
#endif
};
__device__ __forceinline__ const char *row_src_g(const Pool &p, const SelectArgs &a, int k, int i, int ix)
{
    const long long rb = a.g[k].row_bytes;
    if (a.g[k].src) return a.g[k].src + ((long long)ix * p.N + i) * rb;
    const char *slot = (ix == a.new_e) ? pick3(a.new_slot, k) : p.slots[(size_t)k * p.E + ix];
    return slot ? slot + (long long)i * rb : nullptr;
}

template <int G>
__device__ __forceinline__ void gather_rows_g(const Pool &p, const SelectArgs &a, int i, int ix, int parent_reset, int lane, int slot_row = -1)
{
    LZB_CHK(p, ix >= 0 && ix < p.E, 10);
    for (int k = 0; k < a.ng; ++k) {
        const long long rb = a.g[k].row_bytes;
        const char *s;
        const int ks = (slot_row >= 0) ? slot_row : k;
        if (a.g[k].src) s = a.g[k].src + ((long long)ix * p.N + i) * rb;
        else {
            const char *slot = (ix == a.new_e) ? pick3(a.new_slot, ks) : p.slots[(size_t)ks * p.E + ix];
            LZB_CHK(p, slot != nullptr, 11); // a slot that was never registered
#ifdef LZB_CHECKED
            if (slot == nullptr) continue; // reported above: do not fault on it
#endif
            s = slot + (long long)i * rb;
        }
        char *d = a.g[k].dst + (long long)i * rb;
        const bool zero = a.g[k].zero_if_reset && parent_reset == 1;
        if ((rb & 15) == 0 && ((((unsigned long long)s) | ((unsigned long long)d)) & 15ull) == 0) {
            const int4 *s4 = reinterpret_cast<const int4 *>(s);
            int4 *d4 = reinterpret_cast<int4 *>(d);
            const int n4 = (int)(rb >> 4);
            if (zero) {
                for (int t = lane; t < n4; t += G) d4[t] = make_int4(0, 0, 0, 0);
            } else {
#ifdef LZB_X_GATHER_PREFETCH
                // experiment: what the first round of loads does not cover - the tail of this row, and (from row 0) the
                // whole of the tree's other rows - is requested into L2 now (no registers), so that the later rounds of this
                // gather meet L2 instead of DRAM: 12 + 4 vectors per lane are two round trips, EfficientZero's three rows four
                {
                    const int first = (n4 >= 12 * G) ? 12 * G : ((n4 >= 8 * G) ? 8 * G : ((n4 >= 4 * G) ? 4 * G : n4));
                    for (long long off = (long long)first * 16 + (long long)lane * 128; off < rb; off += (long long)G * 128)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(s + off));
                    if (k == 0)
                        for (int kk = 1; kk < a.ng; ++kk) {
                            const long long rb2 = a.g[kk].row_bytes;
                            const int ks2 = (slot_row >= 0) ? slot_row : kk;
                            const char *s2;
                            if (a.g[kk].src) s2 = a.g[kk].src + ((long long)ix * p.N + i) * rb2;
                            else {
                                const char *slot2 = (ix == a.new_e) ? pick3(a.new_slot, ks2) : p.slots[(size_t)ks2 * p.E + ix];
                                if (slot2 == nullptr) continue;
                                s2 = slot2 + (long long)i * rb2;
                            }
                            for (long long off = (long long)lane * 128; off < rb2; off += (long long)G * 128)
                                asm volatile("prefetch.global.L2 [%0];" ::"l"(s2 + off));
                        }
                }
#endif
                int t = lane;
                for (; t + 11 * G < n4; t += 12 * G) { // 12 independent 16-byte loads in flight per lane
                    int4 v[12];
#pragma unroll
                    for (int u = 0; u < 12; ++u) v[u] = LZB_GATHER_LD(s4 + t + u * G);
#pragma unroll
                    for (int u = 0; u < 12; ++u) LZB_GATHER_ST(d4 + t + u * G, v[u]);
                }
#ifndef LZB_X_NO_GATHER8 // rows of 8 vectors per lane (the EfficientZero LSTM rows: 2 KB per 16-lane group) in one round trip instead of
                          // two: ez 29.65 -> 29.45 us, its tree + LSTM rows alone 19.08 -> 18.60 us (round 2; -DLZB_X_NO_GATHER8 = before)
                for (; t + 7 * G < n4; t += 8 * G) {
                    int4 v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) v[u] = LZB_GATHER_LD(s4 + t + u * G);
#pragma unroll
                    for (int u = 0; u < 8; ++u) LZB_GATHER_ST(d4 + t + u * G, v[u]);
                }
#endif
                for (; t + 3 * G < n4; t += 4 * G) {
                    int4 v[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) v[u] = LZB_GATHER_LD(s4 + t + u * G);
#pragma unroll
                    for (int u = 0; u < 4; ++u) LZB_GATHER_ST(d4 + t + u * G, v[u]);
                }
                for (; t < n4; t += G) d4[t] = LZB_GATHER_LD(s4 + t);
            }
        } else {
            const int *s1 = reinterpret_cast<const int *>(s);
            int *d1 = reinterpret_cast<int *>(d);
            const int n1 = (int)(rb >> 2);
            for (int t = lane; t < n1; t += G) d1[t] = zero ? 0 : __ldcs(s1 + t);
        }
    }
}

// Outcome of a chance node (Stochastic MuZero, ctree_stochastic_muzero/lib/cnode.cpp:610-630) for a uniform u in
// [0, 1): libstdc++'s std::discrete_distribution (bits/random.tcc:2657-2714) over the n children's priors - widened
// to double, normalised by their sum, partial sums in index order with the last one forced to 1.0, lower_bound.
// The lane group holds prior j = lane + G*r in c[r].z; sums run strictly in index order (+0.0 beyond n is exact).
template <int G, int R>
__device__ __forceinline__ int chance_draw_g(const int4 (&c)[R], int n, double u, const Grp<G> &grp)
{
    const int lane = grp.lane;
    double pr[R];
    double sum = 0.0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        pr[r] = (lane + G * r < n) ? (double)__int_as_float(c[r].z) : 0.0;
#pragma unroll
        for (int l = 0; l < G; ++l) sum += grp.shfl(pr[r], l);
    }
    double cp = 0.0;
    int below = 0; // partial sums smaller than u == index lower_bound returns
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int j = lane + G * r;
        const double q = (j < n) ? pr[r] / sum : 0.0;
        double mine = 0.0;
#pragma unroll
        for (int l = 0; l < G; ++l) {
            cp += grp.shfl(q, l);
            mine = (lane == l) ? cp : mine;
        }
        const double cj = (j == n - 1) ? 1.0 : mine;
        below += __popc(grp.ballot(j < n && cj < u));
    }
    return (n < 2) ? 0 : min(below, n - 1);
}

// pUCT traverse of tree i by a lane group (semantics and citations: select_puct_tree in lzb_select.cuh).
// `live`: this group has a tree (i < N); dead groups run along on tree 0 with every write predicated off.
// REUSE = cbatch_traverse_with_reuse (ReZero, cnode.cpp:826-928; ez :965-1073): at the root the child on the
// trajectory's true action is scored by carm_score (cnode.cpp:700-753) and, when chosen, ends the walk even if it
// is already expanded; such a tree needs no inference and reports ix = -1.
// VAR = LZB_SMZ (Stochastic MuZero, smz cnode.cpp:596-660,713-787): nodes at odd depth are chance nodes whose outcome
// is drawn from their priors (chance_draw_g) instead of scored; every going group of a warp is at the same depth, so
// the decision / chance branch is warp-uniform.
// experiment LZB_X_CARRY (fused step kernel only): what the backup half already holds in registers and the select half
// would otherwise load again - behind the backup's stores, so as one more exposed round trip at the head of the
// simulation's dependent chain - is handed over instead: the tree's updated MinMax pair and root record, plus the
// per-tree inputs no kernel of the search writes (to_play, the players word, the tie-break seed), which are requested at
// the top of the kernel.
struct Carry {
    float2 mm;
    int4 root;
    int to_play, players_word;
    unsigned long long tie_seed;
};

template <int VAR, int G, int R, bool REUSE = false>
__device__ __forceinline__ void select_puct_g(const Pool &p, const SelectArgs &a, int i, bool live, const Grp<G> &grp, int tl = -1,
                                              const Carry *carry = nullptr)
{
    constexpr bool VP = (VAR == LZB_EZ || VAR == LZB_SEZ);
    const int B = p.B, A = p.A, lane = grp.lane;
    const int4 *tree = p.tree(i);
    int4 *pa = p.path_a + (size_t)i * (p.S + 1);
    int2 *pb = p.path_b + (size_t)i * (p.S + 1);
    float2 mm = carry ? carry->mm : p.minmax[i];
    const int4 rt = carry ? carry->root : p.root[i];
    int vtp = carry ? carry->to_play : a.to_play[i];
    const int players = a.players ? a.players : (((carry ? carry->players_word : p.scratch[0]) == -1) ? 1 : 2); // cnode.cpp:776-781
    const float g = a.discount;
    float mm_delta = mm.y - mm.x;
    float mm_den = (mm_delta < a.delta) ? a.delta : mm_delta; // CMinMaxStats::normalize (cminimax.cpp:33-45)
#ifdef LZB_X_BRANCHFREE_DIV
    const float mm_safe = (mm_delta > 0) ? mm_den : 1.0f;            // the divisor every lane can divide by
#endif
    unsigned hroot = (a.rng_mode == LZB_RNG_PER_ROOT) ? (unsigned)(p.root_offset + i) : 0u;

    int parent_visits = rt.x;
    float parent_reward = __int_as_float(rt.z);
    int e = 0, depth = 0, pos = 0, action = 0, parent_reset = 0;
    int parent_chance = 0; // smz: the node the last choice was made at is a chance node
    float parent_q = 0.0f;
    bool is_root = true;
    bool going = live; // this group is still descending
    bool no_inference = false;
    int level = 0;   // rounds of the loop below == depth of every group that is still going (warp-uniform)
    const unsigned long long tie_seed = carry ? carry->tie_seed : (a.seed_dev ? *a.seed_dev : a.seed); // issued with the other entry loads
    unsigned long long chance_seed = tie_seed;
    if (VAR == LZB_SMZ && a.chance_seed) chance_seed = *a.chance_seed;
    bool is_true[R]; // root position j holds the trajectory's action
    float reuse_v = 0.0f;
    if (REUSE) {
        const int ta = a.true_action[i];
        reuse_v = a.reuse_value[i];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int j = lane + G * r;
            is_true[r] = (j < A) && p.legal[(size_t)i * A + j] == ta;
        }
    }

    LZB_TL(tl, 5);
    int *best_i = p.best + (size_t)i * p.E;
#ifdef LZB_X_CVAL
    const float *const cvt = p.cval + (size_t)i * p.E * B; // the tree's cached child values, indexed like its records
#endif
    // One level of the walk for the tree whose state is in the variables above, by the lanes of group `gx` with Rx child records
    // per lane.  Written once and used twice: by the trees' own lane groups (G lanes x R records), and - "lone walker" - by the
    // WHOLE warp (32 lanes x 1 record) for a tree whose warp sibling has finished its walk: the launch lasts as long as its
    // deepest trees (profiles/r02_step_timeline_L1024.txt), a deep tree's sibling has almost always ended levels earlier, and
    // with one record per lane a level needs no second pass over children 16, 17 of A = 18.
    int ti = i; // index of the tree being walked (the lone walker's tree in the second use)
    auto level_body = [&](const auto &gx, auto rtag) {
        constexpr int Gx = std::remove_reference_t<decltype(gx)>::kLanes, Rx = decltype(rtag)::value;
        // ---- one round of independent reads per level: the header (same address in every gx.lane) and Rx records per gx.lane
        const int4 *blk = tree + e * B; // e * B < 2^31: the whole forest is indexed with size_t only across trees
        const int4 h = blk[0];
#if defined(LZB_X_ROW_PREFETCH) && !defined(LZB_X_ROW_PREFETCH_AT_DESCENT)
        if (VAR != LZB_SMZ && going && !is_root && parent_visits <= LZB_X_ROW_PREFETCH) prefetch_rows_g<Gx>(p, a, ti, e, gx.lane);
#endif
        int4 c[Rx];
#ifdef LZB_X_CVAL
        float cvl[Rx]; // experiment: CNode::value() of each child as the backup stored it, same round as the records
#endif
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const int j = gx.lane + Gx * r;
            c[r] = (j < A) ? blk[1 + j] : make_int4(0, 0, 0, 0);
#ifdef LZB_X_CVAL
            cvl[r] = (j < A) ? cvt[e * B + 1 + j] : 0.0f;
#endif
        }
#ifdef LZB_X_PBC2
        // experiment: pb_c of every (parent visits, child visits) pair from a table built with the same expression
        const float *const pb_row = p.pbc2 ? p.pbc2 + (parent_visits - 1) * p.pbc2_w : nullptr;
        float2 pt = make_float2(0.0f, 0.0f);
        if (!pb_row) pt = p.pbc[parent_visits - 1];
        float pbt[Rx];
#else
        const float2 pt = p.pbc[parent_visits - 1]; // {logf((T + c2 + 1) / c2) + c1, sqrtf(T)}, T = parent visits - 1; same round as the block
#endif
        const int n = h.x;
#ifdef LZB_X_LONE_PREFETCH
        // experiment: the lone walker (few warps left, issue slots to spare) asks L1 for the block of every EXPANDED child of this
        // node - the next level's block is one of them - so that its read is an L1 hit instead of an L2 round trip
        if constexpr (Gx == 32) {
            const int ce0 = child_expansion(c[0].x);
            if (gx.lane < n && ce0 >= 0) {
                const char *nb = reinterpret_cast<const char *>(tree + ce0 * B);
                for (int off = 0; off < (B << 4); off += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(nb + off));
            }
        }
#endif
        LZB_CHK(p, !going || (e >= 0 && e < p.E), 1);
        LZB_CHK(p, !going || (n >= 0 && n <= A), 2);
        LZB_CHK(p, !going || (parent_visits >= 1 && parent_visits - 1 <= p.E), 4);
        if (level == 0) LZB_TLD(tl, 6, n ^ c[0].x ^ c[Rx - 1].x);

        // ---- compute_mean_q: serial fp32 sum over visited children in legal order
        float val[Rx], tr[Rx];
        int nvis = 0;
        float total = 0.0f;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const int j = gx.lane + Gx * r;
            const int vc = child_visits(c[r].x);
            const bool vis = (j < n) && vc > 0;
#ifdef LZB_X_PBC2
            pbt[r] = pb_row ? __ldg(pb_row + vc) : 0.0f; // consumed by cucb_score below, after the mean_q work
#endif
            const float rew = __int_as_float(c[r].w);
            tr[r] = rew;
            if (VP) tr[r] = (h.z == 1) ? rew : rew - parent_reward;
            val[r] = 0.0f;
            const unsigned wvis = __ballot_sync(kFull, vis);
            if (wvis) { // warp-uniform: slices without a visited child in any group add nothing
#if defined(LZB_X_CVAL)
                val[r] = vis ? cvl[r] : 0.0f; // CNode::value() = value_sum / visit_count, divided once by the backup
#elif !defined(LZB_X_BRANCHFREE_DIV) // the compiler branches around the division (BSSY / divergent BRA per slice)
                val[r] = vis ? __int_as_float(c[r].y) / (float)vc : 0.0f; // CNode::value()
#else // experiment (round 2, measured and rejected: 18.2 vs 17.5 us per step; ez 30.1 vs 29.7): every gx.lane divides (by 1
      // where there is nothing to divide) and selects - no divergent branch, but the full-precision division's ~10
      // dependent instructions then sit on every gx.lane's chain instead of on the lanes that need them
                const float vq = __int_as_float(c[r].y) / (float)max(vc, 1);
                val[r] = vis ? vq : 0.0f; // CNode::value()
#endif
                const float qsa = vis ? tr[r] + g * val[r] : 0.0f;
                nvis += __popc((wvis >> gx.shift) & ((Gx == 32) ? 0xffffffffu : ((1u << (Gx & 31)) - 1u)));
                // j = l + Gx*r ascending; +0.0f for the rest is exact; the last slice stops at the action space
#ifdef LZB_X_SPARSE_SUM
                // experiment: below the root a node has one to three visited children, so add only the lanes that hold a
                // visited child in SOME group of the warp (ascending, the other groups add +0.0f there)
                unsigned u = 0;
#pragma unroll
                for (int s = 0; s < 32; s += Gx) u |= wvis >> s;
                u &= (Gx == 32) ? 0xffffffffu : ((1u << (Gx & 31)) - 1u);
                if (__popc(u) <= 4) {
#pragma unroll 1
                    while (u) {
                        const int l = __ffs(u) - 1;
                        u &= u - 1;
                        total += gx.shfl(qsa, l);
                    }
                } else
#endif
                if constexpr (Rx == 1 && Gx == 32) total = gx.ordered_add_upto(total, qsa, A); // lone walker: 18 of 32 lanes hold a child
                else
                total = (r == Rx - 1 && Rx > 1) ? gx.ordered_add(total, qsa, A - Gx * (Rx - 1)) : gx.ordered_add(total, qsa, Gx);
            }
        }
        const bool root_mean = is_root && nvis > 0; // one division: the two forms differ in numerator and count only
        const float mean_q = (root_mean ? total : parent_q + total) / (float)(root_mean ? nvis : nvis + 1);
        if (level == 0) LZB_TLD(tl, 7, __float_as_int(mean_q));

        int newpos = 0;
        if (VAR == LZB_SMZ && (level & 1)) {
            // ---- chance node: draw the outcome (or replay a given one)
            const unsigned long long bits = chance_bits(chance_seed, (unsigned)a.sim, (unsigned)(p.root_offset + ti), (unsigned)depth);
            newpos = chance_draw_g<Gx, Rx>(c, n, (double)(bits >> 11) * (1.0 / 9007199254740992.0), gx);
            if (a.forced_outcome && going && depth < a.forced_stride) {
                const int fo = a.forced_outcome[(size_t)ti * a.forced_stride + depth];
                newpos = (fo >= 0 && fo < n) ? fo : newpos; // an outcome the node does not have is ignored
            }
        } else {
        // ---- cucb_score
        const float pbc0 = pt.x, sqT = pt.y;
        float sc[Rx];
        float lmax = -INFINITY;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const int j = gx.lane + Gx * r;
            const int vc = child_visits(c[r].x);
#ifdef LZB_X_PBC2
            const float pb_c = pb_row ? pbt[r] : pbc0 * (sqT / (float)(vc + 1));
#else
            const float pb_c = pbc0 * (sqT / (float)(vc + 1));
#endif
            float prior_score;
            if (VAR == LZB_SEZ) prior_score = pb_c / (float)n; // uniform branch, sez cnode.cpp:1078
            else prior_score = pb_c * __int_as_float(c[r].z);
            // carm_score: the visited arm takes the next state's searched value for its own and has no prior term
            const bool arm = REUSE && is_root && is_true[r] && vc > 0;
            const float cv = arm ? reuse_v : val[r];
            float vs;
            if (vc == 0) vs = mean_q;
            else vs = (players == 1) ? tr[r] + g * cv : tr[r] + g * (-cv);
#ifndef LZB_X_BRANCHFREE_DIV
            if (mm_delta > 0) vs = (vs - mm.x) / mm_den;
#else
            const float vn = (vs - mm.x) / mm_safe; // CMinMaxStats::normalize; discarded while the bounds are unset
            vs = (mm_delta > 0) ? vn : vs;
#endif
            if (vs < 0) vs = 0;
            if (vs > 1) vs = 1;
            sc[r] = (j < n) ? (arm ? vs : prior_score + vs) : -INFINITY;
            lmax = fmaxf(lmax, sc[r]);
        }
        const float gmax = gx.fmax(lmax);
        if (level == 0) LZB_TLD(tl, 8, __float_as_int(gmax));

        // ---- order-dependent tie set + pinned draw (cselect_child, cnode.cpp:567-596), branch-free:
        // jstar = first index attaining the maximum; tie set = later indices within epsilon; draw k in [0, later]
        const bool any_score = gmax > kFloatMin;
        int jstar = 0;
#pragma unroll
        for (int r = Rx - 1; r >= 0; --r) {
            const unsigned m = gx.ballot(sc[r] == gmax);
            jstar = m ? (__ffs(m) - 1) + Gx * r : jstar;
        }
        const float thr = gmax - 0.000001f;
        unsigned tie[Rx];
        int later = 0;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const int j = gx.lane + Gx * r;
            tie[r] = gx.ballot(j > jstar && sc[r] >= thr); // sc is -inf beyond n
            later += __popc(tie[r]);
        }
        newpos = jstar;
        if (__any_sync(kFull, later > 0)) { // warp-uniform: no group has a tie -> no draw
            // the stream position counts the reference's rand() calls: one per level, per DECISION level for smz
            const unsigned hdepth = (VAR == LZB_SMZ) ? (unsigned)(depth >> 1) : (unsigned)depth;
            const unsigned rnd = tiebreak_hash(tie_seed, (unsigned)a.sim, hroot, (a.rng_mode == LZB_RNG_PER_ROOT) ? hdepth : 0u);
            const int k = (int)(rnd % (unsigned)(later + 1));
            int before = 0; // tie-set members in earlier registers
#pragma unroll
            for (int r = 0; r < Rx; ++r) {
                // the gx.lane whose tie bit is the k-th one of the set announces itself
                const bool mybit = (tie[r] >> gx.lane) & 1u;
                const int rank = before + __popc(tie[r] & ((1u << gx.lane) - 1u)) + 1;
                const unsigned hit = gx.ballot(mybit && rank == k);
                newpos = hit ? (__ffs(hit) - 1) + Gx * r : newpos;
                before += __popc(tie[r]);
            }
        }
        newpos = any_score ? newpos : 0;
        }
        if (level == 0) LZB_TLD(tl, 9, newpos);
        level += 1;

        // ---- the chosen child's record lives in gx.lane newpos % Gx, register newpos / Gx
        const int rsel = newpos / Gx, lsel = newpos - rsel * Gx;
        int4 mine = c[0];
        bool mine_true = REUSE && is_true[0];
#pragma unroll
        for (int r = 1; r < Rx; ++r) {
            mine.x = (rsel == r) ? c[r].x : mine.x;
            mine.y = (rsel == r) ? c[r].y : mine.y;
            mine.w = (rsel == r) ? c[r].w : mine.w;
            if (REUSE) mine_true = (rsel == r) ? is_true[r] : mine_true;
        }
        bool chose_true = false; // cnode.cpp:895-898
        if (REUSE) { // the shuffle is executed by every gx.lane (no short-circuit around a full-mask exchange)
            const int sel_true = gx.shfl((int)mine_true, lsel);
            chose_true = is_root && sel_true != 0;
        }
        int4 ch; // the prior (z) of the chosen child is not needed below
        ch.x = gx.shfl(mine.x, lsel);
        ch.y = gx.shfl(mine.y, lsel);
        ch.w = gx.shfl(mine.w, lsel);
        ch.z = 0;
        LZB_CHK(p, !going || depth <= p.S, 3);
        LZB_CHK(p, !going || (newpos >= 0 && newpos < (n > 0 ? n : 1)), 5);
        if (going) {
            pos = newpos;
            parent_reset = (VAR == LZB_SMZ) ? 0 : h.z; // smz headers carry is_chance in that slot
            if (VAR == LZB_SMZ) parent_chance = h.z;
            if (players > 1) vtp = (vtp == 1) ? 2 : 1; // cnode.cpp:797-804
            if (gx.lane == 0) {
                pa[depth] = make_int4(e * B + 1 + pos, ch.x, ch.y, ch.w);
                pb[depth] = make_int2(h.y, h.z);
                best_i[e] = pos;
            }
            depth += 1;
            const int ce = child_expansion(ch.x);
            if (ce < 0 || chose_true) {
                going = false;
                no_inference = ce >= 0; // cnode.cpp:904-913
            } else {
                parent_visits = child_visits(ch.x);
                parent_reward = __int_as_float(ch.w);
                parent_q = mean_q;
                is_root = false;
                e = ce;
#if defined(LZB_X_ROW_PREFETCH) && defined(LZB_X_ROW_PREFETCH_AT_DESCENT) // the same request from inside the branch that descends
                if (VAR != LZB_SMZ && parent_visits <= LZB_X_ROW_PREFETCH) prefetch_rows_g<Gx>(p, a, ti, e, gx.lane);
#endif
            }
        }
        if (level == 1) LZB_TLD(tl, 10, e ^ depth);
        LZB_TL(tl, 11);
    };
    // Measured (round 2, -DLZB_X_NO_LONE_WALKER = before): mz 17.32 -> 16.70 us per step (tree-only 14.23 -> 13.22), sez 36.9 -> 35.5
    // (28.8 from 31.9 without the latent rows), ez 28.83 -> 28.94 (tree + LSTM rows 18.53 -> 17.91); the stochastic tree is
    // slower with it (18.48 -> 19.54: the chance draw's double-precision shuffles over 32 lanes, spills) and stays without,
    // -DLZB_X_LONE_WALKER_ALL switches it on for every flavour incl. reuse
    constexpr bool kLone = LoneWalker<VAR, G, R, REUSE>::value;
    bool rows_done = false; // lone-walker staging: the tree's row is already in the model batch
    bool lone_entered = false; // the warp's mbarrier has been used (LZB_X_PAIR_SPLIT)
    while (__any_sync(kFull, going)) {
        if constexpr (kLone) {
            const unsigned gm = __ballot_sync(kFull, going);
            if ((gm == 0x0000ffffu || gm == 0xffff0000u) && A <= 32 && level > 0) { // one of the warp's two trees is still walking
                const int src = (gm & 1u) ? 0 : 16;
                const bool walker = going;
                lone_entered = true;
                // the finished tree's results stay in its lanes: set aside what the code after the loop reads
                const int s_e = e, s_depth = depth, s_pos = pos, s_reset = parent_reset, s_vtp = vtp, s_chance = parent_chance;
                const bool s_noinf = no_inference;
#ifndef LZB_X_NO_LONE_STAGE
                bool s_staged = false;
                if (VAR != LZB_SMZ && lone_stage_rows(a)) { // the same in every lane
                    const long long rb0 = a.g[0].row_bytes;
                    char *stage = reinterpret_cast<char *>(lzb_tab_smem + kStageRowWord) + (threadIdx.x >> 5) * rb0;
                    if (!walker && live && !s_noinf && !(a.g[0].zero_if_reset && s_reset == 1)) {
                        const char *ss = row_src_g(p, a, 0, i, s_e);
                        if (ss != nullptr && (((unsigned long long)ss) & 15ull) == 0) {
                            const unsigned sm = (unsigned)__cvta_generic_to_shared(stage);
#ifndef LZB_X_LONE_STAGE_LSU // the same through the bulk-copy engine: one request by one lane, completion on an mbarrier
                            if (lane == 0) {
                                const unsigned mb = (unsigned)__cvta_generic_to_shared(lzb_tab_smem + kStageBarWord + (threadIdx.x >> 5));
                                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
                                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"((unsigned)rb0) : "memory");
                                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sm), "l"(ss),
                                             "r"((unsigned)rb0), "r"(mb)
                                             : "memory");
                            }
#else
                            for (int t = lane; t < (int)(rb0 >> 4); t += G)
                                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sm + t * 16), "l"(ss + (long long)t * 16) : "memory");
#endif
                            s_staged = true;
                        }
                    }
#ifdef LZB_X_LONE_STAGE_LSU
                    asm volatile("cp.async.commit_group;" ::: "memory");
#endif
                }
#endif
                // the walker's state and per-tree constants, in all 32 lanes
                ti = __shfl_sync(kFull, i, src);
                e = __shfl_sync(kFull, e, src); depth = __shfl_sync(kFull, depth, src); pos = __shfl_sync(kFull, pos, src);
                parent_visits = __shfl_sync(kFull, parent_visits, src); parent_reward = __shfl_sync(kFull, parent_reward, src);
                parent_q = __shfl_sync(kFull, parent_q, src); parent_reset = __shfl_sync(kFull, parent_reset, src);
                vtp = __shfl_sync(kFull, vtp, src);
                if (VAR == LZB_SMZ) parent_chance = __shfl_sync(kFull, parent_chance, src);
                mm.x = __shfl_sync(kFull, mm.x, src); mm.y = __shfl_sync(kFull, mm.y, src);
                mm_delta = mm.y - mm.x;
                mm_den = (mm_delta < a.delta) ? a.delta : mm_delta;
                hroot = (a.rng_mode == LZB_RNG_PER_ROOT) ? (unsigned)(p.root_offset + ti) : 0u;
                tree = p.tree(ti);
                pa = p.path_a + (size_t)ti * (p.S + 1);
                pb = p.path_b + (size_t)ti * (p.S + 1);
                best_i = p.best + (size_t)ti * p.E;
                is_root = false; going = true; no_inference = false;
                const Grp<32> g32;
                while (going) level_body(g32, std::integral_constant<int, 1>());
#ifdef LZB_X_LONE_EARLY_RELEASE // experiment: the warp's walks are over here - let the dependent launch go before the row work below
                pdl_release();
#endif
#ifndef LZB_X_NO_LONE_STAGE
                if (VAR != LZB_SMZ && lone_stage_rows(a)) { // every lane holds the walker's leaf here (nothing but s_staged kept across the walk)
                    const long long rb0 = a.g[0].row_bytes;
                    const char *stage = reinterpret_cast<const char *>(lzb_tab_smem + kStageRowWord) + (threadIdx.x >> 5) * rb0;
                    const int l32 = threadIdx.x & 31, n4 = (int)(rb0 >> 4);
                    const bool w_zero = a.g[0].zero_if_reset && parent_reset == 1;
                    const char *ws = (no_inference || w_zero) ? nullptr : row_src_g(p, a, 0, ti, e);
                    const bool w_fast = !no_inference && (w_zero || (ws != nullptr && (((unsigned long long)ws) & 15ull) == 0));
                    int4 v[8];
                    if (w_fast && !w_zero) {
                        const int4 *s4 = reinterpret_cast<const int4 *>(ws);
#pragma unroll
                        for (int u = 0; u < 8; ++u)
                            if (l32 + u * 32 < n4) v[u] = LZB_GATHER_LD(s4 + l32 + u * 32);
                    }
                    if (s_staged) { // the other tree's row: shared memory -> its row of the model batch
#ifndef LZB_X_LONE_STAGE_LSU
                        if (lane == 0) {
                            const unsigned mb = (unsigned)__cvta_generic_to_shared(lzb_tab_smem + kStageBarWord + (threadIdx.x >> 5));
                            const unsigned sm = (unsigned)__cvta_generic_to_shared(stage);
                            unsigned ok = 0;
                            while (!ok)
                                asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0; selp.u32 %0, 1, 0, q; }" : "=r"(ok) : "r"(mb) : "memory");
                            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.g[0].dst + (long long)i * rb0), "r"(sm),
                                         "r"((unsigned)rb0)
                                         : "memory");
                            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); // the block may exit once the engine has read the row
                        }
#else
                        asm volatile("cp.async.wait_all;" ::: "memory");
                        const int4 *m4 = reinterpret_cast<const int4 *>(stage);
                        int4 *d4 = reinterpret_cast<int4 *>(a.g[0].dst + (long long)i * rb0);
                        for (int t = lane; t < n4; t += G) d4[t] = m4[t];
#endif
                    }
                    if (w_fast) {
                        int4 *d4 = reinterpret_cast<int4 *>(a.g[0].dst + (long long)ti * rb0);
#pragma unroll
                        for (int u = 0; u < 8; ++u)
                            if (l32 + u * 32 < n4) LZB_GATHER_ST(d4 + l32 + u * 32, w_zero ? make_int4(0, 0, 0, 0) : v[u]);
                    }
                    rows_done = walker ? w_fast : s_staged;
                }
#endif
                if (!walker) { e = s_e; depth = s_depth; pos = s_pos; parent_reset = s_reset; vtp = s_vtp; no_inference = s_noinf; parent_chance = s_chance; }
                going = false;
                break;
            }
        }
        level_body(grp, std::integral_constant<int, R>());
    }

    // ---- gather of the parent's rows, then the leaf bookkeeping (cnode.cpp:814-822).  The loads the bookkeeping
    // needs (legal-list entry, batch index) are issued first and consumed after the gather, so the row traffic does
    // not queue behind them.
    int root_action = 0, batch_index = i;
    if (VAR != LZB_SEZ && e == 0) root_action = p.legal[(size_t)i * A + pos]; // root children sit in legal-list order
    if (REUSE && !no_inference) batch_index = p.bidx[(size_t)i * p.E + e];
    if (REUSE && a.infer_count) { // one atomic per warp
        const unsigned m = __ballot_sync(kFull, live && lane == 0 && !no_inference);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(a.infer_count, __popc(m));
    }
    pdl_release(); // the walk is over: a dependent launch may begin to take the place of exiting blocks
    LZB_TL(tl, 12);
#if defined(LZB_X_PAIR_SPLIT) && !defined(LZB_X_NO_LONE_STAGE)
    // experiment LZB_X_PAIR_SPLIT: a warp whose two trees ended their walks together (no lone walker) moves the second tree's row
    // through the bulk-copy engine (global -> shared -> global) while all 32 lanes gather the first tree's row in one round trip
    if constexpr (kLone) {
        const unsigned need = __ballot_sync(kFull, live && !no_inference && !rows_done);
        if (need == 0xffffffffu && !lone_entered && VAR != LZB_SMZ && lone_stage_rows(a) && !a.g[0].zero_if_reset) {
            const long long rb0 = a.g[0].row_bytes;
            const int l32 = threadIdx.x & 31, n4 = (int)(rb0 >> 4);
            const char *sb = row_src_g(p, a, 0, i, e);
            if (__all_sync(kFull, sb != nullptr && (((unsigned long long)sb) & 15ull) == 0)) {
                const unsigned sm = (unsigned)__cvta_generic_to_shared(reinterpret_cast<char *>(lzb_tab_smem + kStageRowWord) + (threadIdx.x >> 5) * rb0);
                const unsigned mb = (unsigned)__cvta_generic_to_shared(lzb_tab_smem + kStageBarWord + (threadIdx.x >> 5));
                if (l32 == 16) {
                    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
                    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"((unsigned)rb0) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sm), "l"(sb),
                                 "r"((unsigned)rb0), "r"(mb)
                                 : "memory");
                }
                const int4 *s4 = reinterpret_cast<const int4 *>(__shfl_sync(kFull, (unsigned long long)sb, 0));
                int4 *d4 = reinterpret_cast<int4 *>(a.g[0].dst + (long long)__shfl_sync(kFull, i, 0) * rb0);
                int4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (l32 + u * 32 < n4) v[u] = LZB_GATHER_LD(s4 + l32 + u * 32);
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (l32 + u * 32 < n4) LZB_GATHER_ST(d4 + l32 + u * 32, v[u]);
                if (l32 == 16) {
                    unsigned ok = 0;
                    while (!ok)
                        asm volatile("{ .reg .pred q; mbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0; selp.u32 %0, 1, 0, q; }" : "=r"(ok) : "r"(mb) : "memory");
                    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.g[0].dst + (long long)i * rb0), "r"(sm), "r"((unsigned)rb0)
                                 : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                }
                rows_done = true;
            }
        }
    }
#endif
    if (live && !no_inference) { // nothing to infer: the tree's rows of the model batch are left as they are
        if (VAR == LZB_SEZ) {
            const float *act = p.actions + (((size_t)i * p.E + e) * A + pos) * p.D;
            if (a.last_action_f)
                for (int d = lane; d < p.D; d += G) a.last_action_f[(size_t)i * p.D + d] = act[d];
            if (a.last_action && lane == 0) a.last_action[i] = (long long)act[0]; // discrete sampled: float(index)
        }
        if (!rows_done) gather_rows_g<G>(p, a, i, e, parent_reset, lane, (VAR == LZB_SMZ && a.slot_by_kind) ? (parent_chance ? 1 : 0) : -1);
    }
    LZB_TL(tl, 13);
    if (live && lane == 0) {
        action = (VAR != LZB_SEZ && e == 0) ? root_action : pos;
        p.res[i] = make_int4(depth, no_inference ? -1 : e, pos, vtp);
        if (a.ix) a.ix[i] = no_inference ? -1 : e;
        if (a.iy) a.iy[i] = batch_index;
        if (a.search_len) a.search_len[i] = depth;
        if (a.vtp) a.vtp[i] = vtp;
        if (VAR != LZB_SEZ && a.last_action) a.last_action[i] = (long long)action;
        if (VAR == LZB_SMZ && a.leaf_is_chance) a.leaf_is_chance[i] = depth & 1; // smz cnode.cpp:782
    }
#ifdef LZB_TIMELINE
    {
        unsigned sm;
        asm volatile("mov.u32 %0, %smid;" : "=r"(sm));
        tl_store(tl, 15, (unsigned long long)sm | ((unsigned long long)level << 16) | ((unsigned long long)depth << 32));
    }
#endif
    LZB_TL(tl, 14);
}

// csoftmax (gmz cnode.cpp:904-933) over the n values a lane group holds as x[r] for j = lane + G*r:
// exp(x - max - log(sum)) with the sum taken strictly in index order.
// `need[r]`: this lane's slot r of the result is used afterwards; a slice nobody in the warp needs is left as it is
// (its final expf is skipped), and with no slice needed at all nothing is computed.
template <int G, int R>
__device__ __forceinline__ bool csoftmax_g(float (&x)[R], int n, int A, const Grp<G> &grp, const bool (&need)[R], float *m_out = nullptr,
                                           float *ls_out = nullptr)
{
    const int lane = grp.lane;
    bool any_need[R], some = false; // warp-uniform
#pragma unroll
    for (int r = 0; r < R; ++r) {
        any_need[r] = __any_sync(kFull, need[r]);
        some = some || any_need[r];
    }
    if (!some) return false;
    float m = -INFINITY;
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (lane + G * r < n) m = fmaxf(m, x[r]);
    m = grp.fmax(m);
    float sum = 0.0f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const float ex = (lane + G * r < n) ? glibc_expf_s(x[r] - m) : 0.0f;
        // index order; +0.0f beyond n is exact; the last slice stops at the action space
        sum = grp.template ordered_slice<R>(sum, ex, r, A);
    }
    const float ls = glibc_logf_s(sum);
    if (m_out) { *m_out = m; *ls_out = ls; }
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (any_need[r]) x[r] = glibc_expf_s(x[r] - m - ls);
    return true;
}

template <int G, int R>
__device__ __forceinline__ void csoftmax_g(float (&x)[R], int n, int A, const Grp<G> &grp)
{
    bool need[R];
#pragma unroll
    for (int r = 0; r < R; ++r) need[r] = true;
    csoftmax_g<G, R>(x, n, A, grp, need);
}

// Gumbel MuZero traverse of tree i by a lane group: sequential halving at the root (cselect_root_child, gmz
// cnode.cpp:701-745, score_considered :1097-1132), improved-policy selection below it (cselect_interior_child
// :747-790), both on completed Q values (qtransform_completed_by_mix_value :989-1040 with the header defaults).
// Semantics and evaluation order are those of select_gumbel_tree in lzb_select.cuh; every group of a warp is at
// the root in the first round and below it afterwards, so the root / interior branch is warp-uniform.
template <int G, int R>
__device__ __forceinline__ void select_gumbel_g(const Pool &p, const SelectArgs &a, int i, bool live, const Grp<G> &grp)
{
    const int B = p.B, A = p.A, lane = grp.lane;
    const int4 *tree = p.tree(i);
    int4 *pa = p.path_a + (size_t)i * (p.S + 1);
    int2 *pb = p.path_b + (size_t)i * (p.S + 1);
    const float g = a.discount;
    const int vtp = a.to_play[i]; // never flipped (gmz cnode.cpp:870-885)
    float gum[R];                 // root only: the shared Gumbel noise by position
#pragma unroll
    for (int r = 0; r < R; ++r) gum[r] = (lane + G * r < A) ? p.gumbel[lane + G * r] : 0.0f;
    int e = 0, depth = 0, pos = 0;
    bool going = live, at_root = true;

    // the level body once, used by the tree's own lane group and by the whole warp for a lone walker (see select_puct_g)
    int ti = i;
    auto level_body = [&](const auto &gx, auto rtag) {
        constexpr int Gx = std::remove_reference_t<decltype(gx)>::kLanes, Rx = decltype(rtag)::value;
        const int4 *blk = tree + (size_t)e * B;
        const int4 h = blk[0];
        int4 c[Rx];
#pragma unroll
        for (int r = 0; r < Rx; ++r) c[r] = (gx.lane + Gx * r < A) ? blk[1 + gx.lane + Gx * r] : make_int4(0, 0, 0, 0);
        const int n = h.x;
        LZB_CHK(p, !going || (e >= 0 && e < p.E), 1);
        LZB_CHK(p, !going || (n >= 0 && n <= A), 2);
        LZB_CHK(p, !going || depth <= p.S, 3);
        const float raw_value = __int_as_float(h.w);

        // ---- get_q (:181-197), priors, visit statistics
        float q[Rx], sp[Rx];
        int vc[Rx], vsum_i = 0, vmax_i = 0;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const bool in = gx.lane + Gx * r < n;
            vc[r] = in ? child_visits(c[r].x) : 0;
            const float val = vc[r] ? __int_as_float(c[r].y) / (float)vc[r] : 0.0f;
            q[r] = __int_as_float(c[r].w) + g * val;
            sp[r] = __int_as_float(c[r].z);
            vsum_i += vc[r];
            vmax_i = max(vmax_i, vc[r]);
        }
        vsum_i = gx.isum(vsum_i);
        vmax_i = gx.imax(vmax_i);
        // priors (probabilities) are soft-maxed again (:1014) - reproduced as is; only the entries of visited children
        // are read below (compute_mixed_value sums over them), so the rest is not evaluated
        bool sp_need[Rx];
#pragma unroll
        for (int r = 0; r < Rx; ++r) sp_need[r] = going && vc[r] > 0;
#ifndef LZB_X_NO_GMZ_STATS
        // a node's priors never change after its expansion, so the maximum and the log-sum-exp of this softmax are
        // per-node constants: computed the first time they are needed, kept in gstat, and every later visit evaluates only
        // expf(prior - max - lse) for the visited children (no reduction, no ordered sum, no logf)
        {
            float2 st = make_float2(0.0f, 0.0f);
            bool have = true;
            if (p.gstat) {
                st = p.gstat[(size_t)ti * p.E + e];
                have = !(st.x != st.x);
            }
            bool want = false;
#pragma unroll
            for (int r = 0; r < Rx; ++r) want = want || sp_need[r];
            want = gx.ballot(want) != 0u; // the group needs the softmax at all
            if (p.gstat && !__any_sync(kFull, want && !have)) { // warp-uniform: every group that needs it has it
#pragma unroll
                for (int r = 0; r < Rx; ++r)
                    if (__any_sync(kFull, sp_need[r])) sp[r] = glibc_expf_s(sp[r] - st.x - st.y);
            } else {
                float m_out, ls_out;
                const bool done = csoftmax_g<Gx, Rx>(sp, n, A, gx, sp_need, &m_out, &ls_out);
                if (p.gstat && done && want && !have && gx.lane == 0) p.gstat[(size_t)ti * p.E + e] = make_float2(m_out, ls_out);
            }
        }
#else
        csoftmax_g<Gx, Rx>(sp, n, A, gx, sp_need);
#endif

        // ---- compute_mixed_value (:935-970): serial sums over visited children in index order
        float probs_sum = 0.0f;
#ifdef LZB_X_GMZ_SPARSE
        unsigned vmask[Rx];
#endif
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const float pc = (sp[r] < -10e7f) ? -10e7f : sp[r]; // std::max(prior, min_num)
            const float t = (vc[r] > 0) ? pc : 0.0f;
#ifdef LZB_X_GMZ_SPARSE
            vmask[r] = gx.union_mask(vc[r] > 0); // below the root one to three children are visited: add only those lanes
            if (vmask[r]) {
                if (__popc(vmask[r]) <= 6) probs_sum = gx.ordered_add_sparse(probs_sum, t, vmask[r]);
                else probs_sum = (r == Rx - 1 && Rx > 1) ? gx.ordered_add(probs_sum, t, A - Gx * (Rx - 1)) : gx.ordered_add(probs_sum, t, Gx);
            }
#else
            if (__any_sync(kFull, vc[r] > 0)) // warp-uniform: a slice without a visited child adds only +0.0f
                probs_sum = gx.template ordered_slice<Rx>(probs_sum, t, r, A);
#endif
        }
        float wq = 0.0f;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            const float pc = (sp[r] < -10e7f) ? -10e7f : sp[r];
            const float t = (vc[r] > 0) ? pc * q[r] / probs_sum : 0.0f;
#ifdef LZB_X_GMZ_SPARSE
            if (vmask[r]) {
                if (__popc(vmask[r]) <= 6) wq = gx.ordered_add_sparse(wq, t, vmask[r]);
                else wq = (r == Rx - 1 && Rx > 1) ? gx.ordered_add(wq, t, A - Gx * (Rx - 1)) : gx.ordered_add(wq, t, Gx);
            }
#else
            if (__any_sync(kFull, vc[r] > 0))
                wq = gx.template ordered_slice<Rx>(wq, t, r, A);
#endif
        }
        const float vsum_f = (float)vsum_i;
        const float mixed = (raw_value + vsum_f * wq) / (vsum_f + 1);
        float cq[Rx];
        float mx = -INFINITY, mn = INFINITY;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            cq[r] = (vc[r] > 0) ? q[r] : mixed;
            if (gx.lane + Gx * r < n) { mx = fmaxf(mx, cq[r]); mn = fminf(mn, cq[r]); }
        }
        mx = gx.fmax(mx);
        mn = gx.fmin(mn);
        float gap = mx - mn; // rescale_qvalues (:972-987)
        gap = (gap < 1e-8f) ? 1e-8f : gap;
        const float visit_scale = 50.0f + (float)vmax_i;
#pragma unroll
        for (int r = 0; r < Rx; ++r) cq[r] = (cq[r] - mn) / gap * visit_scale * 0.1f;

        // ---- scores
        float sc[Rx];
        if (at_root) { // cselect_root_child + score_considered
            LZB_CHK(p, !going || (vsum_i >= 0 && vsum_i < p.E), 6);
            const int considered = p.vtable[vsum_i];
            float ml = -INFINITY;
#pragma unroll
            for (int r = 0; r < Rx; ++r)
                if (gx.lane + Gx * r < n) ml = fmaxf(ml, __int_as_float(c[r].z));
            ml = gx.fmax(ml);
#pragma unroll
            for (int r = 0; r < Rx; ++r) {
                const float logit = __int_as_float(c[r].z) - ml;
                float v = gum[r] + logit + cq[r];
                v = (-1e9f < v) ? v : -1e9f;
                v += (vc[r] == considered) ? 0.0f : -INFINITY;
                sc[r] = v;
            }
        } else { // cselect_interior_child
#pragma unroll
            for (int r = 0; r < Rx; ++r) sc[r] = __int_as_float(c[r].z) + cq[r];
            csoftmax_g<Gx, Rx>(sc, n, A, gx);
            const float den = (float)(1 + vsum_i);
#pragma unroll
            for (int r = 0; r < Rx; ++r) sc[r] = sc[r] - (float)vc[r] / den;
        }

        // ---- first index attaining the maximum under a strict '>' scan from -inf (:733-742,778-787); 0 if none
        float lmax = -INFINITY;
#pragma unroll
        for (int r = 0; r < Rx; ++r) {
            sc[r] = (gx.lane + Gx * r < n) ? sc[r] : -INFINITY;
            lmax = fmaxf(lmax, sc[r]);
        }
        const float gmax = gx.fmax(lmax);
        int newpos = 0;
#pragma unroll
        for (int r = Rx - 1; r >= 0; --r) {
            const unsigned m = gx.ballot(sc[r] == gmax && gmax > -INFINITY);
            newpos = m ? (__ffs(m) - 1) + Gx * r : newpos;
        }

        // ---- the chosen child's record lives in gx.lane newpos % Gx, register newpos / Gx
        const int rsel = newpos / Gx, lsel = newpos - rsel * Gx;
        int4 mine = c[0];
#pragma unroll
        for (int r = 1; r < Rx; ++r) {
            mine.x = (rsel == r) ? c[r].x : mine.x;
            mine.y = (rsel == r) ? c[r].y : mine.y;
            mine.w = (rsel == r) ? c[r].w : mine.w;
        }
        int4 ch;
        ch.x = gx.shfl(mine.x, lsel);
        ch.y = gx.shfl(mine.y, lsel);
        ch.w = gx.shfl(mine.w, lsel);
        ch.z = 0;
        if (going) {
            pos = newpos;
            if (gx.lane == 0) {
                pa[depth] = make_int4(e * B + 1 + pos, ch.x, ch.y, ch.w);
                pb[depth] = make_int2(h.y, h.z);
                p.best[(size_t)ti * p.E + e] = pos;
            }
            depth += 1;
            const int ce = child_expansion(ch.x);
            if (ce < 0) going = false;
            else e = ce;
        }
        at_root = false;
    };
    // Gumbel: measured slower with the lone walker (22.59 -> 22.92 us per step), so only with -DLZB_X_LONE_WALKER_ALL
#ifdef LZB_X_LONE_WALKER_ALL
    constexpr bool kLone = G == 16 && R == 2;
#else
    constexpr bool kLone = false;
#endif
    while (__any_sync(kFull, going)) {
        if constexpr (kLone) {
            const unsigned gm = __ballot_sync(kFull, going);
            if ((gm == 0x0000ffffu || gm == 0xffff0000u) && A <= 32 && !at_root) {
                const int src = (gm & 1u) ? 0 : 16;
                const bool walker = going;
                const int s_e = e, s_depth = depth, s_pos = pos;
                ti = __shfl_sync(kFull, i, src);
                e = __shfl_sync(kFull, e, src); depth = __shfl_sync(kFull, depth, src); pos = __shfl_sync(kFull, pos, src);
                tree = p.tree(ti);
                pa = p.path_a + (size_t)ti * (p.S + 1);
                pb = p.path_b + (size_t)ti * (p.S + 1);
                going = true;
                const Grp<32> g32;
                while (going) level_body(g32, std::integral_constant<int, 1>());
                if (!walker) { e = s_e; depth = s_depth; pos = s_pos; }
                going = false;
                break;
            }
        }
        level_body(grp, std::integral_constant<int, R>());
    }

    int root_action = 0;
    if (e == 0) root_action = p.legal[(size_t)i * A + pos];
    pdl_release();
    if (live) gather_rows_g<G>(p, a, i, e, 0, lane);
    if (live && lane == 0) {
        p.res[i] = make_int4(depth, e, pos, vtp);
        if (a.ix) a.ix[i] = e;
        if (a.iy) a.iy[i] = i;
        if (a.search_len) a.search_len[i] = depth;
        if (a.vtp) a.vtp[i] = vtp;
        if (a.last_action) a.last_action[i] = (long long)((e == 0) ? root_action : pos);
    }
}

// REUSE = cbatch_backpropagate_with_reuse (cnode.cpp:502-549; ez :603-644): a tree whose traverse ended on an expanded
// node is not expanded again and backs up reuse_value; a tree that expanded the root's true-action child backs up
// reuse_value instead of the model's value.
template <int VAR, int G, int R, bool REUSE = false>
__device__ __forceinline__ void expand_backup_g(const Pool &p, const BackupArgs &a, int i, bool live, const Grp<G> &grp, int tl = -1,
                                                Carry *carry = nullptr)
{
    constexpr bool VP = (VAR == LZB_EZ || VAR == LZB_SEZ);
    const int A = p.A, lane = grp.lane;
    const int4 res = p.res[i];
    const int depth = res.x; // search_len >= 1
    LZB_CHK(p, !live || (depth >= 1 && depth <= p.S), 7);
    const int e_new = a.latent_index;
    LZB_CHK(p, e_new >= 1 && e_new < p.E, 8);
    const int to_play = a.to_play[i];
    float value = a.value[i];
    float leaf_reward = a.reward[i];
    if (VAR == LZB_SMZ && a.value_c && (depth & 1)) { // a chance leaf takes the other head's outputs
        value = a.value_c[i];
        leaf_reward = a.reward_c[i];
    }
    bool no_inference = false;
    float value_propagate = value;
    if (REUSE) {
        bool reuse;
        if (a.reuse_flags) {
            const int fl = a.reuse_flags[i];
            no_inference = fl & 1;
            reuse = fl & 2;
        } else { // what the reference driver derives from the traverse outputs (mcts_ctree.py:376-384)
            no_inference = res.y == -1;
            reuse = res.y == 0 && p.legal[(size_t)i * A + res.z] == a.true_action[i];
        }
        if (no_inference || reuse) value_propagate = a.reuse_value[i];
    }
    const bool grow = live && !no_inference; // this tree expands its leaf
    // Stochastic MuZero (smz cnode.cpp:96-166): a leaf at odd depth is a chance node - C chance logits, reward 0,
    // decision children; at even depth a decision node with Ad policy logits
    const bool leaf_chance = (VAR == LZB_SMZ) && (depth & 1);
    const int width = (VAR == LZB_SMZ) ? (leaf_chance ? p.C : p.Ad) : A;
    const int4 *pa = p.path_a + (size_t)i * (p.S + 1);
    const int2 *pb = p.path_b + (size_t)i * (p.S + 1);
    const int4 root_rec = p.root[i]; // independent of the depth: issued now
    constexpr bool kPow2G = (G & (G - 1)) == 0;
    (void)kPow2G;
#ifdef LZB_X_EARLY_PATH
    // experiment: the first chunk's path records (one per lane) and the tree's MinMax pair are requested HERE, before the
    // leaf's softmax, whose expf work then hides their round trip (they were requested after it: +1.0 k and +0.5 k cycles
    // of exposed latency in profiles/r01_step_timeline_*.txt).  Nothing below writes what these loads read.
    int4 pre_r = make_int4(0, 0, 0, 0);
    int pre_tp = 0, pre_reset = 0;
    float pre_vp = 0.0f;
    {
        const int q = depth - lane;
        if (VAR != LZB_GMZ && live && to_play != -1 && q >= 0 && q != depth) pre_tp = pb[q].x; // node 0's too (the root's to_play)
        if (live && q >= 1) {
            pre_r = pa[q - 1];
            if (VP) {
                if (q >= 2) pre_vp = __int_as_float(pa[q - 2].w);
                pre_reset = pb[q - 1].y;
            }
        }
    }
    const float2 pre_mm = p.minmax[i];
#endif
    int leaf_reset = 0;
    if (VP) {
        if (a.is_reset) leaf_reset = a.is_reset[i];
        else if (a.horizon > 0) leaf_reset = (depth % a.horizon == 0) ? 1 : 0; // mcts_ctree.py:810
        if (a.is_reset_out && lane == 0 && live) a.is_reset_out[i] = leaf_reset;
        if (REUSE && no_inference && lane == 0 && live) { // results.nodes[i]->is_reset = is_reset_list[i] (ez :640)
            const int ce = child_expansion(pa[depth - 1].y);
            reinterpret_cast<int *>(p.block(i, ce))[2] = leaf_reset;
        }
    }

    // ---- expand the leaf as expansion e_new: masked softmax over all A actions (cnode.cpp:118-140)
    int4 *dst = p.block(i, e_new) + 1;
    if (VAR == LZB_SEZ) {
        if (grow) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (lane + G * r < A) dst[lane + G * r] = make_int4(0, 0, 0, 0);
            const int KD = A * p.D;
            float *tab = p.actions + ((size_t)i * p.E + e_new) * KD;
            for (int t = lane; t < KD; t += G) tab[t] = a.sampled_actions[(size_t)i * KD + t];
        }
    } else {
        const float *logits = a.logits + (size_t)i * ((VAR == LZB_SMZ && a.logits_stride) ? a.logits_stride : A);
        if (VAR == LZB_SMZ && a.logits_c && leaf_chance) logits = a.logits_c + (size_t)i * a.logits_c_stride;
        float lg[R], t[R];
        float mx = kFloatMin; // cnode.cpp:118
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int j = lane + G * r;
            lg[r] = (j < width) ? __ldcs(logits + j) : 0.0f;
            if (j < width) mx = fmaxf(mx, lg[r]);
        }
        mx = grp.fmax(mx);
        LZB_TLD(tl, 16, __float_as_int(mx) ^ depth);
        float sum = 0.0f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int j = lane + G * r;
            t[r] = (j < width) ? glibc_expf_s(lg[r] - mx) : 0.0f;
            if (r == R - 1) LZB_TLD(tl, 20, __float_as_int(t[r]));
            // strictly left to right in action order; +0.0f beyond the width; the last slice stops at the action space
            sum = (r == R - 1 && R > 1) ? grp.ordered_add(sum, t[r], A - G * (R - 1)) : grp.ordered_add(sum, t[r], G);
        }
        LZB_TLD(tl, 17, __float_as_int(sum));
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int j = lane + G * r;
            if (j < width && grow) dst[j] = make_int4(0, 0, __float_as_int(t[r] / sum), 0);
        }
    }
    if (lane == 0 && grow) {
        dst[-1] = make_int4(width, to_play, (VAR == LZB_SMZ) ? (int)leaf_chance : leaf_reset, __float_as_int(value));
        p.best[(size_t)i * p.E + e_new] = -1;
        if (REUSE) p.bidx[(size_t)i * p.E + e_new] = a.batch_index ? a.batch_index[i] : i;
    }
    if (REUSE && lane == 0 && live && no_inference) dst[-1] = make_int4(0, 0, 0, 0); // this tree has no node e_new (tree dumps skip it)

    LZB_TL(tl, 3);
    // ---- backup: node q = depth (leaf) ... 1 are path records q-1, node 0 is the root
    const bool two = (VAR != LZB_GMZ) && (to_play != -1); // cnode.cpp:432; gmz asserts -1 (compiled out)
    const float g = a.discount;
    float b = value_propagate; // bootstrap_value entering the current chunk
    float cmin = kFloatMax, cmax = kFloatMin;
#ifdef LZB_X_CARRY
    int4 new_root = root_rec;
#endif
    for (int qhi = depth; __any_sync(kFull, qhi >= 0); qhi -= G) { // every group runs as many chunks as the deepest one
        const int q = qhi - lane;
        const bool valid = live && q >= 0;
        int x = 0, slot = -1, node_tp = 0, parent_reset = 0;
        float vsum = 0.f, rew = 0.f, parent_vp = 0.f;
        if (valid) {
            if (q == 0) {
                x = root_rec.x; vsum = __int_as_float(root_rec.y); rew = __int_as_float(root_rec.z);
            } else {
#ifdef LZB_X_EARLY_PATH
                const int4 r = (qhi == depth) ? pre_r : pa[q - 1];
#else
                const int4 r = pa[q - 1];
#endif
                slot = r.x; x = r.y; vsum = __int_as_float(r.z); rew = __int_as_float(r.w);
                if (q == depth && !no_inference) rew = leaf_chance ? 0.0f : leaf_reward; // expand() stamps the leaf's reward / value_prefix first
            }
#ifdef LZB_X_EARLY_PATH
            if (two) node_tp = (q == depth) ? to_play : ((qhi == depth) ? pre_tp : pb[q].x);
            if (VP && q >= 1) {
                parent_vp = (q == 1) ? __int_as_float(root_rec.z) : ((qhi == depth) ? pre_vp : __int_as_float(pa[q - 2].w));
                parent_reset = (qhi == depth) ? pre_reset : pb[q - 1].y;
            }
#else
            if (two) node_tp = (q == depth) ? to_play : pb[q].x;
            if (VP && q >= 1) {
                parent_vp = (q == 1) ? __int_as_float(root_rec.z) : __int_as_float(pa[q - 2].w);
                parent_reset = pb[q - 1].y;
            }
#endif
        }
        if (qhi == depth) LZB_TLD(tl, 18, __float_as_int(rew) ^ x);
        float r_pre = rew, r_boot = rew;
        if (VP) {
            r_pre = rew - parent_vp;                    // ez cnode.cpp:515
            r_boot = (parent_reset == 1) ? rew : r_pre; // :518-522
        }
        const bool same = two && (node_tp == to_play);
        const float r_signed = same ? -r_boot : r_boot;
        // serial leaf->root recurrence (cnode.cpp:447,473-475): lane l receives b_q, then b = r + g*b
        float my_b = 0.f;
        const int cnt = min(G, qhi + 1);
        const int cnt_w = __reduce_max_sync(kFull, live ? cnt : 0); // the longest path piece in the warp
#pragma unroll 1
        for (int l = 0; l < cnt_w; ++l) {
            const float rl = grp.shfl(r_signed, l);
            if (lane == l) my_b = b;
            b = (l < cnt) ? rl + g * b : b;
        }
        if (qhi == depth) LZB_TLD(tl, 19, __float_as_int(b));
        if (valid) {
            const float add = (two && !same) ? -my_b : my_b;
            vsum += add;
            const int visits = (q == 0 ? x : child_visits(x)) + 1;
            const float nv = vsum / (float)visits;
            float cand;
            if (VP) cand = r_pre + g * nv;      // ez :516,558 (value not negated)
            else if (two) cand = rew + g * -nv; // mz :470
            else cand = rew + g * nv;           // mz :445, gmz :627
            cmin = fminf(cmin, cand);
            cmax = fmaxf(cmax, cand);
            if (q == 0) {
                p.root[i] = make_int4(visits, __float_as_int(vsum), __float_as_int(rew), 0);
#ifdef LZB_X_CARRY
                new_root = make_int4(visits, __float_as_int(vsum), __float_as_int(rew), 0);
#endif
            } else {
                int4 *rec = p.tree(i) + slot;
                LZB_CHK(p, slot >= 1 && slot < p.E * p.B, 9);
#ifdef LZB_X_CVAL
                if (VAR != LZB_GMZ) p.cval[(size_t)i * p.E * p.B + slot] = nv; // what select would divide out again at every later visit
#endif
                const int exp_idx = (q == depth && !no_inference) ? e_new : child_expansion(x);
                *reinterpret_cast<int2 *>(rec) = make_int2(child_pack(visits, exp_idx), __float_as_int(vsum));
                if (q == depth) reinterpret_cast<int *>(rec)[3] = __float_as_int(rew);
            }
        }
    }
    cmin = grp.fmin(cmin);
    cmax = grp.fmax(cmax);
    LZB_TLD(tl, 21, __float_as_int(cmin) ^ __float_as_int(cmax));
#ifdef LZB_X_CARRY
    {   // every lane holds the pair (same-address load above), so every lane can form the update; lane 0 stores it
        float2 mm = pre_mm;
        mm.x = fminf(mm.x, cmin); // CMinMaxStats::update (cminimax.cpp:23-31)
        mm.y = fmaxf(mm.y, cmax);
        if (lane == 0 && live) p.minmax[i] = mm;
        if (carry) {
            carry->mm = mm;
            const int src = kPow2G ? (depth & (G - 1)) : (depth % G); // the lane that held node 0 in the last chunk
            carry->root = grp.shfl(new_root, src);
        }
    }
#else
    if (lane == 0 && live) {
#ifdef LZB_X_EARLY_PATH
        float2 mm = pre_mm;
#else
        float2 mm = p.minmax[i];
#endif
        mm.x = fminf(mm.x, cmin); // CMinMaxStats::update (cminimax.cpp:23-31)
        mm.y = fmaxf(mm.y, cmax);
        p.minmax[i] = mm;
    }
#endif
    LZB_TL(tl, 4);
}

// ---------------------------------------------------------------------------------------------- kernels
// 128 threads = 4 * (32 / G) trees per block; groups without a tree (tail of the last block, spare lanes of a warp) run
// along on tree 0, muted
template <int VAR, int G, int R, bool REUSE>
__global__ void __launch_bounds__(128, LZB_MIN_BLOCKS) k_select_g(Pool p, SelectArgs a)
{
    fill_shared_tables();
    const Grp<G> grp;
    bool has_tree;
    const int i = group_tree<G>(has_tree);
    const bool live = has_tree && i < p.N;
    pdl_wait();
    if constexpr (VAR == LZB_GMZ) select_gumbel_g<G, R>(p, a, live ? i : 0, live, grp);
    else select_puct_g<VAR, G, R, REUSE>(p, a, live ? i : 0, live, grp);
}

template <int VAR, int G, int R, bool REUSE>
__global__ void __launch_bounds__(128, LZB_MIN_BLOCKS) k_expand_backup_g(Pool p, BackupArgs b)
{
    fill_shared_tables();
    const Grp<G> grp;
    bool has_tree;
    const int i = group_tree<G>(has_tree);
    bool live = has_tree && i < p.N;
    pdl_wait();
    if (b.active && live) live = b.active[i] != 0; // smz leaf_idx_list: the other trees are not touched by this call
    if (blockIdx.x == 0 && threadIdx.x < b.n_slots) p.slots[(size_t)threadIdx.x * p.E + b.latent_index] = pick3(b.new_slot, threadIdx.x);
    expand_backup_g<VAR, G, R, REUSE>(p, b, live ? i : 0, live, grp);
    pdl_release();
}

// fused simulation step (see lzb_step.cuh): expand + backup of simulation s, then select + gather of s + 1
template <int VAR, int G, int R, bool REUSE>
__global__ void __launch_bounds__(128, LZB_MIN_BLOCKS) k_step_g(Pool p, BackupArgs b, SelectArgs a)
{
    fill_shared_tables();
    const Grp<G> grp;
    bool has_tree;
    const int i = group_tree<G>(has_tree);
    const bool live = has_tree && i < p.N;
#ifdef LZB_TIMELINE
    const int tl = b.latent_index;
#else
    const int tl = -1;
#endif
    LZB_TL(tl, 0);
    LZB_TL(tl, 1);
    pdl_wait();
    LZB_TL(tl, 2);
    if (blockIdx.x == 0 && threadIdx.x < b.n_slots) p.slots[(size_t)threadIdx.x * p.E + b.latent_index] = pick3(b.new_slot, threadIdx.x);
#ifdef LZB_X_CARRY
    if constexpr (VAR != LZB_GMZ) {
        Carry carry;
        const int it = live ? i : 0;
        carry.to_play = a.to_play[it];
        carry.players_word = p.scratch[0];
        carry.tie_seed = a.seed_dev ? *a.seed_dev : a.seed;
        expand_backup_g<VAR, G, R, REUSE>(p, b, it, live, grp, tl, &carry);
        __syncwarp();
        select_puct_g<VAR, G, R, REUSE>(p, a, it, live, grp, tl, &carry);
        return;
    }
#endif
    expand_backup_g<VAR, G, R, REUSE>(p, b, live ? i : 0, live, grp, tl);
    __syncwarp(); // the warp's own global writes (records, root, MinMax) are ordered before its reads below
    if constexpr (VAR == LZB_GMZ) select_gumbel_g<G, R>(p, a, live ? i : 0, live, grp);
    else select_puct_g<VAR, G, R, REUSE>(p, a, live ? i : 0, live, grp, tl);
}

} // namespace lzb
