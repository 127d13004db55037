// Node-pool layout shared by all kernels (see DESIGN.md "Data layout in HBM").
//
// One search expands exactly one node per tree per simulation (mcts_ctree.py:255-321), so node
// allocation is arithmetic: the node expanded with current_latent_state_index e (0 = root,
// e = simulation_index + 1 otherwise, cnode.cpp:497) owns block e of its tree.  A block is
// (A + 1) 16-byte records: the node's own header followed by its A child records, so ONE
// dependent read per tree level fetches everything selection needs (lane 0 the header, lanes
// 1..A the children).  A child is addressed by (e, j), j = position in the parent's legal-action
// list (== the action id for every non-root node, cnode.cpp:101-107,138-139).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/lzb200.h"

namespace lzb {

constexpr float kFloatMax = 1000000.0f; // common_lib/cminimax.h:9
constexpr float kFloatMin = -kFloatMax;
constexpr int kMaxGathers = 3;

// child record, 16 B: one LDG.128 per lane per level
//   x: visit_count (low 16 bits) | (expansion index + 1) << 16   (0 in the high half = not expanded)
//   y: value_sum   z: prior   w: reward (mz, gmz) or value_prefix (ez, sez)
// header record, 16 B: {n_children, to_play, is_reset, raw_value bits} of the node expanded as e
__device__ __forceinline__ int child_visits(int x) { return x & 0xffff; }
__device__ __forceinline__ int child_expansion(int x) { return (int)((unsigned)x >> 16) - 1; }
__device__ __forceinline__ int child_pack(int visits, int expansion) { return (visits & 0xffff) | ((expansion + 1) << 16); }

struct Pool {
    int variant, N, A, S, E, D, root_offset;
    int Ad, C;      // stochastic tree: decision actions / chance outcomes per node (A = max of the two); else Ad = A, C = 0
    int B;          // records per block = A + 1
    int4 *node;     // [N][E][B]  block e = {header, child 0 .. child A-1}
    int4 *root;     // [N]     {visit_count, value_sum bits, reward bits, unused}
    int *legal;     // [N][A]  root legal-action ids by position
    float2 *minmax; // [N]     {minimum, maximum} (cminimax.h:14-24)
    int4 *path_a;   // [N][S+1] per level q of the current search path: the chosen child's record as read by
                    //          select {record index within the tree, x, value_sum bits, reward bits} (node q+1)
    int2 *path_b;   // [N][S+1] per level q: {to_play, is_reset} of the node the choice was made AT (node q)
    int4 *res;      // [N]     {search_len, parent expansion (ix), last position, virtual_to_play}
    int *best;      // [N][E]  best_action position chosen at each expanded node (cnode.cpp:806)
    int *bidx;      // [N][E]  batch_index of each expanded node (row in the caller's model batch; == tree index unless a
                    //          ReZero caller compacts its batches, cnode.cpp:536); read and written by the reuse kernels only
    float *actions; // [N][E][A][D] sampled-action table (sez only)
    float *gumbel;  // [A]     gumbel noise shared by every root (gmz cnode.cpp:87-89)
    int *vtable;    // [S]     considered-visit sequence of the current Gumbel config
    int *scratch;   // [4]     {max(to_play), ...}
    float2 *pbc;    // [S+2]   per parent visit count T: {logf((T + c2 + 1) / c2) + c1, sqrtf(T)} (cucb_score, cnode.cpp:673-674)
    float *pbc2;    // experiment LZB_X_PBC2: [E+1][pbc2_w] pb_c of cucb_score per (parent visits - 1, child visits), or NULL
    int pbc2_w;
    float *cval;    // experiment LZB_X_CVAL: [N][E][B] CNode::value() of every child record as the backup last computed it, or NULL
    float2 *gstat;  // [N][E] {max, log-sum-exp} of csoftmax over a node's priors (Gumbel), NaN = not computed yet; or NULL
    const char **slots; // [kMaxGathers][E] base pointer of the [N, row] tensor that holds the rows gathered for
                        //          leaves whose parent was expanded as e (latent, LSTM c, LSTM h)

    __device__ __forceinline__ int4 *block(int i, int e) const { return node + ((size_t)i * E + e) * B; }
    __device__ __forceinline__ int4 *tree(int i) const { return node + (size_t)i * E * B; }
};

// Checked build (-DLZB_CHECKED; profiles/r02_sanitizer.md): every data-dependent index of the tree kernels is tested against
// its bound and the largest violated check's code is left in scratch[1] (lzb_check_word reads it).  compute-sanitizer is
// closed on the GPU pool this was developed on, and most of these indices stay INSIDE the forest's one allocation anyway,
// where memcheck would see nothing.  Compiles to nothing by default.
#ifdef LZB_CHECKED
#define LZB_CHK(pool, cond, code) do { if (!(cond)) atomicMax((pool).scratch + 1, (int)(code)); } while (0)
#else
#define LZB_CHK(pool, cond, code) do { } while (0)
#endif

// k-th of up to kMaxGathers pointers without dynamically indexing kernel-parameter arrays (which costs a stack copy)
__device__ __forceinline__ const char *pick3(const char *const (&v)[kMaxGathers], int k)
{
    return k == 0 ? v[0] : (k == 1 ? v[1] : v[2]);
}

} // namespace lzb
