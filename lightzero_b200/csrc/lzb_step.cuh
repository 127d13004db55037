// Fused simulation step (K4 + K2 + K3 in one launch): expand + backup of simulation s followed by the selection
// (and leaf gather) of simulation s + 1 for the same tree by the same warp.  Between two model calls the
// reference runs cbatch_backpropagate and then cbatch_traverse back to back (mcts_ctree.py:318-321 -> :272-280 of
// the next loop iteration); nothing outside the tree happens in between, so the two are one kernel here: one
// launch per simulation, and the path the backup just updated is still in L1/L2 when selection re-reads it.
#pragma once
#include "lzb_expand.cuh"
#include "lzb_select.cuh"

namespace lzb {

template <int VAR>
__global__ void k_step(Pool p, BackupArgs b, SelectArgs a)
{
    extern __shared__ int4 smem4[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int i = blockIdx.x * wpb + wib;
    // later launches find the new node's slot pointers in the table; this launch takes them from `a.new_slot`
    if (blockIdx.x == 0 && threadIdx.x < b.n_slots) p.slots[(size_t)threadIdx.x * p.E + b.latent_index] = pick3(b.new_slot, threadIdx.x);
    if (i >= p.N) return;
    const int A = p.A;
    int4 *s_child = smem4 + (size_t)wib * A;
    if (VAR == LZB_GMZ) {
        float *fbase = reinterpret_cast<float *>(smem4 + (size_t)wpb * A) + (size_t)wib * 3 * A;
        expand_backup_tree<VAR>(p, b, i, lane, fbase);
        __syncwarp(); // the warp's own global writes (records, root, MinMax) are ordered before its reads below
        select_gumbel_tree(p, a, i, lane, s_child, fbase, fbase + A, fbase + 2 * A);
    } else {
        float *s_score = reinterpret_cast<float *>(smem4 + (size_t)wpb * A) + (size_t)wib * A;
        expand_backup_tree<VAR>(p, b, i, lane, s_score);
        __syncwarp();
        select_puct_tree<VAR>(p, a, i, lane, s_child, s_score);
    }
}

} // namespace lzb
