// End-of-search exchange over NVLink / NVSwitch peer memory: the root outputs of every rank's shard
// (visit counts, values, Gumbel policies, ...) -> the full batch on every rank, as ONE kernel per rank: the rows are
// packed and stored straight into every peer's receive buffer (P2P stores), then a flag barrier over the same
// mapping tells each rank that all shards have landed.  Replaces torch.cat + ncclAllGather for this one small,
// latency-bound exchange (0.6 MB per rank at the headline shape); nothing else in a search crosses GPUs.
//
// Receive buffer of a rank (peer-mapped, allocated by lzb_peer_alloc):
//   [2 epochs][world * rows][row_words] u32   data, double-buffered by epoch parity
//   [world] u32                               flags: flags[r] = last epoch whose shard r has fully arrived
//   [2] u32                                   {block counter, error}
// Reuse is safe with two buffers: rank r can only start epoch k+2 after the barrier of epoch k+1, which needs
// every peer's epoch-(k+1) kernel to be running, i.e. after that peer's stream has consumed the epoch-k result.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace lzb {

constexpr int kMaxPeers = 16;
constexpr int kMaxPieces = 4;

struct PeerArgs {
    uint32_t *data[kMaxPeers];   // this epoch's data area of every rank (self included), as mapped in this process
    uint32_t *flags[kMaxPeers];  // flag array of every rank
    uint32_t *ctl;               // own {block counter, error}
    const uint32_t *piece[kMaxPieces]; // local [rows, words[k]] tensors to pack side by side
    int words[kMaxPieces];
    int n_pieces, row_words, rows, rank, world;
    uint32_t epoch;
    long long timeout_ns;
};

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ long long global_ns()
{
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__global__ void k_peer_allgather(PeerArgs a)
{
    // ---- pack + scatter: word (row, col) of the local shard goes to row rank*rows + row of every rank's buffer
    const long long total = (long long)a.rows * a.row_words;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int row = (int)(t / a.row_words);
        int col = (int)(t - (long long)row * a.row_words);
        uint32_t v = 0;
        int k = 0, c = col;
#pragma unroll
        for (int q = 0; q < kMaxPieces; ++q)
            if (q < a.n_pieces && k == q && c >= a.words[q]) { c -= a.words[q]; k = q + 1; }
#pragma unroll
        for (int q = 0; q < kMaxPieces; ++q)
            if (q == k) v = a.piece[q][(long long)row * a.words[q] + c];
        const long long at = ((long long)a.rank * a.rows + row) * a.row_words + col;
#pragma unroll 1
        for (int w = 0; w < a.world; ++w) a.data[w][at] = v;
    }
    // ---- grid completion, then the flag barrier by the last block
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(&a.ctl[0], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    if (threadIdx.x == 0) a.ctl[0] = 0u; // ready for the next launch
    if (threadIdx.x < a.world) {
        const int w = threadIdx.x;
        st_release_sys(a.flags[w] + a.rank, a.epoch); // my shard of this epoch is complete in rank w's buffer
        const long long t0 = global_ns();
        while ((int32_t)(ld_acquire_sys(a.flags[a.rank] + w) - a.epoch) < 0) { // rank w's shard has not landed here yet
            if (global_ns() - t0 > a.timeout_ns) {
                // a peer never arrived: report it instead of hanging the GPU, and POISON its rows of this epoch's result
                // (0x80000000 = INT_MIN as a visit count, -0.0f as a value) so that a caller who does not read the error
                // word cannot mistake a previous epoch's rows for this search's
                a.ctl[1] = 1u + (uint32_t)w;
                uint32_t *mine = a.data[a.rank] + (long long)w * a.rows * a.row_words;
                for (long long t = 0; t < (long long)a.rows * a.row_words; ++t) mine[t] = 0x80000000u;
                break;
            }
            __nanosleep(64);
        }
    }
}

} // namespace lzb
