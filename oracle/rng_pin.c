/*
 * TEST INFRASTRUCTURE ONLY (oracle/): pins the libc tie-break stream of the compiled reference.
 *
 * The reference tree picks among tied pUCT scores with `rand() % n_ties`
 * (/root/reference/lzero/mcts/ctree/ctree_muzero/lib/cnode.cpp:592) and reseeds with
 * `srand(tv_usec)` on every cbatch_traverse (common_lib/utils.cpp:22-25, cnode.cpp:770), so its
 * outputs are not reproducible as shipped.  `rand`, `srand` are undefined dynamic symbols of the
 * reference translation unit; this file defines them and is linked into oracle/_ref/*.so with
 * -Wl,-Bsymbolic so the reference code binds to these definitions.  No reference source is edited.
 *
 * Stream contract (identical to the one the CUDA kernels implement, see include/lzb200.h):
 *     r(seed, sim, root, depth) = lzb_tiebreak_hash(seed, sim, root, depth)   (31-bit, like rand())
 *   mode 0: pass-through to libc random()                       (timing runs)
 *   mode 1: "batch"    - every rand() between two srand() calls returns r(seed, sim, 0, 0)
 *   mode 2: "per-root" - the k-th rand() after srand() returns r(seed, sim, root, k)
 * (sim, root) are set by the harness before each cbatch_traverse call.
 */
#include <stdint.h>
#include <stdlib.h>

static int g_mode = 0;
static uint64_t g_seed = 0;
static uint32_t g_sim = 0, g_root = 0, g_depth = 0;

static inline uint64_t mix64(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

uint32_t ref_tiebreak_hash(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth)
{
    uint64_t x = mix64(seed + 0x9E3779B97F4A7C15ULL * (uint64_t)(sim + 1u));
    x = mix64(x ^ (((uint64_t)root << 32) | (uint64_t)depth));
    return (uint32_t)(x >> 33); /* 31 bits, non-negative as an int */
}

void ref_rng_mode(int mode, uint64_t seed) { g_mode = mode; g_seed = seed; }
void ref_rng_set(uint32_t sim, uint32_t root) { g_sim = sim; g_root = root; g_depth = 0; }

int rand(void)
{
    if (g_mode == 0) return (int)random();
    if (g_mode == 1) return (int)ref_tiebreak_hash(g_seed, g_sim, 0u, 0u);
    return (int)ref_tiebreak_hash(g_seed, g_sim, g_root, g_depth++);
}

void srand(unsigned int s)
{
    if (g_mode == 0) { srandom(s); return; }
    g_depth = 0;
}
