/*
 * lzb200.h - C ABI of the B200-native batched MCTS tree (liblzb200.so).
 *
 * This is the boundary a LightZero maintainer binds instead of the Cython modules
 *   lzero/mcts/ctree/ctree_muzero/mz_tree.pyx            (Roots, MinMaxStatsList, ResultsWrapper,
 *   lzero/mcts/ctree/ctree_efficientzero/ez_tree.pyx      batch_traverse, batch_backpropagate)
 *   lzero/mcts/ctree/ctree_gumbel_muzero/gmz_tree.pyx
 *   lzero/mcts/ctree/ctree_sampled_efficientzero/ezs_tree.pyx
 *   lzero/mcts/ctree/ctree_stochastic_muzero/stochastic_mz_tree.pyx
 * (citations below are relative to /root/reference).  All pointers named *_dev are DEVICE pointers
 * on the current CUDA device; every call is asynchronous on `stream` (a cudaStream_t passed as
 * void*), takes plain pointers and sizes, never throws, never allocates device memory and keeps no
 * global state besides the thread-local error string.  Return value: 0 on success, negative
 * LZB_E_* on error (lzb_last_error() describes it).
 *
 * The forest (all N search trees of one batch) lives in ONE caller-owned device workspace laid out
 * as a structure-of-arrays node pool (see DESIGN.md "Data layout in HBM").  The caller (PyTorch)
 * owns that memory; the handle only holds offsets.
 *
 * Tie-break stream.  The reference breaks pUCT ties with libc rand() reseeded from the wall clock
 * (ctree_muzero/lib/cnode.cpp:592,770).  Here the stream is an explicit counter-based input:
 *     r(seed, sim, root, depth) = lzb_tiebreak_hash(seed, sim, root, depth)      (31 bits)
 *     pick = tie_set[r % n_ties]
 * LZB_RNG_PER_ROOT keys it by (sim, global root index, depth); LZB_RNG_BATCH uses r(seed, sim, 0, 0)
 * for every draw of one traverse, which is what the unmodified batched reference API can be pinned to.
 */
#ifndef LZB200_H
#define LZB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LZB_ABI_VERSION 6

/* tree flavours */
#define LZB_MZ 0  /* ctree_muzero */
#define LZB_EZ 1  /* ctree_efficientzero: value_prefix + is_reset */
#define LZB_GMZ 2 /* ctree_gumbel_muzero */
#define LZB_SEZ 3 /* ctree_sampled_efficientzero: K sampled actions per node */
#define LZB_SMZ 4 /* ctree_stochastic_muzero: decision and chance levels alternate */

#define LZB_RNG_BATCH 1
#define LZB_RNG_PER_ROOT 2

#define LZB_OK 0
#define LZB_E_ARG (-1)      /* bad argument (null pointer, size out of range) */
#define LZB_E_STATE (-2)    /* call order violated (e.g. more simulations than the pool was sized for) */
#define LZB_E_CUDA (-3)     /* a CUDA runtime call failed */
#define LZB_E_UNSUPPORTED (-4)

typedef struct lzb_forest lzb_forest_t; /* opaque, host memory */

/* pUCT / Gumbel constants of one search: mcts_ctree.py:181-193,246 and :1074. */
typedef struct {
    int32_t pb_c_base;                  /* 19652 */
    float pb_c_init;                    /* 1.25 */
    float discount;                     /* discount_factor */
    int32_t players;                    /* 0 = derive from max(to_play) as cnode.cpp:776-781 does (costs one
                                           device reduction); 1 or 2 = caller states it */
    int32_t num_simulations;            /* Gumbel: sequential-halving horizon */
    int32_t max_num_considered_actions; /* Gumbel */
    int32_t rng_mode;                   /* LZB_RNG_BATCH | LZB_RNG_PER_ROOT */
    uint64_t rng_seed;
    /* May be NULL.  Non-NULL: one 64-bit word in DEVICE memory read by the kernels instead of rng_seed, so that a
     * search captured in a CUDA graph breaks its ties with a fresh stream on every replay (the caller rewrites the
     * word between replays).  The reference reseeds on every traverse (cnode.cpp:770). */
    const uint64_t *rng_seed_dev;
} lzb_search_cfg_t;

const char *lzb_last_error(void);
int lzb_abi_version(void);
uint32_t lzb_tiebreak_hash(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth);
/* Source of the Stochastic-MuZero chance draws: 64 bits per (rng_seed, simulation, global root, depth of the chance
 * node); the draw uses u = (bits >> 11) / 2^53.  The reference seeds an mt19937 from std::random_device for every draw
 * (ctree_stochastic_muzero/lib/cnode.cpp:622-623). */
uint64_t lzb_chance_bits(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth);

/* Bytes of device workspace needed for N trees, children stride A (action space size; K for the
 * sampled tree), room for S simulations, D = action dimension (sampled tree), chance_space_size C (stochastic
 * tree, Roots(root_num, legal_actions_list, chance_space_size), stochastic_mz_tree.pyx:31-33), else 0.
 * Stochastic tree: nodes hold max(A, C) child records and EVERY logits / noise / visits row of this forest has
 * max(A, C) columns; max(A, C) <= 128. */
int lzb_forest_bytes(int variant, int num_roots, int action_space, int max_sims, int action_dim, size_t *bytes);

/* Replaces Roots.__cinit__ (mz_tree.pyx:30-32 -> CRoots::CRoots, cnode.cpp:301-317) and
 * MinMaxStatsList(num) (mz_tree.pyx:8-9).  legal_count_dev[N] / legal_actions_dev[N*A]: per-root
 * legal-action lists (row i holds legal_count[i] action ids, list order preserved); both NULL = every
 * action legal.  root_offset = global index of local root 0 when the batch is sharded over GPUs. */
int lzb_forest_create(lzb_forest_t **out, void *workspace_dev, size_t workspace_bytes, int variant, int num_roots,
                      int action_space, int max_sims, int action_dim, int root_offset, const int32_t *legal_count_dev,
                      const int32_t *legal_actions_dev, void *stream);
void lzb_forest_destroy(lzb_forest_t *f);

/* Roots.clear() followed by a fresh Roots(...) over the SAME workspace (cnode.cpp:360-366 + :301-317): a new
 * batch of roots with new legal-action lists, MinMax bounds cleared; nothing is reallocated, so one
 * workspace serves every search of a collector / reanalyze loop (and a captured CUDA graph can replay it). */
int lzb_forest_reset(lzb_forest_t *f, const int32_t *legal_count_dev, const int32_t *legal_actions_dev, void *stream);

/* MinMaxStatsList.set_delta (mz_tree.pyx:11-12); also clears min/max as a fresh list would be. */
int lzb_set_delta(lzb_forest_t *f, float value_delta_max, void *stream);

/* Roots.prepare / prepare_no_noise (mz_tree.pyx:34-39 -> cnode.cpp:321-358; gumbel adds values,
 * gmz_tree.pyx:36-40).  logits [N,A]; noise [N,A] indexed by position in the legal list (NULL = no
 * noise); reward [N] (value_prefix for EZ/SEZ); value [N] (Gumbel only, else NULL); to_play [N];
 * sampled_actions [N,K,D] (sampled tree only: the injected action table of each root). */
int lzb_prepare(lzb_forest_t *f, const float *logits_dev, const float *noise_dev, float noise_weight,
                const float *reward_dev, const float *value_dev, const int32_t *to_play_dev,
                const float *sampled_actions_dev, void *stream);

/* Where the rows gathered for a leaf live.  The reference keeps one array per expanded node generation
 * (latent_state_batch_in_search_path.append(...), mcts_ctree.py:247,305; reward_hidden_state_{c,h}_batch :727-728,
 * :817-818) and never copies it; neither does this library: up to three "slot tables" hold, per expansion index e,
 * the base pointer of the caller's [N, row] device tensor for e (0: latent, 1: LSTM c, 2: LSTM h). */
typedef struct {
    const void *ptr[3];
    int32_t n; /* how many of ptr[] are used */
} lzb_slots_t;

/* Slot pointers of expansion index `latent_index` (0 = the roots' own latents).  lzb_backprop / lzb_step set
 * the slots of the node they expand themselves (lzb_backprop_io_t.new_slots). */
int lzb_set_slots(lzb_forest_t *f, int latent_index, const lzb_slots_t *slots, void *stream);

/* One gather of leaf inputs: dst[i, :] = row i of the tensor registered for the leaf's parent expansion ix_i
 * (src_dev == NULL: slot table k, k = position of this gather in the array), or src[ix_i, i, :] of one dense
 * [E, N, row_bytes] pool (src_dev != NULL).  row_bytes: multiple of 4.  zero_if_reset: rows whose parent was
 * expanded with is_reset = 1 read as zeros (the LSTM-state reset of mcts_ctree.py:812-816).
 * Replaces the Python loop + np.asarray + H2D copy of mcts_ctree.py:283-289 / :766-775. */
typedef struct {
    const void *src_dev;
    void *dst_dev;
    int64_t row_bytes;
    int32_t zero_if_reset;
} lzb_gather_t;

/* ReZero search-with-reuse (arXiv 2404.16364): batch_traverse_with_reuse / batch_backpropagate_with_reuse
 * (mz_tree.pyx:84-107 -> cnode.cpp:502-549, 826-928; ez_tree.pyx:95-121 -> ez cnode.cpp:603-644, 965-1073).
 * A non-NULL `reuse` in lzb_traverse_io_t / lzb_backprop_io_t selects the reuse flavour of the call; MuZero and
 * EfficientZero trees with action_space <= 128 only (LZB_E_UNSUPPORTED otherwise).
 *   traverse: the root child on true_action[i] is scored by carm_score with reuse_value[i]; choosing it ends the
 *     walk.  A tree that ends on an already expanded child needs no inference: ix[i] = -1, iy[i] = i, nothing is
 *     gathered for it.  infer_count_dev (may be NULL): incremented by the number of trees that DO need inference.
 *     iy[i] of the other trees is the batch_index recorded when their leaf's parent was expanded.
 *   backprop: reward / value / logits stay [N]-row tensors indexed by TREE (rows of no-inference trees are
 *     ignored), so a device-resident caller runs the model on the full batch.  flags_dev[i] bit 0 = tree i is in
 *     the reference's no_inference_lst, bit 1 = in reuse_lst; NULL = derive both from the preceding traverse the
 *     way the reference driver does (mcts_ctree.py:376-384).  batch_index_dev[i] (may be NULL = i): the row tree i
 *     had in the caller's COMPACTED model batch (count_b, cnode.cpp:536), recorded in the new node so that later
 *     traverses report it as iy. */
typedef struct {
    const int32_t *true_action_dev;
    const float *reuse_value_dev;
    int32_t *infer_count_dev;
    const int32_t *flags_dev;
    const int32_t *batch_index_dev;
} lzb_reuse_t;

/* Stochastic MuZero (stochastic_mz_tree.pyx:87-91 -> cbatch_traverse, ctree_stochastic_muzero/lib/cnode.cpp:713-787;
 * cselect_child :596-660).  The root is a decision node; a decision node's children are chance nodes and vice
 * versa, so a node at depth d is a chance node iff d is odd.  Decision levels are pUCT with the tie-break stream
 * (its `depth` counts decision levels, as the reference's rand() calls do); at a chance level the outcome is drawn
 * from the children's priors exactly as libstdc++'s std::discrete_distribution does for u = lzb_chance_bits-derived
 * uniform (priors widened to double, normalised by their sum, partial sums, lower_bound).
 *   forced_outcome_dev [N, forced_stride] (may be NULL): entry [i, d] >= 0 replaces the draw at depth d of tree i
 *     (parity tests replay the outcomes the compiled reference drew).
 *   leaf_is_chance_dev [N] (may be NULL): results.leaf_node_is_chance - which model head the leaf needs
 *     (mcts_ctree_stochastic.py:150-176); equals search_len & 1.
 *   seed_dev (may be NULL): one 64-bit word in DEVICE memory used as the seed of lzb_chance_bits instead of
 *     cfg->rng_seed, so that a search captured in a CUDA graph draws fresh outcomes on every replay (the caller
 *     rewrites the word between replays); NULL: the chance draws use the tie-break stream's seed (cfg->rng_seed, or
 *     the word at cfg->rng_seed_dev when that is given).
 * lzb_backprop / lzb_step expand a chance leaf with the first C entries of its logits row and reward 0
 * (cnode.cpp:115-118), a decision leaf with the first A entries. */
typedef struct {
    const int32_t *forced_outcome_dev;
    int32_t forced_stride;
    int32_t *leaf_is_chance_dev;
    const uint64_t *seed_dev;
} lzb_chance_t;

/* batch_traverse (mz_tree.pyx:95-100 -> cbatch_traverse, cnode.cpp:755-824; ez :886-963; gumbel
 * :834-898; sampled :1110-1187) + the leaf gather.  Outputs [N]: ix (latent_state_index_in_search_path),
 * iy (latent_state_index_in_batch, local index), search_len, vtp (virtual_to_play_batchs),
 * last_action as int64 (what .long() produces, mcts_ctree.py:289); last_action_f [N,D] for the sampled
 * tree.  Any output pointer may be NULL.  gathers/num_gathers may be 0. */
typedef struct {
    const int32_t *to_play_dev;
    int32_t *ix_dev, *iy_dev;
    int64_t *last_action_dev;
    float *last_action_f_dev;
    int32_t *search_len_dev, *vtp_dev;
    const lzb_gather_t *gathers;
    int32_t num_gathers;
    const lzb_reuse_t *reuse; /* NULL = plain batch_traverse */
    const lzb_chance_t *chance; /* stochastic tree only; NULL = draw every outcome, no leaf_is_chance output */
} lzb_traverse_io_t;
int lzb_traverse(lzb_forest_t *f, const lzb_search_cfg_t *cfg, const lzb_traverse_io_t *io, void *stream);

/* batch_backpropagate / batch_back_propagate (mz_tree.pyx:74-82 -> cbatch_backpropagate,
 * cnode.cpp:480-500; ez :577-601; gumbel :633-652).  latent_index = simulation_index + 1.
 * reward/value [N], logits [N,A] straight from the model's device tensors; to_play [N] = the vtp
 * returned by traverse.  EZ/SEZ reset handling (mcts_ctree.py:809-816): if is_reset_dev != NULL it is
 * used as given; else if lstm_horizon_len > 0 is_reset = (search_len % lstm_horizon_len == 0) is
 * computed on device, written to is_reset_out_dev when non-NULL.  sampled_actions [N,K,D]: the new
 * nodes' action table (sampled tree).  new_slots: where the new nodes' latent / LSTM rows live. */
/* Stochastic tree, both model heads run on the whole batch (no compacted sub-batches, nothing leaves the device:
 * mcts_ctree_stochastic.py:165-251 calls the model once per leaf kind on Python-compacted rows).  The tree picks per
 * leaf: a leaf that is a CHANCE node takes reward / value / logits from this struct (the head called with
 * afterstate=False) and its new node's latent rows from latent_dev; a DECISION leaf takes the lzb_backprop_io_t's own
 * reward / value / logits and new_slots.ptr[0].  Later gathers read a node's rows from the tensor of its kind, so no
 * per-row select of the two heads' [N, row] outputs is ever materialised.
 *   logits_stride: floats per row of logits_dev here (>= chance_space_size);
 *   decision_logits_stride: floats per row of lzb_backprop_io_t::logits_dev (>= the decision action space; 0 = the
 *     forest's padded width max(A, C), the layout without this struct). */
typedef struct {
    const float *reward_dev, *value_dev, *logits_dev;
    int32_t logits_stride, decision_logits_stride;
    const void *latent_dev;
} lzb_chance_head_t;

typedef struct {
    int32_t latent_index;
    float discount;
    const float *reward_dev, *value_dev, *logits_dev;
    const int32_t *to_play_dev, *is_reset_dev;
    int32_t lstm_horizon_len;
    int32_t *is_reset_out_dev;
    const float *sampled_actions_dev;
    lzb_slots_t new_slots;
    const lzb_reuse_t *reuse; /* NULL = plain batch_backpropagate */
    /* [N] or NULL.  Non-NULL: only trees with active[i] != 0 are expanded and backed up, the others are left
     * untouched - the leaf_idx_list of the stochastic tree's cbatch_backpropagate (ctree_stochastic_muzero/lib/
     * cnode.cpp:559-587), whose driver backs the chance leaves and the decision leaves up in two calls
     * (mcts_ctree_stochastic.py:232-251).  Lane-group kernels only (every action space <= 128); not in lzb_step. */
    const int32_t *active_dev;
    const lzb_chance_head_t *chance_head; /* stochastic tree only; NULL = one [N] batch holds every leaf's outputs */
} lzb_backprop_io_t;
int lzb_backprop(lzb_forest_t *f, const lzb_backprop_io_t *io, void *stream);

/* One simulation step in ONE launch: lzb_backprop(bio) of simulation s followed by lzb_traverse(tio) of
 * simulation s + 1 (the reference runs the two back to back between model calls: mcts_ctree.py:318-321, then
 * :272-280 of the next iteration).  Results are identical to the two separate calls. */
int lzb_step(lzb_forest_t *f, const lzb_search_cfg_t *cfg, const lzb_backprop_io_t *bio, const lzb_traverse_io_t *tio,
             void *stream);

/* Row copy with optional reset (mcts_ctree.py:809-816): dst[i,:] = is_reset[i] ? 0 : src[i,:]; for callers that keep
 * a dense [E, N, row] pool instead of slot pointers. */
int lzb_store_hidden(const float *src_dev, float *dst_dev, const int32_t *is_reset_dev, int num_roots,
                     int64_t row_floats, void *stream);

/* get_distributions + get_values (cnode.cpp:387-417): visits [N,A] (first legal_count[i] entries of
 * row i are the child visit counts in legal-action order, rest 0), legal_count [N], values [N].
 * Gumbel extras (gmz cnode.cpp:506-541), each [N,A] or NULL: improved policies and completed child values. */
int lzb_root_outputs(lzb_forest_t *f, int32_t *visits_dev, int32_t *legal_count_dev, float *values_dev,
                     float discount, float *gumbel_policies_dev, float *gumbel_children_values_dev, void *stream);

/* get_trajectories (cnode.cpp:237-257,369-385): the greedy best_action chain from each root; row i of
 * actions [N, stride] holds len[i] action ids (entries beyond stride are dropped, len is the true length). */
int lzb_trajectories(lzb_forest_t *f, int32_t *actions_dev, int stride, int32_t *len_dev, void *stream);

/* Introspection for parity tests: raw pool arrays of tree i copied into caller DEVICE buffers laid out
 * as oracle/ref_harness.cpp:ref_dump_tree ([E,A] visit/value_sum/prior/reward/child_e, [E,4] hdr, [E] raw). */
int lzb_dump_tree(lzb_forest_t *f, int tree, int32_t *visit_dev, float *value_sum_dev, float *prior_dev,
                  float *reward_dev, int32_t *child_e_dev, int32_t *hdr_dev, float *raw_value_dev, void *stream);
int lzb_dump_paths(lzb_forest_t *f, int32_t *path_actions_dev, int path_stride, void *stream);
int lzb_minmax(lzb_forest_t *f, float *min_dev, float *max_dev, void *stream);
int lzb_root_stats(lzb_forest_t *f, int32_t *visit_dev, float *value_sum_dev, float *reward_dev, void *stream);

/* The K actions a Sampled-EfficientZero node draws when it is expanded (sez cnode.cpp:236-283 continuous: tanh of a
 * Gaussian per dimension, logits = {mu[D], sigma[D]}; :326-406 discrete: K of A without replacement, keys u^(1/p)).
 * out [N, K, D] (D = 1 and the action id as float for discrete) is what lzb_prepare / lzb_backprop take as
 * sampled_actions.  Counter-based: draws are a function of (seed, counter, root_offset + i, k) only; the reference
 * seeds from the wall clock, so only the distribution is comparable.  uniforms_dev (tests, else NULL): u in (0,1),
 * [N, K, D, 2] continuous (Box-Muller pairs) or [N, A] discrete. */
int lzb_sample_actions(const float *logits_dev, const float *uniforms_dev, float *out_dev, int num_roots, int num_sampled,
                       int action_dim, int action_space, int continuous, int root_offset, uint64_t seed, uint32_t counter,
                       void *stream);
/* Same, with the seed read from device memory when seed_dev != NULL (a captured CUDA graph of a Sampled search then
 * draws fresh numbers on every replay: the caller rewrites *seed_dev between replays). */
int lzb_sample_actions_seeded(const float *logits_dev, const float *uniforms_dev, float *out_dev, int num_roots,
                              int num_sampled, int action_dim, int action_space, int continuous, int root_offset,
                              uint64_t seed, const uint64_t *seed_dev, uint32_t counter, void *stream);

/* PARITY flavour of the sampler: the reference's own draws instead of the counter-based production stream.  Restates what
 * CNode::expand computes with libstdc++ (GCC 13): std::default_random_engine(seed) = minstd_rand0, per action dimension a
 * fresh std::normal_distribution<float>(mu, sigma) (Marsaglia polar over generate_canonical<float, 24>) and tanhf
 * (sez cnode.cpp:251-271), or std::uniform_real_distribution<double> keys pow(u, 1. / p) sorted descending (:336-397).
 * Root i is seeded with (unsigned)(seed_base + i) - the value the reference's i-th expand() of a batched call takes from
 * a wall clock that reads seed_base + i nanoseconds.  Continuous tables are bit-identical to the reference's (glibc
 * 2.39 logf / tanhf restated on the device); discrete tables are identical unless two keys lie within the device
 * pow()'s 2 ulp of each other.  One thread per root: a test / reproducibility mode, not the fast path. */
int lzb_sample_actions_libstdcxx(const float *logits_dev, float *out_dev, int num_roots, int num_sampled, int action_dim,
                                 int action_space, int continuous, uint64_t seed_base, void *stream);

/* ---- InverseScalarTransform of the model's categorical heads (lzero/policy/scaling_transform.py:97-128 with ensure_softmax
 * :36-62), applied by the reference driver to value and reward / value_prefix every simulation (mcts_ctree.py:302-303;
 * :797-798 EfficientZero): logits [N, 2 * support_size + 1] over the support {-s .. s} -> softmax (unless EVERY row of
 * the head already sums to 1 within torch.allclose's atol = rtol = 1e-5, then the rows are taken as probabilities) ->
 * expectation -> h^-1 with `epsilon` -> scalar_dev [N], the tensor lzb_backprop / lzb_step take as value_dev / reward_dev.
 * Up to two heads per launch; every logit is read once.  row_stride: floats between rows (>= 2 * support_size + 1).
 * workspace_dev: lzb_scalar_workspace_bytes(n_heads, num_rows) bytes, 16-byte aligned, ZEROED once before its first use
 * (the call leaves it ready for the next one); one workspace serves one stream at a time.
 * Not bit-pinned (torch fixes no summation order): agrees with the reference class to 1e-5 relative. */
typedef struct {
    const float *logits_dev;
    int64_t row_stride;
    float *scalar_dev;
} lzb_support_head_t;
int lzb_scalar_workspace_bytes(int n_heads, int num_rows, size_t *bytes);
int lzb_inverse_scalar_transform(const lzb_support_head_t *heads, int n_heads, int num_rows, int support_size, float epsilon,
                                 void *workspace_dev, void *stream);

/* lzb_inverse_scalar_transform without a caller workspace and WITHOUT ANY ALLOCATION.  This entry point has no counterpart
 * in the reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel, same results; its scratch (one ticket
 * word, one counter per head and num_rows floats per head) lives in a file-scope `static __device__` array of
 * 256 + 2 * LZB_STATIC_SCALAR_MAX_ROWS * 4 = 524,544 bytes that every GPU which loads the library holds until the process
 * exits.  The call itself performs no cudaMalloc / cudaFree and uses no allocating library routine; inputs and outputs are
 * still the caller's.  num_rows > LZB_STATIC_SCALAR_MAX_ROWS is rejected with LZB_E_ARG (use the workspace flavour above).
 * The array is resolved once per device (cudaGetSymbolAddress with that device current) on the first call made there.
 * EVERY call on a device shares the one array, which keeps the previous call's contents: the words that must start at zero
 * are cleared on `stream` ahead of the kernel, but two calls that OVERLAP on the GPU (different streams, or two host
 * threads launching at once) race on it and give undefined results - serialise them, or give each stream its own
 * workspace through lzb_inverse_scalar_transform. */
#define LZB_STATIC_SCALAR_MAX_ROWS 65536
int lzb_inverse_scalar_transform_static(const lzb_support_head_t *heads, int n_heads, int num_rows, int support_size,
                                        float epsilon, void *stream);

/* ---- Root-batch prologue / epilogue (what the callers of search() do per root in Python) ---- */

/* legal_actions = [a for a, x in enumerate(action_mask[j]) if x == 1] (lzero/policy/muzero.py:671;
 * lzero/mcts/buffer/game_buffer_muzero.py:552): mask [N, A] int8 -> legal_count [N], legal_actions [N, A] (ids in
 * ascending order, row padded with 0), the pair lzb_forest_create / lzb_forest_reset take. */
int lzb_legal_from_mask(const int8_t *mask_dev, int num_roots, int action_space, int32_t *legal_count_dev,
                        int32_t *legal_actions_dev, void *stream);

/* noises[j] = np.random.dirichlet([alpha] * n_legal_j).astype(np.float32) (lzero/policy/muzero.py:674-677) as the
 * [N, A] tensor lzb_prepare takes (row j: n_legal_j draws by legal position, zeros beyond).  legal_count_dev NULL =
 * every action legal.  Counter-based (seed, counter, root_offset + i, position): the reference draws from numpy's
 * global generator, so the distribution is what is comparable. */
int lzb_dirichlet_noise(float alpha, const int32_t *legal_count_dev, float *noise_dev, int num_roots, int action_space,
                        int root_offset, uint64_t seed, uint32_t counter, void *stream);

/* select_action(visit_counts, temperature, deterministic) for every root (lzero/policy/utils.py:515-539) followed by
 * action = np.where(action_mask == 1)[0][action_pos] (lzero/policy/muzero.py:709).  visits / legal_count /
 * legal_actions as lzb_root_outputs and lzb_legal_from_mask produce them (legal_actions NULL = identity).
 * deterministic: first argmax of the visit counts; else one draw from visits^(1/T) normalised, by inverse CDF exactly
 * as np.random.choice does it (uniforms_dev [N] in [0,1) injects the draws; NULL = counter-based generator).
 * entropy [N] f64: base-2 entropy of the distribution.  Outputs may be NULL. */
typedef struct {
    const int32_t *visits_dev, *legal_count_dev, *legal_actions_dev;
    const double *uniforms_dev;
    int32_t *action_pos_dev, *action_dev;
    double *entropy_dev;
    double temperature;
    int32_t num_roots, action_space, deterministic, root_offset;
    uint64_t seed;
    uint32_t counter;
} lzb_select_action_t;
int lzb_select_action(const lzb_select_action_t *args, void *stream);

/* Reanalyze targets (lzero/mcts/buffer/game_buffer_muzero.py:617-656): policy [N, A] f32 = visits / sum(visits)
 * (float64 division, then cast) at the legal action ids (legal_actions NULL = fixed action space: position ==
 * action), all zeros where policy_mask[i] == 0 (NULL = every row valid). */
int lzb_target_policies(const int32_t *visits_dev, const int32_t *legal_count_dev, const int32_t *legal_actions_dev,
                        const int32_t *policy_mask_dev, int num_roots, int action_space, float *policy_dev, void *stream);

/* Reanalyze value / reward targets (lzero/mcts/buffer/game_buffer_muzero.py:463-513, the tail of
 * _compute_target_reward_value): for segment b and unroll step u, with t = b * (U + 1) + u and current = pos[b] + u,
 *   v = root_values[t] * discount ** td_steps[t] (negated for odd td_steps when board_games) * value_mask[t]
 *       + sum_i (+|-) rewards[b][current + i] * discount ** i   over the Python slice [current : current + td_steps[t]]
 *   target_values[t] = v and target_rewards[t] = rewards[b][current] if current < segment_len[b], else 0 and 0.
 * float64 arithmetic in the reference's order.  pow_table_dev[k] = discount ** k for k < pow_len (computed by the caller
 * on the HOST, so that it is the host libm's pow like Python's `**`); pow_len must exceed every td_steps entry.
 * to_play_dev [B, to_play_stride] is read only when board_games != 0 (:494-499, indexed exactly as the reference does).
 * lstm_horizon_len > 0 selects EfficientZero's flavour (game_buffer_efficientzero.py:261-316; ABI v6): target_rewards holds
 * the VALUE PREFIX - the running float64 sum of rewards[b][current] over the steps inside the segment since the last horizon
 * reset, kept (not zeroed) past the segment's end - where a reset happens at every step whose index in the WHOLE transition
 * batch, t, is a multiple of lstm_horizon_len (the reference's horizon_id is not restarted per segment, :278), and a reset
 * also moves the base index of the board-games sign rule to that step (:298-300). */
typedef struct {
    const float *root_values_dev, *value_mask_dev, *rewards_dev;
    const int32_t *td_steps_dev, *reward_len_dev, *segment_len_dev, *pos_dev, *to_play_dev;
    const double *pow_table_dev;
    double *target_values_dev, *target_rewards_dev;
    int32_t num_segments, num_unroll_steps, reward_stride, to_play_stride, pow_len, board_games;
    int32_t lstm_horizon_len;
} lzb_value_targets_t;
int lzb_value_targets(const lzb_value_targets_t *args, void *stream);

/* ---- End-of-search exchange over NVLink / NVSwitch peer memory ----
 * The one place a sharded search crosses GPUs: every rank's root outputs (visit counts [n, A] i32, values [n] f32,
 * Gumbel policies ...) -> the full batch on every rank.  One kernel per rank packs its pieces row-wise, stores them
 * straight into every peer's receive buffer and runs a flag barrier over the same mapping (no NCCL call, no staging
 * copy).  Buffers are cudaMalloc'ed by the library so that they can be shared through CUDA IPC handles:
 *   lzb_peer_bytes   size of one rank's receive buffer for `world` ranks x rows_per_rank rows x row_words 32-bit words
 *   lzb_peer_alloc / lzb_peer_free          the buffer (zeroed)
 *   lzb_peer_export / lzb_peer_import / lzb_peer_unimport   64-byte IPC handle out; a peer's handle in (the caller
 *                    moves the handles between processes, e.g. with torch.distributed.all_gather_object)
 *   lzb_peer_allgather  bufs_host[world]: every rank's buffer as mapped in THIS process (own buffer at [rank]);
 *                    pieces_dev[n_pieces]: local [rows_per_rank, piece_words[k]] 32-bit tensors, concatenated per row;
 *                    epoch: 1, 2, 3 ... the same sequence on every rank; *result_dev: [world * rows_per_rank,
 *                    row_words] in the caller's own buffer, valid until the next-but-one exchange.  A peer that does
 *                    not arrive within timeout_ms (<= 0: 30 s) is reported by lzb_peer_error instead of hanging the
 *                    GPU, and its rows of the result are filled with 0x80000000 (INT_MIN as a count) - never left
 *                    holding an earlier epoch's data.
 *   lzb_peer_error   0, or 1 + the rank that timed out (synchronises the stream); a reported error is cleared. */
int lzb_peer_bytes(int world, int64_t rows_per_rank, int row_words, size_t *bytes);
int lzb_peer_alloc(size_t bytes, void **buf_dev);
int lzb_peer_free(void *buf_dev);
int lzb_peer_export(void *buf_dev, void *handle_64_bytes);
int lzb_peer_import(const void *handle_64_bytes, void **buf_dev);
int lzb_peer_unimport(void *buf_dev);
int lzb_peer_allgather(void *const *bufs_host, int rank, int world, int64_t rows_per_rank, const void *const *pieces_dev,
                       const int32_t *piece_words, int n_pieces, uint32_t epoch, int64_t timeout_ms, void **result_dev,
                       void *stream);
int lzb_peer_error(void *own_buf_dev, int world, int64_t rows_per_rank, int row_words, int *error_host, void *stream);

/* Gumbel constants exposed for tests (gmz_tree.pyx:134-152): host buffers. */
int lzb_gumbel_table(int max_num_considered_actions, int num_simulations, int32_t *out_host);
int lzb_gumbel_noise(float scale, int shape, float *out_host);

/* Single-node Gumbel MuZero helpers: the free functions gmz_tree.pyx:97-149 exposes to the reference's unit tests
 * (softmax -> csoftmax, pcompute_mixed_value, prescale_qvalues, pqtransform_completed_by_mix_value, pscore_considered,
 * select_root_child, select_interior_child; gmz cnode.cpp:701-790,904-1040,1097-1132) over vectors of n floats / ints in
 * DEVICE memory, evaluated by one thread with the tree's own libm.  Vectors and scalars per op:
 *   LZB_GH_SOFTMAX            a = x                                                            -> out[n]
 *   LZB_GH_MIXED_VALUE        s0 = raw value, a = q values, v = child visits, b = child priors -> out[1]
 *   LZB_GH_RESCALE            a = values, s0 = epsilon                                         -> out[n]
 *   LZB_GH_QTRANSFORM         a = the node's get_q(discount), s0 = its raw value, v, b = child priors, s1 = maxvisit_init,
 *                             s2 = value_scale, i0 = rescale_values, s3 = epsilon              -> out[n]
 *   LZB_GH_SCORE_CONSIDERED   i0 = considered visit, a = gumbel, b = logits, c = normalised q, v = visits -> out[n]
 *   LZB_GH_SELECT_ROOT        a = q, s0 = raw value, v, b = child priors, c = the node's gumbel noise, i0 = the considered
 *                             visit count of this simulation (lzb_gumbel_table row min(m, S), column sum(v)) -> out_i[1]: POSITION
 *   LZB_GH_SELECT_INTERIOR    a = q, s0 = raw value, v, b = child priors                       -> out_i[1]: POSITION
 * tmp_dev: 3 n floats of scratch. */
enum { LZB_GH_SOFTMAX = 0, LZB_GH_MIXED_VALUE = 1, LZB_GH_RESCALE = 2, LZB_GH_QTRANSFORM = 3, LZB_GH_SCORE_CONSIDERED = 4,
       LZB_GH_SELECT_ROOT = 5, LZB_GH_SELECT_INTERIOR = 6 };
typedef struct {
    int op, n;
    const float *a_dev, *b_dev, *c_dev;
    const int32_t *v_dev;
    float s0, s1, s2, s3;
    int i0, i1;
    float *out_dev;
    int32_t *out_i_dev;
    float *tmp_dev;
} lzb_gumbel_helper_t;
int lzb_gumbel_helper(const lzb_gumbel_helper_t *args, void *stream);

/* Checked build (-DLZB_CHECKED, profiles/r02_sanitizer.md): the largest code of a violated index check since the forest
 * was created / reset (0 = none; always 0 in the shipped build), and whether this library is a checked build.  No
 * counterpart in the reference.  Synchronises the stream. */
int lzb_check_word(lzb_forest_t *f, int32_t *word_host, int32_t *checked_build_host, void *stream);

/* Device-side libm used by the tree, exposed for the parity tests: out[i] = expf/logf(x[i]). */
int lzb_test_expf(const float *x_dev, float *out_dev, int64_t n, void *stream);
int lzb_test_logf(const float *x_dev, float *out_dev, int64_t n, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* LZB200_H */
