/*
 * TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatement of LightZero's batched ctree MCTS
 * (/root/reference/lzero/mcts/ctree/ctree_{muzero,efficientzero,gumbel_muzero,sampled_efficientzero,
 * stochastic_muzero}).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library; nothing under lightzero_b200/ does.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py drives this restatement and the compiled,
 * unmodified reference (oracle/_ref, built from /root/reference by oracle/build.py) on the same seeded
 * inputs with the tie-break stream pinned by oracle/rng_pin.c and requires bit-identical actions,
 * search paths, visit counts and raw fp32 bits of values / priors / value sums; tests/golden/ holds
 * fixtures generated from the compiled reference for machines where /root/reference is absent.
 * The Sampled-EfficientZero action SAMPLER is the one piece that is parity-unpinned by design
 * (wall-clock seeded std::default_random_engine, ctree_sampled_efficientzero/lib/cnode.cpp:251,336):
 * the compiled reference samples as it likes, oracle/ref_harness_sampled.cpp reads the K actions of every
 * expanded node back out of its tree, and this restatement (and the CUDA path) are handed exactly those
 * tables, so the Sampled tree's selection / backup logic IS pinned bit for bit; the fixtures carry the tables.
 * The Stochastic-MuZero chance DRAW is unpinned in the same way (an mt19937 seeded from std::random_device per
 * draw, ctree_stochastic_muzero/lib/cnode.cpp:622-623): oracle/ref_harness_stochastic.cpp reports the outcome the
 * reference took at every chance level, this restatement and the CUDA path are handed those outcomes, and the draw's
 * arithmetic (orc_discrete_draw) is pinned against libstdc++'s discrete_distribution under a scripted engine.
 */
#ifndef LZB_ORACLE_H
#define LZB_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_MZ = 0, ORC_EZ = 1, ORC_GMZ = 2, ORC_SEZ = 3, ORC_SMZ = 4 };

float orc_expf(float x);
float orc_logf(float x);
uint32_t orc_tiebreak_hash(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth);

typedef struct OrcForest OrcForest;

OrcForest *orc_new(int variant, int n, int A, const int *legal_flat, const int *legal_off, int K, int D,
                   int continuous);
void orc_free(OrcForest *f);
void orc_set_delta(OrcForest *f, float value_delta_max);
void orc_set_rng(OrcForest *f, int mode, uint64_t seed, uint32_t root_offset);
void orc_prepare(OrcForest *f, int with_noise, float noise_w, const float *noise_flat, const int *noise_off,
                 const float *rewards, const float *values, const float *logits, const int *to_play,
                 const float *sampled_actions);
void orc_traverse(OrcForest *f, int pb_c_base, float pb_c_init, float discount, int num_simulations,
                  int max_num_considered_actions, const int *to_play, int *ix, int *iy, int *last_action,
                  float *last_action_f, int *search_len, int *vtp, int *path_actions, int path_stride);
void orc_backprop(OrcForest *f, int latent_index, float discount, const float *rewards, const float *values,
                  const float *logits, const int *is_reset, const int *to_play, const float *sampled_actions);
void orc_traverse_reuse(OrcForest *f, int pb_c_base, float pb_c_init, float discount, const int *to_play,
                        const int *true_action, const float *reuse_value, int *ix, int *iy, int *last_action,
                        int *search_len, int *vtp);
void orc_backprop_reuse(OrcForest *f, int latent_index, float discount, const float *rewards, const float *values,
                        const float *logits, const int *is_reset, const int *to_play, const int *no_inference_lst,
                        const int *reuse_lst, const float *reuse_value);
/* Stochastic MuZero (ORC_SMZ: orc_new's K = chance_space_size C; every logits row has stride max(A, C)) */
uint64_t orc_chance_bits(uint64_t seed, uint32_t sim, uint32_t root, uint32_t depth);
int orc_discrete_draw(const double *probs, int n, double u);
void orc_traverse_chance(OrcForest *f, int pb_c_base, float pb_c_init, float discount, const int *to_play,
                         const int *forced, int forced_stride, int *ix, int *iy, int *last_action, int *search_len,
                         int *vtp, int *leaf_is_chance, int *path_actions, int path_stride);
void orc_distributions(OrcForest *f, int *dist, int *len);
void orc_values(OrcForest *f, float *out);
void orc_minmax(OrcForest *f, float *mn, float *mx);
void orc_gumbel_outputs(OrcForest *f, float discount, float *policies, float *children_values);
void orc_gumbel_table(int max_num_considered_actions, int num_simulations, int *out);
void orc_gumbel_noise(float scale, int shape, float *out);
void orc_softmax(float *x, int n);
void orc_dump_tree(OrcForest *f, int i, int E, int *visit, float *value_sum, float *prior, float *reward,
                   int *child_e, int *hdr, float *raw_value);
void orc_root_stats(OrcForest *f, int *visit, float *value_sum, float *reward);
void orc_sampled_actions(OrcForest *f, float *out /* [n, K, D] */);

#ifdef __cplusplus
}
#endif
#endif
