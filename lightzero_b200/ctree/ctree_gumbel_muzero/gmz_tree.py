"""Drop-in for lzero/mcts/ctree/ctree_gumbel_muzero/gmz_tree.pyx over the CUDA forest (variant LZB_GMZ:
sequential-halving root selection, completed-Q interior selection; gmz cnode.cpp:701-790,935-1132)."""
import ctypes as C

import numpy as np

from ... import _lib
from .._tree_api import MinMaxStatsList, ResultsWrapper, RootsBase, backprop_tensors, set_tie_break  # noqa: F401
from .._tree_api import _search_cfg, traverse_tensors


class Roots(RootsBase):
    """gmz_tree.pyx:28-68: prepare / prepare_no_noise take the root raw values before the logits."""
    variant = "gmz"

    def prepare(self, root_noise_weight, noises, value_prefix_pool, value_pool, policy_logits_pool, to_play_batch):
        self._store_prepare(root_noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch,
                            values=value_pool)

    def prepare_no_noise(self, value_prefix_pool, value_pool, policy_logits_pool, to_play_batch):
        self._store_prepare(0.0, None, value_prefix_pool, policy_logits_pool, to_play_batch, values=value_pool)

    def _gumbel_outputs(self, discount, action_space_size):
        if int(action_space_size) != self.action_space_size:
            raise ValueError("action_space_size %d differs from the prepared logits width %d" % (
                action_space_size, self.action_space_size))
        cached = getattr(self, "_cached_outputs", None)
        if cached is not None and "policies" in cached:
            return cached
        return self._materialize().root_outputs(discount=float(discount), gumbel=True)

    def get_children_values(self, discount, action_space_size):
        """[N][A] completed Q-values, -inf at illegal actions (gmz cnode.cpp:309-339,506-523)."""
        return self._gumbel_outputs(discount, action_space_size)["children_values"].tolist()

    def get_policies(self, discount, action_space_size):
        """[N][A] improved policy softmax(prior + completedQ) over the action space (gmz cnode.cpp:356-385,525-541)."""
        return self._gumbel_outputs(discount, action_space_size)["policies"].tolist()

    def get_policies_tensor(self, discount, action_space_size):
        o = self._gumbel_outputs(discount, action_space_size)
        return o["policies"], o["children_values"]


def batch_traverse(roots, num_simulations, max_num_considered_actions, discount_factor, results, virtual_to_play_batch):
    """gmz_tree.pyx:91-95 -> cbatch_traverse (gmz cnode.cpp:834-898): no MinMax, no pUCT constants."""
    f = roots._materialize(num_simulations)
    cfg = _search_cfg(f, 0, 0.0, discount_factor, virtual_to_play_batch, num_simulations=num_simulations,
                      max_considered=max_num_considered_actions, roots=roots)
    o = traverse_tensors(roots, cfg, virtual_to_play_batch, results)
    return o["ix"].tolist(), o["iy"].tolist(), o["last_action"].tolist(), o["vtp"].tolist()


def batch_back_propagate(current_latent_state_index, discount_factor, value_prefixs, values, policies, min_max_stats_lst,
                         results, to_play_batch):
    """gmz_tree.pyx:81-88 -> cbatch_back_propagate (gmz cnode.cpp:633-652)."""
    if results.out is None:
        raise RuntimeError("batch_back_propagate needs the ResultsWrapper of the preceding batch_traverse")
    backprop_tensors(results, current_latent_state_index, discount_factor, value_prefixs, values, policies, to_play_batch)


# ---- test helpers of gmz_tree.pyx:134-152 that are pure functions of the configuration
def pget_table_of_considered_visits(max_num_considered_actions, num_simulations):
    out = np.zeros((max_num_considered_actions + 1, num_simulations), np.int32)
    _lib.check(_lib.lib().lzb_gumbel_table(int(max_num_considered_actions), int(num_simulations),
                                            out.ctypes.data_as(C.c_void_p)))
    return out.tolist()


def pget_sequence_of_considered_visits(max_num_considered_actions, num_simulations):
    return pget_table_of_considered_visits(max_num_considered_actions, num_simulations)[max_num_considered_actions]


def pgenerate_gumbel(gumbel_scale, gumbel_rng, shape):
    """generate_gumbel (gmz cnode.cpp:1134-1152); the reference seeds mt19937 with (unsigned)gumbel_rng == 0."""
    out = np.zeros(int(shape), np.float32)
    _lib.check(_lib.lib().lzb_gumbel_noise(float(gumbel_scale), int(shape), out.ctypes.data_as(C.c_void_p)))
    return out.tolist()
