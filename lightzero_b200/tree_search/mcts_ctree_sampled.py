"""SampledEfficientZeroMCTSCtree with the reference's signature (lzero/mcts/tree_search/mcts_ctree_sampled.py:15-232),
device resident like the drivers in mcts_ctree.py.  Each expanded node's K sampled actions are drawn on the device
(tree_search/sampling.py) and handed to the tree kernels as a table."""
from .mcts_ctree import EfficientZeroMCTSCtree, _get
from .sampling import sample_actions


def _ezs_module():
    from ..ctree.ctree_sampled_efficientzero import ezs_tree
    return ezs_tree


class SampledEfficientZeroMCTSCtree(EfficientZeroMCTSCtree):
    """Same loop as EfficientZero (value_prefix, LSTM reset); differences (mcts_ctree_sampled.py:150-178,224-231):
    last_actions are [N, D] floats for continuous action spaces, and the new leaves get a sampled-action table."""
    tree = _ezs_module()
    _sampled = True

    @classmethod
    def roots(cls, active_collect_env_num, legal_actions, action_space_size, num_of_sampled_actions,
              continuous_action_space):
        """mcts_ctree_sampled.py:76-94."""
        return cls.tree.Roots(active_collect_env_num, legal_actions, action_space_size, num_of_sampled_actions,
                              continuous_action_space)

    def _continuous(self, roots):
        return bool(_get(_get(self._cfg, "model", {}) or {}, "continuous_action_space", roots.continuous_action_space))

    def _leaf_action(self, roots, o):
        return o["last_action_f"] if self._continuous(roots) else o["last_action"]

    def _expand_inputs(self, roots, net):
        # draws are keyed by (search seed, expansion index, global root, k): csrc/lzb_sample.cuh
        roots._expansions = getattr(roots, "_expansions", 0) + 1
        return dict(sampled_actions=sample_actions(net.policy_logits, roots.env_action_space_size,
                                                   roots.num_of_sampled_actions, self._continuous(roots),
                                                   seed=getattr(roots, "_seed_dev", None) if getattr(roots, "_seed_dev", None) is not None
                                                   else getattr(roots, "sample_seed", 0), counter=roots._expansions,
                                                   root_offset=roots.root_offset))
