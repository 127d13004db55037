"""Drop-in for lzero/mcts/ctree/ctree_sampled_efficientzero/ezs_tree.pyx over the CUDA forest (variant LZB_SEZ).

Each node has K sampled actions (sez cnode.cpp:194-452).  The reference draws them inside expand() from
wall-clock-seeded C++ RNGs; here sampling is a seam: the action table [N, K, D] of every expanded node is an
INPUT of prepare / batch_backpropagate (drawn on the device by tree_search/mcts_ctree_sampled.py), which is
also how the parity tests inject identical tables into the oracle.  Selection uses the uniform-prior pUCT branch
the reference actually executes (sez cnode.cpp:1054-1079)."""
import torch

from ...forest import _dev
from .._tree_api import MinMaxStatsList, NodeBase, ResultsWrapper, RootsBase, backprop_tensors, set_tie_break  # noqa: F401
from .._tree_api import _search_cfg, traverse_tensors


class Roots(RootsBase):
    """ezs_tree.pyx:38-84: Roots(root_num, legal_actions_list, action_space_size, num_of_sampled_actions, continuous)."""
    variant = "sez"

    def __init__(self, root_num, legal_actions_list, action_space_size, num_of_sampled_actions,
                 continuous_action_space, device=None):
        super().__init__(root_num, None, device=device)
        self.sampled_legal = legal_actions_list
        self.env_action_space_size = int(action_space_size)
        self.num_of_sampled_actions = int(num_of_sampled_actions)
        self.continuous_action_space = bool(continuous_action_space)

    def _forest_dims(self):
        return self.num_of_sampled_actions, self.action_dim

    def _store_prepare(self, noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch, values=None,
                       sampled_actions=None):
        if sampled_actions is None:
            from ...tree_search.sampling import new_sample_seed, sample_actions
            self.sample_seed = new_sample_seed(self.tie_break_base)
            sampled_actions = sample_actions(_dev(policy_logits_pool, torch.float32, self._device),
                                             self.env_action_space_size, self.num_of_sampled_actions,
                                             self.continuous_action_space, seed=self.sample_seed, counter=0,
                                             root_offset=self.root_offset)
        sa = _dev(sampled_actions, torch.float32, self._device)
        if sa.dim() != 3 or sa.shape[0] != self.root_num or sa.shape[1] != self.num_of_sampled_actions:
            raise ValueError("sampled_actions must be [root_num, num_of_sampled_actions, action_dim]")
        self.action_dim = sa.shape[2]
        d = self._device
        self._prep = dict(logits=None, noise=None, noise_weight=float(noise_weight),
                          reward=_dev(value_prefix_pool, torch.float32, d).reshape(-1),
                          to_play=_dev(to_play_batch, torch.int32, d).reshape(-1), value=None, sampled_actions=sa)
        self.action_space_size = self.num_of_sampled_actions
        self._cleared = False
        self._cached_outputs = None
        if self.forest is not None:
            self._apply_prepare()

    def prepare(self, root_noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch, sampled_actions=None):
        self._store_prepare(root_noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch,
                            sampled_actions=sampled_actions)

    def prepare_no_noise(self, value_prefix_pool, policy_logits_pool, to_play_batch, sampled_actions=None):
        self._store_prepare(0.0, None, value_prefix_pool, policy_logits_pool, to_play_batch, sampled_actions=sampled_actions)

    def get_sampled_actions(self):
        """[N][K][D] root action tables (sez cnode.cpp:759-801); the reference returns Action objects whose .value
        is the D-vector - plain lists here (policy code reads ``.value`` only through np.array(), which accepts both)."""
        return self._prep["sampled_actions"].tolist()

    def get_sampled_actions_tensor(self):
        return self._prep["sampled_actions"]


class Action:
    """ezs_tree.pyx:26-36: Action(value, is_root_action) - a holder of one sampled action's D-vector."""

    def __init__(self, value=(), is_root_action=0):
        self.value = [float(v) for v in value]
        self.is_root_action = int(is_root_action)


class Node(NodeBase):
    """ezs_tree.pyx:86-101: Node(prior, legal_actions, action_space_size, num_of_sampled_actions, continuous_action_space);
    expand() samples the node's K actions on the device like Roots.prepare_no_noise does."""
    _roots_cls = Roots

    def __init__(self, prior=0.0, legal_actions=None, action_space_size=1, num_of_sampled_actions=1,
                 continuous_action_space=False):
        super().__init__(prior, legal_actions)
        self.action_space_size = int(action_space_size)
        self.num_of_sampled_actions = int(num_of_sampled_actions)
        self.continuous_action_space = bool(continuous_action_space)

    def _make_roots(self):
        return Roots(1, [[]], self.action_space_size, self.num_of_sampled_actions, self.continuous_action_space)

    def sampled_actions(self):
        """[K] Action objects of the expanded node (what CNode::legal_actions holds after expand)."""
        return [Action(v, 1) for v in self.roots.get_sampled_actions()[0]]


def batch_traverse(roots, pb_c_base, pb_c_init, discount_factor, min_max_stats_lst, results, virtual_to_play_batch,
                   continuous_action_space):
    """ezs_tree.pyx:114-119 -> cbatch_traverse (sez cnode.cpp:1110-1187); last_actions is [N][D] floats."""
    f = roots._materialize()
    cfg = _search_cfg(f, pb_c_base, pb_c_init, discount_factor, virtual_to_play_batch, roots=roots)
    o = traverse_tensors(roots, cfg, virtual_to_play_batch, results, min_max_stats_lst)
    return o["ix"].tolist(), o["iy"].tolist(), o["last_action_f"].tolist(), o["vtp"].tolist()


def batch_backpropagate(current_latent_state_index, discount_factor, value_prefixs, values, policies, min_max_stats_lst,
                        results, is_reset_list, to_play_batch, sampled_actions=None):
    """ezs_tree.pyx:103-112 -> cbatch_backpropagate (sez cnode.cpp:924-971).  ``sampled_actions`` [N, K, D] is the new
    leaves' action table; None = draw it on the device from ``policies``."""
    if results.out is None:
        raise RuntimeError("batch_backpropagate needs the ResultsWrapper of the preceding batch_traverse")
    f = results.forest
    if sampled_actions is None:
        from ...tree_search.sampling import sample_actions
        r = results.roots
        sampled_actions = sample_actions(_dev(policies, torch.float32, f.device), r.env_action_space_size, f.A,
                                         r.continuous_action_space, seed=getattr(r, "sample_seed", 0),
                                         counter=int(current_latent_state_index), root_offset=r.root_offset)
    backprop_tensors(results, current_latent_state_index, discount_factor, value_prefixs, values, None, to_play_batch,
                     is_reset=is_reset_list, sampled_actions=sampled_actions)
