"""Drop-in for lzero/mcts/ctree/ctree_stochastic_muzero/stochastic_mz_tree.pyx over the CUDA forest (variant
LZB_SMZ): decision and chance levels alternate below a decision root (ctree_stochastic_muzero/lib/cnode.cpp:96-166,
596-660, 713-787).  Differences from mz_tree that the reference's own surface has: ``Roots`` takes
``chance_space_size``, ``batch_traverse`` returns ``leaf_node_is_chance`` first, and ``batch_backpropagate`` takes
``is_chance_list`` and ``leaf_idx_list`` and expands only the listed trees with COMPACTED model outputs.

Every [N, .] logits / noise / visit row of the forest has max(action_space_size, chance_space_size) columns
(include/lzb200.h); the padding happens here so that callers pass rows of the width their model produces.

The chance outcomes come from the counter-based stream lzb_chance_bits keyed by (seed, simulation, root, depth) -
the reference seeds an mt19937 from std::random_device for every single draw (cnode.cpp:622-623), which nothing can
replay; ``set_tie_break`` seeds both the tie-break and the chance stream."""
import numpy as np
import torch

from .._tree_api import (MinMaxStatsList, ResultsWrapper, RootsBase, _search_cfg, set_tie_break,  # noqa: F401
                         traverse_tensors)
from ...forest import _dev


def pad_columns(t, width):
    """[N, w] -> [N, width] device tensor, zero-filled beyond w (w <= width)."""
    if t is None or t.shape[1] == width:
        return t
    if t.shape[1] > width:
        raise ValueError("row of %d entries exceeds the forest's %d columns" % (t.shape[1], width))
    out = t.new_zeros((t.shape[0], width))
    out[:, :t.shape[1]] = t
    return out


class Roots(RootsBase):
    """stochastic_mz_tree.pyx:27-59: Roots(root_num, legal_actions_list, chance_space_size)."""
    variant = "smz"

    def __init__(self, root_num, legal_actions_list, chance_space_size, device=None):
        super().__init__(root_num, legal_actions_list, device=device)
        self.chance_space_size = int(chance_space_size)
        if self.chance_space_size <= 0:
            raise ValueError("chance_space_size must be positive")

    def _store_prepare(self, noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch, values=None):
        forest, self.forest = self.forest, None  # re-rooting is applied below, once the rows are padded
        try:
            super()._store_prepare(noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch, values)
        finally:
            self.forest = forest
        self.decision_actions = self.action_space_size
        self.action_space_size = max(self.decision_actions, self.chance_space_size)  # columns of every row
        self._prep["logits"] = pad_columns(self._prep["logits"], self.action_space_size)
        self._prep["noise"] = pad_columns(self._prep["noise"], self.action_space_size)
        if self.forest is not None:
            self._apply_prepare()

    def _forest_dims(self):
        return self.decision_actions, 0

    def _forest_kw(self):
        return dict(chance_space=self.chance_space_size)


def batch_traverse(roots, pb_c_base, pb_c_init, discount_factor, min_max_stats_lst, results, virtual_to_play_batch):
    """stochastic_mz_tree.pyx:87-91 -> cbatch_traverse (cnode.cpp:713-787).  Returns FIVE lists:
    (leaf_node_is_chance, ix, iy, last_actions, virtual_to_play)."""
    f = roots._materialize()
    cfg = _search_cfg(f, pb_c_base, pb_c_init, discount_factor, virtual_to_play_batch, roots=roots)
    o = traverse_tensors(roots, cfg, virtual_to_play_batch, results, min_max_stats_lst)
    return ([bool(x) for x in o["leaf_is_chance"].tolist()], o["ix"].tolist(), o["iy"].tolist(),
            o["last_action"].tolist(), o["vtp"].tolist())


def batch_backpropagate(current_latent_state_index, discount_factor, value_prefixs, values, policies, min_max_stats_lst,
                        results, to_play_batch, is_chance_list, leaf_idx_list):
    """stochastic_mz_tree.pyx:75-85 -> cbatch_backpropagate (cnode.cpp:559-587): tree leaf_idx_list[k] is expanded
    with row k of value_prefixs / values / policies and backed up; the other trees are not touched (an empty
    leaf_idx_list means every tree, :575-580).  to_play_batch is indexed by TREE.  is_chance_list is accepted and
    ignored, as the reference's expand ignores its child_is_chance argument (:115-122): a leaf's kind is its depth."""
    if results.out is None:
        raise RuntimeError("batch_backpropagate needs the ResultsWrapper of the preceding batch_traverse")
    f = results.forest
    N, W, dev = f.N, f.A, f.device
    if leaf_idx_list is None or len(leaf_idx_list) == 0:
        leaf_idx_list = range(N)
    idx = np.asarray(list(leaf_idx_list), np.int64)
    vp = np.asarray(value_prefixs, np.float32).reshape(-1)
    va = np.asarray(values, np.float32).reshape(-1)
    po = np.asarray(policies, np.float32)
    po = po.reshape(len(idx), -1) if len(idx) else np.zeros((0, W), np.float32)
    if not (len(vp) == len(va) == len(po) == len(idx)):
        raise ValueError("leaf_idx_list names %d trees but the batches hold %d / %d / %d rows"
                         % (len(idx), len(vp), len(va), len(po)))
    if po.shape[1] > W:
        raise ValueError("policy rows of %d entries exceed max(action_space_size, chance_space_size) = %d" % (po.shape[1], W))
    reward, value, logits = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, W), np.float32)
    active = np.zeros(N, np.int32)
    reward[idx], value[idx], active[idx] = vp, va, 1
    logits[idx, :po.shape[1]] = po
    f.backprop(current_latent_state_index, discount_factor, reward, value, logits, to_play_batch,
               active=_dev(active, torch.int32, dev))
    results._backed_up = getattr(results, "_backed_up", 0) + len(idx)
    if results._backed_up >= N:  # the simulation is complete once every tree of the batch has been backed up
        f.sims_done += 1
