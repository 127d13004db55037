"""Drop-in for lzero/mcts/ctree/ctree_gumbel_muzero/gmz_tree.pyx over the CUDA forest (variant LZB_GMZ:
sequential-halving root selection, completed-Q interior selection; gmz cnode.cpp:701-790,935-1132)."""
import ctypes as C

import numpy as np
import torch

from ... import _lib
from .._tree_api import MinMaxStatsList, NodeBase, ResultsWrapper, RootsBase, backprop_tensors, set_tie_break  # noqa: F401
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


# ---- single-node helpers of gmz_tree.pyx:68-79,97-132,145-149: one-thread device evaluations (csrc/lzb_gumbel_helpers.cuh)
class Node(NodeBase):
    """gmz_tree.pyx:68-79: expand(to_play, current_latent_state_index, batch_index, value_prefix, value, policy_logits)."""
    _roots_cls = Roots

    def expand(self, to_play, current_latent_state_index, batch_index, value_prefix, value, policy_logits):
        self.roots = Roots(1, [[]])
        self.roots.prepare_no_noise([float(value_prefix)], [float(value)], [list(policy_logits)], [int(to_play)])
        self.roots._materialize(1)

    def get_q(self, discount):
        """CNode::get_q (gmz cnode.cpp:181-197): reward + discount * value() per child."""
        c = self.children()
        v = np.where(c["visit"] > 0, c["value_sum"] / np.maximum(c["visit"], 1).astype(np.float32), np.float32(0)).astype(np.float32)
        return (c["reward"].astype(np.float32) + np.float32(discount) * v).astype(np.float32)


def _helper(op, n, a=None, b=None, c=None, v=None, s=(0.0, 0.0, 0.0, 0.0), i=(0, 0), want_int=False):
    dev = torch.device("cuda", torch.cuda.current_device())
    f = lambda x: None if x is None else torch.as_tensor(np.asarray(x, dtype=np.float32), device=dev).contiguous()
    ta, tb, tc = f(a), f(b), f(c)
    tv = None if v is None else torch.as_tensor(np.asarray(v, dtype=np.int32), device=dev).contiguous()
    out = torch.zeros(max(n, 1), dtype=torch.float32, device=dev)
    out_i = torch.zeros(1, dtype=torch.int32, device=dev)
    tmp = torch.zeros(3 * max(n, 1), dtype=torch.float32, device=dev)
    p = lambda t: None if t is None else t.data_ptr()
    args = _lib.GumbelHelper(int(op), int(n), p(ta), p(tb), p(tc), p(tv), float(s[0]), float(s[1]), float(s[2]), float(s[3]),
                             int(i[0]), int(i[1]), p(out), p(out_i), p(tmp))
    _lib.check(_lib.lib().lzb_gumbel_helper(C.byref(args), torch.cuda.current_stream(dev).cuda_stream))
    return int(out_i.item()) if want_int else out.tolist()


def softmax(py_num_list):
    """gmz_tree.pyx:105-111 -> csoftmax (gmz cnode.cpp:904-933); updates the list in place like the reference."""
    py_num_list[:] = _helper(0, len(py_num_list), a=py_num_list)
    return py_num_list


def pcompute_mixed_value(raw_value, py_q_values, py_child_visit, py_child_prior):
    """gmz_tree.pyx:113-117 -> compute_mixed_value (gmz cnode.cpp:935-970)."""
    return _helper(1, len(py_q_values), a=py_q_values, b=py_child_prior, v=py_child_visit, s=(raw_value, 0, 0, 0))[0]


def prescale_qvalues(py_value, epsilon):
    """gmz_tree.pyx:119-124 -> rescale_qvalues (gmz cnode.cpp:972-987); in place."""
    py_value[:] = _helper(2, len(py_value), a=py_value, s=(epsilon, 0, 0, 0))
    return py_value


def pqtransform_completed_by_mix_value(roots, py_child_visit, py_child_prior, discount, maxvisit_init, value_scale,
                                       rescale_values, epsilon):
    """gmz_tree.pyx:126-133 -> qtransform_completed_by_mix_value (gmz cnode.cpp:989-1040) for the expanded Node."""
    q = roots.get_q(discount)
    return _helper(3, len(py_child_visit), a=q, b=py_child_prior, v=py_child_visit,
                   s=(roots.children()["raw_value"], maxvisit_init, value_scale, epsilon), i=(int(bool(rescale_values)), 0))


def pscore_considered(considered_visit, py_gumbel, py_logits, py_normalized_qvalues, py_visit_counts):
    """gmz_tree.pyx:145-150 -> score_considered (gmz cnode.cpp:1097-1132)."""
    return _helper(4, len(py_logits), a=py_gumbel, b=py_logits, c=py_normalized_qvalues, v=py_visit_counts,
                   i=(int(considered_visit), 0))


def select_root_child(roots, discount, num_simulations, max_num_considered_actions):
    """gmz_tree.pyx:97-99 -> cselect_root_child (gmz cnode.cpp:701-745): the ACTION sequential halving picks at this node."""
    c = roots.children()
    n = c["n_children"]
    table = pget_table_of_considered_visits(max_num_considered_actions, num_simulations)
    considered = table[min(int(max_num_considered_actions), int(num_simulations))][int(c["visit"].sum())]
    pos = _helper(5, n, a=roots.get_q(discount), b=c["prior"], c=pgenerate_gumbel(10.0, 0.0, n), v=c["visit"],
                  s=(c["raw_value"], 0, 0, 0), i=(considered, 0), want_int=True)
    return pos  # the default-constructed node's legal actions are 0 .. n-1: position == action


def select_interior_child(roots, discount):
    """gmz_tree.pyx:101-103 -> cselect_interior_child (gmz cnode.cpp:747-790)."""
    c = roots.children()
    return _helper(6, c["n_children"], a=roots.get_q(discount), b=c["prior"], v=c["visit"], s=(c["raw_value"], 0, 0, 0),
                   want_int=True)
