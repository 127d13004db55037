"""StochasticMuZeroMCTSCtree with the reference's name and call signature
(lzero/mcts/tree_search/mcts_ctree_stochastic.py:20-251), rewritten over the device-resident stochastic tree.

The reference moves every leaf latent through numpy, splits the batch into chance and decision leaves with Python
loops, calls the model once per kind, splits the outputs row by row into Python lists and calls
batch_backpropagate twice (:150-251).  Here the leaf rows arrive in a device tensor from the traverse kernel's own
gather, the two kinds are separated with index tensors, the outputs are written back by row into one [N] batch and
ONE fused launch backs every tree up and selects the next leaves.

Two inference modes (``cfg.stochastic_dense_inference``):
  False (default)  the model is called exactly like the reference calls it: chance leaves first
                   (``afterstate=False``), then decision leaves (``afterstate=True``), each on its own compacted
                   sub-batch.  The sub-batch sizes are data-dependent, which costs one device->host read per
                   simulation (the index tensors).
  True             both heads run on the full batch every simulation and a per-row select keeps the one each leaf
                   needs: twice the model rows, but nothing leaves the device inside the loop, so the whole search
                   is one CUDA graph (``cfg.cuda_graph``).  The action handed to the head a row does NOT use is
                   clamped into that head's range.  The select is done by the tree kernels themselves
                   (lzb_chance_head_t: a chance leaf reads the chance head's reward / value / logits, a node's latent
                   rows are gathered from the output tensor of its kind), so nothing of size [N, latent] is copied;
                   ``cfg.stochastic_select_in_torch=True`` keeps the torch.where form (tests compare the two).
"""
import torch

from ..ctree import _tree_api as api
from ..ctree.ctree_stochastic_muzero.stochastic_mz_tree import pad_columns
from .mcts_ctree import _CtreeDriver, _get


def _smz_module():
    from ..ctree.ctree_stochastic_muzero import stochastic_mz_tree
    return stochastic_mz_tree


class StochasticMuZeroMCTSCtree(_CtreeDriver):
    """lzero/mcts/tree_search/mcts_ctree_stochastic.py:20-251."""
    tree = _smz_module()

    def __init__(self, cfg=None):
        super().__init__(cfg)
        self._searches = 0

    @classmethod
    def roots(cls, active_collect_env_num, legal_actions, chance_space_size=2):
        """mcts_ctree_stochastic.py:70-83."""
        return cls.tree.Roots(active_collect_env_num, legal_actions, chance_space_size)

    def search(self, roots, model, latent_state_roots, to_play_batch):
        # every search draws its chance outcomes from its own stream: seed + number of searches this driver has run
        # (the reference seeds from std::random_device per draw, cnode.cpp:622-623); roots.chance_seed overrides
        if getattr(roots, "chance_seed", None) is None:
            roots.chance_seed = roots.tie_break_base + self._searches
        self._searches += 1
        if _get(self._cfg, "cuda_graph", False):
            if not _get(self._cfg, "stochastic_dense_inference", False):
                raise ValueError("cuda_graph needs stochastic_dense_inference=True: compacted sub-batches have "
                                 "data-dependent shapes")
            return self._graphed(roots, model, latent_state_roots, None, to_play_batch)
        return self._search(roots, model, latent_state_roots, None, to_play_batch)

    # ------------------------------------------------------------------ the two inference modes
    def _infer_compact(self, model, leaf, last_action, is_chance, W):
        """Reference call pattern (:165-224): one model call per leaf kind on the rows of that kind."""
        N = leaf.shape[0]
        dev = leaf.device
        reward = torch.empty(N, dtype=torch.float32, device=dev)
        value = torch.empty(N, dtype=torch.float32, device=dev)
        logits = torch.zeros((N, W), dtype=torch.float32, device=dev)
        new_lat = torch.empty_like(leaf)
        chance_rows = torch.nonzero(is_chance, as_tuple=False).squeeze(1)
        n_chance = chance_rows.numel()  # the one host sync of the simulation
        if n_chance == N:
            groups = [(None, True)]
        elif n_chance == 0:
            groups = [(None, False)]
        else:  # chance leaves first, as the reference orders its two calls (:226-227)
            groups = [(chance_rows, True), (torch.nonzero(is_chance == 0, as_tuple=False).squeeze(1), False)]
        for rows, kind in groups:
            whole = rows is None
            lat = leaf if whole else leaf.index_select(0, rows)
            act = last_action if whole else last_action.index_select(0, rows)
            net = model.recurrent_inference(lat, act, afterstate=not kind)
            v = self.inverse_scalar_transform_handle(net.value).reshape(-1).float()
            r = self.inverse_scalar_transform_handle(net.reward).reshape(-1).float()
            pl = net.policy_logits.float()
            nl = net.latent_state.reshape((-1,) + tuple(leaf.shape[1:])).to(leaf.dtype)
            if whole:
                value, reward, new_lat = v, r, nl
                logits[:, :pl.shape[1]] = pl
            else:
                value.index_copy_(0, rows, v)
                reward.index_copy_(0, rows, r)
                new_lat.index_copy_(0, rows, nl)
                logits.index_copy_(0, rows, pad_columns(pl, W))
        return reward, value, logits, new_lat

    def _infer_dense(self, model, leaf, last_action, is_chance, W, A, C):
        """Both heads on the full batch, per-row select; no host sync (CUDA-graph capturable)."""
        m = is_chance != 0
        # a chance leaf was reached by a decision action (< A), a decision leaf by a chance outcome (< C)
        net_c = model.recurrent_inference(leaf, last_action.clamp(max=A - 1), afterstate=False)
        net_d = model.recurrent_inference(leaf, last_action.clamp(max=C - 1), afterstate=True)
        inv = self.inverse_scalar_transform_handle
        value = torch.where(m, inv(net_c.value).reshape(-1).float(), inv(net_d.value).reshape(-1).float())
        reward = torch.where(m, inv(net_c.reward).reshape(-1).float(), inv(net_d.reward).reshape(-1).float())
        logits = torch.where(m.unsqueeze(1), pad_columns(net_c.policy_logits.float(), W),
                             pad_columns(net_d.policy_logits.float(), W))
        mm = m.reshape((-1,) + (1,) * (leaf.dim() - 1))
        new_lat = torch.where(mm, net_c.latent_state.reshape(leaf.shape).to(leaf.dtype),
                              net_d.latent_state.reshape(leaf.shape).to(leaf.dtype))
        return reward, value, logits, new_lat

    def _infer_heads(self, model, leaf, last_action, A, C):
        """Both heads on the full batch; the tree picks per leaf (lzb_chance_head_t).  Returns the decision head's
        (reward, value, logits, latent) and the chance head's dict."""
        # a chance leaf was reached by a decision action (< A), a decision leaf by a chance outcome (< C): the action a
        # head gets for a row it is not meant for is clamped into its range (both clamps in one launch)
        if A == C:  # every action / outcome index is in range for both heads: no clamp, one launch less per simulation
            acts = (last_action, last_action)
        else:
            lim = self.__dict__.get("_act_limits")
            if lim is None or lim[0] != (A, C, last_action.device):
                lim = ((A, C, last_action.device), torch.tensor([[A - 1], [C - 1]], dtype=last_action.dtype, device=last_action.device))
                self._act_limits = lim
            acts = torch.minimum(last_action.unsqueeze(0), lim[1])
        net_c = model.recurrent_inference(leaf, acts[0], afterstate=False)
        net_d = model.recurrent_inference(leaf, acts[1], afterstate=True)
        inv = self.inverse_scalar_transform_handle
        shape = lambda t: t.reshape(leaf.shape).to(leaf.dtype)
        head = dict(reward=inv(net_c.reward).reshape(-1).float(), value=inv(net_c.value).reshape(-1).float(),
                    logits=net_c.policy_logits.float().contiguous(), latent=self._slot_tensor(shape(net_c.latent_state), leaf))
        return (inv(net_d.reward).reshape(-1).float(), inv(net_d.value).reshape(-1).float(),
                net_d.policy_logits.float().contiguous(), shape(net_d.latent_state), head)

    # ------------------------------------------------------------------ the loop
    def _search(self, roots, model, latent_state_roots, hidden_roots, to_play_batch, reuse_lists=None):
        with torch.no_grad():
            forest, scfg, tp, results, S = self._begin(roots, model, to_play_batch)
            discount = self._cfg.discount_factor
            dense = bool(_get(self._cfg, "stochastic_dense_inference", False))
            in_tree = dense and not bool(_get(self._cfg, "stochastic_select_in_torch", False))
            lat0 = self._on_device(forest, latent_state_roots)
            if not self._rows_ok(lat0):
                raise ValueError("latent rows must be a multiple of 4 bytes")
            # the chance stream's seed lives in device memory so that a captured search redraws on replay
            seed_dev = getattr(roots, "_seed_dev", None)
            if seed_dev is None:
                seed = int(getattr(roots, "chance_seed", None) or 0) & 0xffffffffffffffff
                seed_dev = torch.full((1,), seed if seed < (1 << 63) else seed - (1 << 64), dtype=torch.int64,
                                      device=forest.device)
            forest.chance_seed_dev = seed_dev
            W, A, C = forest.A, forest.Ad, forest.C
            forest.set_slots(0, [lat0])
            leaf = torch.empty_like(lat0)
            gathers = [(None, leaf)]
            o = api.traverse_tensors(roots, scfg, tp, results, gathers=gathers)
            for simulation_index in range(S):
                head = None
                if in_tree:
                    reward, value, logits, new_lat, head = self._infer_heads(model, leaf, o["last_action"], A, C)
                elif dense:
                    reward, value, logits, new_lat = self._infer_dense(model, leaf, o["last_action"],
                                                                       o["leaf_is_chance"], W, A, C)
                else:
                    reward, value, logits, new_lat = self._infer_compact(model, leaf, o["last_action"],
                                                                         o["leaf_is_chance"], W)
                new_lat = self._slot_tensor(new_lat, leaf)
                e = simulation_index + 1
                if e < S:
                    o = forest.step(scfg, e, discount, reward, value, logits, o["vtp"], tp, out=o, gathers=gathers,
                                    new_slots=[new_lat], chance_head=head)
                else:
                    forest.backprop(e, discount, reward, value, logits, o["vtp"], new_slots=[new_lat], chance_head=head)
                forest.sims_done += 1
