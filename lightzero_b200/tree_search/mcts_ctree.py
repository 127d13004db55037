"""Search drivers with the reference's names and call signatures (lzero/mcts/tree_search/mcts_ctree.py:172-321
MuZero, :623-827 EfficientZero, :956-1122 Gumbel MuZero), rewritten so that one search issues NO host<->device
round trip: select + leaf gather -> model.recurrent_inference -> inverse scalar transform -> latent store ->
expand + backup all consume and produce device tensors.  The reference moves the latent pool, the model
outputs and four index lists through numpy / Python lists every simulation (mcts_ctree.py:283-310).

`cfg` is the same object the reference takes (an EasyDict or any mapping with attribute access): keys read are
num_simulations, pb_c_base, pb_c_init, discount_factor, value_delta_max, env_type, device, lstm_horizon_len,
max_num_considered_actions, model.{support_scale, categorical_distribution}.  Optional extra keys:
tie_break_seed (int), tie_break_mode ('per_root' | 'batch'), tie_break_advance (bool).

Tie-break stream per search (ctree/_tree_api.py).  Without ``tie_break_seed`` every search breaks its ties with a fresh
stream derived from (LZB_TIE_BREAK_SEED, ordinal), as the reference's per-traverse reseeding does (cnode.cpp:770).  With
``tie_break_seed`` the stream is PINNED to that seed for every search (reproducible; what the parity tests and the
bench use) unless ``tie_break_advance=True``, which derives search k's seed from (tie_break_seed, k).  A search
replayed as a CUDA graph reads its seed from a device word, so the same rules hold across replays.
"""
import copy
import gc

import numpy as np
import torch

from ..ctree import _tree_api as api
from ..forest import _dev
from .scaling import InverseScalarTransform


class AttrDict(dict):
    """Minimal stand-in for easydict.EasyDict (not installed here): nested attribute access over a dict."""

    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        super().__setitem__(k, AttrDict(v) if isinstance(v, dict) and not isinstance(v, AttrDict) else v)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    __setattr__ = __setitem__

    def update(self, other=None, **kw):
        for k, v in dict(other or {}, **kw).items():
            if isinstance(v, dict) and isinstance(self.get(k), dict):
                self[k].update(v)
            else:
                self[k] = v


def _get(cfg, key, default=None):
    try:
        return cfg[key] if key in cfg else default
    except TypeError:
        return getattr(cfg, key, default)


class _CtreeDriver:
    """What the three discrete drivers share: config handling and the device-resident simulation loop."""

    config = dict(root_dirichlet_alpha=0.3, root_noise_weight=0.25, pb_c_base=19652, pb_c_init=1.25,
                  value_delta_max=0.01, env_type='not_board_games')
    tree = None  # the tree module (mz_tree / ez_tree / gmz_tree)

    @classmethod
    def default_config(cls):
        cfg = AttrDict(copy.deepcopy(cls.config))
        cfg.cfg_type = cls.__name__ + 'Dict'
        return cfg

    def __init__(self, cfg=None):
        default_config = self.default_config()
        default_config.update(cfg or {})
        self._cfg = default_config
        model_cfg = _get(self._cfg, "model", {}) or {}
        self.inverse_scalar_transform_handle = InverseScalarTransform(
            _get(model_cfg, "support_scale", 300), _get(self._cfg, "device", "cuda"),
            _get(model_cfg, "categorical_distribution", True))
        self._search_ordinal = 0

    @classmethod
    def roots(cls, active_collect_env_num, legal_actions):
        """mcts_ctree.py:215-225: the CRoots of one batch."""
        return cls.tree.Roots(active_collect_env_num, legal_actions)

    # ------------------------------------------------------------------ pieces of the loop
    def _scalars(self, value_head, reward_head):
        """inverse_scalar_transform_handle of the value and the reward / value_prefix head (mcts_ctree.py:302-303): one
        fused launch for both when the handle is the device transform, else the handle twice (a caller may replace it)."""
        h = self.inverse_scalar_transform_handle
        pair = getattr(h, "pair", None)
        if pair is not None:
            return pair(value_head, reward_head)
        return h(value_head), h(reward_head)

    @staticmethod
    def _on_device(forest, x):
        """Root latents / hidden states as one contiguous device tensor (one H2D copy when they come from the host)."""
        t = x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x))
        if not t.is_cuda and t.numel() * t.element_size() >= (1 << 20) and not t.is_pinned():
            # pageable host memory: through a pinned buffer filled by torch's host threads (see _SearchGraph._h2d)
            pin = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            pin.copy_(t)
            out = pin.to(forest.device, non_blocking=True).contiguous()
            forest.__dict__.setdefault("_pinned_inputs", []).append(pin)  # alive until the forest goes
            return out
        return t.to(forest.device, non_blocking=True).contiguous()

    @staticmethod
    def _slot_tensor(t, leaf):
        """A model output as a slot: contiguous, and never the gather destination itself (an identity dynamics
        network may hand its input back)."""
        t = t.contiguous()
        if t.data_ptr() == leaf.data_ptr():
            t = t.clone()
        if t.numel() != leaf.numel() or t.dtype != leaf.dtype:
            raise ValueError("recurrent_inference returned a latent / hidden state of a different size than the roots'")
        return t

    @staticmethod
    def _rows_ok(t):
        return (t[0].numel() * t.element_size()) % 4 == 0

    def _to_play(self, to_play_batch, forest):
        # board games: the reference deep-copies the list per simulation because Cython mutates nothing here; a
        # device tensor that the kernels only read serves both env types.
        return _dev(to_play_batch, torch.int32, forest.device).reshape(-1)

    def _stamp_tie_break(self, roots):
        """This search's tie-break seed / keying on the roots object (see the module docstring)."""
        if getattr(roots, "_tie_stamped", False):
            return
        cfg = self._cfg
        base = _get(cfg, "tie_break_seed")
        if _get(cfg, "tie_break_mode") is not None:
            roots.tie_break_mode = api._mode_id(cfg.tie_break_mode)
        if base is not None:
            roots.tie_break_base = int(base)
            advance = bool(_get(cfg, "tie_break_advance", False))
            roots.tie_break_seed = api.derive_seed(base, self._search_ordinal) if advance else int(base)
        self._search_ordinal += 1
        roots._tie_stamped = True

    def _begin(self, roots, model, to_play_batch, num_simulations=0, max_considered=0):
        if hasattr(model, "eval"):
            model.eval()
        self._stamp_tie_break(roots)
        roots._cached_outputs = None  # whatever an earlier (graph) search left behind does not describe this one
        cfg = self._cfg
        N, S = roots.num, int(cfg.num_simulations)
        if isinstance(to_play_batch, int):
            to_play_batch = [to_play_batch] * N
        forest = roots._materialize(S)
        mm = api.MinMaxStatsList(N)
        mm.set_delta(cfg.value_delta_max)
        mm._bind(forest)
        scfg = api._search_cfg(forest, _get(cfg, "pb_c_base", 19652), _get(cfg, "pb_c_init", 1.25), cfg.discount_factor,
                               to_play_batch, num_simulations=num_simulations, max_considered=max_considered,
                               players=getattr(roots, "_players", None), roots=roots)
        return forest, scfg, self._to_play(to_play_batch, forest), api.ResultsWrapper(N), S


    # ------------------------------------------------------------------ ReZero search-with-reuse
    def _reuse_args(self, forest, S, true_action_list, reuse_value_list):
        """Device copies of the trajectory's action / next state's searched value per root, plus the per-simulation
        counter the traverse kernels add their inferring trees to."""
        if true_action_list is None or reuse_value_list is None:
            raise ValueError("search_with_reuse needs true_action_list and reuse_value_list")
        dev = forest.device
        return dict(true_action=_dev(true_action_list, torch.int32, dev).reshape(-1),
                    reuse_value=_dev(reuse_value_list, torch.float32, dev).reshape(-1),
                    infer_count=torch.zeros(S, dtype=torch.int32, device=dev))

    @staticmethod
    def _reuse_at(reuse, s):
        """The reuse dict of the call whose traverse half belongs to simulation s (None = no traverse half)."""
        if reuse is None:
            return None
        r = dict(reuse)
        r["infer_count"] = reuse["infer_count"][s:] if s is not None else None
        return r

    @staticmethod
    def _reuse_result(reuse):
        """(length, average_infer) as search_with_reuse returns them (mcts_ctree.py:417-420): trees inferred in the
        last simulation and the mean over simulations.  The one device->host read of the search."""
        cnt = reuse["infer_count"].cpu()
        return int(cnt[-1]), float(cnt.sum()) / cnt.numel()

    # ------------------------------------------------------------------ whole-search CUDA graph (cfg.cuda_graph)
    def _graphed(self, roots, model, latent_state_roots, hidden_roots, to_play_batch):
        """Replays prepare + all S simulation rounds + root outputs as ONE CUDA graph (per chunk).

        A collector or reanalyze loop searches the same shape over and over; the first search of a shape runs eagerly
        once (warm-up) and is then captured over static device buffers.  Later searches copy their inputs into those
        buffers and replay, so the host issues one launch per search instead of ~2 per simulation.  The model must be
        capturable (no host syncs, fixed shapes).  The Roots object receives a copy of the outputs.

        cfg.search_streams = k > 1 cuts the batch into k contiguous chunks of roots, each with its own forest, graph and
        CUDA stream: trees are independent, so chunk c's host->device copies overlap chunk c-1's simulations and the
        latency-bound tree kernels of different chunks share the SMs.  Results do not depend on k (the tie-break stream
        is keyed by the global root index).  A model that is not a pure function of its inputs (table lookups in tests
        and the bench) must offer ``model.chunk(lo, hi)`` returning the model for roots [lo, hi)."""
        if roots.forest is not None:
            raise RuntimeError("cuda_graph search needs roots that have not been traversed yet")
        p = roots._prep
        if p is None:
            raise RuntimeError("Roots.prepare / prepare_no_noise must be called before search")
        self._stamp_tie_break(roots)
        N, S, dev = roots.num, int(self._cfg.num_simulations), roots._device
        if isinstance(to_play_batch, int):
            to_play_batch = [to_play_batch] * N
        lat0 = _as_tensor(latent_state_roots)
        hid = None if hidden_roots is None else [_as_tensor(h) for h in hidden_roots]
        if isinstance(to_play_batch, torch.Tensor):
            tp = to_play_batch
        else:  # a Python list goes through the C loop of lzb_pylist.c (np.asarray walks 8192 ints in 0.5 ms)
            from ..forest import _ints_from_list
            tp = torch.as_tensor(np.asarray(_ints_from_list(to_play_batch) if isinstance(to_play_batch, list) and len(to_play_batch) else to_play_batch), dtype=torch.int32)
        k = max(1, min(int(_get(self._cfg, "search_streams", 1) or 1), N))
        # players is batch-global (cnode.cpp:776-781): with several chunks it cannot be left to a per-chunk device reduction
        players = 0 if (tp.is_cuda and k == 1) else (1 if int(tp.max()) == -1 else 2)
        sa = p.get("sampled_actions")  # sampled tree: the root action tables are one more prepared input
        key = (N, roots.action_space_size, S, tuple(lat0.shape), lat0.dtype, id(model), roots._legal is None,
               p["noise"] is None, players, None if hid is None else tuple(hid[0].shape), str(dev), k,
               None if sa is None else tuple(sa.shape[1:]), roots.tie_break_mode)
        cache = self.__dict__.setdefault("_graphs", {})
        chunks = cache.get(key)
        if chunks is not None and chunks[0].model_ref() is not model:
            chunks = None  # id(model) was reused by a new object: the captured graph calls the old one
        if chunks is None:
            from ..sharding import shard_bounds
            chunks = []
            for c in range(k):
                lo, hi = shard_bounds(N, k, c)
                m = model.chunk(lo, hi) if (k > 1 and hasattr(model, "chunk")) else model
                stream = torch.cuda.Stream(device=dev) if k > 1 else None
                chunks.append(_SearchGraph(self, roots, m, lat0, hid, lo, hi, players, dev, stream, roots_model=model))
            cache[key] = chunks
        main = torch.cuda.current_stream(dev)
        for g in chunks:
            g.run(roots, lat0, hid, tp, main)
        for g in chunks:
            if g.stream is not None:
                main.wait_stream(g.stream)
        if k == 1:
            roots._cached_outputs = {n: v.clone() for n, v in chunks[0].outs.items()}
        else:
            roots._cached_outputs = {n: torch.cat([g.outs[n] for g in chunks], dim=0) for n in chunks[0].outs}

    _sampled = False


class _SearchGraph:
    """Static buffers + captured graph of one search shape for roots [lo, hi) (see _CtreeDriver._graphed)."""

    def __init__(self, driver, roots, model, lat0, hid, lo, hi, players, dev, stream, roots_model=None):
        from ..forest import legal_tensors
        p = roots._prep
        A = roots.action_space_size
        n = hi - lo
        self.driver, self.model, self.dev, self.lo, self.hi, self.stream = driver, model, dev, lo, hi, stream
        import weakref
        try:
            self.model_ref = weakref.ref(roots_model) if roots_model is not None else (lambda: None)
        except TypeError:  # objects without weak references: keep it alive instead
            self.model_ref = (lambda m=roots_model: m)
        # the tie-break seed is a device word rewritten before every replay (lzb_search_cfg_t::rng_seed_dev)
        self.tie_seed_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        keys = ("logits", "noise", "reward", "to_play", "value") + (("sampled_actions",) if "sampled_actions" in p else ())
        self.static = {k: (None if p[k] is None else torch.empty_like(p[k][lo:hi])) for k in keys}
        # sampled tree: the kernels read the sampler seed from this word, rewritten before every replay, so a replayed
        # search draws fresh action tables (lzb_sample_actions_seeded)
        # (stochastic tree: the same for the seed of its chance draws, lzb_chance_t.seed_dev)
        seeded = "sampled_actions" in p or roots.variant == "smz"
        self.seed_dev = torch.zeros(1, dtype=torch.int64, device=dev) if seeded else None
        self.noise_weight = p["noise_weight"]
        self.lat0 = torch.empty((n,) + tuple(lat0.shape[1:]), dtype=lat0.dtype, device=dev)
        # hidden states come as [1, N, H] (mcts_ctree.py:727-728)
        self.hid = None if hid is None else [torch.empty((n, h.numel() // roots.num), dtype=h.dtype, device=dev) for h in hid]
        self.tp = torch.empty(n, dtype=torch.int32, device=dev)
        self.legal = None
        if roots._legal is not None:
            cnt, idx = legal_tensors(self._legal_slice(roots), n, A, roots.variant, dev)
            self.legal = (torch.empty_like(cnt), torch.empty_like(idx))
        self.shadow = type(roots).__new__(type(roots))
        self.shadow.__dict__.update(roots.__dict__)
        self.shadow.root_num = n
        self.shadow.root_offset = roots.root_offset + lo
        self.shadow._legal = self.legal
        self.shadow.forest = None
        self.shadow._prep = dict(self.static, noise_weight=self.noise_weight)
        self.shadow._players = players if players else None
        self.shadow._seed_dev = self.seed_dev
        self.shadow._tie_seed_dev = self.tie_seed_dev
        self.shadow._tie_stamped = True  # the caller's roots carry this search's seed; _load copies it into the word
        self.graph = None
        self.outs = None

    def _legal_slice(self, roots):
        lg = roots._legal
        return (lg[0][self.lo:self.hi], lg[1][self.lo:self.hi]) if isinstance(lg, tuple) else lg[self.lo:self.hi]

    def _load(self, roots, lat0, hid, to_play):
        from ..forest import legal_tensors
        p = roots._prep
        lo, hi = self.lo, self.hi
        if p["noise_weight"] != self.noise_weight:
            raise RuntimeError("root_noise_weight changed between searches of one captured shape")
        for k, t in self.static.items():
            if t is not None:
                t.copy_(p[k][lo:hi], non_blocking=True)
        self._h2d(0, self.lat0, lat0[lo:hi])  # host -> device when the caller hands host latents over
        if self.hid is not None:
            for j, (d, s) in enumerate(zip(self.hid, hid)):
                self._h2d(1 + j, d, s.reshape(roots.num, -1)[lo:hi])
        self.tp.copy_(to_play.reshape(-1)[lo:hi], non_blocking=True)
        ts = int(roots.tie_break_seed) & 0xffffffffffffffff
        self.tie_seed_dev.fill_(ts if ts < (1 << 63) else ts - (1 << 64))
        if self.seed_dev is not None:
            seed = int(getattr(roots, "sample_seed", None) or getattr(roots, "chance_seed", None) or 0) & 0xffffffffffffffff
            self.seed_dev.fill_(seed if seed < (1 << 63) else seed - (1 << 64))
        if self.legal is not None:
            cnt, idx = legal_tensors(self._legal_slice(roots), hi - lo, roots.action_space_size, roots.variant, self.dev)
            self.legal[0].copy_(cnt, non_blocking=True)
            self.legal[1].copy_(idx, non_blocking=True)

    def _h2d(self, slot, dst, src):
        """dst.copy_(src) on the current stream.  Pageable host memory (the numpy arrays the reference's callers pass) goes
        through a pinned staging buffer of this chunk first: the host-side copy into it runs at memory speed on torch's
        threads (0.3 ms for 33 MB) and the transfer out of it is asynchronous at link speed, where a pageable cudaMemcpy
        is synchronous and half as fast - so the chunks' transfers overlap the other chunks' simulations as they do for
        pinned inputs."""
        if src.is_cuda or src.is_pinned() or src.numel() * src.element_size() < (1 << 20):
            dst.copy_(src, non_blocking=True)
            return
        stage = self.__dict__.setdefault("_stage", {})
        ent = stage.get(slot)
        if ent is None or ent[0].shape != src.shape or ent[0].dtype != src.dtype:
            ent = (torch.empty(src.shape, dtype=src.dtype, pin_memory=True), torch.cuda.Event())
            stage[slot] = ent
        else:
            ent[1].synchronize()  # the previous search's transfer out of the buffer has finished
        ent[0].copy_(src)
        dst.copy_(ent[0].view(dst.shape) if ent[0].shape != dst.shape else ent[0], non_blocking=True)
        ent[1].record()

    def _body(self):
        sh = self.shadow
        if sh.forest is None:
            sh._materialize(int(self.driver._cfg.num_simulations))
        else:
            sh.forest.reset()
            sh._apply_prepare(reset=False)
        hidden = None if self.hid is None else tuple(self.hid)
        if self.seed_dev is not None:
            sh._expansions = 0  # the same (seed, expansion) stream as an eager search of these roots
        self.driver._search(sh, self.model, self.lat0, hidden, self.tp)
        gum = sh.variant == "gmz"
        return sh.forest.root_outputs(discount=float(self.driver._cfg.discount_factor) if gum else 0.0, gumbel=gum)

    def run(self, roots, lat0, hid, to_play, main):
        if self.stream is not None:
            self.stream.wait_stream(main)  # the prepared root inputs were produced on the caller's stream
        with torch.cuda.stream(self.stream) if self.stream is not None else _nullctx():
            self._load(roots, lat0, hid, to_play)
            if self.graph is None:
                self._body()  # eager warm-up: allocator, lazy uploads (Gumbel tables), cudaFuncSetAttribute
                torch.cuda.synchronize(self.dev)
                g = torch.cuda.CUDAGraph()
                # no cyclic garbage collection while the stream is capturing: a finaliser that frees device memory (an
                # older search graph's pool, a peer buffer) is a prohibited call in global capture mode and invalidates
                # the capture - seen as an order-dependent cudaErrorStreamCaptureInvalidated in the test suite
                gc_was_on = gc.isenabled()
                gc.collect()
                gc.disable()
                try:
                    with torch.cuda.graph(g):
                        self.outs = self._body()
                finally:
                    if gc_was_on:
                        gc.enable()
                self.graph = g
            self.graph.replay()


class _nullctx:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def _mz_module():
    from ..ctree.ctree_muzero import mz_tree
    return mz_tree


def _ez_module():
    from ..ctree.ctree_efficientzero import ez_tree
    return ez_tree


def _gmz_module():
    from ..ctree.ctree_gumbel_muzero import gmz_tree
    return gmz_tree


class MuZeroMCTSCtree(_CtreeDriver):
    """lzero/mcts/tree_search/mcts_ctree.py:172-321."""
    tree = _mz_module()
    gumbel = False

    def search(self, roots, model, latent_state_roots, to_play_batch):
        if _get(self._cfg, "cuda_graph", False):
            return self._graphed(roots, model, latent_state_roots, None, to_play_batch)
        return self._search(roots, model, latent_state_roots, None, to_play_batch)

    def search_with_reuse(self, roots, model, latent_state_roots, to_play_batch, true_action_list=None,
                          reuse_value_list=None):
        """mcts_ctree.py:323-420 (ReZero): the root child on the trajectory's own action is scored with the value the
        next state's search found and is never searched below.  Returns (length, average_infer).  The model runs on
        the full [N] batch every simulation (rows of trees that need no inference are ignored) so that nothing
        leaves the device inside the loop; the reference's saving in model rows becomes idle rows instead."""
        if self.gumbel:
            raise NotImplementedError("the reference has no search_with_reuse for Gumbel MuZero")
        return self._search(roots, model, latent_state_roots, None, to_play_batch,
                            reuse_lists=(true_action_list, reuse_value_list))

    def _search(self, roots, model, latent_state_roots, hidden_roots, to_play_batch, reuse_lists=None):
        with torch.no_grad():
            cfg = self._cfg
            kw = dict(num_simulations=int(cfg.num_simulations), max_considered=int(cfg.max_num_considered_actions)) if self.gumbel else {}
            forest, scfg, tp, results, S = self._begin(roots, model, to_play_batch, **kw)
            discount = cfg.discount_factor
            lat0 = self._on_device(forest, latent_state_roots)
            reuse = self._reuse_args(forest, S, *reuse_lists) if reuse_lists is not None else None
            if not self._rows_ok(lat0):
                if reuse is not None:
                    raise ValueError("search_with_reuse needs latent rows that are a multiple of 4 bytes")
                return self._search_dense(roots, model, lat0, forest, scfg, tp, results, S)
            # latent_state_batch_in_search_path (mcts_ctree.py:247,305) is a table of S+1 base pointers in HBM: the
            # model's output tensor of simulation s IS slot s+1, nothing is copied into a pool.
            forest.set_slots(0, [lat0])
            leaf = torch.empty_like(lat0)
            gathers = [(None, leaf)]
            # stage 1 of simulation 0: selection + gather of each leaf's parent latent (mcts_ctree.py:272-289)
            o = api.traverse_tensors(roots, scfg, tp, results, gathers=gathers, reuse=self._reuse_at(reuse, 0))
            for simulation_index in range(S):
                # stage 2: expansion inputs from the model, still on the device (mcts_ctree.py:298-303);
                # Gumbel MuZero passes the action as [N, 1] (:1090)
                net = model.recurrent_inference(leaf, o["last_action"].unsqueeze(1) if self.gumbel else o["last_action"])
                value, reward = self._scalars(net.value, net.reward)
                new_lat = self._slot_tensor(net.latent_state, leaf)
                e = simulation_index + 1
                if e < S:
                    # stage 3 (expand + backup, :316-321) fused with stage 1 of the next simulation: one launch
                    o = forest.step(scfg, e, discount, reward, value, net.policy_logits, o["vtp"], tp, out=o,
                                    gathers=gathers, new_slots=[new_lat], reuse=self._reuse_at(reuse, e))
                else:
                    forest.backprop(e, discount, reward, value, net.policy_logits, o["vtp"], new_slots=[new_lat],
                                    reuse=self._reuse_at(reuse, None))
                forest.sims_done += 1
            if reuse is not None:
                return self._reuse_result(reuse)

    def _search_dense(self, roots, model, lat0, forest, scfg, tp, results, S):
        """Fallback for latent rows that are not a multiple of 4 bytes (e.g. odd-sized fp16): dense pool + torch gather."""
        discount = self._cfg.discount_factor
        pool = torch.empty((S + 1,) + tuple(lat0.shape), dtype=lat0.dtype, device=forest.device)
        pool[0].copy_(lat0)
        o = None
        for simulation_index in range(S):
            o = api.traverse_tensors(roots, scfg, tp, results, out=o)
            leaf = pool[o["ix"].long(), o["iy"].long()]
            net = model.recurrent_inference(leaf, o["last_action"].unsqueeze(1) if self.gumbel else o["last_action"])
            value, reward = self._scalars(net.value, net.reward)
            pool[simulation_index + 1].copy_(net.latent_state.reshape(lat0.shape))
            api.backprop_tensors(results, simulation_index + 1, discount, reward, value, net.policy_logits, o["vtp"])


class UniZeroMCTSCtree(MuZeroMCTSCtree):
    """lzero/mcts/tree_search/mcts_ctree.py:19-170: the MuZero tree driven by a world model that is called with the
    search's whole state-action history, ``model.recurrent_inference(state_action_history, simulation_index,
    latent_state_index_in_search_path)``.  The tree work is the fused device path of MuZeroMCTSCtree; what the world
    model's key-value cache lookup needs on the host (the leaf latents as numpy, the parent indices as a list,
    :143,153) is copied out once per simulation, as the reference does.  ``cfg.history_on_device=True`` hands the
    history over as device tensors instead (no device->host copy inside the loop) for models that accept it."""

    def search(self, roots, model, latent_state_roots, to_play_batch):
        return self._search(roots, model, latent_state_roots, None, to_play_batch)

    def search_with_reuse(self, *args, **kw):
        raise NotImplementedError("the reference has no search_with_reuse for UniZero")

    def _search(self, roots, model, latent_state_roots, hidden_roots, to_play_batch, reuse_lists=None):
        with torch.no_grad():
            forest, scfg, tp, results, S = self._begin(roots, model, to_play_batch)
            discount = self._cfg.discount_factor
            on_device = bool(_get(self._cfg, "history_on_device", False))
            lat0 = self._on_device(forest, latent_state_roots)
            if not self._rows_ok(lat0):
                raise ValueError("latent rows must be a multiple of 4 bytes")
            forest.set_slots(0, [lat0])
            leaf = torch.empty_like(lat0)
            gathers = [(None, leaf)]
            state_action_history = []
            o = api.traverse_tensors(roots, scfg, tp, results, gathers=gathers)
            for simulation_index in range(S):
                last_actions = o["last_action"].clone()  # the traverse buffers are reused by the next step
                if on_device:
                    state_action_history.append((leaf.clone(), last_actions))
                    ix = o["ix"].clone()
                else:
                    state_action_history.append((leaf.cpu().numpy(), last_actions))  # mcts_ctree.py:143
                    ix = o["ix"].tolist()
                net = model.recurrent_inference(state_action_history, simulation_index, ix)
                value, reward = self._scalars(net.value, net.reward)
                new_lat = self._slot_tensor(self._on_device(forest, net.latent_state).reshape(lat0.shape), leaf)
                e = simulation_index + 1
                if e < S:
                    o = forest.step(scfg, e, discount, reward, value, net.policy_logits, o["vtp"], tp, out=o,
                                    gathers=gathers, new_slots=[new_lat])
                else:
                    forest.backprop(e, discount, reward, value, net.policy_logits, o["vtp"], new_slots=[new_lat])
                forest.sims_done += 1


class MuZeroRNNFullObsMCTSCtree(MuZeroMCTSCtree):
    """lzero/mcts/tree_search/mcts_ctree.py:422-620: the MuZero tree with a second per-node state, the world model's
    latent history ``[1, N, H]``, gathered next to the latent; ``model.recurrent_inference(latent_states,
    world_model_latent_history, last_actions[, ready_env_id=...])`` returns ``predict_next_latent_state``,
    ``policy_logits``, ``value``, ``value_prefix`` (backed up as the reward, :596) and ``reward_hidden_state`` (the
    next history, :604-605).  Both states are slot tables in HBM; nothing is copied to the host."""

    def search(self, roots, model, latent_state_roots, world_model_latent_history_roots, to_play_batch,
               ready_env_id=None):
        return self._search(roots, model, latent_state_roots, world_model_latent_history_roots, to_play_batch,
                            ready_env_id=ready_env_id)

    def search_with_reuse(self, *args, **kw):
        raise NotImplementedError("the reference has no search_with_reuse for the RNN full-observation driver")

    def _search(self, roots, model, latent_state_roots, hidden_roots, to_play_batch, reuse_lists=None, ready_env_id=None):
        with torch.no_grad():
            forest, scfg, tp, results, S = self._begin(roots, model, to_play_batch)
            discount = self._cfg.discount_factor
            N = roots.num
            lat0 = self._on_device(forest, latent_state_roots)
            hist = hidden_roots[0] if isinstance(hidden_roots, (tuple, list)) else hidden_roots
            h0 = self._on_device(forest, hist).reshape(N, -1)  # world_model_latent_history_batch[ix][0][iy] (:562)
            if not (self._rows_ok(lat0) and self._rows_ok(h0)):
                raise ValueError("latent / history rows must be a multiple of 4 bytes")
            forest.set_slots(0, [lat0, h0])
            leaf, leaf_h = torch.empty_like(lat0), torch.empty_like(h0)
            gathers = [(None, leaf), (None, leaf_h)]
            kw = {} if ready_env_id is None else dict(ready_env_id=ready_env_id)
            o = api.traverse_tensors(roots, scfg, tp, results, gathers=gathers)
            for simulation_index in range(S):
                net = model.recurrent_inference(leaf, leaf_h.unsqueeze(0), o["last_action"], **kw)
                value, value_prefix = self._scalars(net.value, net.value_prefix)
                slots = [self._slot_tensor(net.predict_next_latent_state.reshape(lat0.shape), leaf),
                         self._slot_tensor(net.reward_hidden_state.reshape(N, -1), leaf_h)]
                e = simulation_index + 1
                if e < S:
                    o = forest.step(scfg, e, discount, value_prefix, value, net.policy_logits, o["vtp"], tp, out=o,
                                    gathers=gathers, new_slots=slots)
                else:
                    forest.backprop(e, discount, value_prefix, value, net.policy_logits, o["vtp"], new_slots=slots)
                forest.sims_done += 1


class GumbelMuZeroMCTSCtree(MuZeroMCTSCtree):
    """lzero/mcts/tree_search/mcts_ctree.py:956-1122: same loop, Gumbel tree module, actions as [N, 1]."""
    tree = _gmz_module()
    config = dict(_CtreeDriver.config, max_num_considered_actions=6)
    gumbel = True


class EfficientZeroMCTSCtree(_CtreeDriver):
    """lzero/mcts/tree_search/mcts_ctree.py:623-827 (value_prefix + LSTM hidden state with horizon reset)."""
    tree = _ez_module()
    config = dict(_CtreeDriver.config, lstm_horizon_len=5)

    def search(self, roots, model, latent_state_roots, reward_hidden_state_roots, to_play_batch):
        if _get(self._cfg, "cuda_graph", False):
            return self._graphed(roots, model, latent_state_roots, reward_hidden_state_roots, to_play_batch)
        return self._search(roots, model, latent_state_roots, reward_hidden_state_roots, to_play_batch)

    def _leaf_action(self, roots, o):
        return o["last_action"]

    def _expand_inputs(self, roots, net):
        return dict(logits=net.policy_logits)

    def search_with_reuse(self, roots, model, latent_state_roots, reward_hidden_state_roots, to_play_batch,
                          true_action_list=None, reuse_value_list=None):
        """mcts_ctree.py:829-953 (ReZero).  Returns (length, average_infer); see MuZeroMCTSCtree.search_with_reuse.
        is_reset of every leaf comes from that tree's own search length; the reference builds the list from the
        compacted lengths but ctree_efficientzero/lib/cnode.cpp:640 indexes it by tree, so the two only agree there
        when every tree infers."""
        return self._search(roots, model, latent_state_roots, reward_hidden_state_roots, to_play_batch,
                            reuse_lists=(true_action_list, reuse_value_list))

    def _search(self, roots, model, latent_state_roots, hidden_roots, to_play_batch, reuse_lists=None):
        with torch.no_grad():
            forest, scfg, tp, results, S = self._begin(roots, model, to_play_batch)
            discount, horizon = self._cfg.discount_factor, int(self._cfg.lstm_horizon_len)
            assert horizon > 0
            N = roots.num
            lat0 = self._on_device(forest, latent_state_roots)
            # reward_hidden_state_{c,h}_batch (:727-728): [1, N, H] arrays, one per expanded generation -> slot tables 1, 2
            c0 = self._on_device(forest, hidden_roots[0]).reshape(N, -1)
            h0 = self._on_device(forest, hidden_roots[1]).reshape(N, -1)
            if not all(self._rows_ok(t) for t in (lat0, c0, h0)):
                raise ValueError("latent / hidden-state rows must be a multiple of 4 bytes")
            forest.set_slots(0, [lat0, c0, h0])
            leaf, leaf_c, leaf_h = torch.empty_like(lat0), torch.empty_like(c0), torch.empty_like(h0)
            is_reset = torch.empty(N, dtype=torch.int32, device=forest.device)
            # the LSTM state of a node expanded with is_reset = 1 reads as zeros (mcts_ctree.py:809-816): the gather
            # applies the reset from the node's flag instead of rewriting the model's output tensor
            gathers = [(None, leaf), (None, leaf_c, True), (None, leaf_h, True)]
            reuse = self._reuse_args(forest, S, *reuse_lists) if reuse_lists is not None else None
            o = api.traverse_tensors(roots, scfg, tp, results, gathers=gathers, reuse=self._reuse_at(reuse, 0))
            for simulation_index in range(S):
                net = model.recurrent_inference(leaf, (leaf_c.unsqueeze(0), leaf_h.unsqueeze(0)), self._leaf_action(roots, o))
                value, value_prefix = self._scalars(net.value, net.value_prefix)
                slots = [self._slot_tensor(net.latent_state, leaf),
                         self._slot_tensor(net.reward_hidden_state[0].reshape(N, -1), leaf_c),
                         self._slot_tensor(net.reward_hidden_state[1].reshape(N, -1), leaf_h)]
                x = self._expand_inputs(roots, net)
                e = simulation_index + 1
                # is_reset = (search_len % lstm_horizon_len == 0) is computed by the kernel (mcts_ctree.py:809-814)
                kw = dict(lstm_horizon_len=horizon, is_reset_out=is_reset, sampled_actions=x.get("sampled_actions"),
                          new_slots=slots)
                if e < S:
                    o = forest.step(scfg, e, discount, value_prefix, value, x.get("logits"), o["vtp"], tp, out=o,
                                    gathers=gathers, reuse=self._reuse_at(reuse, e), **kw)
                else:
                    forest.backprop(e, discount, value_prefix, value, x.get("logits"), o["vtp"],
                                    reuse=self._reuse_at(reuse, None), **kw)
                forest.sims_done += 1
            if reuse is not None:
                return self._reuse_result(reuse)


def _as_tensor(x):
    return x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x))
