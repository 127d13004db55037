"""Shared implementation of the reference's tree modules over the CUDA forest.

Mirrors the surface of lzero/mcts/ctree/ctree_*/{mz,ez,gmz,ezs}_tree.pyx (Roots, MinMaxStatsList,
ResultsWrapper, batch_traverse, batch_backpropagate / batch_back_propagate) so that lzero/policy/* and
lzero/mcts/buffer/* call it unchanged.  Every argument the reference takes as a Python list may also be a
numpy array or a torch tensor (host or device); lists are converted with ONE host->device copy per argument.
List-returning getters keep the reference's return types; ``*_tensor`` twins return device tensors without
a host sync and are what the rewritten search() loops use.

A forest needs its capacity (number of simulations) when it is created, which the reference API only
reveals at search() time; Roots therefore keeps the prepared root inputs and materialises the forest on
first use (search() passes cfg.num_simulations; a bare batch_traverse() uses DEFAULT_MAX_SIMULATIONS).
"""
import os

import itertools

import numpy as np
import torch

from .. import _lib
from ..forest import RNG_BATCH, RNG_PER_ROOT, Forest, _dev

DEFAULT_MAX_SIMULATIONS = int(os.environ.get("LZB_DEFAULT_MAX_SIMULATIONS", "256"))
# Tie-break stream (include/lzb200.h).  The reference draws rand() reseeded from the wall clock on every traverse
# (cnode.cpp:592,770): a fresh stream per call, not reproducible.  Here the seed is explicit and belongs to the ROOTS
# object (``roots.tie_break_seed`` / ``roots.tie_break_mode``), i.e. to one search:
#   * default: every new Roots takes derive_seed(base, ordinal) - base from LZB_TIE_BREAK_SEED (default 0), ordinal =
#     how many Roots this process has created - so consecutive searches break their ties differently, like the
#     reference, yet a run is reproducible from its base seed;
#   * pinned: after set_tie_break(seed[, mode]) every new Roots takes exactly ``seed`` (parity tests); a driver whose
#     cfg carries ``tie_break_seed`` pins (or, with ``tie_break_advance=True``, derives per search from) that seed.
_default = dict(seed=int(os.environ.get("LZB_TIE_BREAK_SEED", "0")), mode=RNG_PER_ROOT, pinned=False)
_roots_ordinal = itertools.count(1)
_M64 = 0xffffffffffffffff


def derive_seed(base, ordinal):
    """Seed of the ``ordinal``-th search of a run with base seed ``base`` (splitmix64 of base + golden * ordinal)."""
    z = (int(base) + 0x9E3779B97F4A7C15 * int(ordinal)) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def _mode_id(mode):
    return RNG_BATCH if mode in ("batch", RNG_BATCH) else RNG_PER_ROOT


def set_tie_break(seed, mode="per_root"):
    """Pin the tie-break seed (and keying) every Roots object created from now on starts with; ``seed=None`` returns
    to the default of a fresh derived seed per Roots."""
    _default["pinned"] = seed is not None
    if seed is not None:
        _default["seed"] = int(seed)
    _default["mode"] = _mode_id(mode)


def _device(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("lightzero_b200 needs a CUDA device: the tree has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def _rows_to_padded(rows, width, dtype):
    """list of variable-length rows -> [len(rows), width] zero-padded host array (one pass, no per-element python)."""
    if isinstance(rows, (np.ndarray, torch.Tensor)):
        return rows
    n = len(rows)
    P = _lib.pylist() if dtype == np.float32 else None
    if P is not None:  # C loops over the list objects (lzb_pylist.c): ~10x faster than numpy on lists of lists
        out = np.empty((n, width), dtype=np.float32)
        if P.lzb_rows_to_f32(rows, out.ctypes.data, n, width, None) != 0:
            raise ValueError("could not read the rows")
        return out
    lens = np.fromiter((len(r) for r in rows), dtype=np.int64, count=n)
    flat = np.fromiter(itertools.chain.from_iterable(rows), dtype=dtype, count=int(lens.sum()))  # ~30 % faster than np.asarray(rows)
    if n and (lens == width).all():
        return flat.reshape(n, width)
    out = np.zeros((n, width), dtype=dtype)
    mask = np.arange(width)[None, :] < lens[:, None]
    out[mask] = flat
    return out


class MinMaxStatsList:
    """common_lib/cminimax.h:27-37 through mz_tree.pyx:5-15.  The per-root {min, max} pairs live in the forest's
    workspace; this object carries value_delta_max and is bound to a forest by the first batch_traverse."""

    def __init__(self, num):
        self.num = int(num)
        self.value_delta_max = 0.0
        self._bound = None

    def set_delta(self, value_delta_max):
        self.value_delta_max = float(value_delta_max)
        if self._bound is not None:
            self._bound.set_delta(self.value_delta_max)

    def _bind(self, forest):
        if self._bound is not forest:
            forest.set_delta(self.value_delta_max)  # a fresh list: bounds cleared (cminimax.cpp:7-11)
            self._bound = forest


class ResultsWrapper:
    """tree::CSearchResults (cnode.h:67-79) through mz_tree.pyx:17-24: the per-simulation hand-over from
    batch_traverse to batch_backpropagate.  Holds device tensors; the search path itself stays in the workspace."""

    def __init__(self, num):
        self.num = int(num)
        self.out = None

    def get_search_len(self):
        return [] if self.out is None else self.out["search_len"].tolist()

    def get_search_len_tensor(self):
        return None if self.out is None else self.out["search_len"]


class RootsBase:
    """CRoots (cnode.cpp:292-417) through mz_tree.pyx:26-58.  One object = the N roots of a batch; it owns the
    CUDA forest (workspace in HBM) once materialised."""

    variant = "mz"

    def __init__(self, root_num, legal_actions_list, device=None):
        self.root_num = int(root_num)
        if legal_actions_list is not None and not isinstance(legal_actions_list, tuple) and len(legal_actions_list) != self.root_num:
            raise ValueError("legal_actions_list must have root_num rows")
        self._legal = legal_actions_list
        self._device = _device(device)
        self._prep = None
        self.forest = None
        self.root_offset = 0  # global index of root 0 when the env batch is sharded over GPUs
        self._cleared = False
        # this search's tie-break stream (see the module comment); a driver may overwrite both before the search
        self.tie_break_base = _default["seed"]
        self.tie_break_seed = _default["seed"] if _default["pinned"] else derive_seed(_default["seed"], next(_roots_ordinal))
        self.tie_break_mode = _default["mode"]
        self._tie_seed_dev = None  # [1] int64 device tensor read by the kernels instead (CUDA-graph searches)

    # ------------------------------------------------------------------ reference surface
    @property
    def num(self):
        return self.root_num

    def _store_prepare(self, noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch, values=None):
        d = self._device
        logits = _dev(policy_logits_pool, torch.float32, d)
        if logits.dim() != 2 or logits.shape[0] != self.root_num:
            raise ValueError("policy_logits_pool must be [root_num, action_space_size]")
        A = logits.shape[1]
        noise = None
        if noises is not None:
            noise = _dev(_rows_to_padded(noises, A, np.float32), torch.float32, d)
            if tuple(noise.shape) != (self.root_num, A):
                raise ValueError("noises must have one row per root with at most action_space_size entries")
        to_play = _dev(to_play_batch, torch.int32, d)
        self._prep = dict(logits=logits, noise=noise, noise_weight=float(noise_weight),
                          reward=_dev(value_prefix_pool, torch.float32, d).reshape(-1), to_play=to_play.reshape(-1),
                          value=None if values is None else _dev(values, torch.float32, d).reshape(-1))
        self.action_space_size = A
        self._cleared = False
        self._cached_outputs = None  # outputs of an earlier (graph) search do not describe the new roots
        if self.forest is not None:  # prepare() on a live object re-roots the same workspace
            self._apply_prepare()

    def prepare(self, root_noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch):
        self._store_prepare(root_noise_weight, noises, value_prefix_pool, policy_logits_pool, to_play_batch)

    def prepare_no_noise(self, value_prefix_pool, policy_logits_pool, to_play_batch):
        self._store_prepare(0.0, None, value_prefix_pool, policy_logits_pool, to_play_batch)

    def get_distributions(self):
        """[N][n_legal_i] child visit counts in legal-action order (cnode.cpp:259-277,387-403); [] per root if unexpanded."""
        if self._cleared:
            return []
        if self._prep is None:
            return [[] for _ in range(self.root_num)]
        out = self._outputs()
        visits, cnt = out["visits"].cpu().numpy(), out["legal_count"].cpu().numpy()
        P = _lib.pylist()
        if P is not None and visits.dtype == np.int32 and cnt.dtype == np.int32:
            visits, cnt = np.ascontiguousarray(visits), np.ascontiguousarray(cnt)
            return P.lzb_rows_from_i32(visits.ctypes.data, cnt.ctypes.data, visits.shape[0], visits.shape[1])
        if (cnt == visits.shape[1]).all():
            return visits.tolist()
        return [row[:c] for row, c in zip(visits.tolist(), cnt.tolist())]

    def get_values(self):
        if self._cleared:
            return []
        if self._prep is None:
            return [0.0] * self.root_num
        return self._outputs()["values"].tolist()

    def get_trajectories(self):
        if self._cleared:
            return []
        if self._prep is None:
            return [[] for _ in range(self.root_num)]
        acts, ln = self._materialize().trajectories()
        acts, ln = acts.cpu().numpy(), ln.cpu().numpy()
        return [acts[i, :ln[i]].tolist() for i in range(self.root_num)]

    def clear(self):
        self._cleared = True
        self._prep = None
        self.forest = None
        self._cached_outputs = None

    # ------------------------------------------------------------------ device-tensor twins (no host sync)
    def get_distributions_tensor(self):
        """(visits [N, A] int32 with row i's first legal_count[i] entries valid, legal_count [N] int32)."""
        out = self._outputs()
        return out["visits"], out["legal_count"]

    def get_values_tensor(self):
        return self._outputs()["values"]

    # ------------------------------------------------------------------ internals
    def _forest_dims(self):
        return self.action_space_size, 0

    def _forest_kw(self):
        return {}

    def _materialize(self, max_sims=None):
        if self._prep is None:
            raise RuntimeError("Roots.prepare / prepare_no_noise must be called before the roots are searched")
        if max_sims is not None:
            need = int(max_sims)
        else:
            need = self.forest.S if self.forest is not None else DEFAULT_MAX_SIMULATIONS
        if self.forest is None or self.forest.S < need:
            if self.forest is not None and self.forest.sims_done:
                raise RuntimeError("search needs %d simulations but the forest was created for %d" % (need, self.forest.S))
            A, D = self._forest_dims()
            self.forest = Forest(self.variant, self.root_num, A, need, legal_actions=self._legal, action_dim=D,
                                 root_offset=self.root_offset, device=self._device, **self._forest_kw())
            self.forest.sims_done = 0
            self._apply_prepare(reset=False)
        return self.forest

    def _apply_prepare(self, reset=True):
        p = self._prep
        if reset:
            self.forest.reset()
        self.forest.prepare(p["logits"], p["noise"], p["noise_weight"], p["reward"], p["to_play"], value=p["value"],
                            sampled_actions=p.get("sampled_actions"))
        self.forest.sims_done = 0

    def _outputs(self):
        cached = getattr(self, "_cached_outputs", None)  # set by a CUDA-graph search (tree_search/mcts_ctree.py)
        if cached is not None:
            return cached
        f = self._materialize()
        return f.root_outputs()


class NodeBase:
    """The single `Node` the reference's tree modules expose to their unit tests (mz_tree.pyx:60-72, ez_tree.pyx:70-81,
    gmz_tree.pyx:68-79, ezs_tree.pyx:86-101): `Node(prior, legal_actions).expand(to_play, current_latent_state_index,
    batch_index, value_prefix, policy_logits)`.  The `.pyx` constructors ignore their arguments (`__cinit__: pass`: the
    wrapped CNode is default-constructed, so every action of the logits is legal).  Here a Node is a forest of ONE tree in
    HBM: expand() is that tree's root expansion (the same k_prepare kernel the batched path runs), and `children()` reads
    the expanded node back for tests."""
    _roots_cls = None

    def __init__(self, prior=0.0, legal_actions=None, *extra):
        self.prior = float(prior)
        self.legal_actions = [] if legal_actions is None else list(legal_actions)
        self._extra = extra
        self.roots = None

    def _make_roots(self):
        return self._roots_cls(1, [[]])

    def expand(self, to_play, current_latent_state_index, batch_index, value_prefix, policy_logits):
        self.roots = self._make_roots()
        self.roots.prepare_no_noise([float(value_prefix)], [list(policy_logits)], [int(to_play)])
        self.roots._materialize(1)

    def children(self):
        """dict(prior, visit, value_sum, reward: [A] arrays; n_children; raw_value) of the expanded node."""
        if self.roots is None:
            raise RuntimeError("Node.expand has not been called")
        d = self.roots._materialize().dump_tree(0)
        n = int(d["hdr"][0][0])
        return dict(prior=d["prior"][0][:n], visit=d["visit"][0][:n], value_sum=d["value_sum"][0][:n],
                    reward=d["reward"][0][:n], n_children=n, raw_value=float(d["raw_value"][0]))


def _search_cfg(forest, pb_c_base, pb_c_init, discount, to_play, num_simulations=0, max_considered=0, players=None,
                roots=None):
    """players is batch-global (cnode.cpp:776-781): derived on the host when to_play came as a host object,
    else by one device reduction inside lzb_traverse (players = 0).  The tie-break seed / keying are the roots' own."""
    if players is not None:
        pass
    elif forest.variant == "gmz":
        players = 1
    elif isinstance(to_play, torch.Tensor) and to_play.is_cuda:
        players = 0
    elif not isinstance(to_play, torch.Tensor):
        tp = np.asarray(to_play)
        players = 1 if (tp.size and int(tp.max()) == -1) else 2
    elif not to_play.is_cuda:
        players = 1 if (to_play.numel() and int(to_play.max()) == -1) else 2
    seed, mode, seed_dev = _default["seed"], _default["mode"], None
    if roots is not None:
        seed, mode, seed_dev = roots.tie_break_seed, roots.tie_break_mode, getattr(roots, "_tie_seed_dev", None)
    return forest.make_cfg(pb_c_base, pb_c_init, discount, players=players, num_simulations=num_simulations,
                           max_num_considered_actions=max_considered, rng_mode=mode, rng_seed=seed, rng_seed_dev=seed_dev)


def traverse_tensors(roots, cfg, to_play, results, min_max_stats_lst=None, gathers=(), out=None, max_sims=None,
                     reuse=None, **kw):
    """Device-tensor batch_traverse: returns the dict of [N] device tensors (ix, iy, last_action, search_len, vtp).
    ``reuse`` = dict(true_action, reuse_value[, infer_count]): batch_traverse_with_reuse."""
    f = roots._materialize(max_sims)
    if min_max_stats_lst is not None:
        min_max_stats_lst._bind(f)
    o = f.traverse(cfg, to_play, out=out, gathers=gathers, reuse=reuse, **kw)
    results.out = o
    results.forest = f
    results.roots = roots
    return o


def backprop_tensors(results, latent_index, discount, reward, value, logits, to_play, **kw):
    f = results.forest
    f.backprop(latent_index, discount, reward, value, logits, to_play, **kw)
    f.sims_done += 1


def batch_traverse(roots, pb_c_base, pb_c_init, discount_factor, min_max_stats_lst, results, virtual_to_play_batch):
    """mz_tree.pyx:95-100 / ez_tree.pyx:108-114 -> cbatch_traverse.  Returns four Python lists like the reference
    (this list flavour costs one device->host sync per call; search() uses traverse_tensors instead)."""
    f = roots._materialize()
    cfg = _search_cfg(f, pb_c_base, pb_c_init, discount_factor, virtual_to_play_batch, roots=roots)
    o = traverse_tensors(roots, cfg, virtual_to_play_batch, results, min_max_stats_lst)
    return o["ix"].tolist(), o["iy"].tolist(), o["last_action"].tolist(), o["vtp"].tolist()


def batch_backpropagate(current_latent_state_index, discount_factor, value_prefixs, values, policies, min_max_stats_lst,
                        results, to_play_batch):
    """mz_tree.pyx:74-82 -> cbatch_backpropagate (cnode.cpp:480-500)."""
    if results.out is None:
        raise RuntimeError("batch_backpropagate needs the ResultsWrapper of the preceding batch_traverse")
    backprop_tensors(results, current_latent_state_index, discount_factor, value_prefixs, values, policies, to_play_batch)


def batch_backpropagate_ez(current_latent_state_index, discount_factor, value_prefixs, values, policies,
                           min_max_stats_lst, results, is_reset_list, to_play_batch):
    """ez_tree.pyx:83-93 -> cbatch_backpropagate with is_reset_list (ctree_efficientzero/lib/cnode.cpp:577-601)."""
    if results.out is None:
        raise RuntimeError("batch_backpropagate needs the ResultsWrapper of the preceding batch_traverse")
    backprop_tensors(results, current_latent_state_index, discount_factor, value_prefixs, values, policies, to_play_batch,
                     is_reset=is_reset_list)


def batch_traverse_with_reuse(roots, pb_c_base, pb_c_init, discount_factor, min_max_stats_lst, results,
                              virtual_to_play_batch, true_action, reuse_value):
    """mz_tree.pyx:102-107 / ez_tree.pyx:116-121 -> cbatch_traverse_with_reuse (ReZero, cnode.cpp:826-928).  Same four
    lists as batch_traverse; a tree that needs no inference reports ix = -1 and iy = its own index."""
    f = roots._materialize()
    cfg = _search_cfg(f, pb_c_base, pb_c_init, discount_factor, virtual_to_play_batch, roots=roots)
    o = traverse_tensors(roots, cfg, virtual_to_play_batch, results, min_max_stats_lst,
                         reuse=dict(true_action=true_action, reuse_value=reuse_value))
    return o["ix"].tolist(), o["iy"].tolist(), o["last_action"].tolist(), o["vtp"].tolist()


def _scatter_compacted(N, A, no_inference_lst, reuse_lst, value_prefixs, values, policies):
    """The reference passes model outputs COMPACTED over the trees that inferred (count_b, cnode.cpp:527-545) plus two
    ascending index lists ending in -1.  The library takes rows indexed by tree: scatter, and describe the lists as
    per-tree flags (bit 0 no inference, bit 1 reuse) and the compacted row of every tree."""
    skip = np.zeros(N, bool)
    skip[[i for i in no_inference_lst if i >= 0]] = True
    flags = skip.astype(np.int32)
    flags[[i for i in reuse_lst if i >= 0]] |= 2
    rows = np.flatnonzero(~skip)
    vp = np.asarray(value_prefixs, np.float32).reshape(-1)
    va = np.asarray(values, np.float32).reshape(-1)
    po = np.asarray(policies, np.float32).reshape(-1, A) if len(rows) else np.zeros((0, A), np.float32)
    if not (len(vp) == len(va) == len(po) == len(rows)):
        raise ValueError("with_reuse: %d trees need inference but the batches hold %d / %d / %d rows"
                         % (len(rows), len(vp), len(va), len(po)))
    reward, value, logits = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, A), np.float32)
    reward[rows], value[rows], logits[rows] = vp, va, po
    batch_index = (np.cumsum(~skip) - 1).astype(np.int32)
    return reward, value, logits, flags, batch_index


def batch_backpropagate_with_reuse(current_latent_state_index, discount_factor, value_prefixs, values, policies,
                                   min_max_stats_lst, results, to_play_batch, no_inference_lst, reuse_lst,
                                   reuse_value_lst):
    """mz_tree.pyx:84-93 -> cbatch_backpropagate_with_reuse (cnode.cpp:502-549); arguments as the reference's."""
    if results.out is None:
        raise RuntimeError("batch_backpropagate_with_reuse needs the ResultsWrapper of the preceding traverse")
    f = results.forest
    reward, value, logits, flags, bidx = _scatter_compacted(f.N, f.A, no_inference_lst, reuse_lst, value_prefixs, values,
                                                            policies)
    backprop_tensors(results, current_latent_state_index, discount_factor, reward, value, logits, to_play_batch,
                     reuse=dict(reuse_value=reuse_value_lst, flags=flags, batch_index=bidx))


def batch_backpropagate_with_reuse_ez(current_latent_state_index, discount_factor, value_prefixs, values, policies,
                                      min_max_stats_lst, results, is_reset_list, to_play_batch, no_inference_lst,
                                      reuse_lst, reuse_value_lst):
    """ez_tree.pyx:95-106 -> cbatch_backpropagate_with_reuse (ctree_efficientzero/lib/cnode.cpp:603-644).
    is_reset_list is indexed by TREE, as cnode.cpp:640 reads it."""
    if results.out is None:
        raise RuntimeError("batch_backpropagate_with_reuse needs the ResultsWrapper of the preceding traverse")
    f = results.forest
    if len(is_reset_list) != f.N:
        raise ValueError("is_reset_list must have one entry per tree (cnode.cpp:640 indexes it by tree)")
    reward, value, logits, flags, bidx = _scatter_compacted(f.N, f.A, no_inference_lst, reuse_lst, value_prefixs, values,
                                                            policies)
    backprop_tensors(results, current_latent_state_index, discount_factor, reward, value, logits, to_play_batch,
                     is_reset=is_reset_list, reuse=dict(reuse_value=reuse_value_lst, flags=flags, batch_index=bidx))
