"""Device-tensor front end of liblzb200.so: one ``Forest`` = the N search trees of one batch in HBM.

This is the fast path used by the rewritten ``search()`` loops (lightzero_b200/tree_search); the
list-based reference API (lightzero_b200/ctree/*.py) is a thin layer over it.  PyTorch only provides
device memory and the CUDA stream here.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from ._lib import BackpropIO, Chance, ChanceHead, Gather, Reuse, SearchCfg, Slots, TraverseIO, check

_CHECKED_RUN = bool(os.environ.get("LZB_CHECKED"))  # with a -DLZB_CHECKED library: raise at the end of a search that broke a check

_NVTX = bool(os.environ.get("LZB_NVTX"))  # NVTX ranges around every forest call (nsys / ncu --nvtx): lzb.prepare, lzb.step[e] ...


def _ranged(name):
    """With LZB_NVTX=1 the call runs inside an NVTX range `lzb.<name>` (plus the latent index for the per-simulation
    calls), so a timeline shows tree work and model work apart; without it the method is returned untouched."""
    def wrap(fn):
        if not _NVTX:
            return fn

        def inner(self, *a, **kw):
            tag = "lzb." + name
            if name in ("step", "backprop") and len(a) > (1 if name == "step" else 0):
                tag += "[%s]" % (a[1] if name == "step" else a[0])
            torch.cuda.nvtx.range_push(tag)
            try:
                return fn(self, *a, **kw)
            finally:
                torch.cuda.nvtx.range_pop()
        inner.__name__, inner.__doc__ = fn.__name__, fn.__doc__
        return inner
    return wrap


VARIANT_ID = {"mz": 0, "ez": 1, "gmz": 2, "sez": 3, "smz": 4}
RNG_BATCH, RNG_PER_ROOT = 1, 2


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _dev(x, dtype, device):
    """Bring a list / numpy array / tensor to a contiguous device tensor of ``dtype`` (one H2D copy at most)."""
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        if x.dtype == dtype and x.device == device and x.is_contiguous():
            return x  # the per-simulation case: nothing to do, and .to() / .contiguous() cost ~3 us of host time per tensor
        return x.to(device=device, dtype=dtype, non_blocking=True).contiguous()
    if dtype == torch.float32 and isinstance(x, (list, tuple)) and len(x):
        x = _floats_from_lists(x)
    elif dtype == torch.int32 and isinstance(x, (list, tuple)) and len(x):
        x = _ints_from_list(x)
    return torch.as_tensor(x, dtype=dtype).to(device, non_blocking=True).contiguous()


def _ints_from_list(x):
    """[N] Python list of ints (to_play_batch) -> int32 array through lzb_pylist.c; anything else is returned as it came."""
    P = _lib.pylist()
    if P is None or not isinstance(x[0], int):
        return x
    arr = np.empty(len(x), np.int32)
    if P.lzb_seq_to_i32(x, arr.ctypes.data, len(x)) != 0:
        raise ValueError("could not read the integers")
    return arr


def _floats_from_lists(x):
    """[N] or [N][A] Python lists of numbers -> float32 array through the C loops of lzb_pylist.c (torch.as_tensor walks
    a list of lists at ~70 ns per element); anything else is returned as it came."""
    P = _lib.pylist()
    first = x[0]
    if P is None:
        return x
    if isinstance(first, (list, tuple)):
        n, w = len(x), len(first)
        arr, lens = np.empty((n, w), np.float32), np.empty(n, np.int32)
        if P.lzb_rows_to_f32(x, arr.ctypes.data, n, w, lens.ctypes.data) != 0 or (lens != w).any():
            raise ValueError("rows of different lengths")
        return arr
    if isinstance(first, (int, float)):
        arr = np.empty(len(x), np.float32)
        if P.lzb_seq_to_f32(x, arr.ctypes.data, len(x)) != 0:
            raise ValueError("could not read the numbers")
        return arr
    return x


def legal_tensors(legal_actions, N, A, variant, device):
    """Per-root legal-action lists (or a (count, ids) pair) -> (count [N], ids [N, A]) device tensors;
    (None, None) = every action legal."""
    cnt = idx = None
    if legal_actions is not None and variant != "sez":
        if isinstance(legal_actions, tuple):
            cnt, idx = legal_actions
            cnt, idx = _dev(cnt, torch.int32, device), _dev(idx, torch.int32, device)
        else:
            P = _lib.pylist()
            if P is not None and isinstance(legal_actions, (list, tuple)):
                # C loops over the list objects (lzb_pylist.c): 0.5 ms at 8192 x 18 where torch.as_tensor needs 150 ms
                idx_a, cnt_a = np.empty((N, A), np.int32), np.empty(N, np.int32)
                if P.lzb_rows_to_i32(legal_actions, idx_a.ctypes.data, N, A, cnt_a.ctypes.data) != 0:
                    raise ValueError("could not read the legal-action lists")
                return (torch.from_numpy(cnt_a).to(device, non_blocking=True), torch.from_numpy(idx_a).to(device, non_blocking=True))
            lens = [len(l) for l in legal_actions]
            if max(lens, default=0) > A:
                raise ValueError("a root has more legal actions than the action space")
            cnt_h = torch.tensor(lens, dtype=torch.int32)
            idx_h = torch.zeros((N, A), dtype=torch.int32)
            if lens and min(lens) == max(lens) and lens[0] > 0:
                idx_h[:, :lens[0]] = torch.as_tensor(legal_actions, dtype=torch.int32)
            else:
                for i, l in enumerate(legal_actions):
                    if len(l):
                        idx_h[i, :len(l)] = torch.as_tensor(l, dtype=torch.int32)
            cnt, idx = cnt_h.to(device), idx_h.to(device)
    return cnt, idx


class Forest:
    def __init__(self, variant, num_roots, action_space, max_sims, legal_actions=None, action_dim=0, root_offset=0,
                 device="cuda", workspace=None, chance_space=0):
        """``chance_space``: chance_space_size of the stochastic tree ("smz"); its nodes hold max(action_space,
        chance_space) child records and every [N, A] tensor of this forest has that many columns (``self.A``)."""
        if not torch.cuda.is_available():
            raise RuntimeError("lightzero_b200 needs a CUDA device (there is no CPU fallback)")
        self.L = _lib.lib()
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.variant = variant
        self.N, self.A, self.S, self.D = int(num_roots), int(action_space), int(max_sims), int(action_dim)
        self.Ad, self.C = self.A, int(chance_space)  # decision actions / chance outcomes per node
        if variant == "smz":
            if self.C <= 0:
                raise ValueError("the stochastic tree needs chance_space > 0")
            self.A = max(self.Ad, self.C)
        extra = self.C if variant == "smz" else self.D  # the ABI's action_dim slot carries C for the stochastic tree
        self.root_offset = int(root_offset)
        nbytes = C.c_size_t(0)
        check(self.L.lzb_forest_bytes(VARIANT_ID[variant], self.N, self.Ad, self.S, extra, C.byref(nbytes)))
        if workspace is not None and (workspace.numel() < nbytes.value or workspace.device != self.device):
            raise ValueError("workspace must be a uint8 tensor of >= %d bytes on %s" % (nbytes.value, self.device))
        self.workspace = workspace if workspace is not None else torch.empty(nbytes.value, dtype=torch.uint8, device=self.device)
        cnt, idx = self._legal_tensors(legal_actions)
        self._legal = (cnt, idx)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            check(self.L.lzb_forest_create(C.byref(h), _p(self.workspace), nbytes.value, VARIANT_ID[variant], self.N,
                                           self.Ad, self.S, extra, self.root_offset, _p(cnt), _p(idx),
                                           _stream(self.device)))
        self.h = h
        self._slot_keep = {}

    def _legal_tensors(self, legal_actions):
        return legal_tensors(legal_actions, self.N, self.A, self.variant, self.device)

    def reset(self, legal_actions="same"):
        """A fresh batch of roots over the same workspace (Roots.clear() + Roots(...), cnode.cpp:301-317,360-366)."""
        if not (isinstance(legal_actions, str) and legal_actions == "same"):
            self._legal = self._legal_tensors(legal_actions)
        cnt, idx = self._legal
        check(self.L.lzb_forest_reset(self.h, _p(cnt), _p(idx), _stream(self.device)))
        self._slot_keep = {}

    def __del__(self):
        h = getattr(self, "h", None)
        if h:
            self.L.lzb_forest_destroy(h)
            self.h = None

    # ------------------------------------------------------------------ reference surface, tensor flavoured
    def set_delta(self, value_delta_max):
        check(self.L.lzb_set_delta(self.h, float(value_delta_max), _stream(self.device)))

    @_ranged("prepare")
    def prepare(self, logits, noise, noise_weight, reward, to_play, value=None, sampled_actions=None):
        d = self.device
        logits = _dev(logits, torch.float32, d)
        noise = _dev(noise, torch.float32, d)
        reward = _dev(reward, torch.float32, d)
        value = _dev(value, torch.float32, d)
        to_play = _dev(to_play, torch.int32, d)
        sampled_actions = _dev(sampled_actions, torch.float32, d)
        if logits is not None and tuple(logits.shape) != (self.N, self.A):
            raise ValueError("logits must be [N, A]")
        if noise is not None and tuple(noise.shape) != (self.N, self.A):
            raise ValueError("noise must be [N, A] (row i: noise by position in root i's legal list)")
        check(self.L.lzb_prepare(self.h, _p(logits), _p(noise), float(noise_weight), _p(reward), _p(value), _p(to_play),
                                 _p(sampled_actions), _stream(d)))

    def make_cfg(self, pb_c_base=19652, pb_c_init=1.25, discount=0.997, players=0, num_simulations=0,
                 max_num_considered_actions=0, rng_mode=RNG_PER_ROOT, rng_seed=0, rng_seed_dev=None):
        """``rng_seed_dev``: [1] int64 device tensor the kernels read the tie-break seed from instead of ``rng_seed``
        (a search replayed as a CUDA graph then takes a new seed per replay)."""
        cfg = SearchCfg(int(pb_c_base), float(pb_c_init), float(discount), int(players), int(num_simulations),
                        int(max_num_considered_actions), int(rng_mode), int(rng_seed) & 0xffffffffffffffff,
                        None if rng_seed_dev is None else rng_seed_dev.data_ptr())
        cfg._keep = rng_seed_dev
        return cfg

    def _reuse(self, reuse):
        """lzb_reuse_t from a dict(true_action, reuse_value[, infer_count, flags, batch_index]) of device tensors
        (ReZero search-with-reuse); returns (pointer | None, keep-alive)."""
        if reuse is None:
            return None, None
        d = self.device
        ts = dict(true_action=_dev(reuse.get("true_action"), torch.int32, d),
                  reuse_value=_dev(reuse.get("reuse_value"), torch.float32, d),
                  infer_count=reuse.get("infer_count"),
                  flags=_dev(reuse.get("flags"), torch.int32, d),
                  batch_index=_dev(reuse.get("batch_index"), torch.int32, d))
        for k in ("true_action", "reuse_value", "flags", "batch_index"):
            if ts[k] is not None and ts[k].numel() != self.N:
                raise ValueError("reuse.%s must have N elements" % k)
        r = Reuse(_p(ts["true_action"]), _p(ts["reuse_value"]), _p(ts["infer_count"]), _p(ts["flags"]),
                  _p(ts["batch_index"]))
        return C.pointer(r), (r, ts)

    def _traverse_io(self, to_play, out, gathers, reuse=None, forced_outcome=None):
        d = self.device
        N = self.N
        to_play = _dev(to_play, torch.int32, d)
        # the per-simulation case: the same output tensors, gathers and to_play as the previous call -> the same struct
        # (keyed by every pointer that goes into it; the cached tuple keeps those tensors alive)
        if out and reuse is None and forced_outcome is None:
            try:
                seed = getattr(self, "chance_seed_dev", None)
                key = (to_play.data_ptr(), 0 if seed is None else seed.data_ptr()) + tuple(
                    t.data_ptr() for t in out.values()) + tuple(
                    (0 if g[0] is None else g[0].data_ptr(), g[1].data_ptr(), len(g) > 2 and bool(g[2])) for g in gathers)
            except AttributeError:
                key = None
            hit = self.__dict__.get("_tio_cache")
            if key is not None and hit is not None and hit[0] == key:
                return hit[1], hit[2], dict(out)
        else:
            key = None
        o = dict(out) if out else {}
        for k in ("ix", "iy", "search_len", "vtp"):
            if k not in o:
                o[k] = torch.empty(N, dtype=torch.int32, device=d)
        if "last_action" not in o:
            o["last_action"] = torch.empty(N, dtype=torch.int64, device=d)
        if self.variant == "sez" and "last_action_f" not in o:
            o["last_action_f"] = torch.empty((N, self.D), dtype=torch.float32, device=d)
        if self.variant == "smz" and "leaf_is_chance" not in o:
            o["leaf_is_chance"] = torch.empty(N, dtype=torch.int32, device=d)
        garr = (Gather * max(len(gathers), 1))()
        for k, g in enumerate(gathers):
            src, dst = g[0], g[1]
            zero = int(bool(g[2])) if len(g) > 2 else 0
            if not dst.is_contiguous() or dst.shape[0] != N:
                raise ValueError("gather destination must be a contiguous [N, ...] tensor")
            row = dst[0].numel() * dst.element_size()
            if src is None:  # rows come from slot table k (set_slots / backprop(new_slots=...))
                garr[k] = Gather(None, dst.data_ptr(), row, zero)
                continue
            if not src.is_contiguous():
                raise ValueError("gather tensors must be contiguous")
            if src.shape[1] != N or src[0, 0].numel() * src.element_size() != row:
                raise ValueError("gather pool must be [E, N, row] matching dst [N, row]")
            garr[k] = Gather(src.data_ptr(), dst.data_ptr(), row, zero)
        io = TraverseIO(to_play.data_ptr(), o["ix"].data_ptr(), o["iy"].data_ptr(), o["last_action"].data_ptr(),
                        o["last_action_f"].data_ptr() if "last_action_f" in o else None, o["search_len"].data_ptr(),
                        o["vtp"].data_ptr(), garr, len(gathers))
        rp, rkeep = self._reuse(reuse)
        if rp is not None:
            io.reuse = rp
        ckeep = None
        if self.variant == "smz":
            forced = _dev(forced_outcome, torch.int32, d)
            if forced is not None and (forced.dim() != 2 or forced.shape[0] != N):
                raise ValueError("forced_outcome must be [N, stride]")
            seed = getattr(self, "chance_seed_dev", None)  # [1] int64 device tensor, or None = cfg.rng_seed
            ch = Chance(_p(forced), 0 if forced is None else int(forced.shape[1]), o["leaf_is_chance"].data_ptr(),
                        _p(seed))
            io.chance = C.pointer(ch)
            ckeep = (ch, forced, seed)
        elif forced_outcome is not None:
            raise ValueError("forced_outcome exists for the stochastic tree only")
        keep = (garr, to_play, rkeep, ckeep)
        if key is not None and len(o) == len(out):  # nothing was allocated here: `out` named every output
            self._tio_cache = (key, io, keep + (tuple(o.values()), tuple(gathers)))
        return io, keep, o

    def _slots(self, e, tensors):
        sl = Slots()
        sl.n = 0
        if tensors:
            if len(tensors) > 3:
                raise ValueError("at most 3 slot tensors (latent, LSTM c, LSTM h)")
            for k, t in enumerate(tensors):
                if not (t.is_cuda and t.is_contiguous()):
                    raise ValueError("slot tensors must be contiguous device tensors")
                sl.ptr[k] = t.data_ptr()
            sl.n = len(tensors)
            self._slot_keep[e] = tuple(tensors)  # the tree reads these rows until the search ends
        return sl

    def set_slots(self, latent_index, tensors):
        """Register the [N, row] tensors whose rows leaves below expansion ``latent_index`` gather (0 = the roots)."""
        sl = self._slots(int(latent_index), tensors)
        check(self.L.lzb_set_slots(self.h, int(latent_index), C.byref(sl), _stream(self.device)))

    def _backprop_io(self, latent_index, discount, reward, value, logits, to_play, is_reset, lstm_horizon_len,
                     is_reset_out, sampled_actions, new_slots, reuse=None, chance_head=None):
        d = self.device
        reward = _dev(reward, torch.float32, d).reshape(-1)
        value = _dev(value, torch.float32, d).reshape(-1)
        logits = _dev(logits, torch.float32, d)
        to_play = _dev(to_play, torch.int32, d)
        is_reset = _dev(is_reset, torch.int32, d)
        sampled_actions = _dev(sampled_actions, torch.float32, d)
        if reward.numel() != self.N or value.numel() != self.N:
            raise ValueError("reward / value must have N elements")
        if logits is not None and chance_head is None and (logits.numel() != self.N * self.A and self.variant != "sez"):
            raise ValueError("logits must be [N, A]")
        io = BackpropIO(int(latent_index), float(discount), _p(reward), _p(value), _p(logits), _p(to_play), _p(is_reset),
                        int(lstm_horizon_len), _p(is_reset_out), _p(sampled_actions), self._slots(int(latent_index), new_slots))
        rp, rkeep = self._reuse(reuse)
        if rp is not None:
            io.reuse = rp
        hkeep = None
        if chance_head is not None:  # lzb_chance_head_t: the other head's outputs, used by the trees whose leaf is a chance node
            if self.variant != "smz":
                raise ValueError("chance_head exists for the stochastic tree only")
            hr = _dev(chance_head["reward"], torch.float32, d).reshape(-1)
            hv = _dev(chance_head["value"], torch.float32, d).reshape(-1)
            hl = _dev(chance_head["logits"], torch.float32, d)
            lat = chance_head.get("latent")
            if hr.numel() != self.N or hv.numel() != self.N or hl.dim() != 2 or hl.shape[0] != self.N or logits.dim() != 2:
                raise ValueError("chance_head: reward / value [N], logits [N, >= C]; decision logits [N, >= A]")
            if lat is not None and not (lat.is_cuda and lat.is_contiguous()):
                raise ValueError("chance_head latent must be a contiguous device tensor")
            ch = ChanceHead(_p(hr), _p(hv), _p(hl), int(hl.shape[1]), int(logits.shape[1]), _p(lat))
            io.chance_head = C.pointer(ch)
            if lat is not None:
                self._slot_keep[("chance", int(latent_index))] = lat
            hkeep = (ch, hr, hv, hl, lat)
        return io, (reward, value, logits, to_play, is_reset, sampled_actions, rkeep, hkeep)

    @_ranged("traverse")
    def traverse(self, cfg, to_play, out=None, gathers=(), reuse=None, forced_outcome=None):
        """One batch_traverse. ``out``: optional dict of preallocated tensors (ix, iy, last_action, last_action_f,
        search_len, vtp); missing ones are allocated. ``gathers``: [(pool[E,N,...] | None, dst[N,...][, zero_if_reset])];
        None = rows come from the slot table with the gather's index.  ``reuse``: dict(true_action, reuse_value
        [, infer_count]) of [N] device tensors = batch_traverse_with_reuse (ReZero).  ``forced_outcome`` (stochastic
        tree): [N, stride] outcomes to take at the chance levels instead of drawing (entries < 0 draw)."""
        io, keep, o = self._traverse_io(to_play, out, gathers, reuse, forced_outcome)
        check(self.L.lzb_traverse(self.h, C.byref(cfg), C.byref(io), _stream(self.device)))
        return o

    @_ranged("backprop")
    def backprop(self, latent_index, discount, reward, value, logits, to_play, is_reset=None, lstm_horizon_len=0,
                 is_reset_out=None, sampled_actions=None, new_slots=None, reuse=None, active=None, chance_head=None):
        """One batch_backpropagate; ``reuse``: dict(reuse_value, true_action | flags[, batch_index]) =
        batch_backpropagate_with_reuse with reward / value / logits indexed by tree.  ``active``: [N] int32, trees
        with a zero entry are left untouched (the stochastic tree's leaf_idx_list)."""
        io, keep = self._backprop_io(latent_index, discount, reward, value, logits, to_play, is_reset, lstm_horizon_len,
                                     is_reset_out, sampled_actions, new_slots, reuse, chance_head)
        active = _dev(active, torch.int32, self.device)
        if active is not None:
            if active.numel() != self.N:
                raise ValueError("active must have N elements")
            io.active_dev = active.data_ptr()
        check(self.L.lzb_backprop(self.h, C.byref(io), _stream(self.device)))

    @_ranged("step")
    def step(self, cfg, latent_index, discount, reward, value, logits, vtp, to_play, out=None, gathers=(), is_reset=None,
             lstm_horizon_len=0, is_reset_out=None, sampled_actions=None, new_slots=None, reuse=None,
             forced_outcome=None, chance_head=None):
        """backprop(latent_index, ..., vtp) of simulation s and traverse(to_play) of simulation s + 1 in ONE launch.
        ``chance_head`` (stochastic tree): dict(reward, value, logits[, latent]) = the outputs of the head run on chance
        leaves over the whole batch; reward / value / logits / new_slots[0] are then the decision head's."""
        bio, keep_b = self._backprop_io(latent_index, discount, reward, value, logits, vtp, is_reset, lstm_horizon_len,
                                        is_reset_out, sampled_actions, new_slots, reuse, chance_head)
        tio, keep_t, o = self._traverse_io(to_play, out, gathers, reuse, forced_outcome)
        check(self.L.lzb_step(self.h, C.byref(cfg), C.byref(bio), C.byref(tio), _stream(self.device)))
        return o

    def store_hidden(self, src, dst, is_reset):
        check(self.L.lzb_store_hidden(_p(src), _p(dst), _p(is_reset), self.N, src.numel() // self.N, _stream(self.device)))

    @_ranged("root_outputs")
    def root_outputs(self, discount=0.0, gumbel=False):
        d = self.device
        visits = torch.empty((self.N, self.A), dtype=torch.int32, device=d)
        legal_count = torch.empty(self.N, dtype=torch.int32, device=d)
        values = torch.empty(self.N, dtype=torch.float32, device=d)
        pol = cv = None
        if gumbel:
            pol = torch.empty((self.N, self.A), dtype=torch.float32, device=d)
            cv = torch.empty((self.N, self.A), dtype=torch.float32, device=d)
        check(self.L.lzb_root_outputs(self.h, _p(visits), _p(legal_count), _p(values), float(discount), _p(pol), _p(cv),
                                      _stream(d)))
        if _CHECKED_RUN and not torch.cuda.is_current_stream_capturing():  # scripts/gpu_checked.sh: every search ends here
            word, _ = self.check_word()
            if word:
                raise RuntimeError("checked build: index check %d was violated during this search" % word)
        out = dict(visits=visits, legal_count=legal_count, values=values)
        if gumbel:
            out.update(policies=pol, children_values=cv)
        return out

    def check_word(self):
        """(code, checked_build): the largest code of a violated index check of a -DLZB_CHECKED library since the forest
        was created / reset (0 = none; always 0 with the shipped build), and whether the loaded library is a checked one
        (profiles/r02_sanitizer.md).  Synchronises."""
        w, c = C.c_int32(0), C.c_int32(0)
        check(self.L.lzb_check_word(self.h, C.byref(w), C.byref(c), _stream(self.device)))
        return int(w.value), bool(c.value)

    def trajectories(self):
        """(actions [N, S+1] int32, lengths [N] int32): greedy best_action chains (get_trajectories)."""
        acts = torch.full((self.N, self.S + 1), -1, dtype=torch.int32, device=self.device)
        ln = torch.empty(self.N, dtype=torch.int32, device=self.device)
        check(self.L.lzb_trajectories(self.h, _p(acts), self.S + 1, _p(ln), _stream(self.device)))
        return acts, ln

    # ------------------------------------------------------------------ introspection (tests)
    def dump_tree(self, i):
        d = self.device
        E, A = self.S + 1, self.A
        visit = torch.full((E, A), -1, dtype=torch.int32, device=d)
        vsum = torch.zeros((E, A), dtype=torch.float32, device=d)
        prior = torch.zeros((E, A), dtype=torch.float32, device=d)
        reward = torch.zeros((E, A), dtype=torch.float32, device=d)
        child_e = torch.full((E, A), -2, dtype=torch.int32, device=d)
        hdr = torch.full((E, 4), -9, dtype=torch.int32, device=d)
        raw = torch.zeros(E, dtype=torch.float32, device=d)
        check(self.L.lzb_dump_tree(self.h, int(i), _p(visit), _p(vsum), _p(prior), _p(reward), _p(child_e), _p(hdr),
                                   _p(raw), _stream(d)))
        return dict(visit=visit.cpu().numpy(), value_sum=vsum.cpu().numpy(), prior=prior.cpu().numpy(),
                    reward=reward.cpu().numpy(), child_e=child_e.cpu().numpy(), hdr=hdr.cpu().numpy(),
                    raw_value=raw.cpu().numpy())

    def dump_paths(self, stride):
        out = torch.full((self.N, stride), -1, dtype=torch.int32, device=self.device)
        check(self.L.lzb_dump_paths(self.h, _p(out), int(stride), _stream(self.device)))
        return out.cpu().numpy()

    def minmax(self):
        mn = torch.empty(self.N, dtype=torch.float32, device=self.device)
        mx = torch.empty(self.N, dtype=torch.float32, device=self.device)
        check(self.L.lzb_minmax(self.h, _p(mn), _p(mx), _stream(self.device)))
        return mn.cpu().numpy(), mx.cpu().numpy()

    def root_stats(self):
        d = self.device
        vc = torch.empty(self.N, dtype=torch.int32, device=d)
        vs = torch.empty(self.N, dtype=torch.float32, device=d)
        rw = torch.empty(self.N, dtype=torch.float32, device=d)
        check(self.L.lzb_root_stats(self.h, _p(vc), _p(vs), _p(rw), _stream(d)))
        return vc.cpu().numpy(), vs.cpu().numpy(), rw.cpu().numpy()
