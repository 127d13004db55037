"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the two checker libraries.

``OracleForest``  -> oracle/_build/liboracle.so   (plain-C restatement, oracle/mcts_oracle.c)
``RefForest``     -> oracle/_ref/libref_<v>.so    (compiled, unmodified reference + oracle/ref_harness*.cpp)

Both expose the same numpy-level interface so tests can drive them (and the CUDA path) in lock-step.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
VARIANTS = ("mz", "ez", "gmz", "sez", "smz")
_VID = {"mz": 0, "ez": 1, "gmz": 2, "sez": 3, "smz": 4}

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def _ptr(a, ct):
    return None if a is None else a.ctypes.data_as(C.POINTER(ct))


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _i(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


def _csr(lists):
    off = np.zeros(len(lists) + 1, np.int32)
    for k, l in enumerate(lists):
        off[k + 1] = off[k] + len(l)
    flat = np.zeros(max(int(off[-1]), 1), np.int32)
    for k, l in enumerate(lists):
        flat[off[k]:off[k + 1]] = l
    return flat, off


def _csr_f(lists):
    off = np.zeros(len(lists) + 1, np.int32)
    for k, l in enumerate(lists):
        off[k + 1] = off[k] + len(l)
    flat = np.zeros(max(int(off[-1]), 1), np.float32)
    for k, l in enumerate(lists):
        flat[off[k]:off[k + 1]] = l
    return flat, off


_oracle_lib = None


def oracle_lib():
    global _oracle_lib
    if _oracle_lib is None:
        path = os.path.join(HERE, "_build", "liboracle.so")
        if os.path.exists(os.path.join(HERE, "mcts_oracle.c")):
            path = _build.build_oracle()
        L = C.CDLL(path)
        L.orc_new.restype = C.c_void_p
        L.orc_expf.restype = C.c_float
        L.orc_expf.argtypes = [C.c_float]
        L.orc_logf.restype = C.c_float
        L.orc_logf.argtypes = [C.c_float]
        L.orc_tiebreak_hash.restype = C.c_uint32
        L.orc_tiebreak_hash.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_chance_bits.restype = C.c_uint64
        L.orc_chance_bits.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_discrete_draw.restype = C.c_int
        L.orc_discrete_draw.argtypes = [C.POINTER(C.c_double), C.c_int, C.c_double]
        L.orc_tanhf.restype = C.c_float
        L.orc_tanhf.argtypes = [C.c_float]
        L.orc_tanhf_mismatch_range.restype = C.c_long
        L.orc_tanhf_mismatch_range.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_libstdcxx_normal.argtypes = [C.c_uint32, C.c_float, C.c_float, C.c_int, C.c_void_p]
        L.orc_libstdcxx_uniform.argtypes = [C.c_uint32, C.c_int, C.c_void_p]
        L.orc_sample_libstdcxx.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p]
        for name in ("orc_expf_mismatch_range", "orc_logf_mismatch_range"):
            getattr(L, name).restype = C.c_long
            getattr(L, name).argtypes = [C.c_uint32, C.c_uint32]
        _oracle_lib = L
    return _oracle_lib


def sample_libstdcxx(logits, K, D, A, continuous, seed_base):
    """oracle/libstdcxx_rng.c: the action tables [n, K, D] the reference's expand() draws when its k-th call of a batch
    sees the clock value seed_base + k (continuous: logits [n, 2D]; discrete: logits [n, A], D = 1)."""
    logits = _f(logits)
    n = logits.shape[0]
    out = np.zeros((n, K, D if continuous else 1), np.float32)
    oracle_lib().orc_sample_libstdcxx(logits.ctypes.data, n, int(K), int(D), int(A), int(bool(continuous)),
                                      C.c_uint64(int(seed_base)), out.ctypes.data)
    return out


# ---- single-node Gumbel helpers (gmz_tree.pyx:97-149): the oracle's restatement and the compiled reference, same signatures
def _gh_case(case):
    f = lambda k: np.ascontiguousarray(case[k], np.float32)
    return f("logits"), np.ascontiguousarray(case["child_visit"], np.int32), f("child_prior")


def gumbel_helpers_oracle(case):
    """case: dict(logits [A], to_play, value_prefix, value, discount, num_simulations, max_considered, child_visit [A],
    child_prior [A], maxvisit_init, value_scale, rescale_values, epsilon, x [A], q [A], gumbel [A], nq [A], considered)
    -> dict of what every helper returns, from oracle/mcts_oracle.c (node-taking helpers: a freshly expanded root)."""
    O = oracle_lib()
    O.orc_mixed_value.restype = C.c_float
    logits, cv, cp = _gh_case(case)
    A = logits.size
    f32 = lambda k: np.ascontiguousarray(case[k], np.float32)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    out = {}
    x = f32("x").copy(); O.orc_softmax(P(x), A); out["softmax"] = x
    out["mixed_value"] = np.float32(O.orc_mixed_value(C.c_float(case["value"]), P(f32("q")), P(cv), P(cp), A))
    x = f32("x").copy(); O.orc_rescale(P(x), A, C.c_float(case["epsilon"])); out["rescale"] = x
    sc = np.zeros(A, np.float32)
    O.orc_score_considered(int(case["considered"]), P(f32("gumbel")), P(logits), P(f32("nq")), P(cv), A, P(sc)); out["score_considered"] = sc
    # the expanded node: priors = softmax of the logits (csoftmax form is NOT what expand uses: exp(l - max) / sum)
    pri = node_priors_oracle(logits)
    out["priors"] = pri
    q0 = np.zeros(A, np.float32)  # fresh children: reward 0 + discount * value() 0
    qt = np.zeros(A, np.float32)
    O.orc_qtransform(P(q0), C.c_float(case["value"]), P(cv), P(cp), A, C.c_float(case["maxvisit_init"]), C.c_float(case["value_scale"]),
                     int(case["rescale_values"]), C.c_float(case["epsilon"]), P(qt)); out["qtransform"] = qt
    zero_v = np.zeros(A, np.int32)
    gum = np.zeros(A, np.float32); O.orc_gumbel_noise(C.c_float(10.0), A, P(gum))
    tab = np.zeros((int(case["max_considered"]) + 1, int(case["num_simulations"])), np.int32)
    O.orc_gumbel_table(int(case["max_considered"]), int(case["num_simulations"]), P(tab))
    considered = int(tab[min(int(case["max_considered"]), int(case["num_simulations"]))][0])
    out["root_action"] = int(O.orc_select_root_arrays(P(q0), C.c_float(case["value"]), P(zero_v), P(pri), P(gum), A, considered))
    out["interior_action"] = int(O.orc_select_interior_arrays(P(q0), C.c_float(case["value"]), P(zero_v), P(pri), A))
    return out


def node_priors_oracle(logits):
    """CNode::expand's prior computation (gmz cnode.cpp:112-166: exp(l - max) / sum in fp32, serial) through a 1-root oracle forest."""
    logits = np.ascontiguousarray(logits, np.float32)
    f = OracleForest("gmz", 1, logits.size)
    f.prepare(0.0, None, np.zeros(1, np.float32), logits[None, :], np.full(1, -1, np.int32), values=np.zeros(1, np.float32))
    return f.dump_tree(0, 1)["prior"][0][:logits.size].copy()


def gumbel_helpers_reference(case):
    """The same dict from the compiled reference (oracle/_ref/libref_gmz.so)."""
    R = ref_lib("gmz")
    R.ref_mixed_value.restype = C.c_float
    logits, cv, cp = _gh_case(case)
    A = logits.size
    f32 = lambda k: np.ascontiguousarray(case[k], np.float32)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    out = {}
    x = f32("x").copy(); R.ref_softmax(P(x), A); out["softmax"] = x
    out["mixed_value"] = np.float32(R.ref_mixed_value(C.c_float(case["value"]), P(f32("q")), P(cv), P(cp), A))
    x = f32("x").copy(); R.ref_rescale(P(x), A, C.c_float(case["epsilon"])); out["rescale"] = x
    sc = np.zeros(A, np.float32)
    R.ref_score_considered(int(case["considered"]), P(f32("gumbel")), P(logits), P(f32("nq")), P(cv), A, P(sc)); out["score_considered"] = sc
    pri, qt = np.zeros(A, np.float32), np.zeros(A, np.float32)
    ra, ia = C.c_int(0), C.c_int(0)
    R.ref_node_helpers(int(case["to_play"]), C.c_float(case["value_prefix"]), C.c_float(case["value"]), P(logits), A,
                       C.c_float(case["discount"]), int(case["num_simulations"]), int(case["max_considered"]), P(cv), P(cp),
                       C.c_float(case["maxvisit_init"]), C.c_float(case["value_scale"]), int(case["rescale_values"]),
                       C.c_float(case["epsilon"]), P(pri), P(qt), C.byref(ra), C.byref(ia))
    out.update(priors=pri, qtransform=qt, root_action=int(ra.value), interior_action=int(ia.value))
    return out


_ref_libs = {}


def ref_lib(variant):
    """Load oracle/_ref/libref_<variant>.so (building it first when /root/reference is present)."""
    if variant not in _ref_libs:
        built = _build.build_ref()
        if variant not in built:
            return None
        L = C.CDLL(built[variant])
        L.ref_new.restype = C.c_void_p
        L.ref_rng_mode.argtypes = [C.c_int, C.c_uint64]
        _ref_libs[variant] = L
    return _ref_libs[variant]


def ref_available(variant="mz"):
    try:
        return ref_lib(variant) is not None
    except Exception:
        return False


class _ForestBase:
    """Common numpy-level surface. legal_actions: list of per-root action lists."""

    variant = "mz"
    n = 0
    A = 0

    def traverse(self, pb_c_base, pb_c_init, discount, to_play, num_simulations=0, max_considered=0, path_stride=0):
        raise NotImplementedError

    def run_search(self, model_fn, S, pb_c_base, pb_c_init, discount, to_play, max_considered=0, lstm_horizon_len=5,
                   record=None):
        """Restates the reference driver loop (lzero/mcts/tree_search/mcts_ctree.py:255-321, EZ :733-827,
        Gumbel :1054-1122) over a scripted model: model_fn(sim, ix, iy, last_action) -> dict(reward, value, logits)."""
        tp = np.asarray(to_play, np.int32)
        for s in range(S):
            out = self.traverse(pb_c_base, pb_c_init, discount, tp, num_simulations=S, max_considered=max_considered,
                                path_stride=S + 1 if record is not None else 0)
            m = model_fn(s, out["ix"], out["iy"], out["last_action"])
            is_reset = None
            if self.variant in ("ez", "sez"):
                is_reset = (out["search_len"] % lstm_horizon_len == 0).astype(np.int32)  # mcts_ctree.py:809-814
            if record is not None:
                record.append(dict(out, is_reset=is_reset))
            self.backprop(s + 1, discount, m["reward"], m["value"], m["logits"], is_reset, out["vtp"],
                          m.get("sampled_actions"))


    def run_search_reuse(self, model_fn, S, pb_c_base, pb_c_init, discount, to_play, true_action, reuse_value,
                         lstm_horizon_len=5, record=None):
        """Restates ``search_with_reuse`` (mcts_ctree.py:323-420; EZ :829-953): trees whose traverse stopped on an
        already expanded root child skip inference, model outputs are COMPACTED over the others (``count_b``).
        model_fn returns per-tree tables; row i is what the model would return for tree i's leaf.
        ``is_reset`` is derived per TREE from its own search length; the reference driver builds that list from the
        compacted lengths but cnode.cpp (ez:640) indexes it by tree, which only agrees when every tree infers."""
        tp = np.asarray(to_play, np.int32)
        ta = np.asarray(true_action, np.int32)
        rv = np.asarray(reuse_value, np.float32)
        infer_sum, length = 0, 0
        for s in range(S):
            out = self.traverse_reuse(pb_c_base, pb_c_init, discount, tp, ta, rv)
            infer = out["ix"] != -1
            no_inference_lst = [int(v) for v in out["iy"][~infer]] + [-1]
            reuse_lst = [i for i in range(self.n) if out["ix"][i] == 0 and out["last_action"][i] == ta[i]] + [-1]
            m = model_fn(s, out["ix"], out["iy"], out["last_action"])
            is_reset = None
            if self.variant == "ez":
                is_reset = (out["search_len"] % lstm_horizon_len == 0).astype(np.int32)
            if record is not None:
                record.append(dict(out, is_reset=is_reset))
            self.backprop_reuse(s + 1, discount, m["reward"][infer], m["value"][infer], m["logits"][infer], is_reset,
                                out["vtp"], no_inference_lst, reuse_lst, rv)
            length = int(infer.sum())
            infer_sum += length
        return length, infer_sum / S


    def run_search_chance(self, model_fn, S, pb_c_base, pb_c_init, discount, to_play, forced=None, record=None):
        """Restates the Stochastic MuZero driver loop (lzero/mcts/tree_search/mcts_ctree_stochastic.py:117-251):
        model_fn(sim, ix, iy, last_action, leaf_is_chance) -> dict(reward, value, logits [n, W]) where row i holds the
        chance logits when leaf i is a chance node (afterstate dynamics + prediction) and policy logits otherwise.
        ``forced``: per simulation an [n, stride] table of outcomes for the chance levels (oracle / CUDA side), or None
        to draw; the compiled reference always draws and reports what it took in out["paths"]."""
        tp = np.asarray(to_play, np.int32)
        for s in range(S):
            out = self.traverse_chance(pb_c_base, pb_c_init, discount, tp, forced=None if forced is None else forced[s],
                                       path_stride=S + 1)
            m = model_fn(s, out["ix"], out["iy"], out["last_action"], out["leaf_is_chance"])
            if record is not None:
                record.append(dict(out))
            self.backprop(s + 1, discount, m["reward"], m["value"], m["logits"], None, out["vtp"])


class OracleForest(_ForestBase):
    def __init__(self, variant, n, A, legal_actions=None, K=0, D=0, continuous=False):
        """``K``: sampled actions per node (sez) or chance_space_size (smz)."""
        self.L = oracle_lib()
        self.variant, self.n, self.A, self.K, self.D = variant, n, A, K, D
        self.W = K if variant == "sez" else (max(A, K) if variant == "smz" else A)
        if legal_actions is None:
            legal_actions = [[] for _ in range(n)]
        flat, off = _csr(legal_actions)
        self.h = C.c_void_p(self.L.orc_new(_VID[variant], n, A, _ptr(flat, C.c_int), _ptr(off, C.c_int), K, D,
                                           int(continuous)))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_free(self.h)
            self.h = None

    def set_delta(self, d):
        self.L.orc_set_delta(self.h, C.c_float(d))

    def set_rng(self, mode, seed, root_offset=0):
        self.L.orc_set_rng(self.h, mode, C.c_uint64(seed), C.c_uint32(root_offset))

    def prepare(self, noise_w, noises, rewards, logits, to_play, values=None, sampled_actions=None):
        with_noise = noises is not None
        nf, no = _csr_f(noises) if with_noise else (None, None)
        rewards, values, logits, to_play = _f(rewards), _f(values), _f(logits), _i(to_play)
        sa = _f(sampled_actions)
        self.L.orc_prepare(self.h, int(with_noise), C.c_float(noise_w), _ptr(nf, C.c_float), _ptr(no, C.c_int),
                           _ptr(rewards, C.c_float), _ptr(values, C.c_float), _ptr(logits, C.c_float),
                           _ptr(to_play, C.c_int), _ptr(sa, C.c_float))

    def traverse(self, pb_c_base, pb_c_init, discount, to_play, num_simulations=0, max_considered=0, path_stride=0):
        n = self.n
        tp = _i(to_play)
        o = {k: np.zeros(n, np.int32) for k in ("ix", "iy", "last_action", "search_len", "vtp")}
        paths = np.full((n, max(path_stride, 1)), -1, np.int32)
        laf = np.zeros((n, max(self.D, 1)), np.float32) if self.variant == "sez" else None
        self.L.orc_traverse(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), int(num_simulations),
                            int(max_considered), _ptr(tp, C.c_int), _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int),
                            _ptr(o["last_action"], C.c_int), _ptr(laf, C.c_float), _ptr(o["search_len"], C.c_int),
                            _ptr(o["vtp"], C.c_int), _ptr(paths, C.c_int) if path_stride else None, int(path_stride))
        if path_stride:
            o["paths"] = paths
        if laf is not None:
            o["last_action_f"] = laf
        return o

    def traverse_chance(self, pb_c_base, pb_c_init, discount, to_play, forced=None, path_stride=0):
        n = self.n
        tp, forced = _i(to_play), _i(forced)
        o = {k: np.zeros(n, np.int32) for k in ("ix", "iy", "last_action", "search_len", "vtp", "leaf_is_chance")}
        paths = np.full((n, max(path_stride, 1)), -1, np.int32)
        self.L.orc_traverse_chance(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), _ptr(tp, C.c_int),
                                   _ptr(forced, C.c_int), 0 if forced is None else int(forced.shape[1]),
                                   _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int), _ptr(o["last_action"], C.c_int),
                                   _ptr(o["search_len"], C.c_int), _ptr(o["vtp"], C.c_int),
                                   _ptr(o["leaf_is_chance"], C.c_int), _ptr(paths, C.c_int) if path_stride else None,
                                   int(path_stride))
        if path_stride:
            o["paths"] = paths
        return o

    def backprop(self, latent_index, discount, rewards, values, logits, is_reset, to_play, sampled_actions=None):
        rewards, values, logits = _f(rewards), _f(values), _f(logits)
        is_reset, to_play, sa = _i(is_reset), _i(to_play), _f(sampled_actions)
        self.L.orc_backprop(self.h, int(latent_index), C.c_float(discount), _ptr(rewards, C.c_float),
                            _ptr(values, C.c_float), _ptr(logits, C.c_float), _ptr(is_reset, C.c_int),
                            _ptr(to_play, C.c_int), _ptr(sa, C.c_float))

    def traverse_reuse(self, pb_c_base, pb_c_init, discount, to_play, true_action, reuse_value):
        tp, ta, rv = _i(to_play), _i(true_action), _f(reuse_value)
        o = {k: np.zeros(self.n, np.int32) for k in ("ix", "iy", "last_action", "search_len", "vtp")}
        self.L.orc_traverse_reuse(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), _ptr(tp, C.c_int),
                                  _ptr(ta, C.c_int), _ptr(rv, C.c_float), _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int),
                                  _ptr(o["last_action"], C.c_int), _ptr(o["search_len"], C.c_int), _ptr(o["vtp"], C.c_int))
        return o

    def backprop_reuse(self, latent_index, discount, rewards, values, logits, is_reset, to_play, no_inference_lst,
                       reuse_lst, reuse_value):
        rewards, values, logits = _f(rewards), _f(values), _f(logits)
        is_reset, to_play, rv = _i(is_reset), _i(to_play), _f(reuse_value)
        ni, ru = _i(no_inference_lst), _i(reuse_lst)
        self.L.orc_backprop_reuse(self.h, int(latent_index), C.c_float(discount), _ptr(rewards, C.c_float),
                                  _ptr(values, C.c_float), _ptr(logits, C.c_float), _ptr(is_reset, C.c_int),
                                  _ptr(to_play, C.c_int), _ptr(ni, C.c_int), _ptr(ru, C.c_int), _ptr(rv, C.c_float))

    def distributions(self):
        d = np.zeros((self.n, self.W), np.int32)
        ln = np.zeros(self.n, np.int32)
        self.L.orc_distributions(self.h, _ptr(d, C.c_int), _ptr(ln, C.c_int))
        return d, ln

    def values(self):
        v = np.zeros(self.n, np.float32)
        self.L.orc_values(self.h, _ptr(v, C.c_float))
        return v

    def minmax(self):
        mn, mx = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.orc_minmax(self.h, _ptr(mn, C.c_float), _ptr(mx, C.c_float))
        return mn, mx

    def gumbel_outputs(self, discount):
        p = np.zeros((self.n, self.A), np.float32)
        v = np.zeros((self.n, self.A), np.float32)
        self.L.orc_gumbel_outputs(self.h, C.c_float(discount), _ptr(p, C.c_float), _ptr(v, C.c_float))
        return p, v

    def root_stats(self):
        vc, vs, rw = np.zeros(self.n, np.int32), np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.orc_root_stats(self.h, _ptr(vc, C.c_int), _ptr(vs, C.c_float), _ptr(rw, C.c_float))
        return vc, vs, rw

    def trajectories(self, stride):
        out = np.full((self.n, stride), -1, np.int32)
        ln = np.zeros(self.n, np.int32)
        self.L.orc_trajectories(self.h, _ptr(out, C.c_int), int(stride), _ptr(ln, C.c_int))
        return out, ln

    def sampled_actions(self):
        out = np.zeros((self.n, self.K, self.D), np.float32)
        self.L.orc_sampled_actions(self.h, _ptr(out, C.c_float))
        return out

    def dump_tree(self, i, E):
        return _dump(lambda *a: self.L.orc_dump_tree(self.h, *a), i, E, self.W)


def _dump(fn, i, E, W):
    visit = np.full(E * W, -1, np.int32)
    vsum = np.zeros(E * W, np.float32)
    prior = np.zeros(E * W, np.float32)
    reward = np.zeros(E * W, np.float32)
    child_e = np.full(E * W, -2, np.int32)
    hdr = np.full(E * 4, -9, np.int32)
    raw = np.zeros(E, np.float32)
    fn(int(i), int(E), _ptr(visit, C.c_int), _ptr(vsum, C.c_float), _ptr(prior, C.c_float), _ptr(reward, C.c_float),
       _ptr(child_e, C.c_int), _ptr(hdr, C.c_int), _ptr(raw, C.c_float))
    return dict(visit=visit.reshape(E, W), value_sum=vsum.reshape(E, W), prior=prior.reshape(E, W),
                reward=reward.reshape(E, W), child_e=child_e.reshape(E, W), hdr=hdr.reshape(E, 4), raw_value=raw)


class RefForest(_ForestBase):
    """The compiled reference tree (discrete variants mz / ez / gmz) behind the same surface."""

    def __init__(self, variant, n, A, legal_actions=None, per_root=False):
        self.L = ref_lib(variant)
        if self.L is None:
            raise RuntimeError("oracle/_ref/libref_%s.so is not built and /root/reference is absent" % variant)
        self.variant, self.n, self.A, self.W = variant, n, A, A
        if legal_actions is None:
            legal_actions = [[] for _ in range(n)]
        flat, off = _csr(legal_actions)
        self.h = C.c_void_p(self.L.ref_new(n, A, _ptr(flat, C.c_int), _ptr(off, C.c_int), int(per_root)))
        self.per_root = per_root

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_free(self.h)
            self.h = None

    def set_delta(self, d):
        self.L.ref_set_delta(self.h, C.c_float(d))

    def set_rng(self, mode, seed, root_offset=0):
        """mode 0 = libc random (timing), 1 = batch-pinned, 2 = per-(root, depth) pinned (needs per_root=True)."""
        assert root_offset == 0
        assert mode != 2 or self.per_root
        self.L.ref_rng_mode(int(mode), C.c_uint64(seed))

    def prepare(self, noise_w, noises, rewards, logits, to_play, values=None, sampled_actions=None):
        with_noise = noises is not None
        nf, no = _csr_f(noises) if with_noise else (None, None)
        rewards, values, logits, to_play = _f(rewards), _f(values), _f(logits), _i(to_play)
        if self.variant == "gmz" and values is None:
            values = np.zeros(self.n, np.float32)
        self.L.ref_prepare(self.h, int(with_noise), C.c_float(noise_w), _ptr(nf, C.c_float), _ptr(no, C.c_int),
                           _ptr(rewards, C.c_float), _ptr(values, C.c_float), _ptr(logits, C.c_float),
                           _ptr(to_play, C.c_int))

    def traverse(self, pb_c_base, pb_c_init, discount, to_play, num_simulations=0, max_considered=0, path_stride=0):
        n = self.n
        tp = _i(to_play)
        o = {k: np.zeros(n, np.int32) for k in ("ix", "iy", "last_action", "search_len", "vtp")}
        paths = np.full((n, max(path_stride, 1)), -1, np.int32)
        self.L.ref_traverse(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), int(num_simulations),
                            int(max_considered), _ptr(tp, C.c_int), _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int),
                            _ptr(o["last_action"], C.c_int), _ptr(o["search_len"], C.c_int), _ptr(o["vtp"], C.c_int),
                            _ptr(paths, C.c_int) if path_stride else None, int(path_stride))
        if path_stride:
            o["paths"] = paths
        return o

    def backprop(self, latent_index, discount, rewards, values, logits, is_reset, to_play, sampled_actions=None):
        rewards, values, logits = _f(rewards), _f(values), _f(logits)
        is_reset, to_play = _i(is_reset), _i(to_play)
        self.L.ref_backprop(self.h, int(latent_index), C.c_float(discount), _ptr(rewards, C.c_float),
                            _ptr(values, C.c_float), _ptr(logits, C.c_float), _ptr(is_reset, C.c_int),
                            _ptr(to_play, C.c_int))

    def traverse_reuse(self, pb_c_base, pb_c_init, discount, to_play, true_action, reuse_value):
        tp, ta, rv = _i(to_play), _i(true_action), _f(reuse_value)
        o = {k: np.zeros(self.n, np.int32) for k in ("ix", "iy", "last_action", "search_len", "vtp")}
        self.L.ref_traverse_reuse(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), _ptr(tp, C.c_int),
                                  _ptr(ta, C.c_int), _ptr(rv, C.c_float), _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int),
                                  _ptr(o["last_action"], C.c_int), _ptr(o["search_len"], C.c_int), _ptr(o["vtp"], C.c_int))
        return o

    def backprop_reuse(self, latent_index, discount, rewards, values, logits, is_reset, to_play, no_inference_lst,
                       reuse_lst, reuse_value):
        rewards, values, logits = _f(rewards), _f(values), _f(logits)
        is_reset, to_play, rv = _i(is_reset), _i(to_play), _f(reuse_value)
        ni, ru = _i(no_inference_lst), _i(reuse_lst)
        self.L.ref_backprop_reuse(self.h, int(latent_index), C.c_float(discount), _ptr(rewards, C.c_float),
                                  _ptr(values, C.c_float), _ptr(logits, C.c_float), _ptr(is_reset, C.c_int),
                                  _ptr(to_play, C.c_int), _ptr(ni, C.c_int), _ptr(ru, C.c_int), _ptr(rv, C.c_float),
                                  int(len(values)))

    def distributions(self):
        d = np.zeros((self.n, self.A), np.int32)
        ln = np.zeros(self.n, np.int32)
        self.L.ref_distributions(self.h, _ptr(d, C.c_int), _ptr(ln, C.c_int))
        return d, ln

    def values(self):
        v = np.zeros(self.n, np.float32)
        self.L.ref_values(self.h, _ptr(v, C.c_float))
        return v

    def minmax(self):
        mn, mx = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.ref_minmax(self.h, _ptr(mn, C.c_float), _ptr(mx, C.c_float))
        return mn, mx

    def gumbel_outputs(self, discount):
        p = np.zeros((self.n, self.A), np.float32)
        v = np.zeros((self.n, self.A), np.float32)
        self.L.ref_gumbel_outputs(self.h, C.c_float(discount), _ptr(p, C.c_float), _ptr(v, C.c_float))
        return p, v

    def root_stats(self):
        vc, vs, rw = np.zeros(self.n, np.int32), np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.ref_root_stats(self.h, _ptr(vc, C.c_int), _ptr(vs, C.c_float), _ptr(rw, C.c_float))
        return vc, vs, rw

    def trajectories(self, stride):
        out = np.full((self.n, stride), -1, np.int32)
        ln = np.zeros(self.n, np.int32)
        self.L.ref_trajectories(self.h, _ptr(out, C.c_int), int(stride), _ptr(ln, C.c_int))
        return out, ln

    def dump_tree(self, i, E):
        return _dump(lambda *a: self.L.ref_dump_tree(self.h, *a), i, E, self.A)


class RefSampledForest(_ForestBase):
    """The compiled Sampled-EfficientZero reference tree (oracle/ref_harness_sampled.cpp).  The reference samples each
    node's K actions itself (wall-clock seeded); ``prepare`` and ``backprop`` RETURN the tables it drew ([n, K, D]) so
    that the oracle and the CUDA path can be handed the same ones."""

    variant = "sez"

    def __init__(self, n, A, K, continuous=True, per_root=False):
        self.L = ref_lib("sez")
        if self.L is None:
            raise RuntimeError("oracle/_ref/libref_sez.so is not built and /root/reference is absent")
        self.n, self.A, self.K, self.W = n, A, K, K
        self.D = A if continuous else 1
        self.continuous, self.per_root = continuous, per_root
        # None: the reference seeds its sampler from the real clock.  An int: the clock is pinned for the duration of
        # prepare / backprop so that root i's expand() of expansion e (0 = prepare, s + 1 = simulation s) is seeded with
        # (unsigned)(clock_base + e * n + i) - the convention oracle/libstdcxx_rng.c and the CUDA parity sampler share
        self.clock_base = None
        self.L.ref_clock_pin.argtypes = [C.c_int, C.c_uint64]
        self.h = C.c_void_p(self.L.ref_new(n, A, K, int(continuous), int(per_root)))

    def _pin(self, expansion):
        if self.clock_base is not None:
            self.L.ref_clock_pin(1, C.c_uint64((int(self.clock_base) + int(expansion) * self.n) & 0xffffffffffffffff))

    def _unpin(self):
        if self.clock_base is not None:
            self.L.ref_clock_pin(0, C.c_uint64(0))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_free(self.h)
            self.h = None

    def set_delta(self, d):
        self.L.ref_set_delta(self.h, C.c_float(d))

    def set_rng(self, mode, seed, root_offset=0):
        assert root_offset == 0
        assert mode != 2 or self.per_root
        self.L.ref_rng_mode(int(mode), C.c_uint64(seed))

    def prepare(self, noise_w, noises, rewards, logits, to_play, values=None, sampled_actions=None):
        nz = None if noises is None else _f(np.asarray(noises, np.float32).reshape(self.n, self.K))
        rewards, logits, to_play = _f(rewards), _f(logits), _i(to_play)
        out = np.zeros((self.n, self.K, self.D), np.float32)
        self._pin(0)
        self.L.ref_prepare(self.h, int(nz is not None), C.c_float(noise_w), _ptr(nz, C.c_float), _ptr(rewards, C.c_float),
                           _ptr(logits, C.c_float), int(logits.shape[1]), _ptr(to_play, C.c_int), _ptr(out, C.c_float))
        self._unpin()
        return out

    def traverse(self, pb_c_base, pb_c_init, discount, to_play, num_simulations=0, max_considered=0, path_stride=0):
        n = self.n
        tp = _i(to_play)
        o = {k: np.zeros(n, np.int32) for k in ("ix", "iy", "search_len", "vtp")}
        laf = np.zeros((n, self.D), np.float32)
        self.L.ref_traverse(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), _ptr(tp, C.c_int),
                            _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int), _ptr(laf, C.c_float),
                            _ptr(o["search_len"], C.c_int), _ptr(o["vtp"], C.c_int))
        o["last_action_f"] = laf
        o["last_action"] = laf[:, 0].astype(np.int32)
        return o

    def backprop(self, latent_index, discount, rewards, values, logits, is_reset, to_play, sampled_actions=None):
        rewards, values, logits = _f(rewards), _f(values), _f(logits)
        is_reset, to_play = _i(is_reset), _i(to_play)
        out = np.zeros((self.n, self.K, self.D), np.float32)
        self._pin(latent_index)
        self.L.ref_backprop(self.h, int(latent_index), C.c_float(discount), _ptr(rewards, C.c_float),
                            _ptr(values, C.c_float), _ptr(logits, C.c_float), int(logits.shape[1]),
                            _ptr(is_reset, C.c_int), _ptr(to_play, C.c_int), _ptr(out, C.c_float))
        self._unpin()
        return out

    def distributions(self):
        d = np.zeros((self.n, self.K), np.int32)
        ln = np.zeros(self.n, np.int32)
        self.L.ref_distributions(self.h, _ptr(d, C.c_int), _ptr(ln, C.c_int))
        return d, ln

    def values(self):
        v = np.zeros(self.n, np.float32)
        self.L.ref_values(self.h, _ptr(v, C.c_float))
        return v

    def minmax(self):
        mn, mx = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.ref_minmax(self.h, _ptr(mn, C.c_float), _ptr(mx, C.c_float))
        return mn, mx

    def root_stats(self):
        vc, vs, rw = np.zeros(self.n, np.int32), np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.ref_root_stats(self.h, _ptr(vc, C.c_int), _ptr(vs, C.c_float), _ptr(rw, C.c_float))
        return vc, vs, rw

    def dump_tree(self, i, E):
        return _dump(lambda *a: self.L.ref_dump_tree(self.h, *a), i, E, self.K)


class RefStochasticForest(_ForestBase):
    """The compiled Stochastic-MuZero reference tree (oracle/ref_harness_stochastic.cpp).  The reference draws its
    chance outcomes itself (std::random_device); ``traverse_chance`` reports the action / outcome taken at every level
    in out["paths"] so that the oracle and the CUDA path can be handed the same outcomes.  Every logits array is
    [n, W], W = max(A, C)."""

    variant = "smz"

    def __init__(self, n, A, C_, legal_actions=None, per_root=False):
        self.L = ref_lib("smz")
        if self.L is None:
            raise RuntimeError("oracle/_ref/libref_smz.so is not built and /root/reference is absent")
        self.n, self.A, self.K, self.W = n, A, C_, max(A, C_)
        if legal_actions is None:
            legal_actions = [[] for _ in range(n)]
        flat, off = _csr(legal_actions)
        self.h = C.c_void_p(self.L.ref_new(n, A, C_, _ptr(flat, C.c_int), _ptr(off, C.c_int), int(per_root)))
        self.per_root = per_root

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_free(self.h)
            self.h = None

    def set_delta(self, d):
        self.L.ref_set_delta(self.h, C.c_float(d))

    def set_rng(self, mode, seed, root_offset=0):
        assert root_offset == 0
        assert mode != 2 or self.per_root
        self.L.ref_rng_mode(int(mode), C.c_uint64(seed))

    def prepare(self, noise_w, noises, rewards, logits, to_play, values=None, sampled_actions=None):
        with_noise = noises is not None
        nf, no = _csr_f(noises) if with_noise else (None, None)
        logits = _f(np.asarray(logits, np.float32)[:, :self.A])  # the reference takes [n, A] policy logits at the roots
        rewards, to_play = _f(rewards), _i(to_play)
        self.L.ref_prepare(self.h, int(with_noise), C.c_float(noise_w), _ptr(nf, C.c_float), _ptr(no, C.c_int),
                           _ptr(rewards, C.c_float), _ptr(logits, C.c_float), _ptr(to_play, C.c_int))

    def traverse_chance(self, pb_c_base, pb_c_init, discount, to_play, forced=None, path_stride=0):
        assert forced is None, "the compiled reference draws its own outcomes"
        n = self.n
        tp = _i(to_play)
        o = {k: np.zeros(n, np.int32) for k in ("ix", "iy", "last_action", "search_len", "vtp", "leaf_is_chance")}
        paths = np.full((n, max(path_stride, 1)), -1, np.int32)
        self.L.ref_traverse(self.h, int(pb_c_base), C.c_float(pb_c_init), C.c_float(discount), _ptr(tp, C.c_int),
                            _ptr(o["ix"], C.c_int), _ptr(o["iy"], C.c_int), _ptr(o["last_action"], C.c_int),
                            _ptr(o["search_len"], C.c_int), _ptr(o["vtp"], C.c_int), _ptr(o["leaf_is_chance"], C.c_int),
                            _ptr(paths, C.c_int) if path_stride else None, int(path_stride))
        if path_stride:
            o["paths"] = paths
        return o

    def backprop(self, latent_index, discount, rewards, values, logits, is_reset, to_play, sampled_actions=None):
        rewards, values, logits, to_play = _f(rewards), _f(values), _f(logits), _i(to_play)
        assert logits.shape[1] == self.W
        self.L.ref_backprop(self.h, int(latent_index), C.c_float(discount), _ptr(rewards, C.c_float),
                            _ptr(values, C.c_float), _ptr(logits, C.c_float), int(self.W), _ptr(to_play, C.c_int))

    def discrete(self, probs, w0, w1):
        """libstdc++'s std::discrete_distribution<>(probs) under an engine that returns w0, then w1."""
        p = np.ascontiguousarray(probs, np.float64)
        return int(self.L.ref_discrete(_ptr(p, C.c_double), int(p.size), C.c_uint32(w0), C.c_uint32(w1)))

    def distributions(self):
        d = np.zeros((self.n, self.W), np.int32)
        ln = np.zeros(self.n, np.int32)
        self.L.ref_distributions(self.h, _ptr(d, C.c_int), _ptr(ln, C.c_int))
        return d, ln

    def values(self):
        v = np.zeros(self.n, np.float32)
        self.L.ref_values(self.h, _ptr(v, C.c_float))
        return v

    def minmax(self):
        mn, mx = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.ref_minmax(self.h, _ptr(mn, C.c_float), _ptr(mx, C.c_float))
        return mn, mx

    def root_stats(self):
        vc, vs, rw = np.zeros(self.n, np.int32), np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self.L.ref_root_stats(self.h, _ptr(vc, C.c_int), _ptr(vs, C.c_float), _ptr(rw, C.c_float))
        return vc, vs, rw

    def dump_tree(self, i, E):
        return _dump(lambda *a: self.L.ref_dump_tree(self.h, *a), i, E, self.W)
