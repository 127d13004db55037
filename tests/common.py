"""Shared scenario generator for the parity tests (seeded synthetic 'model' outputs)."""
import numpy as np

PB_C_BASE, PB_C_INIT = 19652, 1.25
TICTACTOE_MASKS = [  # 16 x 9 action mask of lzero/mcts/tests/test_mcts_ctree.py:120-137
    [0, 0, 0, 1, 0, 1, 1, 0, 0], [1, 0, 0, 1, 0, 0, 1, 0, 0], [1, 1, 0, 0, 0, 1, 0, 0, 1], [0, 0, 1, 0, 0, 1, 0, 0, 1],
    [1, 0, 0, 1, 1, 0, 0, 0, 1], [1, 1, 0, 0, 0, 0, 1, 0, 1], [1, 0, 0, 1, 1, 0, 0, 0, 1], [1, 1, 0, 0, 0, 0, 1, 1, 0],
    [1, 0, 1, 0, 0, 0, 1, 0, 1], [0, 1, 0, 0, 1, 1, 0, 0, 1], [1, 1, 0, 1, 0, 0, 0, 1, 0], [0, 0, 1, 0, 0, 1, 1, 1, 0],
    [1, 0, 0, 1, 0, 0, 1, 0, 1], [0, 0, 1, 1, 0, 1, 0, 0, 1], [1, 0, 1, 0, 1, 0, 0, 1, 0], [0, 1, 0, 1, 0, 0, 1, 0, 1],
]


class Scenario:
    """Seeded inputs of one search: root logits / rewards / values / noise and per-simulation model outputs."""

    def __init__(self, variant, n, A, S, seed=0, legal=None, to_play=None, discount=0.997, max_considered=4,
                 reward_scale=0.1, zero_model=False, K=0, D=0, reuse=False, no_noise=False, trajectories=False):
        rng = np.random.default_rng(seed)
        self.variant, self.n, self.A, self.S, self.discount, self.m = variant, n, A, S, discount, max_considered
        self.K, self.D = K, D
        self.no_noise = no_noise          # the evaluation path: Roots.prepare_no_noise (lzero/policy/muzero.py:831)
        self.trajectories = trajectories  # the fixture also carries get_trajectories()
        W = self.W = max(A, K) if variant == "smz" else A  # smz: K = chance_space_size, rows hold policy OR chance logits
        self.legal = legal if legal is not None else [list(range(A)) for _ in range(n)]
        self.to_play = np.asarray(to_play if to_play is not None else [-1] * n, np.int32)
        z = 0.0 if zero_model else 1.0
        self.root_logits = (rng.standard_normal((n, W)) * z).astype(np.float32)
        self.root_reward = (rng.standard_normal(n) * reward_scale * z).astype(np.float32)
        self.root_value = (rng.standard_normal(n) * z).astype(np.float32)
        self.noises = [rng.dirichlet([0.3] * len(l)).astype(np.float32) for l in self.legal]
        self.noise_pad = np.zeros((n, W), np.float32)
        for i, nz in enumerate(self.noises):
            self.noise_pad[i, :len(nz)] = nz
        self.logits = (rng.standard_normal((S, n, W)) * z).astype(np.float32)
        self.reward = (rng.standard_normal((S, n)) * reward_scale * z).astype(np.float32)
        if variant in ("ez", "sez"):  # value prefixes accumulate along a path in a real model; any floats do here
            self.reward = (rng.standard_normal((S, n)) * z).astype(np.float32)
        self.value = (rng.standard_normal((S, n)) * z).astype(np.float32)
        if variant == "sez":
            self.root_actions = np.tanh(rng.standard_normal((n, K, D))).astype(np.float32)
            self.actions = np.tanh(rng.standard_normal((S, n, K, D))).astype(np.float32)

        self.reuse = reuse
        if reuse:  # ReZero: the action the trajectory took at each root and the value searched at the next state
            self.true_action = np.asarray([l[rng.integers(len(l))] for l in self.legal], np.int32)
            self.reuse_value = (rng.standard_normal(n) * z).astype(np.float32)

        self.forced = None  # smz: per simulation, the [n, S+1] outcome tables to take at the chance levels

    def model_fn_chance(self, s, ix, iy, last_action, leaf_is_chance):
        return dict(reward=self.reward[s], value=self.value[s], logits=self.logits[s])

    def model_fn(self, s, ix, iy, last_action):
        out = dict(reward=self.reward[s], value=self.value[s], logits=self.logits[s])
        if self.variant == "sez":
            out["sampled_actions"] = self.actions[s]
        return out


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_same(a, b, what):
    x, y = bits(np.asarray(a)), bits(np.asarray(b))
    assert x.shape == y.shape, "%s: shape %s vs %s" % (what, x.shape, y.shape)
    if not np.array_equal(x, y):
        bad = np.argwhere(x != y)
        raise AssertionError("%s: %d mismatches, first at %s: %r vs %r" % (
            what, len(bad), bad[0], np.asarray(a)[tuple(bad[0])], np.asarray(b)[tuple(bad[0])]))


def run_cpu(forest, sc, rng_mode, seed, record=None, horizon=5):
    """Drive an OracleForest / RefForest through one full search of scenario ``sc``."""
    forest.set_delta(0.01)
    forest.set_rng(rng_mode, seed)
    kw = {}
    if sc.variant == "sez":
        kw["sampled_actions"] = sc.root_actions
    with_noise = sc.variant != "sez" and not getattr(sc, "no_noise", False)
    forest.prepare(0.25 if with_noise or sc.variant == "sez" else 0.0, sc.noises if with_noise else None, sc.root_reward,
                   sc.root_logits, sc.to_play, values=sc.root_value, **kw)
    if sc.variant == "smz":
        forest.run_search_chance(sc.model_fn_chance, sc.S, PB_C_BASE, PB_C_INIT, sc.discount, sc.to_play,
                                 forced=None if forest.__class__.__name__.startswith("Ref") else sc.forced, record=record)
        return forest
    if getattr(sc, "reuse", False):
        forest.infer_stats = forest.run_search_reuse(sc.model_fn, sc.S, PB_C_BASE, PB_C_INIT, sc.discount, sc.to_play,
                                                     sc.true_action, sc.reuse_value, lstm_horizon_len=horizon,
                                                     record=record)
        return forest
    forest.run_search(sc.model_fn, sc.S, PB_C_BASE, PB_C_INIT, sc.discount, sc.to_play, max_considered=sc.m,
                      lstm_horizon_len=horizon, record=record)
    return forest


def sez_logits(x, continuous):
    """Policy head output the Sampled-EfficientZero reference can sample from: continuous heads are {mu[D], sigma[D]}
    and sigma must be positive (std::normal_distribution, sez cnode.cpp:262-266)."""
    x = np.array(x, np.float32, copy=True)
    if continuous:
        D = x.shape[-1] // 2
        x[..., D:] = np.abs(x[..., D:]) * 0.5 + 0.1
    return x


def run_sez_reference(sc, rng_mode, seed, continuous=True, horizon=5, clock_base=None):
    """Drive the compiled Sampled-EfficientZero reference through one search of scenario ``sc`` and RECORD the action
    tables it samples into ``sc.root_actions`` / ``sc.actions`` (what the oracle and the CUDA path are then handed).
    Returns (forest, per-simulation records)."""
    from oracle.binding import RefSampledForest
    env_actions = sc.D if continuous else sc.A
    ref = RefSampledForest(sc.n, env_actions, sc.K, continuous=continuous, per_root=(rng_mode == 2))
    ref.clock_base = clock_base  # not None: the sampler's wall-clock seed is pinned (oracle/binding.py)
    ref.set_delta(0.01)
    ref.set_rng(rng_mode, seed)
    sc.root_actions = ref.prepare(0.25, None, sc.root_reward, sez_logits(sc.root_logits, continuous), sc.to_play)
    sc.actions = np.zeros((sc.S,) + sc.root_actions.shape, np.float32)
    rec = []
    tp = np.asarray(sc.to_play, np.int32)
    for s in range(sc.S):
        out = ref.traverse(PB_C_BASE, PB_C_INIT, sc.discount, tp)
        is_reset = (out["search_len"] % horizon == 0).astype(np.int32)
        rec.append(dict(out, is_reset=is_reset))
        sc.actions[s] = ref.backprop(s + 1, sc.discount, sc.reward[s], sc.value[s], sez_logits(sc.logits[s], continuous),
                                     is_reset, out["vtp"])
    return ref, rec
