"""Pins the CPU oracle (oracle/mcts_oracle.c) against the compiled, unmodified reference ctree
(oracle/_ref, built from /root/reference by oracle/build.py) and against the committed golden fixtures
(tests/golden/*.npz, generated from the compiled reference by tests/golden/make_golden.py).

Everything is compared bit-for-bit: leaf indices, last actions, search paths, visit counts, and the raw
fp32 bits of root values, MinMax bounds, priors and value sums.
"""
import glob
import os

import numpy as np
import pytest

from oracle.binding import OracleForest, RefForest, RefStochasticForest, ref_available
from tests.common import TICTACTOE_MASKS, Scenario, assert_same, run_cpu, run_sez_reference

HERE = os.path.dirname(os.path.abspath(__file__))
LEGAL9 = [[a for a, x in enumerate(row) if x == 1] for row in TICTACTOE_MASKS]

CASES = {
    # name: (scenario kwargs, rng_mode)
    "mz_cartpole": (dict(variant="mz", n=8, A=2, S=25, seed=1), 2),
    "mz_a18_batch": (dict(variant="mz", n=64, A=18, S=50, seed=2), 1),
    "mz_a18_per_root": (dict(variant="mz", n=32, A=18, S=50, seed=3), 2),
    "mz_zero_model": (dict(variant="mz", n=16, A=9, S=40, seed=4, zero_model=True), 2),
    "mz_masks": (dict(variant="mz", n=16, A=9, S=50, seed=5, legal=LEGAL9), 2),
    "mz_two_player": (dict(variant="mz", n=32, A=9, S=25, seed=6, to_play=[1, 2] * 16, discount=1.0), 2),
    "mz_to_play_zero": (dict(variant="mz", n=16, A=9, S=25, seed=7, to_play=[0] * 16), 2),
    "mz_wide": (dict(variant="mz", n=8, A=80, S=60, seed=8), 2),
    "ez_pong": (dict(variant="ez", n=32, A=6, S=50, seed=9), 2),
    "ez_two_player": (dict(variant="ez", n=32, A=9, S=25, seed=10, to_play=[2, 1] * 16, discount=1.0), 1),
    "gmz_tictactoe": (dict(variant="gmz", n=16, A=9, S=30, seed=11, legal=LEGAL9, discount=1.0, max_considered=3), 1),
    "gmz_gomoku": (dict(variant="gmz", n=32, A=36, S=50, seed=12, discount=1.0, max_considered=6), 1),
    "gmz_m16": (dict(variant="gmz", n=16, A=18, S=50, seed=13, max_considered=16), 1),
    # the evaluation path: Roots.prepare_no_noise (lzero/policy/muzero.py:831; cnode.cpp:340-358); these fixtures also
    # carry CRoots::get_trajectories (cnode.cpp:237-257, 369-385)
    "mz_eval_no_noise": (dict(variant="mz", n=32, A=18, S=50, seed=41, no_noise=True, trajectories=True), 2),
    "mz_eval_no_noise_masks_two_player": (dict(variant="mz", n=16, A=9, S=40, seed=42, legal=LEGAL9, to_play=[1, 2] * 8,
                                               discount=1.0, no_noise=True, trajectories=True), 1),
    "ez_eval_no_noise": (dict(variant="ez", n=32, A=6, S=50, seed=43, no_noise=True, trajectories=True), 2),
    "gmz_eval_no_noise": (dict(variant="gmz", n=16, A=9, S=30, seed=44, legal=LEGAL9, discount=1.0, max_considered=3,
                               no_noise=True, trajectories=True), 1),
    # ReZero search_with_reuse (mcts_ctree.py:323-420, 829-953)
    "mz_reuse_a6": (dict(variant="mz", n=32, A=6, S=50, seed=14, reuse=True), 2),
    "mz_reuse_batch": (dict(variant="mz", n=16, A=18, S=50, seed=15, reuse=True), 1),
    "mz_reuse_masks_two_player": (dict(variant="mz", n=16, A=9, S=30, seed=16, legal=LEGAL9, to_play=[1, 2] * 8,
                                       discount=1.0, reuse=True), 2),
    "ez_reuse_a6": (dict(variant="ez", n=32, A=6, S=50, seed=17, reuse=True), 2),
    "ez_reuse_two_player": (dict(variant="ez", n=16, A=9, S=30, seed=18, to_play=[2, 1] * 8, discount=1.0,
                                 reuse=True), 1),
    # Sampled EfficientZero: the reference samples each node's K actions itself; its tables are recorded and handed to
    # the oracle (tests/common.py:run_sez_reference), so the whole tree logic downstream of the sampler is compared
    "sez_continuous_k20_d6": (dict(variant="sez", n=8, A=12, S=50, seed=19, K=20, D=6), 2),
    "sez_continuous_batch_rng": (dict(variant="sez", n=16, A=4, S=40, seed=20, K=8, D=2), 1),
    "sez_continuous_two_player": (dict(variant="sez", n=8, A=6, S=30, seed=21, K=6, D=3, to_play=[1, 2] * 4,
                                       discount=1.0), 2),
    "sez_discrete_k4_of_6": (dict(variant="sez", n=16, A=6, S=40, seed=22, K=4, D=1), 2),
    # Stochastic MuZero (K = chance_space_size): the reference draws its chance outcomes itself; the outcomes it took
    # are recorded per level ("paths") and handed to the oracle, so everything but the draw is compared
    "smz_2048_a4_c32": (dict(variant="smz", n=16, A=4, S=50, seed=23, K=32), 2),
    "smz_a6_c2_batch": (dict(variant="smz", n=32, A=6, S=50, seed=24, K=2), 1),
    "smz_masks_two_player": (dict(variant="smz", n=16, A=9, S=40, seed=25, K=5, legal=LEGAL9, to_play=[1, 2] * 8,
                                  discount=1.0), 2),
}
SEZ_DISCRETE = {"sez_discrete_k4_of_6"}


def sez_continuous(name):
    return name not in SEZ_DISCRETE


def search_summary(forest, sc, rec, trees=3):
    """Everything a parity check looks at, as a flat dict of numpy arrays."""
    out = {}
    keys = ("ix", "iy", "last_action_f", "search_len", "vtp", "is_reset") if sc.variant == "sez" else \
        ("ix", "iy", "last_action", "search_len", "vtp", "paths", "leaf_is_chance")
    for k in keys:
        if k in rec[0]:  # the ReZero traverse reports no paths
            out[k] = np.stack([r[k] for r in rec])
    if sc.variant == "sez":  # the tables the reference sampled: part of the fixture, injected everywhere else
        out["sez_root_actions"], out["sez_actions"] = sc.root_actions, sc.actions
    if getattr(sc, "reuse", False):
        out["infer_stats"] = np.asarray(forest.infer_stats, np.float64)
    d, ln = forest.distributions()
    out["dist"], out["dist_len"] = d, ln
    out["values"] = forest.values()
    out["mm_min"], out["mm_max"] = forest.minmax()
    vc, vs, rw = forest.root_stats()
    out["root_visits"], out["root_value_sum"] = vc, vs
    for i in range(min(trees, sc.n)):
        t = forest.dump_tree(i, sc.S + 1)
        for k in ("visit", "value_sum", "prior", "reward", "child_e", "hdr"):
            if k == "prior" and sc.variant == "sez":
                continue  # log-prob priors never reach selection (uniform branch, sez cnode.cpp:1054-1079): not restated
            out["tree%d_%s" % (i, k)] = t[k]
    if sc.variant == "gmz":
        out["gmz_policies"], out["gmz_child_values"] = forest.gumbel_outputs(sc.discount)
    if getattr(sc, "trajectories", False):
        out["trajectories"], out["trajectory_len"] = forest.trajectories(sc.S + 1)
    return out


def oracle_summary(name, tables=None, forced=None):
    """tables = (root_actions, actions): the Sampled-EfficientZero action tables to inject (from the compiled reference
    or from its golden fixture); forced = [S, n, S+1] Stochastic-MuZero path tables whose chance levels are replayed."""
    kw, mode = CASES[name]
    sc = Scenario(**kw)
    sc.forced = forced
    if tables is not None:
        sc.root_actions, sc.actions = np.asarray(tables[0], np.float32), np.asarray(tables[1], np.float32)
    rec = []
    f = run_cpu(OracleForest(sc.variant, sc.n, sc.A, sc.legal, K=sc.K, D=sc.D), sc, mode, 2024, record=rec)
    return sc, search_summary(f, sc, rec)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_golden_fixture(name):
    path = os.path.join(HERE, "golden", name + ".npz")
    assert os.path.exists(path), "missing fixture %s (run tests/golden/make_golden.py where /root/reference exists)" % path
    want = np.load(path)
    tables = (want["sez_root_actions"], want["sez_actions"]) if name.startswith("sez") else None
    sc, got = oracle_summary(name, tables, forced=want["paths"] if name.startswith("smz") else None)
    assert sorted(want.files) == sorted(got)
    for k in want.files:
        assert_same(got[k], want[k], "%s/%s" % (name, k))
    assert (got["dist"].clip(min=0).sum(1) == sc.S).all()
    assert (got["root_visits"] == sc.S + 1).all()


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_compiled_reference(name):
    kw, mode = CASES[name]
    if not ref_available(kw["variant"]):
        pytest.skip("oracle/_ref not built and /root/reference absent")
    sc = Scenario(**kw)
    rec = []
    if sc.variant == "sez":
        ref, rec = run_sez_reference(sc, mode, 2024, continuous=sez_continuous(name))
        want = search_summary(ref, sc, rec)
        _, got = oracle_summary(name, (sc.root_actions, sc.actions))
        assert (np.abs(sc.root_actions) <= 1.0).all() if sez_continuous(name) else (sc.root_actions == np.round(sc.root_actions)).all()
    elif sc.variant == "smz":
        ref = run_cpu(RefStochasticForest(sc.n, sc.A, sc.K, sc.legal, per_root=(mode == 2)), sc, mode, 2024, record=rec)
        want = search_summary(ref, sc, rec)
        _, got = oracle_summary(name, forced=want["paths"])
        assert want["leaf_is_chance"].any() and not want["leaf_is_chance"].all()
    else:
        ref = run_cpu(RefForest(sc.variant, sc.n, sc.A, sc.legal, per_root=(mode == 2)), sc, mode, 2024, record=rec)
        want = search_summary(ref, sc, rec)
        _, got = oracle_summary(name)
    assert sorted(want) == sorted(got)
    for k in want:
        assert_same(got[k], want[k], "%s/%s" % (name, k))


def test_zero_model_values_are_exactly_zero():
    _, got = oracle_summary("mz_zero_model")
    assert (got["values"] == 0.0).all()


def test_distribution_length_equals_legal_count():  # reference test_mcts_ctree.py:265-270
    sc, got = oracle_summary("mz_masks")
    assert got["dist_len"].tolist() == [len(l) for l in LEGAL9]
    for i, l in enumerate(LEGAL9):
        assert (got["dist"][i, len(l):] == -1).all()


def test_gumbel_constants_match_reference():
    if not ref_available("gmz"):
        pytest.skip("oracle/_ref not built and /root/reference absent")
    import ctypes as C
    from oracle.binding import oracle_lib, ref_lib
    O, R = oracle_lib(), ref_lib("gmz")
    for m, S in ((3, 30), (6, 50), (16, 50), (4, 200), (1, 10), (16, 8)):
        a = np.zeros((m + 1, S), np.int32)
        b = np.zeros((m + 1, S), np.int32)
        O.orc_gumbel_table(m, S, a.ctypes.data_as(C.c_void_p))
        R.ref_gumbel_table(m, S, b.ctypes.data_as(C.c_void_p))
        assert_same(a, b, "considered-visit table m=%d S=%d" % (m, S))
    a = np.zeros(700, np.float32)  # > 624 words: crosses an mt19937 state refill
    b = np.zeros(700, np.float32)
    O.orc_gumbel_noise(C.c_float(10.0), 700, a.ctypes.data_as(C.c_void_p))
    R.ref_gumbel_noise(C.c_float(10.0), C.c_float(0.0), 700, b.ctypes.data_as(C.c_void_p))
    assert_same(a, b, "gumbel noise table")


def test_chance_draw_arithmetic_matches_libstdcxx():
    """orc_discrete_draw (what the oracle and the CUDA kernels do at a chance node for a uniform u) against libstdc++'s
    std::discrete_distribution driven by an engine scripted to produce exactly that u (ref_harness_stochastic.cpp)."""
    if not ref_available("smz"):
        pytest.skip("oracle/_ref not built and /root/reference absent")
    import ctypes as C
    from oracle.binding import oracle_lib
    O = oracle_lib()
    ref = RefStochasticForest(1, 2, 2)
    rng = np.random.default_rng(5)
    for n in (1, 2, 3, 5, 32, 100):
        for _ in range(40):
            p = rng.random(n).astype(np.float32).astype(np.float64) ** rng.integers(1, 6)  # float priors widened, some tiny
            for _ in range(8):
                m = int(rng.integers(0, 1 << 53))  # u = m / 2^53, the engine words that generate_canonical turns into it
                got = O.orc_discrete_draw(p.ctypes.data_as(C.POINTER(C.c_double)), n, C.c_double(m / float(1 << 53)))
                want = ref.discrete(p, (m << 11) & 0xffffffff, m >> 21)
                assert got == want, (n, m, got, want)
    # exact ties with a partial sum: lower_bound takes the first index whose cumulative probability is >= u
    p = np.asarray([0.25, 0.25, 0.5])
    for m, want in ((1 << 51, 0), ((1 << 51) + 1, 1), (1 << 52, 1), ((1 << 52) + 1, 2)):
        assert O.orc_discrete_draw(p.ctypes.data_as(C.POINTER(C.c_double)), 3, C.c_double(m / float(1 << 53))) == want
        assert ref.discrete(p, (m << 11) & 0xffffffff, m >> 21) == want


def test_chance_draws_follow_the_priors():
    """Unforced oracle search: outcome frequencies at the roots' first chance level follow that node's priors."""
    kw, mode = CASES["smz_a6_c2_batch"]
    sc = Scenario(**dict(kw, n=256, S=12, K=3))
    f = run_cpu(OracleForest("smz", sc.n, sc.A, sc.legal, K=3), sc, 2, 99)
    seen = np.zeros(3)
    want = np.zeros(3)
    for i in range(sc.n):
        t = f.dump_tree(i, sc.S + 1)
        for e in range(1, sc.S + 1):
            if t["hdr"][e, 2] == 1 and t["hdr"][e, 3] == 0:  # a chance node directly below the root
                v = t["visit"][e, :3].clip(min=0)
                seen += v
                want += v.sum() * t["prior"][e, :3]
    assert seen.sum() > 500
    assert np.abs(seen - want).max() < 4 * np.sqrt(seen.sum())


def test_no_stray_fixtures():
    have = {os.path.basename(p)[:-4] for p in glob.glob(os.path.join(HERE, "golden", "*.npz"))}
    assert have >= set(CASES)
