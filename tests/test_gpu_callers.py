"""GPU: the caller-shaped device paths (lightzero_b200/callers.py) at the reanalyze shape - 256 segments x (5 unroll
steps + 1) = 1,536 roots - against the reference's own loops restated in oracle/reanalyze_oracle.py
(lzero/mcts/buffer/game_buffer_muzero.py:463-513, :611-656) fed with the CPU checker's search results, and the collect /
eval step (lzero/policy/muzero.py:655-725, :827-850) against oracle/collect_oracle.py.

Bars: visit counts, actions and float32 policy targets bit-exact; float64 value / reward targets bit-exact (the kernel
does the reference's float64 operations in the reference's order, discount ** k from the host's pow)."""
import numpy as np
import pytest

from oracle import collect_oracle as co
from oracle import reanalyze_oracle as ro
from tests.common import PB_C_BASE, PB_C_INIT, Scenario, assert_same
from tests.test_gpu_api import ScenarioModel, identity

pytestmark = pytest.mark.gpu
SEED = 515


def _masks(rng, n, A):
    m = (rng.random((n, A)) < 0.6).astype(np.int8)
    m[np.arange(n), rng.integers(0, A, n)] = 1
    return m


def _oracle_search(sc, noise=False):
    from oracle.binding import OracleForest
    orc = OracleForest(sc.variant, sc.n, sc.A, sc.legal)
    orc.set_delta(0.01)
    orc.set_rng(2, SEED)
    orc.prepare(0.25 if noise else 0.0, sc.noises if noise else None, sc.root_reward, sc.root_logits, sc.to_play, values=sc.root_value)
    orc.run_search(sc.model_fn, sc.S, PB_C_BASE, PB_C_INIT, sc.discount, sc.to_play, max_considered=sc.m)
    return orc


def _driver(sc, cls, graph=False, **extra):
    cfg = dict(num_simulations=sc.S, pb_c_base=PB_C_BASE, pb_c_init=PB_C_INIT, discount_factor=sc.discount, value_delta_max=0.01,
               device="cuda", tie_break_seed=SEED, lstm_horizon_len=5, max_num_considered_actions=sc.m, cuda_graph=graph,
               root_noise_weight=0.25, root_dirichlet_alpha=0.3, reanalyze_noise=False,
               model=dict(support_scale=300, categorical_distribution=False))
    cfg.update(extra)
    return identity(cls(cfg))


@pytest.mark.parametrize("graph", [False, True])
@pytest.mark.parametrize("fixed", [True, False])
def test_reanalyze_policy_targets_at_the_reanalyze_shape(fixed, graph):
    import torch
    import lightzero_b200 as lzb
    from lightzero_b200 import callers
    B, U, S = 256, 5, 50
    A = 6 if fixed else 9
    T = B * (U + 1)
    rng = np.random.default_rng(7 + A)
    mask = np.ones((T, A), np.int8) if fixed else _masks(rng, T, A)
    legal = co.legal_from_mask(mask)
    to_play = [-1] * T if fixed else [1, 2] * (T // 2)
    sc = Scenario("mz", T, A, S, seed=31 + A, legal=legal, to_play=to_play, discount=0.997 if fixed else 1.0)
    policy_mask = (rng.random(T) < 0.8).astype(np.int32)
    orc = _oracle_search(sc)
    dist, ln = orc.distributions()
    want, want_child = ro.policy_targets([dist[i, :ln[i]].tolist() for i in range(T)], legal, policy_mask.tolist(), B, U, A,
                                         fixed_action_space=fixed)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(a).to(dev)
    mcts = _driver(sc, lzb.MuZeroMCTSCtree, graph=graph)
    model = ScenarioModel(sc, dev)
    lat0 = torch.zeros(T, 16, device=dev)
    for _ in range(2 if graph else 1):
        out = callers.reanalyze_policy_targets(mcts, model, lat0, t(sc.root_reward), t(sc.root_logits), t(mask), t(sc.to_play),
                                               t(policy_mask), U, fixed_action_space=fixed)
    assert tuple(out["target_policies"].shape) == (B, U + 1, A)
    assert_same(out["target_policies"].cpu().numpy(), np.asarray(want, np.float64).astype(np.float32), "target policies")
    assert_same(out["root_values"].cpu().numpy(), orc.values(), "searched root values")
    child = out["child_visits"].cpu().numpy()
    for k, row in want_child.items():
        assert_same(child[k, :len(row)], np.asarray(row, np.float64).astype(np.float32), "child_visits[%d]" % k)
    assert (child[policy_mask == 0] == 0).all()


@pytest.mark.parametrize("board_games", [False, True])
def test_reanalyze_value_targets_match_the_reference_loops(board_games):
    import torch
    from lightzero_b200 import callers
    rng = np.random.default_rng(11 + board_games)
    B, U, gamma = 256, 5, 0.997 if not board_games else 1.0
    T = B * (U + 1)
    seg_len = rng.integers(3, 40, B)
    reward_len = seg_len + rng.integers(0, 12, B)            # reward lists run past the segment (padding of the next one)
    pos = np.asarray([rng.integers(0, l + 3) for l in seg_len])  # some windows reach past the segment's end
    rewards = [(rng.standard_normal(l) * (1.0 if not board_games else 0.5)).astype(np.float32) for l in reward_len]
    to_play = [[1 + (k % 2) for k in range(l + 8)] for l in reward_len]
    td = rng.integers(0, 6, T).astype(np.int32)
    vmask = (rng.random(T) < 0.85).astype(np.float32)
    roots_values = rng.standard_normal(T).astype(np.float32)
    want_r, want_v = ro.value_targets(roots_values.tolist(), td.tolist(), vmask.tolist(), rewards, seg_len.tolist(), pos.tolist(),
                                      to_play, U, gamma, board_games=board_games)
    Lr, Lt = int(reward_len.max()), max(len(x) for x in to_play)
    rew_pad = np.zeros((B, Lr), np.float32)
    tp_pad = np.zeros((B, Lt), np.int32)
    for b in range(B):
        rew_pad[b, :reward_len[b]] = rewards[b]
        tp_pad[b, :len(to_play[b])] = to_play[b]
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(a).to(dev)
    tr, tv = callers.reanalyze_value_targets(t(roots_values), td, t(vmask), t(rew_pad), reward_len, seg_len, pos, U, gamma,
                                             to_play_segments=tp_pad if board_games else None, board_games=board_games)
    got_v, got_r = tv.cpu().numpy(), tr.cpu().numpy()
    assert got_v.dtype == np.float64 and got_v.shape == (B, U + 1)
    wv, wr = np.asarray(want_v, np.float64), np.asarray(want_r, np.float64)
    assert np.array_equal(got_v.view(np.uint64), wv.view(np.uint64)), "value targets: %d of %d differ, max |d| %g" % (
        (got_v != wv).sum(), wv.size, np.abs(got_v - wv).max())
    assert np.array_equal(got_r.view(np.uint64), wr.view(np.uint64)), "reward targets differ"


@pytest.mark.parametrize("variant", ["mz", "ez", "gmz"])
def test_collect_and_eval_step_match_the_oracle(variant):
    import torch
    import lightzero_b200 as lzb
    from lightzero_b200 import callers
    N, A, S = 96, 9, 30
    rng = np.random.default_rng(5)
    mask = _masks(rng, N, A)
    legal = co.legal_from_mask(mask)
    sc = Scenario(variant, N, A, S, seed=77, legal=legal, max_considered=4)
    cls = {"mz": lzb.MuZeroMCTSCtree, "ez": lzb.EfficientZeroMCTSCtree, "gmz": lzb.GumbelMuZeroMCTSCtree}[variant]
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(a).to(dev)
    mcts = _driver(sc, cls)
    model = ScenarioModel(sc, dev, hidden=8 if variant == "ez" else 0)
    lat0 = torch.zeros(N, 16, device=dev)
    hid = (torch.zeros(1, N, 8, device=dev), torch.zeros(1, N, 8, device=dev)) if variant == "ez" else None
    # evaluation: no noise, arg-max of the visit counts (muzero.py:827-850)
    out = callers.collect_search(mcts, model, lat0, t(sc.root_reward), t(sc.root_logits), t(mask), t(sc.to_play), eval_mode=True,
                                 root_values=t(sc.root_value), hidden_state_roots=hid)
    orc = _oracle_search(sc)
    dist, ln = orc.distributions()
    assert_same(out["legal_count"].cpu().numpy(), ln, "legal_count")
    assert_same(out["visit_counts"].cpu().numpy(), np.where(dist < 0, 0, dist), "visit counts")
    assert_same(out["searched_value"].cpu().numpy(), orc.values(), "searched values")
    want = [co.select_action(dist[i, :ln[i]].tolist(), 1.0, True) for i in range(N)]
    assert out["action_pos"].cpu().tolist() == [w[0] for w in want]
    assert out["action"].cpu().tolist() == [legal[i][w[0]] for i, w in enumerate(want)]
    np.testing.assert_allclose(out["entropy"].cpu().numpy(), [w[1] for w in want], rtol=1e-12, atol=1e-15)
    # collect: Dirichlet noise drawn on the device, temperature sampling with injected uniforms
    u = rng.random(N)
    out = callers.collect_search(mcts, model, lat0, t(sc.root_reward), t(sc.root_logits), t(mask), t(sc.to_play), temperature=0.5,
                                 root_values=t(sc.root_value), hidden_state_roots=hid, seed=3, uniforms=t(u))
    v = out["visit_counts"].cpu().numpy()
    assert (v.sum(1) == S).all()
    act, posn = out["action"].cpu().numpy(), out["action_pos"].cpu().numpy()
    for i in range(N):
        assert act[i] == legal[i][posn[i]]
        assert posn[i] == co.select_action(v[i, :ln[i]].tolist(), 0.5, False, uniform=u[i])[0]
    assert out["roots"].get_distributions()[3] == v[3, :ln[3]].tolist()


@pytest.mark.parametrize("name", ["value_single_player", "value_board_games"])
def test_value_targets_kernel_matches_the_reference_golden(name):
    """lzb_value_targets against tests/golden/reanalyze_targets.npz: the outputs of the reference's OWN source lines
    (game_buffer_muzero.py:464-512, executed by tests/golden/make_reanalyze_golden.py), float64 bit for bit."""
    import torch
    from lightzero_b200 import callers
    from tests.test_reanalyze_oracle import value_inputs
    g, rewards, to_play = value_inputs(name)
    B = len(rewards)
    rew_pad = np.zeros((B, max(len(r) for r in rewards)), np.float32)
    tp_pad = np.zeros((B, max(len(x) for x in to_play)), np.int32)
    for b in range(B):
        rew_pad[b, :len(rewards[b])] = rewards[b]
        tp_pad[b, :len(to_play[b])] = to_play[b]
    board = name.endswith("board_games")
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(a).to(dev)
    tr, tv = callers.reanalyze_value_targets(t(g("values")), g("td"), t(g("vmask")), t(rew_pad), g("reward_len"), g("seg_len"),
                                             g("pos"), int(g("U")), float(g("gamma")), to_play_segments=tp_pad if board else None,
                                             board_games=board)
    assert np.array_equal(tv.cpu().numpy().view(np.uint64), g("out_values").view(np.uint64)), "value targets"
    assert np.array_equal(tr.cpu().numpy().view(np.uint64), g("out_rewards").view(np.uint64)), "reward targets"


@pytest.mark.parametrize("name", ["policy_fixed", "policy_varied"])
def test_target_policies_kernel_matches_the_reference_golden(name):
    """lzb_target_policies against the outputs of game_buffer_muzero.py:606-656 (float64 there, float32 as the learner
    casts them)."""
    import torch
    from lightzero_b200.collect import legal_from_mask, target_policies
    from tests.test_reanalyze_oracle import GOLD, split
    g = lambda k: GOLD[name + "/" + k]
    mask = g("mask")
    T, A = mask.shape
    lens = mask.sum(1)
    visits = np.zeros((T, A), np.int32)
    for i, d in enumerate(split(g("dists"), lens)):
        visits[i, :len(d)] = d
    dev = torch.device("cuda", 0)
    cnt, ids = legal_from_mask(torch.as_tensor(mask).to(dev))
    pol = target_policies(torch.as_tensor(visits).to(dev), cnt, None if name == "policy_fixed" else ids,
                          torch.as_tensor(g("pmask")).to(dev))
    assert_same(pol.cpu().numpy().reshape(g("out").shape), g("out").astype(np.float32), name)


@pytest.mark.parametrize("name", ["prefix_single_player", "prefix_board_games", "prefix_horizon3"])
def test_value_prefix_targets_kernel_matches_the_reference_golden(name):
    """EfficientZero's flavour of lzb_value_targets (lstm_horizon_len > 0) against tests/golden/reanalyze_prefix_targets.npz:
    the outputs of game_buffer_efficientzero.py:261-316 executed as it stands (value prefixes with the batch-wide horizon
    counter, the base index of the board-games rule moving with the resets), float64 bit for bit."""
    import torch
    from lightzero_b200 import callers
    from tests.test_reanalyze_oracle import GOLD_EZ, split
    g = lambda k: GOLD_EZ[name + "/" + k]
    rewards = split(g("rewards"), g("reward_len"))
    to_play = split(g("to_play"), g("reward_len") + 8)
    B = len(rewards)
    rew_pad = np.zeros((B, max(len(r) for r in rewards)), np.float32)
    tp_pad = np.zeros((B, max(len(x) for x in to_play)), np.int32)
    for b in range(B):
        rew_pad[b, :len(rewards[b])] = rewards[b]
        tp_pad[b, :len(to_play[b])] = to_play[b]
    board = name.endswith("board_games")
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(a).to(dev)
    tp, tv = callers.reanalyze_value_targets(t(g("values")), g("td"), t(g("vmask")), t(rew_pad), g("reward_len"), g("seg_len"),
                                             g("pos"), int(g("U")), float(g("gamma")), to_play_segments=tp_pad if board else None,
                                             board_games=board, lstm_horizon_len=int(g("H")))
    assert np.array_equal(tv.cpu().numpy().view(np.uint64), g("out_values").view(np.uint64)), "value targets"
    assert np.array_equal(tp.cpu().numpy().view(np.uint64), g("out_prefixes").view(np.uint64)), "value-prefix targets"


def test_reanalyze_policy_targets_for_the_sampled_tree():
    """game_buffer_sampled_efficientzero.py:437-585: the Sampled buffer's reanalyze - roots(n, legal, A, K, continuous),
    5-argument search, targets over the K sampled actions, root_sampled_actions returned with them."""
    import torch
    import lightzero_b200 as lzb
    from lightzero_b200 import callers
    B, U, K, D, S, H = 8, 5, 6, 2, 20, 4
    T = B * (U + 1)
    dev = torch.device("cuda", 0)

    class Out:
        pass

    class Model:
        def eval(self):
            pass

        def recurrent_inference(self, latent, hidden, actions):
            o = Out()
            o.latent_state = latent * 0.5 + actions.sum(1, keepdim=True)
            o.policy_logits = torch.cat([torch.tanh(latent[:, :D]), torch.full((T, D), 0.5, device=dev)], 1)
            o.value = latent[:, :1] * 0.1
            o.value_prefix = latent[:, 1:2] * 0.05
            o.reward_hidden_state = (hidden[0] * 0.9, hidden[1] * 0.9)
            return o

    cfg = dict(num_simulations=S, discount_factor=0.997, value_delta_max=0.01, device="cuda", lstm_horizon_len=5, tie_break_seed=SEED,
               reanalyze_noise=False, model=dict(support_scale=300, categorical_distribution=False, continuous_action_space=True,
                                                 num_of_sampled_actions=K, action_space_size=D))
    mcts = identity(lzb.SampledEfficientZeroMCTSCtree(cfg))
    g = torch.Generator(device=dev).manual_seed(9)
    logits = torch.cat([torch.randn(T, D, generator=g, device=dev) * 0.3, torch.full((T, D), 0.4, device=dev)], 1)
    pmask = (torch.rand(T, generator=g, device=dev) < 0.8).to(torch.int32)
    hidden = (torch.zeros(1, T, H, device=dev), torch.zeros(1, T, H, device=dev))
    out = callers.reanalyze_policy_targets(mcts, Model(), torch.randn(T, 3, generator=g, device=dev), torch.zeros(T, device=dev), logits,
                                           None, torch.full((T,), -1, dtype=torch.int32, device=dev), pmask, U,
                                           hidden_state_roots=hidden, roots_args=(D, K, True))
    pol, acts = out["target_policies"], out["root_sampled_actions"]
    assert tuple(pol.shape) == (B, U + 1, K) and tuple(acts.shape) == (T, K, D) and float(acts.abs().max()) <= 1.0
    visits = out["visit_counts"].cpu().numpy()
    assert (visits.sum(1) == S).all()
    want = np.where(pmask.cpu().numpy()[:, None] != 0, (visits.astype(np.float64) / S).astype(np.float32), np.float32(0))
    assert np.array_equal(pol.reshape(T, K).cpu().numpy().view(np.uint32), want.view(np.uint32))
