"""GPU: the reference-shaped public API (Roots / MinMaxStatsList / ResultsWrapper / batch_traverse /
batch_backpropagate modules and the XMCTSCtree.search drivers) against the CPU oracle, plus the shape / legality
checks the reference's own tests make (lzero/mcts/tests/test_mcts_ctree.py:155-179,240-281, test_mcts_sampled_ctree.py)."""
import numpy as np
import pytest

from tests.common import PB_C_BASE, PB_C_INIT, TICTACTOE_MASKS, Scenario, assert_same, run_cpu

pytestmark = pytest.mark.gpu
SEED = 2024


class Out:
    pass


class ScenarioModel:
    """model.recurrent_inference over a Scenario's per-simulation tables (device tensors)."""

    def __init__(self, sc, dev, latent_dim=16, hidden=0):
        import torch
        self.t = lambda a: torch.as_tensor(a).to(dev)
        self.sc, self.s, self.dev, self.L, self.H = sc, 0, dev, latent_dim, hidden
        self.calls = []
        self.logits, self.value, self.reward = self.t(sc.logits), self.t(sc.value), self.t(sc.reward)  # tables in HBM

    def eval(self):
        self.s = 0

    def recurrent_inference(self, latent, *rest):
        import torch
        sc, s = self.sc, self.s
        self.s += 1
        snap = lambda x: tuple(snap(y) for y in x) if isinstance(x, (tuple, list)) else x.clone()
        self.calls.append((latent.clone(), snap(rest)))  # the driver reuses its output buffers every simulation
        o = Out()
        o.latent_state = torch.full_like(latent, float(s + 1))
        o.policy_logits = self.logits[s]
        o.value = self.value[s].unsqueeze(1)
        o.reward = o.value_prefix = self.reward[s].unsqueeze(1)
        if self.H:
            o.reward_hidden_state = (torch.full((1, sc.n, self.H), float(s + 1), device=self.dev),
                                     torch.full((1, sc.n, self.H), -float(s + 1), device=self.dev))
        return o


def driver_cfg(sc, **extra):
    cfg = dict(num_simulations=sc.S, pb_c_base=PB_C_BASE, pb_c_init=PB_C_INIT, discount_factor=sc.discount,
               value_delta_max=0.01, device="cuda", tie_break_seed=SEED, lstm_horizon_len=5,
               max_num_considered_actions=sc.m, model=dict(support_scale=300, categorical_distribution=False))
    cfg.update(extra)
    return cfg


def identity(mcts):
    mcts.inverse_scalar_transform_handle = lambda x: x
    return mcts


def oracle_run(sc):
    from oracle.binding import OracleForest
    orc = OracleForest(sc.variant, sc.n, sc.A, sc.legal, K=sc.K, D=sc.D)
    rec = []
    run_cpu(orc, sc, 2, SEED, record=rec)
    return orc, rec


def check_roots(roots, orc, sc):
    dist, ln = orc.distributions()
    got = roots.get_distributions()
    assert [len(r) for r in got] == ln.tolist()
    for i, r in enumerate(got):
        assert r == dist[i, :ln[i]].tolist(), "root %d visit counts" % i
        assert sum(r) == sc.S
    assert_same(np.asarray(roots.get_values(), np.float32), orc.values(), "root values")


@pytest.mark.parametrize("graph", [False, True])
def test_muzero_driver_matches_oracle(graph):
    import torch
    import lightzero_b200 as lzb
    legal = [[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]
    sc = Scenario("mz", 16, 9, 30, seed=41, legal=legal)
    orc, rec = oracle_run(sc)
    mcts = identity(lzb.MuZeroMCTSCtree(driver_cfg(sc, cuda_graph=graph)))
    model = ScenarioModel(sc, torch.device("cuda"))
    for rep in range(2 if graph else 1):  # second search of a shape replays the captured graph
        roots = lzb.MuZeroMCTSCtree.roots(sc.n, sc.legal)
        roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
        mcts.search(roots, model, np.zeros((sc.n, 16), np.float32), sc.to_play.tolist())
        check_roots(roots, orc, sc)
    if not graph:  # every model call received the parent latents of the oracle's leaves: slot e is filled with e
        for s, (lat, rest) in enumerate(model.calls):
            assert_same(lat[:, 0].cpu().numpy(), rec[s]["ix"].astype(np.float32), "sim %d gathered latents" % s)
            assert_same(rest[0].cpu().numpy().astype(np.int32), rec[s]["last_action"], "sim %d last actions" % s)
            assert rest[0].dtype == torch.int64


def test_muzero_two_player_board_game_driver():
    import torch
    import lightzero_b200 as lzb
    sc = Scenario("mz", 8, 9, 25, seed=42, to_play=[1, 2] * 4, discount=1.0)
    orc, _ = oracle_run(sc)
    mcts = identity(lzb.MuZeroMCTSCtree(driver_cfg(sc, env_type="board_games")))
    roots = lzb.MuZeroMCTSCtree.roots(sc.n, sc.legal)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    tp = sc.to_play.tolist()
    mcts.search(roots, ScenarioModel(sc, torch.device("cuda")), np.zeros((sc.n, 4, 2, 2), np.float32), tp)
    assert tp == sc.to_play.tolist()  # the caller's list is not mutated
    check_roots(roots, orc, sc)


def test_efficientzero_driver_matches_oracle_and_resets_hidden_state():
    import torch
    import lightzero_b200 as lzb
    sc = Scenario("ez", 32, 6, 40, seed=43)
    orc, rec = oracle_run(sc)
    mcts = identity(lzb.EfficientZeroMCTSCtree(driver_cfg(sc)))
    model = ScenarioModel(sc, torch.device("cuda"), hidden=8)
    roots = lzb.EfficientZeroMCTSCtree.roots(sc.n, sc.legal)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    hidden = (np.full((1, sc.n, 8), 0.5, np.float32), np.full((1, sc.n, 8), -0.5, np.float32))
    mcts.search(roots, model, np.zeros((sc.n, 16), np.float32), hidden, sc.to_play.tolist())
    check_roots(roots, orc, sc)
    # hidden state handed to the model: the parent's (c, h), zeroed when the parent was expanded on a horizon boundary
    reset = {0: np.zeros(sc.n, bool)}
    for s in range(sc.S):
        reset[s + 1] = rec[s]["is_reset"].astype(bool)
    for s, (lat, rest) in enumerate(model.calls):
        (c, h), act = rest
        ix = rec[s]["ix"]
        par_reset = np.array([reset[int(e)][i] for i, e in enumerate(ix)])
        want_c = np.where(ix == 0, 0.5, ix.astype(np.float32))
        want_c = np.where(par_reset, 0.0, want_c).astype(np.float32)
        assert c.shape == (1, sc.n, 8) and h.shape == (1, sc.n, 8)
        assert_same(c[0, :, 0].cpu().numpy(), want_c, "sim %d LSTM c" % s)
        assert_same(h[0, :, 3].cpu().numpy(), np.where(par_reset, 0.0, -want_c).astype(np.float32), "sim %d LSTM h" % s)


@pytest.mark.parametrize("graph", [False, True])
def test_gumbel_driver_matches_oracle(graph):
    import torch
    import lightzero_b200 as lzb
    legal = [[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]
    sc = Scenario("gmz", 16, 9, 30, seed=44, legal=legal, discount=1.0, max_considered=3)
    orc, _ = oracle_run(sc)
    mcts = identity(lzb.GumbelMuZeroMCTSCtree(driver_cfg(sc, cuda_graph=graph)))
    model = ScenarioModel(sc, torch.device("cuda"))
    for rep in range(2 if graph else 1):
        roots = lzb.GumbelMuZeroMCTSCtree.roots(sc.n, sc.legal)
        roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_value.tolist(),
                      sc.root_logits.tolist(), sc.to_play.tolist())
        mcts.search(roots, model, np.zeros((sc.n, 16), np.float32), sc.to_play.tolist())
        check_roots(roots, orc, sc)
        p, v = orc.gumbel_outputs(sc.discount)
        assert_same(np.asarray(roots.get_policies(sc.discount, sc.A), np.float32), p, "improved policies")
        assert_same(np.asarray(roots.get_children_values(sc.discount, sc.A), np.float32), v, "children values")
    if not graph:
        assert model.calls[0][1][0].shape == (sc.n, 1)  # Gumbel passes actions as [N, 1] (mcts_ctree.py:1090)


def test_gumbel_roots_prepared_again_forget_the_cached_softmax_constants():
    """The Gumbel kernels memoise, per node, the maximum and log-sum-exp of the softmax over its priors (Pool::gstat).
    prepare() on a LIVE Roots object re-roots the same workspace with other priors: the cache must not survive it."""
    import torch
    import lightzero_b200 as lzb
    legal = [[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]
    first = Scenario("gmz", 16, 9, 30, seed=91, legal=legal, discount=1.0, max_considered=3)
    sc = Scenario("gmz", 16, 9, 30, seed=44, legal=legal, discount=1.0, max_considered=3)
    orc, _ = oracle_run(sc)
    mcts = identity(lzb.GumbelMuZeroMCTSCtree(driver_cfg(sc)))
    roots = lzb.GumbelMuZeroMCTSCtree.roots(sc.n, sc.legal)
    for s_ in (first, sc):  # same object, same forest: searched with `first`, then prepared again and searched with `sc`
        roots.prepare(0.25, [n.tolist() for n in s_.noises], s_.root_reward.tolist(), s_.root_value.tolist(),
                      s_.root_logits.tolist(), s_.to_play.tolist())
        mcts.search(roots, ScenarioModel(s_, torch.device("cuda")), np.zeros((sc.n, 16), np.float32), s_.to_play.tolist())
    check_roots(roots, orc, sc)
    p_, v_ = orc.gumbel_outputs(sc.discount)
    assert_same(np.asarray(roots.get_policies(sc.discount, sc.A), np.float32), p_, "improved policies")


def test_list_api_module_level_calls_match_oracle():
    """The tree modules driven exactly like the reference loop drives the Cython ones (Python lists in and out)."""
    from lightzero_b200.ctree.ctree_muzero import mz_tree
    sc = Scenario("mz", 8, 6, 20, seed=45)
    orc, rec = oracle_run(sc)
    mz_tree.set_tie_break(SEED)
    roots = mz_tree.Roots(sc.n, sc.legal)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    mm = mz_tree.MinMaxStatsList(sc.n)
    mm.set_delta(0.01)
    for s in range(sc.S):
        results = mz_tree.ResultsWrapper(sc.n)
        ix, iy, last_actions, vtp = mz_tree.batch_traverse(roots, PB_C_BASE, PB_C_INIT, sc.discount, mm, results, sc.to_play.tolist())
        assert isinstance(ix, list) and isinstance(last_actions[0], int)
        assert ix == rec[s]["ix"].tolist() and iy == list(range(sc.n)) and last_actions == rec[s]["last_action"].tolist()
        assert results.get_search_len() == rec[s]["search_len"].tolist()
        mz_tree.batch_backpropagate(s + 1, sc.discount, sc.reward[s].tolist(), sc.value[s].tolist(), sc.logits[s].tolist(),
                                    mm, results, vtp)
    check_roots(roots, orc, sc)
    traj = roots.get_trajectories()
    assert len(traj) == sc.n and all(len(t) >= 1 for t in traj)
    roots.clear()
    assert roots.get_distributions() == []


@pytest.mark.parametrize("variant", ["mz", "ez"])
def test_reuse_list_api_driven_like_the_reference_loop(variant):
    """batch_traverse_with_reuse / batch_backpropagate_with_reuse with the reference's own bookkeeping around them
    (mcts_ctree.py:359-416): compacted model batches, no_inference_lst / reuse_lst ending in -1."""
    from lightzero_b200.ctree.ctree_efficientzero import ez_tree
    from lightzero_b200.ctree.ctree_muzero import mz_tree
    tree = mz_tree if variant == "mz" else ez_tree
    sc = Scenario(variant, 24, 6, 40, seed=46, reuse=True)
    orc, rec = oracle_run(sc)
    tree.set_tie_break(SEED)
    roots = tree.Roots(sc.n, sc.legal)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    mm = tree.MinMaxStatsList(sc.n)
    mm.set_delta(0.01)
    true_action, reuse_value = sc.true_action.tolist(), sc.reuse_value.tolist()
    skipped = 0
    for s in range(sc.S):
        results = tree.ResultsWrapper(sc.n)
        ix, iy, last_actions, vtp = tree.batch_traverse_with_reuse(roots, PB_C_BASE, PB_C_INIT, sc.discount, mm, results,
                                                                  sc.to_play.tolist(), true_action, reuse_value)
        assert ix == rec[s]["ix"].tolist() and iy == rec[s]["iy"].tolist()  # iy: the compacted batch index
        assert last_actions == rec[s]["last_action"].tolist()
        search_lens = results.get_search_len()
        rows, no_inference_lst, reuse_lst = [], [], []
        for count, (x, y) in enumerate(zip(ix, iy)):
            if x != -1:
                rows.append(count)
            else:
                no_inference_lst.append(y)
            if x == 0 and last_actions[count] == true_action[count]:
                reuse_lst.append(count)
        skipped += len(no_inference_lst)
        no_inference_lst.append(-1)
        reuse_lst.append(-1)
        args = (s + 1, sc.discount, sc.reward[s][rows].tolist(), sc.value[s][rows].tolist(), sc.logits[s][rows].tolist(),
                mm, results)
        if variant == "ez":
            is_reset = (np.asarray(search_lens) % 5 == 0).astype(np.int32).tolist()
            tree.batch_backpropagate_with_reuse(*args, is_reset, vtp, no_inference_lst, reuse_lst, reuse_value)
        else:
            tree.batch_backpropagate_with_reuse(*args, vtp, no_inference_lst, reuse_lst, reuse_value)
    assert skipped > 0
    check_roots(roots, orc, sc)


@pytest.mark.parametrize("variant", ["mz", "ez"])
def test_search_with_reuse_driver_matches_oracle(variant):
    import torch
    import lightzero_b200 as lzb
    sc = Scenario(variant, 48, 6, 50, seed=47, reuse=True)
    orc, rec = oracle_run(sc)
    dev = torch.device("cuda")
    if variant == "mz":
        mcts = identity(lzb.MuZeroMCTSCtree(driver_cfg(sc)))
        roots = lzb.MuZeroMCTSCtree.roots(sc.n, sc.legal)
        model = ScenarioModel(sc, dev)
    else:
        mcts = identity(lzb.EfficientZeroMCTSCtree(driver_cfg(sc)))
        roots = lzb.EfficientZeroMCTSCtree.roots(sc.n, sc.legal)
        model = ScenarioModel(sc, dev, hidden=8)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    lat = np.zeros((sc.n, 16), np.float32)
    if variant == "mz":
        length, average_infer = mcts.search_with_reuse(roots, model, lat, sc.to_play.tolist(), sc.true_action.tolist(),
                                                       sc.reuse_value.tolist())
    else:
        hid = (np.zeros((1, sc.n, 8), np.float32), np.zeros((1, sc.n, 8), np.float32))
        length, average_infer = mcts.search_with_reuse(roots, model, lat, hid, sc.to_play.tolist(),
                                                       sc.true_action.tolist(), sc.reuse_value.tolist())
    assert (length, average_infer) == tuple(orc.infer_stats)
    assert average_infer < sc.n  # some trees skipped inference
    check_roots(roots, orc, sc)
    for s, (lat_s, rest) in enumerate(model.calls):  # inferring trees received their parent's latent
        need = rec[s]["ix"] != -1
        assert_same(lat_s[:, 0].cpu().numpy()[need], rec[s]["ix"].astype(np.float32)[need], "sim %d gathered latents" % s)


def test_search_with_reuse_is_not_offered_for_gumbel():
    import lightzero_b200 as lzb
    sc = Scenario("gmz", 4, 6, 8, seed=48)
    with pytest.raises(NotImplementedError):
        lzb.GumbelMuZeroMCTSCtree(driver_cfg(sc)).search_with_reuse(None, None, None, None, [0] * 4, [0.0] * 4)


@pytest.mark.parametrize("on_device", [False, True])
def test_unizero_driver_matches_oracle_and_hands_over_the_history(on_device):
    """mcts_ctree.py:19-170: same tree as MuZero; the world model receives (state_action_history, simulation_index,
    latent_state_index_in_search_path)."""
    import torch
    import lightzero_b200 as lzb
    sc = Scenario("mz", 24, 6, 30, seed=60)
    orc, rec = oracle_run(sc)
    dev = torch.device("cuda")
    base = ScenarioModel(sc, dev)

    class WorldModel:
        def __init__(self):
            self.seen = []

        def eval(self):
            base.eval()

        def recurrent_inference(self, state_action_history, simulation_index, latent_state_index_in_search_path):
            assert len(state_action_history) == simulation_index + 1
            latent, actions = state_action_history[-1]
            if on_device:
                assert isinstance(latent, torch.Tensor) and latent.is_cuda and isinstance(latent_state_index_in_search_path, torch.Tensor)
                ix = latent_state_index_in_search_path.cpu().tolist()
                lat = latent
            else:
                assert isinstance(latent, np.ndarray) and isinstance(latent_state_index_in_search_path, list)
                ix = latent_state_index_in_search_path
                lat = torch.from_numpy(latent).to(dev)
            assert actions.dtype == torch.int64 and actions.is_cuda
            self.seen.append((ix, actions.cpu().tolist(), lat[:, 0].cpu().tolist()))
            return base.recurrent_inference(lat, actions)

    model = WorldModel()
    mcts = identity(lzb.UniZeroMCTSCtree(driver_cfg(sc, history_on_device=on_device)))
    roots = lzb.UniZeroMCTSCtree.roots(sc.n, sc.legal)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    mcts.search(roots, model, np.zeros((sc.n, 16), np.float32), sc.to_play.tolist())
    check_roots(roots, orc, sc)
    for s, (ix, actions, lat0) in enumerate(model.seen):
        assert ix == rec[s]["ix"].tolist() and actions == rec[s]["last_action"].tolist()
        assert lat0 == rec[s]["ix"].astype(np.float32).tolist()  # slot e is filled with e: the parents' latents arrived
    # earlier history entries are not overwritten by later simulations
    assert model.seen[0][1] == rec[0]["last_action"].tolist()


def test_rnn_full_obs_driver_matches_oracle_and_gathers_the_history():
    """mcts_ctree.py:422-620: MuZero tree, value_prefix backed up as the reward, world-model history gathered next to
    the latent and replaced by reward_hidden_state."""
    import torch
    import lightzero_b200 as lzb
    sc = Scenario("mz", 24, 6, 30, seed=61)
    orc, rec = oracle_run(sc)
    dev = torch.device("cuda")
    base = ScenarioModel(sc, dev)
    H = 8

    class RNNModel:
        def __init__(self):
            self.seen, self.kw = [], []

        def eval(self):
            base.eval()

        def recurrent_inference(self, latent, history, actions, **kw):
            s = base.s
            assert history.shape == (1, sc.n, H)
            self.seen.append((latent[:, 0].cpu().tolist(), history[0, :, 0].cpu().tolist(), actions.cpu().tolist()))
            self.kw.append(kw)
            o = base.recurrent_inference(latent, actions)
            o.predict_next_latent_state = o.latent_state
            o.reward_hidden_state = torch.full((1, sc.n, H), -float(s + 1), device=dev)
            o.value_prefix = o.reward
            del o.reward, o.latent_state
            return o

    model = RNNModel()
    mcts = identity(lzb.MuZeroRNNFullObsMCTSCtree(driver_cfg(sc)))
    roots = lzb.MuZeroRNNFullObsMCTSCtree.roots(sc.n, sc.legal)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits.tolist(), sc.to_play.tolist())
    mcts.search(roots, model, np.zeros((sc.n, 16), np.float32), np.zeros((1, sc.n, H), np.float32), sc.to_play.tolist(),
                ready_env_id=[3, 4])
    check_roots(roots, orc, sc)
    for s, (lat0, hist0, actions) in enumerate(model.seen):
        want = rec[s]["ix"].astype(np.float32)
        assert lat0 == want.tolist() and hist0 == (-want).tolist() and actions == rec[s]["last_action"].tolist()
    assert all(k == dict(ready_env_id=[3, 4]) for k in model.kw)


def test_reference_unit_test_shapes_with_zero_model():
    """test_mcts_ctree.py:155-179,240-281: all-zero fake model, 16 x 9 mask; output shapes, per-root distribution
    length == legal count, and argmax action is legal."""
    import torch
    import lightzero_b200 as lzb
    legal = [[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]
    sc = Scenario("mz", 16, 9, 50, seed=46, legal=legal, zero_model=True)
    mcts = identity(lzb.MuZeroMCTSCtree(driver_cfg(sc)))
    roots = lzb.MuZeroMCTSCtree.roots(16, legal)
    noises = [np.random.dirichlet([0.3] * len(l)).astype(np.float32).tolist() for l in legal]
    roots.prepare(0.25, noises, [0.0] * 16, np.zeros((16, 9), np.float32).tolist(), [-1] * 16)
    mcts.search(roots, ScenarioModel(sc, torch.device("cuda")), np.zeros((16, 8), np.float32), [-1] * 16)
    dist, vals = roots.get_distributions(), roots.get_values()
    assert len(dist) == 16 and len(vals) == 16 and all(v == 0.0 for v in vals)
    for i, d in enumerate(dist):
        assert len(d) == len(legal[i]) and sum(d) == 50
        assert legal[i][int(np.argmax(d))] in legal[i]


@pytest.mark.parametrize("policy", ["EfficientZero", "GumbelMuZero", "MuZero"])
def test_reference_large_case_256_roots_500_simulations(policy):
    """test_mcts_ctree.py:182-237 (test_mcts_vs_bot_to_play_large): 256 roots x 500 simulations x 20 actions with the
    all-zero fake model and to_play = 0 for every root; output shapes as asserted there, plus the visit-count sum."""
    import torch
    import lightzero_b200 as lzb
    n, A, S = 256, 20, 500
    dev = torch.device("cuda")

    class FakeModel:  # MuZeroModelFake of the reference test: zeros everywhere
        def eval(self):
            pass

        def recurrent_inference(self, latent, *rest):
            o = Out()
            o.latent_state = torch.zeros_like(latent)
            o.policy_logits = torch.zeros(n, A, device=dev)
            o.value = torch.zeros(n, 1, device=dev)
            o.reward = o.value_prefix = torch.zeros(n, 1, device=dev)
            o.reward_hidden_state = (torch.zeros(1, n, 16, device=dev), torch.zeros(1, n, 16, device=dev))
            return o

    cfg = dict(num_simulations=S, pb_c_base=PB_C_BASE, pb_c_init=PB_C_INIT, discount_factor=0.997, value_delta_max=0.01,
               device="cuda", lstm_horizon_len=5, max_num_considered_actions=6,
               model=dict(support_scale=300, categorical_distribution=False))
    legal = [list(range(A)) for _ in range(n)]
    noises = [np.random.dirichlet([0.3] * A).astype(np.float32).tolist() for _ in range(n)]
    zeros, logits, to_play = [0.0] * n, np.zeros((n, A), np.float32).tolist(), [0] * n
    lat = np.zeros((n, 100), np.float32)
    if policy == "EfficientZero":
        mcts = identity(lzb.EfficientZeroMCTSCtree(cfg))
        roots = lzb.EfficientZeroMCTSCtree.roots(n, legal)
        roots.prepare(0.25, noises, zeros, logits, to_play)
        mcts.search(roots, FakeModel(), lat, (np.zeros((1, n, 16), np.float32), np.zeros((1, n, 16), np.float32)), to_play)
    elif policy == "GumbelMuZero":
        mcts = identity(lzb.GumbelMuZeroMCTSCtree(cfg))
        roots = lzb.GumbelMuZeroMCTSCtree.roots(n, legal)
        roots.prepare(0.25, noises, zeros, zeros, logits, to_play)
        mcts.search(roots, FakeModel(), lat, to_play)
    else:
        mcts = identity(lzb.MuZeroMCTSCtree(cfg))
        roots = lzb.MuZeroMCTSCtree.roots(n, legal)
        roots.prepare(0.25, noises, zeros, logits, to_play)
        mcts.search(roots, FakeModel(), lat, to_play)
    dist, vals = roots.get_distributions(), roots.get_values()
    assert np.array(dist).shape == (n, A) and np.array(vals).shape == (n,)
    assert (np.array(dist).sum(1) == S).all()


def test_sampled_driver_continuous_shapes():
    """test_mcts_sampled_ctree.py:61-135: continuous K=6, D=2, 5 roots; distributions are (5, 6)."""
    import torch
    import lightzero_b200 as lzb
    n, K, D, S = 5, 6, 2, 40
    dev = torch.device("cuda")

    class Model:
        def eval(self):
            pass

        def recurrent_inference(self, latent, hidden, actions):
            assert actions.shape == (n, D) and actions.dtype == torch.float32 and float(actions.abs().max()) <= 1.0
            o = Out()
            o.latent_state = latent + 1
            o.policy_logits = torch.cat([torch.zeros(n, D, device=dev), torch.ones(n, D, device=dev)], 1)
            o.value = torch.full((n, 1), 0.1, device=dev)
            o.value_prefix = torch.full((n, 1), 0.05, device=dev)
            o.reward_hidden_state = (hidden[0] + 1, hidden[1] + 1)
            return o

    cfg = dict(num_simulations=S, discount_factor=0.997, value_delta_max=0.01, device="cuda", lstm_horizon_len=5,
               model=dict(support_scale=300, categorical_distribution=False, continuous_action_space=True,
                          num_of_sampled_actions=K, action_space_size=D))
    mcts = identity(lzb.SampledEfficientZeroMCTSCtree(cfg))
    legal = [[-1] * K for _ in range(n)]
    roots = lzb.SampledEfficientZeroMCTSCtree.roots(n, legal, D, K, True)
    logits = np.concatenate([np.zeros((n, D)), np.ones((n, D))], 1).astype(np.float32)
    roots.prepare(0.25, [np.random.dirichlet([0.3] * K).tolist() for _ in range(n)], [0.0] * n, logits.tolist(), [-1] * n)
    hidden = (np.zeros((1, n, 4), np.float32), np.zeros((1, n, 4), np.float32))
    mcts.search(roots, Model(), np.zeros((n, 3), np.float32), hidden, [-1] * n)
    dist = np.asarray(roots.get_distributions())
    assert dist.shape == (n, K) and (dist.sum(1) == S).all()
    acts = np.asarray(roots.get_sampled_actions())
    assert acts.shape == (n, K, D) and np.abs(acts).max() < 1.0
    assert len(roots.get_values()) == n


def test_sampled_driver_cuda_graph_matches_eager_and_redraws():
    """cfg.cuda_graph for the sampled tree: the captured search reads its sampler seed from device memory, so (1) a
    replay under the seed of an eager search reproduces that search exactly, and (2) the next search draws new tables."""
    import torch
    import lightzero_b200 as lzb
    from lightzero_b200.tree_search import sampling
    import itertools
    n, K, D, S, H = 64, 8, 3, 30, 8
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(3)
    tab = dict(logits=torch.cat([torch.randn(S, n, D, generator=g, device=dev),
                                 torch.rand(S, n, D, generator=g, device=dev) * 0.5 + 0.1], -1),
               value=torch.randn(S, n, 1, generator=g, device=dev), prefix=torch.randn(S, n, 1, generator=g, device=dev))

    class Model:
        def eval(self):
            self.s = 0

        def recurrent_inference(self, latent, hidden, actions):
            s = self.s
            self.s += 1
            o = Out()
            o.latent_state = latent * 0.5 + actions.sum(1, keepdim=True)  # depends on the sampled action that was taken
            o.policy_logits, o.value, o.value_prefix = tab["logits"][s], tab["value"][s] + o.latent_state[:, :1], tab["prefix"][s]
            o.reward_hidden_state = (hidden[0] + 1, hidden[1] - 1)
            return o

    root_logits = torch.cat([torch.randn(n, D, generator=g, device=dev), torch.rand(n, D, generator=g, device=dev) * 0.5 + 0.1], -1)
    lat = torch.randn(n, 16, generator=g, device=dev)
    hid = (torch.zeros(1, n, H, device=dev), torch.zeros(1, n, H, device=dev))
    tp = [-1] * n

    def run(graph, ordinal):
        sampling._search_counter = itertools.count(ordinal)  # the search ordinal is half of the sampler seed
        cfg = dict(num_simulations=S, discount_factor=0.997, value_delta_max=0.01, device="cuda", lstm_horizon_len=5,
                   cuda_graph=graph, tie_break_seed=SEED,
                   model=dict(support_scale=300, categorical_distribution=False, continuous_action_space=True,
                              num_of_sampled_actions=K, action_space_size=D))
        mcts = run.cache.setdefault(graph, identity(lzb.SampledEfficientZeroMCTSCtree(cfg)))
        roots = lzb.SampledEfficientZeroMCTSCtree.roots(n, None, D, K, True)
        roots.prepare_no_noise(torch.zeros(n, device=dev), root_logits, tp)
        mcts.search(roots, Model(), lat, hid, tp)
        v, _ = roots.get_distributions_tensor()
        return v.clone(), roots.get_values_tensor().clone(), roots.get_sampled_actions_tensor().clone()

    run.cache = {}
    e1, e2 = run(False, 101), run(False, 102)
    g1 = run(True, 101)   # warm-up + capture + first replay
    g2 = run(True, 102)   # pure replay with a new seed
    g1b = run(True, 101)  # replay again under the first seed
    for a, b in ((e1, g1), (e2, g2), (e1, g1b)):
        assert torch.equal(a[2], b[2]) and torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert not torch.equal(g1[2], g2[2]) and not torch.equal(g1[0], g2[0])  # fresh root tables and a different search
    assert (g2[0].sum(1) == S).all()


# ---------------------------------------------------------------- Stochastic MuZero (chance nodes)
class StochasticScenarioModel:
    """model.recurrent_inference(latent, action, afterstate=...) of the stochastic drivers over a Scenario's tables.
    Sub-batches arrive compacted, so column 1 of every latent row carries the tree index and the model looks its
    rows up by it; every tree is inferred once per simulation (``calls_per_row`` times with dense inference)."""

    def __init__(self, sc, dev, calls_per_row=1, poison=None):
        import torch
        self.sc, self.dev, self.per_sim = sc, dev, sc.n * calls_per_row
        # dense inference runs both heads on every row; ``poison`` = the oracle's per-simulation leaf kinds: the rows a
        # head is NOT meant for then come back as garbage, so a tree that picks the wrong head cannot match the oracle
        self.poison = None if poison is None else [torch.as_tensor(r["leaf_is_chance"].astype(np.bool_)).to(dev) for r in poison]
        t = lambda a: torch.as_tensor(a).to(dev)
        self.logits, self.value, self.reward = t(sc.logits), t(sc.value), t(sc.reward)
        self.rows_seen = 0
        self.calls = []

    def eval(self):
        self.rows_seen = 0

    def chunk(self, lo, hi):
        """The model of roots [lo, hi) (cfg.search_streams): rows are looked up by tree index, only the call counter is per chunk."""
        import copy
        m = copy.copy(self)
        m.per_sim, m.rows_seen, m.calls = (hi - lo) * (self.per_sim // self.sc.n), 0, []
        return m

    def root_latents(self, L=8):
        lat = np.zeros((self.sc.n, L), np.float32)
        lat[:, 1] = np.arange(self.sc.n)
        return lat

    def recurrent_inference(self, latent, action, afterstate=False):
        import torch
        sc = self.sc
        s = self.rows_seen // self.per_sim
        self.rows_seen += latent.shape[0]
        rows = latent[:, 1].long()
        self.calls.append(dict(s=s, rows=rows.clone(), ix=latent[:, 0].clone(), action=action.clone(), afterstate=afterstate))
        width = sc.A if afterstate else sc.K  # afterstate in -> decision node out (policy logits); else chance logits
        o = Out()
        o.latent_state = latent.clone()
        o.latent_state[:, 0] = float(s + 1)
        o.policy_logits = self.logits[s][rows, :width]
        o.value = self.value[s][rows].unsqueeze(1)
        o.reward = self.reward[s][rows].unsqueeze(1)
        if self.poison is not None and s < len(self.poison):
            wrong = self.poison[s][rows] == afterstate  # afterstate=False serves the chance leaves
            o.latent_state[wrong, 0] = -777.0
            o.policy_logits = torch.where(wrong.unsqueeze(1), o.policy_logits.flip(1) * 3.0 + 1.0, o.policy_logits)
            o.value = torch.where(wrong.unsqueeze(1), o.value + 1000.0, o.value)
            o.reward = torch.where(wrong.unsqueeze(1), o.reward - 1000.0, o.reward)
        return o


def test_stochastic_list_api_driven_like_the_reference_loop():
    """stochastic_mz_tree driven the way mcts_ctree_stochastic.py:128-251 drives the Cython module: five lists out of
    batch_traverse, then chance leaves and decision leaves backed up by two calls with compacted rows."""
    from lightzero_b200.ctree.ctree_stochastic_muzero import stochastic_mz_tree as smz
    sc = Scenario("smz", 24, 5, 30, seed=61, K=3)
    orc, rec = oracle_run(sc)
    smz.set_tie_break(SEED)
    roots = smz.Roots(sc.n, sc.legal, sc.K)
    roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits[:, :sc.A].tolist(),
                  sc.to_play.tolist())
    mm = smz.MinMaxStatsList(sc.n)
    mm.set_delta(0.01)
    both = 0
    for s in range(sc.S):
        results = smz.ResultsWrapper(sc.n)
        is_chance, ix, iy, last_actions, vtp = smz.batch_traverse(roots, PB_C_BASE, PB_C_INIT, sc.discount, mm, results,
                                                                  sc.to_play.tolist())
        assert all(isinstance(x, bool) for x in is_chance)
        assert is_chance == [bool(x) for x in rec[s]["leaf_is_chance"]]
        assert ix == rec[s]["ix"].tolist() and last_actions == rec[s]["last_action"].tolist()
        chance = [i for i, c in enumerate(is_chance) if c]
        decision = [i for i, c in enumerate(is_chance) if not c]
        both += bool(chance) and bool(decision)
        for idx, width in ((chance, sc.K), (decision, sc.A)):
            if idx:
                smz.batch_backpropagate(s + 1, sc.discount, sc.reward[s][idx].tolist(), sc.value[s][idx].tolist(),
                                        sc.logits[s][idx, :width].tolist(), mm, results, vtp, is_chance, idx)
    assert both > 5  # mixed simulations happened: the subset calls were exercised
    check_roots(roots, orc, sc)
    assert roots.forest.sims_done == sc.S


@pytest.mark.parametrize("mode", ["compact", "dense", "dense_torch", "dense_graph", "dense_graph_3streams"])
def test_stochastic_driver_matches_oracle(mode):
    import torch
    import lightzero_b200 as lzb
    legal = [[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]
    sc = Scenario("smz", 16, 9, 30, seed=62, K=4, legal=legal)
    orc, rec = oracle_run(sc)
    dense = mode != "compact"
    mcts = identity(lzb.StochasticMuZeroMCTSCtree(driver_cfg(sc, stochastic_dense_inference=dense,
                                                             stochastic_select_in_torch=mode == "dense_torch",
                                                             search_streams=3 if mode == "dense_graph_3streams" else 1,
                                                             cuda_graph=mode.startswith("dense_graph"))))
    # "dense" / "dense_graph": the tree kernels pick each leaf's head (lzb_chance_head_t); "dense_torch": torch.where does
    model = StochasticScenarioModel(sc, torch.device("cuda"), calls_per_row=2 if dense else 1,
                                    poison=rec if mode in ("dense", "dense_torch") else None)
    first = None
    for rep in range(3 if mode.startswith("dense_graph") else 1):
        roots = lzb.StochasticMuZeroMCTSCtree.roots(sc.n, sc.legal, sc.K)
        if rep == 2:
            roots.chance_seed = SEED  # replay of the captured search under the first search's chance stream
        roots.prepare(0.25, [n.tolist() for n in sc.noises], sc.root_reward.tolist(), sc.root_logits[:, :sc.A].tolist(),
                      sc.to_play.tolist())
        mcts.search(roots, model, model.root_latents(), sc.to_play.tolist())
        got = roots.get_distributions()
        assert all(sum(r) == sc.S for r in got) and [len(r) for r in got] == [len(l) for l in legal]
        if rep == 1:  # the next search of the driver draws other outcomes (seed + 1): not the same search again
            assert got != first
        else:
            check_roots(roots, orc, sc)
            first = got
    if mode in ("dense", "dense_torch"):  # every head saw, on the rows meant for it, the latent the oracle's leaf hangs below
        for c in model.calls:
            mine = rec[c["s"]]["leaf_is_chance"].astype(bool) != c["afterstate"]
            assert_same(c["ix"].cpu().numpy()[mine], rec[c["s"]]["ix"][mine].astype(np.float32), "sim %d gathered latents" % c["s"])
    if mode == "compact":  # the model saw the reference's call pattern: chance rows first, each leaf kind by itself
        by_sim = {}
        for c in model.calls:
            by_sim.setdefault(c["s"], []).append(c)
        for s in range(sc.S):
            isc = rec[s]["leaf_is_chance"].astype(bool)
            calls = by_sim[s]
            assert [c["afterstate"] for c in calls] == ([False] if isc.all() else [True] if not isc.any() else [False, True])
            for c in calls:
                rows = c["rows"].cpu().numpy()
                assert_same(rows, np.flatnonzero(isc != c["afterstate"]), "sim %d rows of afterstate=%s" % (s, c["afterstate"]))
                assert_same(c["ix"].cpu().numpy(), rec[s]["ix"][rows].astype(np.float32), "sim %d gathered latents" % s)
                assert_same(c["action"].cpu().numpy().astype(np.int32), rec[s]["last_action"][rows], "sim %d actions" % s)


def test_stochastic_driver_rejects_graph_without_dense_inference():
    import lightzero_b200 as lzb
    sc = Scenario("smz", 4, 3, 5, seed=63, K=2)
    mcts = lzb.StochasticMuZeroMCTSCtree(driver_cfg(sc, cuda_graph=True))
    roots = lzb.StochasticMuZeroMCTSCtree.roots(sc.n, sc.legal)
    assert roots.chance_space_size == 2  # the reference's default (mcts_ctree_stochastic.py:72)
    roots.prepare_no_noise(sc.root_reward.tolist(), sc.root_logits[:, :sc.A].tolist(), sc.to_play.tolist())
    with pytest.raises(ValueError, match="dense"):
        mcts.search(roots, None, np.zeros((sc.n, 4), np.float32), sc.to_play.tolist())
