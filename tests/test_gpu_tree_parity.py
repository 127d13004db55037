"""GPU parity: the CUDA forest (through the C ABI) against the CPU oracle on identical seeded inputs.

Bar (BASELINE.json north_star): selected actions, search paths and visit counts bit-exact;
root values, value sums, priors and MinMax bounds compared as raw fp32 bits (the reference tree is fp32,
so the tolerance is ZERO ulp, stricter than the 1e-5 relative fallback the north_star allows).
"""
import numpy as np
import pytest

from tests.common import PB_C_BASE, PB_C_INIT, TICTACTOE_MASKS, Scenario, assert_same, run_cpu

pytestmark = pytest.mark.gpu


def run_gpu(sc, rng_mode, seed, record=None, horizon=5, root_offset=0, fused=False, compact=False):
    """fused=False: lzb_traverse + lzb_backprop per simulation; fused=True: lzb_step (backprop s + traverse s+1 in one
    launch), with the leaf rows gathered through the slot tables.  Scenarios with ``reuse`` run the ReZero flavour of
    the same calls; compact=True additionally hands over the reference's no_inference / reuse lists (as flags) and the
    compacted batch index of every tree, the way a caller that compacts its model batches would."""
    import torch
    from lightzero_b200.forest import Forest
    f = Forest(sc.variant, sc.n, sc.K if sc.variant == "sez" else sc.A, sc.S, legal_actions=sc.legal, action_dim=sc.D,
               root_offset=root_offset, chance_space=sc.K if sc.variant == "smz" else 0)
    f.set_delta(0.01)
    dev = f.device
    t = lambda a: torch.as_tensor(a).to(dev)
    if sc.variant == "sez":
        f.prepare(None, None, 0.25, t(sc.root_reward), t(sc.to_play), sampled_actions=t(sc.root_actions))
    else:
        no_noise = getattr(sc, "no_noise", False)  # prepare_no_noise: no noise tensor, weight 0
        f.prepare(t(sc.root_logits), None if no_noise else t(sc.noise_pad), 0.0 if no_noise else 0.25, t(sc.root_reward),
                  t(sc.to_play), value=t(sc.root_value) if sc.variant == "gmz" else None)
    players = 1 if int(sc.to_play.max()) == -1 else 2
    cfg = f.make_cfg(PB_C_BASE, PB_C_INIT, sc.discount, players=players, num_simulations=sc.S,
                     max_num_considered_actions=sc.m, rng_mode=rng_mode, rng_seed=seed)
    tp = t(sc.to_play)
    vp = sc.variant in ("ez", "sez")
    # rows to gather: slot e holds rows filled with the value e (+ i / 1024), so a gathered row names its source
    L = 12
    rows = [(torch.arange(sc.n, device=dev).float().unsqueeze(1) / 1024 + e).expand(sc.n, L).contiguous() for e in range(sc.S + 1)]
    leaf, leaf_z = torch.empty(sc.n, L, device=dev), torch.empty(sc.n, L, device=dev)
    gathers = [(None, leaf), (None, leaf_z, True)] if fused else ()
    if fused:
        f.set_slots(0, [rows[0], rows[0]])
    is_reset_out = torch.empty(sc.n, dtype=torch.int32, device=dev) if vp else None
    reset_of = torch.zeros(sc.S + 1, sc.n, dtype=torch.int32, device=dev)

    reuse = None
    infer_count = torch.zeros(sc.S, dtype=torch.int32, device=dev)
    if getattr(sc, "reuse", False):
        reuse = dict(true_action=t(sc.true_action), reuse_value=t(sc.reuse_value))

    def chance_kw(s_traverse):
        """Stochastic MuZero: the outcomes to replay at the chance levels of simulation s_traverse (None = draw)."""
        if sc.variant != "smz" or sc.forced is None:
            return {}
        return dict(forced_outcome=t(np.ascontiguousarray(sc.forced[s_traverse], np.int32)))

    def reuse_for(s_traverse, o_prev=None):
        """reuse dict of one call: the traverse of simulation s_traverse counts its inferring trees; the backprop half
        takes flags / batch_index derived from the previous traverse's outputs (mcts_ctree.py:376-384) when compact."""
        if reuse is None:
            return None
        r = dict(reuse)
        if s_traverse is not None:
            r["infer_count"] = infer_count[s_traverse:]
        if compact and o_prev is not None:
            ix, la = o_prev["ix"].cpu().numpy(), o_prev["last_action"].cpu().numpy()
            infer = ix != -1
            flags = (~infer).astype(np.int32) | (((ix == 0) & (la == sc.true_action)).astype(np.int32) << 1)
            r["flags"] = t(flags)
            r["batch_index"] = t((np.cumsum(infer) - 1).astype(np.int32))
        return r

    def note(o, s):
        if fused:  # the gather delivered the parent's row; reset parents read as zeros through a zero_if_reset gather
            ix = o["ix"].long()
            inf = (ix >= 0).unsqueeze(1)  # trees that need no inference keep whatever their rows held
            ix = ix.clamp(min=0)
            want = ix.float().unsqueeze(1) + torch.arange(sc.n, device=dev).float().unsqueeze(1) / 1024
            assert torch.equal(torch.where(inf, leaf, want), want.expand(sc.n, L)), "sim %d: slot gather delivered the wrong rows" % s
            par_reset = reset_of[ix, torch.arange(sc.n, device=dev)].bool().unsqueeze(1)
            assert torch.equal(torch.where(inf, leaf_z, leaf), torch.where(par_reset & inf, torch.zeros_like(leaf), leaf)), "sim %d: zero_if_reset" % s
        if record is not None:
            rec = {k: v.cpu().numpy() for k, v in o.items()}
            rec["paths"] = f.dump_paths(sc.S + 1)
            record.append(rec)

    o = f.traverse(cfg, tp, gathers=gathers, reuse=reuse_for(0), **chance_kw(0))
    note(o, 0)
    for s in range(sc.S):
        m = sc.model_fn(s, None, None, None)
        kw = {}
        if vp:
            kw["lstm_horizon_len"] = horizon
            kw["is_reset_out"] = is_reset_out
        if sc.variant == "sez":
            kw["sampled_actions"] = t(m["sampled_actions"])
        if fused:
            kw["new_slots"] = [rows[s + 1], rows[s + 1]]
        args = (s + 1, sc.discount, t(m["reward"]), t(m["value"]), t(m["logits"]), o["vtp"])
        last = s + 1 == sc.S
        if fused and not last:
            o = f.step(cfg, *args, tp, out=o, gathers=gathers, reuse=reuse_for(s + 1, o), **chance_kw(s + 1), **kw)
        else:
            f.backprop(*args, reuse=reuse_for(None, o), **kw)
        if vp:
            reset_of[s + 1] = is_reset_out
            if record is not None:
                record[-1]["is_reset"] = is_reset_out.cpu().numpy()
        if not last:
            if not fused:
                o = f.traverse(cfg, tp, out=o, reuse=reuse_for(s + 1), **chance_kw(s + 1))
            note(o, s + 1)
    torch.cuda.synchronize()
    f.infer_count = infer_count.cpu().numpy()
    return f


def cpu_checker(sc, rng_mode, prefer_reference=False):
    """The CPU side of a comparison: the plain-C oracle, or (prefer_reference, discrete trees) the compiled unmodified
    reference tree from oracle/_ref when its library is there (it travels to the GPU box prebuilt)."""
    from oracle.binding import OracleForest, RefForest, ref_available
    if prefer_reference and sc.variant in ("mz", "ez", "gmz") and ref_available(sc.variant):
        return RefForest(sc.variant, sc.n, sc.A, sc.legal, per_root=(rng_mode == 2))
    return OracleForest(sc.variant, sc.n, sc.A, sc.legal, K=sc.K, D=sc.D)


def compare(sc, rng_mode=2, seed=2024, trees=4, fused=False, compact=False, prefer_reference=False):
    orc = cpu_checker(sc, rng_mode, prefer_reference)
    rec_c, rec_g = [], []
    run_cpu(orc, sc, rng_mode, seed, record=rec_c)
    gpu = run_gpu(sc, rng_mode, seed, record=rec_g, fused=fused, compact=compact)
    is_reuse = getattr(sc, "reuse", False)
    for s, (c, g) in enumerate(zip(rec_c, rec_g)):
        for k in ("search_len", "ix", "iy", "vtp", "paths", "leaf_is_chance"):
            if k == "iy" and is_reuse and not compact:  # model batches indexed by tree: every node's batch_index is its tree
                assert_same(g[k], np.arange(sc.n, dtype=np.int32), "sim %d iy" % s)
            elif k in c:
                assert_same(g[k], c[k], "sim %d %s" % (s, k))
        if is_reuse:
            assert gpu.infer_count[s] == int((c["ix"] != -1).sum()), "sim %d: inference count" % s
        if sc.variant == "sez":
            assert_same(g["last_action_f"], c["last_action_f"], "sim %d last_action_f" % s)
        else:
            assert_same(g["last_action"].astype(np.int32), c["last_action"], "sim %d last_action" % s)
        if c.get("is_reset") is not None:
            assert_same(g["is_reset"], c["is_reset"], "sim %d is_reset" % s)
    out = gpu.root_outputs(sc.discount, gumbel=sc.variant == "gmz")
    dist, ln = orc.distributions()
    W = dist.shape[1]
    assert_same(out["legal_count"].cpu().numpy(), ln, "legal_count")
    assert_same(out["visits"].cpu().numpy()[:, :W], np.where(dist < 0, 0, dist), "visit counts")
    assert_same(out["values"].cpu().numpy(), orc.values(), "root values")
    assert (out["visits"].cpu().numpy().sum(1) == sc.S).all()
    mn, mx = gpu.minmax()
    omn, omx = orc.minmax()
    assert_same(mn, omn, "minmax.min")
    assert_same(mx, omx, "minmax.max")
    vc, vs, rw = gpu.root_stats()
    ovc, ovs, orw = orc.root_stats()
    assert_same(vc, ovc, "root visit_count")
    assert (vc == sc.S + 1).all()
    assert_same(vs, ovs, "root value_sum")
    for i in range(min(trees, sc.n)):
        tg, tc = gpu.dump_tree(i), orc.dump_tree(i, sc.S + 1)
        for k in ("visit", "value_sum", "reward", "child_e", "hdr") + (("prior",) if sc.variant != "sez" else ()):
            assert_same(tg[k], tc[k], "tree %d %s" % (i, k))
        if sc.variant == "gmz":
            assert_same(tg["raw_value"], tc["raw_value"], "tree %d raw_value" % i)
    if sc.variant == "gmz":
        p, v = orc.gumbel_outputs(sc.discount)
        assert_same(out["policies"].cpu().numpy(), p, "gumbel improved policy")
        assert_same(out["children_values"].cpu().numpy(), v, "gumbel completed child values")
    return gpu, orc


def test_device_libm_matches_oracle():
    import torch
    from lightzero_b200 import _lib
    from lightzero_b200._lib import check
    from oracle.binding import oracle_lib
    import ctypes as C
    L, O = _lib.lib(), oracle_lib()
    rng = np.random.default_rng(0)
    # softmax arguments are <= 0; pUCT log arguments are >= 1; plus raw bit patterns over the whole range
    xs = np.concatenate([-np.abs(rng.standard_normal(1 << 20) * 8).astype(np.float32),
                         rng.integers(0, 2 ** 32, 1 << 20, dtype=np.uint64).astype(np.uint32).view(np.float32),
                         np.array([0.0, -0.0, -87.9, -88.0, -103.0, -103.9, -104.5, 88.7, 89.0, -np.inf, np.inf], np.float32)])
    ls = np.concatenate([(1 + np.abs(rng.standard_normal(1 << 20)) * 3).astype(np.float32),
                         rng.integers(0, 2 ** 32, 1 << 20, dtype=np.uint64).astype(np.uint32).view(np.float32),
                         np.array([1.0, 1.0000001, 0.0, 1e-40, np.inf, 19653.0 / 19652.0], np.float32)])
    for name, x, fn in (("expf", xs, "orc_expf_array"), ("logf", ls, "orc_logf_array")):
        want = np.empty_like(x)
        getattr(O, fn)(x.ctypes.data_as(C.c_void_p), want.ctypes.data_as(C.c_void_p), C.c_long(len(x)))
        xd = torch.from_numpy(x).cuda()
        od = torch.empty_like(xd)
        check(getattr(L, "lzb_test_" + name)(xd.data_ptr(), od.data_ptr(), len(x), None))
        got = od.cpu().numpy()
        both_nan = np.isnan(got) & np.isnan(want)
        bad = (got.view(np.uint32) != want.view(np.uint32)) & ~both_nan
        assert not bad.any(), "%s: %d mismatches, e.g. x=%r got=%r want=%r" % (name, bad.sum(), x[bad][0], got[bad][0], want[bad][0])


@pytest.mark.parametrize("rng_mode", [1, 2])
def test_muzero_cartpole_shape(rng_mode):  # C1: 8 roots x 25 sims x A=2
    compare(Scenario("mz", 8, 2, 25, seed=1), rng_mode=rng_mode)


def test_muzero_synthetic_a18():  # C5 shape at a size the oracle finishes in seconds
    compare(Scenario("mz", 512, 18, 50, seed=2))


def test_muzero_zero_model_all_ties():  # the reference tests' fake model: every level is one big tie
    gpu, _ = compare(Scenario("mz", 16, 9, 40, seed=3, zero_model=True))
    assert (gpu.root_outputs()["values"].cpu().numpy() == 0.0).all()


def test_muzero_legal_masks():  # the 16 x 9 mask of test_mcts_ctree.py:120-137
    legal = [[a for a, x in enumerate(row) if x == 1] for row in TICTACTOE_MASKS]
    compare(Scenario("mz", 16, 9, 50, seed=4, legal=legal))


def test_muzero_two_player_sign_flips():  # C4': to_play alternation, gamma = 1
    compare(Scenario("mz", 32, 9, 25, seed=5, to_play=[1, 2] * 16, discount=1.0))


def test_muzero_to_play_zero_runs_two_player_branch():  # asserts are compiled out
    compare(Scenario("mz", 16, 9, 25, seed=6, to_play=[0] * 16))


def test_muzero_deep_paths_more_than_one_warp():  # A=2, 200 sims: search paths longer than 32 nodes
    sc = Scenario("mz", 8, 2, 200, seed=7, reward_scale=0.0)
    sc.value[:] = 0.0
    sc.root_logits[:] = 0.0
    sc.logits[:] = np.array([8.0, -8.0], np.float32)  # one dominant prior drives a deep chain
    compare(sc)


def test_muzero_wide_action_space_multi_chunk():  # A > 32 exercises the chunked warp loops
    compare(Scenario("mz", 16, 80, 60, seed=8))


def test_muzero_very_wide_warp_per_tree_kernels():  # A > 128: the shared-memory warp-per-tree kernels
    compare(Scenario("mz", 8, 150, 40, seed=32))
    compare(Scenario("ez", 8, 150, 30, seed=33), fused=True)


def test_warp_per_tree_kernels_on_small_action_space(monkeypatch):
    monkeypatch.setenv("LZB_FORCE_WARP_PER_TREE", "1")
    compare(Scenario("mz", 64, 18, 50, seed=34), fused=True)
    compare(Scenario("ez", 32, 6, 50, seed=35))
    compare(Scenario("gmz", 16, 9, 30, seed=36, discount=1.0, max_considered=3), rng_mode=1, fused=True)
    compare(Scenario("gmz", 8, 18, 40, seed=37, max_considered=16), rng_mode=1)


def test_gumbel_wide_action_spaces():  # lane groups of 32 x 2 / 32 x 4 records, and the warp-per-tree kernels past 128
    compare(Scenario("gmz", 8, 40, 60, seed=38, max_considered=8), rng_mode=1, fused=True)
    compare(Scenario("gmz", 4, 100, 60, seed=39, max_considered=16), rng_mode=1)
    compare(Scenario("gmz", 4, 140, 50, seed=40, max_considered=16), rng_mode=1, fused=True)


def test_gumbel_many_roots():  # more than one block per group width, partial last block
    compare(Scenario("gmz", 1100, 18, 30, seed=41, max_considered=8), rng_mode=1, fused=True)


def test_efficientzero_pong_shape():  # C2: A=6, 50 sims, value_prefix + lstm_horizon_len=5 resets
    compare(Scenario("ez", 64, 6, 50, seed=9))


def test_efficientzero_two_player():
    compare(Scenario("ez", 32, 9, 25, seed=10, to_play=[2, 1] * 16, discount=1.0))


def test_gumbel_tictactoe_masks():  # C4: bot mode, legal masks, m=3, 30 sims, gamma=1
    legal = [[a for a, x in enumerate(row) if x == 1] for row in TICTACTOE_MASKS]
    compare(Scenario("gmz", 16, 9, 30, seed=11, legal=legal, discount=1.0, max_considered=3))


def test_gumbel_gomoku_shape():  # C4: 6x6 board, A=36, m=6, 50 sims
    compare(Scenario("gmz", 32, 36, 50, seed=12, discount=1.0, max_considered=6))


def test_gumbel_m16_a18():
    compare(Scenario("gmz", 64, 18, 50, seed=13, max_considered=16))


def test_sampled_continuous_halfcheetah_shape():  # C3: K=20, D=6, 50 sims
    compare(Scenario("sez", 8, 12, 50, seed=14, K=20, D=6))


# Stochastic MuZero (K = chance_space_size).  sc.forced is None: the oracle and the CUDA path DRAW their chance
# outcomes from the same counter-based stream, so the draws themselves are compared too.
SMZ_CASES = {
    "2048_a4_c32": lambda: Scenario("smz", 64, 4, 50, seed=41, K=32),
    "a6_c2": lambda: Scenario("smz", 200, 6, 50, seed=42, K=2),
    "a18_c18_many_roots": lambda: Scenario("smz", 1200, 18, 30, seed=43, K=18),
    "a3_c40_two_records_per_lane": lambda: Scenario("smz", 32, 3, 40, seed=44, K=40),
    "a100_c70": lambda: Scenario("smz", 8, 100, 40, seed=45, K=70),
    "masks_two_player": lambda: Scenario("smz", 16, 9, 40, seed=46, K=5, to_play=[1, 2] * 8, discount=1.0,
                                         legal=[[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]),
    "single_outcome": lambda: Scenario("smz", 16, 5, 30, seed=47, K=1),
    "zero_model_all_ties": lambda: Scenario("smz", 16, 4, 40, seed=48, K=4, zero_model=True),
    "deep_a1_c1": lambda: Scenario("smz", 4, 1, 120, seed=49, K=1),
}


@pytest.mark.parametrize("rng_mode", [1, 2])
@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("name", sorted(SMZ_CASES))
def test_stochastic_muzero_matches_oracle(name, fused, rng_mode):
    compare(SMZ_CASES[name](), rng_mode=rng_mode, fused=fused)


def test_stochastic_muzero_is_lane_group_only(monkeypatch):
    from lightzero_b200._lib import LzbError
    from lightzero_b200.forest import Forest
    with pytest.raises(LzbError, match="<= 128"):
        Forest("smz", 4, 4, 10, chance_space=200)
    with pytest.raises(ValueError):
        Forest("smz", 4, 4, 10)


FUSED_CASES = {
    "mz_a18": lambda: Scenario("mz", 256, 18, 50, seed=21),
    "mz_cartpole": lambda: Scenario("mz", 8, 2, 25, seed=22),
    "mz_masks": lambda: Scenario("mz", 16, 9, 50, seed=23, legal=[[a for a, x in enumerate(r) if x == 1] for r in TICTACTOE_MASKS]),
    "mz_two_player": lambda: Scenario("mz", 32, 9, 25, seed=24, to_play=[1, 2] * 16, discount=1.0),
    "mz_wide": lambda: Scenario("mz", 16, 80, 40, seed=25),
    "ez_pong": lambda: Scenario("ez", 64, 6, 50, seed=26),
    "ez_two_player": lambda: Scenario("ez", 32, 9, 25, seed=27, to_play=[2, 1] * 16, discount=1.0),
    "gmz_gomoku": lambda: Scenario("gmz", 32, 36, 50, seed=28, discount=1.0, max_considered=6),
    "sez_halfcheetah": lambda: Scenario("sez", 8, 12, 50, seed=29, K=20, D=6),
}


@pytest.mark.parametrize("name", sorted(FUSED_CASES))
def test_fused_step_matches_oracle(name):
    """lzb_step (expand + backup of simulation s and selection + slot gather of s + 1 in one launch)."""
    compare(FUSED_CASES[name](), fused=True)


def test_fused_step_deep_paths():
    sc = Scenario("mz", 8, 2, 120, seed=30, reward_scale=0.0)
    sc.value[:] = 0.0
    sc.root_logits[:] = 0.0
    sc.logits[:] = np.array([8.0, -8.0], np.float32)
    compare(sc, fused=True)


LEGAL9 = [[a for a, x in enumerate(row) if x == 1] for row in TICTACTOE_MASKS]
REUSE_CASES = {  # ReZero search_with_reuse (mcts_ctree.py:323-420, 829-953): scenario kwargs, rng_mode
    "mz_a6": (dict(variant="mz", n=32, A=6, S=50, seed=14, reuse=True), 2),
    "mz_a18_batch_rng": (dict(variant="mz", n=16, A=18, S=50, seed=15, reuse=True), 1),
    "mz_a18_many_roots": (dict(variant="mz", n=700, A=18, S=40, seed=40, reuse=True), 2),
    "mz_masks_two_player": (dict(variant="mz", n=16, A=9, S=30, seed=16, legal=LEGAL9, to_play=[1, 2] * 8, discount=1.0,
                                 reuse=True), 2),
    "mz_zero_model": (dict(variant="mz", n=16, A=9, S=40, seed=41, zero_model=True, reuse=True), 2),
    "mz_a40": (dict(variant="mz", n=16, A=40, S=60, seed=42, reuse=True), 2),
    "mz_a100": (dict(variant="mz", n=8, A=100, S=50, seed=43, reuse=True), 2),
    "ez_a6": (dict(variant="ez", n=32, A=6, S=50, seed=17, reuse=True), 2),
    "ez_two_player": (dict(variant="ez", n=16, A=9, S=30, seed=18, to_play=[2, 1] * 8, discount=1.0, reuse=True), 1),
    "ez_a18": (dict(variant="ez", n=300, A=18, S=50, seed=44, reuse=True), 2),
}


@pytest.mark.parametrize("compact", [False, True])
@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("name", sorted(REUSE_CASES))
def test_reuse_search_matches_oracle(name, fused, compact):
    kw, mode = REUSE_CASES[name]
    gpu, orc = compare(Scenario(**kw), rng_mode=mode, fused=fused, compact=compact)
    assert gpu.infer_count.sum() / kw["S"] == orc.infer_stats[1]  # average_infer of search_with_reuse


def test_reuse_search_is_rejected_where_the_reference_has_none():
    import torch
    from lightzero_b200.forest import Forest
    for variant, A, kw in (("gmz", 9, {}), ("mz", 150, {})):
        f = Forest(variant, 4, A, 8)
        dev = f.device
        f.prepare(torch.zeros(4, A, device=dev), None, 0.0, torch.zeros(4, device=dev),
                  torch.full((4,), -1, dtype=torch.int32, device=dev), value=torch.zeros(4, device=dev))
        cfg = f.make_cfg(num_simulations=8, players=1)
        with pytest.raises(RuntimeError, match="reuse"):
            f.traverse(cfg, torch.full((4,), -1, dtype=torch.int32, device=dev),
                       reuse=dict(true_action=torch.zeros(4, dtype=torch.int32, device=dev),
                                  reuse_value=torch.zeros(4, device=dev)))


def test_chance_head_is_checked_at_the_abi():
    """lzb_chance_head_t: stochastic tree only, non-NULL outputs, strides not below the node widths."""
    import torch
    from lightzero_b200.forest import Forest
    n, A, S = 4, 6, 4
    z = lambda *shape: torch.zeros(*shape, device="cuda")
    tp = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    f = Forest("mz", n, A, S)
    f.prepare(z(n, A), None, 0.0, z(n), tp)
    cfg = f.make_cfg(num_simulations=S, players=1)
    o = f.traverse(cfg, tp)
    head = dict(reward=z(n), value=z(n), logits=z(n, A))
    with pytest.raises(ValueError, match="stochastic"):
        f.backprop(1, 0.997, z(n), z(n), z(n, A), o["vtp"], chance_head=head)
    f = Forest("smz", n, A, S, chance_space=5)
    f.prepare(z(n, f.A), None, 0.0, z(n), tp)
    cfg = f.make_cfg(num_simulations=S, players=1)
    o = f.traverse(cfg, tp)
    with pytest.raises(RuntimeError, match="stride"):  # 3 < 5 chance outcomes
        f.backprop(1, 0.997, z(n), z(n), z(n, A), o["vtp"], chance_head=dict(reward=z(n), value=z(n), logits=z(n, 3)))
    with pytest.raises(RuntimeError, match="stride"):  # decision logits narrower than the decision action space
        f.backprop(1, 0.997, z(n), z(n), z(n, A - 1), o["vtp"], chance_head=dict(reward=z(n), value=z(n), logits=z(n, 5)))
    # exact widths (no padding to max(A, C)) are accepted
    f.backprop(1, 0.997, z(n), z(n), z(n, A), o["vtp"], chance_head=dict(reward=z(n), value=z(n), logits=z(n, 5), latent=z(n, 8)),
               new_slots=[z(n, 8)])


EDGE_SHAPES = {  # smallest / boundary sizes of every kernel flavour: (variant, n, A, S, extra scenario kwargs)
    "one_root_one_action": ("mz", 1, 1, 12, {}),
    "one_simulation": ("mz", 33, 18, 1, {}),
    "two_simulations_ez": ("ez", 5, 6, 2, {}),
    "a8_group_boundary": ("mz", 40, 8, 30, {}),
    "a16_group_boundary": ("ez", 40, 16, 30, {}),
    "a24_group_boundary": ("mz", 40, 24, 30, {}),
    "a32_group_boundary": ("mz", 24, 32, 40, {}),
    "a33": ("ez", 24, 33, 40, {}),
    "a64_group_boundary": ("mz", 12, 64, 70, {}),
    "a128_group_boundary": ("mz", 6, 128, 140, {}),
    "a129_first_generic": ("mz", 6, 129, 140, {}),
    "single_legal_action": ("mz", 9, 9, 20, dict(legal=[[k] for k in range(9)])),
    "ragged_legal_lists": ("ez", 7, 12, 40, dict(legal=[list(range(k + 1)) for k in range(7)])),
    "partial_last_block": ("mz", 129, 18, 20, {}),
    "gumbel_one_root": ("gmz", 1, 5, 16, dict(max_considered=2)),
    "gumbel_more_considered_than_actions": ("gmz", 9, 4, 24, dict(max_considered=16)),
    "sampled_one_dim": ("sez", 9, 5, 20, dict(K=5, D=1)),
}


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("name", sorted(EDGE_SHAPES))
def test_edge_shapes_match_oracle(name, fused):
    variant, n, A, S, kw = EDGE_SHAPES[name]
    compare(Scenario(variant, n, A, S, seed=900 + len(name), **kw), fused=fused)


def test_many_simulations_long_search():  # visit counts well past one byte, 800-simulation AlphaZero-style budget
    compare(Scenario("mz", 4, 5, 800, seed=77), fused=True)


def test_trajectories_follow_best_actions():
    sc = Scenario("mz", 16, 6, 30, seed=31)
    rec = []
    gpu = run_gpu(sc, 2, 5, record=rec)
    acts, ln = gpu.trajectories()
    acts, ln = acts.cpu().numpy(), ln.cpu().numpy()
    last = rec[-1]
    for i in range(sc.n):  # the most recent traverse wrote best_action along its whole path
        d = int(last["search_len"][i])
        assert ln[i] >= d and (acts[i, :d] == last["paths"][i, :d]).all()


def test_root_offset_keys_the_tiebreak_stream():
    """Sharding invariance: roots [4, 8) searched as a shard with root_offset=4 equal the same roots of the full batch."""
    full = Scenario("mz", 8, 6, 30, seed=15, zero_model=True)
    g_full = run_gpu(full, 2, 77)
    part = Scenario("mz", 8, 6, 30, seed=15, zero_model=True)
    for name in ("root_logits", "root_reward", "root_value", "noise_pad", "to_play"):
        setattr(part, name, getattr(full, name)[4:])
    part.logits, part.reward, part.value = full.logits[:, 4:], full.reward[:, 4:], full.value[:, 4:]
    part.legal, part.n = full.legal[4:], 4
    g_part = run_gpu(part, 2, 77, root_offset=4)
    a = g_full.root_outputs()["visits"].cpu().numpy()[4:]
    b = g_part.root_outputs()["visits"].cpu().numpy()
    assert_same(b, a, "sharded visits")
