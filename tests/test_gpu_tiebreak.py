"""GPU: the tie-break stream is per search and safe under CUDA-graph replay.

The reference reseeds rand() from the wall clock on every traverse (ctree_muzero/lib/cnode.cpp:770), so consecutive
searches never share a first-visit order.  Here the stream is explicit: per Roots object by default, pinned when the
caller gives ``tie_break_seed``, derived per search ordinal with ``tie_break_advance``; in graph mode the kernels read
the seed from a device word (lzb_search_cfg_t::rng_seed_dev) that is rewritten before every replay."""
import numpy as np
import pytest

from tests.common import PB_C_BASE, PB_C_INIT, Scenario
from tests.test_gpu_api import ScenarioModel, identity

pytestmark = pytest.mark.gpu


def _searches(cfg_extra, graph, count=2, variant="mz"):
    """`count` consecutive searches of a ZERO model (every level is an all-way tie) by ONE driver;
    returns the visit counts of each."""
    import torch
    import lightzero_b200 as lzb
    sc = Scenario(variant, 64, 9, 24, seed=5, zero_model=True)
    dev = torch.device("cuda", 0)
    cls = {"mz": lzb.MuZeroMCTSCtree, "ez": lzb.EfficientZeroMCTSCtree}[variant]
    cfg = dict(num_simulations=sc.S, pb_c_base=PB_C_BASE, pb_c_init=PB_C_INIT, discount_factor=sc.discount,
               value_delta_max=0.01, device="cuda", lstm_horizon_len=5, cuda_graph=graph,
               model=dict(support_scale=300, categorical_distribution=False))
    cfg.update(cfg_extra)
    mcts = identity(cls(cfg))
    model = ScenarioModel(sc, dev, hidden=8 if variant == "ez" else 0)
    lat0 = torch.zeros(sc.n, 16, device=dev)
    hid = (torch.zeros(1, sc.n, 8, device=dev), torch.zeros(1, sc.n, 8, device=dev))
    tp = torch.as_tensor(sc.to_play).to(dev)
    out = []
    for _ in range(count):
        roots = cls.roots(sc.n, sc.legal)
        roots.prepare_no_noise(sc.root_reward, sc.root_logits, tp)  # uniform priors: the root is an all-way tie too
        if variant == "ez":
            mcts.search(roots, model, lat0, hid, tp)
        else:
            mcts.search(roots, model, lat0, tp)
        v, _ = roots.get_distributions_tensor()
        out.append(v.cpu().numpy().copy())
        assert (out[-1].sum(1) == sc.S).all()
    return out


@pytest.mark.parametrize("variant", ["mz", "ez"])
@pytest.mark.parametrize("graph", [False, True])
def test_default_stream_changes_from_search_to_search(graph, variant):
    a = _searches({}, graph, count=3, variant=variant)  # graph: eager warm-up + capture, then two replays
    assert (a[0] != a[1]).any() and (a[1] != a[2]).any(), "consecutive searches replayed the same first-visit order"


@pytest.mark.parametrize("graph", [False, True])
def test_pinned_seed_repeats_and_advance_follows_the_ordinal(graph):
    pinned = _searches(dict(tie_break_seed=11), graph, count=3)
    assert np.array_equal(pinned[0], pinned[1]) and np.array_equal(pinned[1], pinned[2])
    adv_a = _searches(dict(tie_break_seed=11, tie_break_advance=True), graph, count=3)
    adv_b = _searches(dict(tie_break_seed=11, tie_break_advance=True), graph, count=3)
    assert (adv_a[0] != adv_a[1]).any() and (adv_a[1] != adv_a[2]).any()
    for k in range(3):  # the same (seed, ordinal) reproduces
        assert np.array_equal(adv_a[k], adv_b[k]), "search %d" % k
    # graph replay and eager launches agree for the same (seed, ordinal)
    other = _searches(dict(tie_break_seed=11, tie_break_advance=True), not graph, count=3)
    for k in range(3):
        assert np.array_equal(adv_a[k], other[k]), "graph vs eager, search %d" % k


def test_seed_word_equals_seed_by_value():
    """lzb_search_cfg_t::rng_seed_dev == the same value in rng_seed, at the C ABI."""
    import torch
    from lightzero_b200.forest import Forest
    sc = Scenario("mz", 32, 6, 16, seed=9, zero_model=True)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.as_tensor(a).to(dev)
    outs = []
    for by_word in (False, True):
        f = Forest("mz", sc.n, sc.A, sc.S, legal_actions=sc.legal, device=dev)
        f.set_delta(0.01)
        f.prepare(t(sc.root_logits), None, 0.0, t(sc.root_reward), t(sc.to_play))
        seed = 0x1234567890abcdef
        word = torch.full((1,), seed, dtype=torch.int64, device=dev) if by_word else None
        cfg = f.make_cfg(PB_C_BASE, PB_C_INIT, sc.discount, players=1, rng_mode=2, rng_seed=0 if by_word else seed,
                         rng_seed_dev=word)
        tp = t(sc.to_play)
        for s in range(sc.S):
            o = f.traverse(cfg, tp)
            f.backprop(s + 1, sc.discount, t(sc.reward[s]), t(sc.value[s]), t(sc.logits[s]), o["vtp"])
        outs.append(f.root_outputs()["visits"].cpu().numpy())
    assert np.array_equal(outs[0], outs[1])
