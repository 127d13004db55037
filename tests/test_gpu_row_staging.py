"""The leaf-row gather of the fused step kernel when ONE row per tree is gathered (the MuZero case): every byte of every
gathered row, for row sizes on both sides of the lone walker's staging path (lzb_group.cuh, lone_stage_rows: one row of at
most 4096 bytes, a multiple of 16, goes through the bulk-copy engine for the tree of a warp that ends its walk first and
through a 32-lane gather for the other; anything else takes the ordinary 16-lane path).  The walk itself is compared with
the CPU checker (oracle/, TEST INFRASTRUCTURE) per simulation, so a wrong row cannot hide behind a right index.

Reference behaviour: the Python leaf gather of mcts_ctree.py:283-289 (latent_state_batch_in_search_path[ix][iy])."""
import numpy as np
import pytest

from tests.common import PB_C_BASE, PB_C_INIT, Scenario, assert_same, run_cpu

pytestmark = pytest.mark.gpu


def _search_with_rows(sc, L, seed=31):
    """Fused search (lzb_step) with one int32 row of L words per tree; slot e, tree i, word c holds (e << 22 | i << 11 | c)."""
    import torch
    from lightzero_b200.forest import Forest
    f = Forest(sc.variant, sc.n, sc.A, sc.S, legal_actions=sc.legal)
    f.set_delta(0.01)
    dev = f.device
    t = lambda a: torch.as_tensor(a).to(dev)
    f.prepare(t(sc.root_logits), t(sc.noise_pad), 0.25, t(sc.root_reward), t(sc.to_play))
    players = 1 if int(sc.to_play.max()) == -1 else 2
    cfg = f.make_cfg(PB_C_BASE, PB_C_INIT, sc.discount, players=players, num_simulations=sc.S, rng_mode=2, rng_seed=seed)
    tp = t(sc.to_play)
    base = (torch.arange(sc.n, device=dev, dtype=torch.int32).unsqueeze(1) << 11) | torch.arange(L, device=dev, dtype=torch.int32).unsqueeze(0)
    rows = [(base | (e << 22)).contiguous() for e in range(sc.S + 1)]
    leaf = torch.full((sc.n, L), -1, dtype=torch.int32, device=dev)
    gathers = [(None, leaf)]
    f.set_slots(0, [rows[0]])
    rec = []

    def note(o, s):
        ix = o["ix"].long()
        assert bool((ix >= 0).all()), "sim %d: a tree without a leaf to infer" % s
        want = base | (ix.to(torch.int32).unsqueeze(1) << 22)
        bad = (leaf != want).any(dim=1).nonzero().flatten()
        assert bad.numel() == 0, "sim %d: %d trees received wrong rows (first: tree %d, L=%d)" % (s, bad.numel(), int(bad[0]), L)
        leaf.fill_(-1)  # the next gather must write every word again
        rec.append({k: v.cpu().numpy() for k, v in o.items()})

    o = f.traverse(cfg, tp, gathers=gathers)
    note(o, 0)
    for s in range(sc.S):
        m = sc.model_fn(s, None, None, None)
        args = (s + 1, sc.discount, t(m["reward"]), t(m["value"]), t(m["logits"]), o["vtp"])
        if s + 1 < sc.S:
            o = f.step(cfg, *args, tp, out=o, gathers=gathers, new_slots=[rows[s + 1]])
            note(o, s + 1)
        else:
            f.backprop(*args, new_slots=[rows[s + 1]])
    torch.cuda.synchronize()
    return f, rec


# 16 B (one vector), 400 B (25 vectors: a partial 32-lane round), 4096 B (the staging limit, 8 full rounds of 32 lanes),
# 4092 B (not a multiple of 16: ordinary path), 4112 B (past the limit: ordinary path)
@pytest.mark.parametrize("L", [4, 100, 1024, 1023, 1028])
@pytest.mark.parametrize("n", [257, 64])  # 257: the last warp holds one tree (no sibling to stage for)
def test_single_row_gather_delivers_every_byte(L, n):
    from oracle.binding import OracleForest
    sc = Scenario("mz", n, 18, 24, seed=7 + L)
    orc = OracleForest("mz", sc.n, sc.A, sc.legal)
    rec_c = []
    run_cpu(orc, sc, 2, 31, record=rec_c)
    f, rec_g = _search_with_rows(sc, L)
    for s, (c, g) in enumerate(zip(rec_c, rec_g)):
        for k in ("search_len", "ix"):
            assert_same(g[k], c[k], "sim %d %s" % (s, k))
    out = f.root_outputs(sc.discount)
    dist, _ = orc.distributions()
    assert_same(out["visits"].cpu().numpy()[:, :dist.shape[1]], np.where(dist < 0, 0, dist), "visit counts")
    assert_same(out["values"].cpu().numpy(), orc.values(), "root values")
