"""GPU: the root-batch prologue / epilogue kernels (lightzero_b200/collect.py) against oracle/collect_oracle.py
(itself pinned to numpy / scipy in tests/test_collect_oracle.py) and, for the Dirichlet sampler, against the
distribution's moments.  Integer results bit-exact; float32 targets bit-exact; float64 entropy to 1e-12 (device
log() vs glibc log() may differ in the last bit)."""
import numpy as np
import pytest

from oracle import collect_oracle as co

pytestmark = pytest.mark.gpu


def random_masks(rng, N, A):
    mask = (rng.random((N, A)) < 0.6).astype(np.int8)
    mask[np.arange(N), rng.integers(0, A, N)] = 1  # at least one legal action
    mask[0] = 1
    return mask


@pytest.mark.parametrize("A", [1, 9, 18, 40, 130])
def test_legal_from_mask(A):
    import torch
    from lightzero_b200.collect import legal_from_mask
    rng = np.random.default_rng(A)
    mask = random_masks(rng, 257, A)
    cnt, ids = legal_from_mask(torch.from_numpy(mask).cuda())
    want = co.legal_from_mask(mask)
    assert cnt.cpu().tolist() == [len(l) for l in want]
    got = ids.cpu().numpy()
    for i, l in enumerate(want):
        assert got[i, :len(l)].tolist() == l and (got[i, len(l):] == 0).all()
    # float masks compare equal to 1 the same way (np.where(action_mask == 1.0), policy/muzero.py:709)
    cnt2, ids2 = legal_from_mask(mask.astype(np.float32))
    assert torch.equal(cnt, cnt2) and torch.equal(ids, ids2)


@pytest.mark.parametrize("temperature", [1.0, 0.5, 0.25, 0.7])
def test_select_action_matches_oracle(temperature):
    import torch
    from lightzero_b200.collect import legal_from_mask, select_action
    rng = np.random.default_rng(11)
    N, A, S = 600, 18, 50
    mask = random_masks(rng, N, A)
    legal = co.legal_from_mask(mask)
    dists = [rng.multinomial(S, rng.dirichlet([0.5] * len(l))).tolist() for l in legal]
    visits = np.zeros((N, A), np.int32)
    for i, d in enumerate(dists):
        visits[i, :len(d)] = d
    cnt, ids = legal_from_mask(mask)
    v = torch.from_numpy(visits).cuda()
    # deterministic
    pos, act, ent = select_action(v, cnt, ids, temperature, deterministic=True)
    want = [co.select_action(d, temperature, True) for d in dists]
    assert pos.cpu().tolist() == [w[0] for w in want]
    assert act.cpu().tolist() == [legal[i][w[0]] for i, w in enumerate(want)]
    np.testing.assert_allclose(ent.cpu().numpy(), [w[1] for w in want], rtol=1e-12, atol=1e-15)
    # sampled, with the draws np.random.choice would have made injected
    u = rng.random(N)
    pos, act, _ = select_action(v, cnt, ids, temperature, deterministic=False, uniforms=torch.from_numpy(u).cuda())
    want = [co.select_action(d, temperature, False, uniform=u[i])[0] for i, d in enumerate(dists)]
    if temperature in (1.0, 0.5, 0.25):  # integral exponents: exact powers on both sides
        assert pos.cpu().tolist() == want
    else:  # pow() may differ in the last bit: a draw within 1e-12 of a CDF step may land on either side
        assert (np.asarray(pos.cpu().tolist()) != np.asarray(want)).sum() <= 1
    assert act.cpu().tolist() == [legal[i][p] for i, p in enumerate(pos.cpu().tolist())]


def test_select_action_sampling_distribution_and_reproducibility():
    import torch
    from lightzero_b200.collect import select_action
    visits = torch.tensor([[30, 15, 5, 0]], dtype=torch.int32).cuda().repeat(40000, 1)
    pos, _, ent = select_action(visits, temperature=1.0, deterministic=False, seed=5, counter=1)
    freq = torch.bincount(pos.long(), minlength=4).double() / pos.numel()
    np.testing.assert_allclose(freq.cpu().numpy(), [0.6, 0.3, 0.1, 0.0], atol=0.01)
    assert float(ent[0]) == pytest.approx(co.select_action([30, 15, 5, 0])[1], rel=1e-12)
    again, _, _ = select_action(visits, temperature=1.0, deterministic=False, seed=5, counter=1)
    other, _, _ = select_action(visits, temperature=1.0, deterministic=False, seed=5, counter=2)
    assert torch.equal(pos, again) and not torch.equal(pos, other)
    # sharded batches draw the same numbers: the stream is keyed by the global root index
    part, _, _ = select_action(visits[100:200], temperature=1.0, deterministic=False, seed=5, counter=1, root_offset=100)
    assert torch.equal(part, pos[100:200])


@pytest.mark.parametrize("fixed", [True, False])
def test_target_policies_match_oracle(fixed):
    import torch
    from lightzero_b200.collect import legal_from_mask, target_policies
    rng = np.random.default_rng(12)
    N, A, S = 500, 18, 50
    mask = np.ones((N, A), np.int8) if fixed else random_masks(rng, N, A)
    legal = co.legal_from_mask(mask)
    dists = [rng.multinomial(S, rng.dirichlet([0.5] * len(l))).tolist() for l in legal]
    visits = np.zeros((N, A), np.int32)
    for i, d in enumerate(dists):
        visits[i, :len(d)] = d
    pm = (rng.random(N) < 0.8).astype(np.int32)
    cnt, ids = legal_from_mask(mask)
    got = target_policies(torch.from_numpy(visits).cuda(), cnt, None if fixed else ids, torch.from_numpy(pm).cuda())
    want = co.target_policies(dists, legal, A, policy_mask=pm, fixed_action_space=fixed)
    assert np.array_equal(got.cpu().numpy().view(np.uint32), want.view(np.uint32))


def test_dirichlet_noise_moments_and_support():
    import torch
    from lightzero_b200.collect import dirichlet_noise
    N, A, alpha = 20000, 18, 0.3
    cnt = torch.full((N,), A, dtype=torch.int32).cuda()
    cnt[: N // 2] = 6
    x = dirichlet_noise(alpha, cnt, A, seed=9, counter=1).double()
    assert torch.allclose(x.sum(1), torch.ones(N, dtype=torch.float64, device=x.device), atol=1e-5)
    assert (x[: N // 2, 6:] == 0).all() and (x >= 0).all()
    for rows, n in ((x[: N // 2, :6], 6), (x[N // 2:], A)):
        mean, var = rows.mean().item(), rows.var(0).mean().item()
        a0 = alpha * n
        assert mean == pytest.approx(1.0 / n, rel=1e-6)
        assert var == pytest.approx((1.0 / n) * (1 - 1.0 / n) / (a0 + 1), rel=0.05)  # Dirichlet marginal variance
        # marginal is Beta(alpha, alpha (n - 1)): check its third raw moment too
        m3 = (rows ** 3).mean().item()
        b = alpha * (n - 1)
        want3 = alpha * (alpha + 1) * (alpha + 2) / ((alpha + b) * (alpha + b + 1) * (alpha + b + 2))
        assert m3 == pytest.approx(want3, rel=0.08)
    assert torch.equal(x.float(), dirichlet_noise(alpha, cnt, A, seed=9, counter=1))
    assert not torch.equal(x.float(), dirichlet_noise(alpha, cnt, A, seed=9, counter=2))


def test_collect_round_trip_through_a_search():
    """mask -> legal lists -> noise -> Roots.prepare -> search -> select_action / targets, all device tensors."""
    import torch
    import lightzero_b200 as lzb
    from lightzero_b200.collect import dirichlet_noise, legal_from_mask, select_action, target_policies
    from tests.test_gpu_api import ScenarioModel, driver_cfg, identity
    from tests.common import Scenario
    rng = np.random.default_rng(13)
    sc = Scenario("mz", 64, 9, 30, seed=50)
    mask = random_masks(rng, sc.n, sc.A)
    dev = torch.device("cuda")
    cnt, ids = legal_from_mask(torch.from_numpy(mask).to(dev))
    noise = dirichlet_noise(0.3, cnt, sc.A, seed=1, counter=1)
    roots = lzb.MuZeroMCTSCtree.roots(sc.n, (cnt, ids))
    roots.prepare(0.25, noise, torch.from_numpy(sc.root_reward).to(dev), torch.from_numpy(sc.root_logits).to(dev),
                  torch.from_numpy(sc.to_play).to(dev))
    mcts = identity(lzb.MuZeroMCTSCtree(driver_cfg(sc)))
    mcts.search(roots, ScenarioModel(sc, dev), torch.zeros(sc.n, 16, device=dev), torch.from_numpy(sc.to_play).to(dev))
    visits, count = roots.get_distributions_tensor()
    assert torch.equal(count, cnt) and (visits.sum(1) == sc.S).all()
    pos, act, ent = select_action(visits, count, ids, 1.0, deterministic=True)
    assert (torch.from_numpy(mask).to(dev)[torch.arange(sc.n, device=dev), act.long()] == 1).all()  # chosen actions are legal
    pol = target_policies(visits, count, ids)
    assert torch.allclose(pol.sum(1), torch.ones(sc.n, device=dev)) and (pol[torch.from_numpy(mask).to(dev) == 0] == 0).all()
    lists = roots.get_distributions()
    assert [len(l) for l in lists] == cnt.cpu().tolist()
