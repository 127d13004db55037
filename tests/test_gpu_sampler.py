"""GPU: the Sampled-EfficientZero action sampler kernel (csrc/lzb_sample.cuh).  The reference's sampler is seeded from
the wall clock (sez cnode.cpp:251,336), so parity is statistical: distribution tests, plus exact checks of the
arithmetic with injected uniforms."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_continuous_matches_tanh_gaussian_distribution():
    import torch
    from scipy import stats
    from lightzero_b200.tree_search.sampling import sample_actions_cuda
    N, K, D = 4096, 8, 2
    mu, sg = np.array([0.3, -0.5], np.float32), np.array([0.7, 1.2], np.float32)
    logits = torch.tensor(np.tile(np.concatenate([mu, sg]), (N, 1))).cuda()
    a = sample_actions_cuda(logits, D, K, True, seed=11, counter=3).cpu().numpy()
    assert a.shape == (N, K, D) and np.abs(a).max() <= 1.0
    for d in range(D):
        x = np.clip(a[:, :, d].ravel().astype(np.float64), -1 + 1e-7, 1 - 1e-7)
        cdf = lambda v, d=d: stats.norm.cdf((np.arctanh(v) - mu[d]) / sg[d])
        assert stats.kstest(x, cdf).pvalue > 1e-3
    # draws of different roots / slots are uncorrelated
    assert abs(np.corrcoef(a[:-1, 0, 0], a[1:, 0, 0])[0, 1]) < 0.05 and abs(np.corrcoef(a[:, 0, 0], a[:, 1, 0])[0, 1]) < 0.05


def test_continuous_arithmetic_with_injected_uniforms():
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_cuda
    g = torch.Generator().manual_seed(0)
    N, K, D = 64, 5, 3
    logits = torch.cat([torch.randn(N, D, generator=g), torch.rand(N, D, generator=g) + 0.1], 1)
    u = torch.rand(N, K, D, 2, generator=g).clamp(1e-6, 1 - 1e-6)
    got = sample_actions_cuda(logits.cuda(), D, K, True, uniforms=u.cuda()).cpu()
    z = torch.sqrt(-2 * torch.log(u[..., 0])) * torch.cos(2 * np.pi * u[..., 1])
    want = torch.tanh(logits[:, None, :D] + logits[:, None, D:] * z)
    assert torch.allclose(got, want, atol=2e-5)


def test_discrete_top_k_with_injected_uniforms_is_the_reference_rule():
    """sez cnode.cpp:340-397: p = exp(l) / (sum exp(l) + 1e-6); keys u^(1/p) sorted descending; first K."""
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_cuda
    g = torch.Generator().manual_seed(1)
    N, A, K = 128, 9, 5
    logits = torch.randn(N, A, generator=g)
    u = torch.rand(N, A, generator=g).clamp(1e-6, 1 - 1e-6)
    got = sample_actions_cuda(logits.cuda(), A, K, False, uniforms=u.cuda()).cpu()
    assert got.shape == (N, K, 1)
    e = torch.exp(logits.double())
    p = e / (e.sum(1, keepdim=True) + 1e-6)
    keys = torch.pow(u.double(), 1.0 / p)
    want = torch.argsort(keys, dim=1, descending=True)[:, :K]
    assert torch.equal(got[:, :, 0].long(), want)


def test_discrete_first_pick_follows_the_policy_and_never_repeats():
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_cuda
    N, A, K = 20000, 6, 4
    l = torch.tensor([0.0, 1.0, -1.0, 0.5, 2.0, -2.0])
    got = sample_actions_cuda(l.repeat(N, 1).cuda(), A, K, False, seed=5).cpu()[:, :, 0].long()
    assert all(len(set(r)) == K for r in got[:500].tolist())
    freq = torch.bincount(got[:, 0], minlength=A).double() / N
    p = torch.softmax(l.double(), 0)
    assert float((freq - p).abs().max()) < 0.015  # first of the u^(1/p) order is distributed as p


def test_draws_are_counter_based():
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_cuda
    logits = torch.randn(32, 8).cuda()
    a = sample_actions_cuda(logits, 4, 6, True, seed=9, counter=2)
    assert torch.equal(a, sample_actions_cuda(logits, 4, 6, True, seed=9, counter=2))
    assert not torch.equal(a, sample_actions_cuda(logits, 4, 6, True, seed=9, counter=3))
    assert not torch.equal(a, sample_actions_cuda(logits, 4, 6, True, seed=10, counter=2))
    # sharding: roots [16, 32) drawn as a shard with root_offset 16 equal the same roots of the full batch
    assert torch.equal(a[16:], sample_actions_cuda(logits[16:], 4, 6, True, seed=9, counter=2, root_offset=16))


# ------------------------------------------------------------------------------------------------ parity sampler (a15)
def test_device_tanhf_matches_the_oracle_restatement():
    """glibc_tanhf on the device == oracle/glibc_flt32.c:orc_tanhf (itself pinned against the live libm) bit for bit,
    through the parity sampler: sigma = 0 makes every draw equal mu, so out = tanhf(mu)."""
    import ctypes as C
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_parity
    from oracle.binding import oracle_lib
    rng = np.random.default_rng(3)
    xs = np.concatenate([rng.standard_normal(20000).astype(np.float32) * 3, np.linspace(-25, 25, 5001, dtype=np.float32),
                         np.array([0.0, 1e-10, -1e-10, 0.34, 0.35, 1.0, -1.0, 1.04, 21.9, 22.0, 22.1, 88.0, -88.0], np.float32)])
    D = len(xs)
    # one root: mu = xs, sigma = 0 (draw * 0 + mu == mu for every mu but -0.0, which is therefore not among the inputs)
    logits = np.concatenate([xs, np.zeros(D, np.float32)])[None, :]
    got = sample_actions_parity(torch.from_numpy(logits).cuda(), D, 1, True, seed_base=1).cpu().numpy().reshape(-1)
    want = np.empty_like(xs)
    O = oracle_lib()
    O.orc_tanhf_array.argtypes = [C.c_void_p, C.c_void_p, C.c_long]
    O.orc_tanhf_array(xs.ctypes.data, want.ctypes.data, D)
    bad = got.view(np.uint32) != want.view(np.uint32)
    assert not bad.any(), "%d of %d differ, e.g. tanhf(%r): device %r, glibc restatement %r" % (bad.sum(), D, xs[bad][0], got[bad][0], want[bad][0])


@pytest.mark.parametrize("name", ["sez_sampler_continuous_k20_d6", "sez_sampler_discrete_k4_of_6", "sez_sampler_discrete_k5_of_18"])
def test_cuda_sampler_reproduces_the_reference_tables_and_search(name):
    """Row a15 closed: the CUDA sampler itself - not an injected table - reproduces the action tables the compiled,
    unmodified reference tree sampled under a pinned clock (tests/golden/sez_sampler_*.npz), bit for bit, and the search
    run on the CUDA-sampled tables matches the reference's leaf indices, last actions, visit counts and root values."""
    import os
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_parity
    from tests.common import Scenario, assert_same, sez_logits
    from tests.test_gpu_tree_parity import run_gpu
    from tests.test_sampler_oracle import SAMPLER_CASES
    kw, continuous = SAMPLER_CASES[name]
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    sc = Scenario(**kw)
    clock = int(g["clock_base"])
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    root = sample_actions_parity(t(sez_logits(sc.root_logits, continuous)), sc.A, sc.K, continuous, seed_base=clock)
    sims = torch.stack([sample_actions_parity(t(sez_logits(sc.logits[s], continuous)), sc.A, sc.K, continuous,
                                              seed_base=clock + (s + 1) * sc.n) for s in range(sc.S)])
    assert_same(root.cpu().numpy(), g["root_actions"], name + ": root action tables")
    assert_same(sims.cpu().numpy(), g["actions"], name + ": per-simulation action tables")
    # the search on the tables the CUDA sampler drew
    sc.root_actions, sc.actions = root.cpu().numpy(), sims.cpu().numpy()
    rec = []
    gpu = run_gpu(sc, 1, 2024, record=rec, fused=True)
    for k in ("ix", "search_len", "vtp", "last_action_f", "is_reset"):
        assert_same(np.stack([r[k] for r in rec]), g[k], "%s/%s" % (name, k))
    out = gpu.root_outputs(sc.discount)
    assert_same(out["visits"].cpu().numpy(), np.where(g["dist"] < 0, 0, g["dist"]), name + "/visit counts")
    assert_same(out["values"].cpu().numpy(), g["values"], name + "/root values")


def test_discrete_sampler_never_repeats_an_action_when_probabilities_underflow():
    """K larger than the number of actions with a usable key (masked / extreme logits: p underflows to 0, or log(u) / p
    overflows to -inf): the table still holds K DISTINCT action ids, the finite keys first."""
    import torch
    from lightzero_b200.tree_search.sampling import sample_actions_cuda
    N, A, K = 64, 40, 12
    logits = torch.full((N, A), -200.0)           # exp(-200 - max) underflows: p == 0 for all but three actions
    logits[:, 5], logits[:, 17], logits[:, 33] = 3.0, 2.0, 2.5
    for counter in range(4):
        a = sample_actions_cuda(logits.cuda(), A, K, False, seed=9, counter=counter).cpu().numpy()[:, :, 0].astype(int)
        assert a.shape == (N, K) and a.min() >= 0 and a.max() < A
        for row in a:
            assert len(set(row.tolist())) == K, row
            assert set(row[:3].tolist()) == {5, 17, 33}
            assert row[3:].tolist() == sorted(row[3:].tolist()), "zero-probability actions fill the tail lowest id first"
