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
