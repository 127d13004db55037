"""The action sampler of the Sampled-EfficientZero tree as a device-side seam.

The reference draws K actions per expanded node inside CNode::expand from a wall-clock-seeded
std::default_random_engine (ctree_sampled_efficientzero/lib/cnode.cpp:251,259-283 continuous; :336-406 discrete),
so its draws are not reproducible and cannot be pinned bit-wise.  The DISTRIBUTIONS are restated
here on the device with torch's counter-based generator; the resulting table [N, K, D] is what the tree kernels
take as input (and what parity tests inject into the oracle).
"""
import torch


def sample_actions(policy_logits, action_space_size, num_of_sampled_actions, continuous_action_space, generator=None):
    """policy_logits [N, 2D] (mu | sigma) for continuous, [N, A] for discrete.  Returns float32 [N, K, D] (D=1 discrete).

    continuous: a = tanh(x), x ~ Normal(mu, sigma) per dimension (sez cnode.cpp:262-271).
    discrete:   K actions without replacement, keys U^(1/p) sorted descending with p = exp(l) / (sum exp(l) + 1e-6)
                (sez cnode.cpp:340-397); K is clipped to A by the caller's configuration as in the reference."""
    x = policy_logits.float()
    N, K = x.shape[0], int(num_of_sampled_actions)
    if continuous_action_space:
        D = x.shape[1] // 2
        mu, sigma = x[:, :D], x[:, D:2 * D]
        eps = torch.randn((N, K, D), device=x.device, dtype=torch.float32, generator=generator)
        return torch.tanh(mu.unsqueeze(1) + sigma.unsqueeze(1) * eps).contiguous()
    e = torch.exp(x)
    p = e / (e.sum(1, keepdim=True) + 1e-6)
    u = torch.rand(x.shape, device=x.device, dtype=torch.float64, generator=generator)
    keys = torch.pow(u, 1.0 / p.double())
    idx = torch.argsort(keys, dim=1, descending=True)[:, :K]
    return idx.to(torch.float32).unsqueeze(-1).contiguous()
