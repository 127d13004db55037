"""The action sampler of the Sampled-EfficientZero tree as a device-side seam.

The reference draws K actions per expanded node inside CNode::expand from a wall-clock-seeded
std::default_random_engine (ctree_sampled_efficientzero/lib/cnode.cpp:251,259-283 continuous; :336-406 discrete),
so its draws are not reproducible and cannot be pinned bit-wise.  The DISTRIBUTIONS are restated
here on the device with torch's counter-based generator; the resulting table [N, K, D] is what the tree kernels
take as input (and what parity tests inject into the oracle).
"""
import ctypes as C
import itertools

import torch

_search_counter = itertools.count(1)


def new_sample_seed(base=0):
    """A fresh sampler seed per search: the reference re-seeds from the clock at every expand, i.e. every search sees
    new draws; here (base seed, search ordinal) does the same reproducibly."""
    return ((int(base) & 0xffffffff) << 32) | (next(_search_counter) & 0xffffffff)


def sample_actions_cuda(policy_logits, action_space_size, num_of_sampled_actions, continuous_action_space, seed=0, counter=0,
                        root_offset=0, uniforms=None):
    """The hand-written sampler kernel (csrc/lzb_sample.cuh) through the C ABI: float32 [N, K, D] on the logits' device."""
    from .. import _lib
    x = policy_logits.float().contiguous()
    N, K = x.shape[0], int(num_of_sampled_actions)
    D = x.shape[1] // 2 if continuous_action_space else 1
    out = torch.empty((N, K, D), dtype=torch.float32, device=x.device)
    u = None if uniforms is None else uniforms.float().contiguous()
    seed_dev = seed if isinstance(seed, torch.Tensor) else None  # int64 device scalar: read by the kernel at run time
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().lzb_sample_actions_seeded(
            C.c_void_p(x.data_ptr()), None if u is None else C.c_void_p(u.data_ptr()), C.c_void_p(out.data_ptr()), N, K, D,
            x.shape[1], int(bool(continuous_action_space)), int(root_offset),
            C.c_uint64(0 if seed_dev is not None else int(seed)), None if seed_dev is None else C.c_void_p(seed_dev.data_ptr()),
            C.c_uint32(int(counter)), C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)))
    return out


def sample_actions_parity(policy_logits, action_space_size, num_of_sampled_actions, continuous_action_space, seed_base):
    """The reference's own draws (libstdc++ minstd_rand0 + normal_distribution<float> / uniform_real<double>, C ABI
    lzb_sample_actions_libstdcxx): root i is seeded with (unsigned)(seed_base + i), what the reference's expand() takes from
    a wall clock reading seed_base + i ns.  A test / reproducibility mode (one thread per root)."""
    from .. import _lib
    x = policy_logits.float().contiguous()
    N, K = x.shape[0], int(num_of_sampled_actions)
    D = x.shape[1] // 2 if continuous_action_space else 1
    out = torch.empty((N, K, D), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().lzb_sample_actions_libstdcxx(
            C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), N, K, D, x.shape[1], int(bool(continuous_action_space)),
            C.c_uint64(int(seed_base) & 0xffffffffffffffff), C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)))
    return out


def sample_actions(policy_logits, action_space_size, num_of_sampled_actions, continuous_action_space, generator=None,
                   seed=None, counter=0, root_offset=0):
    """policy_logits [N, 2D] (mu | sigma) for continuous, [N, A] for discrete.  Returns float32 [N, K, D] (D=1 discrete).

    continuous: a = tanh(x), x ~ Normal(mu, sigma) per dimension (sez cnode.cpp:262-271).
    discrete:   K actions without replacement, keys U^(1/p) sorted descending with p = exp(l) / (sum exp(l) + 1e-6)
                (sez cnode.cpp:340-397); K is clipped to A by the caller's configuration as in the reference."""
    x = policy_logits.float()
    if x.is_cuda and generator is None:
        return sample_actions_cuda(x, action_space_size, num_of_sampled_actions, continuous_action_space,
                                   seed=0 if seed is None else seed, counter=counter, root_offset=root_offset)
    # torch restatement of the same distributions (CPU tensors; used by the statistical tests as the comparison)
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
