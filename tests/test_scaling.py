"""InverseScalarTransform (lzero/policy/scaling_transform.py:36-62,97-128): the numpy oracle and the host fallback
against the golden vectors made by the reference's own class (tests/golden/make_scaling_golden.py), then - on the GPU -
the fused CUDA kernel (csrc/lzb_scaling.cuh through lzb_inverse_scalar_transform) against the same vectors.

Tolerance: 1e-5 RELATIVE plus 1e-3 absolute.  This transform is floating point with no pinned summation order (torch's
own reductions), so it is not held to bits.  The absolute term is the reference's OWN fp32 noise: its expectation sums
601 products of magnitude up to 300 in float32 (one rounding of such a term is 300 * 2^-24 = 1.8e-5; the golden vectors
differ from an exact fp64 evaluation by up to 7e-5 in the expectation), and h^-1 has slope 1 .. 130 there, so outputs of
magnitude < 100 cannot be pinned more tightly than ~1e-3 absolute against the reference itself."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "inverse_scalar_transform.npz"))
CASES = sorted({k.split("/")[0] for k in GOLD.files if k.endswith("/logits")})
RTOL, ATOL = 1e-5, 1e-3


def close(got, want, what):
    got, want = np.asarray(got, np.float32).reshape(-1), np.asarray(want, np.float32).reshape(-1)
    assert got.shape == want.shape, what
    bad = ~(np.abs(got - want) <= ATOL + RTOL * np.abs(want))
    assert not bad.any(), "%s: %d of %d beyond tolerance, e.g. got %r want %r" % (what, bad.sum(), bad.size, got[bad][0], want[bad][0])


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_golden(name):
    from oracle.scaling_oracle import inverse_scalar_transform
    close(inverse_scalar_transform(GOLD[name + "/logits"], int(GOLD[name + "/support"])), GOLD[name + "/out"], name)


def test_oracle_scalar_head_matches_reference_golden():
    from oracle.scaling_oracle import inverse_scalar_transform
    close(inverse_scalar_transform(GOLD["scalar/in"], 300, categorical_distribution=False), GOLD["scalar/out"], "scalar")


@pytest.mark.parametrize("name", CASES)
def test_host_fallback_matches_reference_golden(name):
    import torch
    from lightzero_b200.tree_search.scaling import InverseScalarTransform
    t = InverseScalarTransform(int(GOLD[name + "/support"]), "cpu", True)
    close(t(torch.from_numpy(GOLD[name + "/logits"].copy())).numpy(), GOLD[name + "/out"], name)


def test_oracle_matches_reference_class_directly():
    """Where /root/reference is present (the authoring container): the reference class itself on fresh random input."""
    ref = os.path.join(os.environ.get("LZB_REFERENCE_ROOT", "/root/reference"), "lzero", "policy", "scaling_transform.py")
    if not os.path.exists(ref):
        pytest.skip("reference tree not present on this machine; the golden vectors stand in")
    import importlib.util
    import torch
    from oracle.scaling_oracle import inverse_scalar_transform
    spec = importlib.util.spec_from_file_location("ref_scaling_transform", ref)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    rng = np.random.default_rng(3)
    for support in (5, 50, 300):
        x = (rng.standard_normal((40, 2 * support + 1)) * 2).astype(np.float32)
        want = mod.InverseScalarTransform(support, "cpu", True)(torch.from_numpy(x.copy())).numpy()
        close(inverse_scalar_transform(x, support), want, "support %d" % support)


# ------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_kernel_matches_reference_golden(name):
    import torch
    from lightzero_b200.tree_search.scaling import InverseScalarTransform
    dev = torch.device("cuda", 0)
    t = InverseScalarTransform(int(GOLD[name + "/support"]), dev, True)
    x = torch.from_numpy(GOLD[name + "/logits"].copy()).to(dev)
    assert t._kernel_ok(x)
    for _ in range(2):  # the second call finds the workspace as the first one left it
        y = t(x)
        assert tuple(y.shape) == (x.shape[0], 1)
        close(y.cpu().numpy(), GOLD[name + "/out"], name)


@pytest.mark.gpu
def test_cuda_kernel_two_heads_one_launch_and_mixed_normalisation():
    """pair(): value and reward heads in one launch; ensure_softmax is decided PER HEAD - one head holds probabilities
    (taken as they are), the other raw logits (soft-maxed) - and a head with one un-normalised row is soft-maxed whole."""
    import torch
    from lightzero_b200.tree_search.scaling import InverseScalarTransform
    dev = torch.device("cuda", 0)
    t = InverseScalarTransform(300, dev, True)
    n = 16
    probs = torch.from_numpy(GOLD["probabilities_s300/logits"][:n].copy()).to(dev)
    raw = torch.from_numpy(GOLD["atari_s300/logits"][:n].copy()).to(dev)
    for _ in range(3):
        v, r = t.pair(probs, raw)
        close(v.cpu().numpy(), GOLD["probabilities_s300/out"][:n], "probabilities head")
        close(r.cpu().numpy(), GOLD["atari_s300/out"][:n], "logits head")
        v, r = t.pair(raw, probs)
        close(v.cpu().numpy(), GOLD["atari_s300/out"][:n], "logits head first")
        close(r.cpu().numpy(), GOLD["probabilities_s300/out"][:n], "probabilities head second")
    t10 = InverseScalarTransform(10, dev, True)
    x = torch.from_numpy(GOLD["one_row_unnormalised_s10/logits"].copy()).to(dev)
    close(t10(x).cpu().numpy(), GOLD["one_row_unnormalised_s10/out"], "one row off")


@pytest.mark.gpu
def test_cuda_kernel_headline_batch_strided_rows_and_wide_support():
    """8192 rows x 601 bins (every block of the grid finishes a share; the last one runs the ensure_softmax epilogue)
    against the oracle; rows taken from a wider tensor (row stride > width); a support wider than the register path."""
    import torch
    from lightzero_b200.tree_search.scaling import InverseScalarTransform
    from oracle.scaling_oracle import inverse_scalar_transform
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(11)
    x = torch.randn(8192, 601, generator=g, device=dev) * 2
    t = InverseScalarTransform(300, dev, True)
    v, r = t.pair(x, x.flip(1).contiguous())
    close(v.cpu().numpy(), inverse_scalar_transform(x.cpu().numpy(), 300), "8192 x 601")
    close(r.cpu().numpy(), inverse_scalar_transform(x.flip(1).cpu().numpy(), 300), "8192 x 601 flipped")
    big = torch.randn(64, 700, generator=g, device=dev)
    view = big[:, :601]  # stride(0) = 700
    assert t._kernel_ok(view)
    close(t(view).cpu().numpy(), inverse_scalar_transform(view.cpu().numpy(), 300), "strided rows")
    wide = torch.randn(32, 2 * 400 + 1, generator=g, device=dev)
    close(InverseScalarTransform(400, dev, True)(wide).cpu().numpy(), inverse_scalar_transform(wide.cpu().numpy(), 400), "support 400")
    probs = torch.softmax(torch.randn(8192, 601, generator=g, device=dev), dim=1)
    close(t(probs).cpu().numpy(), inverse_scalar_transform(probs.cpu().numpy(), 300), "8192 probability rows")
