"""TEST INFRASTRUCTURE ONLY: numpy restatement of the reference's InverseScalarTransform
(/root/reference/lzero/policy/scaling_transform.py:97-128) with ensure_softmax (:36-62).

Parity: PINNED against tests/golden/inverse_scalar_transform.npz, which tests/golden/make_scaling_golden.py produced
by calling the reference class itself (loaded from its file).  Floating point, not bit-pinned: torch fixes no
summation order, so the bar is 1e-5 relative (stated in the tests)."""
import numpy as np


def ensure_softmax(logits):
    """scaling_transform.py:36-62: ONE decision for the whole batch - torch.allclose(row sums, 1, atol=1e-5) with the
    default rtol=1e-5 - then softmax over dim 1 or the rows as they are."""
    x = np.asarray(logits, np.float32)
    sums = x.sum(axis=1, dtype=np.float32)
    if np.all(np.abs(sums - 1.0) <= 1e-5 + 1e-5 * 1.0):
        return x
    z = x - x.max(axis=1, keepdims=True)
    e = np.exp(z.astype(np.float64))
    return (e / e.sum(axis=1, keepdims=True)).astype(np.float32)


def inverse_scalar_transform(logits, support_size, epsilon=0.001, categorical_distribution=True):
    """scaling_transform.py:118-128: expectation over the support {-s .. s}, then h^-1."""
    if categorical_distribution:
        probs = ensure_softmax(logits)
        support = np.arange(-support_size, support_size + 1, dtype=np.float64)[None, :]
        value = (probs.astype(np.float64) * support).sum(axis=1, keepdims=True).astype(np.float32)
    else:
        value = np.asarray(logits, np.float32)
    eps = np.float32(epsilon)
    tmp = (np.sqrt(np.float32(1) + np.float32(4) * eps * (np.abs(value) + np.float32(1) + eps)) - np.float32(1)) / (np.float32(2) * eps)
    return (np.sign(value) * (tmp * tmp - np.float32(1))).astype(np.float32)
