"""Device-side h^-1 of the reference driver: what lzero/policy/scaling_transform.py:97-128 (InverseScalarTransform,
with ensure_softmax :36-62) computes, kept on the model's device and free of host syncs so value / reward never
leave HBM between recurrent_inference and the expand+backup kernel."""
import torch


class InverseScalarTransform:
    def __init__(self, support_size, device="cpu", categorical_distribution=True):
        self.value_support = torch.arange(-support_size, support_size + 1, dtype=torch.float32, device=device).unsqueeze(0)
        self.categorical_distribution = categorical_distribution

    def __call__(self, logits, epsilon=0.001):
        if self.categorical_distribution:
            # ensure_softmax: rows that already sum to 1 (allclose, atol 1e-5 + rtol 1e-5) are taken as probabilities;
            # the branch is a device-side select instead of the reference's host-synchronising `if`.
            total = logits.sum(dim=1, keepdim=True)
            normalized = ((total - 1).abs() <= 2e-5).all()
            probs = torch.where(normalized, logits, torch.softmax(logits, dim=1))
            value = (probs * self.value_support.to(probs.device)).sum(1, keepdim=True)
        else:
            value = logits
        tmp = (torch.sqrt(1 + 4 * epsilon * (torch.abs(value) + 1 + epsilon)) - 1) / (2 * epsilon)
        return torch.sign(value) * (tmp * tmp - 1)
