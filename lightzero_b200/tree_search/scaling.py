"""Device-side h^-1 of the reference driver: what lzero/policy/scaling_transform.py:97-128 (InverseScalarTransform,
with ensure_softmax :36-62) computes, kept on the model's device and free of host syncs so value / reward never
leave HBM between recurrent_inference and the expand+backup kernel.

Categorical heads on a CUDA device go through ONE hand-written kernel (csrc/lzb_scaling.cuh, C ABI
``lzb_inverse_scalar_transform``): softmax -> expectation over the support -> h^-1, both heads of a simulation in one
launch, every logit read once.  Scalar heads (``categorical_distribution=False``) and CPU tensors take the few torch
ops below."""
import ctypes as C

import torch


class InverseScalarTransform:
    def __init__(self, support_size, device="cpu", categorical_distribution=True):
        self.support_size = int(support_size)
        self.value_support = torch.arange(-support_size, support_size + 1, dtype=torch.float32, device=device).unsqueeze(0)
        self.categorical_distribution = categorical_distribution
        self._ws = {}  # (device, rows, heads) -> zeroed workspace of the kernel

    # ------------------------------------------------------------------ the CUDA path
    def _kernel_ok(self, *heads):
        return self.categorical_distribution and all(
            isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.dim() == 2
            and t.shape[1] == 2 * self.support_size + 1 and t.stride(1) == 1 and t.shape[0] == heads[0].shape[0]
            for t in heads)

    def _launch(self, heads, epsilon):
        from .. import _lib
        L = _lib.lib()
        dev, n, k = heads[0].device, heads[0].shape[0], len(heads)
        key = (dev, n, k)
        ws = self._ws.get(key)
        if ws is None:
            nbytes = C.c_size_t()
            _lib.check(L.lzb_scalar_workspace_bytes(k, n, C.byref(nbytes)))
            ws = self._ws[key] = torch.zeros(nbytes.value, dtype=torch.uint8, device=dev)
        outs = [torch.empty((n, 1), dtype=torch.float32, device=dev) for _ in heads]
        arr = (_lib.SupportHead * k)()
        for j, (t, o) in enumerate(zip(heads, outs)):
            arr[j] = _lib.SupportHead(t.data_ptr(), t.stride(0), o.data_ptr())
        with torch.cuda.device(dev):
            _lib.check(L.lzb_inverse_scalar_transform(arr, k, n, self.support_size, float(epsilon), C.c_void_p(ws.data_ptr()),
                                                      C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
        return outs

    def pair(self, value_logits, reward_logits, epsilon=0.001):
        """Both heads of one simulation in ONE launch -> (value [N, 1], reward [N, 1])."""
        if self._kernel_ok(value_logits, reward_logits):
            return tuple(self._launch([value_logits, reward_logits], epsilon))
        return self(value_logits, epsilon), self(reward_logits, epsilon)

    # ------------------------------------------------------------------ reference surface
    def __call__(self, logits, epsilon=0.001):
        if self._kernel_ok(logits):
            return self._launch([logits], epsilon)[0]
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
