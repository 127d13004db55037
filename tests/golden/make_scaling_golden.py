"""Generates tests/golden/inverse_scalar_transform.npz from the REFERENCE's own InverseScalarTransform
(/root/reference/lzero/policy/scaling_transform.py:97-128, loaded by file path: the module imports only torch and
numpy).  Run in the authoring container only; the fixture travels, /root/reference does not.

    python tests/golden/make_scaling_golden.py
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("LZB_REFERENCE_ROOT", "/root/reference")


def reference_module():
    spec = importlib.util.spec_from_file_location("ref_scaling_transform", os.path.join(REF, "lzero", "policy", "scaling_transform.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def cases():
    rng = np.random.default_rng(20260511)
    f = lambda *s: rng.standard_normal(s).astype(np.float32)
    out = {}
    out["atari_s300"] = (300, f(64, 601) * 3.0)                      # every shipped Atari config: support 300 -> 601 bins
    out["small_s10"] = (10, f(32, 21))
    out["ez_s50"] = (50, f(24, 101) * 2.0)
    p = torch.softmax(torch.from_numpy(f(16, 601) * 4.0), dim=1).numpy()
    out["probabilities_s300"] = (300, p.astype(np.float32))           # rows already sum to 1: ensure_softmax's else branch
    peaked = f(16, 601)
    peaked[np.arange(16), rng.integers(0, 601, 16)] += 40.0           # one dominant bin, some at the far ends
    peaked[0, 0] += 80.0
    peaked[1, 600] += 80.0
    out["peaked_s300"] = (300, peaked)
    mixed = torch.softmax(torch.from_numpy(f(8, 21)), dim=1).numpy().astype(np.float32)
    mixed[3] *= 1.5                                                    # ONE row off: the whole batch is soft-maxed again
    out["one_row_unnormalised_s10"] = (10, mixed)
    out["zeros_s300"] = (300, np.zeros((4, 601), np.float32))
    return out


def main():
    mod = reference_module()
    data = {}
    for name, (support, x) in cases().items():
        t = mod.InverseScalarTransform(support, "cpu", True)
        y = t(torch.from_numpy(x.copy())).numpy().astype(np.float32)
        data[name + "/support"] = np.int32(support)
        data[name + "/logits"] = x
        data[name + "/out"] = y
    rng = np.random.default_rng(7)
    v = np.concatenate([rng.standard_normal(60).astype(np.float32) * 50, np.array([0.0, -0.0, 300.0, -300.0], np.float32)])
    t = mod.InverseScalarTransform(300, "cpu", False)
    data["scalar/in"] = v.reshape(-1, 1)
    data["scalar/out"] = t(torch.from_numpy(v.reshape(-1, 1).copy())).numpy().astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "inverse_scalar_transform.npz"), **data)
    print("wrote", len(data), "arrays")


if __name__ == "__main__":
    main()
