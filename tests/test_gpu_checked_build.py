"""The checked build's surface (profiles/r02_sanitizer.md): lzb_check_word reads 0 after a search with the shipped library;
with a -DLZB_CHECKED library (scripts/gpu_checked.sh sets LZB200_LIB and LZB_CHECKED=1) a deliberately unregistered latent
slot is REPORTED (check 11) instead of read, which shows the checks are live in that build."""
import numpy as np
import pytest
import torch

from lightzero_b200.forest import Forest

pytestmark = pytest.mark.gpu


def _searched_forest(register_root_slot):
    dev = torch.device("cuda", 0)
    N, A, S, L = 64, 6, 4, 32
    g = torch.Generator(device=dev).manual_seed(3)
    f = Forest("mz", N, A, S, device=dev)
    tp = torch.full((N,), -1, dtype=torch.int32, device=dev)
    f.prepare(torch.randn(N, A, generator=g, device=dev), None, 0.0, torch.zeros(N, device=dev), tp)
    lat0 = torch.randn(N, L, generator=g, device=dev)
    if register_root_slot:
        f.set_slots(0, [lat0])
    dst = torch.zeros(N, L, device=dev)
    cfg = f.make_cfg(rng_seed=5)
    f.traverse(cfg, tp, gathers=[(None, dst)])
    torch.cuda.synchronize()
    return f, lat0, dst


def test_check_word_is_zero_after_a_clean_search():
    f, lat0, dst = _searched_forest(True)
    word, checked = f.check_word()
    assert word == 0
    assert torch.equal(dst, lat0)  # every first walk ends at the root: its row is what the gather delivers


def test_checked_build_reports_an_unregistered_slot():
    f, lat0, dst = _searched_forest(True)
    if not f.check_word()[1]:
        pytest.skip("the loaded library is not a -DLZB_CHECKED build (scripts/gpu_checked.sh runs this test with one)")
    f2, _, dst2 = _searched_forest(False)
    word, _ = f2.check_word()
    assert word == 11, word
    assert float(dst2.abs().sum()) == 0.0  # reported, not read


def test_nvtx_ranges_do_not_change_a_search():
    """LZB_NVTX=1 wraps every forest call in an NVTX range (lzb.prepare, lzb.traverse, lzb.step[e], lzb.backprop[e],
    lzb.root_outputs) for nsys / ncu --nvtx timelines; a search under it gives the same visit counts."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import torch\n"
        "from lightzero_b200 import forest as F\n"
        "assert F._NVTX == bool(__import__('os').environ.get('LZB_NVTX'))\n"
        "dev = torch.device('cuda', 0); N, A, S = 32, 5, 6\n"
        "g = torch.Generator(device=dev).manual_seed(1)\n"
        "f = F.Forest('mz', N, A, S, device=dev)\n"
        "tp = torch.full((N,), -1, dtype=torch.int32, device=dev)\n"
        "f.prepare(torch.randn(N, A, generator=g, device=dev), None, 0.0, torch.zeros(N, device=dev), tp)\n"
        "cfg = f.make_cfg(rng_seed=3)\n"
        "o = f.traverse(cfg, tp)\n"
        "for e in range(1, S + 1):\n"
        "    r, v, l = (torch.randn(N, generator=g, device=dev) * 0.1, torch.randn(N, generator=g, device=dev), torch.randn(N, A, generator=g, device=dev))\n"
        "    if e < S: o = f.step(cfg, e, 0.997, r, v, l, o['vtp'], tp, out=o)\n"
        "    else: f.backprop(e, 0.997, r, v, l, o['vtp'])\n"
        "print(f.root_outputs()['visits'].cpu().numpy().tolist())\n" % root)
    outs = []
    for flag in ("", "1"):
        env = dict(os.environ, LZB_NVTX=flag)
        if not flag:
            env.pop("LZB_NVTX")
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout.strip().splitlines()[-1])
    assert outs[0] == outs[1] and outs[0].startswith("[[")
