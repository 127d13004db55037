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
