"""GPU: lzb_inverse_scalar_transform_static - the no-allocation flavour of the inverse scalar transform (scratch in a
file-scope device array, include/lzb200.h) - against the existing entry point lzb_inverse_scalar_transform, through the
C ABI, on two calls in a row.  Named test_zz_* so that pytest collects it last: a failure here must not hide other tests
under `pytest -x`.  Passes alone in a fresh process and after the rest of the suite (it keeps no state of its own; the
static array is shared by design and every call clears the words that must start at zero).

Tolerance: the one tests/test_scaling.py uses for this transform (1e-5 relative + 1e-3 absolute); both entry points run
the same kernel, so the outputs are in fact expected to be identical."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 1e-3


def _heads(_lib, pairs):
    arr = (_lib.SupportHead * len(pairs))()
    for j, (x, out) in enumerate(pairs):
        arr[j] = _lib.SupportHead(x.data_ptr(), x.stride(0), out.data_ptr())
    return arr


def _close(got, want, what):
    got, want = got.cpu().numpy().reshape(-1), want.cpu().numpy().reshape(-1)
    assert np.isfinite(got).all(), what
    bad = ~(np.abs(got - want) <= ATOL + RTOL * np.abs(want))
    assert not bad.any(), "%s: %d of %d beyond tolerance, e.g. %r vs %r" % (what, bad.sum(), bad.size, got[bad][0], want[bad][0])


@pytest.mark.parametrize("rows,support", [(8192, 300), (777, 10), (64, 400), (65536, 10)])
def test_static_scratch_entry_point_matches_the_workspace_one(rows, support):
    import torch
    from lightzero_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    W = 2 * support + 1
    g = torch.Generator(device=dev).manual_seed(1000 + rows)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    nbytes = C.c_size_t()
    _lib.check(L.lzb_scalar_workspace_bytes(2, rows, C.byref(nbytes)))
    ws = torch.zeros(nbytes.value, dtype=torch.uint8, device=dev)
    for call in range(2):  # two calls in a row: the second finds the static array as the first left it
        logits = torch.randn(rows, W, generator=g, device=dev) * 2          # raw logits: the soft-max branch
        probs = torch.softmax(torch.randn(rows, W, generator=g, device=dev), dim=1)  # rows sum to 1: taken as they are
        want = [torch.empty(rows, device=dev), torch.empty(rows, device=dev)]
        got = [torch.full((rows,), float("nan"), device=dev), torch.full((rows,), float("nan"), device=dev)]
        # the second call swaps the heads, so that stale "un-normalised row" counters of call 1 would show
        ins = (logits, probs) if call == 0 else (probs, logits)
        _lib.check(L.lzb_inverse_scalar_transform(_heads(_lib, list(zip(ins, want))), 2, rows, support, 0.001,
                                                  C.c_void_p(ws.data_ptr()), stream))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        _lib.check(L.lzb_inverse_scalar_transform_static(_heads(_lib, list(zip(ins, got))), 2, rows, support, 0.001, stream))
        torch.cuda.synchronize()
        if call == 1:  # (the first call on a device resolves the symbol; nothing may be allocated after that)
            assert torch.cuda.mem_get_info(dev)[0] == free_before, "the static entry point changed the device's free memory"
        for h in range(2):
            _close(got[h], want[h], "call %d head %d" % (call, h))
    # one head only, twice more on the same array
    for call in range(2):
        x = torch.randn(rows, W, generator=g, device=dev)
        want, got = torch.empty(rows, device=dev), torch.full((rows,), float("nan"), device=dev)
        _lib.check(L.lzb_inverse_scalar_transform(_heads(_lib, [(x, want)]), 1, rows, support, 0.001, C.c_void_p(ws.data_ptr()), stream))
        _lib.check(L.lzb_inverse_scalar_transform_static(_heads(_lib, [(x, got)]), 1, rows, support, 0.001, stream))
        torch.cuda.synchronize()
        _close(got, want, "single head, call %d" % call)


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    import torch
    from lightzero_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    x = torch.zeros(4, 21, device=dev)
    out = torch.zeros(4, device=dev)
    heads = _heads(_lib, [(x, out)])
    limit = 65536  # LZB_STATIC_SCALAR_MAX_ROWS (include/lzb200.h)
    assert L.lzb_inverse_scalar_transform_static(heads, 1, limit + 1, 10, 0.001, None) == -1
    assert b"LZB_STATIC_SCALAR_MAX_ROWS" in L.lzb_last_error()
    assert L.lzb_inverse_scalar_transform_static(heads, 3, 4, 10, 0.001, None) == -1
    assert L.lzb_inverse_scalar_transform_static(None, 1, 4, 10, 0.001, None) == -1
    assert L.lzb_inverse_scalar_transform_static(heads, 1, 4, 11, 0.001, None) == -1  # row stride 21 < 2 * 11 + 1
    torch.cuda.synchronize()
