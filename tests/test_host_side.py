"""CPU-side checks: the C-ABI library loads and exports every symbol include/lzb200.h declares (no compute calls),
host-side helpers of the reference-shaped API, and the multi-rank exchange over gloo (world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "lzb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lzb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from lightzero_b200 import _lib
    path = _lib.build()
    L = ctypes.CDLL(path)
    names = header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), "liblzb200.so does not export %s" % n
    assert sorted(_lib.exported_symbols()) == names, "ctypes prototypes and include/lzb200.h disagree"
    assert L.lzb_abi_version() == _lib.ABI_VERSION


def test_tiebreak_hash_host_matches_oracle():
    from lightzero_b200 import _lib
    from oracle.binding import oracle_lib
    L, O = _lib.lib(), oracle_lib()
    for seed, sim, root, depth in [(0, 0, 0, 0), (2024, 3, 17, 2), (2 ** 63 + 5, 49, 65535, 40), (7, 199, 8191, 0)]:
        assert L.lzb_tiebreak_hash(seed, sim, root, depth) == O.orc_tiebreak_hash(seed, sim, root, depth)
        assert L.lzb_tiebreak_hash(seed, sim, root, depth) < 2 ** 31


def test_argument_errors_are_reported_not_thrown():
    from lightzero_b200 import _lib
    L = _lib.lib()
    n = ctypes.c_size_t(0)
    assert L.lzb_forest_bytes(9, 8, 4, 10, 0, ctypes.byref(n)) == -1 and b"variant" in L.lzb_last_error()
    assert L.lzb_forest_bytes(0, 0, 4, 10, 0, ctypes.byref(n)) == -1
    assert L.lzb_forest_bytes(3, 8, 4, 10, 0, ctypes.byref(n)) == -1  # sampled tree needs action_dim
    assert L.lzb_forest_bytes(0, 8, 4, 70000, 0, ctypes.byref(n)) == -4
    assert L.lzb_forest_bytes(0, 8192, 18, 50, 0, ctypes.byref(n)) == 0 and n.value > 8192 * 51 * 18 * 16
    assert L.lzb_set_delta(None, 0.01, None) == -1


def test_gumbel_constants_match_oracle():
    from lightzero_b200.ctree.ctree_gumbel_muzero import gmz_tree
    from oracle.binding import oracle_lib
    O = oracle_lib()
    for m, S in ((3, 30), (6, 50), (16, 50), (4, 200)):
        want = np.zeros((m + 1, S), np.int32)
        O.orc_gumbel_table(m, S, want.ctypes.data_as(ctypes.c_void_p))
        assert gmz_tree.pget_table_of_considered_visits(m, S) == want.tolist()
        assert gmz_tree.pget_sequence_of_considered_visits(m, S) == want[m].tolist()
    want = np.zeros(40, np.float32)
    O.orc_gumbel_noise(ctypes.c_float(10.0), 40, want.ctypes.data_as(ctypes.c_void_p))
    assert np.array_equal(np.asarray(gmz_tree.pgenerate_gumbel(10.0, 0.0, 40), np.float32), want)


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from lightzero_b200 import MuZeroMCTSCtree
    with pytest.raises(RuntimeError, match="CUDA"):
        MuZeroMCTSCtree.roots(4, [[0, 1]] * 4)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under lightzero_b200/ may import, load or link it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "lightzero_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+\.*oracle", src, flags=re.M), f
                assert "liboracle" not in src and "oracle/_ref" not in src and "libref_" not in src, f


def test_rows_to_padded_and_attrdict():
    from lightzero_b200.ctree._tree_api import _rows_to_padded
    from lightzero_b200.tree_search.mcts_ctree import AttrDict
    out = _rows_to_padded([[1.0, 2.0], [3.0], []], 3, np.float32)
    assert out.tolist() == [[1, 2, 0], [3, 0, 0], [0, 0, 0]]
    assert _rows_to_padded([[1, 2], [3, 4]], 2, np.float32).shape == (2, 2)
    c = AttrDict(dict(a=1, model=dict(support_scale=300)))
    c.update(dict(model=dict(categorical_distribution=False), b=2))
    assert c.model.support_scale == 300 and c.model.categorical_distribution is False and c.b == 2


def test_inverse_scalar_transform_matches_formula():
    from lightzero_b200.tree_search.scaling import InverseScalarTransform
    t = InverseScalarTransform(5, "cpu", True)
    logits = torch.randn(7, 11)
    v = (torch.softmax(logits, 1) * torch.arange(-5, 6).float()).sum(1, keepdim=True)
    eps = 0.001
    want = torch.sign(v) * (((torch.sqrt(1 + 4 * eps * (v.abs() + 1 + eps)) - 1) / (2 * eps)) ** 2 - 1)
    assert torch.allclose(t(logits), want, atol=1e-6)
    probs = torch.softmax(logits, 1)
    assert torch.allclose(t(probs), want, atol=1e-5)  # already-normalised rows are not soft-maxed again
    s = InverseScalarTransform(5, "cpu", False)
    assert torch.allclose(s(v), want, atol=1e-6)


def test_sampler_shapes_and_support():
    from lightzero_b200.tree_search.sampling import sample_actions
    g = torch.Generator().manual_seed(0)
    a = sample_actions(torch.cat([torch.zeros(5, 2), torch.ones(5, 2)], 1), 2, 6, True, generator=g)
    assert a.shape == (5, 6, 2) and float(a.abs().max()) < 1.0
    d = sample_actions(torch.randn(4, 9), 9, 5, False, generator=g)
    assert d.shape == (4, 5, 1)
    for row in d[:, :, 0].long().tolist():
        assert len(set(row)) == 5 and all(0 <= x < 9 for x in row)  # without replacement


def test_shard_bounds_cover_the_batch():
    from lightzero_b200.sharding import shard_bounds
    for n, w in ((8192, 8), (10, 4), (3, 8), (65536, 2)):
        b = [shard_bounds(n, w, r) for r in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        assert max(h - l for l, h in b) - min(h - l for l, h in b) <= 1


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
from lightzero_b200.sharding import gather_search_outputs, shard_bounds
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
full_v = torch.arange(10 * 4, dtype=torch.int32).reshape(10, 4)
full_x = torch.arange(10, dtype=torch.float32) * 0.5
for n in (10, 8):                       # ragged (10 over 2 is even; use 3 ranks' worth of raggedness via n=... below) and even
    counts = [shard_bounds(n, w, k)[1] - shard_bounds(n, w, k)[0] for k in range(w)]
    lo, hi = shard_bounds(n, w, r)
    v, x = gather_search_outputs(full_v[lo:hi], full_x[lo:hi], counts)
    assert torch.equal(v, full_v[:n]) and torch.equal(x, full_x[:n]), (r, n)
counts = [7, 3]                          # explicitly ragged
lo = sum(counts[:r]); hi = lo + counts[r]
v, x = gather_search_outputs(full_v[lo:hi], full_x[lo:hi], counts)
assert torch.equal(v, full_v) and torch.equal(x, full_x)
# extras travel in the same exchange: Gumbel policies [n, A] f32, sampled action tables [n, K, D] f32
pol = torch.rand(10, 4); tab = torch.rand(10, 3, 2)
v, x, p2, t2 = gather_search_outputs(full_v[lo:hi], full_x[lo:hi], counts, extras=(pol[lo:hi], tab[lo:hi]))
torch.manual_seed(0)
assert torch.equal(v, full_v) and torch.equal(x, full_x) and p2.shape == pol.shape and t2.shape == tab.shape
assert torch.equal(p2[lo:hi], pol[lo:hi]) and torch.equal(t2[lo:hi], tab[lo:hi]) and p2.dtype == torch.float32
# a 64-bit extra cannot be packed with the 32-bit pieces: falls back to one collective per tensor
v, x, q = gather_search_outputs(full_v[lo:hi], full_x[lo:hi], counts, extras=(full_x[lo:hi].double(),))
assert torch.equal(q, full_x.double())
dist.destroy_process_group()
print("rank", r, "ok")
"""


def test_end_of_search_exchange_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER % ROOT)
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29631", str(script)],
                         capture_output=True, text=True, timeout=240)
    assert res.returncode == 0, res.stdout + res.stderr
    assert res.stdout.count("ok") == 2


def test_list_marshalling_helper_matches_numpy():
    """lzb_pylist.c (the C loops behind list-of-lists arguments and get_distributions) against the numpy walk."""
    from lightzero_b200 import _lib
    from lightzero_b200.ctree import _tree_api as api
    P = _lib.pylist()
    assert P is not None, "the list helper did not build (gcc + Python.h are in this image)"
    rng = np.random.default_rng(3)
    a = rng.standard_normal((257, 18)).astype(np.float32)
    rows = [r[: i % 19] for i, r in enumerate(a.tolist())]  # ragged, including empty rows
    got = api._rows_to_padded(rows, 18, np.float32)
    want = np.zeros_like(a)
    for i, r in enumerate(rows):
        want[i, : len(r)] = r
    assert got.dtype == np.float32 and np.array_equal(got, want)
    assert np.array_equal(api._rows_to_padded([(1, 2.5), [3], ()], 3, np.float32), np.array([[1, 2.5, 0], [3, 0, 0], [0, 0, 0]], np.float32))
    with pytest.raises(ValueError):
        api._rows_to_padded([[1.0, 2.0, 3.0, 4.0]], 3, np.float32)  # longer than the action space
    with pytest.raises(TypeError):
        api._rows_to_padded([[1.0, "x"]], 3, np.float32)
    v = rng.integers(0, 1000, (129, 9)).astype(np.int32)
    c = rng.integers(0, 10, 129).astype(np.int32)
    out = P.lzb_rows_from_i32(v.ctypes.data, c.ctypes.data, 129, 9)
    assert out == [r[:k] for r, k in zip(v.tolist(), c.tolist())] and all(type(x) is int for r in out for x in r)


def test_list_marshalling_parallel_paths_and_int_lists():
    """The threaded fast paths of lzb_pylist.c (lists of lists of exact floats / small ints, large enough to be split over
    threads) against numpy, the serial fallbacks for every other shape, and the int readers behind legal-action lists and
    to_play (Roots(root_num, legal_actions_list), lzero/policy/muzero.py:671,680)."""
    import gc
    from lightzero_b200 import _lib
    from lightzero_b200.forest import _floats_from_lists, _ints_from_list
    P = _lib.pylist()
    _list_marshalling_checks(P, _floats_from_lists, _ints_from_list, gc)
    os.environ["LZB_PYLIST_THREADS"] = "4"  # the opt-in threaded readers (read per call)
    try:
        _list_marshalling_checks(P, _floats_from_lists, _ints_from_list, gc)
    finally:
        del os.environ["LZB_PYLIST_THREADS"]


def _list_marshalling_checks(P, _floats_from_lists, _ints_from_list, gc):
    rng = np.random.default_rng(5)
    N, A = 4096, 18
    x = rng.standard_normal((N, A)).astype(np.float32)
    L = x.tolist()
    assert np.array_equal(_floats_from_lists(L), x)                       # threaded
    assert np.array_equal(_floats_from_lists([tuple(r) for r in L]), x)  # serial: tuples
    mixed = [list(r) for r in L]
    mixed[77][3] = 2                                                      # an int among the floats: serial fallback
    got = _floats_from_lists(mixed)
    assert got[77, 3] == 2.0 and np.array_equal(got[78], x[78])
    ragged = [r[: 1 + i % A] for i, r in enumerate(L)]
    out, lens = np.empty((N, A), np.float32), np.empty(N, np.int32)
    assert P.lzb_rows_to_f32(ragged, out.ctypes.data, N, A, lens.ctypes.data) == 0
    assert lens.tolist() == [len(r) for r in ragged]
    assert all(np.array_equal(out[i, :len(r)], np.asarray(r, np.float32)) and (out[i, len(r):] == 0).all() for i, r in enumerate(ragged))
    # legal-action lists
    legal = [sorted(rng.choice(A, size=1 + i % A, replace=False).tolist()) for i in range(N)]
    idx, cnt = np.full((N, A), -7, np.int32), np.empty(N, np.int32)
    assert P.lzb_rows_to_i32(legal, idx.ctypes.data, N, A, cnt.ctypes.data) == 0     # threaded
    assert cnt.tolist() == [len(l) for l in legal]
    assert all(idx[i, :len(l)].tolist() == l and (idx[i, len(l):] == 0).all() for i, l in enumerate(legal))
    idx2, cnt2 = np.empty((N, A), np.int32), np.empty(N, np.int32)
    assert P.lzb_rows_to_i32([tuple(l) for l in legal], idx2.ctypes.data, N, A, cnt2.ctypes.data) == 0  # serial
    assert np.array_equal(idx, idx2) and np.array_equal(cnt, cnt2)
    floats = [[float(v) for v in l] for l in legal]                      # a float mask's .tolist(): integers stored as floats
    assert P.lzb_rows_to_i32(floats, idx2.ctypes.data, N, A, cnt2.ctypes.data) == 0 and np.array_equal(idx, idx2)
    with pytest.raises(ValueError):
        P.lzb_rows_to_i32([list(range(A + 1))] * N, idx2.ctypes.data, N, A, cnt2.ctypes.data)
    assert _ints_from_list([-1] * 100).tolist() == [-1] * 100 and _ints_from_list([1, 2] * 50).dtype == np.int32
    # building the output lists pauses the cyclic collector and puts it back the way it was
    v = rng.integers(0, 60, (N, A)).astype(np.int32)
    for enabled in (True, False):
        (gc.enable if enabled else gc.disable)()
        try:
            rows = P.lzb_rows_from_i32(v.ctypes.data, cnt.ctypes.data, N, A)
            assert gc.isenabled() == enabled
        finally:
            gc.enable()
        assert rows[5] == v[5, :cnt[5]].tolist()


def test_every_experiment_flag_of_the_index_exists_in_the_sources():
    """EXPERIMENTS.md lists each measured experiment with the build flag that re-creates it: every -DLZB_* flag named there
    must still be tested for somewhere under csrc/ (scripts/check_flags.sh compiles them)."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "EXPERIMENTS.md")).read()
    flags = set(re.findall(r"-D(LZB_[A-Z0-9_]+)", text))
    assert len(flags) >= 25, sorted(flags)
    src = ""
    csrc = os.path.join(root, "lightzero_b200", "csrc")
    for f in os.listdir(csrc):
        src += open(os.path.join(csrc, f)).read()
    missing = sorted(f for f in flags if f not in src and not any(f.startswith(x) and x in src for x in ("LZB_X_GATHER_", "LZB_X_ROW_PREFETCH")))
    assert not missing, missing


def test_strong_scaling_flag_splits_the_total_over_the_ranks(monkeypatch):
    """bench.py --scaling strong: --roots-total roots divided over WORLD_SIZE ranks (strong-scaling row)."""
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench
    monkeypatch.setenv("WORLD_SIZE", "8")
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "8", "--scaling", "strong"])
    a = bench.parse_args()
    assert a.roots == 8192 and a.scaling == "strong" and bench.workload_config(a, 8)["roots_total"] == 65536
    monkeypatch.setenv("WORLD_SIZE", "1")
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    a = bench.parse_args()
    assert a.roots == 8192 and a.scaling == "weak"
