"""Host-only probe of the list marshalling helpers (no GPU work): per-piece milliseconds at 8192 x 18."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightzero_b200 import _lib
P = _lib.pylist()
N, A = 8192, 18
rng = np.random.default_rng(0)
x = rng.standard_normal((N, A)).astype(np.float32)
L, Lt = x.tolist(), [tuple(r) for r in x.tolist()]
legal, legal_t = [list(range(A)) for _ in range(N)], [tuple(range(A)) for _ in range(N)]
out, idx, cnt = np.empty((N, A), np.float32), np.empty((N, A), np.int32), np.empty(N, np.int32)
vis = rng.integers(0, 50, (N, A)).astype(np.int32); full = np.full(N, A, np.int32)
def T(f, n=30):
    f(); t = time.perf_counter()
    for _ in range(n): f()
    return (time.perf_counter() - t) / n * 1e3
print("cpus", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
print("floats threaded %.3f  serial(tuples) %.3f" % (T(lambda: P.lzb_rows_to_f32(L, out.ctypes.data, N, A, None)), T(lambda: P.lzb_rows_to_f32(Lt, out.ctypes.data, N, A, None))))
print("ints   threaded %.3f  serial(tuples) %.3f" % (T(lambda: P.lzb_rows_to_i32(legal, idx.ctypes.data, N, A, cnt.ctypes.data)), T(lambda: P.lzb_rows_to_i32(legal_t, idx.ctypes.data, N, A, cnt.ctypes.data))))
print("rows_from_i32 %.3f" % T(lambda: P.lzb_rows_from_i32(vis.ctypes.data, full.ctypes.data, N, A)))
