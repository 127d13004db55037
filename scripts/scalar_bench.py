"""Tuning aid: device time of the fused inverse-scalar-transform kernel alone (lzb_inverse_scalar_transform, both heads per
launch) at the headline batch: 8192 rows x 2 heads x 601 bins, inputs rotated through a set larger than L2 so every launch
reads HBM.  Usage (GPU box): [LZB200_LIB=...] python scripts/scalar_bench.py [--rows 8192] [--support 300]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from lightzero_b200.tree_search.scaling import InverseScalarTransform  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=8192)
    ap.add_argument("--support", type=int, default=300)
    ap.add_argument("--sets", type=int, default=10)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--probe", action="store_true", help="also time torch reductions over the same tensors")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    W = 2 * a.support + 1
    g = torch.Generator(device=dev).manual_seed(7)
    val = [torch.randn(a.rows, W, generator=g, device=dev) for _ in range(a.sets)]
    rew = [torch.randn(a.rows, W, generator=g, device=dev) for _ in range(a.sets)]
    handle = InverseScalarTransform(a.support, dev, categorical_distribution=True)
    for s in range(a.sets):
        handle.pair(val[s], rew[s])
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()  # the launches back to back on the device: the eager loop is bound by the host (~25 us per call)
    with torch.cuda.graph(graph):
        for s in range(a.sets):
            handle.pair(val[s], rew[s])
    best = None
    for _ in range(a.reps):
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        graph.replay()
        t1.record()
        torch.cuda.synchronize()
        us = 1e3 * t0.elapsed_time(t1) / a.sets
        best = us if best is None else min(best, us)
    if a.probe:  # reference points on the same bytes: torch's row sum (two launches) and a whole-tensor max (two launches)
        for name, fn in (("torch.sum(dim=1) x2", lambda s: (val[s].sum(1), rew[s].sum(1))), ("torch.amax() x2", lambda s: (val[s].amax(), rew[s].amax())),
                         ("torch.softmax(dim=1) x2 (read + write)", lambda s: (torch.softmax(val[s], 1), torch.softmax(rew[s], 1)))):
            for s in range(a.sets):
                fn(s)
            torch.cuda.synchronize()
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2):
                for s in range(a.sets):
                    fn(s)
            b2 = None
            for _ in range(a.reps):
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                g2.replay()
                t1.record()
                torch.cuda.synchronize()
                us = 1e3 * t0.elapsed_time(t1) / a.sets
                b2 = us if b2 is None else min(b2, us)
            print("probe %-42s %.2f us per pair of heads" % (name, b2))
    nbytes = 2 * a.rows * W * 4 + 2 * a.rows * 4
    print("scalar transform: %.2f us per launch, %.1f MB, %.0f GB/s (%s)" % (best, nbytes / 1e6, nbytes / best / 1e3,
                                                                         os.environ.get("LZB200_LIB", "default build")))


if __name__ == "__main__":
    main()
