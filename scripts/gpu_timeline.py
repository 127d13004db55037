"""Tuning aid: in-kernel phase timeline of the fused step kernel at the headline shape.

Builds a second copy of the library with -DLZB_TIMELINE (lightzero_b200/liblzb200_tl.so, not the shipped one), runs
eager searches (8192 roots x 50 sims x A=18, latent 1024) and prints, per phase, the mean / p50 / p99 / max cycles a
warp spends in it, plus the spread of warp start and end times per launch.  Usage (GPU box):
    python scripts/gpu_timeline.py [--latent 1024] [--roots 8192] [--build-only]
"""
import argparse
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TL_LIB = os.path.join(ROOT, "lightzero_b200", "liblzb200_tl.so")


def build_tl():
    from lightzero_b200 import _lib
    extra = os.environ.get("LZB_TL_FLAGS", "").split()  # e.g. LZB_TL_FLAGS="-DLZB_X_CARRY": the timeline of an experiment build
    cmd = ["nvcc"] + _lib.NVCC_FLAGS + ["-DLZB_TIMELINE"] + extra + [os.path.join(_lib.CSRC, "lzb_api.cu"), "-o", TL_LIB]
    subprocess.run(cmd, check=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--latent", type=int, default=1024)
    ap.add_argument("--roots", type=int, default=8192)
    ap.add_argument("--sims", type=int, default=50)
    ap.add_argument("--variant", default="mz", help="mz | ez (the pUCT step kernels carry the stamps)")
    ap.add_argument("--build-only", action="store_true")
    ap.add_argument("--no-save", action="store_true", help="print the summary only (the raw stamps are ~40 MB)")
    ap.add_argument("--analyse", default=None, help="only analyse a saved .npy")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "timeline.npy"))
    a = ap.parse_args()
    if a.analyse:
        import numpy as np
        analyse(np.load(a.analyse), a.sims)
        return
    if a.build_only or not os.path.exists(TL_LIB):
        build_tl()
    if a.build_only:
        return
    os.environ["LZB200_LIB"] = TL_LIB
    import ctypes as C
    import numpy as np
    import torch
    import bench
    from lightzero_b200 import _lib
    L = _lib.lib()
    L.lzb_debug_timeline.restype = C.c_int
    L.lzb_debug_timeline.argtypes = [C.c_void_p]
    sys.argv = ["bench.py", "--roots", str(a.roots), "--sims", str(a.sims), "--latent", str(a.latent), "--no-graph",
                "--variant", a.variant]
    ba = bench.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    inp = bench.make_inputs(ba, 0, dev)
    if ba.variant in bench.HIDDEN:
        inp["hidden_c"] = torch.zeros(1, a.roots, bench.HIDDEN[ba.variant], device=dev)
        inp["hidden_h"] = torch.zeros(1, a.roots, bench.HIDDEN[ba.variant], device=dev)
    mcts = bench.make_driver(ba, dev, False)
    model = bench.TableModel(inp, ba.variant, ba.actions)
    for _ in range(3):
        bench.one_search(ba, mcts, model, inp)
    torch.cuda.synchronize()
    nwarp = ((a.roots + 7) // 8) * 4  # >= warps of any lane-group launch (G >= 8: 128 / G trees per 4-warp block)
    buf = torch.zeros((a.sims + 2) * nwarp * 24, dtype=torch.int64, device=dev)
    assert L.lzb_debug_timeline(C.c_void_p(buf.data_ptr())) == 0
    bench.one_search(ba, mcts, model, inp)
    torch.cuda.synchronize()
    assert L.lzb_debug_timeline(C.c_void_p(0)) == 0
    t = buf.cpu().numpy().astype(np.uint64)
    if not a.no_save:
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        np.save(a.out, t)
    analyse(t, a.sims)


def analyse(t, sims):
    import numpy as np
    # the grid of the step launches is not known here: find it from the non-zero rows of launch 2
    flat = t.reshape(-1, 24)
    nz = np.nonzero(flat[:, 0])[0]
    print("stamped warp rows:", len(nz))
    # rows are [launch][warp]; warps per launch = stride between first rows of consecutive launches
    first = nz[0]
    gaps = np.diff(nz)
    big = gaps[gaps > 1]
    W = int(first)  # the first step launch has latent_index 1 and stamps its warp 0: row = 1 * W + 0
    print("assumed warps per launch:", W, "first row", first, "gaps>1:", big[:5])
    rows = flat[nz].astype(np.int64)
    launch = nz // W
    label = {2: "pdl_wait done", 16: "x: inputs arrived, logits max", 20: "x: expf done", 17: "x: ordered sum done", 3: "x: expansion stored",
             18: "b: path records arrived", 19: "b: recurrence done", 21: "b: records updated", 4: "backup done (MinMax)",
             5: "select loop entry", 6: "L1: block in registers", 7: "L1: mean_q", 8: "L1: scores+max", 9: "L1: pick", 10: "L1: done",
             11: "all levels done", 12: "before the row gather", 13: "rows moved", 14: "end"}
    for lo, hi, name in ((2, 6, "sims 2-5"), (20, 30, "sims 20-29"), (40, sims, "sims 40+")):
        sel = rows[(launch >= lo) & (launch < hi)]
        if not len(sel):
            continue
        print("== %s: %d warp-launches; total start->end cycles mean %.0f p50 %.0f p99 %.0f max %.0f" % (
            (name, len(sel)) + stats(sel[:, 14] - sel[:, 0])))
        lvl = (sel[:, 15] >> 16) & 0xffff
        print("   levels walked by the warp: mean %.2f max %d" % (lvl.mean(), lvl.max()))
        for which, m0 in (("all warps", np.ones(len(sel), bool)), ("slowest 1%", (sel[:, 14] - sel[:, 0]) >= np.percentile(sel[:, 14] - sel[:, 0], 99))):
            a = sel[m0]
            off = {k: (a[:, k] - a[:, 0]).mean() for k in label if (a[:, k] > 0).all()}
            print("   -- %s (n=%d): mean cycles since entry" % (which, len(a)))
            prev = 0
            for k, v in sorted(off.items(), key=lambda kv: kv[1]):
                print("      %-28s %7.0f  (+%.0f)" % (label[k], v, v - prev))
                prev = v
    # the row gather as a function of WHEN a warp reaches it: how many warps start in each 2 k-cycle window after kernel entry
    # and how long their gather takes - a gather that is bandwidth-bound in the crowded window shows as durations that follow
    # the count, a latency-bound one as a flat line
    sel = rows[(launch >= 20) & (launch < 30)]
    if len(sel) and (sel[:, 12] > 0).all() and (sel[:, 13] > 0).all():
        start, dur = sel[:, 12] - sel[:, 0], sel[:, 13] - sel[:, 12]
        print("== row gather by start window (sims 20-29): window | warps (share) | mean gather cycles | p90")
        for lo in range(0, 44000, 2000):
            m = (start >= lo) & (start < lo + 2000)
            if m.sum() >= 20:
                print("   %5d-%5d | %6d (%4.1f%%) | %6.0f | %6.0f" % (lo, lo + 2000, m.sum(), 100.0 * m.mean(), dur[m].mean(), np.percentile(dur[m], 90)))
    # per launch: spread of starts / ends by globaltimer is not available per phase; use slot 1 (globaltimer at entry)
    for l in (3, 25, 45):
        sel = rows[launch == l]
        if len(sel):
            g = sel[:, 1]
            print("launch %d: globaltimer entry spread %.2f us over %d warps" % (l, (g.max() - g.min()) / 1e3, len(sel)))


def stats(d):
    import numpy as np
    return (float(np.mean(d)), float(np.percentile(d, 50)), float(np.percentile(d, 99)), float(np.max(d)))


if __name__ == "__main__":
    main()
