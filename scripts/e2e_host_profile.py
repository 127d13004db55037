"""Where the host time of the e2e leg goes: cProfile of one_search() with pinned host inputs (bench.py's e2e step)."""
import argparse
import cProfile
import pstats
import sys
import time

import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    import torch
    a = bench.parse_args()
    device = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    inp = bench.make_inputs(a, 0, device)
    model = bench.TableModel(inp, a.variant, a.actions, a.chance_space, calls_per_sim=2 if a.variant == "smz" else 1)
    mcts = bench.make_driver(a, device, graph=not a.no_graph)
    pin = lambda t: t.detach().cpu().pin_memory()
    host = {k: pin(inp[k]) for k in ("root_logits", "noise", "root_reward", "to_play", "latent_roots")}
    host["visits_out"] = torch.empty((a.roots, a.actions), dtype=torch.int32).pin_memory()
    host["values_out"] = torch.empty(a.roots, dtype=torch.float32).pin_memory()

    def step():
        bench.one_search(a, mcts, model, inp, host=host)
        torch.cuda.synchronize()

    for _ in range(5):
        step()
    t0 = time.perf_counter()
    for _ in range(20):
        step()
    print("e2e ms/step: %.3f" % ((time.perf_counter() - t0) / 20 * 1e3))
    # host time only (no sync inside): how long the host needs to enqueue one search
    t0 = time.perf_counter()
    for _ in range(20):
        bench.one_search(a, mcts, model, inp, host=host)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print("host enqueue ms/step: %.3f" % ((t1 - t0) / 20 * 1e3))
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(20):
        step()
    pr.disable()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(28)




def timeline():
    """Device-side timeline of one chunked e2e search: when each chunk's inputs are loaded and when its replay ends."""
    import torch
    from lightzero_b200.tree_search import mcts_ctree as mc
    a = bench.parse_args()
    device = torch.device("cuda", 0)
    inp = bench.make_inputs(a, 0, device)
    model = bench.TableModel(inp, a.variant, a.actions, a.chance_space)
    mcts = bench.make_driver(a, device, graph=True)
    pin = lambda t: t.detach().cpu().pin_memory()
    host = {k: pin(inp[k]) for k in ("root_logits", "noise", "root_reward", "to_play", "latent_roots")}
    host["visits_out"] = torch.empty((a.roots, a.actions), dtype=torch.int32).pin_memory()
    host["values_out"] = torch.empty(a.roots, dtype=torch.float32).pin_memory()
    for _ in range(5):
        bench.one_search(a, mcts, model, inp, host=host)
        torch.cuda.synchronize()
    marks = []
    orig_load, orig_replay = mc._SearchGraph._load, torch.cuda.CUDAGraph.replay

    def ev(tag):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        marks.append((tag, e, time.perf_counter()))

    def load(self, *args):
        ev("c%d load begin" % self.lo)
        orig_load(self, *args)
        ev("c%d load end" % self.lo)

    def replay(self):
        orig_replay(self)
        ev("replay end")

    mc._SearchGraph._load, torch.cuda.CUDAGraph.replay = load, replay
    t0 = torch.cuda.Event(enable_timing=True)
    t0.record()
    h0 = time.perf_counter()
    bench.one_search(a, mcts, model, inp, host=host)
    torch.cuda.synchronize()
    h1 = time.perf_counter()
    for tag, e, h in marks:
        print("%-22s device %.3f ms   host %.3f ms" % (tag, t0.elapsed_time(e), (h - h0) * 1e3))
    print("total host wall %.3f ms" % ((h1 - h0) * 1e3))


if __name__ == "__main__":
    if os.environ.get("TIMELINE"):
        timeline()
    else:
        main()
