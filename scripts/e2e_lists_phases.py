"""Tuning aid: where one lists-in / lists-out search call spends its wall time (bench.py's e2e_lists leg), phase by phase,
with a device synchronisation after each phase (so overlap is lost: the sum exceeds the real call).  GPU box only."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    import torch
    a = bench.parse_args()
    dev = torch.device("cuda", 0)
    inp = bench.make_inputs(a, 0, dev)
    model = bench.TableModel(inp, a.variant, a.actions)
    mcts = bench.make_driver(a, dev, graph=True, streams=a.e2e_streams)
    N, A = a.roots, a.actions
    cpu = lambda t: t.detach().cpu()
    L = dict(noise=cpu(inp["noise"]).tolist(), reward=cpu(inp["root_reward"]).tolist(), logits=cpu(inp["root_logits"]).tolist(),
             to_play=cpu(inp["to_play"]).tolist(), legal=[list(range(A)) for _ in range(N)], latent=cpu(inp["latent_roots"]).numpy())
    acc = {}

    def phase(name, fn, sync=True):
        t0 = time.perf_counter()
        r = fn()
        if sync:
            torch.cuda.synchronize()
        acc[name] = acc.get(name, 0.0) + time.perf_counter() - t0
        return r

    for it in range(13):
        if it == 3:
            acc.clear()
        roots = phase("roots()", lambda: mcts.roots(N, L["legal"]))
        phase("prepare()", lambda: roots.prepare(bench.NOISE_W, L["noise"], L["reward"], L["logits"], L["to_play"]))
        phase("search()", lambda: mcts.search(roots, model, L["latent"], L["to_play"]))
        phase("get_distributions()", lambda: roots.get_distributions())
        phase("get_values()", lambda: roots.get_values())
    tot = sum(acc.values())
    for k, v in acc.items():
        print("%-22s %.3f ms" % (k, v / 10 * 1e3))
    print("%-22s %.3f ms (phases synchronised)" % ("sum", tot / 10 * 1e3))
    t0 = time.perf_counter()
    for _ in range(10):
        roots = mcts.roots(N, L["legal"])
        roots.prepare(bench.NOISE_W, L["noise"], L["reward"], L["logits"], L["to_play"])
        mcts.search(roots, model, L["latent"], L["to_play"])
        roots.get_distributions(), roots.get_values()
    torch.cuda.synchronize()
    print("%-22s %.3f ms" % ("whole call", (time.perf_counter() - t0) / 10 * 1e3))
    import cProfile
    import pstats
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(5):
        roots = mcts.roots(N, L["legal"])
        roots.prepare(bench.NOISE_W, L["noise"], L["reward"], L["logits"], L["to_play"])
        mcts.search(roots, model, L["latent"], L["to_play"])
        roots.get_distributions(), roots.get_values()
    pr.disable()
    pstats.Stats(pr).sort_stats("tottime").print_stats(18)


if __name__ == "__main__":
    main()
