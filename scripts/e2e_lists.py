"""What the unmodified-policy call pattern costs: one search call with Python LISTS in and out (lzero/policy/muzero.py:
672-719 calls .tolist() before prepare and reads lists back), next to the same call with host arrays.  GPU box only."""
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
    mcts = bench.make_driver(a, dev, graph=not a.no_graph, streams=a.e2e_streams)
    host = {k: inp[k].cpu() for k in ("root_logits", "noise", "root_reward", "to_play", "latent_roots")}
    lists = {k: host[k].tolist() for k in ("root_logits", "noise", "root_reward", "to_play")}
    lat = host["latent_roots"].numpy()

    def call(as_lists):
        roots = mcts.roots(a.roots, None)
        if as_lists:
            roots.prepare(bench.NOISE_W, lists["noise"], lists["root_reward"], lists["root_logits"], lists["to_play"])
            mcts.search(roots, model, lat, lists["to_play"])
            d, v = roots.get_distributions(), roots.get_values()
            return len(d), len(v)
        roots.prepare(bench.NOISE_W, host["noise"], host["root_reward"], host["root_logits"], host["to_play"])
        mcts.search(roots, model, lat, host["to_play"])
        d, _ = roots.get_distributions_tensor()
        return d.cpu(), roots.get_values_tensor().cpu()

    for as_lists in (False, True):
        for _ in range(3):
            call(as_lists)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            call(as_lists)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / 10 * 1e3
        print("%s in/out: %.2f ms per search call = %.1f M simulations/s" % ("lists" if as_lists else "host arrays", ms, a.roots * a.sims / ms / 1e3))


if __name__ == "__main__":
    main()
