"""Reanalyze-shaped measurement: 256 segments x (5 unroll steps + 1) = 1,536 roots x 50 sims,
A = 6 (Atari Pong), latent 1024 f32: callers.reanalyze_policy_targets + reanalyze_value_targets on the device (search as
a replayed CUDA graph) next to the compiled reference tree + the reference's Python target loops on one host core.
Prints one JSON line.  GPU box:  python scripts/reanalyze_bench.py [--segments 256]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--segments", type=int, default=256)
    ap.add_argument("--unroll", type=int, default=5)
    ap.add_argument("--steps", type=int, default=20)
    x = ap.parse_args()
    import torch
    from lightzero_b200 import callers
    from oracle import reanalyze_oracle as ro
    B, U = x.segments, x.unroll
    T = B * (U + 1)
    sys.argv = ["bench.py", "--roots", str(T), "--actions", "6", "--sims", "50", "--latent", "1024"]
    a = bench.parse_args()
    dev = torch.device("cuda", 0)
    inp = bench.make_inputs(a, 0, dev)
    model = bench.TableModel(inp, "mz", a.actions)
    mcts = bench.make_driver(a, dev, graph=True)
    rng = np.random.default_rng(0)
    mask = torch.ones((T, a.actions), dtype=torch.int8, device=dev)
    pmask = torch.as_tensor((rng.random(T) < 0.9).astype(np.int32)).to(dev)
    seg_len = rng.integers(20, 400, B)
    pos = np.asarray([rng.integers(0, l) for l in seg_len])
    rewards = [rng.standard_normal(l + 10).astype(np.float32) for l in seg_len]
    rew_pad = np.zeros((B, max(len(r) for r in rewards)), np.float32)
    for b, r in enumerate(rewards):
        rew_pad[b, :len(r)] = r
    td = rng.integers(1, 6, T).astype(np.int32)
    vmask = np.ones(T, np.float32)
    dv = lambda z: torch.as_tensor(z).to(dev)
    rew_d, rl_d, sl_d, pos_d, td_d, vm_d = dv(rew_pad), dv(np.asarray([len(r) for r in rewards], np.int32)), dv(seg_len.astype(np.int32)), dv(pos.astype(np.int32)), dv(td), dv(vmask)

    def step():
        out = callers.reanalyze_policy_targets(mcts, model, inp["latent_roots"], inp["root_reward"], inp["root_logits"], mask,
                                               inp["to_play"], pmask, U)
        tr, tv = callers.reanalyze_value_targets(out["root_values"], td, vm_d, rew_d, rl_d, sl_d, pos_d, U, bench.DISCOUNT)
        return out["target_policies"], tv

    for _ in range(3):
        pol, tv = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(x.steps):
        pol, tv = step()
    e1.record()
    torch.cuda.synchronize()
    wall_ms = 1e3 * (time.perf_counter() - t0) / x.steps
    dev_ms = e0.elapsed_time(e1) / x.steps
    # reference side: compiled reference tree (one core) + the reference's target loops in Python
    d = bench.cpu_inputs(a, T, 1234)
    bench.cpu_search(a, d)
    t0 = time.perf_counter()
    dt, kind, dist, vals = bench.cpu_search(a, d, want_values=True)
    legal = [list(range(a.actions))] * T
    ro.policy_targets([r.tolist() for r in dist], legal, pmask.cpu().tolist(), B, U, a.actions)
    ro.value_targets(vals.tolist(), td.tolist(), vmask.tolist(), rewards, seg_len.tolist(), pos.tolist(), [[-1] * len(r) for r in rewards], U, bench.DISCOUNT)
    ref_ms = 1e3 * (time.perf_counter() - t0)
    print(json.dumps({"workload": "reanalyze: %d segments x %d = %d roots x 50 sims, A=6, latent 1024 f32, device-resident inputs" % (B, U + 1, T),
                      "b200_ms_per_batch_wall": round(wall_ms, 3), "b200_ms_per_batch_device": round(dev_ms, 3),
                      "b200_sims_per_s": round(T * 50 / (wall_ms * 1e-3), 1),
                      "reference_ms_per_batch_1core": round(ref_ms, 2), "reference_kind": kind,
                      "reference_sims_per_s_1core": round(T * 50 / (ref_ms * 1e-3), 1), "ratio_vs_1core": round(ref_ms / wall_ms, 1)}))


if __name__ == "__main__":
    main()
