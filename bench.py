#!/usr/bin/env python
"""Headline benchmark: MCTS simulations/sec of the batched tree search (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--roots 8192] [--sims 50] [--actions 18] [--latent 1024]
    python bench.py --impl reference ...      # the reference ctree on the host cores (oracle/_ref), same JSON line

A "step" is ONE whole search: Roots -> prepare (Dirichlet noise as an input tensor) -> S x [select + leaf
gather -> synthetic model (fixed table lookup, no NN time) -> latent store -> expand + backup] -> root visit
counts + values.  `value` = roots * sims / step time with every input resident in HBM (the S rounds replayed
as one CUDA graph); `e2e` = the same search through the reference-shaped public API
(lightzero_b200.MuZeroMCTSCtree.search) with HOST inputs in pinned memory and the results read back to the host
inside the timed region.  Multi-GPU: roots shard by env batch, one forest per rank, no collective inside the
search loop, one exchange of visit counts / values at the end of each search: the NVLink peer-memory
kernel of csrc/lzb_peer.cuh (--nccl-exchange: one packed NCCL all-gather instead); every rank's gathered rows are checked
against an NCCL all-gather of the shards (`exchange_check`).  `--scaling strong` splits --roots-total over the ranks.
Every N=1 line also carries `parity`: the CUDA search against the compiled reference at the bench shape, bit for bit.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

PB_C_BASE, PB_C_INIT, DISCOUNT, DELTA, NOISE_W, ALPHA = 19652, 1.25, 0.997, 0.01, 0.25, 0.3
TIE_SEED = 2024
L2_FLUSH_BYTES = 512 << 20
HIDDEN = {"ez": 512, "sez": 256}  # LSTM reward-hidden-state width of the EfficientZero flavours


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--roots", type=int, default=8192, help="roots PER GPU (weak scaling)")
    ap.add_argument("--sims", type=int, default=50)
    ap.add_argument("--actions", type=int, default=18)
    ap.add_argument("--latent", type=int, default=1024, help="latent floats per root gathered / stored per simulation")
    ap.add_argument("--variant", default="mz", choices=["mz", "ez", "gmz", "sez", "smz"],
                    help="sez = Sampled EfficientZero, continuous actions: --actions is K (sampled actions per node), --action-dim is D; "
                         "smz = Stochastic MuZero: --chance-space is the number of chance outcomes per chance node")
    ap.add_argument("--chance-space", type=int, default=18)
    ap.add_argument("--action-dim", type=int, default=6)
    ap.add_argument("--two-player", action="store_true", help="to_play alternates 1/2 (board-game sign flips)")
    ap.add_argument("--streams", type=int, default=1, help="chunks of roots searched on concurrent CUDA streams (graph mode)")
    ap.add_argument("--e2e-streams", type=int, default=4,
                    help="search_streams of the e2e leg's driver: chunk c's host->device copies overlap chunk c-1's simulations")
    ap.add_argument("--no-graph", action="store_true", help="value leg: launch per call instead of one CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--nccl-exchange", action="store_true", help="end-of-search exchange through NCCL instead of the peer-memory kernel")
    ap.add_argument("--cpu-sample-roots", type=int, default=1024)
    ap.add_argument("--categorical", action="store_true",
                    help="the model emits CATEGORICAL value / reward heads ([N, 2 * support + 1] logits, every shipped Atari config) and "
                         "the driver applies InverseScalarTransform every simulation (mcts_ctree.py:302-303) - the fused CUDA kernel")
    ap.add_argument("--support", type=int, default=300, help="support_scale of the categorical heads (300 -> 601 bins)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --roots per GPU whatever N is (the contract's default); strong: --roots-total roots split over the N GPUs")
    ap.add_argument("--roots-total", type=int, default=65536, help="roots of the whole job under --scaling strong")
    a = ap.parse_args()
    if a.scaling == "strong":
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if a.roots_total % world:
            ap.error("--roots-total must be a multiple of the number of GPUs")
        a.roots = a.roots_total // world
    return a


def workload_name(a):
    return "synthetic %s search: %d roots/GPU x %d sims x A=%d, latent %d f32" % (
        {"mz": "MuZero", "ez": "EfficientZero", "gmz": "Gumbel MuZero (m=16)",
         "sez": "Sampled EfficientZero (continuous, D=%d, A = K sampled actions)" % a.action_dim,
         "smz": "Stochastic MuZero (%d chance outcomes per chance node, both model heads on the full batch)" % a.chance_space}[a.variant] + (" 2-player" if a.two_player else ""),
        a.roots, a.sims, a.actions, a.latent)


def workload_config(a, world):
    """`config` of the JSON line: names the WORKLOAD only, so both arms (--impl b200 / reference) print the same dict;
    how an arm ran it (graphs, streams, exchange, host CPU) goes to the line's `run` object."""
    return {"workload": workload_name(a), "roots_total": a.roots * world, "roots_per_gpu": a.roots, "sims": a.sims,
            "actions": a.actions, "latent_f32": a.latent, "model": "table lookup (no NN time)",
            "scalar_transform": ("categorical heads, support %d (%d bins): InverseScalarTransform of value and reward every simulation"
                                 % (a.support, 2 * a.support + 1)) if getattr(a, "categorical", False) else "identity"}


# --------------------------------------------------------------------------------------------- synthetic inputs
def make_inputs(a, rank, device):
    """seeded N(0,1) logits / values, N(0,0.1^2) rewards, Dirichlet(0.3) root noise, to_play=-1."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(1234 + rank)
    N, S, A, L = a.roots, a.sims, a.actions, a.latent
    W = max(A, a.chance_space) if a.variant == "smz" else A  # smz: a row holds policy OR chance logits
    rn = lambda *shape: torch.randn(*shape, generator=g, device=device, dtype=torch.float32)
    gam = torch._standard_gamma(torch.full((N, A), ALPHA, device=device), generator=g)
    inp = dict(
        root_logits=rn(N, A), root_reward=rn(N) * 0.1, root_value=rn(N),
        noise=(gam / gam.sum(1, keepdim=True)).float().contiguous(),
        to_play=(torch.arange(N, device=device, dtype=torch.int32) % 2 + 1) if a.two_player
        else torch.full((N,), -1, dtype=torch.int32, device=device),
        logits=rn(S, N, W), reward=rn(S, N) * 0.1, value=rn(S, N),
        latent_roots=rn(N, max(L, 1)), model_latent=rn(S, N, max(L, 1)),
    )
    if a.variant in ("ez", "sez"):
        inp["reward"] = rn(S, N)  # value prefixes
    if a.variant == "sez":  # policy head = {mu[D], sigma[D] > 0} (sez cnode.cpp:239-247)
        D = a.action_dim
        head = lambda *shape: torch.cat([rn(*shape, D), rn(*shape, D).abs() * 0.5 + 0.1], dim=-1)
        inp["root_logits"], inp["logits"] = head(N), head(S, N)
    if getattr(a, "categorical", False):
        # categorical heads: N(0,1) logits over the support with one +8 bump per row at a bin drawn around the centre, so
        # the expectations spread over a few tens of support units (2 x S x N x (2s+1) floats: ~2 GB at the headline shape)
        Wb = 2 * a.support + 1

        def head_table(spread):
            t = rn(S, N, Wb)
            centre = (rn(S, N) * spread).round().long().clamp(-a.support, a.support) + a.support
            t.scatter_add_(2, centre.unsqueeze(2), torch.full((S, N, 1), 8.0, device=device))
            return t

        inp["value_logits"], inp["reward_logits"] = head_table(20.0), head_table(5.0)
    return inp


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- the synthetic "network"
class NetOut:
    __slots__ = ("latent_state", "policy_logits", "value", "reward", "value_prefix", "reward_hidden_state")


class TableModel:
    """recurrent_inference as a fixed table lookup (NN time excluded and stated as such): call s of a
    search returns row s of pre-generated logits / value / reward tables (a distinct next-latent tensor per simulation), all in HBM."""

    def __init__(self, inp, variant, actions=0, chance_space=0, calls_per_sim=1):
        self.inp, self.variant, self.s = inp, variant, 0
        self.A, self.C, self.calls_per_sim = actions, chance_space, calls_per_sim
        N = inp["model_latent"].shape[1]
        if variant in ("ez", "sez"):
            H = HIDDEN[variant]
            self.hidden = (inp["model_latent"][0, :, :1].expand(N, H).contiguous().unsqueeze(0),
                           inp["model_latent"][0, :, :1].expand(N, H).contiguous().unsqueeze(0))

    def eval(self):
        self.s = 0  # search() calls model.eval() once per search (mcts_ctree.py:242)

    def chunk(self, lo, hi):
        """The same lookup restricted to roots [lo, hi) (a table is not a function of the latent it is handed)."""
        sub = dict(self.inp)
        for k in ("logits", "value", "reward", "model_latent", "value_logits", "reward_logits"):
            if k in self.inp:
                sub[k] = self.inp[k][:, lo:hi]
        m = TableModel.__new__(TableModel)
        m.inp, m.variant, m.s = sub, self.variant, 0
        m.A, m.C, m.calls_per_sim = self.A, self.C, self.calls_per_sim
        if hasattr(self, "hidden"):
            m.hidden = tuple(h[:, lo:hi].contiguous() for h in self.hidden)
        return m

    def recurrent_inference(self, latent, *rest, afterstate=None):
        i, s = self.inp, self.s // self.calls_per_sim
        self.s += 1
        o = NetOut()
        o.latent_state, o.policy_logits, o.value = i["model_latent"][s], i["logits"][s], i["value"][s].unsqueeze(1)
        if afterstate is not None:  # stochastic model: afterstate in -> policy logits [A]; latent in -> chance logits [C]
            o.policy_logits = o.policy_logits[:, :self.A if afterstate else self.C]
        o.reward = o.value_prefix = i["reward"][s].unsqueeze(1)
        if "value_logits" in i:  # categorical heads: [N, 2 * support + 1] logits, transformed by the driver
            o.value, o.reward = i["value_logits"][s], i["reward_logits"][s]
            o.value_prefix = o.reward
        o.reward_hidden_state = getattr(self, "hidden", None)
        return o


def measure_depth(a, mcts, model, inp, rank):
    """Mean search_len over all simulations of one search (an untimed eager search with the depth read per call)."""
    import torch
    from lightzero_b200.forest import Forest
    acc = []
    orig_t, orig_s = Forest.traverse, Forest.step

    def grab(fn):
        def inner(self_, *args, **kw):
            o = fn(self_, *args, **kw)
            acc.append(o["search_len"].float().mean())
            return o
        return inner

    Forest.traverse, Forest.step = grab(orig_t), grab(orig_s)
    try:
        one_search(a, mcts, model, inp, rank=rank)
    finally:
        Forest.traverse, Forest.step = orig_t, orig_s
    return float(torch.stack(acc).mean())


def make_driver(a, device, graph, streams=None, tie_mode="per_root"):
    import lightzero_b200 as lzb
    cls = {"mz": lzb.MuZeroMCTSCtree, "ez": lzb.EfficientZeroMCTSCtree, "gmz": lzb.GumbelMuZeroMCTSCtree,
           "sez": lzb.SampledEfficientZeroMCTSCtree, "smz": lzb.StochasticMuZeroMCTSCtree}[a.variant]
    mcts = cls(dict(num_simulations=a.sims, max_num_considered_actions=16, pb_c_base=PB_C_BASE, pb_c_init=PB_C_INIT, discount_factor=DISCOUNT,
                    value_delta_max=DELTA, device=str(device), cuda_graph=graph, tie_break_seed=TIE_SEED, tie_break_mode=tie_mode,
                    search_streams=(streams or a.streams) if graph else 1, stochastic_dense_inference=True,
                    lstm_horizon_len=5,
                    model=dict(support_scale=a.support, categorical_distribution=bool(a.categorical),
                               continuous_action_space=a.variant == "sez",
                               action_space_size=a.action_dim if a.variant == "sez" else a.actions)))
    if not a.categorical:
        mcts.inverse_scalar_transform_handle = lambda x: x  # the synthetic model emits scalars, not categorical supports
    return mcts


def one_search(a, mcts, model, inp, host=None, rank=0):
    """The call sequence of lzero/policy/muzero.py:680-690: roots -> prepare -> search -> distributions / values.
    host=None: every input already in HBM, outputs stay in HBM.  host=dict of pinned tensors: inputs come from the
    host and the results are read back to the host (the e2e leg)."""
    src = host if host is not None else inp
    if a.variant == "sez":
        roots = mcts.roots(a.roots, None, a.action_dim, a.actions, True)
    elif a.variant == "smz":
        roots = mcts.roots(a.roots, None, a.chance_space)
    else:
        roots = mcts.roots(a.roots, None)
    roots.root_offset = rank * a.roots  # global root index keys the tie-break stream
    if a.variant == "gmz":
        roots.prepare(NOISE_W, src["noise"], src["root_reward"], src["root_value"], src["root_logits"], src["to_play"])
    else:
        roots.prepare(NOISE_W, src["noise"], src["root_reward"], src["root_logits"], src["to_play"])
    if a.variant in ("ez", "sez"):
        mcts.search(roots, model, src["latent_roots"], (src["hidden_c"], src["hidden_h"]), src["to_play"])
    else:
        mcts.search(roots, model, src["latent_roots"], src["to_play"])
    visits, _ = roots.get_distributions_tensor()
    values = roots.get_values_tensor()
    if host is not None:
        host["visits_out"].copy_(visits, non_blocking=True)
        host["values_out"].copy_(values, non_blocking=True)
    return roots, visits, values


# --------------------------------------------------------------------------------------------- per-kernel timing
class KernelTimer:
    """CUDA-event timing of our per-simulation kernels on torch's current stream (the stream the C ABI launches on).

    A search issues S-1 fused step launches back to back (the table-lookup model launches nothing in between), so the
    step kernel is timed as ONE event pair around the whole run of step launches divided by their count: an event
    pair around every single launch costs more than the ~20 us kernel itself.  The caller first queues a spin kernel
    (torch.cuda._sleep) so the host runs ahead and the launches execute back to back; the figure therefore includes
    the launch-to-launch gap, i.e. it is an upper bound of the kernel's own duration.  The first select and the last
    expand+backup are bracketed individually.  Mean search depth feeds the roofline byte model."""
    NAMES = {"step": "step (expand+backup s, select+gather s+1)", "traverse": "select+gather", "backprop": "expand+backup"}

    def __init__(self, steps_per_search):
        self.per_search = steps_per_search
        self.ev = {k: [] for k in self.NAMES}
        self.depth = []
        self._n = 0

    def install(self):
        import torch
        from lightzero_b200.forest import Forest
        self._orig = {k: getattr(Forest, k) for k in self.NAMES}
        timer = self

        def wrap(fn, name):
            def inner(self_, *args, **kw):
                first = last = True
                if name == "step":
                    first, last = timer._n == 0, timer._n == timer.per_search - 1
                    timer._n = (timer._n + 1) % timer.per_search
                if first:
                    timer._e0 = torch.cuda.Event(enable_timing=True)
                    timer._e0.record()
                r = fn(self_, *args, **kw)
                if last:
                    e1 = torch.cuda.Event(enable_timing=True)
                    e1.record()
                    timer.ev[name].append((timer._e0, e1))
                return r
            return inner

        for k, fn in self._orig.items():
            setattr(Forest, k, wrap(fn, k))

    def remove(self):
        from lightzero_b200.forest import Forest
        for k, fn in self._orig.items():
            setattr(Forest, k, fn)

    def summary(self):
        import torch
        torch.cuda.synchronize()
        out = {k: ([a.elapsed_time(b) / (self.per_search if k == "step" else 1) for a, b in v]) for k, v in self.ev.items()}
        return out


def roofline(a, per_kernel_ms, depth, peak_gbs, peak_src):
    """Algorithmic bytes per launch (per root per simulation, x N roots per launch).  The mean depth
    is measured over the whole search; note the leaf rows are read in place from the model's output tensors, so the
    gather moves 8 L bytes per root (4 L read + 4 L written) and there is no separate store into a pool."""
    N, A, Lb = a.roots, a.actions, 8 * max(a.latent, 1) + (16 * HIDDEN[a.variant] if a.variant in HIDDEN else 0)
    if a.variant == "smz":  # decision levels hold A children, chance levels C, alternating: the mean of the two per level
        A = (a.actions + a.chance_space) / 2
    sel = depth * (16 * A + 16) + 8 + 4 * (depth + 1) + Lb
    bak = 4 * A + 16 * A + 28 + (depth + 1) * 20 + 8
    if a.variant == "sez":  # K children per node; the new node's [K, D] action table is read and stored, the leaf's action read and written
        sel += 8 * a.action_dim
        bak += 8 * A * a.action_dim - 4 * A
    alg = {"step": N * (sel + bak), "traverse": N * sel, "backprop": N * bak}
    launches = {"step": a.sims - 1, "traverse": 1, "backprop": 1}  # per search
    avg = {k: sum(v) / len(v) for k, v in per_kernel_ms.items() if v}
    tot = {k: avg[k] * launches[k] for k in avg}
    top = max(tot, key=tot.get)
    avg_ms = avg[top]
    achieved = alg[top] / (avg_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    headline = (a.variant, a.roots, a.sims, a.actions, a.latent, a.two_player) == ("mz", 8192, 50, 18, 1024, False)
    if os.path.exists(tpath) and headline:  # the ncu capture is of the headline workload only
        try:
            traffic = json.load(open(tpath)).get(top)
        except Exception:
            traffic = None
    share = {k: round(v / sum(tot.values()), 4) for k, v in tot.items()}
    return {"bound": "hbm", "kernel": KernelTimer.NAMES[top], "achieved": round(achieved, 2), "peak": peak_gbs, "unit": "GB/s",
            "frac": round(achieved / peak_gbs, 4), "traffic": traffic, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": int(alg[top]), "avg_launch_us": round(avg_ms * 1e3, 2),
            "mean_search_depth": round(depth, 3), "kernel_time_share": share,
            "avg_launch_us_all": {k: round(1e3 * avg[k], 2) for k in avg},
            "timing": "CUDA events on the launch stream; step = one event pair around the S-1 back-to-back step launches / (S-1)"}


# --------------------------------------------------------------------------------------------- CPU side (checker only)
def cpu_inputs(a, n, seed):
    """Host copy of the same synthetic distributions for n roots (numpy; the CPU arm never touches the GPU)."""
    rng = np.random.default_rng(seed)
    S, A, L = a.sims, a.actions, max(a.latent, 1)
    f = lambda *s: rng.standard_normal(s).astype(np.float32)
    if a.variant == "sez":
        D = a.action_dim
        head = lambda *s: np.concatenate([f(*s, D), np.abs(f(*s, D)) * np.float32(0.5) + np.float32(0.1)], axis=-1)
        return dict(root_logits=head(n), root_reward=f(n) * np.float32(0.1), root_value=f(n), noises=None,
                    to_play=np.full(n, -1, np.int32), logits=head(S, n), reward=f(S, n), value=f(S, n),
                    latent_roots=f(n, L), model_latent=f(S, n, L))
    W = max(A, a.chance_space) if a.variant == "smz" else A
    return dict(root_logits=f(n, A), root_reward=f(n) * np.float32(0.1), root_value=f(n),
                noises=[r for r in rng.dirichlet([ALPHA] * A, size=n).astype(np.float32)],
                to_play=(np.arange(n, dtype=np.int32) % 2 + 1) if getattr(a, "two_player", False) else np.full(n, -1, np.int32),
                logits=f(S, n, W),
                reward=(f(S, n) if a.variant == "ez" else f(S, n) * np.float32(0.1)), value=f(S, n),
                latent_roots=f(n, L), model_latent=f(S, n, L))


def cpu_search(a, d, rng_mode=0, want_values=False):
    """One whole search of the reference ctree on ONE host thread, driven the way mcts_ctree.py:241-321 drives it:
    batch_traverse -> Python gather of leaf latents + np.asarray (:283-286) -> (table-lookup model) ->
    batch_backpropagate.  Uses the compiled unmodified reference (oracle/_ref) when present, else the C port.
    rng_mode 0: the reference's own libc rand() ties (timing); 1: the pinned batch stream r(TIE_SEED, sim, 0, 0) that
    the CUDA path reproduces (parity).  Returns (seconds, kind, visit counts[, root values])."""
    from oracle.binding import OracleForest, RefForest, RefSampledForest, ref_available
    n = len(d["to_play"])
    kind = "reference" if ref_available(a.variant) else "port"
    t0 = time.perf_counter()
    if a.variant == "sez":  # K = a.actions sampled actions of dimension D; the reference samples them itself
        D, K = a.action_dim, a.actions
        if kind == "reference":
            f = RefSampledForest(n, D, K, continuous=True)
        else:
            f = OracleForest("sez", n, 2 * D, None, K=K, D=D, continuous=True)
        f.set_delta(DELTA)
        f.set_rng(0 if kind == "reference" else 2, TIE_SEED)
        table = lambda: np.tanh(np.random.default_rng(0).standard_normal((n, K, D))).astype(np.float32)
        f.prepare(NOISE_W, None, d["root_reward"], d["root_logits"], d["to_play"], sampled_actions=table() if kind == "port" else None)
        pool = [d["latent_roots"]]
        for s in range(a.sims):
            o = f.traverse(PB_C_BASE, PB_C_INIT, DISCOUNT, d["to_play"])
            leaf = np.asarray([pool[ix][iy] for ix, iy in zip(o["ix"].tolist(), o["iy"].tolist())])  # noqa: F841
            pool.append(d["model_latent"][s])
            is_reset = (o["search_len"] % 5 == 0).astype(np.int32)
            f.backprop(s + 1, DISCOUNT, d["reward"][s], d["value"][s], d["logits"][s], is_reset, o["vtp"],
                       sampled_actions=table() if kind == "port" else None)
        dist, _ = f.distributions()
        f.values()
        return time.perf_counter() - t0, kind, dist
    legal = [list(range(a.actions))] * n  # the policies always pass explicit legal-action lists (muzero.py:671)
    if a.variant == "smz":  # mcts_ctree_stochastic.py:117-251; the model outputs of both leaf kinds share a [n, W] row table
        from oracle.binding import RefStochasticForest
        f = RefStochasticForest(n, a.actions, a.chance_space, legal) if kind == "reference" else \
            OracleForest("smz", n, a.actions, legal, K=a.chance_space)
        f.set_delta(DELTA)
        f.set_rng(0 if kind == "reference" else 2, TIE_SEED)
        f.prepare(NOISE_W, d["noises"], d["root_reward"], d["root_logits"], d["to_play"])
        pool = [d["latent_roots"]]
        for s in range(a.sims):
            o = f.traverse_chance(PB_C_BASE, PB_C_INIT, DISCOUNT, d["to_play"])
            leaf = np.asarray([pool[ix][iy] for ix, iy in zip(o["ix"].tolist(), o["iy"].tolist())])  # noqa: F841
            chance = np.flatnonzero(o["leaf_is_chance"])  # noqa: F841  (the driver's split into two model batches, :165-176)
            pool.append(d["model_latent"][s])
            f.backprop(s + 1, DISCOUNT, d["reward"][s], d["value"][s], d["logits"][s], None, o["vtp"])
        dist, _ = f.distributions()
        f.values()
        return time.perf_counter() - t0, kind, dist
    f = RefForest(a.variant, n, a.actions, legal) if kind == "reference" else OracleForest(a.variant, n, a.actions, legal)
    f.set_delta(DELTA)
    f.set_rng(rng_mode if rng_mode else (0 if kind == "reference" else 2), TIE_SEED)  # mode 0 = the reference's own libc rand() stream
    f.prepare(NOISE_W, d["noises"], d["root_reward"], d["root_logits"], d["to_play"], values=d["root_value"])
    pool = [d["latent_roots"]]
    for s in range(a.sims):
        o = f.traverse(PB_C_BASE, PB_C_INIT, DISCOUNT, d["to_play"], num_simulations=a.sims, max_considered=16)
        leaf = np.asarray([pool[ix][iy] for ix, iy in zip(o["ix"].tolist(), o["iy"].tolist())])  # noqa: F841
        pool.append(d["model_latent"][s])
        is_reset = (o["search_len"] % 5 == 0).astype(np.int32) if a.variant == "ez" else None
        f.backprop(s + 1, DISCOUNT, d["reward"][s], d["value"][s], d["logits"][s], is_reset, o["vtp"])
    dist, _ = f.distributions()
    vals = f.values()
    dt = time.perf_counter() - t0
    return (dt, kind, dist, vals) if want_values else (dt, kind, dist)


def parity_check(a, inp, model, device):
    """Bit-exact comparison AT THE BENCH SHAPE: the CUDA path (reference-shaped driver, eager launches) against the
    compiled unmodified reference tree on the very inputs the timed GPU arm searches (copied to the host), both under
    the pinned batch tie-break stream r(TIE_SEED, sim, 0, 0) - the keying the unmodified batched reference API can be
    pinned to (include/lzb200.h).  Visit counts must be identical and root values equal as raw fp32 bits."""
    import torch
    if a.variant not in ("mz", "ez", "gmz"):
        return {"roots_compared": 0, "skipped": "the reference draws its own %s inside the tree; tree parity for this flavour is "
                "pinned in tests/ with those draws injected" % ("action tables" if a.variant == "sez" else "chance outcomes")}
    mcts = make_driver(a, device, graph=False, tie_mode="batch")
    _, visits, values = one_search(a, mcts, model, inp, rank=0)
    torch.cuda.synchronize()
    h = lambda t: t.detach().cpu().numpy()
    n, S = a.roots, a.sims
    value, reward = inp["value"], inp["reward"]
    if a.categorical:  # the tree is compared, not the transform (tests/test_scaling.py pins that): the reference gets the
        pairs = [mcts.inverse_scalar_transform_handle.pair(inp["value_logits"][s], inp["reward_logits"][s]) for s in range(S)]
        value = torch.stack([v.reshape(-1) for v, _ in pairs])   # scalars the CUDA search consumed, simulation by simulation
        reward = torch.stack([r.reshape(-1) for _, r in pairs])
    d = dict(root_logits=h(inp["root_logits"]), root_reward=h(inp["root_reward"]), root_value=h(inp["root_value"]),
             noises=list(h(inp["noise"])), to_play=h(inp["to_play"]), logits=h(inp["logits"]), reward=h(reward),
             value=h(value), latent_roots=np.zeros((n, 1), np.float32), model_latent=np.zeros((S, n, 1), np.float32))
    dt, kind, dist, vals = cpu_search(a, d, rng_mode=1, want_values=True)
    gv, gx = h(visits), h(values)
    dist = np.where(dist < 0, 0, dist)
    return {"roots_compared": int(n), "sims_per_root": int(S), "visit_mismatches": int((gv[:, :dist.shape[1]] != dist).sum()),
            "value_bit_mismatches": int((gx.view(np.uint32) != np.ascontiguousarray(vals, np.float32).view(np.uint32)).sum()),
            "checker": kind, "checker_seconds": round(dt, 2),
            "stream": "pinned batch keying r(seed=%d, sim, 0, 0) on both sides" % TIE_SEED}


def _cpu_worker(a, n, seed, conn):
    d = cpu_inputs(a, n, seed)
    conn.send("ready")
    while True:
        msg = conn.recv()
        if msg == "stop":
            return
        dt, kind, _ = cpu_search(a, d)
        conn.send((dt, kind))


def run_reference_arm(a):
    """--impl reference: the reference ctree on every host core (roots are independent, so the batch is sharded over
    P processes, one single-threaded reference tree each); same metric / config / unit as the B200 arm."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    P = max(1, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    # bounded sample: at most the full per-GPU batch, at least 64 roots per process
    total = a.roots
    per = max(1, total // P)
    P = min(P, total // per)
    # the compiled reference is loaded HERE, before the fork: the workers inherit the mapping and the driver's list of
    # native libraries of this process shows oracle/_ref/libref_<variant>.so
    from oracle.binding import ref_available
    ref_available(a.variant)
    ctx = mp.get_context("fork")
    procs = []
    for r in range(P):
        pc, cc = ctx.Pipe()
        p = ctx.Process(target=_cpu_worker, args=(a, per, 1234 + r, cc), daemon=True)
        p.start()
        procs.append((p, pc))
    for _, pc in procs:
        pc.recv()
    kind = "port"

    def step():
        nonlocal kind
        t0 = time.perf_counter()
        for _, pc in procs:
            pc.send("go")
        for _, pc in procs:
            _, kind = pc.recv()
        return time.perf_counter() - t0

    for _ in range(min(a.warmup, 1)):
        step()
    times = [step() for _ in range(a.steps)]
    for p, pc in procs:
        pc.send("stop")
        p.join(timeout=5)
    sims = per * P * a.sims
    ms = 1e3 * sum(times) / len(times)
    v = sims / (ms * 1e-3)
    cpu = ""
    try:
        cpu = [l.split(":", 1)[1].strip() for l in open("/proc/cpuinfo") if l.startswith("model name")][0]
    except Exception:
        pass
    sample = "%d roots (%d processes x %d) x %d sims per step, tree + Python leaf gather, no NN, libc rand() ties" % (per * P, P, per, a.sims)
    print(json.dumps({
        "impl": "reference", "metric": "mcts_simulations_per_sec", "value": round(v, 1), "unit": "simulations/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
        "scaling": a.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(a, a.gpus),
        "run": {"host_cpu": cpu, "processes": P, "marshalling": "flat arrays through oracle/ref_harness.cpp: Cython's list -> "
                "std::vector copies (~9 % of the reference's time, SURVEY 6) are NOT included, which flatters the reference"},
        "cpu_baseline": {"value": round(v, 1), "unit": "simulations/s", "cores": P, "kind": kind, "sample": sample},
        "e2e": {"value": round(v, 1), "unit": "simulations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# --------------------------------------------------------------------------------------------- B200 arm
def run_b200_arm(a):
    import torch
    import torch.distributed as dist
    from lightzero_b200 import sharding

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    N, S, A = a.roots, a.sims, a.actions
    inp = make_inputs(a, rank, device)
    if a.variant in HIDDEN:
        inp["hidden_c"] = torch.zeros(1, N, HIDDEN[a.variant], device=device)
        inp["hidden_h"] = torch.zeros(1, N, HIDDEN[a.variant], device=device)
    model = TableModel(inp, a.variant, a.actions, a.chance_space, calls_per_sim=2 if a.variant == "smz" else 1)
    W = max(A, a.chance_space) if a.variant == "smz" else A  # columns of the visit-count rows
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=device)
    counts = [N] * world

    # the one exchange of a search; identity at world == 1.  Default: one NVLink peer-memory kernel per
    # rank (lzb_peer_allgather); --nccl-exchange: one packed NCCL all-gather
    peer = None
    if world > 1 and not a.nccl_exchange:
        try:
            peer = sharding.PeerExchange(N, W + 1)  # raises on every rank together if any rank cannot map its peers
        except RuntimeError as exc:  # e.g. CUDA IPC not permitted in this container: the NCCL route, stated in config
            print("rank %d: %s; using NCCL for the exchange" % (rank, exc), file=sys.stderr)
            a.nccl_exchange = True

    def exchange(visits, values):
        return sharding.gather_search_outputs(visits, values, counts, peer=peer, check=False) if world > 1 else (visits, values)  # error word read once, after the timed loop

    # ---- value leg: inputs resident in HBM, the whole search replayed as one CUDA graph
    mcts = make_driver(a, device, graph=not a.no_graph)

    mids = []

    local_out = {}

    def dev_step(timed=False):
        _, visits, values = one_search(a, mcts, model, inp, rank=rank)
        local_out["visits"], local_out["values"] = visits, values
        if timed and world > 1:  # where the search ends and the exchange begins (diagnostic: config.exchange_ms_per_step)
            mids.append(torch.cuda.Event(enable_timing=True))
            mids[-1].record()
        return exchange(visits, values)

    for _ in range(max(a.warmup, 3)):
        dev_step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    align = torch.zeros(1, device=device)
    t_wall = time.perf_counter()
    for e0, e1 in ev:
        flush.fill_(1)  # L2 flush between timed steps (512 MB write > 126 MB L2), outside the event pair
        if world > 1:  # ranks leave the (untimed) flush together (stream-ordered, the host keeps queueing), so the
            dist.all_reduce(align)  # exchange inside the timed region does not absorb the flush skew between ranks
        e0.record()
        visits, values = dev_step(timed=True)
        e1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall
    dev_ms = sum(x.elapsed_time(y) for x, y in ev)
    exchange_ms = sum(m.elapsed_time(y) for m, (_, y) in zip(mids, ev)) / a.steps if mids else 0.0
    # sanity: a search spends exactly S visits below each root - every row of the gathered batch, not just this shard
    assert bool((visits.sum(1) == S).all()), "visit counts do not sum to the number of simulations"
    # correctness of the exchange on real NVLink: the rows every rank RECEIVED (peer-memory kernel or packed NCCL
    # all-gather) against a plain per-tensor NCCL all-gather of each rank's own shard, on every rank
    exchange_check = None
    if world > 1:
        ref_v = torch.empty((N * world, W), dtype=torch.int32, device=device)
        ref_x = torch.empty(N * world, dtype=torch.float32, device=device)
        dist.all_gather_into_tensor(ref_v, local_out["visits"].contiguous())
        dist.all_gather_into_tensor(ref_x, local_out["values"].contiguous())
        bad = torch.stack([(visits != ref_v).any(1).sum(), (values.view(torch.int32) != ref_x.view(torch.int32)).sum()]).to(torch.int64)
        dist.all_reduce(bad)  # summed over ranks: every rank checked every row it received
        own = slice(rank * N, (rank + 1) * N)
        assert torch.equal(visits[own], local_out["visits"]) and torch.equal(values[own], local_out["values"])
        exchange_check = {"rows_checked_per_rank": N * world, "ranks": world, "visit_row_mismatches": int(bad[0]),
                          "value_bit_mismatches": int(bad[1]),
                          "against": "torch.distributed all_gather_into_tensor (NCCL) of every rank's own shard"}

    # ---- e2e leg: reference-shaped API, HOST inputs (pinned), outputs read back to the host, wall clock
    e2e_ms = None
    h2d = d2h = 0
    if not a.no_e2e:
        pin = lambda t: t.detach().cpu().pin_memory()
        host = {k: pin(inp[k]) for k in ("root_logits", "noise", "root_reward", "root_value", "to_play", "latent_roots")}
        if a.variant in HIDDEN:
            host["hidden_c"], host["hidden_h"] = pin(inp["hidden_c"]), pin(inp["hidden_h"])
        host["visits_out"] = torch.empty((N, W), dtype=torch.int32).pin_memory()
        host["values_out"] = torch.empty(N, dtype=torch.float32).pin_memory()
        if a.variant != "gmz":
            host.pop("root_value")
        h2d = sum(host[k].numel() * host[k].element_size() for k in host if not k.endswith("_out")) + 4 * N  # to_play twice
        d2h = host["visits_out"].numel() * 4 + host["values_out"].numel() * 4

        # the e2e leg's own driver: the batch is searched in chunks on concurrent streams so that the host->device copy of
        # chunk c overlaps the simulations of chunk c-1 (same results; `value` above is the single-stream figure)
        e2e_streams = max(a.streams, a.e2e_streams) if (not a.no_graph and a.variant in ("mz", "ez", "gmz")) else a.streams
        if a.no_graph:
            e2e_streams = 1
        mcts_e2e = mcts if e2e_streams == a.streams else make_driver(a, device, graph=not a.no_graph, streams=e2e_streams)

        def e2e_step():
            _, visits, values = one_search(a, mcts_e2e, model, inp, host=host, rank=rank)
            exchange(visits, values)
            torch.cuda.synchronize()

        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.steps):
            e2e_step()
        barrier()
        e2e_ms = 1e3 * (time.perf_counter() - t0)
        assert int(host["visits_out"].sum()) == N * S
        # context for the e2e figure: what this box's host->device link delivers for the largest input of a step
        big = max((host[k] for k in host if not k.endswith("_out")), key=lambda t: t.numel() * t.element_size())
        dst = torch.empty(big.shape, dtype=big.dtype, device=device)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dst.copy_(big, non_blocking=True)
        ev0.record()
        for _ in range(3):
            dst.copy_(big, non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        h2d_gbs = 3 * big.numel() * big.element_size() / (ev0.elapsed_time(ev1) * 1e-3) / 1e9

    # ---- e2e_lists leg: the call pattern of an UNMODIFIED lzero policy (lzero/policy/muzero.py:671-719): per-root legal
    # action lists, noises / logits / rewards / to_play as Python lists (the policy calls .tolist() before prepare),
    # root latents as a pageable numpy array, and Python lists back from get_distributions() / get_values().
    lists_ms = None
    if not a.no_e2e and world == 1 and a.variant in ("mz", "ez", "gmz"):
        cpu = lambda t: t.detach().cpu()
        L = dict(noise=cpu(inp["noise"]).tolist(), reward=cpu(inp["root_reward"]).tolist(), logits=cpu(inp["root_logits"]).tolist(),
                 value=cpu(inp["root_value"]).tolist(), to_play=cpu(inp["to_play"]).tolist(),
                 legal=[list(range(A)) for _ in range(N)], latent=cpu(inp["latent_roots"]).numpy())
        if a.variant == "ez":
            L["hidden"] = (cpu(inp["hidden_c"]).numpy(), cpu(inp["hidden_h"]).numpy())

        def lists_step():
            roots = mcts_e2e.roots(N, L["legal"])
            if a.variant == "gmz":
                roots.prepare(NOISE_W, L["noise"], L["reward"], L["value"], L["logits"], L["to_play"])
            else:
                roots.prepare(NOISE_W, L["noise"], L["reward"], L["logits"], L["to_play"])
            if a.variant == "ez":
                mcts_e2e.search(roots, model, L["latent"], L["hidden"], L["to_play"])
            else:
                mcts_e2e.search(roots, model, L["latent"], L["to_play"])
            return roots.get_distributions(), roots.get_values()

        for _ in range(3):
            d_, v_ = lists_step()
        assert len(d_) == N and len(v_) == N and all(sum(r) == S for r in d_[:64])
        t0 = time.perf_counter()
        for _ in range(a.steps):
            lists_step()
        torch.cuda.synchronize()
        lists_ms = 1e3 * (time.perf_counter() - t0) / a.steps
    clk = clocks.stop() if rank == 0 else None

    # ---- per-kernel pass (eager launches bracketed by events) for the roofline of the dominant kernel
    timer = KernelTimer(S - 1)
    eager = make_driver(a, device, graph=False)
    one_search(a, eager, model, inp, rank=rank)  # warm the eager path
    depth = measure_depth(a, eager, model, inp, rank)
    timer.install()
    for _ in range(3):
        flush.fill_(1)
        torch.cuda._sleep(int(3e7))  # ~15 ms of spin: the host queues the whole search behind it
        one_search(a, eager, model, inp, rank=rank)
        torch.cuda.synchronize()
    timer.remove()
    per_kernel = timer.summary()
    transform = None
    if a.categorical:
        # the scalar-transform kernel alone: S launches back to back over the S distinct (HBM-cold) table rows, one event
        # pair around them.  In the eager pass above a transform launch sits between every two step launches, so that
        # pass's "step" figure is step + transform; the transform's own time is taken off it here.
        handle = eager.inverse_scalar_transform_handle
        for rep in range(3):
            flush.fill_(1)
            torch.cuda._sleep(int(3e7))
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for s_ in range(S):
                handle.pair(inp["value_logits"][s_], inp["reward_logits"][s_])
            t1.record()
            torch.cuda.synchronize()
            us = 1e3 * t0.elapsed_time(t1) / S
            transform = us if transform is None else min(transform, us)
        per_kernel["step"] = [max(x - transform * 1e-3, 1e-6) for x in per_kernel["step"]]
        tbytes = 2 * N * (2 * a.support + 1) * 4 + 2 * N * 4
        transform = {"kernel": "k_inverse_scalar_transform (value + reward heads, one launch per simulation)",
                     "avg_launch_us": round(transform, 2), "algorithmic_bytes_per_launch": tbytes,
                     "achieved": round(tbytes / (transform * 1e-6) / 1e9, 1), "unit": "GB/s",
                     "note": "the step figure of this line is the eager per-simulation time minus this kernel's"}

    # ---- reduce over ranks (max time), print on rank 0
    t = torch.tensor([dev_ms, e2e_ms or 0.0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_tot = float(t[0]), float(t[1])
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        ms_step = dev_ms / a.steps
        total_sims = N * world * S
        line = {
            "metric": "mcts_simulations_per_sec", "value": round(total_sims / (ms_step * 1e-3), 1), "unit": "simulations/s",
            "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": round(ms_step, 4),
            "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(a, world),
            "run": {"cuda_graph": not a.no_graph, "search_streams": 1 if a.no_graph else a.streams,
                    "l2": "512 MB flush between timed steps; working set > L2",
                    "parallelism": "roots sharded x%d, all-gather of root outputs per search (%s)" % (
                        world, "single GPU" if world == 1 else ("NCCL" if a.nccl_exchange else "NVLink peer-memory kernel")),
                    "wall_ms_per_step_incl_flush": round(1e3 * t_wall / a.steps, 3),
                    "exchange_ms_per_step": round(exchange_ms, 4)},
            # init, minmax, prepare, root slots, select, S-1 steps, backup, outputs (+ the exchange kernel when sharded)
            "gpu_launches": a.steps * ((5 + S + 1) * (1 if a.no_graph else a.streams) + (1 if peer is not None else 0)
                                       + (S + 1 if a.variant == "sez" else 0)   # + one action-sampling kernel per expansion
                                       + (S * (2 if a.variant == "smz" else 1) * (1 if a.no_graph else a.streams) if a.categorical else 0)),  # + the scalar-transform kernel
            "clocks": clk,
            "roofline": roofline(a, per_kernel, depth, peak, src),
        }
        if exchange_check is not None:
            line["exchange_check"] = exchange_check
        if transform is not None:
            transform["frac"] = round(transform["achieved"] / peak, 4)
            try:  # DRAM bytes per launch of this kernel from its ncu --set full capture (profiles/r02_scalar_ncu_raw.csv)
                transform["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("scalar_transform") if a.support == 300 and N == 8192 else None
            except Exception:
                transform["traffic"] = None
            line["roofline"]["scalar_transform"] = transform
        if e2e_ms is not None:
            line["e2e"] = {"value": round(total_sims / (e2e_tot / a.steps * 1e-3), 1), "unit": "simulations/s",
                           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                           "ms_per_step": round(e2e_tot / a.steps, 4), "clock": "host wall clock incl. Python, max over ranks",
                           "h2d_link_gbs": round(h2d_gbs, 1), "search_streams": e2e_streams}
        if lists_ms is not None:
            line["e2e_lists"] = {"value": round(total_sims / (lists_ms * 1e-3), 1), "unit": "simulations/s", "ms_per_step": round(lists_ms, 4),
                                 "what": "unmodified-policy call pattern: Python lists of lists in (legal actions, noises, logits, rewards, to_play), "
                                         "pageable numpy latents, Python lists out (get_distributions / get_values); host wall clock"}
        if world == 1 and not a.no_cpu_baseline:
            n = min(a.cpu_sample_roots, N)
            d = cpu_inputs(a, n, 1234)
            cpu_search(a, d)  # warm-up
            best, kind = None, "port"
            for _ in range(3):
                dt, kind, _ = cpu_search(a, d)
                best = dt if best is None else min(best, dt)
            line["cpu_baseline"] = {"value": round(n * S / best, 1), "unit": "simulations/s", "cores": 1, "kind": kind,
                                    "sample": "%d of %d roots x %d sims, best of 3, tree + Python leaf gather, no NN" % (n, N, S)}
            line["parity"] = parity_check(a, inp, model, device)
        print(json.dumps(line))
        bad = line.get("parity", {})
        if bad.get("visit_mismatches") or bad.get("value_bit_mismatches"):
            raise SystemExit("parity: the CUDA search differs from the reference at the bench shape: %r" % bad)
        if exchange_check and (exchange_check["visit_row_mismatches"] or exchange_check["value_bit_mismatches"]):
            raise SystemExit("exchange: gathered rows differ from the NCCL all-gather of the shards: %r" % exchange_check)
    if peer is not None:
        if peer.error():
            raise SystemExit("peer exchange: rank %d did not arrive" % (peer.error() - 1))
        dist.barrier()
        peer.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)
