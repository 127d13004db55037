"""Golden vectors for the single-node Gumbel MuZero helpers (gmz_tree.pyx:97-149), produced by CALLING the compiled,
unmodified reference (oracle/_ref/libref_gmz.so: ref_softmax / ref_mixed_value / ref_rescale / ref_score_considered /
ref_node_helpers in oracle/ref_harness.cpp) on seeded inputs.  Needs /root/reference; writes tests/golden/gumbel_helpers.npz.

    python tests/golden/make_gumbel_helper_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import binding as B  # noqa: E402

KEYS_IN = ("logits", "child_visit", "child_prior", "x", "q", "gumbel", "nq")
SCAL_IN = ("to_play", "value_prefix", "value", "discount", "num_simulations", "max_considered", "maxvisit_init", "value_scale",
           "rescale_values", "epsilon", "considered")
KEYS_OUT = ("softmax", "mixed_value", "rescale", "score_considered", "priors", "qtransform", "root_action", "interior_action")


def cases(seed=11, count=48):
    rng = np.random.default_rng(seed)
    out = []
    for t in range(count):
        A = int(rng.choice([2, 3, 6, 9, 18, 33, 64]))
        visits = rng.integers(0, 4, size=A)
        if t % 7 == 0:
            visits[:] = 0  # nothing visited: probs_sum = 0 -> 0 / 0 inside compute_mixed_value is skipped, value = raw / 1
        if t % 11 == 0:
            visits[:] = 2  # every child visited
        out.append(dict(logits=(rng.normal(size=A) * rng.choice([0.1, 1.0, 5.0])).astype(np.float32), to_play=int(rng.choice([-1, 1, 2])),
                        value_prefix=float(np.float32(rng.normal() * 0.1)), value=float(np.float32(rng.normal())), discount=0.997,
                        num_simulations=int(rng.integers(8, 60)), max_considered=int(rng.integers(1, 17)),
                        child_visit=visits.astype(np.int32), child_prior=rng.random(A).astype(np.float32),
                        maxvisit_init=float(rng.choice([50.0, 10.0])), value_scale=float(rng.choice([0.1, 1.0])),
                        rescale_values=int(rng.integers(0, 2)), epsilon=1e-8, x=(rng.normal(size=A) * 3).astype(np.float32),
                        q=rng.normal(size=A).astype(np.float32), gumbel=(rng.normal(size=A) * 10).astype(np.float32),
                        nq=rng.random(A).astype(np.float32), considered=int(rng.integers(0, 3))))
    return out


def main():
    if not B.ref_available("gmz"):
        raise SystemExit("needs the compiled reference (oracle/_ref/libref_gmz.so, built from /root/reference)")
    pack = {}
    cs = cases()
    for t, c in enumerate(cs):
        r = B.gumbel_helpers_reference(c)
        for k in KEYS_IN:
            pack["c%02d_in_%s" % (t, k)] = np.asarray(c[k])
        pack["c%02d_in_scalars" % t] = np.asarray([float(c[k]) for k in SCAL_IN], np.float64)
        for k in KEYS_OUT:
            pack["c%02d_out_%s" % (t, k)] = np.asarray(r[k])
    pack["count"] = np.asarray(len(cs))
    np.savez_compressed(os.path.join(HERE, "gumbel_helpers.npz"), **pack)
    print("wrote %d cases" % len(cs))


def load(path=os.path.join(HERE, "gumbel_helpers.npz")):
    """-> list of (case dict, reference outputs dict)"""
    z = np.load(path)
    out = []
    for t in range(int(z["count"])):
        c = {k: z["c%02d_in_%s" % (t, k)] for k in KEYS_IN}
        sc = z["c%02d_in_scalars" % t]
        for k, v in zip(SCAL_IN, sc):
            c[k] = int(v) if k in ("to_play", "num_simulations", "max_considered", "rescale_values", "considered") else float(v)
        out.append((c, {k: z["c%02d_out_%s" % (t, k)] for k in KEYS_OUT}))
    return out


if __name__ == "__main__":
    main()
