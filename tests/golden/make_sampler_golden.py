"""Generates tests/golden/sez_sampler_*.npz: the action tables the compiled, UNMODIFIED Sampled-EfficientZero reference
tree samples inside CNode::expand (ctree_sampled_efficientzero/lib/cnode.cpp:251-283, :336-406) when the wall-clock seed
of its std::default_random_engine is pinned (oracle/ref_harness_sampled.cpp: system_clock::now()), together with what
the search then did with them (per-simulation leaf indices / last actions, final visit counts and root values).
Authoring container only (needs /root/reference):   python tests/golden/make_sampler_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from tests.common import Scenario, run_sez_reference  # noqa: E402
from tests.test_sampler_oracle import CLOCK, SAMPLER_CASES  # noqa: E402


def main():
    for name, (kw, continuous) in sorted(SAMPLER_CASES.items()):
        sc = Scenario(**kw)
        ref, rec = run_sez_reference(sc, 1, 2024, continuous=continuous, clock_base=CLOCK)
        dist, ln = ref.distributions()
        mn, mx = ref.minmax()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), clock_base=np.uint64(CLOCK), root_actions=sc.root_actions,
                            actions=sc.actions, dist=dist, dist_len=ln, values=ref.values(), mm_min=mn, mm_max=mx,
                            **{k: np.stack([r[k] for r in rec]) for k in ("ix", "iy", "search_len", "vtp", "last_action_f", "is_reset")})
        print(name, sc.root_actions.shape, sc.actions.shape)


if __name__ == "__main__":
    main()
