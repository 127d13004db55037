"""Generates tests/golden/*.npz from the COMPILED, UNMODIFIED reference ctree (oracle/_ref, built from
/root/reference by oracle/build.py) under the pinned tie-break stream of oracle/rng_pin.c.

Run in the build container (where /root/reference exists):  python tests/golden/make_golden.py
The fixtures are what tests/test_oracle_vs_reference.py::test_oracle_matches_golden_fixture checks the
oracle against on machines without the reference (e.g. the GPU box).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.binding import RefForest, RefStochasticForest  # noqa: E402
from tests.common import Scenario, run_cpu, run_sez_reference  # noqa: E402
from tests.test_oracle_vs_reference import CASES, search_summary, sez_continuous  # noqa: E402

if __name__ == "__main__":
    out_dir = os.path.dirname(os.path.abspath(__file__))
    only = sys.argv[1:]
    for name, (kw, mode) in sorted(CASES.items()):
        if only and not any(name.startswith(o) for o in only):
            continue
        sc = Scenario(**kw)
        rec = []
        if sc.variant == "sez":  # the fixture carries the action tables the reference sampled in this run
            ref, rec = run_sez_reference(sc, mode, 2024, continuous=sez_continuous(name))
        elif sc.variant == "smz":  # the fixture's "paths" carry the chance outcomes the reference drew in this run
            ref = run_cpu(RefStochasticForest(sc.n, sc.A, sc.K, sc.legal, per_root=(mode == 2)), sc, mode, 2024, record=rec)
        else:
            ref = run_cpu(RefForest(sc.variant, sc.n, sc.A, sc.legal, per_root=(mode == 2)), sc, mode, 2024, record=rec)
        summ = search_summary(ref, sc, rec)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **summ)
        print(name, "mean depth %.2f" % summ["search_len"].mean(), "bytes", os.path.getsize(os.path.join(out_dir, name + ".npz")))
