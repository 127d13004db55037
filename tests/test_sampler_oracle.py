"""The Sampled-EfficientZero SAMPLER pinned: oracle/libstdcxx_rng.c + orc_tanhf against
(1) libstdc++'s own default_random_engine / normal_distribution<float> / uniform_real_distribution<double> (through
    oracle/_ref/libref_sez.so, which links the real library), bit for bit;
(2) the live libm's tanhf over a sample of the whole float range, bit for bit;
(3) the action tables the compiled, unmodified reference tree samples inside CNode::expand when its wall-clock seed is
    pinned (ctree_sampled_efficientzero/lib/cnode.cpp:251-283 continuous, :336-406 discrete), bit for bit;
(4) the committed golden tables (tests/golden/sez_sampler_*.npz, made by (3) in the authoring container).
(1)-(3) need /root/reference (skipped elsewhere); (2) and (4) run everywhere."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle.binding import oracle_lib, ref_available, ref_lib, sample_libstdcxx
from tests.common import Scenario, assert_same, run_sez_reference, sez_logits

HERE = os.path.dirname(os.path.abspath(__file__))
CLOCK = 1_762_300_000_123_456_789  # some nanosecond count; its low 32 bits are what the reference keeps
SAMPLER_CASES = {
    "sez_sampler_continuous_k20_d6": (dict(variant="sez", n=24, A=12, S=12, seed=61, K=20, D=6), True),
    "sez_sampler_discrete_k4_of_6": (dict(variant="sez", n=24, A=6, S=12, seed=62, K=4, D=1), False),
    "sez_sampler_discrete_k5_of_18": (dict(variant="sez", n=16, A=18, S=10, seed=63, K=5, D=1), False),
}
needs_ref = pytest.mark.skipif(not ref_available("sez"), reason="oracle/_ref/libref_sez.so not built (no /root/reference here)")


def expected_tables(sc, continuous, clock):
    """What oracle/libstdcxx_rng.c says the reference samples: expansion e of root i is seeded with clock + e * n + i."""
    root = sample_libstdcxx(sez_logits(sc.root_logits, continuous), sc.K, sc.D, sc.A, continuous, clock)
    sims = [sample_libstdcxx(sez_logits(sc.logits[s], continuous), sc.K, sc.D, sc.A, continuous, clock + (s + 1) * sc.n)
            for s in range(sc.S)]
    return root, np.stack(sims)


@needs_ref
@pytest.mark.parametrize("seed", [0, 1, 2147483647, 2147483646, 123456789, 4294967295])
def test_engine_and_distributions_match_libstdcxx(seed):
    R, O = ref_lib("sez"), oracle_lib()
    R.ref_libstdcxx_normal.argtypes = [C.c_uint, C.c_float, C.c_float, C.c_int, C.c_void_p]
    R.ref_libstdcxx_uniform.argtypes = [C.c_uint, C.c_int, C.c_void_p]
    n = 20000
    for mu, sigma in ((0.0, 1.0), (-0.7, 0.35), (3.0, 0.1)):
        a, b = np.zeros(n, np.float32), np.zeros(n, np.float32)
        R.ref_libstdcxx_normal(seed, mu, sigma, n, a.ctypes.data)
        O.orc_libstdcxx_normal(seed, mu, sigma, n, b.ctypes.data)
        assert_same(b, a, "normal_distribution<float>(%g, %g), seed %d" % (mu, sigma, seed))
    a, b = np.zeros(n, np.float64), np.zeros(n, np.float64)
    R.ref_libstdcxx_uniform(seed, n, a.ctypes.data)
    O.orc_libstdcxx_uniform(seed, n, b.ctypes.data)
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), "uniform_real_distribution<double>"
    assert 0.0 <= a.min() and a.max() < 1.0


def test_tanhf_restatement_matches_live_libm():
    O = oracle_lib()
    assert O.orc_tanhf_mismatch_range(0, 0xffffffff, 97) == 0          # every 97th bit pattern of the whole range
    assert O.orc_tanhf_mismatch_range(0x3c000000, 0x42000000, 5) == 0   # dense over 2^-7 .. 32, where tanh varies
    assert O.orc_tanhf_mismatch_range(0xbc000000, 0xc2000000, 5) == 0


@needs_ref
@pytest.mark.parametrize("name", sorted(SAMPLER_CASES))
def test_oracle_sampler_matches_the_reference_tree_under_a_pinned_clock(name):
    kw, continuous = SAMPLER_CASES[name]
    sc = Scenario(**kw)
    run_sez_reference(sc, 1, 2024, continuous=continuous, clock_base=CLOCK)  # fills sc.root_actions / sc.actions
    root, sims = expected_tables(sc, continuous, CLOCK)
    assert_same(root, sc.root_actions, name + ": root action tables")
    assert_same(sims, sc.actions, name + ": per-simulation action tables")


@pytest.mark.parametrize("name", sorted(SAMPLER_CASES))
def test_oracle_sampler_matches_the_golden_tables(name):
    kw, continuous = SAMPLER_CASES[name]
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    sc = Scenario(**kw)
    root, sims = expected_tables(sc, continuous, int(g["clock_base"]))
    assert_same(root, g["root_actions"], name + ": root action tables")
    assert_same(sims, g["actions"], name + ": per-simulation action tables")
