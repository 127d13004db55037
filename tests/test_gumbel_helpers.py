"""The single-node Gumbel MuZero helpers of gmz_tree.pyx:97-149 (softmax, pcompute_mixed_value, prescale_qvalues,
pqtransform_completed_by_mix_value, pscore_considered, select_root_child, select_interior_child) and the `Node` / `Action`
test helpers of the other tree modules.

CPU: the oracle's restatement (oracle/mcts_oracle.c: orc_mixed_value ...) against golden vectors produced by calling the
compiled reference (tests/golden/make_gumbel_helper_golden.py) and, where /root/reference is present, against the compiled
reference itself on fresh random cases.  GPU: the module functions (one-thread device evaluations,
csrc/lzb_gumbel_helpers.cuh, through the C ABI's lzb_gumbel_helper) against the same goldens, bit for bit."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_gumbel_helper_golden as G  # noqa: E402
from oracle import binding as B  # noqa: E402

FLOAT_KEYS = ("softmax", "mixed_value", "rescale", "score_considered", "priors", "qtransform")


def _same_bits(a, b):
    a, b = np.atleast_1d(np.asarray(a, np.float32)), np.atleast_1d(np.asarray(b, np.float32))
    return a.shape == b.shape and bool((a.view(np.uint32) == b.view(np.uint32)).all())


def test_oracle_helpers_match_the_reference_golden():
    cases = G.load()
    assert len(cases) == 48
    for t, (c, want) in enumerate(cases):
        got = B.gumbel_helpers_oracle(c)
        for k in FLOAT_KEYS:
            assert _same_bits(got[k], want[k]), (t, k, got[k], want[k])
        assert got["root_action"] == int(want["root_action"]) and got["interior_action"] == int(want["interior_action"]), t


def test_oracle_helpers_match_the_compiled_reference():
    if not B.ref_available("gmz"):
        pytest.skip("oracle/_ref not built and /root/reference absent")
    for t, c in enumerate(G.cases(seed=77, count=120)):
        got, want = B.gumbel_helpers_oracle(c), B.gumbel_helpers_reference(c)
        for k in FLOAT_KEYS:
            assert _same_bits(got[k], want[k]), (t, k)
        assert got["root_action"] == want["root_action"] and got["interior_action"] == want["interior_action"], t


def test_golden_is_what_the_maker_generates():
    if not B.ref_available("gmz"):
        pytest.skip("oracle/_ref not built and /root/reference absent")
    for (c, want), c2 in zip(G.load(), G.cases()):
        assert all(np.array_equal(np.asarray(c[k]), np.asarray(c2[k])) for k in G.KEYS_IN)
        again = B.gumbel_helpers_reference(c2)
        assert all(_same_bits(again[k], want[k]) for k in FLOAT_KEYS)


@pytest.mark.gpu
def test_cuda_helpers_match_the_reference_golden():
    from lightzero_b200.ctree.ctree_gumbel_muzero import gmz_tree as T
    for t, (c, want) in enumerate(G.load()):
        x = [float(v) for v in c["x"]]
        assert T.softmax(x) is x and _same_bits(x, want["softmax"]), t  # in place, like the .pyx
        mv = T.pcompute_mixed_value(c["value"], list(c["q"]), list(c["child_visit"]), list(c["child_prior"]))
        assert _same_bits(mv, want["mixed_value"]), (t, mv, want["mixed_value"])
        x = [float(v) for v in c["x"]]
        assert _same_bits(T.prescale_qvalues(x, c["epsilon"]), want["rescale"]), t
        sc = T.pscore_considered(c["considered"], list(c["gumbel"]), list(c["logits"]), list(c["nq"]), list(c["child_visit"]))
        assert _same_bits(sc, want["score_considered"]), t
        node = T.Node(0.0, list(range(c["logits"].size)))
        node.expand(c["to_play"], 0, 0, c["value_prefix"], c["value"], [float(v) for v in c["logits"]])
        assert _same_bits(node.children()["prior"], want["priors"]), t
        qt = T.pqtransform_completed_by_mix_value(node, list(c["child_visit"]), list(c["child_prior"]), c["discount"], c["maxvisit_init"],
                                                  c["value_scale"], bool(c["rescale_values"]), c["epsilon"])
        assert _same_bits(qt, want["qtransform"]), (t, qt, want["qtransform"])
        assert T.select_root_child(node, c["discount"], c["num_simulations"], c["max_considered"]) == int(want["root_action"]), t
        assert T.select_interior_child(node, c["discount"]) == int(want["interior_action"]), t


@pytest.mark.gpu
def test_node_and_action_helpers_of_the_other_trees():
    """mz / ez `Node(prior, legal_actions).expand(...)` (mz_tree.pyx:60-72): the expanded node's priors are the softmax the
    batched root expansion computes; ezs `Action` / `Node` (ezs_tree.pyx:26-36,86-101): K sampled actions of dimension D."""
    from lightzero_b200.ctree.ctree_efficientzero import ez_tree
    from lightzero_b200.ctree.ctree_muzero import mz_tree
    from lightzero_b200.ctree.ctree_sampled_efficientzero import ezs_tree
    logits = np.asarray([0.3, -1.2, 2.0, 0.0, 0.7], np.float32)
    want = B.OracleForest("mz", 1, 5)
    want.prepare(0.0, None, np.zeros(1, np.float32), logits[None, :], np.full(1, -1, np.int32))
    want = want.dump_tree(0, 1)["prior"][0]
    for mod in (mz_tree, ez_tree):
        n = mod.Node(0.0, [0, 1, 2, 3, 4])
        n.expand(-1, 0, 0, 0.25, [float(v) for v in logits])
        c = n.children()
        assert c["n_children"] == 5 and _same_bits(c["prior"], want) and (c["visit"] == 0).all()
    a = ezs_tree.Action([0.5, -0.25], 1)
    assert a.value == [0.5, -0.25] and a.is_root_action == 1
    n = ezs_tree.Node(0.0, [], 3, 7, True)
    n.expand(-1, 0, 0, 0.0, [0.1, -0.2, 0.3, 0.5, 0.4, 0.6])  # mu[3] | sigma[3]
    acts = n.sampled_actions()
    assert len(acts) == 7 and all(len(x.value) == 3 and all(-1.0 <= v <= 1.0 for v in x.value) for x in acts)
