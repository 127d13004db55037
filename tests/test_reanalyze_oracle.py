"""oracle/reanalyze_oracle.py against tests/golden/reanalyze_targets.npz - outputs of the reference's OWN source lines
(game_buffer_muzero.py:464-512 and :606-656, executed by tests/golden/make_reanalyze_golden.py) - bit for bit in
float64.  The GPU twin is tests/test_gpu_callers.py."""
import os

import numpy as np
import pytest

from oracle import reanalyze_oracle as ro

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reanalyze_targets.npz"))


def split(flat, lens):
    out, at = [], 0
    for l in lens:
        out.append(flat[at:at + l])
        at += l
    return out


def value_inputs(name):
    g = lambda k: GOLD[name + "/" + k]
    rewards = split(g("rewards"), g("reward_len"))
    to_play = split(g("to_play"), g("reward_len") + 8)
    return g, rewards, to_play


@pytest.mark.parametrize("name", ["value_single_player", "value_board_games"])
def test_value_targets_match_the_reference_lines(name):
    g, rewards, to_play = value_inputs(name)
    board = name.endswith("board_games")
    r, v = ro.value_targets(g("values").tolist(), g("td").tolist(), g("vmask").tolist(), rewards, g("seg_len").tolist(),
                            g("pos").tolist(), [x.tolist() for x in to_play], int(g("U")), float(g("gamma")), board_games=board)
    assert np.array_equal(np.asarray(v, np.float64).view(np.uint64), g("out_values").view(np.uint64))
    assert np.array_equal(np.asarray(r, np.float64).view(np.uint64), g("out_rewards").view(np.uint64))


@pytest.mark.parametrize("name", ["policy_fixed", "policy_varied"])
def test_policy_targets_match_the_reference_lines(name):
    g = lambda k: GOLD[name + "/" + k]
    mask = g("mask")
    legal = [[i for i, x in enumerate(row) if x == 1] for row in mask]
    dists = [d.tolist() for d in split(g("dists"), [len(l) for l in legal])]
    U = int(g("U"))
    got, _ = ro.policy_targets(dists, legal, g("pmask").tolist(), len(legal) // (U + 1), U, mask.shape[1],
                               fixed_action_space=name == "policy_fixed")
    assert np.array_equal(np.asarray(got, np.float64).view(np.uint64), g("out").view(np.uint64))


GOLD_EZ = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reanalyze_prefix_targets.npz"))


@pytest.mark.parametrize("name", ["prefix_single_player", "prefix_board_games", "prefix_horizon3"])
def test_value_prefix_targets_match_the_reference_lines(name):
    """EfficientZero (game_buffer_efficientzero.py:261-316 executed by the golden maker): value-prefix targets with the
    batch-wide horizon counter."""
    g = lambda k: GOLD_EZ[name + "/" + k]
    rewards = split(g("rewards"), g("reward_len"))
    to_play = split(g("to_play"), g("reward_len") + 8)
    p, v = ro.value_prefix_targets(g("values").tolist(), g("td").tolist(), g("vmask").tolist(), rewards, g("seg_len").tolist(),
                                   g("pos").tolist(), [x.tolist() for x in to_play], int(g("U")), float(g("gamma")), int(g("H")),
                                   board_games=name.endswith("board_games"))
    assert np.array_equal(np.asarray(v, np.float64).view(np.uint64), g("out_values").view(np.uint64))
    assert np.array_equal(np.asarray(p, np.float64).view(np.uint64), g("out_prefixes").view(np.uint64))
