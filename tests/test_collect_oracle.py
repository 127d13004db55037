"""Pins oracle/collect_oracle.py against the library calls the reference itself makes (scipy.stats.entropy,
numpy RandomState.choice) — lzero/policy/utils.py:515-539."""
import numpy as np
import pytest

from oracle import collect_oracle as co


def reference_select_action(visit_counts, temperature, deterministic):
    """lzero/policy/utils.py:530-538 verbatim in behaviour, with the two library calls it makes."""
    from scipy.stats import entropy
    action_probs = [v ** (1 / temperature) for v in visit_counts]
    action_probs = [x / sum(action_probs) for x in action_probs]
    if deterministic:
        action_pos = np.argmax([v for v in visit_counts])
    else:
        action_pos = np.random.choice(len(visit_counts), p=action_probs)
    return action_pos, entropy(action_probs, base=2)


@pytest.mark.parametrize("temperature", [1.0, 0.5, 0.25, 0.7])
def test_select_action_restatement_matches_numpy_and_scipy(temperature):
    rng = np.random.default_rng(3)
    for trial in range(300):
        n = int(rng.integers(1, 20))
        visits = rng.multinomial(50, rng.dirichlet([0.5] * n)).tolist()
        pos_ref, ent_ref = reference_select_action(visits, temperature, True)
        pos, ent = co.select_action(visits, temperature, True)
        assert pos == pos_ref
        assert ent == pytest.approx(ent_ref, rel=1e-12, abs=1e-15)
        np.random.seed(1000 + trial)
        u = np.random.random_sample()  # what choice() will draw
        np.random.seed(1000 + trial)
        pos_ref, _ = reference_select_action(visits, temperature, False)
        pos, _ = co.select_action(visits, temperature, False, uniform=u)
        assert pos == pos_ref


def test_target_policies_and_masks():
    mask = np.array([[1, 0, 1, 1], [0, 0, 1, 0], [1, 1, 1, 1]], np.int8)
    legal = co.legal_from_mask(mask)
    assert legal == [[0, 2, 3], [2], [0, 1, 2, 3]]
    dist = [[5, 0, 3], [8], [1, 2, 3, 2]]
    p = co.target_policies(dist, legal, 4, policy_mask=[1, 1, 0], fixed_action_space=False)
    assert p.dtype == np.float32 and p.shape == (3, 4)
    assert p[0].tolist() == [0.625, 0, 0, 0.375] and p[1].tolist() == [0, 0, 1, 0]
    assert (p[2] == 0).all()
