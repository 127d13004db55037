"""TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).  CPU restatement of the per-root Python code that brackets a
search in LightZero's policies / reanalyze buffers, for checking lightzero_b200/collect.py:

  legal_from_mask   lzero/policy/muzero.py:671; lzero/mcts/buffer/game_buffer_muzero.py:552
  dirichlet         lzero/policy/muzero.py:674-677 (np.random.dirichlet: distribution only, see below)
  select_action     lzero/policy/utils.py:515-539 (+ the action-id lookup of lzero/policy/muzero.py:709)
  target_policies   lzero/mcts/buffer/game_buffer_muzero.py:617-656

Parity status: select_action / target_policies / legal_from_mask are PINNED: tests/test_collect_oracle.py checks this
restatement against scipy.stats.entropy and numpy's RandomState.choice themselves (the two library calls the
reference makes) on seeded inputs.  The Dirichlet draw is parity-unpinned by design: the reference samples from
numpy's process-global Mersenne Twister; only its distribution is restated (moments checked in the tests).
"""
import math

import numpy as np


def legal_from_mask(action_mask):
    return [[i for i, x in enumerate(row) if x == 1] for row in action_mask]


def select_action(visit_counts, temperature=1, deterministic=True, uniform=None):
    """utils.py:530-538 with np.random.choice and scipy.stats.entropy written out.
    ``uniform``: the value ``random_sample()`` would have returned inside np.random.choice."""
    action_probs = [visit_count_i ** (1 / temperature) for visit_count_i in visit_counts]
    total = sum(action_probs)
    action_probs = [x / total for x in action_probs]
    if deterministic:
        action_pos = int(np.argmax([v for v in visit_counts]))
    else:  # numpy/random/mtrand.pyx RandomState.choice: cdf = p.cumsum(); cdf /= cdf[-1]; searchsorted(side='right')
        cdf = np.asarray(action_probs, np.float64).cumsum()
        cdf /= cdf[-1]
        action_pos = int(cdf.searchsorted(uniform, side="right"))
    # scipy.stats.entropy(pk, base=2): pk = pk / sum(pk); S = sum(entr(pk)) / log(2)
    pk_total = 0.0
    for p in action_probs:
        pk_total += p
    s = 0.0
    for p in action_probs:
        pk = p / pk_total
        if pk > 0:
            s += -pk * math.log(pk)
    return action_pos, s / math.log(2.0)


def target_policies(distributions, legal_actions, action_space_size, policy_mask=None, fixed_action_space=True):
    """game_buffer_muzero.py:617-656 for one flat batch of roots -> float32 [N, A] (the learner casts to float32)."""
    out = []
    for k, dist in enumerate(distributions):
        if policy_mask is not None and policy_mask[k] == 0:
            out.append([0 for _ in range(action_space_size)])
            continue
        sum_visits = sum(dist)
        policy = [visit_count / sum_visits for visit_count in dist]
        if fixed_action_space:
            out.append(policy + [0] * (action_space_size - len(policy)))
        else:
            policy_tmp = [0 for _ in range(action_space_size)]
            for index, legal_action in enumerate(legal_actions[k]):
                policy_tmp[legal_action] = policy[index]
            out.append(policy_tmp)
    return np.array(out, dtype=np.float64).astype(np.float32)
