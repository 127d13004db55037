"""Generates tests/golden/reanalyze_targets.npz by EXECUTING the reference's own source lines - not a restatement -
on seeded inputs: /root/reference/lzero/mcts/buffer/game_buffer_muzero.py:464-512 (value / reward targets), :606-656
(policy targets) and game_buffer_efficientzero.py:261-316 (value-prefix targets -> tests/golden/reanalyze_prefix_targets.npz), the code after the search call in _compute_target_reward_value / _compute_target_policy_reanalyzed.
The module itself cannot be imported here (it needs DI-engine); the two line ranges only use numpy, `self._cfg`, and the
search results, which this script supplies (a stand-in `roots` whose getters return the given lists).
Run in the authoring container only:   python tests/golden/make_reanalyze_golden.py
"""
import os
import textwrap
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(os.environ.get("LZB_REFERENCE_ROOT", "/root/reference"), "lzero", "mcts", "buffer", "game_buffer_muzero.py")


def reference_lines(first, last):
    lines = open(SRC).read().split("\n")[first - 1:last]
    return textwrap.dedent("\n".join(lines))


def cfg(**kw):
    c = types.SimpleNamespace(**kw)
    c.model = types.SimpleNamespace(action_space_size=kw.get("action_space_size"))
    return c


def value_case(rng, board_games, B=48, U=5):
    gamma = 1.0 if board_games else 0.997
    T = B * (U + 1)
    seg_len = rng.integers(3, 40, B)
    reward_len = seg_len + rng.integers(0, 12, B)
    pos = np.asarray([rng.integers(0, l + 3) for l in seg_len])
    rewards = [rng.standard_normal(l).astype(np.float32) for l in reward_len]
    to_play = [[1 + (k % 2) for k in range(l + 8)] for l in reward_len]
    td = rng.integers(0, 6, T).astype(np.int64)
    vmask = (rng.random(T) < 0.85).astype(np.float32)
    values = rng.standard_normal(T).astype(np.float32)
    ns = dict(np=np, self=types.SimpleNamespace(_cfg=cfg(env_type="board_games" if board_games else "not_board_games",
                                                         discount_factor=gamma, num_unroll_steps=U)),
              value_list=np.array([float(v) for v in values]), td_steps_list=td.tolist(), transition_batch_size=T,
              value_mask=vmask.tolist(), to_play_segment=to_play if board_games else [[-1] * len(x) for x in to_play],
              game_segment_lens=seg_len.tolist(), rewards_list=[[float(r) for r in x] for x in rewards],
              pos_in_game_segment_list=pos.tolist(), batch_target_values=[], batch_rewards=[])
    exec(reference_lines(464, 512), ns)
    return dict(values=values, td=td.astype(np.int32), vmask=vmask, seg_len=seg_len.astype(np.int32), reward_len=reward_len.astype(np.int32),
                pos=pos.astype(np.int32), rewards=np.concatenate(rewards), to_play=np.concatenate([np.asarray(x, np.int32) for x in to_play]),
                U=np.int32(U), gamma=np.float64(gamma), out_values=np.asarray(ns["batch_target_values"], np.float64),
                out_rewards=np.asarray(ns["batch_rewards"], np.float64))


SRC_EZ = os.path.join(os.path.dirname(SRC), "game_buffer_efficientzero.py")


def value_prefix_case(rng, board_games, B=40, U=5, H=5):
    """EfficientZero: game_buffer_efficientzero.py:261-316 executed as it stands (value-prefix targets, horizon resets)."""
    gamma = 1.0 if board_games else 0.997
    T = B * (U + 1)
    seg_len = rng.integers(3, 40, B)
    reward_len = seg_len + rng.integers(0, 12, B)
    pos = np.asarray([rng.integers(0, max(l - 2, 1)) for l in seg_len])  # reward_list[current_index] is read while inside the segment
    rewards = [rng.standard_normal(l).astype(np.float32) for l in reward_len]
    to_play = [[1 + (k % 2) for k in range(l + 8)] for l in reward_len]
    td = rng.integers(0, 6, T).astype(np.int64)
    vmask = (rng.random(T) < 0.85).astype(np.float32)
    values = rng.standard_normal(T).astype(np.float32)
    ns = dict(np=np, self=types.SimpleNamespace(_cfg=cfg(env_type="board_games" if board_games else "not_board_games",
                                                         discount_factor=gamma, num_unroll_steps=U, lstm_horizon_len=H)),
              value_list=np.array([float(v) for v in values]), td_steps_list=td.tolist(), transition_batch_size=T,
              value_mask=vmask.tolist(), to_play_segment=to_play if board_games else [[-1] * len(x) for x in to_play],
              game_segment_lens=seg_len.tolist(), rewards_list=[[float(r) for r in x] for x in rewards],
              pos_in_game_segment_list=pos.tolist(), batch_target_values=[], batch_value_prefixs=[])
    lines = open(SRC_EZ).read().split("\n")[261 - 1:316]
    exec(textwrap.dedent("\n".join(lines)), ns)
    return dict(values=values, td=td.astype(np.int32), vmask=vmask, seg_len=seg_len.astype(np.int32), reward_len=reward_len.astype(np.int32),
                pos=pos.astype(np.int32), rewards=np.concatenate(rewards), to_play=np.concatenate([np.asarray(x, np.int32) for x in to_play]),
                U=np.int32(U), H=np.int32(H), gamma=np.float64(gamma), out_values=np.asarray(ns["batch_target_values"], np.float64),
                out_prefixes=np.asarray(ns["batch_value_prefixs"], np.float64))


def policy_case(rng, fixed, B=32, U=5, S=50):
    A = 6 if fixed else 9
    T = B * (U + 1)
    mask = np.ones((T, A), np.int8)
    if not fixed:
        mask = (rng.random((T, A)) < 0.6).astype(np.int8)
        mask[np.arange(T), rng.integers(0, A, T)] = 1
    legal = [[i for i, x in enumerate(row) if x == 1] for row in mask]
    dists = [rng.multinomial(S, rng.dirichlet([0.5] * len(l))).tolist() for l in legal]
    pmask = (rng.random(T) < 0.8).astype(np.int32)
    child_visits = [dict() for _ in range(B)]  # child_visit[current_index] = ... : any indexable that accepts assignment
    pos = rng.integers(0, 20, B).tolist()
    roots = types.SimpleNamespace(get_distributions=lambda: dists, get_values=lambda: [0.0] * T)
    ns = dict(np=np, self=types.SimpleNamespace(_cfg=cfg(num_unroll_steps=U, action_space_size=A,
                                                         action_type="fixed_action_space" if fixed else "varied_action_space")),
              legal_actions=legal, roots=roots, pos_in_game_segment_list=pos, child_visits=child_visits,
              batch_index_list=list(range(B)), policy_mask=pmask.tolist(), batch_target_policies_re=[])
    exec(reference_lines(606, 656), ns)
    return dict(mask=mask, pmask=pmask, dists=np.concatenate([np.asarray(d, np.int32) for d in dists]), U=np.int32(U),
                out=np.asarray(ns["batch_target_policies_re"], np.float64))


def main():
    rng = np.random.default_rng(20261105)
    data = {}
    for name, case in (("value_single_player", value_case(rng, False)), ("value_board_games", value_case(rng, True)),
                       ("policy_fixed", policy_case(rng, True)), ("policy_varied", policy_case(rng, False))):
        for k, v in case.items():
            data[name + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "reanalyze_targets.npz"), **data)
    print("wrote", sorted(data))
    # EfficientZero's value-prefix targets: their own file and generator, so that the MuZero fixtures above stay byte-identical
    rng = np.random.default_rng(20261106)
    data = {}
    for name, case in (("prefix_single_player", value_prefix_case(rng, False)), ("prefix_board_games", value_prefix_case(rng, True)),
                       ("prefix_horizon3", value_prefix_case(rng, False, U=7, H=3))):
        for k, v in case.items():
            data[name + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "reanalyze_prefix_targets.npz"), **data)
    print("wrote", sorted(data))


if __name__ == "__main__":
    main()
