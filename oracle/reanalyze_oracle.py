"""TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).  Restatement of the two reanalyze functions of the MuZero game
buffer AROUND the search call, for checking lightzero_b200/callers.py:

  policy_targets   lzero/mcts/buffer/game_buffer_muzero.py:611-656 (_compute_target_policy_reanalyzed after the search)
  value_targets    lzero/mcts/buffer/game_buffer_muzero.py:463-513 (_compute_target_reward_value after the search)

Both are the reference's own Python loops over Python floats and lists, restated line by line with the search results
(roots_distributions, roots_values) as inputs - the search itself is pinned elsewhere.  Parity: the arithmetic is the
reference's own Python / numpy float64 arithmetic (no third-party algorithm to restate); with numpy >= 2 a float32 scalar
times a Python float stays float32, so rewards are widened to Python floats first, which is what numpy 1.x (the version
LightZero pins) computes."""
import numpy as np


def policy_targets(roots_distributions, roots_legal_actions_list, policy_mask, batch_size, num_unroll_steps,
                   action_space_size, fixed_action_space=True):
    """game_buffer_muzero.py:611-656.  Returns (batch_target_policies_re [B][U+1][A], child_visits {policy_index: list})."""
    batch_target_policies_re, child_visit = [], {}
    policy_index = 0
    for _ in range(batch_size):
        target_policies = []
        for _current in range(num_unroll_steps + 1):
            distributions = roots_distributions[policy_index]
            if policy_mask[policy_index] == 0:
                target_policies.append([0 for _ in range(action_space_size)])  # :620-622
            else:
                sum_visits = sum(distributions)
                child_visit[policy_index] = [visit_count / sum_visits for visit_count in distributions]  # :625-626
                if fixed_action_space:  # :640-644
                    policy = [visit_count / sum_visits for visit_count in distributions]
                    target_policies.append(policy)
                else:  # :645-653
                    policy_tmp = [0 for _ in range(action_space_size)]
                    policy = [visit_count / sum_visits for visit_count in distributions]
                    for index, legal_action in enumerate(roots_legal_actions_list[policy_index]):
                        policy_tmp[legal_action] = policy[index]
                    target_policies.append(policy_tmp)
            policy_index += 1
        batch_target_policies_re.append(target_policies)
    return batch_target_policies_re, child_visit


def value_targets(roots_values, td_steps_list, value_mask, rewards_list, game_segment_lens, pos_in_game_segment_list,
                  to_play_segment, num_unroll_steps, discount_factor, board_games=False):
    """game_buffer_muzero.py:463-513.  Returns (batch_rewards [B][U+1], batch_target_values [B][U+1]) of Python floats."""
    transition_batch_size = len(roots_values)
    value_list = np.array([float(v) for v in roots_values])
    if board_games:  # :466-473
        value_list = value_list.reshape(-1) * np.array([
            discount_factor ** td_steps_list[i] if int(td_steps_list[i]) % 2 == 0 else -discount_factor ** td_steps_list[i]
            for i in range(transition_batch_size)])
    else:  # :474-477
        value_list = value_list.reshape(-1) * (
            np.array([discount_factor for _ in range(transition_batch_size)]) ** np.asarray(td_steps_list))
    value_list = value_list * np.array(value_mask)  # :479
    value_list = value_list.tolist()
    value_index = 0
    batch_rewards, batch_target_values = [], []
    for game_segment_len_non_re, reward_list, state_index, to_play_list in zip(game_segment_lens, rewards_list,
                                                                               pos_in_game_segment_list, to_play_segment):
        target_values, target_rewards = [], []
        base_index = state_index
        reward_list = [float(r) for r in reward_list]
        for current_index in range(state_index, state_index + num_unroll_steps + 1):
            bootstrap_index = current_index + td_steps_list[value_index]
            for i, reward in enumerate(reward_list[current_index:bootstrap_index]):  # :491
                if board_games:
                    if to_play_list[base_index] == to_play_list[i]:  # :496
                        value_list[value_index] += reward * discount_factor ** i
                    else:
                        value_list[value_index] += -reward * discount_factor ** i
                else:
                    value_list[value_index] += reward * discount_factor ** i  # :501
            if current_index < game_segment_len_non_re:  # :505-510
                target_values.append(value_list[value_index])
                target_rewards.append(reward_list[current_index])
            else:
                target_values.append(0)
                target_rewards.append(0.0)
            value_index += 1
        batch_rewards.append(target_rewards)
        batch_target_values.append(target_values)
    return batch_rewards, batch_target_values


def value_prefix_targets(roots_values, td_steps_list, value_mask, rewards_list, game_segment_lens, pos_in_game_segment_list,
                         to_play_segment, num_unroll_steps, discount_factor, lstm_horizon_len, board_games=False):
    """EfficientZero: game_buffer_efficientzero.py:261-316 (_compute_target_reward_value after the search).  The reward
    target is the VALUE PREFIX: the running sum of rewards since the last horizon reset, where the reset counter
    ``horizon_id`` runs over the whole transition batch (it is not restarted per game segment, :278) and a reset also moves
    ``base_index`` of the board-games sign rule.  Returns (batch_value_prefixs [B][U+1], batch_target_values [B][U+1])."""
    transition_batch_size = len(roots_values)
    value_list = np.array([float(v) for v in roots_values])
    if board_games:  # :262-270
        value_list = value_list.reshape(-1) * np.array([
            discount_factor ** td_steps_list[i] if int(td_steps_list[i]) % 2 == 0 else -discount_factor ** td_steps_list[i]
            for i in range(transition_batch_size)])
    else:  # :271-274
        value_list = value_list.reshape(-1) * (
            np.array([discount_factor for _ in range(transition_batch_size)]) ** np.asarray(td_steps_list))
    value_list = value_list * np.array(value_mask)  # :276
    value_list = value_list.tolist()
    horizon_id, value_index = 0, 0  # :278
    batch_value_prefixs, batch_target_values = [], []
    for game_segment_len_non_re, reward_list, state_index, to_play_list in zip(game_segment_lens, rewards_list,
                                                                               pos_in_game_segment_list, to_play_segment):
        target_values, target_value_prefixs = [], []
        value_prefix = 0.0
        base_index = state_index
        reward_list = [float(r) for r in reward_list]
        for current_index in range(state_index, state_index + num_unroll_steps + 1):
            bootstrap_index = current_index + td_steps_list[value_index]
            for i, reward in enumerate(reward_list[current_index:bootstrap_index]):  # :287
                if board_games:
                    if to_play_list[base_index] == to_play_list[i]:  # :290
                        value_list[value_index] += reward * discount_factor ** i
                    else:
                        value_list[value_index] += -reward * discount_factor ** i
                else:
                    value_list[value_index] += reward * discount_factor ** i  # :295
            if horizon_id % lstm_horizon_len == 0:  # :298-300 reset every lstm_horizon_len
                value_prefix = 0.0
                base_index = current_index
            horizon_id += 1
            if current_index < game_segment_len_non_re:  # :303-311
                target_values.append(value_list[value_index])
                value_prefix += reward_list[current_index]
                target_value_prefixs.append(value_prefix)
            else:
                target_values.append(0)
                target_value_prefixs.append(value_prefix)
            value_index += 1
        batch_value_prefixs.append(target_value_prefixs)
        batch_target_values.append(target_values)
    return batch_value_prefixs, batch_target_values
