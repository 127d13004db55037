"""Caller-shaped device paths: what LightZero's two callers of ``search()`` do AROUND it, from the outputs of
``initial_inference`` to actions / training targets, as device tensors end to end.

    collect_search            lzero/policy/muzero.py:655-725 (_forward_collect; eval :800-860 with eval_mode=True)
    reanalyze_policy_targets  lzero/mcts/buffer/game_buffer_muzero.py:519-660 (_compute_target_policy_reanalyzed)
    reanalyze_value_targets   lzero/mcts/buffer/game_buffer_muzero.py:367-518 (_compute_target_reward_value)

The reference builds Python lists of lists for every root on both sides of the search (legal actions from the mask,
Dirichlet noise, ``.tolist()`` of logits and rewards, visit-count lists, per-root ``select_action``, per-segment target
loops).  Here the mask, the logits, the rewards and the root latents stay where ``initial_inference`` left them (HBM),
the bracketing work is the kernels of csrc/lzb_collect.cuh, and nothing is copied to the host unless the caller asks.
The search itself is the unchanged driver (``mcts.search``), eager or as a replayed CUDA graph (cfg.cuda_graph)."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .collect import dirichlet_noise, legal_from_mask, select_action, target_policies
from .forest import _dev, _p, _stream


def _cfg(mcts, key, default=None):
    cfg = mcts._cfg
    try:
        return cfg[key] if key in cfg else default
    except TypeError:
        return getattr(cfg, key, default)


def _search(mcts, model, action_mask, policy_logits, reward_roots, latent_state_roots, to_play, noise_weight, alpha,
            root_values, hidden_state_roots, seed, counter, root_offset, extra_roots_args=()):
    """mask -> legal lists -> Roots -> prepare(+noise) -> search; returns (roots, legal_count, legal_ids)."""
    dev = policy_logits.device
    N, A = policy_logits.shape
    if action_mask is None:  # no mask (continuous / sampled action spaces): the tree has no legal-action list
        cnt = ids = None
        roots = mcts.roots(N, None, *extra_roots_args)
    else:
        cnt, ids = legal_from_mask(action_mask, device=dev)
        roots = mcts.roots(N, (cnt, ids), *extra_roots_args)
    roots.root_offset = int(root_offset)
    to_play = _dev(to_play, torch.int32, dev).reshape(-1)
    reward = _dev(reward_roots, torch.float32, dev).reshape(-1)
    gumbel = getattr(roots, "variant", "mz") == "gmz"
    vals = (() if not gumbel else (_dev(root_values, torch.float32, dev).reshape(-1),))
    if noise_weight and noise_weight > 0:
        width = int(getattr(roots, "num_of_sampled_actions", A))  # the Sampled tree's noise is over its K sampled actions
        if cnt is None:
            cnt = torch.full((N,), width, dtype=torch.int32, device=dev)
        noise = dirichlet_noise(alpha, cnt, width, seed=seed, counter=counter, root_offset=root_offset, device=dev)
        roots.prepare(noise_weight, noise, reward, *vals, policy_logits, to_play)
    else:
        roots.prepare_no_noise(reward, *vals, policy_logits, to_play)
    if hidden_state_roots is not None:
        mcts.search(roots, model, latent_state_roots, hidden_state_roots, to_play)
    else:
        mcts.search(roots, model, latent_state_roots, to_play)
    return roots, cnt, ids


def collect_search(mcts, model, latent_state_roots, reward_roots, policy_logits, action_mask, to_play, temperature=1.0,
                   deterministic=False, eval_mode=False, root_values=None, hidden_state_roots=None, seed=0, counter=None,
                   root_offset=0, uniforms=None):
    """One collect (or, with ``eval_mode``, evaluation) step for a batch of environments, from the unpacked output of
    ``model.initial_inference`` to the actions the environments take.

    muzero.py:671-677  legal actions from ``action_mask``, Dirichlet(root_dirichlet_alpha) noise per root -> kernels
    muzero.py:680-690  Roots / prepare / search                                                          -> the driver
    muzero.py:693-719  per env: select_action(distributions, temperature) + action-id lookup + entropy    -> one kernel
    Evaluation (muzero.py:827-850): no noise, deterministic arg-max.

    Returns a dict of DEVICE tensors: action [N] i32 (ids in the whole action space), action_pos [N] i32 (index within
    the legal set), visit_counts [N, A] i32 (row i: legal_count[i] valid entries), legal_count [N] i32, searched_value
    [N] f32, entropy [N] f64 - and ``roots`` for callers that want the list getters."""
    policy_logits = _dev(policy_logits, torch.float32, torch.device("cuda", torch.cuda.current_device())
                         if not (isinstance(policy_logits, torch.Tensor) and policy_logits.is_cuda) else policy_logits.device)
    w = 0.0 if eval_mode else float(_cfg(mcts, "root_noise_weight", 0.25))
    roots, cnt, ids = _search(mcts, model, action_mask, policy_logits, reward_roots, latent_state_roots, to_play, w,
                              float(_cfg(mcts, "root_dirichlet_alpha", 0.3)), root_values, hidden_state_roots, seed, counter,
                              root_offset)
    visits, legal_count = roots.get_distributions_tensor()
    pos, act, ent = select_action(visits, legal_count, ids, temperature=temperature,
                                  deterministic=bool(deterministic or eval_mode), seed=seed, counter=counter,
                                  root_offset=root_offset, uniforms=uniforms)
    return dict(action=act, action_pos=pos, visit_counts=visits, legal_count=legal_count,
                searched_value=roots.get_values_tensor(), entropy=ent, roots=roots)


def reanalyze_policy_targets(mcts, model, latent_state_roots, reward_pool, policy_logits_pool, action_mask, to_play,
                             policy_mask, num_unroll_steps, fixed_action_space=True, reanalyze_noise=None,
                             root_values=None, hidden_state_roots=None, seed=0, counter=None, root_offset=0, roots_args=()):
    """_compute_target_policy_reanalyzed (game_buffer_muzero.py:519-660) for a transition batch of
    T = batch_size * (num_unroll_steps + 1) roots whose ``initial_inference`` outputs are on the device.

    :552       legal actions from the mask            -> lzb_legal_from_mask
    :580-591   Dirichlet noise (cfg.reanalyze_noise), Roots, prepare / prepare_no_noise, search -> the driver
    :611-656   per root: visits / sum(visits) placed at the legal action ids, zero rows where policy_mask == 0
                                                      -> lzb_target_policies (float64 division, float32 result)
    Returns dict(target_policies [B, U+1, A] f32, child_visits [T, A] f32 (the normalised distributions the reference
    writes back into ``child_visit_segment``, in legal-action order), root_values [T] f32, visit_counts, legal_count).

    EfficientZero (game_buffer_efficientzero.py:337-464): pass ``hidden_state_roots``.  Sampled EfficientZero
    (game_buffer_sampled_efficientzero.py:437-585): ``roots_args=(action_space_size, num_of_sampled_actions,
    continuous_action_space)`` as its ``roots()`` takes them, ``action_mask=None`` for continuous actions; the targets are
    then over the K sampled actions and ``root_sampled_actions`` [T, K, D] comes back with them (:526-533)."""
    logits = policy_logits_pool
    dev = logits.device
    T_, A = logits.shape
    per = int(num_unroll_steps) + 1
    if T_ % per:
        raise ValueError("the transition batch (%d roots) is not a multiple of num_unroll_steps + 1 = %d" % (T_, per))
    if reanalyze_noise is None:
        reanalyze_noise = bool(_cfg(mcts, "reanalyze_noise", False))
    w = float(_cfg(mcts, "root_noise_weight", 0.25)) if reanalyze_noise else 0.0
    roots, cnt, ids = _search(mcts, model, action_mask, logits, reward_pool, latent_state_roots, to_play, w,
                              float(_cfg(mcts, "root_dirichlet_alpha", 0.3)), root_values, hidden_state_roots, seed, counter,
                              root_offset, extra_roots_args=tuple(roots_args))
    visits, legal_count = roots.get_distributions_tensor()
    pm = _dev(policy_mask, torch.int32, dev)
    pol = target_policies(visits, legal_count, None if fixed_action_space else ids, pm)
    child = target_policies(visits, legal_count, None, pm)  # in legal-action order, as child_visit stores it (:625-626)
    out = dict(target_policies=pol.reshape(T_ // per, per, pol.shape[1]), child_visits=child, root_values=roots.get_values_tensor(),
               visit_counts=visits, legal_count=legal_count, roots=roots)
    if hasattr(roots, "get_sampled_actions_tensor"):  # Sampled EfficientZero (game_buffer_sampled_efficientzero.py:526-533)
        out["root_sampled_actions"] = roots.get_sampled_actions_tensor()
    return out


def reanalyze_value_targets(root_values, td_steps, value_mask, rewards, reward_lens, game_segment_lens,
                            pos_in_game_segment, num_unroll_steps, discount_factor, to_play_segments=None,
                            board_games=False, lstm_horizon_len=0):
    """The tail of _compute_target_reward_value (game_buffer_muzero.py:463-513): bootstrapped n-step value targets and
    reward targets of a transition batch, one kernel (lzb_value_targets), float64 like the reference's Python floats.

    root_values [T] f32: the searched root values (cfg.use_root_value: ``roots.get_values_tensor()`` of a search run as
    in reanalyze_policy_targets) or the model's predicted values.  td_steps [T] int, value_mask [T]; rewards [B, Lmax]
    f32 = each game segment's reward list zero-padded, reward_lens [B] their true lengths; game_segment_lens [B];
    pos_in_game_segment [B]; to_play_segments [B, Lmax] int for board games (env_type == 'board_games' and to_play in
    {1, 2}).  Returns (target_rewards [B, U+1] f64, target_values [B, U+1] f64) on the device.

    ``lstm_horizon_len`` > 0 is EfficientZero's flavour (game_buffer_efficientzero.py:261-316; the Sampled buffer's is the
    same code): the first result is then the target VALUE PREFIX - the running sum of rewards since the last horizon reset,
    with the reference's batch-wide reset counter - instead of the per-step reward."""
    rv = root_values if isinstance(root_values, torch.Tensor) and root_values.is_cuda else _dev(
        root_values, torch.float32, torch.device("cuda", torch.cuda.current_device()))
    dev = rv.device
    rv = rv.to(torch.float32).reshape(-1).contiguous()
    per = int(num_unroll_steps) + 1
    T_ = rv.numel()
    B = T_ // per
    if B * per != T_:
        raise ValueError("root_values must hold batch_size * (num_unroll_steps + 1) entries")
    td_host = np.asarray(td_steps.cpu() if isinstance(td_steps, torch.Tensor) else td_steps).reshape(-1)
    td = _dev(td_steps, torch.int32, dev).reshape(-1)
    vm = _dev(value_mask, torch.float32, dev).reshape(-1)
    rew = _dev(rewards, torch.float32, dev)
    if rew.dim() != 2 or rew.shape[0] != B:
        raise ValueError("rewards must be [batch_size, Lmax] (zero-padded reward lists)")
    rl, sl, pos = (_dev(x, torch.int32, dev).reshape(-1) for x in (reward_lens, game_segment_lens, pos_in_game_segment))
    if not (td.numel() == vm.numel() == T_ and rl.numel() == sl.numel() == pos.numel() == B):
        raise ValueError("td_steps / value_mask need T entries, reward_lens / game_segment_lens / pos_in_game_segment B")
    tp = None
    if board_games:
        tp = _dev(to_play_segments, torch.int32, dev)
        if tp.dim() != 2 or tp.shape[0] != B:
            raise ValueError("to_play_segments must be [batch_size, Lmax]")
    # discount ** k on the HOST (Python's pow == the host libm's, what the reference evaluates)
    kmax = max(int(td_host.max(initial=0)), int(rew.shape[1])) + 1
    table = torch.tensor([float(discount_factor) ** k for k in range(kmax + 1)], dtype=torch.float64).to(dev)
    tv = torch.empty((B, per), dtype=torch.float64, device=dev)
    tr = torch.empty((B, per), dtype=torch.float64, device=dev)
    args = _lib.ValueTargets(rv.data_ptr(), vm.data_ptr(), rew.data_ptr(), td.data_ptr(), rl.data_ptr(), sl.data_ptr(),
                             pos.data_ptr(), None if tp is None else tp.data_ptr(), table.data_ptr(), tv.data_ptr(),
                             tr.data_ptr(), B, int(num_unroll_steps), int(rew.shape[1]),
                             0 if tp is None else int(tp.shape[1]), int(table.numel()), int(bool(board_games)),
                             int(lstm_horizon_len))
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().lzb_value_targets(C.byref(args), _stream(dev)))
    return tr, tv
