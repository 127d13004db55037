"""Root-batch prologue / epilogue on the device: the per-root Python loops that bracket every ``search`` call in
LightZero's policies and reanalyze buffers, as batched calls over device tensors.

    legal = legal_from_mask(action_mask)                     # policy/muzero.py:671, game_buffer_muzero.py:552
    noise = dirichlet_noise(alpha, legal[0], A, seed=...)    # policy/muzero.py:674-677
    roots = MCTSCtree.roots(N, legal); roots.prepare(w, noise, rewards, logits, to_play); mcts.search(...)
    visits, count = roots.get_distributions_tensor()
    action_pos, action, entropy = select_action(visits, count, legal[1], temperature, deterministic)   # utils.py:515-539
    policy = target_policies(visits, count, legal[1])        # game_buffer_muzero.py:617-656

Everything stays a device tensor; ``Roots`` accepts the ``(count, ids)`` pair and ``[N, A]`` noise directly.
"""
import ctypes as C
import itertools

import torch

from . import _lib
from ._lib import SelectAction, check
from .forest import _dev, _p, _stream

_counter = itertools.count(1)


def _cuda(device=None):
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def legal_from_mask(action_mask, device=None):
    """``[[a for a, x in enumerate(mask_j) if x == 1] for j]`` as ``(count [N] i32, ids [N, A] i32)`` device tensors."""
    m = action_mask if isinstance(action_mask, torch.Tensor) else torch.as_tensor(action_mask)
    d = m.device if m.is_cuda else _cuda(device)
    m = (m.to(d) == 1).to(torch.int8).contiguous()
    if m.dim() != 2:
        raise ValueError("action_mask must be [N, A]")
    N, A = m.shape
    cnt = torch.empty(N, dtype=torch.int32, device=d)
    ids = torch.empty((N, A), dtype=torch.int32, device=d)
    check(_lib.lib().lzb_legal_from_mask(_p(m), N, A, _p(cnt), _p(ids), _stream(d)))
    return cnt, ids


def dirichlet_noise(alpha, legal_count, action_space, num_roots=None, seed=0, counter=None, root_offset=0, device=None):
    """Dirichlet(alpha) over each root's legal actions as the ``[N, A]`` tensor ``Roots.prepare`` takes (row j: one
    draw of dimension n_legal_j by legal position, zeros beyond).  ``legal_count`` None = every action legal."""
    d = legal_count.device if isinstance(legal_count, torch.Tensor) and legal_count.is_cuda else _cuda(device)
    cnt = _dev(legal_count, torch.int32, d)
    N = int(num_roots if cnt is None else cnt.numel())
    out = torch.empty((N, int(action_space)), dtype=torch.float32, device=d)
    c = next(_counter) if counter is None else int(counter)
    check(_lib.lib().lzb_dirichlet_noise(float(alpha), _p(cnt), _p(out), N, int(action_space), int(root_offset),
                                         int(seed), c & 0xffffffff, _stream(d)))
    return out


def select_action(visit_counts, legal_count=None, legal_actions=None, temperature=1.0, deterministic=True, seed=0,
                  counter=None, root_offset=0, uniforms=None):
    """Batched ``select_action`` (lzero/policy/utils.py:515-539): returns ``(action_pos [N] i32, action [N] i32,
    entropy [N] f64)`` — the index within the legal set, the action id it maps to, and the base-2 entropy of the
    visit-count distribution."""
    v = visit_counts
    if not (isinstance(v, torch.Tensor) and v.is_cuda):
        raise ValueError("visit_counts must be a [N, A] device tensor (Roots.get_distributions_tensor)")
    d = v.device
    v = v.to(torch.int32).contiguous()
    N, A = v.shape
    cnt, ids, u = _dev(legal_count, torch.int32, d), _dev(legal_actions, torch.int32, d), _dev(uniforms, torch.float64, d)
    pos = torch.empty(N, dtype=torch.int32, device=d)
    act = torch.empty(N, dtype=torch.int32, device=d)
    ent = torch.empty(N, dtype=torch.float64, device=d)
    c = next(_counter) if counter is None else int(counter)
    args = SelectAction(v.data_ptr(), _p(cnt), _p(ids), _p(u), pos.data_ptr(), act.data_ptr(), ent.data_ptr(),
                        float(temperature), N, A, int(bool(deterministic)), int(root_offset), int(seed), c & 0xffffffff)
    check(_lib.lib().lzb_select_action(C.byref(args), _stream(d)))
    return pos, act, ent


def target_policies(visit_counts, legal_count=None, legal_actions=None, policy_mask=None):
    """Reanalyze policy targets (game_buffer_muzero.py:617-656): ``[N, A]`` f32, ``visits / sum`` at the legal action
    ids (``legal_actions`` None = fixed action space), zero rows where ``policy_mask == 0``."""
    v = visit_counts
    if not (isinstance(v, torch.Tensor) and v.is_cuda):
        raise ValueError("visit_counts must be a [N, A] device tensor")
    d = v.device
    v = v.to(torch.int32).contiguous()
    N, A = v.shape
    cnt, ids, pm = _dev(legal_count, torch.int32, d), _dev(legal_actions, torch.int32, d), _dev(policy_mask, torch.int32, d)
    out = torch.empty((N, A), dtype=torch.float32, device=d)
    check(_lib.lib().lzb_target_policies(v.data_ptr(), _p(cnt), _p(ids), _p(pm), N, A, out.data_ptr(), _stream(d)))
    return out
