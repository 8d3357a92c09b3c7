"""Tensor-level wrappers of the handle-free C-ABI entry points (K5, K7, K8, a11)."""
from __future__ import annotations

import ctypes as C

from . import _C


def _f32(n, like):
    import torch
    return torch.empty(n, dtype=torch.float32, device=like.device)


def td_workspace(B, device):
    import torch
    n = C.c_int64()
    _C.call("apexb200_td_workspace_bytes", int(B), C.byref(n))
    return torch.empty(int(n.value), dtype=torch.uint8, device=device)


def td_forward(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, alpha, eps=0.0, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0,
               workspace=None, weights=None):
    """learner.py:43-52 on device: returns (delta[B], prio[B], loss[B], loss_mean[1])."""
    B, nA = q_on_t.shape
    delta, prio, loss, mean = _f32(B, q_on_t), _f32(B, q_on_t), _f32(B, q_on_t), _f32(1, q_on_t)
    ws = workspace if workspace is not None else td_workspace(B, q_on_t.device)
    _C.call("apexb200_td_fwd", _C.ptr(q_on_t), _C.ptr(q_on_tpn), _C.ptr(q_tg_tpn), _C.ptr(A), _C.ptr(R), _C.ptr(Gamma),
            int(B), int(nA), float(alpha), float(eps), int(loss_kind), float(kappa), _C.ptr(weights), _C.ptr(delta),
            _C.ptr(prio), _C.ptr(loss), _C.ptr(mean), _C.ptr(ws), int(ws.numel()), _C.stream_ptr())
    return delta, prio, loss, mean


def is_weights(leaf, total, n_records, beta):
    """PER importance weights from the leaf / total outputs of a sample call -> f32 [B], max-normalised."""
    B = leaf.numel()
    w = _f32(B, leaf)
    _C.call("apexb200_is_weights", _C.ptr(leaf), _C.ptr(total), int(B), int(n_records), float(beta), _C.ptr(w),
            _C.stream_ptr())
    return w


def td_forward_noalloc(q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, alpha, eps=0.0, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0,
                       out=None):
    """td_forward for callers that cannot provide (or allocate) a workspace: the reduction scratch
    is a small static buffer inside the library, shared by every call on the GPU.  At most
    _C.TD_NOALLOC_MAX_B rows; calls that may overlap on the GPU must be ordered by the caller.
    `out` = (delta, prio, loss, loss_mean) tensors to write into (allocated here if omitted)."""
    B, nA = q_on_t.shape
    delta, prio, loss, mean = out if out is not None else (_f32(B, q_on_t), _f32(B, q_on_t), _f32(B, q_on_t), _f32(1, q_on_t))
    _C.call("apexb200_td_fwd_noalloc", _C.ptr(q_on_t), _C.ptr(q_on_tpn), _C.ptr(q_tg_tpn), _C.ptr(A), _C.ptr(R),
            _C.ptr(Gamma), int(B), int(nA), float(alpha), float(eps), int(loss_kind), float(kappa), _C.ptr(delta),
            _C.ptr(prio), _C.ptr(loss), _C.ptr(mean), _C.stream_ptr())
    return delta, prio, loss, mean


def td_backward(delta, A, grad_loss_mean, num_actions, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0, count_dev=None,
                mean_divisor=0.0, weights=None):
    import torch
    B = delta.numel()
    grad_q = torch.empty((B, num_actions), dtype=torch.float32, device=delta.device)
    _C.call("apexb200_td_bwd", _C.ptr(delta), _C.ptr(A), _C.ptr(grad_loss_mean), int(B), int(num_actions),
            int(loss_kind), float(kappa), _C.ptr(count_dev), float(mean_divisor), _C.ptr(weights), _C.ptr(grad_q),
            _C.stream_ptr())
    return grad_q


def nstep_returns(rewards, dones, n, gamma):
    """K7 (actor.py:29-43, intended semantics): rewards f32 [E,T], dones u8 [E,T] -> (R, Gamma, span)."""
    import torch
    E, T = rewards.shape
    R, G = torch.empty_like(rewards), torch.empty_like(rewards)
    span = torch.empty((E, T), dtype=torch.int32, device=rewards.device)
    _C.call("apexb200_nstep", _C.ptr(rewards), _C.ptr(dones), int(E), int(T), int(n), float(gamma), _C.ptr(R),
            _C.ptr(G), _C.ptr(span), _C.stream_ptr())
    return R, G, span


def initial_priority(q_t, q_tpn, A, R, Gamma):
    """actor.py:138-140 on device: |R + Gamma max_a q_tpn - q_t[A]| -> f32 [count]."""
    count, nA = q_t.shape
    out = _f32(count, q_t)
    _C.call("apexb200_initial_priority", _C.ptr(q_t), _C.ptr(q_tpn), _C.ptr(A), _C.ptr(R), _C.ptr(Gamma), int(count),
            int(nA), _C.ptr(out), _C.stream_ptr())
    return out


def shard_targets(totals, rank, u):
    """K8: (local_targets u64-as-int64 [B], local_pos int32 [B], local_count int32 [1])."""
    import torch
    B = u.numel()
    lt = torch.empty(B, dtype=torch.int64, device=u.device)
    lp = torch.empty(B, dtype=torch.int32, device=u.device)
    lc = torch.empty(1, dtype=torch.int32, device=u.device)
    _C.call("apexb200_shard_targets", _C.ptr(totals), int(totals.numel()), int(rank), _C.ptr(u), int(B), _C.ptr(lt),
            _C.ptr(lp), _C.ptr(lc), _C.stream_ptr())
    return lt, lp, lc
