"""CPU oracle: learner loss / TD / priority arithmetic (TEST INFRASTRUCTURE).

Restates /root/reference/learner.py:43-52.  Pinned by tests/golden/td_reference.npz:
learner.py:44 raises TypeError under this container's torch 2.11 / numpy 2.3 as
written, so tests/golden/make_golden.py executes the unmodified
Learner.compute_loss_and_priorities under an operator shim (ndarray operands of
Tensor.__rmul__/__radd__ coerced with torch.from_numpy, as torch 0.4 did) and
records its outputs; this file is also checked against an independent torch fp64
evaluation (tests/test_oracle_cpu.py).

    a*      = argmax_a Q_online(S_tpn)[b, a]                learner.py:45
    G_b     = R_b + Gamma_b * Q_target(S_tpn)[b, a*]        learner.py:44-45
    delta_b = G_b - Q_online(S_t)[b, A_b]                   learner.py:46-47
    loss_b  = 1/2 delta_b^2                                 learner.py:48
    prio_b  = |delta_b|                                     learner.py:50 (intended, Q12)
    loss    = mean_b loss_b                                 learner.py:52
"""
from __future__ import annotations

import numpy as np

LOSS_HALF_SQ = 0   # reference (learner.py:48)
LOSS_HUBER = 1     # north_star option, kappa = huber_kappa


def td_forward(q_on_t, q_on_tpn, q_tg_tpn, A, R, G, alpha=0.6, loss_kind=LOSS_HALF_SQ,
               kappa=1.0, eps=0.0, dtype=np.float32, weights=None):
    """Returns dict(delta, abs_td, prio (|delta|^alpha), loss, loss_mean)."""
    f = dtype
    q_on_t, q_on_tpn, q_tg_tpn = (np.asarray(x, f) for x in (q_on_t, q_on_tpn, q_tg_tpn))
    A = np.asarray(A, np.int64)
    R, G = np.asarray(R, f), np.asarray(G, f)
    rows = np.arange(A.shape[0])
    a_star = np.argmax(q_on_tpn, axis=1)                    # first maximum on ties
    target = R + G * q_tg_tpn[rows, a_star]
    delta = (target - q_on_t[rows, A]).astype(f)
    ad = np.abs(delta)
    if loss_kind == LOSS_HALF_SQ:
        loss = f(0.5) * delta * delta
    else:
        k = f(kappa)
        loss = np.where(ad <= k, f(0.5) * delta * delta, k * (ad - f(0.5) * k))
    if weights is not None:
        loss = loss * np.asarray(weights, f)               # importance-sampling correction (not in the reference)
    base = np.maximum(ad, f(eps)) if eps > 0 else ad
    prio = np.power(base, f(alpha)).astype(f)
    return dict(delta=delta, abs_td=ad, prio=prio, loss=loss.astype(f),
                loss_mean=f(np.sum(loss.astype(np.float64)) / A.shape[0]), a_star=a_star)


def is_weights(leaf_u64, total, n_records, beta):
    """PER importance weights w = (N * P)^(-beta) / max w, P = leaf / total (parameters.json:30, unused upstream)."""
    p = np.asarray(leaf_u64, np.float64) / float(total)
    w = np.where(p > 0, (n_records * p) ** (-float(beta)), 0.0)
    return (w / w.max()).astype(np.float32)


def td_backward(delta, A, num_actions, grad_out=1.0, loss_kind=LOSS_HALF_SQ, kappa=1.0,
                dtype=np.float32, weights=None):
    """d loss_mean / d Q_online(S_t): -dloss/ddelta / B at [b, A_b], 0 elsewhere."""
    f = dtype
    delta = np.asarray(delta, f)
    B = delta.shape[0]
    g = delta if loss_kind == LOSS_HALF_SQ else np.clip(delta, -f(kappa), f(kappa))
    out = np.zeros((B, num_actions), f)
    col = -g * f(grad_out) / f(B)
    if weights is not None:
        col = col * np.asarray(weights, f)
    out[np.arange(B), np.asarray(A, np.int64)] = col
    return out
