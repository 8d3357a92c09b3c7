"""CPU oracle: actor-side n-step assembly and initial priority (TEST INFRASTRUCTURE).

Follows /root/reference/actor.py:
  * ExperienceBuffer.add / update_buffer / construct_nstep_transition
    (actor.py:29-70)  -> LiteralExperienceBuffer  (literal semantics incl. the
    double counting Q8, gamma^(B-1) Q9 and chunking Q10 of SURVEY §2.1; PINNED
    against the real class by tests/golden/nstep_literal.json)
  * intended n-step return  R = sum_{k<m} gamma^k r_{t+k}, Gamma = gamma^m,
    Gamma = 0 on terminal (SURVEY Q8-Q11 "intended")      -> nstep_intended
  * Actor.compute_priorities (actor.py:127-143, intended per-row |TD|, Q12)
                                                          -> initial_priority
"""
from __future__ import annotations

from collections import namedtuple

import numpy as np

Step = namedtuple("Step", ["S", "A", "R", "Gamma", "q"])
NStep = namedtuple("NStep", ["S_t", "A_t", "R_ttpB", "Gamma_ttpB", "qS_t", "S_tpn", "qS_tpn", "key"])


class LiteralExperienceBuffer:
    """Literal behaviour of actor.py:15-93, restated (not imported)."""

    def __init__(self, n, actor_id, gamma=0.99):
        self.n, self.actor_id, self.gamma = int(n), actor_id, gamma
        self.pending = []        # 1-step records of the open chunk
        self.ready = []          # emitted n-step records
        self.seq = 0

    def _reaccumulate(self):
        # actor.py:35-43 — every earlier record adds the discounted *current* R of
        # all later records on top of what it already holds (=> Q8 double counting);
        # the last record keeps the Gamma it was given.
        cnt = len(self.pending)
        for i in range(cnt - 1):
            acc, disc = self.pending[i].R, 1
            for k in range(i + 1, cnt):
                disc *= self.gamma
                acc += disc * self.pending[k].R
            self.pending[i] = self.pending[i]._replace(R=acc, Gamma=disc)

    def flush(self, nxt):
        # actor.py:44-55 — emit head-of-chunk with the next step's S/q; drop the rest.
        if not self.pending:
            return
        head = self.pending[0]
        self.ready.append(NStep(head.S, head.A, head.R, head.Gamma, head.q, nxt.S, nxt.q,
                                str(self.actor_id) + str(self.seq)))
        self.seq += 1
        self.pending = []

    def add(self, step):
        # actor.py:57-70
        if len(self.pending) < self.n:
            self.pending.append(step)
            self._reaccumulate()
        else:
            self.flush(step)

    def get(self, k):
        assert k <= len(self.ready)
        out, self.ready = self.ready[:k], self.ready[k:]
        return out


def nstep_intended(rewards, dones, n, gamma, dtype=np.float32):
    """Sliding-window n-step return for E environments x T steps.

    Returns (R, Gamma, span): span[e,t] = number of steps accumulated (S_tpn is
    the observation `span` steps after t); span == 0 marks a window cut by the
    end of the array before n steps or a terminal (not yet a valid transition).
    gamma^k is built by repeated fp32 multiplication, as the kernel does.
    """
    f = dtype
    rewards = np.asarray(rewards, f)
    dones = np.asarray(dones).astype(bool)
    E, T = rewards.shape
    R = np.zeros((E, T), f)
    G = np.zeros((E, T), f)
    span = np.zeros((E, T), np.int32)
    g = f(gamma)
    for e in range(E):
        for t in range(T):
            acc, disc, m, term = f(0), f(1), 0, False
            for k in range(n):
                if t + k >= T:
                    break
                acc = f(acc + f(disc * rewards[e, t + k]))
                disc = f(disc * g)
                m += 1
                if dones[e, t + k]:
                    term = True
                    break
            if term or m == n:
                R[e, t], G[e, t], span[e, t] = acc, (f(0) if term else disc), m
    return R, G, span


def initial_priority(q_t, q_tpn, A, R, G):
    """|R + Gamma * max_a qS_tpn - qS_t[A]|  (actor.py:138-140), fp64."""
    q_t, q_tpn = np.asarray(q_t, np.float64), np.asarray(q_tpn, np.float64)
    A = np.asarray(A, np.int64)
    tgt = np.asarray(R, np.float64) + np.asarray(G, np.float64) * q_tpn.max(axis=1)
    return np.abs(tgt - q_t[np.arange(A.shape[0]), A])
