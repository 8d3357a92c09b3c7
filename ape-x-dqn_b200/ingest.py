"""Device-side ingest for vectorised actors: frame streams in, n-step records out.

Replaces, for E environments stepped in lock-step, the chain ExperienceBuffer.add /
update_buffer / construct_nstep_transition (actor.py:29-70) -> Actor.compute_priorities
(actor.py:127-143) -> queue -> ReplayMemory.add (main.py:21-25, replay.py:59-69) with
K7 (apexb200_nstep) + K6 (apexb200_replay_insert); nothing is pickled and the frames are
stored once.

One environment owns one lane of a stream-layout replay.  The frame observed at env step
tau is stored at lane position tau together with the n-step record of step t = tau - n
(its S_tpn ends at that frame): A_t, R = sum_{k<m} gamma^k r_{t+k}, Gamma = gamma^m or 0
when a terminal lies inside the window, initial priority |R + Gamma max q_tau - q_t[A_t]|
computed by the insert kernel.  Everything a record needs lies in the past of its frame,
so a chunk of T env steps emits T records immediately; only the last n steps are carried
to the next chunk.  The first n frames of a stream carry no record (priority 0).
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .replay import stream_offsets


class VectorActorIngest:
    def __init__(self, mem, num_envs, n_step=None, gamma=0.99):
        mem._ensure()
        if mem.layout != "stream" or mem.lanes != num_envs:
            raise ValueError("VectorActorIngest needs a stream-layout replay with one lane per environment")
        self.mem, self.E = mem, int(num_envs)
        self.n = int(n_step if n_step is not None else mem.n_step)
        if mem.lookback < mem.C - 1 + self.n:
            raise ValueError("replay lookback %d < frames_per_state - 1 + n_step" % mem.lookback)
        self.gamma = float(gamma)
        E, n, dev = self.E, self.n, mem.device
        # carry of the last n steps per env (device) and episode ages (host)
        self.c_act = torch.zeros((E, n), dtype=torch.int32, device=dev)
        self.c_rew = torch.zeros((E, n), dtype=torch.float32, device=dev)
        self.c_done = torch.ones((E, n), dtype=torch.uint8, device=dev)      # "terminal": no record before the stream starts
        self.c_q = None                                                      # sized by the first chunk's q
        self.seen = 0                                                        # frames pushed per env so far
        self.age = np.zeros(E, np.int64)                                     # frames since episode start, at the next frame

    def push(self, frames, actions, rewards, dones, q):
        """One chunk of T env steps for all E envs.

        frames [E, T, F] uint8: observation s_tau seen at step tau; actions/rewards/dones [E, T]:
        what was done at s_tau and what came back (dones[e, tau] = episode ended by that step, so
        s_{tau+1} starts a new episode); q [E, T, A] = Q(s_tau).  Device tensors or host arrays.
        Returns the number of records (E * T) written.
        """
        mem, E, n, dev = self.mem, self.E, self.n, self.mem.device
        frames = torch.as_tensor(frames).to(dev)
        actions = torch.as_tensor(actions).to(dev, torch.int32)
        rewards = torch.as_tensor(rewards).to(dev, torch.float32)
        dones_h = np.asarray(dones.cpu() if isinstance(dones, torch.Tensor) else dones).astype(bool)
        dones_d = torch.as_tensor(dones_h.astype(np.uint8)).to(dev)
        q = torch.as_tensor(q).to(dev, torch.float32)
        T = frames.shape[1]
        # steps tau-n .. tau-1 for every new frame tau: carry + this chunk
        act = torch.cat([self.c_act, actions], 1)
        rew = torch.cat([self.c_rew, rewards], 1).contiguous()
        don = torch.cat([self.c_done, dones_d], 1).contiguous()
        if self.c_q is None:
            self.c_q = torch.zeros((E, n, q.shape[-1]), dtype=torch.float32, device=dev)
        qq = torch.cat([self.c_q, q], 1)
        R, G, _span = ops.nstep_returns(rew, don, n, self.gamma)             # K7 over [E, n + T]
        A_rec, R_rec, G_rec = act[:, :T].contiguous(), R[:, :T].contiguous(), G[:, :T].contiguous()
        q_t, q_tpn = qq[:, :T].contiguous(), qq[:, n:n + T].contiguous()
        # episode ages on the host (tiny): age of frame tau and of step t = tau - n
        age_tau = np.zeros((E, T), np.int64)
        a = self.age.copy()
        for t in range(T):
            age_tau[:, t] = a
            a = np.where(dones_h[:, t], 0, a + 1)
        self.age = a
        hist = getattr(self, "_age_hist", np.zeros((E, n), np.int64))
        age_all = np.concatenate([hist, age_tau], 1)                         # ages of frames tau-n .. tau
        self._age_hist = age_all[:, -n:]
        age_t = age_all[:, :T]                                               # age of frame t = tau - n
        prio_mask = torch.ones((E, T), dtype=torch.float32, device=dev)
        warm = max(0, n - self.seen)                                         # first n frames of the stream: no record yet
        if warm:
            prio_mask[:, :min(warm, T)] = 0.0
        C = mem.C
        back = np.arange(C - 1, -1, -1, dtype=np.int64)[None, None, :]
        s_tpn = -np.minimum(back, age_tau[:, :, None])
        s_t = -(n + np.minimum(back, age_t[:, :, None]))
        rel = torch.as_tensor(np.concatenate([s_t, s_tpn], 2).astype(np.int32)).to(dev)
        prio = ops.initial_priority(q_t.reshape(E * T, -1), q_tpn.reshape(E * T, -1), A_rec.reshape(-1), R_rec.reshape(-1),
                                    G_rec.reshape(-1)).view(E, T) * prio_mask
        for e in range(E):                                                   # K6, one lane per env
            mem.insert(e, frames[e].reshape(T, -1), rel[e], A_rec[e], R_rec[e], G_rec[e], prio=prio[e].contiguous())
        self.c_act, self.c_rew, self.c_done, self.c_q = (act[:, -n:].contiguous(), rew[:, -n:].contiguous(),
                                                         don[:, -n:].contiguous(), qq[:, -n:].contiguous())
        self.seen += T
        return E * T


__all__ = ["VectorActorIngest", "stream_offsets"]
