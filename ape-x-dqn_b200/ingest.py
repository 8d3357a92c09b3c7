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
    def __init__(self, mem, num_envs, n_step=None, gamma=0.99, lanes=None):
        mem._ensure()
        self.lanes = list(range(num_envs)) if lanes is None else [int(x) for x in lanes]
        if mem.layout != "stream" or len(self.lanes) != num_envs or max(self.lanes) >= mem.lanes:
            raise ValueError("VectorActorIngest needs a stream-layout replay and one lane per environment")
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
        mem.insert_lanes(self.lanes, frames.reshape(E, T, -1), rel, A_rec, R_rec, G_rec, prio=prio.contiguous())   # K6, all lanes
        self.c_act, self.c_rew, self.c_done, self.c_q = (act[:, -n:].contiguous(), rew[:, -n:].contiguous(),
                                                         don[:, -n:].contiguous(), qq[:, -n:].contiguous())
        self.seen += T
        return E * T


class ActorStagingRing:
    """Single-producer / single-consumer ring of step chunks in named shared host memory: the
    actor -> replay transport (replaces the Manager queue + pickle of main.py:21-25 /
    actor.py:181-187, 566 KB of pickle per 5 records).  An actor process writes chunks of T env
    steps (uint8 frames, actions, rewards, dones, Q rows) with plain memcpys; the process that
    owns the replay maps the same memory, registers it with CUDA (page-locked) and copies each
    chunk to the device asynchronously before handing it to VectorActorIngest.

    Layout: int64 control[8] = (produced, consumed, closed, ...), then `slots` chunks of
    frames [T, F] u8 | actions [T] i32 | rewards [T] f32 | dones [T] u8 (padded) | q [T, nA] f32.
    """

    def __init__(self, name, slots, chunk_steps, frame_bytes, num_actions, create):
        from . import _C
        self.slots, self.T, self.F, self.nA = int(slots), int(chunk_steps), int(frame_bytes), int(num_actions)
        T, F, nA = self.T, self.F, self.nA
        self.off_act = T * F
        self.off_rew = self.off_act + 4 * T
        self.off_done = self.off_rew + 4 * T
        self.off_q = self.off_done + ((T + 15) // 16) * 16
        self.slot_bytes = ((self.off_q + 4 * T * nA + 255) // 256) * 256
        total = 256 + self.slots * self.slot_bytes
        self.shm = _C.open_shared_memory(name, create, total)
        self.ctrl = np.ndarray((8,), np.int64, self.shm.buf, 0)
        if create:
            self.ctrl[:] = 0
        self._data = np.ndarray((self.slots, self.slot_bytes), np.uint8, self.shm.buf, 256)
        self._registered = False

    # views of slot k
    def _views(self, k):
        raw, T, F, nA = self._data[k], self.T, self.F, self.nA
        return (raw[:T * F].reshape(T, F), raw[self.off_act:self.off_act + 4 * T].view(np.int32),
                raw[self.off_rew:self.off_rew + 4 * T].view(np.float32), raw[self.off_done:self.off_done + T],
                raw[self.off_q:self.off_q + 4 * T * nA].view(np.float32).reshape(T, nA))

    # ---- producer side (actor process, no CUDA)
    def push(self, frames, actions, rewards, dones, q):
        """Write one chunk of T steps; returns False (nothing written) when the ring is full."""
        produced, consumed = int(self.ctrl[0]), int(self.ctrl[1])
        if produced - consumed >= self.slots:
            return False
        f, a, r, d, qq = self._views(produced % self.slots)
        f[...] = np.asarray(frames, np.uint8).reshape(self.T, self.F)
        a[...] = actions
        r[...] = rewards
        d[...] = np.asarray(dones).astype(np.uint8)
        qq[...] = q
        self.ctrl[0] = produced + 1                       # publish after the payload (x86 store order; SPSC)
        return True

    # ---- consumer side (process that owns the replay)
    def register_cuda(self):
        """Page-lock the shared memory for this process so that H2D copies out of it are asynchronous."""
        if not self._registered:
            from . import _C
            self._reg_base = self._data.ctypes.data - 256
            _C.host_register(self._reg_base, 256 + self.slots * self.slot_bytes)
            self._registered = True
        return self._registered

    def pop(self):
        """-> (frames, actions, rewards, dones, q) host views of the oldest unread chunk, or None.
        Call release() after the copies out of the views have been issued and completed."""
        produced, consumed = int(self.ctrl[0]), int(self.ctrl[1])
        if produced == consumed:
            return None
        return self._views(consumed % self.slots)

    def release(self):
        self.ctrl[1] = int(self.ctrl[1]) + 1

    def close(self, unlink=False):
        if self._registered:                              # before the mapping goes away: the range may be reused
            from . import _C
            _C.host_unregister(self._reg_base)
            self._registered = False
        self.ctrl = self._data = None
        self.shm.close()
        if unlink:
            from . import _C
            _C.unlink_shared_memory(self.shm)


class RingIngest:
    """Drains a set of ActorStagingRings (one per actor = one lane each) into a stream-layout replay."""

    def __init__(self, mem, rings, n_step=None, gamma=0.99):
        self.mem, self.rings = mem, list(rings)
        self.lane_ingest = [VectorActorIngest(mem, 1, n_step=n_step, gamma=gamma, lanes=[k]) for k in range(len(self.rings))]
        for r in self.rings:
            r.register_cuda()

    def poll(self):
        """Move every chunk that is ready; returns the number of records inserted."""
        n = 0
        dev = self.mem.device
        for ring, ing in zip(self.rings, self.lane_ingest):
            while True:
                chunk = ring.pop()
                if chunk is None:
                    break
                f, a, r, d, q = chunk
                tensors = [torch.from_numpy(x).to(dev, non_blocking=True) for x in (f, a, r, q)]
                dones = d.copy()
                torch.cuda.current_stream().synchronize()         # payload is on the device: the slot can be reused
                ring.release()
                n += ing.push(tensors[0][None], tensors[1][None], tensors[2][None], dones[None], tensors[3][None])
        return n


__all__ = ["VectorActorIngest", "ActorStagingRing", "RingIngest", "stream_offsets"]
