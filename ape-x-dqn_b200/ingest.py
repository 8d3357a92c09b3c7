"""Device-side ingest for vectorised actors: frame streams in, n-step records out.

Replaces, for E environments, the chain ExperienceBuffer.add / update_buffer /
construct_nstep_transition (actor.py:29-70) -> Actor.compute_priorities (actor.py:127-143) ->
queue -> ReplayMemory.add (main.py:21-25, replay.py:59-69) with the apexb200_ingest_* entry points:
nothing is pickled, the frames are stored once, and the n-step carry, the episode ages and the ring
heads of every environment live on the device — a push is three kernel launches (frame copy, record
build + insert, state roll) with no host arithmetic in between.

One environment owns one lane of a stream-layout replay.  The frame observed at env step tau is
stored at lane position tau together with the n-step record of step t = tau - n (its S_tpn ends at
that frame): A_t, R = sum_{k<m} gamma^k r_{t+k}, Gamma = gamma^m or 0 when a terminal lies inside
the window, initial priority |R + Gamma max q_tau - q_t[A_t]| computed by the insert kernel.
Everything a record needs lies in the past of its frame, so a chunk of T env steps emits T records
immediately; only the last n steps are carried to the next chunk.  The first n frames of a stream
carry no record (priority 0).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _C
from .replay import stream_offsets


class VectorActorIngest:
    def __init__(self, mem, num_envs, n_step=None, gamma=0.99, lanes=None, num_actions=None, max_chunk_steps=0):
        mem._ensure()
        self.lanes = list(range(num_envs)) if lanes is None else [int(x) for x in lanes]
        if mem.layout != "stream" or len(self.lanes) != num_envs or max(self.lanes) >= mem.lanes:
            raise ValueError("VectorActorIngest needs a stream-layout replay and one lane per environment")
        self.mem, self.E = mem, int(num_envs)
        self.n = int(n_step if n_step is not None else mem.n_step)
        if mem.lookback < mem.C - 1 + self.n:
            raise ValueError("replay lookback %d < frames_per_state - 1 + n_step" % mem.lookback)
        self.gamma = float(gamma)
        self.max_chunk_steps = int(max_chunk_steps)
        self._g = None
        self.nA = None
        if num_actions is not None:
            self._create(int(num_actions))

    def _create(self, num_actions):
        g = C.c_void_p()
        lanes = (C.c_int32 * self.E)(*self.lanes)
        with torch.cuda.device(self.mem.device):
            _C.call("apexb200_ingest_create", self.mem._h, self.E, lanes, self.n, self.gamma, int(num_actions),
                    self.max_chunk_steps, C.byref(g))
        self._g, self.nA = g, int(num_actions)

    def __del__(self):
        g, self._g = getattr(self, "_g", None), None
        if g is not None:
            try:
                _C.lib().apexb200_ingest_destroy(g)
            except Exception:
                pass

    close = __del__

    def _field(self, x, dtype, np_dtype):
        """Device tensor of the wanted dtype: CUDA tensors are used in place, host data goes through the replay's
        pinned staging (asynchronous copy)."""
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.to(dtype).contiguous()
        a = x.numpy() if isinstance(x, torch.Tensor) else np.asarray(x)
        return self.mem._dev(a, np_dtype)

    def push(self, frames, actions, rewards, dones, q, envs=None):
        """One chunk of T env steps for all E envs (or for the environments listed in `envs`).

        frames [E, T, F] uint8: observation s_tau seen at step tau; actions/rewards/dones [E, T]:
        what was done at s_tau and what came back (dones[e, tau] = episode ended by that step, so
        s_{tau+1} starts a new episode); q [E, T, A] = Q(s_tau).  Device tensors or host arrays.
        Returns the number of records (E * T) written.
        """
        q = self._field(q, torch.float32, np.float32)
        if self._g is None:
            self._create(int(q.shape[-1]))
        frames = self._field(frames, torch.uint8, np.uint8)
        actions = self._field(actions, torch.int32, np.int32)
        rewards = self._field(rewards, torch.float32, np.float32)
        dones = self._field(dones, torch.uint8, np.uint8)
        E, T = int(actions.shape[0]), int(actions.shape[1])
        if envs is None:
            if E != self.E:
                raise ValueError("push: inputs for %d environments, ingest has %d" % (E, self.E))
            ids = None
        else:
            ids = (C.c_int32 * E)(*[int(e) for e in envs])
        _C.call("apexb200_ingest_push", self._g, ids, E, T, _C.ptr(frames), _C.ptr(actions), _C.ptr(rewards), _C.ptr(dones),
                _C.ptr(q), _C.stream_ptr())
        return E * T

    def sync_heads(self):
        """After replaying a CUDA graph that contains push(): bring the host-side ring heads (size(), checkpoints,
        plain insert calls) up to date with the device-side ones."""
        if self._g is not None:
            _C.call("apexb200_ingest_sync_heads", self._g, _C.stream_ptr())


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
        """Page-lock the shared memory for this process and map it for kernels: copies out of it are asynchronous
        and the ingest's pack kernel can read the chunks in place (self.dev_base = device-visible address)."""
        if not self._registered:
            self._reg_base = self._data.ctypes.data - 256
            self.dev_base = _C.host_register_mapped(self._reg_base, 256 + self.slots * self.slot_bytes)
            self._registered = True
        return self._registered

    def blob_address(self, k):
        """Device-visible address of chunk slot k (register_cuda() first)."""
        return self.dev_base + 256 + (k % self.slots) * self.slot_bytes

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
            _C.host_unregister(self._reg_base)
            self._registered = False
        self.ctrl = self._data = None
        self.shm.close()
        if unlink:
            _C.unlink_shared_memory(self.shm)


class RingIngest:
    """Drains a set of ActorStagingRings (one per actor = one lane each) into a stream-layout replay.

    Every poll() collects ALL rings that hold an unread chunk and moves them with one apexb200_ingest_push_blobs
    call: a pack kernel reads the chunks in place from the page-locked, device-mapped shared memory (no per-ring
    copy calls), then the three ingest launches build and insert the records of all those actors.  A ring slot is
    handed back to its actor when a CUDA event recorded behind that work has completed — the host never waits
    for the stream."""

    def __init__(self, mem, rings, n_step=None, gamma=0.99):
        self.mem, self.rings = mem, list(rings)
        r0 = self.rings[0]
        if any((r.T, r.F, r.nA) != (r0.T, r0.F, r0.nA) for r in self.rings):
            raise ValueError("RingIngest: all rings must carry chunks of the same shape")
        self.ingest = VectorActorIngest(mem, len(self.rings), n_step=n_step, gamma=gamma, num_actions=r0.nA,
                                        max_chunk_steps=r0.T)
        for r in self.rings:
            r.register_cuda()
        self._taken = [int(r.ctrl[1]) for r in self.rings]                # chunks handed to the device per ring
        self._inflight = []                                               # (event, [ring indices]) in stream order

    def _reap(self, wait=False):
        while self._inflight:
            ev, ks = self._inflight[0]
            if wait:
                ev.synchronize()
            elif not ev.query():
                break
            for k in ks:
                self.rings[k].release()                                   # the actor may overwrite the slot now
            self._inflight.pop(0)

    def poll(self):
        """Move every chunk that is ready; returns the number of records inserted (enqueued)."""
        self._reap()
        n, r0 = 0, self.rings[0]
        while True:
            ready = [k for k, r in enumerate(self.rings) if int(r.ctrl[0]) > self._taken[k]]
            if not ready:
                break
            envs = (C.c_int32 * len(ready))(*ready)
            blobs = (C.c_uint64 * len(ready))(*[self.rings[k].blob_address(self._taken[k]) for k in ready])
            _C.call("apexb200_ingest_push_blobs", self.ingest._g, envs, blobs, len(ready), r0.T, r0.off_act, r0.off_rew,
                    r0.off_done, r0.off_q, _C.stream_ptr())
            ev = torch.cuda.Event()
            ev.record()
            self._inflight.append((ev, ready))
            for k in ready:
                self._taken[k] += 1
            n += len(ready) * r0.T
        return n

    def flush(self):
        """Wait for everything handed to the device and release the slots."""
        self._reap(wait=True)


__all__ = ["VectorActorIngest", "ActorStagingRing", "RingIngest", "stream_offsets"]
