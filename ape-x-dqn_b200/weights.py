"""Learner -> actor weight publication and the batched actor policy.

The reference publishes its weights by assigning a whole state_dict to a Manager dict on every
learner step (learner.py:24, :74: 13.3 MB pickled through a proxy) and every actor unpickles it
every Q_network_sync_freq steps (actor.py:106, :189-191).  Here the weights live on a *board*:

    named shared host memory
      int64 ctrl[32]   [0] completed version  [1] begun version  [2] flat_bytes  [3] tensors
      slot 0, slot 1   two flat copies of the state_dict (every tensor at a 256-byte offset)

WeightPublisher (learner, needs CUDA) packs the model's tensors into one device snapshot with
one kernel (K9, apexb200_params_publish) and lets the library copy the snapshot into slot
version & 1 on its own copy stream; a host callback publishes the version when the bytes have
landed.  WeightSubscriber (a CPU actor, no CUDA) and DeviceWeightSubscriber (an actor process
with a GPU model) poll the version and read the slot — a memcpy instead of a pickle, and only
when the version moved.  BatchedActorPolicy evaluates E actors' epsilon-greedy policies in one
forward + one kernel (K10, apexb200_actor_policy).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _C

_CTRL_WORDS = 32
_HEADER = _CTRL_WORDS * 8


def param_layout(state_dict):
    """[(name, shape, dtype, flat_offset, nbytes)], flat_bytes — the same for every process that
    builds the same model (state_dict order is deterministic)."""
    spans, off = [], 0
    for name, t in state_dict.items():
        nbytes = t.numel() * t.element_size()
        spans.append((name, tuple(t.shape), t.dtype, off, nbytes))
        off += (nbytes + 255) // 256 * 256
    return spans, max(off, 256)


def _board_bytes(flat_bytes):
    slot = (flat_bytes + 4095) // 4096 * 4096
    return slot, _HEADER + 2 * slot


class _Board:
    """The shared-memory side, common to publisher and subscribers."""

    def __init__(self, name, flat_bytes, n_tensors, create):
        self.flat_bytes = int(flat_bytes)
        self.slot_bytes, total = _board_bytes(self.flat_bytes)
        self.shm = _C.open_shared_memory(name, create, total)
        self.name = self.shm.name
        self.ctrl = np.ndarray((_CTRL_WORDS,), np.int64, self.shm.buf, 0)
        if create:
            self.ctrl[:] = 0
            self.ctrl[2], self.ctrl[3] = self.flat_bytes, n_tensors
        elif int(self.ctrl[2]) != self.flat_bytes or int(self.ctrl[3]) != n_tensors:
            raise ValueError("weight board %r holds %d bytes / %d tensors, this model has %d / %d"
                             % (name, int(self.ctrl[2]), int(self.ctrl[3]), self.flat_bytes, n_tensors))
        self.slots = [np.ndarray((self.flat_bytes,), np.uint8, self.shm.buf, _HEADER + k * self.slot_bytes) for k in (0, 1)]

    @property
    def version(self):
        return int(self.ctrl[0])

    def read_consistent(self, copy_out):
        """Seqlock read: copy_out(slot_bytes_view) runs on the slot of the completed version and is
        repeated if a newer publication began overwriting that slot meanwhile.  -> version read."""
        while True:
            v = int(self.ctrl[0])
            if v == 0:
                return 0
            copy_out(self.slots[v & 1])
            if int(self.ctrl[1]) <= v + 1:
                return v

    def close(self, unlink=False):
        self.ctrl = self.slots = None
        self.shm.close()
        if unlink:
            _C.unlink_shared_memory(self.shm)


def _spans_of(tensors, layout):
    arr = (_C.ParamSpan * len(layout))()
    for k, ((_n, _s, _d, off, nbytes), t) in enumerate(zip(layout, tensors)):
        if not t.is_cuda or not t.is_contiguous():
            raise ValueError("weight board tensors must be contiguous CUDA tensors")
        arr[k].tensor, arr[k].flat_offset, arr[k].bytes = t.data_ptr(), off, nbytes
    return arr


class _CudaBytes:
    """Borrowed device memory, exposed to torch through the CUDA array interface."""

    def __init__(self, address, nbytes, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (address, False), "version": 2}
        self._owner = owner


class _DeviceBoard:
    """apexb200_params handle over the tensors of one CUDA module."""

    def __init__(self, module):
        sd = module.state_dict()
        self.layout, self.flat_bytes = param_layout(sd)
        self.tensors = list(sd.values())                        # keeps the storages alive
        dev = self.tensors[0].device
        if dev.type != "cuda":
            raise RuntimeError("the module must be on a CUDA device (no CPU fallback)")
        self.device = dev
        self._h = C.c_void_p()
        spans = _spans_of(self.tensors, self.layout)
        _C.call("apexb200_params_create", dev.index if dev.index is not None else torch.cuda.current_device(),
                C.cast(spans, C.c_void_p), len(self.layout), self.flat_bytes, C.byref(self._h))

    def flat(self):
        """The device snapshot as a uint8 tensor view (valid while this object lives)."""
        p, n = C.c_void_p(), C.c_uint64()
        _C.call("apexb200_params_flat", self._h, C.byref(p), C.byref(n))
        return torch.as_tensor(_CudaBytes(p.value, int(n.value), self), device=self.device)

    def destroy(self):
        if self._h:
            _C.call("apexb200_params_destroy", self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class WeightPublisher(_DeviceBoard):
    """Learner side (learner.py:24, :74)."""

    def __init__(self, module, name=None, create=True):
        super().__init__(module)
        self.board = _Board(name, self.flat_bytes, len(self.layout), create)
        self.name = self.board.name
        base = self.board.slots[0].ctypes.data - _HEADER
        _C.host_register(base, _HEADER + 2 * self.board.slot_bytes)
        self._base = base
        self._issued = int(self.board.ctrl[1])

    def publish(self, stream=None):
        """Snapshot the module's tensors as they are on `stream` and publish them; returns the version
        the snapshot will carry.  Does not wait for the copy."""
        self._issued += 1
        v = self._issued
        slot = self.board.slots[v & 1]
        _C.call("apexb200_params_publish", self._h, C.c_void_p(slot.ctypes.data), C.c_void_p(self.board.ctrl.ctypes.data),
                v, _C.stream_ptr(stream))
        return v

    def snapshot(self, stream=None):
        """Pack only (device subscribers on the same node read flat() directly)."""
        _C.call("apexb200_params_publish", self._h, None, None, 0, _C.stream_ptr(stream))

    @property
    def version(self):
        return self.board.version

    def wait(self, version, timeout=10.0):
        import time
        t0 = time.time()
        while self.board.version < version:
            if time.time() - t0 > timeout:
                raise TimeoutError("weight publication %d did not complete" % version)
            time.sleep(0.0002)

    def close(self, unlink=True):
        self.destroy()                                          # drains the copy stream
        if self._base is not None:
            _C.host_unregister(self._base)
            self._base = None
        self.board.close(unlink=unlink)


class WeightSubscriber:
    """CPU actor side (actor.py:106, :189-191): no CUDA in this process."""

    def __init__(self, module, name):
        self.module = module
        self.layout, self.flat_bytes = param_layout(module.state_dict())
        self.board = _Board(name, self.flat_bytes, len(self.layout), create=False)
        self.seen = 0

    def _copy_into_module(self, slot):
        sd = self.module.state_dict()
        with torch.no_grad():
            for (name, shape, dtype, off, nbytes) in self.layout:
                sd[name].copy_(torch.from_numpy(slot[off:off + nbytes]).view(dtype).view(shape))

    def poll(self):
        """Load the newest published weights into the module if they changed; -> True when loaded.
        The slot is copied straight into the parameters (one pass over the bytes) and the copy is
        repeated if the publisher began overwriting that slot meanwhile."""
        if self.board.version <= self.seen:
            return False
        self.seen = self.board.read_consistent(self._copy_into_module)
        return True

    def close(self):
        self.board.close()


class DeviceWeightSubscriber(_DeviceBoard):
    """Actor process whose model lives on a GPU (batched inference)."""

    def __init__(self, module, name):
        super().__init__(module)
        self.board = _Board(name, self.flat_bytes, len(self.layout), create=False)
        self.seen = 0
        base = self.board.slots[0].ctypes.data - _HEADER          # page-lock: H2D straight out of the board
        _C.host_register(base, _HEADER + 2 * self.board.slot_bytes)
        self._base = base

    def poll(self, stream=None):
        """Newest published weights -> the module's tensors (H2D from the slot + K9 scatter on `stream`).
        Waits for the copy so that the seqlock can be checked; repeated if the slot was being rewritten."""
        if self.board.version <= self.seen:
            return False
        st = stream if stream is not None else torch.cuda.current_stream()

        def load(slot):
            _C.call("apexb200_params_load", self._h, C.c_void_p(slot.ctypes.data), 1, _C.stream_ptr(st))
            st.synchronize()
        self.seen = self.board.read_consistent(load)
        return True

    def load_from(self, flat_dev, stream=None):
        """Scatter a device (or NVLink-peer) snapshot, e.g. WeightPublisher.flat() of the same node."""
        _C.call("apexb200_params_load", self._h, C.c_void_p(flat_dev.data_ptr()), 0, _C.stream_ptr(stream))

    def close(self):
        self.destroy()
        if self._base is not None:
            _C.host_unregister(self._base)
            self._base = None
        self.board.close()


def actor_epsilons(actor_params, actor_ids):
    """actor.py:111-114: eps_i = eps^(1 + alpha * i / (N - 1))."""
    eps, N, alpha = actor_params["epsilon"], actor_params["num_actors"], actor_params["alpha"]
    return np.array([eps ** (1 + alpha * i / (N - 1)) if N > 1 else eps for i in actor_ids], np.float32)


def actor_policy(advantage_or_q, eps, value=None, u_explore=None, u_action=None, seed=0, step=0, want_q=True,
                 want_explored=False):
    """K10.  -> (actions int32 [E], q float32 [E, A] or None, explored uint8 [E] or None)."""
    x = advantage_or_q.contiguous().float()
    E, A = x.shape
    dev = x.device
    if dev.type != "cuda":
        raise RuntimeError("actor_policy needs CUDA tensors (no CPU fallback)")
    actions = torch.empty(E, dtype=torch.int32, device=dev)
    q = torch.empty((E, A), dtype=torch.float32, device=dev) if want_q else None
    ex = torch.empty(E, dtype=torch.uint8, device=dev) if want_explored else None
    v = value.contiguous().float().reshape(E) if value is not None else None
    ue = u_explore.contiguous().double() if u_explore is not None else None
    ua = u_action.contiguous().double() if u_action is not None else None
    _C.call("apexb200_actor_policy", _C.ptr(v), _C.ptr(x), _C.ptr(eps.contiguous().float()), _C.ptr(ue), _C.ptr(ua),
            int(seed) & (2 ** 64 - 1), int(step), E, A, _C.ptr(q), _C.ptr(actions), _C.ptr(ex), _C.stream_ptr())
    return actions, q, ex


class BatchedActorPolicy:
    """E reference actors' policies in one call: Q forward for [E, C, H, W] observations, duelling
    combine per row (what each batch-1 actor computes, actor.py:160-161), per-actor epsilon
    (actor.py:114), epsilon-greedy (actor.py:121-125)."""

    def __init__(self, q_network, actor_params, actor_ids, seed=0):
        self.Q = q_network
        dev = next(q_network.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("BatchedActorPolicy needs the Q network on a CUDA device")
        self.device = dev
        self.eps = torch.as_tensor(actor_epsilons(actor_params, list(actor_ids)), device=dev)
        self.seed, self.steps = int(seed), 0

    @torch.no_grad()
    def act(self, obs, u_explore=None, u_action=None):
        """obs uint8/float [E, C, H, W] (device or host) -> (actions int32 [E], q float32 [E, A])."""
        x = torch.as_tensor(obs).to(self.device).float()
        value, adv, _q_coupled = self.Q(x)
        actions, q, _ = actor_policy(adv, self.eps, value=value, u_explore=u_explore, u_action=u_action,
                                     seed=self.seed, step=self.steps)
        self.steps += 1
        return actions, q


__all__ = ["param_layout", "WeightPublisher", "WeightSubscriber", "DeviceWeightSubscriber", "BatchedActorPolicy",
           "actor_policy", "actor_epsilons"]
