"""GPU-resident prioritized replay memory with the reference's ReplayMemory interface.

Mirror of /root/reference/replay.py:8-83.  Two tiers on one class:

* compat tier — exactly the reference's six public methods and record type
  (``add(priorities, xp_batch)``, ``sample(n) -> list[N_Step_Transition]``,
  ``set_priorities(dict)``, ``update_sample_probabilities()``,
  ``remove_to_fit()``, ``size()``); works through a ``BaseManager`` proxy the way
  main.py:18,42 uses it; host records in, host records out.
* fast tier — tensor in / tensor out on the CUDA stream of the learner
  (``insert``, ``sample_tensors``, ``update_priorities``, ``td_update``); the
  batch never leaves HBM.

Semantics are the *intended* ones fixed by SURVEY.md §2.1 (proportional
sampling, per-record priorities, ring eviction); see DESIGN.md.  Every device
operation goes through libapexb200.so — there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from collections import namedtuple

import numpy as np

from . import _C

N_Step_Transition = namedtuple(
    "N_Step_Transition", ["S_t", "A_t", "R_ttpB", "Gamma_ttpB", "qS_t", "S_tpn", "qS_tpn", "key"])   # replay.py:5

Batch = namedtuple("Batch", ["idx", "S_t", "S_tpn", "A", "R", "Gamma", "fsel"])


def stream_offsets(frames_per_state, span, age):
    """Relative frame offsets for the frame-once ("stream") layout.

    One record is stored per environment step, owning the newest frame x_{t+m}
    (m = span, the number of steps its n-step return covers).  S_tpn is the C
    frames ending at offset 0, S_t the C frames ending at offset -m; offsets
    older than the episode start (``age`` = frames since the episode began,
    counted at the record's own frame) repeat the episode's first frame.
    Returns int32 [len(span), 2C].
    """
    span = np.asarray(span, np.int64).reshape(-1, 1)
    age = np.asarray(age, np.int64).reshape(-1, 1)
    back = np.arange(frames_per_state - 1, -1, -1, dtype=np.int64)[None, :]
    s_tpn = -np.minimum(back, age)
    s_t = -np.minimum(back + span, age)
    return np.concatenate([s_t, s_tpn], axis=1).astype(np.int32)


class KeyIndex:
    """Host bookkeeping for the compat tier: reference keys (actor.py:47) <-> ring slots.

    Reproduces dict.update semantics of replay.py:40 for keys that are not in
    memory yet (the priority is kept and applied when the record arrives) and
    drops the key of a record when the ring overwrites its slot.
    """

    def __init__(self):
        self.slot_of = {}
        self.key_of = {}
        self.pending = {}

    def assign(self, key, slot):
        old = self.key_of.get(slot)
        if old is not None and self.slot_of.get(old) == slot:
            del self.slot_of[old]
        prev = self.slot_of.get(key)
        if prev is not None and prev != slot:
            self.key_of.pop(prev, None)    # key collision: newest record owns the key
        self.key_of[slot] = key
        self.slot_of[key] = slot

    def split(self, priorities):
        """dict -> (slots, values) for known keys; unknown keys are remembered."""
        slots, vals = [], []
        for k, v in priorities.items():
            s = self.slot_of.get(k)
            if s is None:
                self.pending[k] = float(v)
            else:
                slots.append(s)
                vals.append(float(v))
        return slots, vals


class ReplayMemory(object):
    def __init__(self, soft_capacity, params):
        # replay.py:9-16.  No CUDA call here: main.py:40-42 constructs this object inside a
        # manager server process and main.py:45-47 forks after constructing the learner.
        self.soft_capacity = int(soft_capacity)
        self.params = dict(params)
        self.alpha = float(params["priority_exponent"])
        self.layout = params.get("layout", "materialised")
        self.lanes = int(params.get("lanes", 1))
        self.n_step = int(params.get("n_step", 3))
        self.num_actions = int(params.get("num_actions", 0))
        self.device_index = params.get("device", None)
        self.state_shape = tuple(params["state_shape"]) if "state_shape" in params else None
        self.eps = float(params.get("priority_eps", 0.0))
        self.seed = int(params.get("seed", 0x5EED))
        self._h = None
        self._keys = KeyIndex()
        self._last_q = None
        self._out_cache = {}
        self._pinned = {}
        self._ws = None

    # ------------------------------------------------------------------ set-up
    def _ensure(self, state_shape=None):
        if self._h is not None:
            return
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("apex_b200.ReplayMemory needs a CUDA device (no CPU fallback)")
        if self.state_shape is None:
            if state_shape is None:
                raise RuntimeError("replay used before any record was added and no params['state_shape'] given")
            self.state_shape = tuple(int(x) for x in state_shape)
        self.C = int(self.state_shape[0])
        self.F = int(np.prod(self.state_shape[1:]))
        if self.device_index is None:
            self.device_index = torch.cuda.current_device()
        self.device = torch.device("cuda", int(self.device_index))
        if self.layout == "stream":
            fpr, lookback = 1, self.C - 1 + self.n_step
        else:
            fpr, lookback = 2 * self.C, 0
        cap = self.soft_capacity
        if cap % self.lanes:
            cap += self.lanes - cap % self.lanes
        self.capacity, self.fpr, self.lookback = cap, fpr, lookback
        cfg = _C.ReplayConfig(cap, self.lanes, fpr, lookback, self.C, self.F, self.num_actions, self.alpha,
                              int(self.device_index))
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _C.call("apexb200_replay_create", C.byref(h), C.byref(cfg))
        self._h = h
        self._own_rel = np.arange(-(2 * self.C - 1), 1, dtype=np.int32)[None, :]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h is not None:
            try:
                _C.lib().apexb200_replay_destroy(h)
            except Exception:
                pass

    close = __del__

    def _stream(self):
        return _C.stream_ptr()

    def _dev(self, arr, dtype):
        """Host array -> device tensor through a reusable pinned staging buffer."""
        import torch
        a = np.ascontiguousarray(arr, dtype=dtype)
        key = (a.dtype.str, a.shape)
        ent = self._pinned.get(key)
        if ent is None:
            ent = [torch.from_numpy(np.empty(a.shape, a.dtype)).pin_memory(), None]
            if len(self._pinned) < 64:
                self._pinned[key] = ent
        pin, ev = ent
        if ev is not None:
            ev.synchronize()                      # the previous async copy out of this buffer has finished
        pin.numpy()[...] = a
        out = pin.to(self.device, non_blocking=True)
        ent[1] = torch.cuda.Event()
        ent[1].record()
        return out

    # ------------------------------------------------------------ compat tier
    def add(self, priorities, xp_batch):
        """replay.py:59-69: append records, then merge the priority dict."""
        xp_batch = list(xp_batch)
        priorities = dict(priorities)
        if xp_batch:
            first = np.asarray(xp_batch[0].S_t)
            self._ensure(first.shape if first.ndim == 3 else (1,) + first.shape)
            if self.layout != "materialised":
                raise RuntimeError("add(priorities, xp_batch) needs layout='materialised'; use insert() for streams")
            m = len(xp_batch)
            frames = np.empty((m, 2, self.C, self.F), np.uint8)
            for j, xp in enumerate(xp_batch):
                frames[j, 0] = _to_u8(xp.S_t).reshape(self.C, self.F)
                frames[j, 1] = _to_u8(xp.S_tpn).reshape(self.C, self.F)
            A = np.array([int(xp.A_t) for xp in xp_batch], np.int32)
            R = np.array([float(xp.R_ttpB) for xp in xp_batch], np.float32)
            G = np.array([float(xp.Gamma_ttpB) for xp in xp_batch], np.float32)
            own = []
            for xp in xp_batch:
                p = priorities.pop(xp.key, None)
                if p is None:
                    p = self._keys.pending.pop(xp.key, 0.0)
                own.append(float(p))
            q_t = q_tpn = None
            if self.num_actions and xp_batch[0].qS_t is not None:
                q_t = np.stack([np.asarray(xp.qS_t, np.float32) for xp in xp_batch])
                q_tpn = np.stack([np.asarray(xp.qS_tpn, np.float32) for xp in xp_batch])
            slots = self.insert(0, frames.reshape(m * 2 * self.C, self.F), np.repeat(self._own_rel, m, 0), A, R, G,
                                prio=np.array(own, np.float32), q_t=q_t, q_tpn=q_tpn)
            for xp, s in zip(xp_batch, slots):
                self._keys.assign(xp.key, int(s))
        if priorities:
            self.set_priorities(priorities)

    def set_priorities(self, new_priorities):
        """replay.py:32-42: add-or-overwrite by key (K2 point update)."""
        slots, vals = self._keys.split(dict(new_priorities))
        if slots:
            self.update_priorities(self._dev(slots, np.int32), self._dev(vals, np.float32))

    def update_sample_probabilities(self):
        """replay.py:18-30: with a sum tree this is a full rebuild of the inner levels (K1)."""
        if self._h is not None:
            _C.call("apexb200_tree_build", self._h, self._stream())

    def sample(self, sample_size):
        """replay.py:44-57: `sample_size` i.i.d. proportional draws, records returned in draw order.

        The uniforms come from numpy's global RNG, the stream np.random.choice
        (replay.py:52) consumes, one per draw.
        """
        import torch
        self._ensure()
        B = int(sample_size)
        u = np.random.random_sample(B)
        b = self.sample_tensors(B, u=self._dev(u, np.float64), with_q=bool(self.num_actions))
        host = {k: v.cpu() for k, v in b._asdict().items() if v is not None and k != "fsel"}
        torch.cuda.current_stream().synchronize()
        self.check_status()
        S_t = host["S_t"].numpy().reshape((B,) + self.state_shape)
        S_tpn = host["S_tpn"].numpy().reshape((B,) + self.state_shape)
        idx, A, R, G = (host[k].numpy() for k in ("idx", "A", "R", "Gamma"))
        q_t = self._last_q[0].cpu().numpy() if self._last_q is not None else [None] * B
        q_tpn = self._last_q[1].cpu().numpy() if self._last_q is not None else [None] * B
        return [N_Step_Transition(S_t[j], int(A[j]), float(R[j]), float(G[j]), q_t[j], S_tpn[j], q_tpn[j],
                                  self._keys.key_of.get(int(idx[j]))) for j in range(B)]

    def remove_to_fit(self):
        """replay.py:71-80: the ring evicts on insert, so there is never an excess to trim."""
        return None

    def size(self):
        """replay.py:82-83."""
        if self._h is None:
            return 0
        n = C.c_int64()
        _C.call("apexb200_replay_size", self._h, C.byref(n))
        return int(n.value)

    # -------------------------------------------------------------- fast tier
    def insert(self, lane, frames, rel, A, R, Gamma, prio=None, q_t=None, q_tpn=None):
        """K6.  Arguments may be host arrays (staged through pinned memory) or CUDA tensors.

        Returns the record indices (numpy int64) the new records occupy.
        """
        import torch
        self._ensure()

        def dev(x, dt):
            if x is None or isinstance(x, torch.Tensor):
                return x
            return self._dev(x, dt)
        count = int(len(A))
        frames, rel, A, R, Gamma = dev(frames, np.uint8), dev(rel, np.int32), dev(A, np.int32), dev(R, np.float32), dev(Gamma, np.float32)
        prio, q_t, q_tpn = dev(prio, np.float32), dev(q_t, np.float32), dev(q_tpn, np.float32)
        first = C.c_int64()
        _C.call("apexb200_replay_insert", self._h, int(lane), count, _C.ptr(frames), _C.ptr(rel), _C.ptr(A), _C.ptr(R),
                _C.ptr(Gamma), _C.ptr(prio), _C.ptr(q_t), _C.ptr(q_tpn), C.byref(first), self._stream())
        lane_cap = self.capacity // self.lanes
        base = int(lane) * lane_cap
        return base + (first.value - base + np.arange(count, dtype=np.int64)) % lane_cap

    def insert_lanes(self, lanes, frames, rel, A, R, Gamma, prio=None, q_t=None, q_tpn=None):
        """K6 for several lanes in one launch pair (lock-stepped actors).  lanes: distinct lane ids [E];
        every other argument carries a leading lane dimension: frames [E, T*frames_per_record, F],
        rel [E, T, 2C], A/R/Gamma/prio [E, T], q_t/q_tpn [E, T, num_actions] — contiguous CUDA tensors
        or host arrays."""
        import torch
        self._ensure()

        def dev(x, dt):
            if x is None:
                return None
            return x.contiguous() if isinstance(x, torch.Tensor) else self._dev(np.ascontiguousarray(x), dt)
        lanes = (C.c_int32 * len(lanes))(*[int(x) for x in lanes])
        E, T = int(A.shape[0]), int(A.shape[1])
        if E != len(lanes):
            raise ValueError("insert_lanes: %d lanes but inputs for %d" % (len(lanes), E))
        frames, rel, A, R, Gamma = dev(frames, np.uint8), dev(rel, np.int32), dev(A, np.int32), dev(R, np.float32), dev(Gamma, np.float32)
        prio, q_t, q_tpn = dev(prio, np.float32), dev(q_t, np.float32), dev(q_tpn, np.float32)
        _C.call("apexb200_replay_insert_lanes", self._h, lanes, E, T, _C.ptr(frames), _C.ptr(rel), _C.ptr(A), _C.ptr(R),
                _C.ptr(Gamma), _C.ptr(prio), _C.ptr(q_t), _C.ptr(q_tpn), self._stream())

    def _outputs(self, B, with_q, f32=False):
        import torch
        key = (B, with_q) if not f32 else (B, with_q, "f32")
        out = self._out_cache.get(key)
        if out is None:
            d = self.device
            sdt = torch.float32 if f32 else torch.uint8
            out = dict(idx=torch.empty(B, dtype=torch.int32, device=d),
                       fsel=torch.empty((B, 2 * self.C), dtype=torch.int32, device=d),
                       S_t=torch.empty((B, self.C, self.F), dtype=sdt, device=d),
                       S_tpn=torch.empty((B, self.C, self.F), dtype=sdt, device=d),
                       A=torch.empty(B, dtype=torch.int32, device=d),
                       R=torch.empty(B, dtype=torch.float32, device=d),
                       Gamma=torch.empty(B, dtype=torch.float32, device=d),
                       q_t=torch.empty((B, self.num_actions), dtype=torch.float32, device=d) if with_q else None,
                       q_tpn=torch.empty((B, self.num_actions), dtype=torch.float32, device=d) if with_q else None)
            self._out_cache[key] = out
        return out

    def sample_indices(self, u, idx_out=None, fsel_out=None, leaf_out=None, total_out=None):
        """K3 only: u (CUDA float64 [B]) -> record indices (CUDA int32 [B])."""
        import torch
        self._ensure()
        B = int(u.numel())
        if idx_out is None:
            idx_out = torch.empty(B, dtype=torch.int32, device=self.device)
        _C.call("apexb200_sample", self._h, _C.ptr(u), B, _C.ptr(idx_out), _C.ptr(leaf_out), _C.ptr(total_out),
                _C.ptr(fsel_out), self._stream())
        return idx_out

    def sample_targets(self, targets, count_dev, idx_out=None, fsel_out=None, leaf_out=None):
        """K3 from explicit integer targets (sharded sampling): targets int64-viewed u64 [max], count_dev int32 [1]."""
        import torch
        self._ensure()
        n = int(targets.numel())
        if idx_out is None:
            idx_out = torch.empty(n, dtype=torch.int32, device=self.device)
        _C.call("apexb200_sample_targets", self._h, _C.ptr(targets), _C.ptr(count_dev), n, _C.ptr(idx_out),
                _C.ptr(leaf_out), _C.ptr(fsel_out), self._stream())
        return idx_out

    def gather(self, idx, fsel=None, with_q=False, out=None, count_dev=None):
        """K4 only: record indices -> Batch of device tensors (rows >= *count_dev are left untouched)."""
        self._ensure()
        B = int(idx.numel())
        o = out if out is not None else self._outputs(B, with_q)
        _C.call("apexb200_gather", self._h, _C.ptr(idx), _C.ptr(fsel), B, _C.ptr(count_dev), _C.ptr(o["S_t"]), _C.ptr(o["S_tpn"]),
                _C.ptr(o["A"]), _C.ptr(o["R"]), _C.ptr(o["Gamma"]), _C.ptr(o.get("q_t")), _C.ptr(o.get("q_tpn")),
                self._stream())
        self._last_q = (o["q_t"], o["q_tpn"]) if with_q and o.get("q_t") is not None else None
        return Batch(idx, o["S_t"], o["S_tpn"], o["A"], o["R"], o["Gamma"], fsel)

    def sample_stratified(self, batch_size, idx_out=None, fsel_out=None, u_out=None, leaf_out=None, total_out=None):
        """K3 with the stratified draws generated in the kernel from (params['seed'], device call counter)."""
        import torch
        self._ensure()
        B = int(batch_size)
        if idx_out is None:
            idx_out = torch.empty(B, dtype=torch.int32, device=self.device)
        _C.call("apexb200_sample_stratified", self._h, self.seed, B, _C.ptr(idx_out), _C.ptr(leaf_out),
                _C.ptr(total_out), _C.ptr(fsel_out), _C.ptr(u_out), self._stream())
        return idx_out

    def sample_tensors(self, batch_size, u=None, with_q=False, u_out=None, fused=True, f32=False):
        """ReplayMemory.sample on the device: draw `batch_size` records and gather them into
        [B, C, F] uint8 batch tensors (K3 + K4, one fused launch unless fused=False).

        u: CUDA float64 uniforms in [0,1); None = stratified draws generated inside the kernel
        (written to u_out if given).  f32=True: S_t / S_tpn come back as float32 [B, C, F] — the dtype
        the network consumes (learner.py:37-38), every byte widened exactly by the gather kernel itself
        (fused kernel only).  The returned tensors are reused by the next call with the same batch size.
        """
        self._ensure()
        B = int(batch_size)
        o = self._outputs(B, with_q, f32)
        if f32 and not fused:
            raise ValueError("float32 batches are written by the fused sample+gather kernel only")
        if fused:
            _C.call("apexb200_sample_gather", self._h, _C.ptr(u), self.seed, B, _C.ptr(o["idx"]), _C.ptr(o["fsel"]),
                    None, None, _C.ptr(u_out), _C.ptr(o["S_t"]), _C.ptr(o["S_tpn"]), _C.ptr(o["A"]), _C.ptr(o["R"]),
                    _C.ptr(o["Gamma"]), _C.ptr(o.get("q_t")), _C.ptr(o.get("q_tpn")),
                    _C.FMT_F32 if f32 else _C.FMT_U8, self._stream())
            self._last_q = (o["q_t"], o["q_tpn"]) if with_q and o.get("q_t") is not None else None
            return Batch(o["idx"], o["S_t"], o["S_tpn"], o["A"], o["R"], o["Gamma"], o["fsel"])
        if u is None:
            self.sample_stratified(B, idx_out=o["idx"], fsel_out=o["fsel"], u_out=u_out)
        else:
            self.sample_indices(u, idx_out=o["idx"], fsel_out=o["fsel"])
        return self.gather(o["idx"], fsel=o["fsel"], with_q=with_q, out=o)

    def update_priorities(self, idx, abs_td):
        """K2: leaves[idx] = max(|td|, eps)^alpha, last writer wins on duplicates."""
        _C.call("apexb200_tree_update", self._h, _C.ptr(idx), _C.ptr(abs_td), int(idx.numel()), self.eps, self._stream())

    def set_leaves(self, idx, leaves):
        _C.call("apexb200_tree_set_leaves", self._h, _C.ptr(idx), _C.ptr(leaves), int(idx.numel()), self._stream())

    def workspace(self, B):
        import torch
        n = C.c_int64()
        _C.call("apexb200_td_workspace_bytes", int(B), C.byref(n))
        if self._ws is None or self._ws.numel() < n.value:
            self._ws = torch.empty(int(n.value), dtype=torch.uint8, device=self.device)
        return self._ws

    def td_update(self, idx, q_on_t, q_on_tpn, q_tg_tpn, A, R, Gamma, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0, out=None,
                  count_dev=None, mean_divisor=0.0, weights=None):
        """K5 + K2 fused (learner.py:70 + :76): returns (delta, prio, loss, loss_mean) device tensors.
        count_dev / mean_divisor: live-row count on the device and global batch size for sharded batches."""
        import torch
        B, nA = q_on_t.shape
        if out is None:
            out = tuple(torch.empty(B, dtype=torch.float32, device=self.device) for _ in range(3)) + \
                (torch.empty(1, dtype=torch.float32, device=self.device),)
        ws = self.workspace(B)
        _C.call("apexb200_td_fwd_update", self._h, _C.ptr(idx), _C.ptr(q_on_t), _C.ptr(q_on_tpn), _C.ptr(q_tg_tpn),
                _C.ptr(A), _C.ptr(R), _C.ptr(Gamma), int(B), int(nA), self.eps, int(loss_kind), float(kappa),
                _C.ptr(weights), _C.ptr(out[0]), _C.ptr(out[1]), _C.ptr(out[2]), _C.ptr(out[3]), _C.ptr(ws), int(ws.numel()),
                _C.ptr(count_dev), float(mean_divisor), self._stream())
        return out

    def total(self):
        """Tree total as a python int (synchronises)."""
        import torch
        self._ensure()
        t = torch.zeros(1, dtype=torch.int64, device=self.device)
        _C.call("apexb200_tree_total", self._h, _C.ptr(t), self._stream())
        return int(t.item()) & ((1 << 64) - 1)

    def total_tensor(self, out):
        _C.call("apexb200_tree_total", self._h, _C.ptr(out), self._stream())
        return out

    def check_status(self):
        st = C.c_int32()
        _C.call("apexb200_replay_status", self._h, C.byref(st), self._stream())
        if st.value:
            names = [n for bit, n in ((1, "sample on empty tree"), (2, "record/frame index out of range"),
                                      (4, "target beyond tree total"), (8, "action out of range"),
                                      (16, "peer exchange timed out"), (32, "shard quota exceeded (max_rows)"))
                     if st.value & bit]
            raise _C.ApexError(_C.EINVAL, "device status: " + ", ".join(names))

    def buffer(self, which):
        """Raw device view of an internal array as a torch tensor (tests / checkpoint)."""
        import torch
        p, n = C.c_void_p(), C.c_int64()
        _C.call("apexb200_replay_buffer", self._h, int(which), C.byref(p), C.byref(n))
        dt = {_C.BUF_FRAMES: torch.uint8, _C.BUF_FIDX: torch.int32, _C.BUF_ACTION: torch.int32,
              _C.BUF_LEAVES: torch.int64}.get(which, torch.float32)
        if not p.value:
            return None
        return torch.as_tensor(_RawCuda(p.value, n.value, self.device.index), device=self.device).view(dt)

    def pack_frames(self, first_slot=0, count=None):
        """Lossless packing of frame slots [first_slot, first_slot + count) on the device (XOR with the previous
        frame, key frame every 64, only non-zero 16-byte blocks stored; apexb200_frames_pack*).  Returns
        (sizes int32-viewed u32 [count], offsets int64 [count], packed uint8 [total]) as CUDA tensors."""
        import torch
        self._ensure()
        if count is None:
            count = self.capacity * self.fpr - int(first_slot)
        count = int(count)
        sizes = torch.empty(count, dtype=torch.int32, device=self.device)
        _C.call("apexb200_frames_pack_sizes", self._h, int(first_slot), count, _C.ptr(sizes), self._stream())
        ends = torch.cumsum(sizes.to(torch.int64), 0)
        offsets = (ends - sizes).contiguous()
        total = int(ends[-1].item()) if count else 0
        packed = torch.empty(max(total, 16), dtype=torch.uint8, device=self.device)
        _C.call("apexb200_frames_pack", self._h, int(first_slot), count, _C.ptr(offsets), _C.ptr(packed), self._stream())
        return sizes, offsets, packed[:total]

    def unpack_frames(self, offsets, packed, first_slot=0):
        """Inverse of pack_frames: writes the frames back into the store (device tensors, or host tensors that are
        copied first)."""
        self._ensure()
        offsets = offsets.to(self.device).contiguous()
        packed = packed.to(self.device).contiguous()
        _C.call("apexb200_frames_unpack", self._h, int(first_slot), int(offsets.numel()), _C.ptr(offsets), _C.ptr(packed),
                self._stream())

    def state_dict(self, pack_frames=False):
        """Checkpoint of the whole shard (device -> host copies of ring, metadata, leaves, heads).
        pack_frames=True: the frame store travels losslessly packed (see pack_frames; frames of consecutive env
        steps differ in a few sprites, so Atari-like streams shrink by an order of magnitude)."""
        heads = (C.c_int64 * self.lanes)()
        _C.call("apexb200_replay_get_heads", self._h, heads)
        out = {"heads": list(heads), "config": dict(capacity=self.capacity, lanes=self.lanes, layout=self.layout,
                                                    state_shape=self.state_shape, alpha=self.alpha)}
        for name, which in (("frames", _C.BUF_FRAMES), ("fidx", _C.BUF_FIDX), ("A", _C.BUF_ACTION), ("R", _C.BUF_REWARD),
                            ("Gamma", _C.BUF_GAMMA), ("q_t", _C.BUF_QT), ("q_tpn", _C.BUF_QTPN), ("leaves", _C.BUF_LEAVES)):
            if name == "frames" and pack_frames:
                chunks, step = [], 1 << 16                       # a chunk of frame slots at a time: bounded scratch
                for lo in range(0, self.capacity * self.fpr, step):
                    n = min(step, self.capacity * self.fpr - lo)
                    _sizes, offsets, packed = self.pack_frames(lo, n)
                    chunks.append((lo, offsets.cpu(), packed.cpu()))
                out["frames"] = None
                out["frames_packed"] = chunks
                continue
            t = self.buffer(which)
            out[name] = None if t is None else t.cpu()
        return out

    def load_state_dict(self, sd):
        self._ensure(sd["config"]["state_shape"])
        for name, which in (("frames", _C.BUF_FRAMES), ("fidx", _C.BUF_FIDX), ("A", _C.BUF_ACTION), ("R", _C.BUF_REWARD),
                            ("Gamma", _C.BUF_GAMMA), ("q_t", _C.BUF_QT), ("q_tpn", _C.BUF_QTPN), ("leaves", _C.BUF_LEAVES)):
            t = self.buffer(which)
            if t is not None and sd.get(name) is not None:
                t.copy_(sd[name])
        for lo, offsets, packed in sd.get("frames_packed") or ():
            self.unpack_frames(offsets, packed, lo)
        heads = (C.c_int64 * self.lanes)(*sd["heads"])
        _C.call("apexb200_replay_set_heads", self._h, heads)
        self.update_sample_probabilities()


class _RawCuda:
    """__cuda_array_interface__ shim so torch can view library-owned device memory."""

    def __init__(self, ptr, nbytes, device):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def _to_u8(x):
    """Reference observations are float64 arrays; the ring stores uint8."""
    a = np.asarray(x)
    if a.dtype == np.uint8:
        return a
    return np.clip(np.rint(a), 0, 255).astype(np.uint8)
