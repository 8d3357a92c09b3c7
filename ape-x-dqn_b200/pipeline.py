"""Host-facing step: draws and Q-values arrive in HOST memory, loss / priorities / indices
are read back on the host — the same K3+K4 -> K5+K2 step as the tensor API, issued by one
C-ABI call (apexb200_host_step: pinned H2D copy, two kernels, one D2H copy, event record);
use_graph=True replays a torch CUDA graph of the tensor-API calls instead.

Replaces, for a caller that keeps its numbers on the host like the reference does, the
round trip learner.py:68 (sample) -> :70 (loss, priorities) -> :76 (set_priorities).
`depth` buffer sets (default 3, at most _C.HOST_STEP_SLOTS) are used in turn, so the host can fill and
submit steps k+1 .. k+depth-1 while step k runs; with depth 2 the device idles whenever the
host's wake-up + submit latency exceeds one step (measured: 5 us of every 22).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _C


class HostStepPipeline:
    DEPTH = 3

    def __init__(self, mem, batch_size, num_actions, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0, use_graph=False, depth=None):
        mem._ensure()
        if depth is not None:
            if not 1 <= int(depth) <= _C.HOST_STEP_SLOTS:
                raise ValueError("depth must be in [1, %d]" % _C.HOST_STEP_SLOTS)
            self.DEPTH = int(depth)
        self.mem, self.B, self.nA = mem, int(batch_size), int(num_actions)
        B, nA, dev = self.B, self.nA, mem.device
        self.in_bytes = B * 8 + 3 * B * nA * 4
        self.out_words = 4 * B + 4                       # delta | prio | loss | idx(int32 bits) | loss_mean
        self.slots = []
        mem._outputs(B, False)
        mem.workspace(B)
        for _ in range(self.DEPTH):
            h_in = torch.empty(self.in_bytes, dtype=torch.uint8).pin_memory()
            h_out = torch.empty(self.out_words, dtype=torch.float32).pin_memory()
            d_in = torch.empty(self.in_bytes, dtype=torch.uint8, device=dev)
            d_out = torch.empty(self.out_words, dtype=torch.float32, device=dev)
            slot = dict(h_in=h_in, h_out=h_out, d_in=d_in, d_out=d_out, event=torch.cuda.Event(), graph=None,
                        u_host=h_in[:B * 8].view(torch.float64).numpy(),
                        q_host=h_in[B * 8:].view(torch.float32).numpy().reshape(3, B, nA),
                        u=d_in[:B * 8].view(torch.float64), q=d_in[B * 8:].view(torch.float32).view(3, B, nA),
                        td_out=(d_out[0:B], d_out[B:2 * B], d_out[2 * B:3 * B], d_out[4 * B:4 * B + 1]),
                        idx_bits=d_out[3 * B:4 * B])
            self.slots.append(slot)
        self.loss_kind, self.kappa = loss_kind, kappa
        self._k = 0
        o = mem._outputs(B, False)
        self._batch = o
        self._lib = _C.lib()
        for k, s in enumerate(self.slots):                 # everything apexb200_host_step needs, prebuilt
            s["event"].record()                            # materialises the cudaEvent_t
            s["args"] = (mem._h, k, _C.ptr(s["h_in"]), _C.C.c_void_p(s["h_in"].data_ptr() + B * 8), B, nA, float(mem.eps),
                         int(loss_kind), float(kappa), _C.ptr(o["S_t"]), _C.ptr(o["S_tpn"]), _C.ptr(o["A"]), _C.ptr(o["R"]),
                         _C.ptr(o["Gamma"]), _C.ptr(s["h_out"]), _C.C.c_void_p(s["event"].cuda_event),
                         _C.stream_ptr())                  # kernels run on the stream current at construction
            out = s["h_out"].numpy()
            s["views"] = (out[4 * B:4 * B + 1], out[B:2 * B], out[3 * B:4 * B].view(np.int32))
            s["busy"] = False
        self._host_step = self._lib.apexb200_host_step
        torch.cuda.current_stream().synchronize()
        if use_graph:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for slot in self.slots:                    # nothing allocates in _enqueue: capture directly
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        self._enqueue(slot)
                    slot["graph"] = g
            torch.cuda.current_stream().wait_stream(side)

    def _enqueue(self, s):
        s["d_in"].copy_(s["h_in"], non_blocking=True)
        b = self.mem.sample_tensors(self.B, u=s["u"])
        self.mem.td_update(b.idx, s["q"][0], s["q"][1], s["q"][2], b.A, b.R, b.Gamma, self.loss_kind, self.kappa,
                           out=s["td_out"])
        s["idx_bits"].copy_(b.idx.view(torch.float32))
        s["h_out"].copy_(s["d_out"], non_blocking=True)

    def submit(self, u, q):
        """u: float64 [B] host array of draws; q: float32 [3, B, nA] host array (Q(S_t), Q(S_tpn), Q_target(S_tpn)).
        Returns a ticket for result()."""
        s = self.slots[self._k % self.DEPTH]
        if s["busy"]:
            s["event"].synchronize()                      # the slot's previous step has left the device
        s["u_host"][:] = u
        s["q_host"][:] = q
        if s["graph"] is not None:
            s["graph"].replay()
            s["event"].record()
        else:                                              # one C call: H2D, K3+K4, K5+K2, D2H, event record
            rc = self._host_step(*s["args"])
            if rc:
                _C.check(rc)
        s["busy"] = True
        self._k += 1
        return self._k - 1

    def result(self, ticket):
        """Blocks until step `ticket` has finished; returns (loss_mean, priorities[B], indices[B]) host views."""
        s = self.slots[ticket % self.DEPTH]
        s["event"].synchronize()
        s["busy"] = False
        mean, prio, idx = s["views"]
        return float(mean[0]), prio, idx


class ShardedHostStepPipeline:
    """HostStepPipeline for one rank of a sharded replay: Q rows arrive from pinned HOST memory, this rank's
    share of the global batch is sampled + gathered (in-kernel stratified draws, identical on every rank), and
    loss part / priorities / record indices / global draw positions / the row count are read back on the host.
    One apexb200_host_step_sharded call per step.  `max_rows` (default: the whole global batch, so that no
    draw can ever be dropped) is the number of rows the host supplies Q values for.  Collective discipline as
    for ShardedSampler: every rank submits the same number of steps; peers connected, push-on-update enabled."""
    DEPTH = 3

    def __init__(self, sampler, global_count, num_actions, max_rows=None, loss_kind=_C.LOSS_HALF_SQ, kappa=1.0, depth=None):
        mem = sampler.mem
        mem._ensure()
        if depth is not None:
            if not 1 <= int(depth) <= _C.HOST_STEP_SLOTS:
                raise ValueError("depth must be in [1, %d]" % _C.HOST_STEP_SLOTS)
            self.DEPTH = int(depth)
        self.mem, self.sampler = mem, sampler
        self.Bg, self.nA = int(global_count), int(num_actions)
        self.M = M = int(max_rows) if max_rows else self.Bg
        nA = self.nA
        self.in_bytes = 3 * M * nA * 4
        self.out_words = 5 * M + 4
        o = mem._outputs(M, False)
        self._batch = o
        self.slots = []
        for k in range(self.DEPTH):
            h_in = torch.empty(3 * M * nA, dtype=torch.float32).pin_memory()
            h_out = torch.empty(self.out_words, dtype=torch.float32).pin_memory()
            ev = torch.cuda.Event()
            ev.record()
            out = h_out.numpy()
            self.slots.append(dict(
                h_in=h_in, h_out=h_out, event=ev, busy=False, q_host=h_in.numpy().reshape(3, M, nA),
                views=(out[5 * M:5 * M + 1], out[M:2 * M], out[3 * M:4 * M].view(np.int32), out[4 * M:5 * M].view(np.int32),
                       out[5 * M + 1:5 * M + 2].view(np.int32)),
                args=(mem._h, k, int(mem.seed), self.Bg, M, _C.ptr(h_in), nA, float(mem.eps), int(loss_kind), float(kappa),
                      _C.ptr(o["S_t"]), _C.ptr(o["S_tpn"]), _C.ptr(o["A"]), _C.ptr(o["R"]), _C.ptr(o["Gamma"]),
                      _C.ptr(h_out), _C.C.c_void_p(ev.cuda_event), _C.stream_ptr())))
        self._host_step = _C.lib().apexb200_host_step_sharded
        self._k = 0
        torch.cuda.current_stream().synchronize()

    def submit(self, q, rows=None):
        """q: float32 [3, rows, nA] host array — Q(S_t), Q(S_tpn), Q_target(S_tpn) for the first `rows` rows this
        rank will own (rows defaults to max_rows).  Returns a ticket for result()."""
        s = self.slots[self._k % self.DEPTH]
        if s["busy"]:
            s["event"].synchronize()
        if rows is None:
            s["q_host"][:] = q
        else:
            s["q_host"][:, :rows] = q
        rc = self._host_step(*s["args"])
        if rc:
            _C.check(rc)
        s["busy"] = True
        self._k += 1
        return self._k - 1

    def result(self, ticket):
        """Blocks until step `ticket` has finished; returns (loss part, priorities[n], record indices[n],
        global draw positions[n], n) as host views, n = rows this rank owned in that step."""
        s = self.slots[ticket % self.DEPTH]
        s["event"].synchronize()
        s["busy"] = False
        mean, prio, idx, pos, cnt = s["views"]
        n = int(cnt[0])
        return float(mean[0]), prio[:n], idx[:n], pos[:n], n
