"""Multi-GPU replay: one shard per rank, every rank samples its share of one global batch.

No counterpart in the reference (single process).  Design from SURVEY §8(e): every rank owns a
replay shard; the only coupling is the global normaliser, the sum of the shard totals.  With exact
u64 totals, identical global draws on every rank and each rank descending its own tree for the
targets that fall into its prefix interval, the union of the ranks' rows is precisely what a single
tree over the concatenated shards would return.

Default path (peer memory, no NCCL on the step): the shard totals travel as 8-byte words written
straight into the peers' exchange buffers over NVLink — by `exchange()` (one launch) or, with
`push_on_update()`, by the TD/priority-update kernel itself — and `sample_fused()` is ONE kernel
that waits for the words, finds the rank's run of the in-kernel stratified draws, searches and
gathers.  A rank receives every draw it owns whatever the split of priority mass (outputs are sized
for the whole global batch; the kernel runs extra waves for an over-sized share).

Alternative path for caller-supplied (non-stratified) draws: `gather_totals()` (NCCL all-gather of
the totals) + `route()` (apexb200_shard_targets) + `sample()`.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import ops


class ShardedSampler:
    def __init__(self, mem, group=None, rank=None, world_size=None, route=None):
        self.mem = mem
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world_size is None else world_size
        self._route = route or ops.shard_targets
        self._totals = None
        self._mine = None

    def gather_totals(self):
        """All-gather of the per-shard totals -> int64 [world] (bit pattern of the u64 totals)."""
        dev = self.mem.device
        if self._totals is None:
            self._totals = torch.empty(self.world, dtype=torch.int64, device=dev)
            self._mine = torch.empty(1, dtype=torch.int64, device=dev)
        self.mem.total_tensor(self._mine)
        if self.world > 1:
            dist.all_gather_into_tensor(self._totals, self._mine, group=self.group)
        else:
            self._totals.copy_(self._mine)
        return self._totals

    def route(self, u_global):
        """-> (local_targets [Bg], local_pos [Bg], local_count [1]) for this rank; identical u on every rank."""
        return self._route(self.gather_totals(), self.rank, u_global)

    # ---- peer-memory path: no NCCL call, no host round trip on the per-step path -----------
    def connect_peers(self):
        """Exchange CUDA-IPC handles of the shards' exchange buffers (once, at start-up) and map
        every peer's buffer.  Collective over the group."""
        import ctypes as C
        from . import _C
        self.mem._ensure()
        mine = (C.c_ubyte * _C.IPC_HANDLE_BYTES)()
        _C.call("apexb200_peer_export", self.mem._h, mine)
        handles = [None] * self.world
        if self.world > 1:
            dist.all_gather_object(handles, bytes(mine), group=self.group)
        else:
            handles = [bytes(mine)]
        blob = (C.c_ubyte * (_C.IPC_HANDLE_BYTES * self.world)).from_buffer_copy(b"".join(handles))
        _C.call("apexb200_peer_connect", self.mem._h, int(self.rank), int(self.world), blob)
        if self.world > 1:
            dist.barrier(group=self.group)
        self.connected = True

    def push_on_update(self, enabled=True):
        """Let every td_update publish the new shard total to the peers from inside its kernel
        (fused compute + collective); the step loop then needs no exchange() call."""
        from . import _C
        _C.call("apexb200_peer_push_on_update", self.mem._h, 1 if enabled else 0)

    def connect_local(self, samplers):
        """Single-process variant of connect_peers: `samplers` are the ShardedSamplers of ALL shards, in rank
        order, living in this process (one GPU or several with peer access).  No IPC, no collective."""
        import ctypes as C
        from . import _C
        for s_ in samplers:
            s_.mem._ensure()
        arr = (C.c_void_p * len(samplers))(*[s_.mem._h.value for s_ in samplers])
        _C.call("apexb200_peer_connect_local", self.mem._h, int(self.rank), len(samplers), arr)
        self.connected = True

    def push(self):
        """Publish this shard's total as a new exchange round without waiting for the peers' (the publishing
        half of exchange(); for shards driven in lock step from one process)."""
        from . import _C
        _C.call("apexb200_shard_push", self.mem._h, _C.stream_ptr())

    def peer_wait_stats(self, reset=True):
        """(mean microseconds at 1965 MHz, launches): how long the fused sharded kernels waited in-kernel for the
        slowest peer's total since the last reset (synchronises)."""
        import ctypes as C
        from . import _C
        st = (C.c_uint64 * 2)()
        _C.call("apexb200_peer_wait_stats", self.mem._h, st, 1 if reset else 0, _C.stream_ptr())
        n = int(st[1])
        return (int(st[0]) / n / 1965.0 if n else 0.0), n

    def exchange(self, totals_out=None):
        """One-launch all-gather of the shard totals over peer memory (collective: every rank
        calls it once per step, before sample_fused)."""
        from . import _C
        _C.call("apexb200_shard_exchange", self.mem._h, _C.ptr(totals_out), _C.stream_ptr())

    def sample_fused(self, global_count, max_rows=None, with_q=False, u_out=None, f32=False):
        """Fused routing + search + gather of this rank's share of `global_count` stratified draws
        generated in the kernel (identical on every rank).  Returns (Batch over max_rows rows,
        pos [max_rows] global draw index per row, count [1] rows owned).  max_rows defaults to
        global_count — the outputs can then hold the whole batch and no draw is ever dropped, however
        uneven the shards' priority mass; only the live rows are written (and touch HBM)."""
        import torch
        from . import _C
        from .replay import Batch
        mem = self.mem
        if max_rows is None:
            max_rows = int(global_count)
        o = mem._outputs(int(max_rows), with_q, f32)               # f32: float32 batches, as in sample_tensors
        if "pos" not in o:
            o["pos"] = torch.empty(int(max_rows), dtype=torch.int32, device=mem.device)
            o["count"] = torch.zeros(1, dtype=torch.int32, device=mem.device)
        _C.call("apexb200_sample_gather_sharded", mem._h, mem.seed, int(global_count), int(max_rows), _C.ptr(o["idx"]),
                _C.ptr(o["fsel"]), _C.ptr(o["pos"]), _C.ptr(u_out), _C.ptr(o["count"]), _C.ptr(o["S_t"]),
                _C.ptr(o["S_tpn"]), _C.ptr(o["A"]), _C.ptr(o["R"]), _C.ptr(o["Gamma"]), _C.ptr(o.get("q_t")),
                _C.ptr(o.get("q_tpn")), _C.FMT_F32 if f32 else _C.FMT_U8, _C.stream_ptr())
        return Batch(o["idx"], o["S_t"], o["S_tpn"], o["A"], o["R"], o["Gamma"], o["fsel"]), o["pos"], o["count"]

    def sample(self, u_global, out=None, with_q=False):
        """Global draws -> this rank's part of the batch (rows beyond local_count are stale)."""
        lt, lp, lc = self.route(u_global)
        Bg = int(u_global.numel())
        o = out if out is not None else self.mem._outputs(Bg, with_q)
        self.mem.sample_targets(lt, lc, idx_out=o["idx"], fsel_out=o["fsel"])
        batch = self.mem.gather(o["idx"], fsel=o["fsel"], with_q=with_q, out=o, count_dev=lc)
        return batch, lp, lc
