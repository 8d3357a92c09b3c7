"""Multi-GPU replay: one shard per rank, quotas from an all-gather of shard totals.

No counterpart in the reference (single process).  Design from SURVEY §8(e): every
rank owns a replay shard; per step the ranks all-gather their tree totals (8 bytes
each, NCCL over NVLink), derive the same global integer targets from the same
uniform draws, and each rank descends its own tree for the targets that fall into
its prefix interval.  With exact u64 totals this returns precisely the records a
single tree over the concatenated shards would return.
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

    def sample(self, u_global, out=None, with_q=False):
        """Global draws -> this rank's part of the batch (rows beyond local_count are stale)."""
        lt, lp, lc = self.route(u_global)
        Bg = int(u_global.numel())
        o = out if out is not None else self.mem._outputs(Bg, with_q)
        self.mem.sample_targets(lt, lc, idx_out=o["idx"], fsel_out=o["fsel"])
        batch = self.mem.gather(o["idx"], fsel=o["fsel"], with_q=with_q, out=o, count_dev=lc)
        return batch, lp, lc
