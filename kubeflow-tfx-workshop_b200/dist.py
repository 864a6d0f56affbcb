"""Row sharding and the analyzer-state exchange between ranks (one process per GPU).

The reference merges analyzer accumulators inside Beam (``CombineGlobally`` / ``CombinePerKey``,
SURVEY.md 3.1); here the per-GPU partial states are merged with ``torch.distributed`` collectives
(NCCL over NVLink on GPUs, gloo in the CPU tests): integer histograms/cell counts by all-reduce,
Welford triples and vocabulary entries by all-gather.  Everything here works on CPU tensors too.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_total: int, rank: int, world: int, align: int = 8) -> Tuple[int, int]:
    """Contiguous row range of `rank`; boundaries are multiples of `align` rows (bitmap bytes)."""
    per = -(-n_total // world)
    per = -(-per // align) * align
    lo = min(rank * per, n_total)
    hi = min(lo + per, n_total)
    return lo, hi


class Comm:
    """Thin wrapper over a torch.distributed process group (None -> single process)."""

    def __init__(self, group: Optional[object] = None):
        self.enabled = dist.is_available() and dist.is_initialized()
        self.group = group
        self.rank = dist.get_rank(group) if self.enabled else 0
        self.world = dist.get_world_size(group) if self.enabled else 1

    @property
    def active(self) -> bool:
        return self.enabled and self.world > 1

    def all_gather(self, t: torch.Tensor) -> torch.Tensor:
        """-> [world, *t.shape], rank order."""
        if not self.active:
            return t.unsqueeze(0)
        t = t.contiguous()
        out = torch.empty((self.world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t, group=self.group)
        return out.view((self.world,) + tuple(t.shape))

    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.active:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_max_(self, t: torch.Tensor) -> torch.Tensor:
        if self.active:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t

    def all_gather_var(self, t: torch.Tensor, count: int) -> Tuple[torch.Tensor, List[int]]:
        """Concatenate the first `count` elements of every rank's 1-D tensor, in rank order."""
        if not self.active:
            return t[:count], [count]
        counts_t = torch.tensor([count], dtype=torch.int64, device=t.device)
        counts = [int(c) for c in self.all_gather(counts_t).flatten().tolist()]
        cap = max(max(counts), 1)
        pad = torch.zeros(cap, dtype=t.dtype, device=t.device)
        pad[:count] = t[:count]
        g = self.all_gather(pad)
        return torch.cat([g[r, :counts[r]] for r in range(self.world)]), counts

    def barrier(self):
        if self.active:
            dist.barrier(group=self.group)
