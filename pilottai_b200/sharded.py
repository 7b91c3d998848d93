"""Row-sharded semantic store: one process per GPU, torch.distributed for the plumbing.

SURVEY.md section 8e: global insertion index i lives on rank i % world at local sequence
i // world (round-robin keeps streaming inserts and ring eviction balanced).  A search is
  local fused search (exact top-k of the shard, fp64 scores, GLOBAL indices)
  -> all_gather of the [Q, k] (index, score) lists  (NCCL on GPUs, gloo in CPU tests)
  -> pilott_topk_merge: global top-k by (score desc, index asc)
which is identical to what one store holding every row returns.  There is no reference
counterpart: the reference is single-process.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from .store import SemanticStore, topk_merge


def shard_rows(first_global: int, n: int, rank: int, world: int) -> np.ndarray:
    """Positions (0..n-1) within a batch starting at global index first_global that rank owns."""
    start = (rank - first_global) % world
    return np.arange(start, n, world, dtype=np.int64)


class ShardedSemanticStore:
    def __init__(self, dim: int, capacity: int, rank: int, world: int, device: int,
                 group: Optional[dist.ProcessGroup] = None):
        self.rank, self.world, self.group = rank, world, group
        self.local = SemanticStore(dim, (capacity + world - 1) // world, device)
        self.local.set_sequence_map(world, rank)
        self.dim, self.device = dim, device
        self.total = 0                                  # global rows ever appended
        self._gi = self._gs = None

    def append(self, x, prio=None, tagmask=None, expires_us=None) -> int:
        """Every rank calls this with the SAME global batch (host numpy or CUDA tensor); each
        keeps its own rows.  No collective on the insert path."""
        n = x.shape[0]
        first = self.total
        rows = shard_rows(first, n, self.rank, self.world)
        if rows.size:
            if isinstance(x, np.ndarray):
                sel = lambda a: None if a is None else np.ascontiguousarray(a[rows])
            else:
                ridx = torch.from_numpy(rows).to(x.device)
                sel = lambda a: None if a is None else a.index_select(0, ridx)
            self.local.append(sel(x), sel(prio), sel(tagmask), sel(expires_us))
        self.total += n
        return first

    def search(self, q: torch.Tensor, k: int, **flt):
        """q: CUDA fp32 [Q, dim], same on every rank.  Returns (idx, score, score64) on every rank,
        plus the local inexact counter (callers check it, see SemanticStore.search)."""
        idx, sc, sc64, inexact, flags = self.local.search(q, k, want_score64=True, **flt)
        if self.world == 1:
            return idx, sc, sc64, inexact
        nq = q.shape[0]
        if self._gi is None or self._gi.shape[1:] != (nq, k):
            self._gi = torch.empty((self.world, nq, k), dtype=torch.int64, device=q.device)
            self._gs = torch.empty((self.world, nq, k), dtype=torch.float64, device=q.device)
        dist.all_gather_into_tensor(self._gi, idx, group=self.group)
        dist.all_gather_into_tensor(self._gs, sc64, group=self.group)
        m_idx, m_sc, m_sc64 = topk_merge(self._gi, self._gs)
        return m_idx, m_sc, m_sc64, inexact

    def close(self):
        self.local.close()
