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

from .store import SemanticStore, topk_merge_packed


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
        self._bufs = {}

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

    def _buffers(self, nq: int, k: int, dev, slot: int):
        key = (nq, k, slot)
        if key not in self._bufs:
            mine = torch.empty((2, nq, k), dtype=torch.int64, device=dev)          # [0] indices, [1] fp64 score bits
            self._bufs[key] = {
                "mine": mine, "idx": mine[0], "sc64": mine[1].view(torch.float64),
                "sc": torch.empty((nq, k), dtype=torch.float32, device=dev),
                "flags": torch.empty(nq, dtype=torch.uint8, device=dev),
                "inexact": torch.zeros(1, dtype=torch.int32, device=dev),
                "gathered": torch.empty((self.world, 2, nq, k), dtype=torch.int64, device=dev),
            }
        return self._bufs[key]

    def search_begin(self, q: torch.Tensor, k: int, slot: int = 0, resolve: bool = False, **flt):
        """Local search + ONE asynchronous all-gather of the packed (index, score) block.
        Returns a ticket for search_end(); with two slots a caller overlaps the exchange of
        batch i with the local search of batch i+1.  resolve=True reads the local inexact
        counter (one host sync) and repairs flagged queries before the exchange; with
        resolve=False the caller checks the counter search_end() returns."""
        b = self._buffers(q.shape[0], k, q.device, slot)
        self.local.search_into(q, k, b["idx"], b["sc"], b["sc64"], b["inexact"], b["flags"], **flt)
        if resolve and int(b["inexact"].item()):
            self.local.refine(q, k, b["flags"].nonzero(), (b["idx"], b["sc"], b["sc64"]), **flt)
            b["inexact"].zero_()
        work = None
        if self.world > 1:
            work = dist.all_gather_into_tensor(b["gathered"], b["mine"], group=self.group, async_op=True)
        return b, work

    def search_end(self, ticket):
        b, work = ticket
        if self.world == 1:
            return b["idx"], b["sc"], b["sc64"], b["inexact"]
        work.wait()
        m_idx, m_sc, m_sc64 = topk_merge_packed(b["gathered"])
        return m_idx, m_sc, m_sc64, b["inexact"]

    def search(self, q: torch.Tensor, k: int, **flt):
        """q: CUDA fp32 [Q, dim], same on every rank.  Returns (idx, score, score64) on every rank,
        plus the local inexact counter (0: flagged queries are repaired locally before the exchange)."""
        return self.search_end(self.search_begin(q, k, 0, True, **flt))

    def close(self):
        self.local.close()
