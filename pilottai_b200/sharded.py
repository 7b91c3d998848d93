"""Row-sharded semantic store: one process per GPU, torch.distributed for the plumbing.

SURVEY.md section 8e: global insertion index i lives on rank i % world at local sequence
i // world (round-robin keeps streaming inserts and ring eviction balanced).  A search is
  local fused search (exact top-k of the shard, fp64 scores, GLOBAL indices)
  -> exchange of the [Q, k] (index, score) blocks
  -> global top-k by (score desc, index asc)
which is identical to what one store holding every row returns.  There is no reference
counterpart: the reference is single-process.

Three exchanges:
  "copy" (default) -- pilott_xchg_publish_copy: the block and a flag are written into every peer's mapped
         receive buffer by the copy engines (cudaMemcpyAsync on a side stream), the merge kernel
         acquires the flags.  No kernel runs for the exchange, so nothing competes with the scan of
         the next batch for SMs: the pair scan fills 72 of a B200's 74 SM pairs with one CTA per
         SM, and a collective's kernel that lands on one of them holds that pair back for the
         length of the collective.  Measured on 8 B200, 1 M x 768, 1024-query batches, 0.177 ms of local search
         per step: 0.194 ms per step with this exchange, 0.230 with the NCCL one below.
         If peer mapping fails on any rank, every rank falls back to "nccl".
  "nccl" -- one all_gather_into_tensor of the packed block on NCCL's own stream (a process group
         of its own, held to one CTA) + pilott_topk_merge_packed; the exchange of batch i overlaps
         the local searches of the batches behind it.
  "p2p"  -- pilott_xchg_*: every rank maps every peer's receive buffer (CUDA IPC);
         finish_kernel stores each result row straight into the peers over NVLink and the last
         CTA releases a flag, the merge kernel acquires the flags and ranks.  No collective
         library on the data path.  Measured on 8 B200 (1 M x 768, 1024-query batches) it is the
         slowest of the three for this 160 KB payload (round 1: 0.254 ms per step): the remote stores,
         the system-scope fence and the flag release sit inside the finish kernel on the search stream.
  Select with exchange= or PILOTT_EXCHANGE=copy|nccl|p2p.  Whatever the exchange, the merge kernel
  runs on a side stream (search_end).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from ._lib import check, ptr
from .store import SemanticStore, topk_merge_packed

IPC_HANDLE_BYTES = 64


def shard_rows(first_global: int, n: int, rank: int, world: int) -> np.ndarray:
    """Positions (0..n-1) within a batch starting at global index first_global that rank owns."""
    start = (rank - first_global) % world
    return np.arange(start, n, world, dtype=np.int64)


class ShardedSemanticStore:
    def __init__(self, dim: int, capacity: int, rank: int, world: int, device: int,
                 group: Optional[dist.ProcessGroup] = None, exchange: str = "auto"):
        self.rank, self.world, self.group = rank, world, group
        if group is None and world > 1 and (exchange == "nccl" or (exchange == "auto" and os.environ.get("PILOTT_EXCHANGE", "copy") == "nccl")):
            self.group = self._small_nccl_group(world)
        if exchange == "auto":
            exchange = os.environ.get("PILOTT_EXCHANGE", "copy")
        if exchange not in ("p2p", "nccl", "copy"):
            raise ValueError("exchange must be 'auto', 'copy', 'p2p' or 'nccl'")
        self.exchange = exchange if world > 1 else "none"
        self._xchg = {}                                 # (nq, k) -> [handle, next step]
        self._ridx = {}                                 # cached row-selection indices for device appends
        if capacity % world:
            # with a remainder the shards together would hold more rows than one ring of `capacity`,
            # and a search could return rows a single store had already evicted
            raise ValueError(f"capacity ({capacity}) must be a multiple of the number of shards ({world})")
        self.local = SemanticStore(dim, capacity // world, device)
        self.local.set_sequence_map(world, rank)
        self.dim, self.device = dim, device
        self.total = 0                                  # global rows ever appended
        self._bufs = {}
        self._merge_stream = None

    @staticmethod
    def _small_nccl_group(world: int):
        """A process group of its own for the exchange, with NCCL held to one CTA.  The block
        exchanged per step is ~160 KB, so bandwidth does not matter -- where the collective's kernel
        RUNS does: the scan of the next batch wants 72 SM pairs of the 74, and every pair an all-gather
        with the default channel count takes from it is delayed by the whole all-gather
        (measured on 4 B200, 0.306 ms of local search per step: 0.358 ms per step with NCCL's default, 0.315 with one CTA).  PILOTT_NCCL_MAX_CTAS overrides (0 = default
        group, NCCL's own choice).  Collective: every rank constructs its store at the same point."""
        n = int(os.environ.get("PILOTT_NCCL_MAX_CTAS", "1"))
        if n <= 0:
            return None
        try:
            opts = dist.ProcessGroupNCCL.Options()
            opts.config.max_ctas = n
            opts.config.min_ctas = 1
            return dist.new_group(list(range(world)), pg_options=opts)
        except Exception:
            return None

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
                key = (int(rows[0]), n, x.device)                 # the pattern only depends on first % world and n
                ridx = self._ridx.get(key)
                if ridx is None:
                    ridx = self._ridx[key] = torch.from_numpy(rows).to(x.device)
                sel = lambda a: None if a is None else a.index_select(0, ridx)
            self.local.append(sel(x), sel(prio), sel(tagmask), sel(expires_us))
        self.total += n
        return first

    def _buffers(self, nq: int, k: int, dev, slot: int):
        key = (nq, k, slot)
        if key not in self._bufs:
            mine = torch.empty((2, nq, k), dtype=torch.int64, device=dev)          # [0] indices, [1] fp64 score bits
            # merged results + the local inexact counter in ONE allocation, so that a host-side
            # consumer fetches a batch with a single D2H copy: [idx int64 | score fp32 | counter int32 | pad]
            ib, sb = nq * k * 8, nq * k * 4
            pack = torch.zeros(ib + sb + 8, dtype=torch.uint8, device=dev)
            self._bufs[key] = {
                "mine": mine, "idx": mine[0], "sc64": mine[1].view(torch.float64),
                "sc": torch.empty((nq, k), dtype=torch.float32, device=dev),
                "flags": torch.empty(nq, dtype=torch.uint8, device=dev),
                "pack": pack,
                "inexact": pack[ib + sb: ib + sb + 4].view(torch.int32),
                "gathered": torch.empty((self.world, 2, nq, k), dtype=torch.int64, device=dev),
                "m_idx": pack[:ib].view(torch.int64).view(nq, k),
                "m_sc": pack[ib: ib + sb].view(torch.float32).view(nq, k),
                "m_sc64": torch.empty((nq, k), dtype=torch.float64, device=dev),
                "begun": torch.cuda.Event(), "merged": torch.cuda.Event(), "merge_pending": False,
            }
        return self._bufs[key]

    def _exchange_for(self, nq: int, k: int):
        """Peer-mapped exchange buffers for [2, nq, k] blocks (created collectively, once per shape).
        Falls back to the NCCL exchange for good if any rank cannot map its peers."""
        key = (nq, k)
        if key in self._xchg:
            return self._xchg[key]
        lib = _lib.load()
        handle = C.create_string_buffer(IPC_HANDLE_BYTES)
        x = C.c_void_p()
        rc = lib.pilott_xchg_create(self.device, self.rank, self.world, 2 * nq * k * 8, C.byref(x), handle)
        if os.environ.get("PILOTT_XCHG_FORCE_FAIL") == str(self.rank):     # test hook: this rank cannot map its peers
            lib.pilott_xchg_destroy(x)
            rc = -1
        gathered = [None] * self.world
        dist.all_gather_object(gathered, handle.raw if rc == 0 else None, group=self.group)
        ok = rc == 0 and all(h is not None for h in gathered)
        if ok:
            ok = lib.pilott_xchg_connect(x, b"".join(gathered)) == 0
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=torch.device("cuda", self.device))
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 0:
            if rc == 0:
                lib.pilott_xchg_destroy(x)
            self.exchange = "nccl"
            if self.group is None:                   # (collective: every rank takes this branch together)
                self.group = self._small_nccl_group(self.world)
            return None
        self._xchg[key] = [x, 0]
        return self._xchg[key]

    def search_begin(self, q: torch.Tensor, k: int, slot: int = 0, resolve: bool = False, **flt):
        """Local search, then the block is sent on its way (publish kernel over NVLink, or ONE
        asynchronous all-gather).  Returns a ticket for search_end(), which merges.  With two
        buffer slots a caller overlaps the exchange of batch i with the local search of batch i+1
        (at most THREE search_begin ahead of the oldest outstanding search_end with the "copy" and "p2p"
        exchanges -- their buffers have 8 slots, see csrc/xchg.cu -- and with distinct `slot`s).
        resolve=True reads the local inexact counter (one host sync) and repairs flagged queries
        before the exchange.  With resolve=False flagged rows are exchanged as they are, and every
        rank only sees its OWN counter: sum `self.local.flagged_total()` over the ranks (one
        all-reduce, at any later point -- the library counts across searches) before trusting the
        merged results, as bench.py does."""
        b = self._buffers(q.shape[0], k, q.device, slot)
        if b.get("merge_pending"):                        # this slot's previous merge may still read its blocks
            if not b["merged"].query():                   # (normally long done: no wait, and the launches stay back to back)
                torch.cuda.current_stream(self.device).wait_event(b["merged"])
            b["merge_pending"] = False
        xs = self._exchange_for(q.shape[0], k) if self.exchange in ("p2p", "copy") else None
        by_copy = xs is not None and self.exchange == "copy"
        if by_copy:                                       # the copy engines may still be reading this slot's block
            check(_lib.load().pilott_xchg_wait_buffer(xs[0], slot % 8, self.local._stream()))
        large = k > _lib.MAX_FUSED_K                      # local top-k through the threshold-collect path (exact, synchronous)
        if xs is not None and not resolve and not large and not by_copy:
            # fused: finish_kernel stores the result rows into every rank's buffer as it produces them
            step = xs[1]
            xs[1] += 1
            self.local.search_publish_into(xs[0], step, q, k, b["idx"], b["sc"], b["sc64"], b["inexact"], b["flags"], **flt)
            b["begun"].record(torch.cuda.current_stream(self.device))
            return b, ("p2p", xs[0], step, q.shape[0], k)
        if large:
            self.local.search_large(q, k, out=(b["idx"], b["sc"], b["sc64"]), **flt)
            b["inexact"].zero_()
        else:
            self.local.search_into(q, k, b["idx"], b["sc"], b["sc64"], b["inexact"], b["flags"], **flt)
        if not large and resolve and int(b["inexact"].item()):
            self.local.refine(q, k, b["flags"].nonzero(), (b["idx"], b["sc"], b["sc64"]), **flt)
            b["inexact"].zero_()
        work = None
        if xs is not None:                               # the block goes out by copy engine, or (repaired rows) by the publish kernel
            step = xs[1]
            xs[1] += 1
            if by_copy:
                check(_lib.load().pilott_xchg_publish_copy(xs[0], ptr(b["mine"]), b["mine"].numel() * 8, step, slot % 8,
                                                           self.local._stream()))
                # (no event of our own on the search stream: the merge stream waits for this slot's copies
                #  instead, which are ordered behind the search; one record fewer between finish and prepare)
                return b, ("copy", xs[0], step, q.shape[0], k, slot % 8)
            check(_lib.load().pilott_xchg_publish(xs[0], ptr(b["mine"]), b["mine"].numel() * 8, step, self.local._stream()))
            b["begun"].record(torch.cuda.current_stream(self.device))
            return b, ("p2p", xs[0], step, q.shape[0], k)
        if self.world > 1:
            work = dist.all_gather_into_tensor(b["gathered"], b["mine"], group=self.group, async_op=True)
            b["begun"].record(torch.cuda.current_stream(self.device))
        return b, work

    def search_end(self, ticket, wait: bool = True):
        """Merges the exchanged blocks of one search_begin.  The merge kernel runs on a side stream of
        the store: it is small enough to sit beside the scan of a later batch on the same SMs, so it
        costs the search stream nothing.  wait=True (default) makes the CURRENT stream wait for it, so
        the returned tensors can be used like any other; wait=False returns at once and leaves the
        ordering to the caller (`ticket[0]["merged"]` is the torch event to wait for) -- what a
        pipeline does that only hands the results to a copy stream."""
        b, work = ticket
        if self.world == 1:
            return b["idx"], b["sc"], b["sc64"], b["inexact"]
        main = torch.cuda.current_stream(self.device)
        if self._merge_stream is None:
            self._merge_stream = torch.cuda.Stream(device=self.device)
        if isinstance(work, tuple) and work[0] == "copy":
            # this rank's outgoing copies of the step are ordered behind its local search: waiting for them
            # orders the merge behind both (and a rank never parks its wait kernel before its own block is out)
            check(_lib.load().pilott_xchg_wait_buffer(work[1], work[5], C.c_void_p(self._merge_stream.cuda_stream)))
        else:
            self._merge_stream.wait_event(b["begun"])    # recorded by search_begin behind its local search / publish
        with torch.cuda.stream(self._merge_stream):
            if isinstance(work, tuple):                  # peer-memory exchange: the block is already on its way
                _, x, step, nq, k = work[:5]
                check(_lib.load().pilott_xchg_merge(x, step, nq, k, ptr(b["m_idx"]), ptr(b["m_sc"]), ptr(b["m_sc64"]),
                                                    self.local._stream()))
            else:
                work.wait()
                topk_merge_packed(b["gathered"], out=(b["m_idx"], b["m_sc"], b["m_sc64"]))
            b["merged"].record(self._merge_stream)
        b["merge_pending"] = True
        if wait:
            main.wait_event(b["merged"])
            b["merge_pending"] = False
        return b["m_idx"], b["m_sc"], b["m_sc64"], b["inexact"]

    def check_exchange(self) -> None:
        """Synchronises and raises if a p2p merge ever gave up waiting for a peer."""
        lib = _lib.load()
        for x, _ in self._xchg.values():
            check(lib.pilott_xchg_check(x))

    def search(self, q: torch.Tensor, k: int, **flt):
        """q: CUDA fp32 [Q, dim], same on every rank.  Returns (idx, score, score64) on every rank,
        plus the local inexact counter (0: flagged queries are repaired locally before the exchange)."""
        return self.search_end(self.search_begin(q, k, 0, True, **flt))

    def close(self):
        lib = _lib.load()
        for x, _ in self._xchg.values():
            lib.pilott_xchg_destroy(x)
        self._xchg = {}
        self.local.close()
