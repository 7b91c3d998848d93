"""EnhancedMemory -- drop-in for pilott.memory.enhanced_memory.EnhancedMemory whose
semantic store lives on a B200.

Same class name, constructor, coroutines, argument meaning and error behaviour as the
reference (pilott/memory/enhanced_memory.py); written from scratch.  What changes:

* ``semantic_search`` ranks by cosine similarity of embeddings on the GPU instead of the
  reference's case-insensitive substring test (:110), and returns the ``limit`` BEST
  matches rather than the first ``limit`` found (:111-113).
* keyword-only extras: ``embedding=`` / ``query_embedding=`` / ``order=`` and the batched
  ``store_semantic_batch`` / ``semantic_search_batch``.
* reference defects deliberately NOT reproduced: stale tag/priority
  indexes after the deque overflows (:86-87 vs :133-144), ``max()`` of an empty index on
  an empty store (:124), first-found early break.
* concurrent ``semantic_search`` calls are COALESCED: the reference's callers issue one search per
  coroutine (:93-99); here each call parks its query and a drain task sends everything that is
  waiting with the same filters to the device as ONE batch (one kernel sequence for up to
  ``coalesce_max`` queries), then hands every caller its own rows.  Signatures are unchanged and so
  are the results: a call gets exactly what it would have got alone.
* tags are unbounded, as in the reference (its tag index is a dict of sets, :133-138): the device
  row carries a 64-bit tag mask, the first 63 distinct tags get a bit each, every further tag shares
  the last bit, and a search that names such a tag is filtered on that bit by the device and on
  the exact tag sets by the host (asking the device for more rows until ``limit`` of them survive).

Host side keeps the ``MemoryItem`` objects (text, metadata) in insertion order; the device
keeps one row per item (bf16 vector, inverse norm, priority, tag mask, expiry).  The two are
joined by the insertion sequence number the search returns.
"""
from __future__ import annotations

import asyncio
from collections import deque
from datetime import datetime, timedelta
from typing import Any, Dict, Iterable, List, Optional, Sequence, Set, Tuple

import numpy as np
from pydantic import BaseModel, ConfigDict, Field

from ._lib import MAX_TAGS
from .embedder import HashEmbedder
from .store import SemanticStore

_EPOCH = datetime(1970, 1, 1)


def _to_us(t: datetime) -> int:
    """Naive local datetime -> integer microseconds (monotone, exact)."""
    d = t - _EPOCH
    return (d.days * 86400 + d.seconds) * 1_000_000 + d.microseconds


class MemoryItem(BaseModel):
    """One stored memory (field-compatible with the reference's MemoryItem, :9-21)."""
    model_config = ConfigDict(arbitrary_types_allowed=True)
    text: str
    metadata: Dict[str, Any] = Field(default_factory=dict)
    timestamp: datetime = Field(default_factory=datetime.now)
    tags: Set[str] = Field(default_factory=set)
    priority: int = Field(ge=0, default=0)
    expires_at: Optional[datetime] = None
    version: int = 1

    def is_expired(self) -> bool:
        return bool(self.expires_at and datetime.now() > self.expires_at)


class EnhancedMemory:
    def __init__(self, max_size: int = 10000, cleanup_interval: int = 3600, *, dim: int = 768,
                 device: int = 0, embedder=None, coalesce_max: int = 256, coalesce_window_s: float = 0.0):
        self.coalesce_max = int(coalesce_max)            # queries per device batch of coalesced single searches
        self.coalesce_window_s = float(coalesce_window_s)  # extra wait for company before a batch leaves (0: one loop turn)
        self._waiting: List[tuple] = []                  # parked single searches: (vector, k, mask, min_priority, future)
        self._drain_task: Optional[asyncio.Task] = None
        self.coalesced_batches = 0                       # device batches the drain task has sent (bookkeeping for tests)
        self.max_size = int(max_size)
        self.dim = int(dim)
        # MemoryItems by insertion sequence number: item of sequence s sits at _ring[s % max_size] (the
        # reference keeps a deque(maxlen) and only ever iterates it, :27; here every search result is an
        # index lookup, and a deque of a million items is O(n) to index).  _total mirrors the device's
        # count of rows ever appended, _count the rows held.
        self._ring: List[Optional[MemoryItem]] = []
        self._total = 0
        self._count = 0
        self._store = SemanticStore(self.dim, self.max_size, device)
        self._embed = embedder or HashEmbedder(self.dim)
        self._tag_bits: Dict[str, int] = {}
        self._task_history: Dict[str, deque] = {}
        self._agent_interactions: Dict[str, Dict[str, Any]] = {}
        self._pattern_store: Dict[str, Dict[str, Any]] = {}
        self.max_task_history = 1000
        self.cleanup_interval = cleanup_interval
        self.last_cleanup = datetime.now()
        self._semantic_lock = asyncio.Lock()
        self._task_lock = asyncio.Lock()
        self._interaction_lock = asyncio.Lock()
        self._pattern_lock = asyncio.Lock()
        self._cleanup_task = None

    # ---- lifecycle (reference :49-58) ---------------------------------------------------
    async def start(self):
        self._cleanup_task = asyncio.create_task(self._periodic_cleanup())

    async def stop(self):
        if self._cleanup_task:
            self._cleanup_task.cancel()
            try:
                await self._cleanup_task
            except asyncio.CancelledError:
                pass

    def close(self):
        self._store.close()

    # ---- tags -> bits ---------------------------------------------------------------------
    _OVERFLOW_BIT = MAX_TAGS - 1                          # shared by every tag after the first 63

    def _mask_for_store(self, tags: Iterable[str]) -> int:
        m = 0
        for t in tags:
            b = self._tag_bits.get(t)
            if b is None:
                b = self._tag_bits[t] = min(len(self._tag_bits), self._OVERFLOW_BIT)
            m |= 1 << b
        return m

    def _mask_for_query(self, tags: Optional[Iterable[str]]) -> Optional[int]:
        """None when some requested tag was never stored: no row can match (:126-130)."""
        m = 0
        for t in tags or ():
            b = self._tag_bits.get(t)
            if b is None:
                return None
            m |= 1 << b
        return m

    def _needs_host_filter(self, tags: Optional[Iterable[str]]) -> bool:
        """True when a requested tag shares the overflow bit: the device filter is then only necessary."""
        return any(self._tag_bits.get(t) == self._OVERFLOW_BIT and len(self._tag_bits) > MAX_TAGS - 1 for t in tags or ())

    # ---- store (reference :60-91) ---------------------------------------------------------
    async def store_semantic(self, text: str, metadata: Optional[Dict[str, Any]] = None,
                             tags: Optional[Set[str]] = None, priority: int = 0,
                             ttl: Optional[int] = None, *, embedding=None) -> None:
        if not text:
            raise ValueError("Text cannot be empty")
        try:
            await self.store_semantic_batch([text], None if embedding is None else np.asarray(embedding)[None],
                                            metadata=[metadata], tags=[tags], priority=[priority], ttl=[ttl])
        except Exception as e:
            raise ValueError(f"Failed to store content: {str(e)}")

    async def store_semantic_batch(self, texts: Sequence[str], embeddings=None, *, metadata=None, tags=None,
                                   priority=None, ttl=None) -> int:
        """Append len(texts) items in one device call.  Returns the first sequence number."""
        n = len(texts)
        if any(not t for t in texts):
            raise ValueError("Text cannot be empty")
        now = datetime.now()
        items = []
        for i in range(n):
            t = None if ttl is None else ttl[i]
            items.append(MemoryItem(
                text=texts[i], metadata=(metadata[i] if metadata else None) or {},
                tags=(tags[i] if tags else None) or set(),
                priority=(priority[i] if priority is not None else 0),
                expires_at=(now + timedelta(seconds=t)) if t is not None else None))
        if embeddings is None:
            embeddings = self._embed.batch(list(texts))
        emb = np.ascontiguousarray(embeddings, dtype=np.float32).reshape(n, self.dim)
        prio = np.array([it.priority for it in items], dtype=np.int32)
        exp = np.array([_to_us(it.expires_at) if it.expires_at else 0 for it in items], dtype=np.int64)
        async with self._semantic_lock:
            mask = np.array([self._mask_for_store(it.tags) for it in items], dtype=np.uint64)
            first = -1
            for lo in range(0, n, self.max_size):          # a device batch may not exceed the ring
                hi = min(n, lo + self.max_size)
                s = await asyncio.to_thread(self._store.append, emb[lo:hi], prio[lo:hi], mask[lo:hi], exp[lo:hi])
                first = s if first < 0 else first
                for it in items[lo:hi]:
                    slot = self._total % self.max_size
                    if slot < len(self._ring):
                        self._ring[slot] = it
                    else:
                        self._ring.append(it)
                    self._total += 1
                self._count = min(self._count + (hi - lo), self.max_size)
        return first

    # ---- search (reference :93-131) -------------------------------------------------------
    def _item_for_seq(self, seq: int) -> MemoryItem:
        if not (self._total - self._count <= seq < self._total):
            raise IndexError(f"sequence number {seq} is not held (rows {self._total - self._count}..{self._total - 1})")
        return self._ring[seq % self.max_size]

    @property
    def _items(self) -> List[MemoryItem]:
        """The held items, oldest first (what the reference's deque iterates)."""
        return [self._ring[s % self.max_size] for s in range(self._total - self._count, self._total)]

    async def _device_search(self, q: np.ndarray, k: int, req: int, min_priority: int):
        async with self._semantic_lock:
            now_us = _to_us(datetime.now())
            return await asyncio.to_thread(self._store.search, q, k, min_priority, req, now_us)

    async def semantic_search_batch(self, queries, k: int = 5, tags: Optional[Set[str]] = None,
                                    min_priority: int = 0) -> Tuple[np.ndarray, np.ndarray]:
        """queries [Q, dim] fp32 -> (insertion sequence numbers [Q,k] (-1 = none), scores [Q,k])."""
        q = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, self.dim)
        req = self._mask_for_query(tags)
        if req is None or k < 1:
            return (np.full((q.shape[0], max(k, 0)), -1, np.int64), np.full((q.shape[0], max(k, 0)), -np.inf, np.float32))
        if not self._needs_host_filter(tags):
            return await self._device_search(q, k, req, min_priority)
        # A requested tag shares the overflow bit: the device narrows to rows carrying SOME overflow tag,
        # the exact tag sets decide.  Ask for more rows until k survive (or the store is exhausted).
        want = set(tags)
        out_i = np.full((q.shape[0], k), -1, np.int64)
        out_s = np.full((q.shape[0], k), -np.inf, np.float32)
        todo = np.arange(q.shape[0])
        kk = max(4 * k, 64)
        while todo.size:
            kk = min(kk, max(self._count, 1))
            idx, sc = await self._device_search(q[todo], kk, req, min_priority)
            again = []
            for r, qi in enumerate(todo):
                keep = [j for j in range(kk) if idx[r, j] >= 0 and want <= self._item_for_seq(int(idx[r, j])).tags]
                exhausted = kk >= self._count or idx[r, kk - 1] < 0
                if len(keep) >= k or exhausted:
                    keep = keep[:k]
                    out_i[qi, :len(keep)] = idx[r, keep]
                    out_s[qi, :len(keep)] = sc[r, keep]
                else:
                    again.append(qi)
            todo = np.array(again, dtype=np.int64)
            kk *= 4
        return out_i, out_s

    # -- coalescing of single searches ------------------------------------------------------
    async def _search_one(self, vec: np.ndarray, k: int, tags: Optional[Set[str]], min_priority: int):
        """One query through the shared batch: parks it, makes sure the drain task runs, awaits its rows."""
        req = self._mask_for_query(tags)
        if req is None:
            return np.full(k, -1, np.int64), np.full(k, -np.inf, np.float32)
        if self._needs_host_filter(tags):                # rare path: exact tag sets are checked on the host
            idx, sc = await self.semantic_search_batch(vec[None], k, tags, min_priority)
            return idx[0], sc[0]
        fut = asyncio.get_running_loop().create_future()
        self._waiting.append((vec, k, req, min_priority, fut))
        if self._drain_task is None or self._drain_task.done():
            self._drain_task = asyncio.create_task(self._drain())
        return await fut

    async def _drain(self):
        """Sends whatever is parked, grouped by filter, `coalesce_max` queries at a time.  Callers that
        arrive while a batch is on the device are picked up by the next turn of the loop."""
        await asyncio.sleep(0)                            # one loop turn: every coroutine that is ready parks its query
        if self.coalesce_window_s > 0 and len(self._waiting) < self.coalesce_max:
            await asyncio.sleep(self.coalesce_window_s)
        while self._waiting:
            batch, self._waiting = self._waiting, []
            groups: Dict[tuple, list] = {}
            for w in batch:                               # same filters, same side of the fused-path limit
                groups.setdefault((w[2], w[3], w[1] > 32), []).append(w)
            for (req, min_priority, _), ws in groups.items():
                for lo in range(0, len(ws), self.coalesce_max):
                    part = ws[lo:lo + self.coalesce_max]
                    try:
                        k = max(w[1] for w in part)
                        idx, sc = await self._device_search(np.stack([w[0] for w in part]), k, req, min_priority)
                        self.coalesced_batches += 1
                        for r, w in enumerate(part):      # the top-k of a total order is a prefix of the top-k'
                            if not w[4].done():
                                w[4].set_result((idx[r, :w[1]], sc[r, :w[1]]))
                    except Exception as e:
                        for w in part:
                            if not w[4].done():
                                w[4].set_exception(e)

    async def semantic_search(self, query: str, tags: Optional[Set[str]] = None, min_priority: int = 0,
                              limit: int = 5, *, query_embedding=None, order: str = "priority") -> List[MemoryItem]:
        if not query:
            raise ValueError("Query cannot be empty")
        try:
            if limit < 1:
                return []
            q = self._embed(query) if query_embedding is None else np.asarray(query_embedding, np.float32)
            idx, _ = await self._search_one(np.ascontiguousarray(q, np.float32).reshape(self.dim), limit, tags, min_priority)
            found = [self._item_for_seq(int(s)) for s in idx if s >= 0]
            if order == "score":
                return found
            return sorted(found, key=lambda x: (-x.priority, x.timestamp))       # reference :114
        except Exception as e:
            raise ValueError(f"Search failed: {str(e)}")

    # ---- task history / interactions / patterns: host-only, same contracts (:146-246) ------
    async def store_task(self, task_id: str, task_data: Dict[str, Any]) -> None:
        if not task_id or not task_data:
            raise ValueError("Task ID and data required")
        async with self._task_lock:
            hist = self._task_history.setdefault(task_id, deque(maxlen=self.max_task_history))
            hist.append({"data": dict(task_data), "timestamp": datetime.now(), "version": len(hist) + 1})

    async def store_interaction(self, agent_id: str, interaction_type: str, data: Dict[str, Any]) -> None:
        if not agent_id or not interaction_type or not data:
            raise ValueError("Agent ID, interaction type, and data required")
        async with self._interaction_lock:
            log = self._agent_interactions.setdefault(agent_id, {})
            ts = datetime.now()
            log[ts.isoformat()] = {"type": interaction_type, "data": dict(data), "timestamp": ts,
                                   "version": len(log) + 1}

    async def store_pattern(self, name: str, data: Any, ttl: Optional[int] = None) -> None:
        if not name:
            raise ValueError("Pattern name required")
        async with self._pattern_lock:
            now = datetime.now()
            self._pattern_store[name] = {"data": data, "timestamp": now,
                                         "expires_at": now + timedelta(seconds=ttl) if ttl is not None else None}

    async def get_pattern(self, name: str) -> Optional[Any]:
        if not name:
            raise ValueError("Pattern name required")
        async with self._pattern_lock:
            p = self._pattern_store.get(name)
            if p is None:
                return None
            if p["expires_at"] and datetime.now() > p["expires_at"]:
                del self._pattern_store[name]
                return None
            return p["data"]

    async def get_recent_tasks(self, limit: int = 10, task_type: Optional[str] = None) -> List[Dict[str, Any]]:
        if limit < 1:
            raise ValueError("Limit must be positive")
        async with self._task_lock:
            rows = [e for h in self._task_history.values() for e in h
                    if not task_type or e["data"].get("type") == task_type]
            return sorted(rows, key=lambda x: x["timestamp"], reverse=True)[:limit]

    # ---- cleanup (reference :248-292) -----------------------------------------------------
    async def _periodic_cleanup(self):
        while True:
            try:
                await asyncio.sleep(self.cleanup_interval)
                await self.cleanup()
            except asyncio.CancelledError:
                break
            except Exception as e:                                  # pragma: no cover
                print(f"Periodic cleanup error: {str(e)}")

    async def cleanup(self):
        try:
            async with self._semantic_lock, self._pattern_lock:
                now = datetime.now()
                kept, keep = await asyncio.to_thread(self._store.compact, _to_us(now))
                # the device renumbers the survivors total - kept .. total - 1 in age order (pilott_mem_compact)
                survivors = [it for it, k in zip(self._items, keep) if k]
                self._count = len(survivors)
                for i, it in enumerate(survivors):
                    slot = (self._total - self._count + i) % self.max_size
                    if slot < len(self._ring):
                        self._ring[slot] = it
                    else:                                  # (cannot happen: the ring only grows to max_size)
                        self._ring.append(it)
                for name in [n for n, p in self._pattern_store.items() if p["expires_at"] and now > p["expires_at"]]:
                    del self._pattern_store[name]
            self.last_cleanup = datetime.now()
        except Exception as e:
            raise ValueError(f"Cleanup failed: {str(e)}")

    def clear(self) -> None:
        self._ring, self._total, self._count = [], 0, 0
        self._store.clear()
        self._tag_bits.clear()
        self._task_history.clear()
        self._agent_interactions.clear()
        self._pattern_store.clear()

    def __len__(self) -> int:
        return self._count
