"""KnowledgeManager drop-in (pilott/knowledge/knowledge_manager.py) with the GPU semantic
memory registered as a knowledge source.

The reference's manager is host-side control flow -- an LRU + TTL result cache in front of
pluggable async sources, with per-source retry and timeout -- and has no arithmetic; it stays
Python here.  Same class and method names, same limits
(`MAX_CACHE_SIZE >= 100`, `DEFAULT_CACHE_TTL >= 60 s`, 30 s query timeout), same error
behaviour (`ValueError` on empty query / empty source list, `[]` on timeout or failure).
What is new is `SemanticMemorySource`: a source whose `query()` is one batched top-k search on
the device, so repeated `query_knowledge` calls become cache hits and misses become one
`pilott_mem_search_host` call.  The reference's `KnowledgeSource` model has no `connect` /
`query` methods although the manager calls them (knowledge_manager.py:16-26 vs :133,:150;
those live on pilott/tools/knowledge.py:14-47): here the model carries them.
"""
from __future__ import annotations

import asyncio
import json
import logging
from collections import OrderedDict
from datetime import datetime
from typing import Any, Dict, List, Optional

from pydantic import BaseModel, ConfigDict, Field


class CacheEntry(BaseModel):
    value: Any
    timestamp: datetime
    ttl: int
    access_count: int = 0
    last_access: datetime = Field(default_factory=datetime.now)


class KnowledgeSource(BaseModel):
    model_config = ConfigDict(arbitrary_types_allowed=True, extra="allow")
    name: str
    type: str
    connection: Dict[str, Any] = Field(default_factory=dict)
    last_access: datetime = Field(default_factory=datetime.now)
    access_count: int = 0
    error_count: int = 0
    is_connected: bool = False
    max_retries: int = 3
    retry_delay: int = 5
    timeout: int = 30

    async def connect(self) -> bool:                       # tools/knowledge.py:14-28
        return self.type in ("database", "api", "file")

    async def query(self, query: str) -> Any:              # tools/knowledge.py:30-47
        return {} if self.type in ("database", "api", "file") else None


class SemanticMemorySource(KnowledgeSource):
    """The device-resident semantic store as a knowledge source."""
    type: str = "semantic_memory"
    limit: int = 5
    min_priority: int = 0

    def __init__(self, memory, name: str = "semantic_memory", **kw):
        super().__init__(name=name, **kw)
        object.__setattr__(self, "_memory", memory)

    async def connect(self) -> bool:
        return self._memory is not None

    async def query(self, query: str) -> Any:
        items = await self._memory.semantic_search(query, min_priority=self.min_priority, limit=self.limit, order="score")
        return [{"text": it.text, "metadata": it.metadata, "tags": sorted(it.tags), "priority": it.priority} for it in items]


class _ResultCache(OrderedDict):
    """Query results by key, youngest at the end: a hit is moved to the young end, an insert evicts
    from the old end until there is room, every entry dies `ttl` seconds after it was stored."""

    @staticmethod
    def alive(entry: CacheEntry) -> bool:
        return (datetime.now() - entry.timestamp).total_seconds() < entry.ttl

    def lookup(self, key: str) -> Optional[Any]:
        entry = self.get(key)
        if entry is None:
            return None
        if not self.alive(entry):
            del self[key]
            return None
        self.move_to_end(key)
        entry.access_count += 1
        entry.last_access = datetime.now()
        return entry.value

    def insert(self, key: str, value: Any, ttl: int, limit: int) -> None:
        while len(self) >= limit:
            self.popitem(last=False)
        self[key] = CacheEntry(value=value, timestamp=datetime.now(), ttl=ttl)
        self.move_to_end(key)

    def drop(self, doomed) -> None:
        for key in [k for k in self if doomed(k, self[k])]:
            del self[key]


class KnowledgeManager:
    QUERY_TIMEOUT_S = 30            # whole query_knowledge call (:96)
    CLEANUP_EVERY_S = 3600          # background sweep (:226)

    def __init__(self, cache_size: int = 1000, cache_ttl: int = 3600):
        self.MAX_CACHE_SIZE = max(100, cache_size)               # floors (:36-37)
        self.DEFAULT_CACHE_TTL = max(60, cache_ttl)
        self.cache = _ResultCache()
        self.cache_lock = asyncio.Lock()
        self.sources: Dict[str, KnowledgeSource] = {}
        self.source_locks: Dict[str, asyncio.Lock] = {}
        self.last_updated: Dict[str, datetime] = {}
        self.logger = logging.getLogger("KnowledgeManager")
        self._cleanup_task = None

    # ---- lifecycle ------------------------------------------------------------------------------
    async def start(self):
        self._cleanup_task = asyncio.create_task(self._periodic_cleanup())

    async def stop(self):
        task = self._cleanup_task
        if task:
            task.cancel()
            try:
                await task
            except asyncio.CancelledError:
                pass

    async def _periodic_cleanup(self):
        while True:
            try:
                await asyncio.sleep(self.CLEANUP_EVERY_S)
                await self.cleanup()
            except asyncio.CancelledError:
                break
            except Exception as e:                               # pragma: no cover
                self.logger.error(f"Periodic cleanup error: {str(e)}")

    async def cleanup(self):                                     # :226-249
        """Expired results go; a source that failed more often than it may retry is re-tested and,
        if it answers, starts over with a clean error count."""
        async with self.cache_lock:
            self.cache.drop(lambda _k, entry: not self._is_cache_entry_valid(entry))
            for src in self.sources.values():
                if src.error_count <= src.max_retries:
                    continue
                src.is_connected = False
                if await self._test_connection(src):
                    src.error_count, src.is_connected = 0, True

    # ---- sources (:62-77, :120-155) ---------------------------------------------------------------
    async def add_source(self, source: KnowledgeSource) -> bool:
        name = source.name
        try:
            if name in self.sources:
                raise ValueError(f"Source {name} already exists")
            self.sources[name], self.source_locks[name] = source, asyncio.Lock()
            self.last_updated[name] = datetime.now()
            source.is_connected = await self._test_connection(source)
            return source.is_connected
        except Exception as e:
            self.logger.error(f"Error adding source {name}: {str(e)}")
            return False

    async def _test_connection(self, source: KnowledgeSource) -> bool:
        try:
            async with asyncio.timeout(source.timeout):
                return await source.connect()             # the flag itself is set by add_source / cleanup only
        except Exception as e:
            self.logger.error(f"Connection test failed for {source.name}: {str(e)}")
            return False

    async def _query_source_with_retry(self, source: KnowledgeSource, query: str) -> Optional[Any]:
        """Up to max_retries attempts; every failed one counts as an error, a pause of retry_delay
        separates attempts that failed with anything but a timeout."""
        last = source.max_retries - 1
        for attempt in range(source.max_retries):
            pause = False
            try:
                if not (source.is_connected or await self._test_connection(source)):
                    raise ConnectionError(f"Source {source.name} is not connected")
                async with asyncio.timeout(source.timeout):
                    answer = await source.query(query)
                    source.access_count += 1
                    source.last_access = datetime.now()
                    return answer
            except asyncio.TimeoutError:
                source.error_count += 1
            except Exception as e:
                self.logger.error(f"Query failed for source {source.name}, attempt {attempt + 1}: {str(e)}")
                source.error_count += 1
                pause = attempt < last
            if pause:
                await asyncio.sleep(source.retry_delay)
        return None

    # ---- the query (:79-118) ----------------------------------------------------------------------
    async def query_knowledge(self, query: str, source_types: List[str], ttl: Optional[int] = None,
                              force_refresh: bool = False) -> List[Any]:
        if not query:
            raise ValueError("Query cannot be empty")
        if not source_types:
            raise ValueError("Source types cannot be empty")
        key = self._generate_cache_key(query, source_types)
        try:
            if not force_refresh:
                cached = await self._get_from_cache(key)
                if cached is not None:
                    return cached
            async with asyncio.timeout(self.QUERY_TIMEOUT_S):
                answers = []
                for name in source_types:                        # unknown names are skipped
                    if name in self.sources:
                        answer = await self._ask(name, query)
                        if answer is not None:
                            answers.append(answer)
                if answers:
                    await self._add_to_cache(key, answers, ttl)
                return answers
        except asyncio.TimeoutError:
            self.logger.error("Query timed out")
        except Exception as e:
            self.logger.error(f"Error during knowledge query: {str(e)}")
        return []

    async def _ask(self, name: str, query: str) -> Optional[Any]:
        source = self.sources[name]
        try:
            async with self.source_locks[name]:
                return await self._query_source_with_retry(source, query)
        except Exception as e:
            self.logger.error(f"Error querying source {name}: {str(e)}")
            source.error_count += 1
            return None

    # ---- result cache (:157-221) --------------------------------------------------------------------
    def _generate_cache_key(self, query: str, source_types: List[str]) -> str:
        return json.dumps({"query": query, "sources": sorted(source_types)}, sort_keys=True)

    def _is_cache_entry_valid(self, entry: CacheEntry) -> bool:
        return _ResultCache.alive(entry)

    async def _get_from_cache(self, key: str) -> Optional[Any]:
        async with self.cache_lock:
            return self.cache.lookup(key)

    async def _add_to_cache(self, key: str, value: Any, ttl: Optional[int] = None):
        async with self.cache_lock:
            self.cache.insert(key, value, ttl or self.DEFAULT_CACHE_TTL, self.MAX_CACHE_SIZE)

    async def invalidate_cache(self, source_name: Optional[str] = None, pattern: Optional[str] = None):
        """By source (every result that source contributed to), else by substring of the key, else all."""
        async with self.cache_lock:
            if source_name:
                self.cache.drop(lambda key, _e: source_name in json.loads(key)["sources"])
            elif pattern:
                self.cache.drop(lambda key, _e: pattern in key)
            else:
                self.cache.clear()

    # ---- introspection ------------------------------------------------------------------------------
    def get_source_stats(self) -> Dict[str, Dict[str, Any]]:
        stats = {}
        for name, src in self.sources.items():
            stats[name] = {"access_count": src.access_count, "error_count": src.error_count,
                           "last_access": src.last_access.isoformat(), "is_connected": src.is_connected}
        return stats

    def get_cache_stats(self) -> Dict[str, Any]:
        n = len(self.cache)
        return {"size": n, "max_size": self.MAX_CACHE_SIZE,
                "hit_ratio": sum(e.access_count for e in self.cache.values()) / max(1, n)}
