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


class KnowledgeManager:
    def __init__(self, cache_size: int = 1000, cache_ttl: int = 3600):
        self.sources: Dict[str, KnowledgeSource] = {}
        self.cache: "OrderedDict[str, CacheEntry]" = OrderedDict()
        self.last_updated: Dict[str, datetime] = {}
        self.source_locks: Dict[str, asyncio.Lock] = {}
        self.cache_lock = asyncio.Lock()
        self.MAX_CACHE_SIZE = max(100, cache_size)
        self.DEFAULT_CACHE_TTL = max(60, cache_ttl)
        self.logger = logging.getLogger("KnowledgeManager")
        self._cleanup_task = None

    async def start(self):
        self._cleanup_task = asyncio.create_task(self._periodic_cleanup())

    async def stop(self):
        if self._cleanup_task:
            self._cleanup_task.cancel()
            try:
                await self._cleanup_task
            except asyncio.CancelledError:
                pass

    async def add_source(self, source: KnowledgeSource) -> bool:            # :62-77
        try:
            if source.name in self.sources:
                raise ValueError(f"Source {source.name} already exists")
            self.source_locks[source.name] = asyncio.Lock()
            self.sources[source.name] = source
            self.last_updated[source.name] = datetime.now()
            source.is_connected = await self._test_connection(source)
            return source.is_connected
        except Exception as e:
            self.logger.error(f"Error adding source {source.name}: {str(e)}")
            return False

    async def query_knowledge(self, query: str, source_types: List[str], ttl: Optional[int] = None,
                              force_refresh: bool = False) -> List[Any]:     # :79-118
        if not query:
            raise ValueError("Query cannot be empty")
        if not source_types:
            raise ValueError("Source types cannot be empty")
        key = self._generate_cache_key(query, source_types)
        results: List[Any] = []
        try:
            if not force_refresh:
                hit = await self._get_from_cache(key)
                if hit is not None:
                    return hit
            async with asyncio.timeout(30):
                for st in source_types:
                    src = self.sources.get(st)
                    if src is None:
                        continue
                    try:
                        async with self.source_locks[st]:
                            r = await self._query_source_with_retry(src, query)
                            if r is not None:
                                results.append(r)
                    except Exception as e:
                        self.logger.error(f"Error querying source {st}: {str(e)}")
                        src.error_count += 1
                if results:
                    await self._add_to_cache(key, results, ttl)
                return results
        except asyncio.TimeoutError:
            self.logger.error("Query timed out")
            return []
        except Exception as e:
            self.logger.error(f"Error during knowledge query: {str(e)}")
            return []

    async def _query_source_with_retry(self, source: KnowledgeSource, query: str) -> Optional[Any]:   # :120-147
        for attempt in range(source.max_retries):
            try:
                if not source.is_connected and not await self._test_connection(source):
                    raise ConnectionError(f"Source {source.name} is not connected")
                async with asyncio.timeout(source.timeout):
                    result = await source.query(query)
                    source.access_count += 1
                    source.last_access = datetime.now()
                    return result
            except asyncio.TimeoutError:
                source.error_count += 1
            except Exception as e:
                self.logger.error(f"Query failed for source {source.name}, attempt {attempt + 1}: {str(e)}")
                source.error_count += 1
                if attempt < source.max_retries - 1:
                    await asyncio.sleep(source.retry_delay)
        return None

    async def _test_connection(self, source: KnowledgeSource) -> bool:
        try:
            async with asyncio.timeout(source.timeout):
                ok = bool(await source.connect())
                source.is_connected = ok or source.is_connected
                return ok
        except Exception as e:
            self.logger.error(f"Connection test failed for {source.name}: {str(e)}")
            return False

    # ---- LRU + TTL cache (:157-197) -----------------------------------------------------------
    def _is_cache_entry_valid(self, entry: CacheEntry) -> bool:
        return (datetime.now() - entry.timestamp).total_seconds() < entry.ttl

    def _generate_cache_key(self, query: str, source_types: List[str]) -> str:
        return json.dumps({"query": query, "sources": sorted(source_types)}, sort_keys=True)

    async def _get_from_cache(self, key: str) -> Optional[Any]:
        async with self.cache_lock:
            entry = self.cache.get(key)
            if entry is None:
                return None
            if not self._is_cache_entry_valid(entry):
                del self.cache[key]
                return None
            self.cache.move_to_end(key)
            entry.access_count += 1
            entry.last_access = datetime.now()
            return entry.value

    async def _add_to_cache(self, key: str, value: Any, ttl: Optional[int] = None):
        async with self.cache_lock:
            while len(self.cache) >= self.MAX_CACHE_SIZE:
                self.cache.popitem(last=False)
            self.cache[key] = CacheEntry(value=value, timestamp=datetime.now(), ttl=ttl or self.DEFAULT_CACHE_TTL)
            self.cache.move_to_end(key)

    async def invalidate_cache(self, source_name: Optional[str] = None, pattern: Optional[str] = None):
        async with self.cache_lock:
            if source_name:
                doomed = [k for k in self.cache if source_name in json.loads(k)["sources"]]
            elif pattern:
                doomed = [k for k in self.cache if pattern in k]
            else:
                self.cache.clear()
                return
            for k in doomed:
                del self.cache[k]

    async def _periodic_cleanup(self):
        while True:
            try:
                await asyncio.sleep(3600)
                await self.cleanup()
            except asyncio.CancelledError:
                break
            except Exception as e:                               # pragma: no cover
                self.logger.error(f"Periodic cleanup error: {str(e)}")

    async def cleanup(self):                                     # :226-249
        async with self.cache_lock:
            for k in [k for k, v in self.cache.items() if not self._is_cache_entry_valid(v)]:
                del self.cache[k]
            for source in self.sources.values():
                if source.error_count > source.max_retries:
                    source.is_connected = False
                    if await self._test_connection(source):
                        source.error_count = 0
                        source.is_connected = True

    def get_source_stats(self) -> Dict[str, Dict[str, Any]]:
        return {n: {"access_count": s.access_count, "error_count": s.error_count,
                    "last_access": s.last_access.isoformat(), "is_connected": s.is_connected}
                for n, s in self.sources.items()}

    def get_cache_stats(self) -> Dict[str, Any]:
        return {"size": len(self.cache), "max_size": self.MAX_CACHE_SIZE,
                "hit_ratio": sum(e.access_count for e in self.cache.values()) / max(1, len(self.cache))}
