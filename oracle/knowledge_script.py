"""One scripted session with a KnowledgeManager -- the same coroutine is run against the
reference's class (oracle/gen_golden_knowledge.py, which writes the fixture) and against the
drop-in (tests/test_knowledge_host.py, which compares).  TEST INFRASTRUCTURE ONLY.

The script touches everything the reference defines for this step of the path
(pilott/knowledge/knowledge_manager.py): duplicate sources (:64-77), empty-argument errors
(:88-91), cache miss / hit / force_refresh (:95-113), unknown source types skipped (:101),
retries with error counting (:122-149), a source whose connect() fails, LRU eviction at
MAX_CACHE_SIZE (:182-184), cache invalidation by source / pattern / everything (:203-221),
cleanup() re-testing unhealthy sources (:233-249), the stats dictionaries (:251-...)."""
import json


def make_sources(base_cls):
    """Scripted sources on top of either implementation's KnowledgeSource model."""

    class Scripted(base_cls):
        fail_first: int = 0            # this many query() calls raise before one succeeds
        connect_ok: bool = True
        calls: int = 0

        async def connect(self) -> bool:
            return self.connect_ok

        async def query(self, query: str):
            self.calls += 1
            if self.calls <= self.fail_first:
                raise RuntimeError("transient")
            return {"source": self.name, "q": query, "n": self.calls}

    return Scripted


async def session(km_cls, source_cls):
    log = []
    km = km_cls(cache_size=5, cache_ttl=1)
    km.logger.disabled = True
    Scripted = make_sources(source_cls)

    def snap(tag):
        stats = {n: {k: v for k, v in s.items() if k != "last_access"} for n, s in km.get_source_stats().items()}
        log.append({"at": tag, "cache_keys": list(km.cache.keys()), "stats": stats})

    log.append({"limits": [km.MAX_CACHE_SIZE, km.DEFAULT_CACHE_TTL]})
    a = Scripted(name="a", type="api", connection={}, retry_delay=0)
    b = Scripted(name="b", type="api", connection={}, retry_delay=0, fail_first=2)
    c = Scripted(name="c", type="api", connection={}, retry_delay=0, fail_first=99, max_retries=2)
    d = Scripted(name="d", type="api", connection={}, retry_delay=0, connect_ok=False)
    for s in (a, b, c, d, a):
        log.append({"add": s.name, "ret": await km.add_source(s)})
    for args in (("", ["a"]), ("x", [])):
        try:
            await km.query_knowledge(*args)
            log.append({"error": None})
        except ValueError as e:
            log.append({"error": str(e)})
    log.append({"q1": await km.query_knowledge("alpha", ["a"])}); snap("after q1")
    log.append({"q1_again": await km.query_knowledge("alpha", ["a"])}); snap("after cached q1")
    log.append({"q1_refresh": await km.query_knowledge("alpha", ["a"], force_refresh=True)})
    log.append({"q2": await km.query_knowledge("beta", ["b", "a", "nobody"])}); snap("after q2 (b needed retries)")
    log.append({"q3": await km.query_knowledge("gamma", ["c"])}); snap("after q3 (c never answers)")
    log.append({"q4": await km.query_knowledge("delta", ["d", "a"])}); snap("after q4 (d cannot connect)")
    log.append({"key": km._generate_cache_key("zeta", ["b", "a"])})
    for i in range(103):                                   # LRU: MAX_CACHE_SIZE floors at 100
        await km.query_knowledge(f"fill {i}", ["a"])
    log.append({"cache_len": len(km.cache), "oldest": next(iter(km.cache)), "newest": next(reversed(km.cache))})
    await km.query_knowledge("fill 10", ["a"])             # a hit moves the entry to the young end
    log.append({"newest_after_hit": next(reversed(km.cache))})
    await km.invalidate_cache(pattern="fill 2")
    log.append({"after_pattern": len(km.cache)})
    await km.query_knowledge("both", ["a", "b"])
    await km.invalidate_cache(source_name="b")
    log.append({"after_source_b": len(km.cache), "has_both": any(json.loads(k)["query"] == "both" for k in km.cache)})
    await km.invalidate_cache()
    log.append({"after_all": len(km.cache)})
    await km.cleanup(); snap("after cleanup")
    log.append({"cache_stats_keys": sorted(km.get_cache_stats().keys()) if hasattr(km, "get_cache_stats") else None})
    return log
