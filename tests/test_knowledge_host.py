"""CPU tests of the KnowledgeManager drop-in: cache key, LRU + TTL behaviour, retry / timeout,
error contract -- the reference's rules (pilott/knowledge/knowledge_manager.py)."""
import asyncio
import json

import pytest

from pilottai_b200.knowledge import KnowledgeManager, KnowledgeSource


class CountingSource(KnowledgeSource):
    type: str = "api"
    calls: int = 0
    fail_first: int = 0

    async def query(self, query: str):
        self.calls += 1
        if self.calls <= self.fail_first:
            raise RuntimeError("transient")
        return {"q": query, "n": self.calls}


def test_cache_key_and_limits():
    km = KnowledgeManager(cache_size=5, cache_ttl=1)
    assert km.MAX_CACHE_SIZE == 100 and km.DEFAULT_CACHE_TTL == 60          # floors :36-37
    k = km._generate_cache_key("q", ["b", "a"])
    assert json.loads(k) == {"query": "q", "sources": ["a", "b"]}
    assert k == km._generate_cache_key("q", ["a", "b"])


def test_query_contract_cache_hit_and_refresh():
    async def go():
        km = KnowledgeManager()
        src = CountingSource(name="api", connection={})
        assert await km.add_source(src) is True
        assert await km.add_source(src) is False                            # duplicate name -> False (:66,:75-77)
        with pytest.raises(ValueError, match="Query cannot be empty"):
            await km.query_knowledge("", ["api"])
        with pytest.raises(ValueError, match="Source types cannot be empty"):
            await km.query_knowledge("x", [])
        a = await km.query_knowledge("x", ["api", "missing"])
        b = await km.query_knowledge("x", ["api", "missing"])
        assert a == b == [{"q": "x", "n": 1}] and src.calls == 1            # second call is a cache hit
        c = await km.query_knowledge("x", ["api", "missing"], force_refresh=True)
        assert c == [{"q": "x", "n": 2}]
        assert await km.query_knowledge("y", ["missing"]) == []             # nothing cached for empty results
        assert km.get_cache_stats()["size"] == 1 and km.get_source_stats()["api"]["access_count"] == 2
        await km.invalidate_cache(source_name="api")
        assert km.get_cache_stats()["size"] == 0
    asyncio.run(go())


def test_lru_eviction_ttl_and_retry(monkeypatch):
    async def go():
        km = KnowledgeManager()
        km.MAX_CACHE_SIZE = 3
        for i in range(5):
            await km._add_to_cache(f"k{i}", i)
        assert list(km.cache) == ["k2", "k3", "k4"]
        assert await km._get_from_cache("k2") == 2 and list(km.cache)[-1] == "k2"      # LRU touch
        km.cache["k3"].ttl = 0
        assert await km._get_from_cache("k3") is None and "k3" not in km.cache         # expired -> dropped
        slept = []

        async def fake_sleep(s):
            slept.append(s)
        monkeypatch.setattr(asyncio, "sleep", fake_sleep)
        src = CountingSource(name="flaky", connection={}, fail_first=2, retry_delay=7)
        await km.add_source(src)
        assert await km.query_knowledge("z", ["flaky"]) == [{"q": "z", "n": 3}]
        assert slept == [7, 7] and src.error_count == 2
    asyncio.run(go())


def test_scripted_session_matches_the_reference_log(golden_dir):
    """tests/golden/knowledge_session.json is the log of oracle/knowledge_script.py run against the
    UNMODIFIED reference KnowledgeManager (oracle/gen_golden_knowledge.py); the drop-in must produce
    the same log: return values, cache keys in LRU order, per-source counters, error strings."""
    import os
    from oracle.knowledge_script import session
    with open(os.path.join(golden_dir, "knowledge_session.json")) as f:
        want = json.load(f)
    got = json.loads(json.dumps(asyncio.run(session(KnowledgeManager, KnowledgeSource))))
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g == w, (g, w)
