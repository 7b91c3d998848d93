"""Host logic of the EnhancedMemory drop-in on CPU: the device store is replaced by the CPU oracle
(tests may use oracle/; the product never does), so what is tested is the coalescing of concurrent
single searches and the unbounded-tag handling around it.  The same scenarios run against the real
device store in tests/test_gpu_dropins.py."""
import asyncio

import numpy as np
import pytest

from oracle import retrieval_c
from oracle.retrieval_oracle import OracleStore, bf16_round_bits, exact_scores
import pilottai_b200.memory as memory_mod


class OracleBackedStore:
    """SemanticStore's surface over oracle.OracleStore (numpy in, numpy out)."""

    def __init__(self, dim, capacity, device=0):
        self.o = OracleStore(dim, capacity)
        self.searches = 0

    total = property(lambda self: self.o.total)
    count = property(lambda self: self.o.count)

    def append(self, x, prio=None, tagmask=None, expires_us=None):
        return self.o.append(x, prio, tagmask, expires_us)

    def search(self, q, k, min_priority=0, req_tagmask=0, now_us=0, want_score64=False):
        self.searches += 1
        idx, sc = retrieval_c.search(self.o, q, k, min_priority=min_priority, req_tagmask=req_tagmask, now_us=now_us, threads=2)
        return idx, sc.astype(np.float32)

    def clear(self):
        self.o = OracleStore(self.o.dim, self.o.capacity)

    def close(self):
        pass


@pytest.fixture()
def mem(monkeypatch):
    monkeypatch.setattr(memory_mod, "SemanticStore", OracleBackedStore)
    return memory_mod.EnhancedMemory(max_size=4096, dim=32)


def test_concurrent_single_searches_are_coalesced_and_exact(mem):
    async def run():
        rng = np.random.default_rng(0)
        texts = [f"note {i}" for i in range(600)]
        emb = rng.standard_normal((600, 32)).astype(np.float32)
        await mem.store_semantic_batch(texts, emb, tags=[{"even"} if i % 2 == 0 else {"odd"} for i in range(600)],
                                       priority=[i % 4 for i in range(600)])
        qs = rng.standard_normal((1000, 32)).astype(np.float32)
        kw = [dict(limit=1 + i % 7, tags=({"even"} if i % 3 == 0 else None), min_priority=i % 2) for i in range(1000)]
        before = mem._store.searches
        got = await asyncio.gather(*[mem.semantic_search("q", query_embedding=qs[i], order="score", **kw[i]) for i in range(1000)])
        batched = mem._store.searches - before
        assert batched <= 12, f"1000 concurrent searches took {batched} device calls"      # 4 filter groups x ceil(n/256)
        for i in range(0, 1000, 37):                                     # alone, each call gets the same rows
            alone = await mem.semantic_search("q", query_embedding=qs[i], order="score", **kw[i])
            assert [x.text for x in alone] == [x.text for x in got[i]]
        assert all(len(g) == kw[i]["limit"] for i, g in enumerate(got))
    asyncio.run(run())


def test_a_failing_batch_fails_every_waiter_with_the_reference_error(mem):
    async def run():
        await mem.store_semantic("x", embedding=np.ones(32, np.float32))
        def boom(*a, **k):
            raise RuntimeError("device gone")
        mem._store.search = boom
        res = await asyncio.gather(*[mem.semantic_search("q", query_embedding=np.ones(32, np.float32)) for _ in range(5)],
                                   return_exceptions=True)
        assert all(isinstance(r, ValueError) and str(r).startswith("Search failed: ") for r in res)
    asyncio.run(run())


def test_more_than_64_distinct_tags(mem):
    """The reference's tag index is unbounded (enhanced_memory.py:133-138); so is the drop-in's."""
    async def run():
        rng = np.random.default_rng(1)
        n = 300
        emb = rng.standard_normal((n, 32)).astype(np.float32)
        tags = [{f"tag{i % 100}", f"grp{i % 3}"} for i in range(n)]       # 103 distinct tags
        await mem.store_semantic_batch([f"item {i}" for i in range(n)], emb, tags=tags)
        assert len(mem._tag_bits) == 103
        q = rng.standard_normal(32).astype(np.float32)
        sims = exact_scores(bf16_round_bits(q[None]), bf16_round_bits(emb))[0]
        for want in ({"tag5"}, {"tag80"}, {"tag99", "grp0"}, {"tag70", "tag71"}, {"grp2"}, {"never"}):
            got = await mem.semantic_search("q", tags=want, limit=50, query_embedding=q, order="score")
            expect = [i for i in range(n) if want <= tags[i]]
            assert len(got) == min(50, len(expect))
            # the best `limit` of the matching rows by the stated score rule, best first
            order = [int(g.text.split()[1]) for g in got]
            assert order == sorted(expect, key=lambda i: (-sims[i], i))[:50]
    asyncio.run(run())
