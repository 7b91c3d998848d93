"""Writes tests/golden/memory_filters.json by running the UNMODIFIED reference
(`pilott.memory.enhanced_memory.EnhancedMemory`, /root/reference) -- test infrastructure.

The reference scores nothing (its match is a substring test, enhanced_memory.py:110), but it does
define WHICH rows a search may return and in what order: priority >= min_priority, every requested
tag present (:118-131), not expired at search time (:20-21,108), result sorted by
(-priority, timestamp) (:114).  Every text below contains the query word, so the substring test
passes everywhere and the returned lists pin exactly those rules.  Stores stay below max_size: once
the reference's deque wraps, its position indexes go stale (:86-87 record positions that shift),
which this build does not reproduce.

usage (container with /root/reference): python oracle/gen_golden_memory.py
"""
import asyncio
import json
import os
import random
import sys

sys.path.insert(0, "/root/reference")
from pilott.memory.enhanced_memory import EnhancedMemory          # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAGS = ["alpha", "beta", "gamma", "delta", "eps", "zeta"]


def scenario(seed, n):
    rng = random.Random(seed)
    items = []
    for i in range(n):
        tags = sorted(t for t in TAGS if rng.random() < 0.4)
        ttl = rng.choice([None, None, None, 3600, 0])            # ttl 0: expired by the time anything searches
        items.append({"text": f"note {i:03d} about the needle", "tags": tags, "priority": rng.randrange(4), "ttl": ttl})
    queries = []
    for min_priority in (0, 1, 2, 3):
        for tags in ([], ["alpha"], ["beta", "gamma"], ["alpha", "delta", "zeta"], ["never-stored"]):
            queries.append({"tags": tags, "min_priority": min_priority, "limit": 1000})
    return items, queries


async def run(items, queries):
    mem = EnhancedMemory(max_size=10000)
    for it in items:
        await mem.store_semantic(it["text"], tags=set(it["tags"]), priority=it["priority"], ttl=it["ttl"])
        await asyncio.sleep(0.0005)                                # distinct timestamps
    await asyncio.sleep(0.01)
    out = []
    for q in queries:
        try:
            got = await mem.semantic_search("NEEDLE", tags=set(q["tags"]) or None, min_priority=q["min_priority"],
                                            limit=q["limit"])
            out.append(dict(q, expect=[g.text for g in got]))
        except ValueError as e:                                   # min_priority above every stored priority, etc.
            out.append(dict(q, error=str(e)))
    return out


def main():
    doc = {}
    for name, seed, n in (("small", 1, 28), ("mid", 2, 60)):
        items, queries = scenario(seed, n)
        doc[name] = {"items": items, "queries": asyncio.run(run(items, queries))}
    path = os.path.join(ROOT, "tests", "golden", "memory_filters.json")
    with open(path, "w") as f:
        json.dump(doc, f, indent=1)
    print(path, {k: len(v["queries"]) for k, v in doc.items()})


if __name__ == "__main__":
    main()
