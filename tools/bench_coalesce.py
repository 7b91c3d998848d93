"""Single-query callers, the reference's pattern (one `semantic_search(query)` per coroutine,
pilott/memory/enhanced_memory.py:93-99), against the drop-in: 1000 coroutines in flight are coalesced
into a few device batches; the same calls one after the other pay one device search each.
usage: python tools/bench_coalesce.py [rows] [dim] [n_callers]"""
import asyncio, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pilottai_b200 import EnhancedMemory

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
nc = int(sys.argv[3]) if len(sys.argv) > 3 else 1000


async def main():
    rng = np.random.default_rng(0)
    mem = EnhancedMemory(max_size=rows, dim=dim)
    for lo in range(0, rows, 100_000):
        m = min(100_000, rows - lo)
        await mem.store_semantic_batch([f"note {lo + i}" for i in range(m)], rng.standard_normal((m, dim), dtype=np.float32))
    qs = rng.standard_normal((nc, dim), dtype=np.float32)
    await mem.semantic_search("warm", query_embedding=qs[0], limit=5)
    b0, t0 = mem.coalesced_batches, time.perf_counter()
    got = await asyncio.gather(*[mem.semantic_search("q", query_embedding=qs[i], limit=5, order="score") for i in range(nc)])
    dt = time.perf_counter() - t0
    batches = mem.coalesced_batches - b0
    n_seq = min(nc, 200)
    t0 = time.perf_counter()
    for i in range(n_seq):
        alone = await mem.semantic_search("q", query_embedding=qs[i], limit=5, order="score")
        assert [x.text for x in alone] == [x.text for x in got[i]]
    dt_seq = (time.perf_counter() - t0) / n_seq
    print(json.dumps({"rows": rows, "dim": dim, "callers": nc, "concurrent_q_per_s": nc / dt, "device_batches": batches,
                      "one_at_a_time_q_per_s": 1 / dt_seq, "same_results": True}))
    mem.close()

asyncio.run(main())
