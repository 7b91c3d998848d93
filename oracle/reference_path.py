"""Times the reference's OWN CPU code for BASELINE.json configs[0] -- TEST / BENCH INFRASTRUCTURE ONLY
(bench.py's cpu_baseline leg; nothing under pilottai_b200/ imports this).

What runs is the unmodified reference package, imported from the first of
  $PILOTT_REFERENCE, <repo>/baseline/_ref (the offline install that travels to the GPU box),
  /root/reference (the build container)
through its public API:
  EnhancedMemory(max_size=10000).store_semantic x 10 000        pilott/memory/enhanced_memory.py:60-91
  EnhancedMemory.semantic_search, limit 10: full scan (no text matches) and early hit     :93-116
  TaskRouter.route_task over 4 BaseAgent workers (1 manager routes to them), 1 000 task dicts
                                                                pilott/core/router.py:34-49
with the harness shims oracle/gen_golden.py already uses (psutil replaced inside the reference's agent
module by per-agent synthetic readings, an adapter that lets `pilott.agents` be iterated and indexed).
Single-threaded by construction: one asyncio loop under the GIL, so cores = 1.
"""
from __future__ import annotations

import asyncio
import os
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def locate():
    for cand in (os.environ.get("PILOTT_REFERENCE"), os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if cand and os.path.isdir(os.path.join(cand, "pilott")):
            return cand
    return None


async def _memory(n_rows: int, n_queries: int):
    from pilott.memory.enhanced_memory import EnhancedMemory
    mem = EnhancedMemory(max_size=n_rows)
    t0 = time.perf_counter()
    for i in range(n_rows):
        await mem.store_semantic(f"note {i:05d}: agent observed event {i * 7919 % 10007} in stream {i % 13}",
                                 tags={"t%d" % (i % 16)}, priority=i % 4)
    t_ins = time.perf_counter() - t0
    t0 = time.perf_counter()
    for i in range(n_queries):                        # no stored text holds this word: every row is tested
        got = await mem.semantic_search(f"zebra{i}", limit=10)
        assert got == []
    t_full = time.perf_counter() - t0
    t0 = time.perf_counter()
    for i in range(n_queries):                        # matches among the first rows: the scan breaks at `limit`
        got = await mem.semantic_search("observed event", limit=10)
        assert len(got) == 10
    t_hit = time.perf_counter() - t0
    return n_rows / t_ins, n_queries / t_full, n_queries / t_hit


async def _routing(n_tasks: int):
    from oracle import gen_golden
    from pilottai_b200 import synth
    from pilott.core.router import TaskRouter
    rows = synth.agent_table(4, seed=42, ties=False, edge=False)
    for r in rows:                                    # everybody viable: below the 0.8 load threshold
        r["queue_size"] = min(r["queue_size"], r["max_queue_size"] // 2)
        r["status"] = "idle"
    agents = gen_golden.build_reference_agents(rows)
    holder = types.SimpleNamespace(agents=gen_golden._Agents(agents))
    router = TaskRouter(holder, {"load_threshold": 0.8})
    router.logger.disabled = True
    tasks = synth.task_table(n_tasks, seed=43)
    t0 = time.perf_counter()
    for t in tasks:                                   # as shipped: the 5 s per-agent score cache stays warm (router.py:70-77)
        await router.route_task(t)
    t_warm = time.perf_counter() - t0
    t0 = time.perf_counter()
    for t in tasks:                                   # every task scored against every agent
        router.agent_scores.clear(); router.last_check.clear()
        await router.route_task(t)
    t_cold = time.perf_counter() - t0
    return n_tasks / t_warm, n_tasks / t_cold


def run(n_rows: int = 10_000, n_queries: int = 100, n_tasks: int = 1000):
    """Returns the cpu_baseline.reference_path block, or {"unavailable": why}."""
    ref = locate()
    if ref is None:
        return {"unavailable": "no reference package found (PILOTT_REFERENCE, baseline/_ref, /root/reference)"}
    if ref not in sys.path:
        sys.path.insert(0, ref)
    os.environ["PILOTT_REFERENCE"] = ref              # oracle/gen_golden.py reads it at import
    import logging
    logging.disable(logging.CRITICAL)
    try:
        ins, q_full, q_hit = asyncio.run(_memory(n_rows, n_queries))
        r_warm, r_cold = asyncio.run(_routing(n_tasks))
    except Exception as e:                            # a reference dependency missing on this box
        return {"unavailable": f"{type(e).__name__}: {e}"}
    finally:
        logging.disable(logging.NOTSET)
    return {
        "source": ref, "cores": 1, "rows": n_rows,
        "inserts_per_s": ins, "queries_per_s": q_full, "queries_per_s_early_hit": q_hit,
        "routes_per_s": r_warm, "routes_per_s_cache_cleared": r_cold,
        "what": f"unmodified reference: EnhancedMemory.store_semantic x {n_rows}, semantic_search limit 10 x {n_queries} "
                f"(full scan / early hit), TaskRouter.route_task x {n_tasks} over 4 workers (score cache warm / cleared per task)",
    }


if __name__ == "__main__":
    import json
    sys.path.insert(0, ROOT)
    print(json.dumps(run(), indent=1))
