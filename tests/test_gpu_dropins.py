"""GPU tests of the reference-facing drop-in classes (same names/signatures as the reference):
EnhancedMemory, TaskRouter, LoadBalancer, TaskDelegator.  Routing results are compared with the
fixtures produced by the unmodified reference; memory semantics with the rules the reference
specifies (filters, expiry, ring capacity, error strings)."""
import asyncio
import json
import os
import types

import numpy as np
import pytest

from oracle import routing_oracle as ro

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, f"routing_{name}.json")) as f:
        return json.load(f)


class FakeAgent:
    """Stands in for BaseAgent: the attributes and the get_metrics() the scorers read."""

    def __init__(self, row):
        self.row, self.id, self.status = row, row["id"], row["status"]
        cfg = {"max_queue_size": row["max_queue_size"]}
        if row["caps"] is not None:
            cfg["required_capabilities"] = list(row["caps"])
        self.config = types.SimpleNamespace(**cfg)
        if row["specializations"] is not None:
            self.specializations = list(row["specializations"])
        if row["max_concurrent_tasks"] is not None:
            self.max_concurrent_tasks = row["max_concurrent_tasks"]
        self.tasks = {}

    async def get_metrics(self):
        return ro.agent_metrics(self.row)


def test_task_router_dropin_matches_reference(torch_cuda, golden_dir, monkeypatch):
    from pilottai_b200 import TaskRouter, TaskPriority
    g = _load(golden_dir, "mid")
    agents = [FakeAgent(r) for r in g["agents"]]
    for container in (agents, {f"role{i}": a for i, a in enumerate(agents)}):       # list or Serve-style dict
        router = TaskRouter(types.SimpleNamespace(agents=container), {"load_threshold": g["load_threshold"]})
        got = asyncio.run(router.route_tasks_batch(g["tasks"]))
        assert got == [r["agent"] for r in g["router"]]
    assert asyncio.run(router.route_task(g["tasks"][0])) == g["router"][0]["agent"]
    assert [router.get_task_priority(t).value for t in g["tasks"]] == [r["priority"] for r in g["router"]]
    assert router.get_task_priority({"urgent": True}) is TaskPriority.CRITICAL
    with pytest.raises(Exception):
        TaskRouter(None, {"load_threshold": 1.5})                                      # RouterConfig range kept

    # nobody viable -> None after max_retry_attempts attempts (sleep patched out)
    g2 = _load(golden_dir, "overloaded")
    slept = []

    async def fake_sleep(s):
        slept.append(s)
    monkeypatch.setattr(asyncio, "sleep", fake_sleep)
    r2 = TaskRouter(types.SimpleNamespace(agents=[FakeAgent(r) for r in g2["agents"]]))
    assert asyncio.run(r2.route_task(g2["tasks"][0])) is None
    assert slept == [1, 1, 1]


def test_load_balancer_dropin_matches_reference(torch_cuda, golden_dir):
    from pilottai_b200 import LoadBalancer, LoadBalancerConfig, LoadMetrics
    g = _load(golden_dir, "mid")
    lbg = g["load_balancer"]
    kids = {r["id"]: FakeAgent(r) for r in g["agents"]}

    class Orch:
        child_agents = kids
        verbose = False
    orch = Orch()
    lb = LoadBalancer(orch, g["lb_config"])
    assert isinstance(lb.config, LoadBalancerConfig) and lb.config.balance_batch_size == 3
    metrics = {r["id"]: LoadMetrics(**ro.lb_metrics(r, lbg["system_cpu_percent"], lbg["system_mem_percent"]))
               for r in g["agents"]}
    got = asyncio.run(lb._find_best_agents(g["tasks"], lbg["candidates"], metrics))
    assert got == [r["agent"] for r in lbg["results"]]
    assert asyncio.run(lb._find_best_agent(g["tasks"][3], lbg["candidates"], metrics)) == lbg["results"][3]["agent"]
    for aid, want in lbg["composite"].items():
        assert float(lb._calculate_composite_load(metrics[aid])).hex() == want
    for aid, h in lbg["history"].items():
        lb._metrics_history[aid] = [LoadMetrics(**m) for m in h]
        assert float(lb._calculate_load_trend(aid)).hex() == lbg["trends"][aid]
    current = {aid: lb._metrics_history[aid][-1] for aid in lbg["history"]}
    over, under = asyncio.run(lb._analyze_agent_loads(current))
    assert over == lbg["overloaded"] and under == lbg["underloaded"]

    async def start_stop():
        await lb.start(); await lb.stop()
    asyncio.run(start_stop())
    assert not lb.running                                  # what the reference's own test asserts


def test_task_delegator_dropin_matches_reference(torch_cuda, golden_dir):
    from pilottai_b200 import TaskDelegator
    g = _load(golden_dir, "mid")
    kids = {r["id"]: FakeAgent(r) for r in g["agents"]}
    mgr = types.SimpleNamespace(id="mgr", child_agents=kids)
    dg = TaskDelegator(mgr)
    for r in g["agents"]:
        for j in range(r["active_delegations"]):
            dg.active_delegations[f"{r['id']}-d{j}"] = {"agent_id": r["id"]}
    got = asyncio.run(dg._find_best_agents(g["tasks"]))
    assert [a.id if a is not None else None for a in got] == [r["agent"] for r in g["delegator"]]


def test_enhanced_memory_dropin_semantics(torch_cuda):
    from pilottai_b200 import EnhancedMemory, MemoryItem

    async def scenario():
        mem = EnhancedMemory(max_size=50, dim=128)
        assert await mem.semantic_search("anything") == []                 # empty store: no crash
        with pytest.raises(ValueError, match="Text cannot be empty"):
            await mem.store_semantic("")
        with pytest.raises(ValueError, match="Query cannot be empty"):
            await mem.semantic_search("")
        with pytest.raises(ValueError, match="Failed to store content"):
            await mem.store_semantic("x", priority=-1)                     # MemoryItem validation (ge=0)
        for i in range(80):                                                # ring: 80 into 50
            await mem.store_semantic(f"document number {i} about topic {i % 7}", metadata={"i": i},
                                     tags={"even" if i % 2 == 0 else "odd", f"t{i % 3}"}, priority=i % 4)
        assert len(mem) == 50
        hit = await mem.semantic_search("document number 63 about topic 0", limit=1)
        assert isinstance(hit[0], MemoryItem) and hit[0].metadata["i"] == 63
        assert await mem.semantic_search("document number 5 about topic 5", limit=1, order="score") != [] \
            and (await mem.semantic_search("document number 5 about topic 5", limit=50))[0].metadata["i"] >= 30  # evicted
        res = await mem.semantic_search("document about topic", tags={"even", "t0"}, min_priority=2, limit=50)
        assert res and all(("even" in r.tags and "t0" in r.tags and r.priority >= 2 and r.metadata["i"] >= 30) for r in res)
        want = sorted((i for i in range(30, 80) if i % 2 == 0 and i % 3 == 0 and i % 4 >= 2))
        assert sorted(r.metadata["i"] for r in res) == want               # filters exact after overflow
        keys = [(-r.priority, r.timestamp) for r in res]
        assert keys == sorted(keys)                                        # reference ordering (:114)
        assert await mem.semantic_search("document", tags={"never-stored"}) == []
        # expiry: ttl in the past -> filtered at query time, dropped by cleanup()
        await mem.store_semantic("short lived note", ttl=-5, metadata={"i": "x"})
        assert all(r.metadata["i"] != "x" for r in await mem.semantic_search("short lived note", limit=50))
        n_before = len(mem)
        await mem.cleanup()
        assert len(mem) == n_before - 1
        hit = await mem.semantic_search("document number 79 about topic 2", limit=1)
        assert hit[0].metadata["i"] == 79                                  # host/device still aligned
        # batched vector API
        rng = np.random.default_rng(0)
        emb = rng.standard_normal((10, 128)).astype(np.float32)
        first = await mem.store_semantic_batch([f"vec {i}" for i in range(10)], emb)
        idx, sc = await mem.semantic_search_batch(emb[[3, 7]], k=2)
        assert idx[0, 0] == first + 3 and idx[1, 0] == first + 7 and np.allclose(sc[:, 0], 1.0, atol=1e-3)
        mem.clear()
        assert len(mem) == 0 and await mem.semantic_search("document") == []
        # the other stores keep their contracts
        await mem.store_task("t1", {"type": "a"}); await mem.store_pattern("p", 5)
        assert (await mem.get_recent_tasks(task_type="a"))[0]["data"]["type"] == "a" and await mem.get_pattern("p") == 5
        mem.close()
    asyncio.run(scenario())


def test_knowledge_manager_fronts_the_semantic_store(torch_cuda):
    from pilottai_b200 import EnhancedMemory
    from pilottai_b200.knowledge import KnowledgeManager, SemanticMemorySource

    async def scenario():
        mem = EnhancedMemory(max_size=200, dim=128)
        for i in range(40):
            await mem.store_semantic(f"note {i}: the {['red', 'green', 'blue'][i % 3]} widget ships in batch {i}",
                                     metadata={"i": i}, priority=i % 3)
        km = KnowledgeManager()
        src = SemanticMemorySource(mem, limit=3)
        assert await km.add_source(src) is True
        r1 = await km.query_knowledge("note 17: the blue widget ships in batch 17", ["semantic_memory"])
        assert r1[0][0]["metadata"]["i"] == 17 and len(r1[0]) == 3
        launches = mem._store.launch_count()
        r2 = await km.query_knowledge("note 17: the blue widget ships in batch 17", ["semantic_memory"])
        assert r2 == r1 and mem._store.launch_count() == launches          # cache hit: no device work
        assert km.get_source_stats()["semantic_memory"]["access_count"] == 1
        mem.close()
    asyncio.run(scenario())


def test_serve_adapter_wires_router_and_memory(torch_cuda, golden_dir):
    from pilottai_b200.serve_adapter import attach, task_to_dict
    g = _load(golden_dir, "mid")
    agents = {f"role{i}": FakeAgent(r) for i, r in enumerate(g["agents"])}

    class FakeServe:                      # the attributes of pilott.Serve the adapter touches
        def __init__(self):
            self.agents, self.stored, self.logger = agents, [], types.SimpleNamespace(error=lambda *a: None)

        async def _select_agent(self, task):            # reference default: first non-busy (pilott.py:494-496)
            return next(a for a in self.agents.values() if a.status != "busy")

        async def _update_memory(self, task, result):
            self.stored.append(task)

    async def scenario():
        serve = FakeServe()
        first_non_busy = await serve._select_agent(None)
        router, mem = attach(serve, device=0)
        picks = [await serve._select_agent(t) for t in g["tasks"][:25]]
        assert [a.id for a in picks] == [r["agent"] for r in g["router"][:25]]       # routed, not first-non-busy
        assert any(a.id != first_non_busy.id for a in picks)
        task = types.SimpleNamespace(id="t-1", description="summarise the quarterly pdf report",
                                     metadata={"type": "spec03", "tags": ["spec01"]}, complexity=9, dependencies=[])
        assert task_to_dict(task)["type"] == "spec03" and task_to_dict(task)["complexity"] == 9
        await serve._update_memory(task, types.SimpleNamespace(success=True))
        assert serve.stored == [task] and len(mem) == 1
        hit = await agents["role0"].enhanced_memory.semantic_search("quarterly pdf report summary", limit=1)
        assert hit[0].metadata["task_id"] == "t-1" and hit[0].priority == 1
        mem.close()
    asyncio.run(scenario())


@pytest.mark.parametrize("name", ["small", "mid"])
def test_enhanced_memory_dropin_returns_the_reference_lists(torch_cuda, golden_dir, name):
    """tests/golden/memory_filters.json holds what the UNMODIFIED reference returned (every text
    matches its substring test).  The drop-in gets the same items with one and the same embedding,
    so every cosine ties and its answer is decided by the same rules alone: priority floor, tag
    AND, expiry, (-priority, timestamp) order.  `small` (28 items, limit 32) goes through the fused
    path and its repair step (all scores tie), `mid` (60 items, limit 100) through the large-k path."""
    from pilottai_b200.memory import EnhancedMemory
    with open(os.path.join(golden_dir, "memory_filters.json")) as f:
        sc = json.load(f)[name]
    limit = 32 if name == "small" else 100
    e = np.ones(64, np.float32)

    async def run():
        mem = EnhancedMemory(max_size=10000, dim=64)
        for it in sc["items"]:
            await mem.store_semantic(it["text"], tags=set(it["tags"]), priority=it["priority"], ttl=it["ttl"], embedding=e)
            await asyncio.sleep(0.0003)                        # distinct timestamps, as in the fixture run
        await asyncio.sleep(0.01)
        for q in sc["queries"]:
            got = await mem.semantic_search("needle", tags=set(q["tags"]) or None, min_priority=q["min_priority"],
                                            limit=limit, query_embedding=e)
            assert [g.text for g in got] == q["expect"], q
        mem.close()
    asyncio.run(run())


def test_load_balancer_redistribution_matches_reference(torch_cuda, golden_dir):
    """tests/golden/rebalance.json: which tasks the UNMODIFIED reference's _redistribute_tasks
    (load_balancer.py:180-266) moved off three overloaded agents, and where to.  A move grows the
    target's queue and the next task's base suitability sees it (agent.py:566-571), so the
    batched drop-in has to re-score after every move to land on the same assignments."""
    import copy
    from pilottai_b200 import LoadBalancer, LoadMetrics
    with open(os.path.join(golden_dir, "rebalance.json")) as f:
        g = json.load(f)

    class DictTaskAgent(FakeAgent):
        def __init__(self, row, queue):
            super().__init__(row)
            own = [{"id": f"{row['id']}-own{j}", "status": "running"} for j in range(row["queue_size"])]
            self.tasks = {t["id"]: t for t in copy.deepcopy(queue if queue is not None else own)}

        async def get_metrics(self):
            return ro.agent_metrics(dict(self.row, queue_size=len(self.tasks)))     # agent.py:535-547

        async def add_task(self, task):
            if task["id"] in self.tasks:
                raise ValueError(f"Task {task['id']} already exists")
            self.tasks[task["id"]] = task

        async def remove_task(self, task_id):
            self.tasks.pop(task_id, None)

    kids = {r["id"]: DictTaskAgent(r, g["queues"].get(r["id"])) for r in g["agents"]}

    class Orch:
        child_agents = kids
        verbose = False
    orch = Orch()                                          # the balancer keeps a weak reference, like the reference
    lb = LoadBalancer(orch, g["lb_config"])
    metrics = {r["id"]: LoadMetrics(**ro.lb_metrics(r, g["system_cpu_percent"], g["system_mem_percent"]))
               for r in g["agents"]}
    asyncio.run(lb._redistribute_tasks(list(g["overloaded"]), list(g["underloaded"]), metrics))
    moved = sorted(({"task": t["id"], "from": t["moved_from"], "to": t["moved_to"]}
                    for a in kids.values() for t in a.tasks.values() if "moved_to" in t), key=lambda m: m["task"])
    assert moved == g["result"]["moved"]
    assert {aid: list(a.tasks.keys()) for aid, a in kids.items()} == g["result"]["queues_after"]


def _reference_or_skip():
    """The unmodified reference package: /root/reference in the build container, baseline/_ref
    (oracle/install_reference.sh) on the GPU box."""
    import sys
    from oracle import reference_path
    ref = reference_path.locate()
    if ref is None:
        pytest.skip("no reference package on this machine (baseline/_ref not installed)")
    os.environ["PILOTT_REFERENCE"] = ref
    if ref not in sys.path:
        sys.path.insert(0, ref)
    return ref


def test_adapter_on_the_real_reference_serve(torch_cuda, golden_dir):
    """SURVEY section 8(f) N1 on the reference's own classes: a real `pilott.Serve` over real
    `BaseAgent`s (psutil readings replaced per agent, as in the fixture run), real `Task` /
    `TaskResult` models.  After attach(), `Serve._select_agent` must return the agent the
    unmodified reference router picked for the same task (tests/golden/routing_mid.json), or the
    reference's own default pick where the router finds nobody (pilott.py:494-496), and
    `Serve._update_memory` must land in both memories."""
    _reference_or_skip()
    import logging
    from oracle import gen_golden
    from pilott import Serve
    from pilott.core.task import Task, TaskResult
    from pilottai_b200.serve_adapter import attach
    g = _load(golden_dir, "mid")
    workers = gen_golden.build_reference_agents(g["agents"])
    logging.disable(logging.CRITICAL)

    async def scenario():
        serve = Serve(agents=workers, manager_llm=object())
        assert list(serve.agents) == [r["id"] for r in g["agents"]]          # dict role -> agent (pilott.py:93)
        default = await serve._select_agent(None)                              # reference behaviour before attach
        router, mem = attach(serve, device=0)
        n, routed = 60, 0
        for t, want in zip(g["tasks"][:n], g["router"][:n]):
            meta = {k: t[k] for k in ("type", "required_capabilities", "tags", "urgent") if k in t}
            # (dependencies stay empty: the reference's cycle check rejects any non-empty list of
            #  plain ids, task.py:121-134; they do not enter the routing score)
            task = Task(description=f"task {t['id']}", metadata=meta, complexity=t.get("complexity"))
            got = await serve._select_agent(task)
            if want["agent"] is None:
                assert got is default
            else:
                assert got.id == want["agent"]
                routed += 1
        assert routed > n // 2
        task = Task(description="summarise the quarterly pdf report", metadata={"type": "spec03"})
        await serve._update_memory(task, TaskResult(success=True, output="done", execution_time=0.1))
        assert len(mem) == 1 and workers[0].enhanced_memory is mem
        kept = serve.memory.retrieve({"type": "task_execution"})                # the reference's own memory got it too
        if asyncio.iscoroutine(kept):
            kept = await kept
        assert len(kept) == 1
        hit = await workers[3].enhanced_memory.semantic_search("quarterly pdf report summary", limit=1)
        assert hit[0].metadata["task_id"] == task.id and hit[0].priority == 1
        mem.close()
    try:
        asyncio.run(scenario())
    finally:
        logging.disable(logging.NOTSET)


def test_thousand_concurrent_single_searches(torch_cuda):
    """Reference callers issue one semantic_search per coroutine (enhanced_memory.py:93-99).  1000 of
    them in flight are served by a handful of device batches and each gets exactly what it gets alone."""
    import time
    from pilottai_b200 import EnhancedMemory

    async def scenario():
        rng = np.random.default_rng(3)
        n, dim = 20_000, 256
        mem = EnhancedMemory(max_size=n, dim=dim)
        emb = rng.standard_normal((n, dim)).astype(np.float32)
        await mem.store_semantic_batch([f"note {i}" for i in range(n)], emb,
                                       tags=[{"even"} if i % 2 == 0 else {"odd"} for i in range(n)],
                                       priority=[i % 4 for i in range(n)])
        qs = rng.standard_normal((1000, dim)).astype(np.float32)
        kw = [dict(limit=1 + i % 9, tags=({"even"} if i % 3 == 0 else None), min_priority=i % 2) for i in range(1000)]
        await mem.semantic_search("warm", query_embedding=qs[0])
        b0, t0 = mem.coalesced_batches, time.perf_counter()
        got = await asyncio.gather(*[mem.semantic_search("q", query_embedding=qs[i], order="score", **kw[i]) for i in range(1000)])
        dt = time.perf_counter() - t0
        assert mem.coalesced_batches - b0 <= 12
        t1 = time.perf_counter()
        for i in range(0, 1000, 10):
            alone = await mem.semantic_search("q", query_embedding=qs[i], order="score", **kw[i])
            assert [x.text for x in alone] == [x.text for x in got[i]] and len(alone) == kw[i]["limit"]
        dt_alone = (time.perf_counter() - t1) / 100
        print(f"\\n1000 concurrent semantic_search: {1000 / dt:.0f} q/s in {mem.coalesced_batches - b0} device batches; "
              f"one at a time: {1 / dt_alone:.0f} q/s")
        mem.close()
    asyncio.run(scenario())


def test_more_than_64_tags_on_the_device_store(torch_cuda):
    from pilottai_b200 import EnhancedMemory

    async def scenario():
        rng = np.random.default_rng(4)
        n, dim = 500, 64
        mem = EnhancedMemory(max_size=1000, dim=dim)
        emb = rng.standard_normal((n, dim)).astype(np.float32)
        tags = [{f"tag{i % 120}", f"grp{i % 3}"} for i in range(n)]
        await mem.store_semantic_batch([f"item {i}" for i in range(n)], emb, tags=tags)
        q = rng.standard_normal(dim).astype(np.float32)
        for want in ({"tag5"}, {"tag100"}, {"tag119", "grp2"}, {"tag70", "tag71"}, {"never"}):
            got = await mem.semantic_search("q", tags=want, limit=20, query_embedding=q)
            expect = {i for i in range(n) if want <= tags[i]}
            assert {int(x.text.split()[1]) for x in got} <= expect and len(got) == min(20, len(expect))
        mem.close()
    asyncio.run(scenario())
