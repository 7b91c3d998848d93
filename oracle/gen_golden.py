"""Generate tests/golden/routing_*.json by running the UNMODIFIED reference.

TEST INFRASTRUCTURE ONLY.  Run in the build container (where /root/reference
exists):  ``python -m oracle.gen_golden``.  The GPU box never needs it: the
fixtures it writes are committed.

What runs is the reference's own code:
  pilott.core.agent.BaseAgent.get_metrics / evaluate_task_suitability
  pilott.core.router.TaskRouter._attempt_routing / _calculate_agent_scores / get_task_priority
  pilott.orchestration.load_balancer.LoadBalancer._find_best_agent / _analyze_agent_loads
  pilott.delegation.task_delegator.TaskDelegator._find_best_agent
Harness shims: psutil is replaced inside the reference
modules by a stand-in that returns the calling agent's synthetic readings;
agents get the rich AgentConfig so ``required_capabilities`` exists; an adapter
lets ``pilott.agents`` be iterated and indexed by id; the router's 5 s score
cache is cleared between tasks.
"""
from __future__ import annotations

import asyncio
import inspect
import json
import os
import sys
import types

REF = os.environ.get("PILOTT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")


def _hex(x):
    return None if x is None else float(x).hex()


class _FakePsutil(types.SimpleNamespace):
    """Returns the synthetic readings of whichever agent is calling get_metrics()."""

    @staticmethod
    def _agent():
        f = inspect.currentframe().f_back.f_back
        while f is not None:
            s = f.f_locals.get("self")
            if s is not None and hasattr(s, "_syn_cpu"):
                return s
            f = f.f_back
        raise RuntimeError("psutil stand-in called outside an agent")

    def cpu_percent(self, interval=None):
        return self._agent()._syn_cpu

    def virtual_memory(self):
        return types.SimpleNamespace(percent=self._agent()._syn_mem)


class _Orch:
    """weakref-able stand-in for the orchestrator (load_balancer.py:34)."""

    def __init__(self, **kw):
        self.__dict__.update(kw)


class _Agents:
    """`for a in pilott.agents` and `pilott.agents[agent_id]` (router.py:67,110)."""

    def __init__(self, agents):
        self._list, self._by_id = list(agents), {a.id: a for a in agents}

    def __iter__(self):
        return iter(self._list)

    def __getitem__(self, k):
        return self._by_id[k]

    def items(self):
        return self._by_id.items()

    def values(self):
        return self._by_id.values()

    def get(self, k, d=None):
        return self._by_id.get(k, d)


def build_reference_agents(rows):
    sys.path.insert(0, REF)
    import pilott.core.agent as ref_agent
    from pilott.core.config import AgentConfig as RichConfig
    from pilott.core.status import AgentStatus

    ref_agent.psutil = _FakePsutil()
    agents = []
    for r in rows:
        a = ref_agent.BaseAgent(role=r["id"], goal="g", llm=object())
        a.id = r["id"]
        if r["caps"] is None:
            a.config.max_queue_size = r["max_queue_size"]          # simple config: no required_capabilities
        else:
            a.config = RichConfig(role=r["id"], goal="g", description="d",
                                  required_capabilities=list(r["caps"]),
                                  max_queue_size=r["max_queue_size"])
        if r["specializations"] is not None:
            a.specializations = list(r["specializations"])
        a.status = AgentStatus(r["status"])
        a.tasks = {f"t{j}": None for j in range(r["queue_size"])}
        a.active_tasks = set(range(r["active_tasks"]))
        a.task_metrics = {"completed": r["completed"], "failed": r["failed"], "timeout": 0}
        a._syn_cpu, a._syn_mem = r["cpu_percent"], r["mem_percent"]
        if r["max_concurrent_tasks"] is not None:
            a.max_concurrent_tasks = r["max_concurrent_tasks"]
        a.logger.disabled = True
        agents.append(a)
    return agents


async def _router_golden(agents, tasks, load_threshold):
    from pilott.core.router import TaskRouter
    holder = types.SimpleNamespace(agents=_Agents(agents))
    router = TaskRouter(holder, {"load_threshold": load_threshold})
    router.logger.disabled = True
    out = []
    for t in tasks:
        router.agent_scores.clear(); router.last_check.clear()
        scores = await router._calculate_agent_scores(t)
        router.agent_scores.clear(); router.last_check.clear()
        best = await router._attempt_routing(t)
        out.append({"agent": best, "score": _hex(scores[best]) if best else None,
                    "priority": router.get_task_priority(t).value,
                    "scores": {k: _hex(v) for k, v in scores.items()}})
    return out


async def _lb_golden(agents, rows, tasks, cfg):
    from pilott.orchestration.load_balancer import LoadBalancer, LoadMetrics
    from oracle.routing_oracle import lb_metrics
    orch = _Orch(child_agents=_Agents(agents), verbose=False)
    lb = LoadBalancer(orch, cfg)
    lb.logger.disabled = True
    metrics = {r["id"]: LoadMetrics(**lb_metrics(r, 12.5, 33.0)) for r in rows}
    cands = [r["id"] for i, r in enumerate(rows) if i % 3 != 1]
    out = []
    for t in tasks:
        best = await lb._find_best_agent(t, cands, metrics)
        sc = None
        if best is not None:
            sc = await lb._calculate_agent_suitability(orch.child_agents[best], t, metrics[best])
        out.append({"agent": best, "score": _hex(sc)})
    composite = {r["id"]: _hex(lb._calculate_composite_load(metrics[r["id"]])) for r in rows}
    # trend / classification on a small synthetic history (:143-170)
    hist_rows = {}
    for i, r in enumerate(rows[:24]):
        h = []
        for step in range(1 + i % 6):
            rr = dict(r); rr["queue_size"] = max(0, r["queue_size"] + (step - 2) * ((i % 3) - 1) * 3)
            rr["cpu_percent"] = min(100.0, max(0.0, r["cpu_percent"] + (step - 2) * ((i % 5) - 2) * 4.5))
            h.append(lb_metrics(rr, 12.5, 33.0))
        hist_rows[r["id"]] = h
        lb._metrics_history[r["id"]] = [LoadMetrics(**m) for m in h]
    current = {aid: lb._metrics_history[aid][-1] for aid in hist_rows}
    over, under = await lb._analyze_agent_loads(current)
    trends = {aid: _hex(lb._calculate_load_trend(aid)) for aid in hist_rows}
    return {"candidates": cands, "system_cpu_percent": 12.5, "system_mem_percent": 33.0,
            "results": out, "composite": composite, "history": hist_rows,
            "overloaded": over, "underloaded": under, "trends": trends}


async def _delegator_golden(agents, rows, tasks):
    from pilott.delegation.task_delegator import TaskDelegator
    mgr = types.SimpleNamespace(id="mgr", child_agents={a.id: a for a in agents})
    dg = TaskDelegator(mgr)
    dg.logger.disabled = True
    for r in rows:
        for j in range(r["active_delegations"]):
            dg.active_delegations[f"{r['id']}-d{j}"] = {"agent_id": r["id"]}
    out = []
    for t in tasks:
        best = await dg._find_best_agent(t)
        sc = await dg._calculate_total_score(best, t) if best is not None else None
        out.append({"agent": best.id if best is not None else None, "score": _hex(sc)})
    return out


def main():
    sys.path.insert(0, os.path.dirname(HERE))
    from pilottai_b200 import synth
    os.makedirs(GOLD, exist_ok=True)
    cases = [("small", 4, 40, 1), ("mid", 64, 400, 42), ("wide", 300, 150, 7), ("overloaded", 12, 30, 3)]
    for name, na, nt, seed in cases:
        rows = synth.agent_table(na, seed=seed, ties=name != "overloaded", edge=name != "overloaded")
        if name == "overloaded":            # nobody viable for the router; a few viable for LB/delegator
            for i, r in enumerate(rows):
                r["queue_size"] = r["max_queue_size"] * 4 // 5 + (i % 3)
        tasks = synth.task_table(nt, seed=seed + 1)
        agents = build_reference_agents(rows)
        cfg = {"max_tasks_per_agent": 60, "overload_threshold": 0.8, "underload_threshold": 0.3}
        gold = {
            "generator": "oracle/gen_golden.py (unmodified reference at /root/reference)",
            "agents": rows, "tasks": tasks, "load_threshold": 0.8, "lb_config": cfg,
            "router": asyncio.run(_router_golden(agents, tasks, 0.8)),
            "load_balancer": asyncio.run(_lb_golden(agents, rows, tasks, cfg)),
            "delegator": asyncio.run(_delegator_golden(agents, rows, tasks)),
        }
        if name != "small":
            for r in gold["router"]:
                r.pop("scores")                 # keep the committed fixture small
        path = os.path.join(GOLD, f"routing_{name}.json")
        with open(path, "w") as f:
            json.dump(gold, f, separators=(",", ":"))
        n_none = sum(1 for r in gold["router"] if r["agent"] is None)
        print(f"{path}: {na} agents, {nt} tasks, router None={n_none}, "
              f"lb None={sum(1 for r in gold['load_balancer']['results'] if r['agent'] is None)}, "
              f"dg None={sum(1 for r in gold['delegator'] if r['agent'] is None)}")

    # SURVEY.md section 7 known-answer test: 4 idle workers, worker 2 specialised in 'extract'
    rows = []
    for i in range(4):
        rows.append({"id": f"w{i}", "status": "idle", "caps": None,
                     "specializations": ["extract"] if i == 2 else None, "queue_size": 0,
                     "max_queue_size": 100, "active_tasks": 0, "completed": 0, "failed": 0,
                     "cpu_percent": 25.0, "mem_percent": 50.0, "active_delegations": 0,
                     "max_concurrent_tasks": None})
    agents = build_reference_agents(rows)
    tasks = [{"type": "extract"}, {"type": "extract", "required_capabilities": ["x"]}, {"type": "other"}]
    kat = {"agents": rows, "tasks": tasks, "load_threshold": 0.8,
           "router": asyncio.run(_router_golden(agents, tasks, 0.8))}
    with open(os.path.join(GOLD, "routing_kat.json"), "w") as f:
        json.dump(kat, f, separators=(",", ":"))
    for t, r in zip(tasks, kat["router"]):
        print(t, r["agent"], {k: float.fromhex(v) for k, v in r["scores"].items()})


if __name__ == "__main__":
    main()
