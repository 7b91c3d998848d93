"""Routing oracle -- scalar CPU restatement of the reference's agent scoring.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Parity status: PINNED.  ``oracle/gen_golden.py`` runs the unmodified reference
(/root/reference/pilott/...) on seeded agent tables and tasks and commits the
outputs under ``tests/golden/``; ``tests/test_oracle_routing.py`` checks this
restatement against those fixtures bit for bit (scores compared as IEEE-754
hex strings).

Every function evaluates Python floats (IEEE fp64) in the reference's own
operation order, one rounding per operation.  Citations are
/root/reference/pilott/<file>:<line>.

An *agent row* is a plain dict::

    {"id": str, "status": "idle|busy|waiting|error|stopped",
     "caps": [str] | None,          # config.required_capabilities; None = attribute missing
     "specializations": [str] | None,   # None = attribute missing (hasattr false)
     "queue_size": int, "max_queue_size": int, "active_tasks": int,
     "completed": int, "failed": int,
     "cpu_percent": float, "mem_percent": float,     # what psutil would return
     "active_delegations": int, "max_concurrent_tasks": int | None}
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

NEG_INF = float("-inf")


# --- BaseAgent.get_metrics, core/agent.py:535-547 ---------------------------------------
def agent_metrics(a: dict) -> dict:
    done = a["completed"] + a["failed"]
    return {
        "queue_size": a["queue_size"],
        "active_tasks": a["active_tasks"],
        "success_rate": a["completed"] / done if done > 0 else 0.0,
        "queue_utilization": a["queue_size"] / a["max_queue_size"],
        "cpu_usage": a["cpu_percent"] / 100,
        "memory_usage": a["mem_percent"] / 100,
    }


# --- BaseAgent.evaluate_task_suitability, core/agent.py:549-575 -------------------------
def evaluate_task_suitability(a: dict, task: dict) -> float:
    if "required_capabilities" in task:
        if a.get("caps") is None:
            return 0.0            # AttributeError on config.required_capabilities -> except -> 0.0 (:573-575)
        if set(task["required_capabilities"]) - set(a["caps"]):
            return 0.0            # :553-556
    score = 0.7                   # :559
    if "type" in task and a.get("specializations") is not None:
        if task["type"] in a["specializations"]:
            score += 0.2          # :562-564
    load_penalty = agent_metrics(a)["queue_utilization"] * 0.3   # :567-568
    score = max(0.0, score - load_penalty)                        # :569
    return min(1.0, score)                                        # :571


# --- TaskRouter, core/router.py ---------------------------------------------------------
def router_load_penalty(a: dict) -> float:                       # :97-105
    m = agent_metrics(a)
    return min(1.0, m["queue_utilization"] * 0.5 + m["cpu_usage"] * 0.3 + m["memory_usage"] * 0.2)


def router_spec_bonus(a: dict, task: dict) -> float:              # :116-126
    if a.get("specializations") is None:
        return 0.0
    task_type = task.get("type", "")
    task_tags = set(task.get("tags", []))
    type_match = task_type in a["specializations"]
    tag_matches = len(task_tags & set(a["specializations"]))
    return 0.3 if type_match else (0.1 * tag_matches)


def router_scores(agents: List[dict], task: dict) -> Dict[str, float]:
    """_calculate_agent_scores with a cold cache, core/router.py:64-95."""
    scores = {}
    for a in agents:
        if a["status"] == "busy":                                 # :68-69
            continue
        base = evaluate_task_suitability(a, task)                 # :79
        load = router_load_penalty(a)                             # :80
        spec = router_spec_bonus(a, task)                         # :81
        perf = agent_metrics(a)["success_rate"]                   # :82, :128-133
        scores[a["id"]] = base * 0.4 + (1 - load) * 0.3 + spec * 0.2 + perf * 0.1   # :83-88
    return scores


def router_route(agents: List[dict], task: dict, load_threshold: float = 0.8
                 ) -> Tuple[Optional[str], Optional[float]]:
    """_attempt_routing, core/router.py:51-62: strict `<` mask, first max wins."""
    scores = router_scores(agents, task)
    if not scores:
        return None, None
    by_id = {a["id"]: a for a in agents}
    viable = {i: s for i, s in scores.items()
              if agent_metrics(by_id[i])["queue_utilization"] < load_threshold}   # :55-59,:107-114
    if not viable:
        return None, None
    best = max(viable.items(), key=lambda x: x[1])                # :62 first max wins
    return best


def task_priority(task: dict) -> str:                             # core/router.py:135-145
    if task.get("urgent", False):
        return "critical"
    complexity = task.get("complexity", 1)
    deps = len(task.get("dependencies", []))
    if complexity > 8 or deps > 5:
        return "high"
    if complexity > 5 or deps > 3:
        return "medium"
    return "low"


# --- LoadBalancer, orchestration/load_balancer.py ----------------------------------------
def lb_metrics(a: dict, system_cpu_percent: float = 0.0, system_mem_percent: float = 0.0) -> dict:
    """LoadMetrics as _collect_system_metrics builds them, :96-111."""
    m = agent_metrics(a)
    return {
        "cpu_usage": max(system_cpu_percent / 100.0, m["cpu_usage"]),
        "memory_usage": max(system_mem_percent / 100.0, m["memory_usage"]),
        "queue_size": m["queue_size"],
        "active_tasks": m["active_tasks"],
        "total_tasks": 0,                       # get_metrics has no 'total_tasks' -> default 0 (:109)
        "error_rate": 1 - m["success_rate"],
    }


def lb_composite_load(lm: dict, max_tasks_per_agent: int) -> float:   # :172-178
    return (0.3 * lm["cpu_usage"] + 0.3 * lm["memory_usage"]
            + 0.2 * (lm["queue_size"] / max_tasks_per_agent) + 0.2 * lm["error_rate"])


def lb_can_accept(a: dict, lm: dict, max_tasks_per_agent: int, overload_threshold: float) -> bool:  # :294-304
    return (a["status"] != "stopped" and lm["queue_size"] < max_tasks_per_agent
            and lb_composite_load(lm, max_tasks_per_agent) < overload_threshold)


def lb_suitability(a: dict, task: dict, lm: dict, max_tasks_per_agent: int) -> float:   # :306-332
    base = evaluate_task_suitability(a, task)
    load_score = 1 - lb_composite_load(lm, max_tasks_per_agent)
    perf_score = 1 - lm["error_rate"]
    resource_score = 1 - max(lm["cpu_usage"], lm["memory_usage"])
    return base * 0.4 + load_score * 0.3 + perf_score * 0.2 + resource_score * 0.1


def lb_find_best(agents: List[dict], metrics: Dict[str, dict], task: dict, candidates: List[str],
                 max_tasks_per_agent: int = 10, overload_threshold: float = 0.8
                 ) -> Tuple[Optional[str], float]:
    """_find_best_agent, :268-292: strict `>` => first best candidate wins."""
    by_id = {a["id"]: a for a in agents}
    best_id, best = None, NEG_INF
    for cid in candidates:
        a, lm = by_id[cid], metrics[cid]
        if not lb_can_accept(a, lm, max_tasks_per_agent, overload_threshold):
            continue
        s = lb_suitability(a, task, lm, max_tasks_per_agent)
        if s > best:
            best, best_id = s, cid
    return best_id, best


def lb_load_trend(history: List[dict], max_tasks_per_agent: int) -> float:   # :161-170
    h = history[-5:]
    if len(h) < 2:
        return 0.0
    loads = [lb_composite_load(m, max_tasks_per_agent) for m in h]
    return (loads[-1] - loads[0]) / len(loads)


def lb_classify(current: Dict[str, dict], history: Dict[str, List[dict]], max_tasks_per_agent: int,
                overload_threshold: float, underload_threshold: float):   # :143-159
    over, under = [], []
    for aid, lm in current.items():
        trend = lb_load_trend(history.get(aid, []), max_tasks_per_agent) if aid in history else 0.0
        load = lb_composite_load(lm, max_tasks_per_agent)
        if load > overload_threshold and trend > 0:
            over.append(aid)
        elif load < underload_threshold and trend < 0:
            under.append(aid)
    return over, under


# --- TaskDelegator, delegation/task_delegator.py -----------------------------------------
def delegator_available(a: dict) -> bool:                          # :308-314
    mc = a.get("max_concurrent_tasks")
    return a["status"] not in ("stopped", "error") and a["active_delegations"] < (5 if mc is None else mc)


def delegator_can_accept(a: dict) -> bool:                         # :316-326
    m = agent_metrics(a)
    return (a["status"] not in ("stopped", "error") and m["queue_utilization"] < 0.8
            and m["cpu_usage"] < 0.8 and m["memory_usage"] < 0.8)


def delegator_score(a: dict, task: dict) -> float:                 # :92-113
    base = evaluate_task_suitability(a, task)
    m = agent_metrics(a)
    load_score = 1 - m["queue_utilization"]
    error_rate = 1 - m["success_rate"]
    resource_score = 1 - max(m["cpu_usage"], m["memory_usage"])
    return base * 0.4 + load_score * 0.3 + (1 - error_rate) * 0.2 + resource_score * 0.1


def delegator_find_best(agents: List[dict], task: dict) -> Tuple[Optional[str], Optional[float]]:
    """_find_best_agent, :60-90: keep score > 0, max() => first max wins."""
    scores = {}
    for a in agents:
        if not delegator_available(a) or not delegator_can_accept(a):
            continue
        s = delegator_score(a, task)
        if s > 0:
            scores[a["id"]] = s
    if not scores:
        return None, None
    return max(scores.items(), key=lambda x: x[1])
