"""Drop-ins for the reference's routing / balancing / delegation classes, scored on the GPU.

Same class names, constructor arguments, coroutine names, config fields/ranges and error
behaviour as
  pilott/core/router.py                     TaskRouter, RouterConfig, TaskPriority
  pilott/orchestration/load_balancer.py     LoadBalancer, LoadBalancerConfig, LoadMetrics
  pilott/delegation/task_delegator.py       TaskDelegator
written from scratch around `RoutingEngine` (pilottai_b200/routing.py).  Each scoring pass
takes ONE metrics snapshot per agent and scores every task of the batch against it in one
kernel call; the reference never mutates agent load inside a pass either.
Not reproduced on purpose: the task-independent 5 s score cache (router.py:70-77) and the
blocking `psutil.cpu_percent(interval=1)` per agent (load_balancer.py:101).
"""
from __future__ import annotations

import asyncio
import logging
import weakref
from collections import defaultdict
from datetime import datetime, timedelta
from enum import Enum
from typing import Any, Dict, List, Optional, Sequence, Tuple

import psutil
from pydantic import BaseModel, Field

from .routing import (PRIORITY_NAMES, RoutingEngine, features_from_agent, lb_composite,
                      task_priority_host)


class TaskPriority(str, Enum):
    LOW = "low"
    MEDIUM = "medium"
    HIGH = "high"
    CRITICAL = "critical"


class RouterConfig(BaseModel):                                     # router.py:15-20
    load_check_interval: int = Field(ge=1, default=5)
    max_queue_size: int = Field(gt=0, default=100)
    routing_timeout: int = Field(gt=0, default=30)
    max_retry_attempts: int = Field(ge=0, default=3)
    load_threshold: float = Field(ge=0.0, le=1.0, default=0.8)


def _agent_list(container) -> List[Any]:
    """`Serve.agents` is a dict role -> agent (pilott.py:93); the reference router iterates it
    as if it were a list (router.py:67).  Accept both."""
    if container is None:
        return []
    if isinstance(container, dict):
        return list(container.values())
    if hasattr(container, "values") and not isinstance(container, (list, tuple)):
        return list(container.values())
    return list(container)


class _EngineCache:
    """Keeps a RoutingEngine per agent population; metrics are re-snapshotted every pass."""

    def __init__(self, device: int):
        self.device, self._eng, self._sig = device, None, None

    async def snapshot(self, agents: Sequence[Any], extra=None) -> Tuple[RoutingEngine, List[dict]]:
        rows = [await features_from_agent(a) for a in agents]
        if extra:
            for r in rows:
                r.update(extra.get(r["id"], {}))
        # The device-resident part of the engine (vocabularies, capability / specialization bit sets)
        # depends only on who the agents are and what they can do; it is rebuilt when that changes.
        # The metrics are new every pass: they only enter the per-agent score terms, which the
        # engine derives from `agents` on the first call of each pass.
        sig = tuple((r["id"], None if r["caps"] is None else tuple(r["caps"]),
                     None if r["specializations"] is None else tuple(r["specializations"])) for r in rows)
        if self._eng is None or self._sig != sig:
            self._eng, self._sig = RoutingEngine(rows, self.device), sig
        else:
            self._eng.agents = rows
            self._eng._tables.clear()
        return self._eng, rows


class TaskRouter:
    def __init__(self, pilott: Any, config: Optional[Dict] = None, *, device: int = 0):
        self.pilott = pilott
        self.config = RouterConfig(**(config or {}))
        self._router_lock = asyncio.Lock()
        self._cache = _EngineCache(device)
        self.logger = logging.getLogger("TaskRouter")

    async def _attempt_batch(self, tasks: Sequence[Dict]) -> List[Optional[str]]:
        agents = _agent_list(getattr(self.pilott, "agents", None))
        if not agents or not tasks:
            return [None] * len(tasks)
        eng, rows = await self._cache.snapshot(agents)
        idx, _ = await asyncio.to_thread(eng.route, list(tasks), "router", load_threshold=self.config.load_threshold)
        return [rows[i]["id"] if i >= 0 else None for i in idx]

    async def route_tasks_batch(self, tasks: Sequence[Dict]) -> List[Optional[str]]:
        """One snapshot, one kernel call: agent id (or None) per task.  No retries."""
        async with self._router_lock:
            return await self._attempt_batch(tasks)

    async def route_task(self, task: Dict) -> Optional[str]:      # router.py:34-49
        try:
            async with asyncio.timeout(self.config.routing_timeout):
                async with self._router_lock:
                    for _ in range(self.config.max_retry_attempts):
                        agent_id = (await self._attempt_batch([task]))[0]
                        if agent_id:
                            return agent_id
                        await asyncio.sleep(1)
                    return None
        except asyncio.TimeoutError:
            self.logger.error("Task routing timed out")
            raise RuntimeError("Task routing timed out")
        except Exception as e:
            self.logger.error(f"Routing error: {str(e)}")
            raise

    def get_task_priority(self, task: Dict) -> TaskPriority:       # router.py:135-145
        return TaskPriority(PRIORITY_NAMES[task_priority_host(task)])


# ------------------------------------------------------------------------------------------------
class LoadMetrics(BaseModel):                                      # load_balancer.py:13-20
    cpu_usage: float = 0.0
    memory_usage: float = 0.0
    queue_size: int = 0
    active_tasks: int = 0
    total_tasks: int = 0
    error_rate: float = 0.0
    timestamp: datetime = Field(default_factory=datetime.now)


class LoadBalancerConfig(BaseModel):                               # load_balancer.py:22-30
    check_interval: int = Field(ge=1, default=30)
    overload_threshold: float = Field(ge=0.0, le=1.0, default=0.8)
    underload_threshold: float = Field(ge=0.0, le=1.0, default=0.2)
    max_tasks_per_agent: int = Field(ge=1, default=10)
    balance_batch_size: int = Field(ge=1, default=3)
    min_load_difference: float = Field(ge=0.0, le=1.0, default=0.3)
    metrics_retention_period: int = Field(ge=0, default=3600)
    task_move_timeout: int = Field(ge=1, default=30)


def _lm_dict(m: LoadMetrics) -> dict:
    return {"cpu_usage": m.cpu_usage, "memory_usage": m.memory_usage, "queue_size": m.queue_size,
            "active_tasks": m.active_tasks, "total_tasks": m.total_tasks, "error_rate": m.error_rate}


class LoadBalancer:
    def __init__(self, orchestrator, config: Optional[Dict] = None, *, device: int = 0):
        try:
            self.orchestrator = weakref.proxy(orchestrator)
        except TypeError:
            self.orchestrator = orchestrator
        self.config = LoadBalancerConfig(**(config or {}))
        self.logger = logging.getLogger("LoadBalancer")
        self.running = False
        self.balancing_task: Optional[asyncio.Task] = None
        self._balance_lock = asyncio.Lock()
        self._metrics_history: Dict[str, List[LoadMetrics]] = {}
        self._last_balance_time = datetime.now()
        self._safe_mode = True
        self._cache = _EngineCache(device)

    async def start(self):
        if self.running:
            return
        self.running = True
        self.balancing_task = asyncio.create_task(self._balancing_loop())

    async def stop(self):
        if not self.running:
            return
        self.running = False
        if self.balancing_task:
            self.balancing_task.cancel()
            try:
                await self.balancing_task
            except asyncio.CancelledError:
                pass

    async def _balancing_loop(self):
        while self.running:
            try:
                async with self._balance_lock:
                    await self._balance_system_load()
                await asyncio.sleep(self.config.check_interval)
            except asyncio.CancelledError:
                break
            except Exception as e:
                self.logger.error(f"Error in balancing loop: {str(e)}")
                await asyncio.sleep(self.config.check_interval)

    def _children(self) -> Dict[str, Any]:
        kids = getattr(self.orchestrator, "child_agents", {})
        return dict(kids.items()) if hasattr(kids, "items") else {a.id: a for a in kids}

    async def _get_available_agents(self) -> List[Any]:
        return [a for a in self._children().values() if getattr(a.status, "value", a.status) != "stopped"]

    async def _balance_system_load(self):                          # load_balancer.py:85-94
        try:
            current = await self._collect_system_metrics()
            self._update_metrics_history(current)
            over, under = await self._analyze_agent_loads(current)
            if not over or not under:
                return
            await self._redistribute_tasks(over, under, current)
        except Exception as e:
            self.logger.error(f"Load balancing error: {str(e)}")

    async def _collect_system_metrics(self) -> Dict[str, LoadMetrics]:     # :96-119, one psutil sample per pass
        sys_cpu = psutil.cpu_percent(interval=None) / 100.0
        sys_mem = psutil.virtual_memory().percent / 100.0
        out: Dict[str, LoadMetrics] = {}
        for agent in await self._get_available_agents():
            try:
                m = await agent.get_metrics()
                out[agent.id] = LoadMetrics(
                    cpu_usage=max(sys_cpu, m.get("cpu_usage", 0.0)), memory_usage=max(sys_mem, m.get("memory_usage", 0.0)),
                    queue_size=m.get("queue_size", 0), active_tasks=m.get("active_tasks", 0),
                    total_tasks=m.get("total_tasks", 0), error_rate=1 - m.get("success_rate", 0.0))
                lm = out[agent.id]
                if lm.cpu_usage > self.config.overload_threshold or lm.memory_usage > self.config.overload_threshold:
                    await self._handle_overload(agent.id)
            except Exception as e:
                self.logger.error(f"Error collecting metrics for agent {agent.id}: {str(e)}")
        return out

    async def _handle_overload(self, agent_id: str):
        try:
            await self._children()[agent_id].pause_task_acceptance()
        except Exception as e:
            self.logger.error(f"Error handling overload for agent {agent_id}: {str(e)}")

    def _update_metrics_history(self, metrics: Dict[str, LoadMetrics]):    # :129-141
        keep_after = datetime.now() - timedelta(seconds=self.config.metrics_retention_period)
        for aid, m in metrics.items():
            hist = self._metrics_history.setdefault(aid, [])
            hist.append(m)
            self._metrics_history[aid] = [x for x in hist if x.timestamp > keep_after]

    def _calculate_composite_load(self, metrics: LoadMetrics) -> float:    # :172-178
        return lb_composite(_lm_dict(metrics), self.config.max_tasks_per_agent)

    def _calculate_load_trend(self, agent_id: str) -> float:               # :161-170
        hist = self._metrics_history.get(agent_id, [])[-5:]
        if len(hist) < 2:
            return 0.0
        loads = [self._calculate_composite_load(m) for m in hist]
        return (loads[-1] - loads[0]) / len(loads)

    async def _analyze_agent_loads(self, current_metrics: Dict[str, LoadMetrics]) -> Tuple[List[str], List[str]]:   # :143-159
        """Classification of every agent in one pass of array arithmetic (the reference walks the
        agents one by one): composite load of the current sample and of the first / last of the
        five newest history samples, trend = (last - first) / samples, over = load > overload
        threshold and rising, under = load < underload threshold and falling.  numpy evaluates the
        same fp64 operations in the same order as the scalar code (lb_composite), so the two
        lists are the reference's, in its order."""
        import numpy as np
        ids = list(current_metrics)
        if not ids:
            return [], []
        mt = self.config.max_tasks_per_agent

        def comp(ms):
            a = np.array([(m.cpu_usage, m.memory_usage, m.queue_size, m.error_rate) for m in ms], dtype=np.float64).reshape(-1, 4)
            return 0.3 * a[:, 0] + 0.3 * a[:, 1] + 0.2 * (a[:, 2] / mt) + 0.2 * a[:, 3]

        hists = [self._metrics_history.get(aid, [])[-5:] for aid in ids]
        n = np.array([len(h) for h in hists], dtype=np.int64)
        have = n >= 2
        first = comp([h[0] for h, ok in zip(hists, have) if ok])
        last = comp([h[-1] for h, ok in zip(hists, have) if ok])
        trend = np.zeros(len(ids), np.float64)
        trend[have] = (last - first) / n[have]
        load = comp([current_metrics[aid] for aid in ids])
        over = (load > self.config.overload_threshold) & (trend > 0)
        under = (load < self.config.underload_threshold) & (trend < 0) & ~over
        return [ids[i] for i in np.nonzero(over)[0]], [ids[i] for i in np.nonzero(under)[0]]

    async def _find_best_agents(self, tasks: Sequence[Dict], candidates: List[str],
                                current_metrics: Dict[str, LoadMetrics]) -> List[Optional[str]]:
        """Batched `_find_best_agent`: every task against the candidate list, one kernel call."""
        kids = self._children()
        cands = [c for c in candidates if c in kids and c in current_metrics]
        if not cands or not tasks:
            return [None] * len(tasks)
        eng, rows = await self._cache.snapshot([kids[c] for c in cands])
        lbm = {c: _lm_dict(current_metrics[c]) for c in cands}
        idx, _ = await asyncio.to_thread(
            eng.route, list(tasks), "load_balancer", lb_metrics=lbm,
            max_tasks_per_agent=self.config.max_tasks_per_agent, overload_threshold=self.config.overload_threshold)
        return [rows[i]["id"] if i >= 0 else None for i in idx]

    async def _find_best_agent(self, task: Dict, candidates: List[str],
                               current_metrics: Dict[str, LoadMetrics]) -> Optional[str]:         # :268-292
        return (await self._find_best_agents([task], candidates, current_metrics))[0]

    def _is_task_moveable(self, task) -> bool:                              # :260-266
        get = task.get if hasattr(task, "get") else (lambda k, d=None: getattr(task, k, d))
        return get("status") == "pending" and not get("locked", False) and not get("unmoveable", False)

    async def _get_moveable_tasks(self, agent_id: str) -> List[Dict]:
        try:
            return [t for t in self._children()[agent_id].tasks.values() if self._is_task_moveable(t)]
        except Exception as e:
            self.logger.error(f"Error getting moveable tasks: {str(e)}")
            return []

    async def _redistribute_tasks(self, overloaded: List[str], underloaded: List[str],
                                  current_metrics: Dict[str, LoadMetrics]):                     # :180-218
        for over_id in overloaded:
            try:
                tasks = await self._get_moveable_tasks(over_id)
                tasks.sort(key=lambda x: x.get("priority", 0), reverse=True)
                # The reference scores one task at a time (:192-213), and a move changes the target's
                # queue (agent.tasks grows), which the next task's base suitability sees
                # (agent.py:566-571).  So: score the tasks still to be placed in one device batch
                # against the current agent state, walk them in order up to the first successful
                # move, then re-snapshot and re-score what is left.  At most balance_batch_size + 1
                # device calls per overloaded agent, same assignments as the one-by-one loop.
                moves, pos = 0, 0
                chunk = max(1, self.config.balance_batch_size * 4)
                while pos < len(tasks) and moves < self.config.balance_batch_size:
                    part = tasks[pos: pos + chunk]
                    targets = await self._find_best_agents(part, underloaded, current_metrics)
                    moved = False
                    for task, best in zip(part, targets):
                        pos += 1
                        if not best:
                            continue
                        try:
                            async with asyncio.timeout(self.config.task_move_timeout):
                                await self._move_task(task, over_id, best)
                                moves += 1
                                moved = True
                        except asyncio.TimeoutError:
                            self.logger.error(f"Task move timed out for task {task['id']}")
                        except Exception as e:
                            self.logger.error(f"Failed to move task {task['id']}: {str(e)}")
                        if moved:
                            break                                   # agent state changed: re-score the rest
                if moves > 0:
                    self._last_balance_time = datetime.now()
            except Exception as e:
                self.logger.error(f"Error redistributing tasks for agent {over_id}: {str(e)}")

    async def _move_task(self, task: Dict, from_agent_id: str, to_agent_id: str):                # :220-251
        kids = self._children()
        src, dst = kids[from_agent_id], kids[to_agent_id]
        task["locked"] = True
        backup = task.copy() if self._safe_mode else None
        try:
            await src.remove_task(task["id"])
            await dst.add_task(task)
            task["moved_at"] = datetime.now().isoformat()
            task["moved_from"], task["moved_to"] = from_agent_id, to_agent_id
        except Exception as e:
            self.logger.error(f"Task movement failed: {str(e)}")
            if backup is not None:
                try:
                    await src.add_task(backup)
                except Exception as restore_error:
                    self.logger.error(f"Failed to restore task {task['id']}: {str(restore_error)}")
            raise
        finally:
            task["locked"] = False

    async def get_metrics(self) -> Dict[str, Any]:
        current = await self._collect_system_metrics()
        loads = {a: self._calculate_composite_load(m) for a, m in current.items()}
        return {"active_agents": len(await self._get_available_agents()),
                "overloaded_agents": sum(1 for v in loads.values() if v > self.config.overload_threshold),
                "underloaded_agents": sum(1 for v in loads.values() if v < self.config.underload_threshold),
                "last_balance_time": self._last_balance_time.isoformat(),
                "metrics_history": {a: [m.model_dump() for m in h] for a, h in self._metrics_history.items()}}


# ------------------------------------------------------------------------------------------------
class DelegationMetrics(BaseModel):                                 # task_delegator.py:8-15
    success_count: int = 0
    failure_count: int = 0
    total_execution_time: float = 0
    avg_execution_time: float = 0
    last_success: Optional[datetime] = None
    last_failure: Optional[datetime] = None
    error_types: Dict[str, int] = Field(default_factory=dict)


class TaskDelegator:
    def __init__(self, agent, *, device: int = 0):
        self.agent = agent
        self.delegation_history: Dict[str, List[Dict]] = defaultdict(list)
        self.agent_metrics: Dict[str, DelegationMetrics] = {}
        self.active_delegations: Dict[str, Dict] = {}
        self.logger = logging.getLogger(f"TaskDelegator_{getattr(agent, 'id', 'agent')}")
        self._delegation_lock = asyncio.Lock()
        self.MAX_HISTORY_PER_AGENT = 1000
        self.HISTORY_CLEANUP_INTERVAL = 3600
        self._last_cleanup = datetime.now()
        self._cleanup_task = None
        self._cache = _EngineCache(device)

    async def start(self):                                          # task_delegator.py:30-31
        self._cleanup_task = asyncio.create_task(self._periodic_cleanup())

    async def stop(self):                                           # :33-39
        if self._cleanup_task:
            self._cleanup_task.cancel()
            try:
                await self._cleanup_task
            except asyncio.CancelledError:
                pass

    async def _should_delegate(self, task: Dict) -> bool:           # task_delegator.py:328-349
        try:
            cfg = self.agent.config
            if not cfg.allow_delegation:
                return False
            if (await self.agent.get_metrics())["queue_utilization"] > 0.8:
                return True
            if task.get("complexity", 1) > cfg.max_task_complexity:
                return True
            if task.get("required_capabilities"):
                if set(task["required_capabilities"]) - set(cfg.required_capabilities):
                    return True
            return False
        except Exception as e:
            self.logger.error(f"Error checking delegation: {str(e)}")
            return False

    async def _find_best_agents(self, tasks: Sequence[Dict]) -> List[Optional[Any]]:
        kids = dict(self.agent.child_agents.items())
        if not kids or not tasks:
            return [None] * len(tasks)
        counts = defaultdict(int)
        for d in self.active_delegations.values():
            counts[d["agent_id"]] += 1
        agents = list(kids.values())
        extra = {a.id: {"active_delegations": counts[a.id]} for a in agents}
        eng, rows = await self._cache.snapshot(agents, extra)
        idx, _ = await asyncio.to_thread(eng.route, list(tasks), "delegator")
        return [agents[i] if i >= 0 else None for i in idx]

    async def _find_best_agent(self, task: Dict) -> Optional[Any]:  # :60-90
        try:
            async with asyncio.timeout(10):
                return (await self._find_best_agents([task]))[0]
        except asyncio.TimeoutError:
            self.logger.error("Agent selection timed out")
            return None
        except Exception as e:
            self.logger.error(f"Error finding best agent: {str(e)}")
            return None

    async def evaluate_delegation(self, task: Dict) -> Tuple[bool, Optional[str]]:   # :41-58
        try:
            async with self._delegation_lock:
                if not await self._should_delegate(task):
                    return False, None
                best = await self._find_best_agent(task)
                if best:
                    self.active_delegations[task["id"]] = {"agent_id": best.id, "started_at": datetime.now(), "task": task}
                    return True, best.id
                return False, None
        except Exception as e:
            self.logger.error(f"Delegation evaluation failed: {str(e)}")
            return False, None

    async def record_delegation(self, agent_id: str, task: Dict, result: Dict):      # :183-219
        """Book-keeping after a delegated task came back: history entry (capped per agent), the
        agent's counters and running mean, and the delegation is no longer active."""
        try:
            async with self._delegation_lock:
                ok = result.get("status") == "completed"
                entry = {"task_id": task["id"], "task": task, "timestamp": datetime.now().isoformat(), "success": ok,
                         "execution_time": result.get("execution_time", 0), "error": result.get("error"),
                         "error_type": result.get("error_type")}
                hist = self.delegation_history[agent_id]
                hist.append(entry)
                if len(hist) > self.MAX_HISTORY_PER_AGENT:
                    hist.pop(0)
                m = self.agent_metrics.setdefault(agent_id, DelegationMetrics())
                if ok:
                    m.success_count += 1
                    m.last_success = datetime.now()
                else:
                    m.failure_count += 1
                    m.last_failure = datetime.now()
                    et = entry.get("error_type", "unknown")             # key exists: a missing type is counted under None (:206)
                    m.error_types[et] = m.error_types.get(et, 0) + 1
                m.total_execution_time += entry["execution_time"]
                m.avg_execution_time = m.total_execution_time / (m.success_count + m.failure_count)
                self.active_delegations.pop(task["id"], None)
        except Exception as e:
            self.logger.error(f"Error recording delegation: {str(e)}")

    def get_agent_metrics(self, agent_id: str) -> Optional[Dict[str, Any]]:          # :221-238
        try:
            m = self.agent_metrics.get(agent_id)
            if not m:
                return None
            total = m.success_count + m.failure_count
            return {"success_rate": m.success_count / total if total > 0 else 0,
                    "avg_execution_time": m.avg_execution_time, "total_tasks": total,
                    "last_success": m.last_success, "last_failure": m.last_failure,
                    "error_distribution": m.error_types}
        except Exception:
            return None

    async def _periodic_cleanup(self):                                               # :270-278
        while True:
            try:
                await asyncio.sleep(self.HISTORY_CLEANUP_INTERVAL)
                await self._cleanup_old_history()
            except asyncio.CancelledError:
                break
            except Exception as e:
                self.logger.error(f"Cleanup error: {str(e)}")

    async def _cleanup_old_history(self):                                            # :280-306
        """History older than a day goes, delegations that never reported back after an hour too."""
        try:
            async with self._delegation_lock:
                now = datetime.now()
                day_ago, hour = now - timedelta(hours=24), timedelta(hours=1)
                for agent_id in list(self.delegation_history.keys()):
                    kept = [e for e in self.delegation_history[agent_id] if datetime.fromisoformat(e["timestamp"]) > day_ago]
                    self.delegation_history[agent_id] = kept[-self.MAX_HISTORY_PER_AGENT:]
                for task_id in [t for t, info in self.active_delegations.items() if now - info["started_at"] > hour]:
                    self.active_delegations.pop(task_id)
                self._last_cleanup = now
        except Exception as e:
            self.logger.error(f"Error cleaning up history: {str(e)}")

    async def get_metrics(self) -> Dict[str, Any]:
        return {"active_delegations": len(self.active_delegations),
                "agent_metrics": {a: m.model_dump() for a, m in self.agent_metrics.items()},
                "history_size": sum(len(h) for h in self.delegation_history.values()),
                "last_cleanup": self._last_cleanup.isoformat()}
