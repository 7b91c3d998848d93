"""Wiring the accelerated path into a reference `Serve` instance without editing `pilott/`.

In the reference none of the hot-path classes are reachable from `Serve` (SURVEY.md section
0.3): `Serve._select_agent` takes the manager's pick or the first non-busy agent
(pilott/pilott.py:488-504) and `Serve._update_memory` writes to `core.memory.Memory`
(pilott/pilott.py:653-666).  `attach()` replaces those two bound methods on ONE instance:

* agent selection -> `TaskRouter.route_task` (scored on the GPU); the original selection is
  the fallback when the router finds no viable agent or fails;
* memory update   -> the original store, plus `EnhancedMemory.store_semantic` of the task
  description so that later `semantic_search` calls can retrieve it; every agent also gets
  `agent.enhanced_memory` (README.md:154 usage).

Works with any object that has the same attributes (`agents`, `_select_agent`,
`_update_memory`), which is how the tests exercise it without the reference installed.
"""
from __future__ import annotations

import types
from typing import Any, Dict


def task_to_dict(task: Any) -> Dict[str, Any]:
    """The reference routes plain dicts (router.py:34); `Serve` carries `Task` models."""
    if isinstance(task, dict):
        return task
    meta = dict(getattr(task, "metadata", {}) or {})
    d: Dict[str, Any] = {"id": getattr(task, "id", None), "description": getattr(task, "description", "")}
    for key in ("type", "required_capabilities", "tags", "urgent"):
        if key in meta:
            d[key] = meta[key]
    if getattr(task, "complexity", None) is not None:
        d["complexity"] = task.complexity
    d["dependencies"] = list(getattr(task, "dependencies", []) or [])
    return d


def attach(serve: Any, *, router=None, memory=None, device: int = 0):
    """Returns (router, memory) after wiring them into `serve`."""
    from .memory import EnhancedMemory
    from .orchestration import TaskRouter

    if router is None:
        router = TaskRouter(serve, device=device)
    if memory is None:
        memory = EnhancedMemory(device=device)
    original_select, original_update = serve._select_agent, serve._update_memory

    async def _select_agent(self, task):
        try:
            agent_id = (await router.route_tasks_batch([task_to_dict(task)]))[0]
            if agent_id is not None:
                for agent in (self.agents.values() if isinstance(self.agents, dict) else self.agents):
                    if agent.id == agent_id:
                        return agent
        except Exception as e:                                   # same stance as pilott.py:499-503
            getattr(self, "logger", None) and self.logger.error(f"Agent selection failed: {str(e)}")
        return await original_select(task)

    async def _update_memory(self, task, result):
        await original_update(task, result)
        try:
            text = getattr(task, "description", None) or str(task)
            ok = getattr(result, "success", None)
            await memory.store_semantic(text, metadata={"task_id": getattr(task, "id", None), "success": ok},
                                        tags={"task_execution"}, priority=1 if ok else 0)
        except Exception as e:                                   # pilott.py:665-666: log, do not raise
            getattr(self, "logger", None) and self.logger.error(f"Error updating memory: {str(e)}")

    serve._select_agent = types.MethodType(_select_agent, serve)
    serve._update_memory = types.MethodType(_update_memory, serve)
    for agent in (serve.agents.values() if isinstance(serve.agents, dict) else serve.agents):
        agent.enhanced_memory = memory
    serve.task_router, serve.enhanced_memory = router, memory
    return router, memory
