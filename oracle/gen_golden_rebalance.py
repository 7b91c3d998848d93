"""Writes tests/golden/rebalance.json: LoadBalancer._redistribute_tasks of the UNMODIFIED reference
(/root/reference/pilott/orchestration/load_balancer.py:180-266) on a seeded scenario -- which tasks
leave which overloaded agent, in what order, and where they land.  Test infrastructure.

Harness notes (adapters around the reference, none of its code is edited): agents are real
BaseAgent objects built as in oracle/gen_golden.py; their add_task / remove_task are replaced by
dict-keyed versions, because the load balancer hands over task DICTS (task['id'], :229-233) while
BaseAgent.add_task reads task.id (core/agent.py:472-477) -- with the stock pair every move fails.
usage (container with /root/reference): python oracle/gen_golden_rebalance.py"""
import asyncio
import copy
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
from oracle.gen_golden import _Agents, _Orch, build_reference_agents      # noqa: E402
from oracle.routing_oracle import lb_metrics                               # noqa: E402
from pilottai_b200 import synth                                            # noqa: E402


def scenario(seed=11, na=24):
    rows = synth.agent_table(na, seed=seed, ties=True, edge=False)
    for r in rows:
        r["status"] = "idle"
        r["queue_size"] = min(r["queue_size"], 20)
    tasks = synth.task_table(60, seed=seed + 1)
    over = [rows[i]["id"] for i in (2, 9, 17)]
    under = [rows[i]["id"] for i in (0, 1, 4, 5, 8, 12, 13, 20, 21)]
    queues = {}
    for j, aid in enumerate(over):
        q = []
        for n, t in enumerate(tasks[j * 20: j * 20 + 20]):
            t = dict(t, id=f"{aid}-task{n}", priority=(n * 7 + j) % 5,
                     status="pending" if n % 6 != 5 else "running")
            if n % 9 == 4:
                t["locked"] = True
            if n % 11 == 7:
                t["unmoveable"] = True
            q.append(t)
        queues[aid] = q
    return rows, over, under, queues


def install_dict_tasks(agent, queue):
    agent.tasks = {t["id"]: t for t in queue}

    async def add_task(task):
        if task["id"] in agent.tasks:
            raise ValueError(f"Task {task['id']} already exists")
        agent.tasks[task["id"]] = task

    async def remove_task(task_id):
        agent.tasks.pop(task_id, None)
    agent.add_task, agent.remove_task = add_task, remove_task


async def run(rows, over, under, queues, cfg):
    from pilott.orchestration.load_balancer import LoadBalancer, LoadMetrics
    agents = build_reference_agents(rows)
    by_id = {a.id: a for a in agents}
    for a in agents:
        install_dict_tasks(a, copy.deepcopy(queues.get(a.id, [{"id": f"{a.id}-own{j}", "status": "running"}
                                                               for j in range(len(a.tasks))])))
    orch = _Orch(child_agents=_Agents(agents), verbose=False)
    lb = LoadBalancer(orch, cfg)
    lb.logger.disabled = True
    metrics = {r["id"]: LoadMetrics(**lb_metrics(r, 12.5, 33.0)) for r in rows}
    await lb._redistribute_tasks(list(over), list(under), metrics)
    moved = []
    for a in agents:
        for t in a.tasks.values():
            if "moved_to" in t:
                moved.append({"task": t["id"], "from": t["moved_from"], "to": t["moved_to"]})
    return {"moved": sorted(moved, key=lambda m: m["task"]),
            "queues_after": {a.id: list(a.tasks.keys()) for a in agents}}


def main():
    rows, over, under, queues = scenario()
    cfg = {"max_tasks_per_agent": 60, "overload_threshold": 0.8, "underload_threshold": 0.3, "balance_batch_size": 4}
    out = {"generator": "oracle/gen_golden_rebalance.py (unmodified reference at /root/reference)",
           "agents": rows, "overloaded": over, "underloaded": under, "queues": queues, "lb_config": cfg,
           "system_cpu_percent": 12.5, "system_mem_percent": 33.0,
           "result": asyncio.run(run(rows, over, under, queues, cfg))}
    path = os.path.join(ROOT, "tests", "golden", "rebalance.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(path, out["result"]["moved"])


if __name__ == "__main__":
    main()
