"""One load-balancing pass at 1024 agents: metrics collection, classification,
batched re-scoring and redistribution through the LoadBalancer drop-in.  Prints one JSON line.
usage: python tools/bench_balancer.py [n_agents] [tasks_per_overloaded_agent]"""
import asyncio, json, os, sys, time, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pilottai_b200 import LoadBalancer, LoadMetrics, synth
from pilottai_b200.routing import metrics_of

na = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
per = int(sys.argv[2]) if len(sys.argv) > 2 else 20


class Agent:
    def __init__(self, row):
        self.row, self.id, self.status = row, row["id"], row["status"]
        self.config = types.SimpleNamespace(max_queue_size=row["max_queue_size"], required_capabilities=list(row["caps"] or []))
        self.specializations = list(row["specializations"] or [])
        self.tasks = {}

    async def get_metrics(self):
        m = dict(metrics_of(self.row))
        m["queue_size"] = len(self.tasks) or m["queue_size"]
        return m

    async def remove_task(self, tid):
        self.tasks.pop(tid)

    async def add_task(self, task):
        self.tasks[task["id"]] = task

    async def pause_task_acceptance(self):
        pass


rows = synth.agent_table(na, seed=42)
for r in rows:
    r["status"] = "idle"
agents = {r["id"]: Agent(r) for r in rows}
orch = types.SimpleNamespace(child_agents=agents, verbose=False)
lb = LoadBalancer(orch, {"max_tasks_per_agent": 60, "overload_threshold": 0.6, "underload_threshold": 0.45, "balance_batch_size": 3})
rng = np.random.default_rng(7)
tasks = synth.task_table_fast(per * na // 8, seed=9)


async def main():
    t0 = time.perf_counter(); cur = await lb._collect_system_metrics(); t_collect = time.perf_counter() - t0
    for aid, m in cur.items():                      # histories: half rising, half falling
        up = rng.random() < 0.5
        lb._metrics_history[aid] = [LoadMetrics(**{**m.model_dump(exclude={"timestamp"}), "cpu_usage": max(0.0, min(1.0, m.cpu_usage + (s - 4) * (0.02 if up else -0.02)))})
                                    for s in range(5)]
    cur = {aid: h[-1] for aid, h in lb._metrics_history.items()}
    t0 = time.perf_counter(); over, under = await lb._analyze_agent_loads(cur); t_class = time.perf_counter() - t0
    ti = iter(tasks)
    for aid in over:
        for _ in range(per):
            t = next(ti, None)
            if t is None:
                break
            t = dict(t, status="pending"); agents[aid].tasks[t["id"]] = t
    await lb._find_best_agents(tasks[:8], under, cur)                       # warm: engine tables on the device
    t0 = time.perf_counter(); picks = await lb._find_best_agents(tasks[:4096], under, cur); t_score = time.perf_counter() - t0
    t0 = time.perf_counter(); await lb._redistribute_tasks(over, under, cur); t_move = time.perf_counter() - t0
    moved = sum(1 for a in agents.values() for t in a.tasks.values() if t.get("moved_to"))
    print(json.dumps({"agents": na, "overloaded": len(over), "underloaded": len(under), "collect_metrics_s": t_collect,
                      "classify_s": t_class, "score_4096_tasks_s": t_score, "assigned": sum(p is not None for p in picks),
                      "redistribute_s": t_move, "tasks_moved": moved,
                      "reference": "1 s per agent in _collect_system_metrics alone (psutil.cpu_percent(interval=1), load_balancer.py:101)"}))

asyncio.run(main())
