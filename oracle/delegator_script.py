"""One scripted session with a TaskDelegator's book-keeping (no scoring, so no GPU): run against
the reference's class by oracle/gen_golden_delegator.py (fixture) and against the drop-in by
tests/test_routing_host.py (comparison).  TEST INFRASTRUCTURE ONLY.
Covers pilott/delegation/task_delegator.py: record_delegation (:183-219) -- history cap, counters,
running mean, error types, removal of the active delegation --, get_agent_metrics (:221-238),
_cleanup_old_history (:280-306), get_metrics (:351-360)."""
import types
from datetime import datetime, timedelta


def _plain(x):
    if isinstance(x, datetime):
        return "<datetime>"
    if isinstance(x, dict):
        return {str(k): _plain(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return [_plain(v) for v in x]
    return x


async def session(delegator_cls):
    agent = types.SimpleNamespace(id="mgr", child_agents={}, config=types.SimpleNamespace(allow_delegation=False),
                                  logger=None)
    dg = delegator_cls(agent)
    dg.logger.disabled = True
    dg.MAX_HISTORY_PER_AGENT = 5
    log = []
    results = [("a", {"status": "completed", "execution_time": 2.0}),
               ("a", {"status": "failed", "execution_time": 1.0, "error": "boom", "error_type": "timeout"}),
               ("a", {"status": "failed", "error": "no type given"}),
               ("b", {"status": "completed", "execution_time": 0.5}),
               ("a", {"status": "completed", "execution_time": 4.0}),
               ("a", {"status": "completed", "execution_time": 3.0}),
               ("a", {"status": "completed", "execution_time": 1.5}),
               ("a", {"status": "cancelled", "execution_time": 0.25, "error_type": "timeout"})]
    for n, (aid, res) in enumerate(results):
        task = {"id": f"t{n}", "type": "extract" if n % 2 else "load", "tags": ["x"] if n % 3 == 0 else []}
        dg.active_delegations[task["id"]] = {"agent_id": aid, "started_at": datetime.now(), "task": task}
        await dg.record_delegation(aid, task, res)
        log.append({"after": n, "active": sorted(dg.active_delegations), "a": _plain(dg.get_agent_metrics("a")),
                    "b": _plain(dg.get_agent_metrics("b")), "history_a": [e["task_id"] for e in dg.delegation_history["a"]],
                    "entry_keys": sorted(dg.delegation_history[aid][-1].keys()),
                    "last_entry": _plain({k: v for k, v in dg.delegation_history[aid][-1].items() if k != "timestamp"})})
    log.append({"unknown_agent": dg.get_agent_metrics("nobody")})
    # age two history entries past the 24 h cut and one delegation past the 1 h cut
    old = (datetime.now() - timedelta(hours=30)).isoformat()
    dg.delegation_history["a"][0]["timestamp"] = old
    dg.delegation_history["a"][2]["timestamp"] = old
    dg.active_delegations["stale"] = {"agent_id": "a", "started_at": datetime.now() - timedelta(hours=2), "task": {"id": "stale"}}
    dg.active_delegations["fresh"] = {"agent_id": "b", "started_at": datetime.now(), "task": {"id": "fresh"}}
    await dg._cleanup_old_history()
    m = await dg.get_metrics()
    log.append({"after_cleanup": {"history_a": [e["task_id"] for e in dg.delegation_history["a"]],
                                  "active": sorted(dg.active_delegations), "metrics_keys": sorted(m.keys()),
                                  "active_delegations": m["active_delegations"], "history_size": m["history_size"],
                                  "agent_metrics": _plain(m["agent_metrics"])}})
    await dg.start()
    running = dg._cleanup_task is not None and not dg._cleanup_task.done()
    await dg.stop()
    log.append({"cleanup_task_ran": running, "stopped": dg._cleanup_task.done()})
    return log
