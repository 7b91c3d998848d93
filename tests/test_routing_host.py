"""CPU checks of the routing host logic: the per-agent fp64 terms, vocabularies and task
encoding in pilottai_b200/routing.py, combined by a straight-line numpy emulation of what
pilott_route_score does on the device, must reproduce the reference fixtures bit for bit."""
import json
import os

import numpy as np
import pytest

from pilottai_b200 import routing as R

CASES = ["small", "mid", "wide", "overloaded", "kat"]


def emulate_kernel(enc, caps, spec, attr_flags, s12, t4, viable, u, max_tags, mode):
    """What route_kernel computes, in plain Python over numpy scalars (fp64, one add each)."""
    nt, na = enc["type"].shape[0], caps.shape[0]
    out_a, out_s = np.full(nt, -1, np.int32), np.full(nt, np.nan)
    for t in range(nt):
        best, best_a = 0.0, -1
        has_req, has_type = bool(enc["flags"][t] & 1), bool(enc["flags"][t] & 2)
        ty = int(enc["type"][t])
        for a in range(na):
            if not viable[a]:
                continue
            missing = bool(np.any(enc["req"][t] & ~caps[a]))
            tagc = int(sum(bin(int(x)).count("1") for x in (enc["tags"][t] & spec[a])))
            tbit = ty >= 0 and bool((int(spec[a][ty >> 6]) >> (ty & 63)) & 1)
            zero = has_req and (not (attr_flags[a] & 2) or missing or bool(enc["flags"][t] & 4))
            sel = 0 if zero else (2 if (has_type and tbit) else 1)
            sc = float(s12[a][sel])
            if mode == 0:
                sc = (sc + float(u[1 if tbit else 0][tagc])) + float(t4[a])
            if mode == 2 and not sc > 0.0:
                continue
            if best_a < 0 or sc > best:
                best, best_a = sc, a
        out_a[t] = best_a
        if best_a >= 0:
            out_s[t] = best
    return out_a, out_s


def host_route(agents, tasks, mode, **cfg):
    vocab = R.Vocab(agents)
    enc = R.encode_tasks(tasks, vocab)
    caps = np.stack([vocab.caps_mask(a.get("caps")) for a in agents])
    spec = np.stack([vocab.spec_mask(a.get("specializations")) for a in agents])
    attr = np.array([(2 if a.get("caps") is not None else 0) for a in agents], np.uint8)
    max_tags = 64 * vocab.words
    u = np.zeros((2, max_tags + 1))
    for c in range(max_tags + 1):
        u[0, c], u[1, c] = (0.1 * c) * 0.2, 0.3 * 0.2
    s12, t4, viable = [], [], []
    cand = cfg.get("candidates")
    for a in agents:
        if mode == "router":
            s, t, ok = R.router_terms(a, cfg["load_threshold"])
        elif mode == "load_balancer":
            s, t, ok = R.lb_terms(a, cfg["lb_metrics"][a["id"]], cfg["max_tasks_per_agent"], cfg["overload_threshold"])
        else:
            s, t, ok = R.delegator_terms(a)
        if cand is not None and a["id"] not in cand:
            ok = False
        s12.append(s); t4.append(t); viable.append(ok)
    return emulate_kernel(enc, caps, spec, attr, np.array(s12), np.array(t4), viable, u, max_tags, R.MODE[mode])


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, f"routing_{name}.json")) as f:
        return json.load(f)


def _ids(agents, idx):
    return [agents[i]["id"] if i >= 0 else None for i in idx]


@pytest.mark.parametrize("name", CASES)
def test_router_host_terms_match_reference(golden_dir, name):
    g = _load(golden_dir, name)
    idx, sc = host_route(g["agents"], g["tasks"], "router", load_threshold=g["load_threshold"])
    assert _ids(g["agents"], idx) == [r["agent"] for r in g["router"]]
    for s, r in zip(sc, g["router"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]
    if name != "kat":
        assert [R.PRIORITY_NAMES[R.task_priority_host(t)] for t in g["tasks"]] == [r["priority"] for r in g["router"]]


@pytest.mark.parametrize("name", CASES[:-1])
def test_load_balancer_host_terms_match_reference(golden_dir, name):
    g = _load(golden_dir, name)
    lbg, cfg = g["load_balancer"], g["lb_config"]
    metrics = {a["id"]: R.lb_metrics_of(a, lbg["system_cpu_percent"], lbg["system_mem_percent"]) for a in g["agents"]}
    for a in g["agents"]:
        assert float(R.lb_composite(metrics[a["id"]], cfg["max_tasks_per_agent"])).hex() == lbg["composite"][a["id"]]
    # candidates keep their list order == table order here (fixture candidates are a subsequence)
    idx, sc = host_route(g["agents"], g["tasks"], "load_balancer", lb_metrics=metrics, candidates=set(lbg["candidates"]),
                         max_tasks_per_agent=cfg["max_tasks_per_agent"], overload_threshold=cfg["overload_threshold"])
    assert _ids(g["agents"], idx) == [r["agent"] for r in lbg["results"]]
    for s, r in zip(sc, lbg["results"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]


@pytest.mark.parametrize("name", CASES[:-1])
def test_delegator_host_terms_match_reference(golden_dir, name):
    g = _load(golden_dir, name)
    idx, sc = host_route(g["agents"], g["tasks"], "delegator")
    assert _ids(g["agents"], idx) == [r["agent"] for r in g["delegator"]]
    for s, r in zip(sc, g["delegator"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]


def test_vocab_and_unknown_capability_flag():
    agents = [{"caps": ["a", "b"], "specializations": ["x"]}, {"caps": None, "specializations": None}]
    v = R.Vocab(agents)
    assert v.cap == {"a": 0, "b": 1} and v.spec == {"x": 0} and v.words == 1
    assert int(v.req_mask(["a", "zzz"])[0]) == 0b01           # unknown capability: not a bit ...
    enc = R.encode_tasks([{"required_capabilities": ["a", "zzz"]}, {"required_capabilities": ["a"]}, {}], v)
    assert list(enc["flags"]) == [1 | 4, 1, 0]                # ... but task flag bit 2
    assert int(v.tag_mask(["x", "x", "nope"])[0]) == 1


def test_task_delegator_bookkeeping_matches_the_reference_log(golden_dir):
    """tests/golden/delegator_session.json is the log of oracle/delegator_script.py run against the
    UNMODIFIED reference TaskDelegator (oracle/gen_golden_delegator.py): history cap, counters,
    running mean, error types (a missing type is counted under None, as there), active-delegation
    removal, day-old history and hour-old delegations dropped by the cleanup, metrics keys."""
    import asyncio
    import json
    import os
    from oracle.delegator_script import session
    from pilottai_b200.orchestration import TaskDelegator
    with open(os.path.join(golden_dir, "delegator_session.json")) as f:
        want = json.load(f)
    got = json.loads(json.dumps(asyncio.run(session(TaskDelegator))))
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g == w, (g, w)


def test_c_task_encoder_equals_the_python_definition():
    """pilottai_b200._taskenc (csrc/task_encode.c) against routing.encode_tasks, the definition:
    every column equal on a seeded table with unknown capabilities / types / tags, tuples and sets
    as containers, missing keys, fractional complexity; a task the fast path declines (a str where
    a list belongs) sends the whole batch through the definition."""
    import numpy as np
    from pilottai_b200 import routing, synth
    if routing._taskenc is None:
        import pytest
        pytest.skip("extension not built (python -m pilottai_b200.build)")
    agents = synth.agent_table(300, seed=7)                      # > 64 capabilities? no: 64 + 64 names, 1 word
    vocab = routing.Vocab(agents)
    tasks = synth.task_table(500, seed=8) + synth.task_table_fast(500, seed=9)
    tasks[3]["required_capabilities"] = tuple(tasks[3].get("required_capabilities", ["cap01"]))
    tasks[4]["tags"] = {"spec01", "spec07", "nope"}
    tasks[5]["complexity"] = 8.5
    tasks[6].pop("complexity", None)
    tasks[7]["required_capabilities"] = []
    fast = routing.encode_task_batch(tasks, vocab)
    slow = routing.encode_tasks(tasks, vocab)
    for k in slow:
        assert np.array_equal(fast[k], slow[k]), k
    assert [routing.task_priority_host(t) for t in tasks] == [
        3 if u else 2 if (c > 8 or d > 5) else 1 if (c > 5 or d > 3) else 0
        for u, c, d in zip(fast["urgent"], fast["cx"], fast["ndeps"])]
    odd = tasks[:10] + [{"required_capabilities": "cap01"}]      # a str iterates as characters in the reference
    got = routing.encode_task_batch(odd, vocab)
    want = routing.encode_tasks(odd, vocab)
    for k in want:
        assert np.array_equal(got[k], want[k]), k


import pytest as _pytest


@_pytest.mark.parametrize("case", ["mid", "wide", "small"])
def test_vectorised_load_classification_matches_reference_fixture(golden_dir, case):
    """LoadBalancer._analyze_agent_loads as array arithmetic against what the
    unmodified reference returned for the same histories (tests/golden/routing_*.json), plus 1024
    agents against the scalar statement of the same rule (load_balancer.py:143-170)."""
    import asyncio, json, os
    import numpy as np
    from pilottai_b200 import LoadBalancer, LoadMetrics
    from pilottai_b200.routing import lb_composite
    with open(os.path.join(golden_dir, f"routing_{case}.json")) as f:
        g = json.load(f)
    lbg = g["load_balancer"]

    class Orch:
        child_agents = {}
    lb = LoadBalancer(Orch(), g["lb_config"])
    for aid, h in lbg["history"].items():
        lb._metrics_history[aid] = [LoadMetrics(**m) for m in h]
    current = {aid: lb._metrics_history[aid][-1] for aid in lbg["history"]}
    over, under = asyncio.run(lb._analyze_agent_loads(current))
    assert over == lbg["overloaded"] and under == lbg["underloaded"]
    if case != "mid":
        return
    rng = np.random.default_rng(11)
    lb2 = LoadBalancer(Orch(), {"max_tasks_per_agent": 60, "overload_threshold": 0.6, "underload_threshold": 0.4})
    cur = {}
    for i in range(1024):
        hist = [LoadMetrics(cpu_usage=float(rng.random()), memory_usage=float(rng.random()), queue_size=int(rng.integers(0, 70)),
                            error_rate=float(rng.random())) for _ in range(int(rng.integers(1, 8)))]
        lb2._metrics_history[f"a{i}"] = hist
        cur[f"a{i}"] = hist[-1]
    over, under = asyncio.run(lb2._analyze_agent_loads(cur))
    want_over, want_under = [], []
    for aid, m in cur.items():
        load, trend = lb2._calculate_composite_load(m), lb2._calculate_load_trend(aid)
        if load > 0.6 and trend > 0:
            want_over.append(aid)
        elif load < 0.4 and trend < 0:
            want_under.append(aid)
    assert over == want_over and under == want_under and over and under
