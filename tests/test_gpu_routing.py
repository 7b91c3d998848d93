"""GPU parity for the routing kernel (through the C ABI): assignments and fp64 scores must
equal the reference fixtures and the oracle restatement bit for bit."""
import json
import os

import numpy as np
import pytest

from oracle import routing_oracle as ro
from pilottai_b200 import synth

pytestmark = pytest.mark.gpu
CASES = ["small", "mid", "wide", "overloaded", "kat"]


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, f"routing_{name}.json")) as f:
        return json.load(f)


def _ids(agents, idx):
    return [agents[i]["id"] if i >= 0 else None for i in idx]


@pytest.mark.parametrize("name", CASES)
def test_router_matches_reference_fixtures(torch_cuda, golden_dir, name):
    from pilottai_b200.routing import RoutingEngine, PRIORITY_NAMES
    g = _load(golden_dir, name)
    eng = RoutingEngine(g["agents"], 0)
    idx, sc = eng.route(g["tasks"], "router", load_threshold=g["load_threshold"])
    assert _ids(g["agents"], idx) == [r["agent"] for r in g["router"]]
    for s, r in zip(sc, g["router"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]
        else:
            assert np.isnan(s)
    if name != "kat":
        assert [PRIORITY_NAMES[p] for p in eng.priorities(g["tasks"])] == [r["priority"] for r in g["router"]]


@pytest.mark.parametrize("name", CASES[:-1])
def test_load_balancer_and_delegator_match_reference_fixtures(torch_cuda, golden_dir, name):
    from pilottai_b200.routing import RoutingEngine, lb_metrics_of
    g = _load(golden_dir, name)
    lbg, cfg = g["load_balancer"], g["lb_config"]
    eng = RoutingEngine(g["agents"], 0)
    metrics = {a["id"]: lb_metrics_of(a, lbg["system_cpu_percent"], lbg["system_mem_percent"]) for a in g["agents"]}
    idx, sc = eng.route(g["tasks"], "load_balancer", lb_metrics=metrics, candidates=lbg["candidates"],
                        max_tasks_per_agent=cfg["max_tasks_per_agent"], overload_threshold=cfg["overload_threshold"])
    assert _ids(g["agents"], idx) == [r["agent"] for r in lbg["results"]]
    for s, r in zip(sc, lbg["results"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]
    idx, sc = eng.route(g["tasks"], "delegator")
    assert _ids(g["agents"], idx) == [r["agent"] for r in g["delegator"]]
    for s, r in zip(sc, g["delegator"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]


def test_router_large_table_matches_oracle(torch_cuda):
    """1024 agents (capability vocabulary of 64 -> one word) x 20k tasks; a 400-task sample is
    checked against the oracle restatement, the rest for range."""
    from pilottai_b200.routing import RoutingEngine
    agents = synth.agent_table(1024, seed=42)
    tasks = synth.task_table(20_000, seed=43)
    eng = RoutingEngine(agents, 0)
    idx, sc = eng.route(tasks, "router", load_threshold=0.8)
    assert all(-1 <= i < 1024 for i in idx)
    for t in range(0, 20_000, 50):
        want, ws = ro.router_route(agents, tasks[t], 0.8)
        assert (agents[idx[t]]["id"] if idx[t] >= 0 else None) == want
        assert want is None or float(sc[t]).hex() == float(ws).hex()


def test_router_wide_vocabulary_generic_words(torch_cuda):
    """> 128 distinct capabilities -> 3 mask words -> the generic (non-templated) kernel path."""
    from pilottai_b200.routing import RoutingEngine
    rng = np.random.default_rng(5)
    agents = synth.agent_table(40, seed=9)
    big = [f"xcap{i}" for i in range(150)]
    for a in agents:
        if a["caps"] is not None:
            a["caps"] = a["caps"] + [c for c in big if rng.random() < 0.3]
    tasks = synth.task_table(300, seed=10)
    for t in tasks[::3]:
        t["required_capabilities"] = [str(c) for c in rng.choice(big, 2, replace=False)]
    eng = RoutingEngine(agents, 0)
    assert eng.vocab.words == 4          # 64 + 150 capabilities -> 214 bits
    idx, sc = eng.route(tasks, "router", load_threshold=0.8)
    for t, i, s in zip(tasks, idx, sc):
        want, ws = ro.router_route(agents, t, 0.8)
        assert (agents[i]["id"] if i >= 0 else None) == want
        assert want is None or float(s).hex() == float(ws).hex()


@pytest.mark.parametrize("seed", range(12))
def test_random_tables_match_oracle_in_all_three_modes(torch_cuda, seed):
    """Seeded fuzz: agent counts that are not multiples of the kernel's unroll, vocabularies of one,
    two and three+ mask words, agents without the capability attribute or without specializations,
    tasks with unknown capabilities / types / tags; router, load balancer and delegator modes."""
    from pilottai_b200.routing import RoutingEngine
    rng = np.random.default_rng(500 + seed)
    na = int(rng.choice([1, 2, 3, 5, 17, 63, 130, 257]))
    extra = int(rng.choice([0, 0, 40, 90, 200]))                 # extra capabilities/specializations -> more words
    agents = synth.agent_table(na, seed=seed + 60)
    xc = [f"xc{i}" for i in range(extra)]
    xs = [f"xs{i}" for i in range(extra)]
    for a in agents:
        if a["caps"] is not None:
            a["caps"] = a["caps"] + [c for c in xc if rng.random() < 0.4]
        if a["specializations"] is not None:
            a["specializations"] = a["specializations"] + [s for s in xs if rng.random() < 0.2]
    tasks = synth.task_table(200, seed=seed + 90)
    for t in tasks[::2]:
        if extra and "required_capabilities" in t:
            t["required_capabilities"] = t["required_capabilities"] + [str(c) for c in rng.choice(xc, 1)]
        if extra and rng.random() < 0.5:
            t["type"] = str(rng.choice(xs))
    eng = RoutingEngine(agents, 0)
    thr = float(rng.choice([0.5, 0.8, 1.0]))
    idx, sc = eng.route(tasks, "router", load_threshold=thr)
    for t, i, s in zip(tasks, idx, sc):
        want, ws = ro.router_route(agents, t, thr)
        assert (agents[i]["id"] if i >= 0 else None) == want, (seed, na, extra)
        assert want is None or float(s).hex() == float(ws).hex()
    lbm = {a["id"]: ro.lb_metrics(a, 10.0, 20.0) for a in agents}
    cands = [a["id"] for a in agents]
    idx, sc = eng.route(tasks, "load_balancer", lb_metrics=lbm, max_tasks_per_agent=60, overload_threshold=thr)
    for t, i, s in zip(tasks, idx, sc):
        want, ws = ro.lb_find_best(agents, lbm, t, cands, 60, thr)
        assert (agents[i]["id"] if i >= 0 else None) == want, (seed, na, extra)
        assert want is None or float(s).hex() == float(ws).hex()
    idx, sc = eng.route(tasks, "delegator")
    for t, i, s in zip(tasks, idx, sc):
        want, ws = ro.delegator_find_best(agents, t)
        assert (agents[i]["id"] if i >= 0 else None) == want, (seed, na, extra)
        assert want is None or float(s).hex() == float(ws).hex()


@pytest.mark.parametrize("seed", range(10))
def test_tensor_core_form_equals_cuda_core_form(torch_cuda, seed):
    """pilott_route_score_tc (tcgen05 GEMM of packed counts + epilogue from TMEM) against
    pilott_route_score (set arithmetic): same agent and bit-identical fp64 score for every task, in
    all three modes, for vocabularies of one to three mask words, agent counts around the 256-column
    tile, agents without attributes, unknown capabilities / types / tags, tasks without keys -- and
    both against the oracle on a sample."""
    torch = torch_cuda
    from pilottai_b200.routing import RoutingEngine
    rng = np.random.default_rng(900 + seed)
    na = int(rng.choice([1, 7, 130, 255, 256, 257, 600, 1024]))
    extra = int(rng.choice([0, 0, 40, 100]))                      # -> 1, 2 or 3 mask words
    agents = synth.agent_table(na, seed=seed + 30)
    xc, xs = [f"xc{i}" for i in range(extra)], [f"xs{i}" for i in range(extra)]
    for a in agents:
        if a["caps"] is not None:
            a["caps"] = a["caps"] + [c for c in xc if rng.random() < 0.4]
        if a["specializations"] is not None:
            a["specializations"] = a["specializations"] + [s for s in xs if rng.random() < 0.2]
    tasks = synth.task_table(700, seed=seed + 40)
    for t in tasks[::2]:
        if extra and "required_capabilities" in t:
            t["required_capabilities"] = t["required_capabilities"] + [str(c) for c in rng.choice(xc, 1)]
        if extra and rng.random() < 0.5:
            t["type"] = str(rng.choice(xs))
    eng = RoutingEngine(agents, 0)
    assert eng.vocab.words <= 3
    thr = float(rng.choice([0.5, 0.8, 1.0]))
    lbm = {a["id"]: ro.lb_metrics(a, 10.0, 20.0) for a in agents}
    for mode, cfg in (("router", dict(load_threshold=thr)),
                      ("load_balancer", dict(lb_metrics=lbm, max_tasks_per_agent=60, overload_threshold=thr)),
                      ("delegator", {})):
        a1, s1 = eng.route(tasks, mode, as_numpy=True, kernel="cuda", **cfg)
        a2, s2 = eng.route(tasks, mode, as_numpy=True, kernel="tc", **cfg)
        assert np.array_equal(a1, a2), (seed, mode, na, extra, np.nonzero(a1 != a2)[0][:5])
        assert np.array_equal(s1.view(np.int64), s2.view(np.int64)), (seed, mode)
    for t in range(0, 700, 35):
        want, ws = ro.router_route(agents, tasks[t], thr)
        a2, s2 = eng.route(tasks[t:t + 1], "router", as_numpy=True, kernel="tc", load_threshold=thr)
        assert (agents[a2[0]]["id"] if a2[0] >= 0 else None) == want
        assert want is None or float(s2[0]).hex() == float(ws).hex()


def test_tensor_core_form_on_reference_fixtures(torch_cuda, golden_dir):
    from pilottai_b200.routing import RoutingEngine
    with open(os.path.join(golden_dir, "routing_mid.json")) as f:
        g = json.load(f)
    eng = RoutingEngine(g["agents"], 0)
    idx, sc = eng.route(g["tasks"], "router", kernel="tc", load_threshold=g["load_threshold"])
    assert [g["agents"][i]["id"] if i >= 0 else None for i in idx] == [r["agent"] for r in g["router"]]
    for s, r in zip(sc, g["router"]):
        if r["agent"] is not None:
            assert float(s).hex() == r["score"]
