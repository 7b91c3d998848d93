"""The routing oracle restatement vs fixtures produced by the unmodified reference
(oracle/gen_golden.py).  Scores are compared as IEEE-754 hex strings: bit-exact."""
import json
import os

import pytest

from oracle import routing_oracle as ro

CASES = ["small", "mid", "wide", "overloaded"]


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, f"routing_{name}.json")) as f:
        return json.load(f)


def _hex(x):
    return None if x is None else float(x).hex()


@pytest.mark.parametrize("name", CASES)
def test_router_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    for task, want in zip(g["tasks"], g["router"]):
        agent, score = ro.router_route(g["agents"], task, g["load_threshold"])
        assert agent == want["agent"]
        assert _hex(score) == want["score"]
        assert ro.task_priority(task) == want["priority"]
        if "scores" in want:
            got = {k: _hex(v) for k, v in ro.router_scores(g["agents"], task).items()}
            assert got == want["scores"]


@pytest.mark.parametrize("name", CASES)
def test_load_balancer_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    lbg, cfg = g["load_balancer"], g["lb_config"]
    metrics = {a["id"]: ro.lb_metrics(a, lbg["system_cpu_percent"], lbg["system_mem_percent"])
               for a in g["agents"]}
    for a in g["agents"]:
        assert _hex(ro.lb_composite_load(metrics[a["id"]], cfg["max_tasks_per_agent"])) == lbg["composite"][a["id"]]
    for task, want in zip(g["tasks"], lbg["results"]):
        agent, score = ro.lb_find_best(g["agents"], metrics, task, lbg["candidates"],
                                       cfg["max_tasks_per_agent"], cfg["overload_threshold"])
        assert agent == want["agent"]
        if agent is not None:
            assert _hex(score) == want["score"]
    hist = lbg["history"]
    for aid, h in hist.items():
        assert _hex(ro.lb_load_trend(h, cfg["max_tasks_per_agent"])) == lbg["trends"][aid]
    current = {aid: h[-1] for aid, h in hist.items()}
    over, under = ro.lb_classify(current, hist, cfg["max_tasks_per_agent"],
                                 cfg["overload_threshold"], cfg["underload_threshold"])
    assert over == lbg["overloaded"] and under == lbg["underloaded"]


@pytest.mark.parametrize("name", CASES)
def test_delegator_matches_reference(golden_dir, name):
    g = _load(golden_dir, name)
    for task, want in zip(g["tasks"], g["delegator"]):
        agent, score = ro.delegator_find_best(g["agents"], task)
        assert agent == want["agent"]
        assert _hex(score) == want["score"]


def test_known_answers(golden_dir):
    """SURVEY.md section 7 hand-checked values, reproduced by the reference itself."""
    g = _load(golden_dir, "kat")
    want_scores = [
        {"w0": 0.5275, "w1": 0.5275, "w2": 0.6675, "w3": 0.5275},
        {"w0": 0.24749999999999997, "w1": 0.24749999999999997, "w2": 0.3075, "w3": 0.24749999999999997},
        {"w0": 0.5275, "w1": 0.5275, "w2": 0.5275, "w3": 0.5275},
    ]
    want_agent = ["w2", "w2", "w0"]          # all-equal scores -> first agent wins
    for task, ref, ws, wa in zip(g["tasks"], g["router"], want_scores, want_agent):
        assert ref["agent"] == wa
        assert {k: float.fromhex(v) for k, v in ref["scores"].items()} == ws
        assert ro.router_scores(g["agents"], task) == ws
        assert ro.router_route(g["agents"], task)[0] == wa
