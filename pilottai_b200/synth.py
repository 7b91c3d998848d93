"""Seeded synthetic inputs: corpora, queries, agent tables, tasks.

Pure data generators shared by tests, bench.py and the golden-fixture script.
They contain no scoring logic.
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np

CAP_VOCAB = [f"cap{i:02d}" for i in range(64)]
SPEC_VOCAB = [f"spec{i:02d}" for i in range(64)]


def corpus(n: int, dim: int, seed: int = 1234, dup_frac: float = 0.0, chunk: int = 65536):
    """L2-normalised standard-normal fp32 rows, generated chunk-wise (constant memory)."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, dim), dtype=np.float32)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        x = rng.standard_normal((hi - lo, dim), dtype=np.float32)
        x /= np.linalg.norm(x, axis=1, keepdims=True)
        out[lo:hi] = x
    if dup_frac > 0:
        m = max(1, int(n * dup_frac))
        src = rng.integers(0, n, m)
        dst = rng.integers(0, n, m)
        out[dst] = out[src]
    return out


def queries(q: int, dim: int, seed: int = 5678):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((q, dim), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x


def metadata(n: int, seed: int = 77, n_tags: int = 16, tag_p: float = 0.1, expired_frac: float = 0.01,
             now_us: int = 1_000_000):
    """priority ~ U{0..3}; n_tags tags each w.p. tag_p; expired_frac rows already expired."""
    rng = np.random.default_rng(seed)
    prio = rng.integers(0, 4, n).astype(np.int32)
    bits = rng.random((n, n_tags)) < tag_p
    tagmask = (bits * (np.uint64(1) << np.arange(n_tags, dtype=np.uint64))).sum(axis=1).astype(np.uint64)
    expires = np.zeros(n, dtype=np.int64)
    dead = rng.random(n) < expired_frac
    expires[dead] = now_us - 1 - rng.integers(0, 1000, dead.sum())
    live = (~dead) & (rng.random(n) < expired_frac)
    expires[live] = now_us + 1 + rng.integers(0, 1000, live.sum())
    return prio, tagmask, expires


def agent_table(n_agents: int, seed: int = 42, ties: bool = True, edge: bool = True) -> List[Dict]:
    """Agent rows in the oracle's dict format (oracle/routing_oracle.py docstring)."""
    rng = np.random.default_rng(seed)
    rows = []
    for i in range(n_agents):
        caps = [c for c in CAP_VOCAB if rng.random() < 0.3]
        spec = [s for s in SPEC_VOCAB if rng.random() < 0.1]
        maxq = int(rng.choice([10, 50, 100, 100, 100, 128]))
        qsize = int(rng.integers(0, int(1.2 * maxq) + 1))
        r = rng.random()
        status = "busy" if r < 0.05 else "stopped" if r < 0.07 else "error" if r < 0.09 else \
            "waiting" if r < 0.15 else "idle"
        completed = int(rng.integers(0, 200))
        failed = int(rng.integers(0, 50))
        if rng.random() < 0.05:
            completed = failed = 0
        rows.append({
            "id": f"agent{i:05d}", "status": status, "caps": caps, "specializations": spec,
            "queue_size": qsize, "max_queue_size": maxq, "active_tasks": int(rng.integers(0, 6)),
            "completed": completed, "failed": failed,
            "cpu_percent": float(np.round(rng.random() * 100, 1)),
            "mem_percent": float(np.round(rng.random() * 100, 1)),
            "active_delegations": int(rng.integers(0, 7)),
            "max_concurrent_tasks": None if rng.random() < 0.5 else int(rng.integers(1, 8)),
        })
    if n_agents >= 8:
        if rng.random() < 0.5:
            rows[3]["specializations"] = None            # hasattr(agent,'specializations') false
        rows[5]["caps"] = None                           # simple AgentConfig: attribute missing
    if ties and n_agents >= 16:                          # exact score ties: clone feature rows
        for dst, src in ((7, 2), (11, 2), (13, 9)):
            keep = rows[dst]["id"]
            rows[dst] = dict(rows[src]); rows[dst]["id"] = keep
    if edge and n_agents >= 16:                          # threshold edges: 0.8 exactly is NOT viable
        rows[10].update(queue_size=80, max_queue_size=100, status="idle")
        rows[12].update(queue_size=79, max_queue_size=100, status="idle")
        rows[14].update(cpu_percent=80.0, status="idle")
    return rows


def task_table(n_tasks: int, seed: int = 43) -> List[Dict]:
    rng = np.random.default_rng(seed)
    tasks = []
    for i in range(n_tasks):
        t: Dict = {"id": f"task{i:07d}"}
        r = rng.random()
        if r < 0.85:
            t["required_capabilities"] = [str(c) for c in rng.choice(CAP_VOCAB, int(rng.integers(0, 4)), replace=False)]
        if rng.random() < 0.9:
            t["type"] = str(rng.choice(SPEC_VOCAB)) if rng.random() < 0.95 else "unknown_type"
        ntag = int(rng.integers(0, 5))
        if ntag or rng.random() < 0.5:
            tags = [str(s) for s in rng.choice(SPEC_VOCAB, ntag, replace=False)]
            if rng.random() < 0.05:
                tags.append("not_in_vocab")
            t["tags"] = tags
        t["complexity"] = int(rng.integers(1, 11))
        t["dependencies"] = [f"d{j}" for j in range(int(rng.integers(0, 8)))]
        if rng.random() < 0.1:
            t["urgent"] = True
        t["priority"] = int(rng.integers(0, 4))
        tasks.append(t)
    return tasks


def task_table_fast(n_tasks: int, seed: int = 43) -> List[Dict]:
    """Same distribution as task_table, drawn column-wise (a million dicts in seconds instead of a
    minute; a different random stream, so the two tables differ row by row)."""
    rng = np.random.default_rng(seed)
    caps, specs = np.array(CAP_VOCAB, dtype=object), np.array(SPEC_VOCAB, dtype=object)
    has_req = rng.random(n_tasks) < 0.85
    nreq = rng.integers(0, 4, n_tasks)
    req_pick = np.argsort(rng.random((n_tasks, len(CAP_VOCAB))), axis=1)[:, :3] if n_tasks else np.zeros((0, 3), np.int64)
    has_type = rng.random(n_tasks) < 0.9
    known = rng.random(n_tasks) < 0.95
    type_pick = rng.integers(0, len(SPEC_VOCAB), n_tasks)
    ntag = rng.integers(0, 5, n_tasks)
    tag_pick = np.argsort(rng.random((n_tasks, len(SPEC_VOCAB))), axis=1)[:, :4] if n_tasks else np.zeros((0, 4), np.int64)
    empty_tags = rng.random(n_tasks) < 0.5
    odd_tag = rng.random(n_tasks) < 0.05
    cx = rng.integers(1, 11, n_tasks)
    ndep = rng.integers(0, 8, n_tasks)
    urgent = rng.random(n_tasks) < 0.1
    prio = rng.integers(0, 4, n_tasks)
    deps = [[f"d{j}" for j in range(m)] for m in range(8)]
    cap_l, spec_l = caps.tolist(), specs.tolist()
    tasks = []
    for i in range(n_tasks):
        t: Dict = {"id": f"task{i:07d}"}
        if has_req[i]:
            t["required_capabilities"] = [cap_l[c] for c in req_pick[i, :nreq[i]]]
        if has_type[i]:
            t["type"] = spec_l[type_pick[i]] if known[i] else "unknown_type"
        if ntag[i] or empty_tags[i]:
            tg = [spec_l[c] for c in tag_pick[i, :ntag[i]]]
            if odd_tag[i]:
                tg.append("not_in_vocab")
            t["tags"] = tg
        t["complexity"] = int(cx[i])
        t["dependencies"] = deps[ndep[i]]
        if urgent[i]:
            t["urgent"] = True
        t["priority"] = int(prio[i])
        tasks.append(t)
    return tasks
