"""Host side of the routing kernel (pilott_route_score): vocabularies, per-agent fp64 terms,
task encoding, and the RoutingEngine that owns the device tables.

The reference scores one (task, agent) pair at a time in Python floats:
  TaskRouter._calculate_agent_scores / _attempt_routing      pilott/core/router.py:51-95
  LoadBalancer._find_best_agent / _calculate_agent_suitability
                                                pilott/orchestration/load_balancer.py:268-336
  TaskDelegator._find_best_agent / _calculate_total_score
                                                pilott/delegation/task_delegator.py:60-113
  BaseAgent.evaluate_task_suitability / get_metrics            pilott/core/agent.py:535-575
Everything that depends only on the agent is evaluated HERE, once per agent, with Python
floats in the reference's own operation order (so every intermediate rounding is the
reference's); the device adds the task-dependent term(s) with single IEEE fp64 additions
and takes the first maximum.  Results are therefore bit-identical to the reference.

An agent row is a plain dict (same format as the test oracle uses):
  id, status, caps (list|None), specializations (list|None), queue_size, max_queue_size,
  active_tasks, completed, failed, cpu_percent, mem_percent, active_delegations,
  max_concurrent_tasks (int|None)
`features_from_agent()` builds one from a live BaseAgent-like object.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Iterable, List, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import check, ptr

MODE = {"router": 0, "load_balancer": 1, "delegator": 2}
# kernel="auto": which form of K5 runs (set from the measurements in DESIGN.md K5)
AUTO_TC = False
AUTO_TC_MIN_TASKS = 65536
PRIORITY_NAMES = ["low", "medium", "high", "critical"]


# ---- agent snapshot ------------------------------------------------------------------------
def metrics_of(a: dict) -> dict:
    """BaseAgent.get_metrics on a snapshot row (pilott/core/agent.py:535-547)."""
    if "queue_utilization" in a:             # row built from a live get_metrics() call
        return a
    done = a["completed"] + a["failed"]
    return {
        "queue_size": a["queue_size"], "active_tasks": a["active_tasks"],
        "success_rate": a["completed"] / done if done > 0 else 0.0,
        "queue_utilization": a["queue_size"] / a["max_queue_size"],
        "cpu_usage": a["cpu_percent"] / 100, "memory_usage": a["mem_percent"] / 100,
    }


async def features_from_agent(agent) -> dict:
    """Snapshot a live agent: ONE get_metrics() call (the reference makes four per score)."""
    m = await agent.get_metrics()
    status = getattr(agent.status, "value", agent.status)
    cfg = getattr(agent, "config", None)
    caps = getattr(cfg, "required_capabilities", None)
    spec = getattr(agent, "specializations", None)
    return {
        "id": agent.id, "status": str(status),
        "caps": None if caps is None else list(caps),
        "specializations": None if spec is None else list(spec),
        "queue_size": m.get("queue_size", 0), "active_tasks": m.get("active_tasks", 0),
        "success_rate": m.get("success_rate", 0.0), "queue_utilization": m["queue_utilization"],
        "cpu_usage": m["cpu_usage"], "memory_usage": m["memory_usage"],
        "active_delegations": 0, "max_concurrent_tasks": getattr(agent, "max_concurrent_tasks", None),
    }


def _base_options(q_util: float) -> Tuple[float, float, float]:
    """evaluate_task_suitability's three possible values (agent.py:549-575):
    [0] a required capability is missing (or the attribute is) -> 0.0
    [1] no type match, [2] task type in specializations."""
    pen = q_util * 0.3
    s1 = 0.7
    s2 = 0.7
    s2 += 0.2
    return 0.0, min(1.0, max(0.0, s1 - pen)), min(1.0, max(0.0, s2 - pen))


def router_terms(a: dict, load_threshold: float):
    """TaskRouter: s12[sel] = base*0.4 + (1-load)*0.3 ; t4 = perf*0.1 ; viability flag."""
    m = metrics_of(a)
    load = min(1.0, m["queue_utilization"] * 0.5 + m["cpu_usage"] * 0.3 + m["memory_usage"] * 0.2)   # router.py:97-105
    t2 = (1 - load) * 0.3
    s12 = [b * 0.4 + t2 for b in _base_options(m["queue_utilization"])]
    t4 = m["success_rate"] * 0.1                                                                     # :128-133
    viable = a["status"] != "busy" and m["queue_utilization"] < load_threshold                       # :68-69, :55-59
    return s12, t4, viable


def lb_metrics_of(a: dict, system_cpu_percent: float = 0.0, system_mem_percent: float = 0.0) -> dict:
    """LoadMetrics as _collect_system_metrics builds them (load_balancer.py:96-111)."""
    m = metrics_of(a)
    return {"cpu_usage": max(system_cpu_percent / 100.0, m["cpu_usage"]),
            "memory_usage": max(system_mem_percent / 100.0, m["memory_usage"]),
            "queue_size": m["queue_size"], "active_tasks": m["active_tasks"], "total_tasks": 0,
            "error_rate": 1 - m["success_rate"]}


def lb_composite(lm: dict, max_tasks_per_agent: int) -> float:                     # load_balancer.py:172-178
    return (0.3 * lm["cpu_usage"] + 0.3 * lm["memory_usage"]
            + 0.2 * (lm["queue_size"] / max_tasks_per_agent) + 0.2 * lm["error_rate"])


def lb_terms(a: dict, lm: dict, max_tasks_per_agent: int, overload_threshold: float):
    comp = lb_composite(lm, max_tasks_per_agent)
    load_score = 1 - comp
    perf_score = 1 - lm["error_rate"]
    resource = 1 - max(lm["cpu_usage"], lm["memory_usage"])
    q_util = metrics_of(a)["queue_utilization"]
    full = [b * 0.4 + load_score * 0.3 + perf_score * 0.2 + resource * 0.1 for b in _base_options(q_util)]   # :306-332
    viable = a["status"] != "stopped" and lm["queue_size"] < max_tasks_per_agent and comp < overload_threshold  # :294-304
    return full, 0.0, viable


def delegator_terms(a: dict):
    m = metrics_of(a)
    load_score = 1 - m["queue_utilization"]
    error_rate = 1 - m["success_rate"]
    resource = 1 - max(m["cpu_usage"], m["memory_usage"])
    full = [b * 0.4 + load_score * 0.3 + (1 - error_rate) * 0.2 + resource * 0.1
            for b in _base_options(m["queue_utilization"])]                          # task_delegator.py:92-113
    mc = a.get("max_concurrent_tasks")
    available = a["status"] not in ("stopped", "error") and a.get("active_delegations", 0) < (5 if mc is None else mc)  # :308-314
    accept = (a["status"] not in ("stopped", "error") and m["queue_utilization"] < 0.8
              and m["cpu_usage"] < 0.8 and m["memory_usage"] < 0.8)                  # :316-326
    return full, 0.0, available and accept


def task_priority_host(task: dict) -> int:
    """TaskRouter.get_task_priority as an integer level (router.py:135-145)."""
    if task.get("urgent", False):
        return 3
    c, d = task.get("complexity", 1), len(task.get("dependencies", []))
    return 2 if (c > 8 or d > 5) else 1 if (c > 5 or d > 3) else 0


# ---- vocabularies and encoding -----------------------------------------------------------------
class Vocab:
    """String -> bit position.  A required capability no agent has is carried as task flag bit 2."""

    def __init__(self, agents: Sequence[dict]):
        self.cap: Dict[str, int] = {}
        self.spec: Dict[str, int] = {}
        for a in agents:
            for c in a.get("caps") or ():
                self.cap.setdefault(c, len(self.cap))
            for s in a.get("specializations") or ():
                self.spec.setdefault(s, len(self.spec))
        self.words = max(1, (len(self.cap) + 63) // 64, (len(self.spec) + 63) // 64)

    def _mask(self, bits: Iterable[int]) -> np.ndarray:
        out = np.zeros(self.words, dtype=np.uint64)
        for b in bits:
            out[b >> 6] |= np.uint64(1) << np.uint64(b & 63)
        return out

    def caps_mask(self, caps) -> np.ndarray:
        return self._mask(self.cap[c] for c in caps or ())

    def spec_mask(self, specs) -> np.ndarray:
        return self._mask(self.spec[s] for s in specs or ())

    def req_mask(self, required) -> np.ndarray:
        return self._mask(self.cap[c] for c in set(required) if c in self.cap)

    def tag_mask(self, tags) -> np.ndarray:
        return self._mask(self.spec[t] for t in set(tags) if t in self.spec)


def encode_tasks(tasks: Sequence[dict], vocab: Vocab) -> Dict[str, np.ndarray]:
    n, w = len(tasks), vocab.words
    req = np.zeros((n, w), np.uint64)
    tags = np.zeros((n, w), np.uint64)
    typ = np.full(n, -1, np.int32)
    flags = np.zeros(n, np.uint8)
    for i, t in enumerate(tasks):
        if "required_capabilities" in t:
            flags[i] |= 1
            req[i] = vocab.req_mask(t["required_capabilities"])
            if any(c not in vocab.cap for c in t["required_capabilities"]):
                flags[i] |= 4                      # a capability nobody has: every agent misses it
        if "type" in t:
            flags[i] |= 2
        typ[i] = vocab.spec.get(t.get("type", ""), -1)
        tags[i] = vocab.tag_mask(t.get("tags", []))
    return {"req": req, "tags": tags, "type": typ, "flags": flags}


try:                                   # C-speed encoder (pilottai_b200/csrc/task_encode.c, built by pilottai_b200.build)
    from . import _taskenc
except ImportError:                    # not built: the Python definition above does the same, ~50x slower
    _taskenc = None


def encode_task_batch(tasks: Sequence[dict], vocab: Vocab) -> Dict[str, np.ndarray]:
    """encode_tasks plus the three get_task_priority inputs, through the C extension when it is there.
    Keys: req, tags, type, flags (what pilott_route_score takes) and urgent, cx, ndeps (what
    pilott_task_priority takes; cx is the complexity class 9 / 6 / 0 for > 8 / > 5 / else, so a
    fractional complexity compares exactly like the reference's float does, router.py:139-142)."""
    n, w = len(tasks), vocab.words
    if _taskenc is not None and isinstance(tasks, list):
        out = {"req": np.zeros((n, w), np.uint64), "tags": np.zeros((n, w), np.uint64), "type": np.empty(n, np.int32),
               "flags": np.empty(n, np.uint8), "urgent": np.empty(n, np.uint8), "cx": np.empty(n, np.int32),
               "ndeps": np.empty(n, np.int32)}
        done = _taskenc.encode(tasks, vocab.cap, vocab.spec, w, out["req"], out["tags"], out["type"], out["flags"],
                               out["urgent"], out["cx"], out["ndeps"])
        if done == n:
            return out
    out = encode_tasks(tasks, vocab)               # a task the fast path declines: the definition handles the batch
    cxs = [t.get("complexity", 1) for t in tasks]
    out["urgent"] = np.array([1 if t.get("urgent", False) else 0 for t in tasks], np.uint8)
    out["cx"] = np.array([9 if c > 8 else 6 if c > 5 else 0 for c in cxs], np.int32)
    out["ndeps"] = np.array([len(t.get("dependencies", [])) for t in tasks], np.int32)
    return out


# ---- the engine ------------------------------------------------------------------------------------
class RoutingEngine:
    """Device-resident agent table + one-call batched scoring/assignment."""

    def __init__(self, agents: Sequence[dict], device: int = 0):
        import torch
        self._torch = torch
        self._lib = _lib.load()
        self.device = torch.device("cuda", device)
        self.agents = list(agents)
        self.vocab = Vocab(self.agents)
        w = self.vocab.words
        na = len(self.agents)
        caps = np.zeros((na, w), np.uint64)
        spec = np.zeros((na, w), np.uint64)
        self._attr_flags = np.zeros(na, np.uint8)
        for i, a in enumerate(self.agents):
            caps[i] = self.vocab.caps_mask(a.get("caps"))
            spec[i] = self.vocab.spec_mask(a.get("specializations"))
            self._attr_flags[i] = (2 if a.get("caps") is not None else 0) | (4 if a.get("specializations") is not None else 0)
        self.max_tags = 64 * w
        u = np.zeros((2, self.max_tags + 1), np.float64)
        for c in range(self.max_tags + 1):
            u[0, c] = (0.1 * c) * 0.2            # router.py:126 then :86
            u[1, c] = 0.3 * 0.2
        self._d = lambda arr: torch.from_numpy(np.ascontiguousarray(arr).view(np.int64) if arr.dtype == np.uint64
                                               else np.ascontiguousarray(arr)).to(self.device)
        self.d_caps, self.d_spec, self.d_u = self._d(caps), self._d(spec), self._d(u)
        self._tables: Dict[tuple, tuple] = {}
        self._ws = None                      # workspace of the tensor-core form (bf16 feature matrices)

    def _table(self, mode: str, **cfg):
        key = (mode, tuple(sorted((k, repr(v)) for k, v in cfg.items() if k != "lb_metrics")), id(cfg.get("lb_metrics")))
        if key in self._tables:
            return self._tables[key]
        na = len(self.agents)
        s12 = np.zeros((na, 3), np.float64)
        t4 = np.zeros(na, np.float64)
        flags = self._attr_flags.copy()
        cand = cfg.get("candidates")
        cand = None if cand is None else set(cand)
        for i, a in enumerate(self.agents):
            if mode == "router":
                s, t, ok = router_terms(a, cfg.get("load_threshold", 0.8))
            elif mode == "load_balancer":
                lm = cfg["lb_metrics"][a["id"]] if cfg.get("lb_metrics") else lb_metrics_of(a)
                s, t, ok = lb_terms(a, lm, cfg.get("max_tasks_per_agent", 10), cfg.get("overload_threshold", 0.8))
            else:
                s, t, ok = delegator_terms(a)
            if cand is not None and a["id"] not in cand:
                ok = False
            s12[i], t4[i] = s, t
            flags[i] |= 1 if ok else 0
        # Only viable agents can win, so only they go to the device (table order is kept: the
        # first maximum among viable agents is the first maximum overall).
        keep = np.nonzero(flags & 1)[0].astype(np.int64)
        tab = (self._d(s12[keep]), self._d(t4[keep]), self._d(flags[keep]), self.d_caps[keep] if keep.size else self.d_caps[:0],
               self.d_spec[keep] if keep.size else self.d_spec[:0], self._torch.from_numpy(keep.astype(np.int32)).to(self.device))
        self._tables[key] = tab
        return tab

    def route_encoded(self, enc: Dict[str, "np.ndarray"], mode: str = "router", kernel: str = "auto", **cfg):
        """enc: arrays from encode_tasks (numpy) or the same keys as CUDA tensors.
        Returns (agent index int32 [T] (-1 = none), score fp64 [T]) as CUDA tensors.
        kernel: "cuda" = route_kernel (set arithmetic on CUDA cores, any vocabulary width), "tc" = the
        tcgen05 GEMM form (pilott_route_score_tc: up to 192-bit vocabularies), "auto" = the faster of
        the two as measured on B200.  Same results from both, bit for bit."""
        torch = self._torch
        d_s12, d_t4, d_flags, d_caps, d_spec, d_keep = self._table(mode, **cfg)
        t = {k: (self._d(v) if isinstance(v, np.ndarray) else v) for k, v in enc.items()}
        nt = t["type"].shape[0]
        out_a = torch.empty(nt, dtype=torch.int32, device=self.device)
        out_s = torch.empty(nt, dtype=torch.float64, device=self.device)
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        na = int(d_keep.numel())
        if kernel == "auto":
            kernel = "tc" if (AUTO_TC and self.vocab.words <= 3 and 0 < na <= 1280 and nt >= AUTO_TC_MIN_TASKS) else "cuda"
        if kernel == "tc":
            need = int(self._lib.pilott_route_tc_workspace(nt, na, self.vocab.words))
            if self._ws is None or self._ws.numel() < need:
                self._ws = torch.empty(need + need // 8, dtype=torch.uint8, device=self.device)
            check(self._lib.pilott_route_score_tc(ptr(t["req"]), ptr(t["tags"]), ptr(t["type"]), ptr(t["flags"]), nt,
                                                  ptr(d_caps), ptr(d_spec), ptr(d_flags), ptr(d_s12), ptr(d_t4), na,
                                                  self.vocab.words, ptr(self.d_u), self.max_tags, MODE[mode], ptr(out_a),
                                                  ptr(out_s), ptr(self._ws), self._ws.numel(), st))
        else:
            check(self._lib.pilott_route_score(ptr(t["req"]), ptr(t["tags"]), ptr(t["type"]), ptr(t["flags"]), nt,
                                               ptr(d_caps), ptr(d_spec), ptr(d_flags), ptr(d_s12), ptr(d_t4),
                                               na, self.vocab.words, ptr(self.d_u), self.max_tags,
                                               MODE[mode], ptr(out_a), ptr(out_s), st))
        if 0 < d_keep.numel() < len(self.agents):      # map compacted positions back to table indices
            out_a = torch.where(out_a >= 0, d_keep[out_a.clamp(min=0).long()], out_a)
        return out_a, out_s

    def encode(self, tasks: Sequence[dict]) -> Dict[str, np.ndarray]:
        """Task dicts -> host columns (C extension when built; see encode_task_batch)."""
        return encode_task_batch(tasks, self.vocab)

    def to_device(self, arr: np.ndarray):
        return self._d(arr)

    def route(self, tasks: Sequence[dict], mode: str = "router", as_numpy: bool = False, kernel: str = "auto", **cfg):
        """Dict tasks -> (agent index, score) per task on the host: lists, or numpy arrays
        (int32 / fp64) with as_numpy=True -- the form a million-task batch wants."""
        if not tasks:
            return (np.empty(0, np.int32), np.empty(0, np.float64)) if as_numpy else ([], [])
        enc = self.encode(tasks)
        a, s = self.route_encoded({k: enc[k] for k in ("req", "tags", "type", "flags")}, mode, kernel, **cfg)
        if as_numpy:
            return a.cpu().numpy(), s.cpu().numpy()
        return a.cpu().tolist(), s.cpu().tolist()

    def priorities(self, tasks: Sequence[dict]) -> List[int]:
        """get_task_priority for a batch, on the device (pilott_task_priority)."""
        torch = self._torch
        if not tasks:
            return []
        enc = self.encode(tasks)
        urgent, cx, nd = self._d(enc["urgent"]), self._d(enc["cx"]), self._d(enc["ndeps"])
        out = torch.empty(len(tasks), dtype=torch.uint8, device=self.device)
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        check(self._lib.pilott_task_priority(ptr(urgent), ptr(cx), ptr(nd), len(tasks), ptr(out), st))
        return out.cpu().tolist()
