"""BASELINE.json configs[3]: router / load-balancer scoring, 1M pending tasks x 1024 agents,
overload_threshold 0.8.  Prints one JSON line (secondary benchmark; bench.py is the contract one).
usage: python tools/bench_router.py [n_tasks] [n_agents] [iters]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pilottai_b200 import synth
from pilottai_b200.routing import RoutingEngine, lb_metrics_of

nt = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
na = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
agents = synth.agent_table(na, seed=42)
eng = RoutingEngine(agents, 0)
w = eng.vocab.words
rng = np.random.default_rng(43)
# encoded synthetic tasks (same distribution as synth.task_table, generated vectorised):
# 0-3 required capabilities out of 64, one type, 0-4 tags
req = np.zeros((nt, w), np.uint64); tags = np.zeros((nt, w), np.uint64)
ncap, nspec = len(eng.vocab.cap), len(eng.vocab.spec)
for _ in range(3):
    on = rng.random(nt) < 0.5
    bit = rng.integers(0, ncap, nt)
    np.bitwise_or.at(req, (np.nonzero(on)[0], bit[on] >> 6), np.uint64(1) << (bit[on] & 63).astype(np.uint64))
for _ in range(4):
    on = rng.random(nt) < 0.5
    bit = rng.integers(0, nspec, nt)
    np.bitwise_or.at(tags, (np.nonzero(on)[0], bit[on] >> 6), np.uint64(1) << (bit[on] & 63).astype(np.uint64))
typ = rng.integers(-1, nspec, nt).astype(np.int32)
flags = (rng.random(nt) < 0.85).astype(np.uint8) | ((rng.random(nt) < 0.9).astype(np.uint8) << 1)
dev = torch.device("cuda:0")
enc = {"req": torch.from_numpy(req.view(np.int64)).to(dev), "tags": torch.from_numpy(tags.view(np.int64)).to(dev),
       "type": torch.from_numpy(typ).to(dev), "flags": torch.from_numpy(flags).to(dev)}
out = {}
for mode, cfg in (("router", {"load_threshold": 0.8}),
                  ("load_balancer", {"lb_metrics": {a["id"]: lb_metrics_of(a) for a in agents}, "max_tasks_per_agent": 60, "overload_threshold": 0.8}),
                  ("delegator", {})):
    for _ in range(3):
        a, s = eng.route_encoded(enc, mode, **cfg)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        a, s = eng.route_encoded(enc, mode, **cfg)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    out[mode] = {"ms": ms, "routes_per_s": nt / (ms * 1e-3), "pair_evals_per_s": nt * na / (ms * 1e-3),
                 "assigned_frac": float((a >= 0).float().mean().item())}
# end-to-end: host-encoded numpy tasks in, host results out (H2D + D2H inside)
encn = {"req": req, "tags": tags, "type": typ, "flags": flags}
t0 = time.perf_counter()
for _ in range(3):
    a, s = eng.route_encoded(encn, "router", load_threshold=0.8); a.cpu(); s.cpu()
e2e = nt * 3 / (time.perf_counter() - t0)
print(json.dumps({"workload": f"router scoring {nt} tasks x {na} agents, overload/load threshold 0.8", "words": w,
                  "modes": out, "e2e_router_routes_per_s": e2e,
                  "algorithmic_bytes_per_task": 16 * w + 5 + 12}))
