"""K5 in its two forms on BASELINE configs[3] (1M tasks x 1024 agents): route_kernel (CUDA cores) against
route_tc_kernel (tcgen05), for 64-bit and 192-bit vocabularies.  usage: python tools/bench_router_forms.py [n_tasks] [iters]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pilottai_b200 import synth
from pilottai_b200.routing import RoutingEngine

nt = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
out = {}
for label, extra in (("64-bit vocabulary (1 word)", 0), ("192-bit vocabulary (3 words)", 120)):
    rng = np.random.default_rng(3)
    agents = synth.agent_table(1024, seed=42)
    xc, xs = [f"xc{i}" for i in range(extra)], [f"xs{i}" for i in range(extra)]
    for a in agents:
        if a["caps"] is not None:
            a["caps"] = a["caps"] + [c for c in xc if rng.random() < 0.3]
        if a["specializations"] is not None:
            a["specializations"] = a["specializations"] + [s for s in xs if rng.random() < 0.1]
    tasks = synth.task_table_fast(nt, seed=43)
    eng = RoutingEngine(agents, 0)
    enc = eng.encode(tasks)
    dev = {k: eng.to_device(enc[k]) for k in ("req", "tags", "type", "flags")}
    res = {"words": eng.vocab.words}
    ref = None
    for kern in ("cuda", "tc"):
        for _ in range(2):
            a, s = eng.route_encoded(dev, "router", kern, load_threshold=0.8)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            a, s = eng.route_encoded(dev, "router", kern, load_threshold=0.8)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        res[kern] = {"ms": ms, "routes_per_s": nt / (ms * 1e-3)}
        if ref is None:
            ref = (a.clone(), s.clone())
        else:
            res["same_assignments"] = bool(torch.equal(a, ref[0]))
            res["same_scores"] = bool(torch.equal(s.view(torch.int64), ref[1].view(torch.int64)))
    out[label] = res
print(json.dumps(out))
