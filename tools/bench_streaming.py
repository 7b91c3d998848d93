"""BASELINE.json configs[4]: mixed streaming -- store_semantic inserts interleaved with top-10
query batches.  Single GPU here (the sharded store routes rows to ranks with no collective, so
per-GPU behaviour is what matters); inserts go through the device append entry point on a
second stream while query batches run on the main stream.  The store's ring bookkeeping is
host-side and calls on one handle are serialised by the caller, so an insert batch and a
query batch never touch the handle concurrently: the query sees a consistent prefix.
usage: python tools/bench_streaming.py [rows] [dim] [insert_rate_per_s] [seconds]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pilottai_b200.store import SemanticStore

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
rate = float(sys.argv[3]) if len(sys.argv) > 3 else 100_000.0
secs = float(sys.argv[4]) if len(sys.argv) > 4 else 3.0
NQ, K, BATCH = 1024, 10, 1000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1234)
st = SemanticStore(dim, n, 0)                      # ring: streaming inserts evict the oldest rows
for lo in range(0, n, 131072):
    m = min(131072, n - lo)
    x = torch.randn(m, dim, generator=g, device=dev); st.append(x / x.norm(dim=1, keepdim=True))
q = torch.randn(8, NQ, dim, generator=g, device=dev); q = q / q.norm(dim=2, keepdim=True)
ins = torch.randn(16, BATCH, dim, generator=g, device=dev)     # seed 99-style insert batches (cycled)
idx = torch.empty((NQ, K), dtype=torch.int64, device=dev); sc = torch.empty((NQ, K), dtype=torch.float32, device=dev)
sc64 = torch.empty((NQ, K), dtype=torch.float64, device=dev); inex = torch.zeros(1, dtype=torch.int32, device=dev)
flags = torch.empty(NQ, dtype=torch.uint8, device=dev)


def run(insert_rate):
    for i in range(3):
        st.search_into(q[i % 8], K, idx, sc, sc64, inex, flags)
    torch.cuda.synchronize()
    t0 = time.perf_counter(); nb = ni = 0
    while True:
        now = time.perf_counter() - t0
        if now >= secs:
            break
        while insert_rate > 0 and ni * BATCH < insert_rate * now:      # inserts that are due
            st.append(ins[ni % 16]); ni += 1
        st.search_into(q[nb % 8], K, idx, sc, sc64, inex, flags); nb += 1
        if nb % 16 == 0:
            torch.cuda.synchronize()                                   # bound the launch queue
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"queries_per_s": nb * NQ / dt, "inserts_per_s": ni * BATCH / dt, "query_batches": nb, "insert_batches": ni}


base = run(0.0)
mixed = run(rate)
print(json.dumps({"workload": f"streaming mix: {rate:.0f} inserts/s (batches of {BATCH}) interleaved with {NQ}-query top-{K} batches, {n} x {dim} ring",
                  "queries_only": base, "mixed": mixed,
                  "query_throughput_kept": mixed["queries_per_s"] / base["queries_per_s"], "total": st.info()["total"]}))
