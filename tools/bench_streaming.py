"""BASELINE.json configs[4]: mixed streaming -- store_semantic inserts interleaved with top-10
query batches, on one GPU or (under torchrun) on a row-sharded store over several.

Schedule (deterministic, so that every rank of a sharded run issues the same collectives):
phase 1 runs `nb` query batches alone and measures the query rate; phase 2 runs `nb` query batches
again and appends one 1000-row insert batch every `period` query batches, `period` chosen from the
phase-1 rate so that the insert rate lands on the target.  Inserts go through the device append
entry point; the sharded store routes rows to ranks with no collective (every rank sees the batch
and keeps the rows it owns).  Calls on one handle are serialised by the caller, so a query batch
always sees a consistent prefix of the inserts.
usage: [torchrun --nproc-per-node N] python tools/bench_streaming.py [rows] [dim] [insert_rate_per_s] [query_batches]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from pilottai_b200.sharded import ShardedSemanticStore

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
rate = float(sys.argv[3]) if len(sys.argv) > 3 else 100_000.0
nb = int(sys.argv[4]) if len(sys.argv) > 4 else 1500
NQ, K, BATCH = 1024, 10, 1000
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
sys.stdout.flush(); json_fd = os.dup(1); os.dup2(2, 1)          # keep library banners off stdout
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
g = torch.Generator(device=dev); g.manual_seed(1234)
st = ShardedSemanticStore(dim, n, rank, world, local)           # ring: streaming inserts evict the oldest rows
for lo in range(0, n, 131072):
    m = min(131072, n - lo)
    x = torch.randn(m, dim, generator=g, device=dev); st.append(x / x.norm(dim=1, keepdim=True))
q = torch.randn(8, NQ, dim, generator=g, device=dev); q = q / q.norm(dim=2, keepdim=True)
g.manual_seed(99)
ins = torch.randn(16, BATCH, dim, generator=g, device=dev)      # insert batches (cycled)


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def run(period):
    pending, ni = [], 0
    for i in range(4):
        st.search_end(st.search_begin(q[i % 8], K, slot=i % 2))
    sync()
    t0 = time.perf_counter()
    for i in range(nb):
        if period and i % period == 0:
            st.append(ins[ni % 16]); ni += 1
        pending.append(st.search_begin(q[i % 8], K, slot=i % 2))
        if len(pending) > 1:
            st.search_end(pending.pop(0))
        if i % 16 == 15:
            torch.cuda.synchronize()                             # bound the launch queue
    while pending:
        st.search_end(pending.pop(0))
    sync()
    dt = time.perf_counter() - t0
    flagged = sum(int(b["inexact"].item()) for b in st._bufs.values())
    return {"queries_per_s": nb * NQ / dt, "inserts_per_s": ni * BATCH / dt, "query_batches": nb, "insert_batches": ni,
            "flagged_in_last_batches": flagged}


base = run(0)
period = max(1, round((base["queries_per_s"] / NQ) / (rate / BATCH)))
mixed = run(period)
if rank == 0:
    os.write(json_fd, (json.dumps({
        "workload": f"streaming mix: one {BATCH}-row insert batch every {period} query batches (target {rate:.0f} inserts/s) "
                    f"interleaved with {NQ}-query top-{K} batches, {n} x {dim} ring on {world} GPU(s)",
        "queries_only": base, "mixed": mixed, "query_throughput_kept": mixed["queries_per_s"] / base["queries_per_s"],
        "rows_ever_appended": st.total}) + "\n").encode())
st.close()
if world > 1:
    dist.destroy_process_group()
