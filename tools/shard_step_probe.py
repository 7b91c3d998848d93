"""Where a row-sharded search step spends its time (developer probe; run under torch.distributed.run).
usage: torchrun --nproc-per-node N tools/shard_step_probe.py [rows] [dim] [steps]
Prints, per variant, device ms/step (CUDA events, max over ranks) and host enqueue ms/step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from pilottai_b200.sharded import ShardedSemanticStore

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
NQ, K = 1024, 10


def build(exchange):
    st = ShardedSemanticStore(dim, rows, rank, world, local, exchange=exchange)
    g = torch.Generator(device=dev); g.manual_seed(1234)
    for lo in range(0, rows, 131072):
        m = min(131072, rows - lo)
        x = torch.randn(m, dim, generator=g, device=dev)
        st.append(x / x.norm(dim=1, keepdim=True))
    return st


g = torch.Generator(device=dev); g.manual_seed(5678)
qpool = torch.randn(8, NQ, dim, generator=g, device=dev)
qpool = qpool / qpool.norm(dim=2, keepdim=True)


def run(st, nslot, local_only=False):
    pending = []
    loc = st.local
    idx = torch.empty((NQ, K), dtype=torch.int64, device=dev); sc = torch.empty((NQ, K), dtype=torch.float32, device=dev)
    sc64 = torch.empty((NQ, K), dtype=torch.float64, device=dev); fl = torch.empty(NQ, dtype=torch.uint8, device=dev)
    ix = torch.zeros(1, dtype=torch.int32, device=dev)

    def step(i):
        if local_only or world == 1:
            loc.search_into(qpool[i % 8], K, idx, sc, sc64, ix, fl)
            return
        pending.append(st.search_begin(qpool[i % 8], K, slot=i % nslot))
        if len(pending) >= nslot:
            t = pending.pop(0)
            if not os.environ.get("PROBE_NO_MERGE"):
                st.search_end(t, wait=False)

    for i in range(10):
        step(i)
    while pending:
        st.search_end(pending.pop(0))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    while pending:
        st.search_end(pending.pop(0))
    host = time.perf_counter() - t0
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps, host * 1e3 / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.tolist()


# (the peer-memory exchange allows ONE search_begin ahead of its search_end: two slots only)
variants = [("copy", "1", 4, 0), ("copy", "1", 2, 0), ("nccl", "1", 4, 0), ("nccl", "1", 4, 140), ("p2p", "1", 2, 0)] if world > 1 else [("nccl", "0", 2, 0)]
if len(sys.argv) > 4:
    variants = [v for v in variants if v[0] in sys.argv[4].split(",")]
for exchange, ctas, nslot, scan_ctas in variants:
    os.environ["PILOTT_NCCL_MAX_CTAS"] = ctas
    st = build(exchange)
    if scan_ctas:
        st.local.set_ctas(scan_ctas)
    # a broken exchange must not cost 2 s per step: three steps first, then ask
    tk = [st.search_begin(qpool[i], K, slot=i) for i in range(2)]
    for t in tk:
        st.search_end(t)
    try:
        st.check_exchange()
    except Exception as e:
        if rank == 0:
            print(f"world {world} exchange {exchange}: BROKEN ({e}), skipped", flush=True)
        st.close()
        continue
    for label, kw in (("local search only", dict(nslot=2, local_only=True)), (f"{nslot} slots", dict(nslot=nslot))):
        d, h = run(st, **kw)
        if rank == 0:
            print(f"world {world} rows {rows} exchange {st.exchange:5s} nccl_ctas {ctas} scan_ctas {scan_ctas or 'all'} {label:18s}: device {d:.4f} ms/step, host enqueue {h:.4f} ms/step", flush=True)
    st.close()
if world > 1:
    dist.destroy_process_group()
