"""Host-side cost of one sharded search step (enqueue only, no waiting for the GPU), per exchange.
usage: torchrun --nproc-per-node N tools/host_overhead.py [rows]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from pilottai_b200.sharded import ShardedSemanticStore

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
g = torch.Generator(device=dev); g.manual_seed(1)
for ex in ("nccl", "p2p"):
    st = ShardedSemanticStore(768, n, rank, world, local, exchange=ex)
    for lo in range(0, n, 131072):
        m = min(131072, n - lo); x = torch.randn(m, 768, generator=g, device=dev); st.append(x / x.norm(dim=1, keepdim=True))
    q = torch.randn(1024, 768, generator=g, device=dev)
    for i in range(6):
        st.search_end(st.search_begin(q, 10, slot=i % 2))
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    steps, pending = 40, []
    t0 = time.perf_counter()
    for i in range(steps):
        pending.append(st.search_begin(q, 10, slot=i % 2))
        if len(pending) > 1:
            st.search_end(pending.pop(0))
    st.search_end(pending.pop(0))
    t_enq = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    if rank == 0:
        print(f"exchange {ex}: host enqueue {t_enq / steps * 1e6:.0f} us per step, wall {t_all / steps * 1e6:.0f} us per step", flush=True)
    st.close()
dist.destroy_process_group()
