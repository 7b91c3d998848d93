"""Timing sweep over shapes and filters, to find performance cliffs away from the headline shape.
Prints one line per case: ms per batch through the device entry point, flagged queries, and the
effective TFLOP/s and GB/s (corpus bytes / time).   usage: python tools/sweep_shapes.py [quick]"""
import itertools, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pilottai_b200.store import SemanticStore

quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(3)


def fill(n, dim, cap):
    st = SemanticStore(dim, cap, 0)
    for lo in range(0, n, 65536):
        m = min(65536, n - lo)
        x = torch.randn(m, dim, generator=g, device=dev)
        prio = torch.randint(0, 4, (m,), generator=g, device=dev, dtype=torch.int32)
        tags = torch.zeros(m, dtype=torch.int64, device=dev)
        for b in range(8):
            tags |= (torch.rand(m, generator=g, device=dev) < 0.1).long() << b
        st.append(x / x.norm(dim=1, keepdim=True), prio, tags)
    return st


def timed(st, q, k, iters, **flt):
    nq = q.shape[0]
    idx = torch.empty((nq, k), dtype=torch.int64, device=dev); sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
    sc64 = torch.empty((nq, k), dtype=torch.float64, device=dev)
    inex = torch.zeros(1, dtype=torch.int32, device=dev); flags = torch.empty(nq, dtype=torch.uint8, device=dev)
    for _ in range(2):
        st.search_into(q, k, idx, sc, sc64, inex, flags, **flt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        st.search_into(q, k, idx, sc, sc64, inex, flags, **flt)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters, int(inex.item())


shapes = [(1_000_000, 768, 1_000_000), (1_000_000, 128, 1_000_000), (300_000, 1536, 300_000),
          (100_000, 768, 100_000), (10_000, 768, 10_000), (1_500_000, 384, 1_000_000), (200_000, 4096, 200_000)]
if quick:
    shapes = shapes[:1]
filters = [("none", {}), ("prio>=3", {"min_priority": 3}), ("2 tags", {"req_tagmask": 0b11}), ("4 tags", {"req_tagmask": 0b1111})]
print(f"{'rows':>9s} {'dim':>5s} {'nq':>5s} {'k':>3s} {'filter':>8s} {'ms':>9s} {'flag':>5s} {'TFLOP/s':>8s} {'GB/s':>7s}")
for n, dim, cap in shapes:
    st = fill(n, dim, cap)
    rows = min(n, cap)
    for nq, k in itertools.product((1, 37, 256, 1024, 3000), (1, 10, 32)):
        if quick and (nq, k) not in ((1024, 10), (37, 10), (3000, 32)):
            continue
        q = torch.randn(nq, dim, generator=g, device=dev)
        for name, flt in filters:
            if name != "none" and k != 10:
                continue
            ms, fl = timed(st, q, k, 5 if nq >= 1024 else 10, **flt)
            print(f"{rows:9d} {dim:5d} {nq:5d} {k:3d} {name:>8s} {ms:9.3f} {fl:5d} {2.0 * rows * dim * nq / ms / 1e9:8.1f} {rows * st.info()['row_stride'] * 2 / ms / 1e6:7.0f}", flush=True)
    st.close()
