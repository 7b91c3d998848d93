"""Developer timing of each scan kernel on the headline shape (not the contract bench).
usage: python tools/quick_bench.py [n] [dim] [nq] [iters] [kernels,comma]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pilottai_b200.store import SemanticStore

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 10
kernels = sys.argv[5].split(",") if len(sys.argv) > 5 else ["generic", "tc", "tc_pair"]
k = int(sys.argv[6]) if len(sys.argv) > 6 else 10
torch.cuda.init()
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1234)
st = SemanticStore(dim, n, 0)
st.set_finish_form(int(os.environ.get("QB_FINISH", "0")))
t0 = time.time()
for lo in range(0, n, 131072):
    m = min(131072, n - lo)
    x = torch.randn(m, dim, generator=g, device=dev)
    x = x / x.norm(dim=1, keepdim=True)
    st.append(x)
torch.cuda.synchronize()
print(f"fill {n}x{dim}: {time.time()-t0:.2f}s", flush=True)
q = torch.randn(nq, dim, generator=g, device=dev); q = q / q.norm(dim=1, keepdim=True)
idx = torch.empty((nq, k), dtype=torch.int64, device=dev); sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
sc64 = torch.empty((nq, k), dtype=torch.float64, device=dev)
inex = torch.zeros(1, dtype=torch.int32, device=dev); flags = torch.empty(nq, dtype=torch.uint8, device=dev)
ref = None
for kern in kernels:
    try:
        dbg = 0; ll = 0
        if "/" in kern:                      # tc_pair/16 forces the candidate-list length
            kern, ll = kern.split("/"); 
            if ":" in ll: ll, d2 = ll.split(":"); kern = kern + ":" + d2
            ll = int(ll)
        if ":" in kern:
            kern, dbg = kern.split(":"); dbg = int(dbg)
        st.set_list_len(ll)
        if dbg:
            st.set_debug(dbg)
        st.set_kernel(kern)
        for _ in range(2):
            st.search_into(q, k, idx, sc, sc64, inex, flags)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            st.search_into(q, k, idx, sc, sc64, inex, flags)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        tf = 2.0 * n * dim * nq / (ms * 1e-3) / 1e12
        same = "" if ref is None else f" same_as_{kernels[0]}={bool((idx == ref).all())}"
        if ref is None:
            ref = idx.clone()
        print(f"{kern:8s} dbg={dbg} list={ll}: {ms:8.3f} ms/batch  {nq/(ms*1e-3):12.0f} q/s  {tf:7.1f} TFLOP/s  inexact={int(inex.item())}{same}", flush=True)
    except Exception as ex:
        print(f"{kern}: FAILED {ex}", flush=True)
        break
# append throughput (K1)
nb = min(131072, n)
xb = torch.randn(nb, dim, generator=g, device=dev)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    st.append(xb)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
byts = nb * (4 * dim + 2 * st.info()["row_stride"] + 32)
print(f"append  : {ms:8.3f} ms / {nb} rows  {nb/(ms*1e-3):12.0f} rows/s  {byts/(ms*1e-3)/1e9:7.1f} GB/s")
