"""How the fused search behaves on a BURSTY corpus: rows arrive in runs of `burst` near-duplicates
of one topic vector (what an agent storing related facts one after the other produces), and every
query asks for one of the topics, so a query's whole top-k sits in one or two 256-row tiles -- the
worst case for per-list candidate lists.  Prints fallback counts and time per batch for each
candidate-list length.   usage: python tools/bench_clustered.py [n] [dim] [nq] [burst] [k] [list length of the device-path timing]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pilottai_b200.store import SemanticStore

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 768
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
burst = int(sys.argv[4]) if len(sys.argv) > 4 else 12
k = int(sys.argv[5]) if len(sys.argv) > 5 else 10
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(7)
st = SemanticStore(dim, n, 0)
chunk = 131072 // burst * burst
nchunks = (n + chunk - 1) // chunk
per = (nq + nchunks - 1) // nchunks
qsrc = []
for lo in range(0, n, chunk):
    m = min(chunk, n - lo)
    t = torch.randn((m + burst - 1) // burst, dim, generator=g, device=dev)
    x = t.repeat_interleave(burst, dim=0)[:m] + 0.05 * torch.randn(m, dim, generator=g, device=dev)
    st.append(x / x.norm(dim=1, keepdim=True))
    qsrc.append(t[torch.randint(0, t.shape[0], (per,), generator=g, device=dev)])
# queries: noisy copies of stored topics, so each has `burst` strong neighbours stored side by side
q = torch.cat(qsrc)[:nq]
q = q + 0.02 * torch.randn(q.shape[0], dim, generator=g, device=dev)
q = (q / q.norm(dim=1, keepdim=True)).contiguous()
nq = q.shape[0]
torch.cuda.synchronize()
qh = q.cpu().numpy()
ref = None
for ll in (32, 16, 8, 0):
    st.set_list_len(ll)
    for _ in range(2):
        out = st.search(qh, k)
    t0 = time.perf_counter()
    it = 10
    for _ in range(it):
        out = st.search(qh, k)
    ms = (time.perf_counter() - t0) / it * 1e3
    same = "" if ref is None else f" same_as_32={bool((out[0] == ref).all())}"
    if ref is None:
        ref = out[0].copy()
    print(f"list_len={ll:2d}: {ms:8.3f} ms/batch through host buffers, {st.last_fallbacks} of {nq} queries re-run by the fallback{same}", flush=True)

# device entry points, timed separately with CUDA events: fused search, then the repair step
st.set_list_len(int(sys.argv[6]) if len(sys.argv) > 6 else 8)
idx, sc, sc64, inexact, flags = st.search(q, k, want_score64=True)
for name in ("fused", "refine"):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(5):
        if name == "fused":
            e0.record(); st.search_into(q, k, idx, sc, sc64, inexact, flags); e1.record()
        else:
            st.search_into(q, k, idx, sc, sc64, inexact, flags)
            ids = flags.nonzero()
            e0.record(); nb = st.refine(q, k, ids, (idx, sc, sc64)); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    extra = "" if name == "fused" else f" ({ids.numel()} flagged, {nb} to brute force)"
    print(f"device {name}: {tot / 5:8.3f} ms{extra}", flush=True)
