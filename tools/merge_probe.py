"""topk_merge_kernel alone on one GPU: [shards, 2, nq, k] packed blocks -> global top-k (what every rank
runs after the all-gather).  usage: python tools/merge_probe.py [shards] [nq] [k] [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pilottai_b200.store import topk_merge_packed
shards, nq, k, iters = (int(sys.argv[i]) if len(sys.argv) > i else d for i, d in ((1, 8), (2, 1024), (3, 10), (4, 50)))
g = torch.Generator(device="cuda"); g.manual_seed(1)
sc = torch.rand(shards, nq, k, generator=g, device="cuda", dtype=torch.float64).sort(dim=2, descending=True).values
ix = torch.randperm(shards * nq * k, generator=g, device="cuda").view(shards, nq, k)
packed = torch.stack([ix, sc.view(torch.int64)], dim=1).contiguous()
out = topk_merge_packed(packed)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    topk_merge_packed(packed, out=out)
e1.record(); torch.cuda.synchronize()
flat_s, flat_i = sc.permute(1, 0, 2).reshape(nq, -1), ix.permute(1, 0, 2).reshape(nq, -1)
want = flat_s.topk(k, dim=1)
assert torch.equal(out[2], want.values) and torch.equal(out[0], flat_i.gather(1, want.indices))
print(f"topk_merge {shards} shards x {nq} queries x k={k}: {e0.elapsed_time(e1) / iters * 1e3:.1f} us per launch (checked against torch.topk)")
