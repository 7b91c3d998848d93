"""Rows built to stress the error bound the exactness proof rests on.

The fused path ranks by an APPROXIMATE cosine (bf16 products, fp32 accumulation on the tensor cores,
fp32 inverse norm) and proves its answer with `approximate + err < exact k-th`, err =
max(1e-4, 2.4e-7 * row_stride) (DESIGN.md K4 derives it: the accumulation error is relative to
|q| |m|, not to the dot product).  If the bound were too small somewhere, a wrong row could be
returned WITHOUT being flagged; so every case here is compared with the oracle bit for bit:
heavy cancellation, 1e+-15 dynamic range inside a row, dimension 2048, rows that are nearly equal
in score to the k-th, large and small (but representable) norms, and the expiry boundary.
"""
import numpy as np
import pytest

from oracle import retrieval_c
from oracle.retrieval_oracle import OracleStore

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _both(x, dim, cap, **meta):
    from pilottai_b200.store import SemanticStore
    gs, os_ = SemanticStore(dim, cap, 0), OracleStore(dim, cap)
    for lo in range(0, x.shape[0], 4096):
        sl = slice(lo, min(x.shape[0], lo + 4096))
        m = {k: v[sl] for k, v in meta.items()}
        gs.append(x[sl], **m)
        os_.append(x[sl], **m)
    return gs, os_


def _same(gs, os_, q, k, **flt):
    idx, sc, sc64 = gs.search(q, k, want_score64=True, **flt)
    oidx, osc = retrieval_c.search(os_, q, k, **flt)
    assert np.array_equal(idx, oidx), np.nonzero((idx != oidx).any(axis=1))[0][:5]
    assert np.array_equal(sc64, osc)
    return gs.last_fallbacks


@pytest.mark.parametrize("dim", [768, 2048])
def test_cancellation_and_dynamic_range(torch_cuda, dim):
    rng = np.random.default_rng(dim)
    n = 20_000
    x = rng.standard_normal((n, dim)).astype(np.float32)
    # heavy cancellation against the queries: +a / -a pairs, so the dot product is tiny next to |q||m|
    x[:4000, 1::2] = -x[:4000, 0::2]
    # 1e+15 .. 1e-15 inside one row (the small half vanishes in the fp32 norm, not in the fp64 score)
    scale = np.exp(rng.uniform(np.log(1e-15), np.log(1e15), (4000, dim))).astype(np.float32)
    x[4000:8000] *= scale / scale.max(axis=1, keepdims=True) * 1e15
    # rows that differ from each other in a few low-order coordinates: scores closer than err
    base = rng.standard_normal(dim).astype(np.float32)
    x[8000:9000] = base + 1e-3 * rng.standard_normal((1000, dim)).astype(np.float32)
    # large and small norms well inside fp32's range
    x[9000:9500] *= 1e12
    x[9500:10000] *= 1e-12
    gs, os_ = _both(x, dim, n)
    q = rng.standard_normal((96, dim)).astype(np.float32)
    q[:16, 1::2] = q[:16, 0::2]                       # makes the cancelling rows cancel exactly-ish
    q[16:32] = base + 1e-3 * rng.standard_normal((16, dim)).astype(np.float32)   # top-k inside the near-tie cluster
    q[32:40] = x[4000:4008]                           # queries with the wide dynamic range themselves
    q[40:48] *= 1e10
    repaired = _same(gs, os_, q, 10)
    _same(gs, os_, q, 32)
    _same(gs, os_, q[:20], 100)
    assert repaired >= 8, "the near-tie cluster must be caught by the gap check, not slip through"
    gs.close()


def test_expiry_boundary_on_the_device(torch_cuda):
    """now_us == expires_us is kept (strict '>' in MemoryItem.is_expired, enhanced_memory.py:20-21);
    one microsecond later the row is gone -- on the fused path, the brute-force path and in compact()."""
    torch = torch_cuda
    rng = np.random.default_rng(3)
    n, dim, now = 3000, 128, 5_000_000
    x = rng.standard_normal((n, dim)).astype(np.float32)
    exp = np.zeros(n, np.int64)
    exp[0::3] = now          # kept at `now`
    exp[1::3] = now - 1      # expired
    gs, os_ = _both(x, dim, n, expires_us=exp)
    q = x[:30].copy()        # each query's best hit is a row of every expiry class in turn
    _same(gs, os_, q, 5, now_us=now)
    idx, _ = gs.search(q, 1, now_us=now)
    assert [int(i) for i in idx[:, 0][0::3]] == list(range(0, 30, 3))            # rows expiring exactly now: found
    assert all(int(i) != j for j, i in zip(range(1, 30, 3), idx[:, 0][1::3]))    # rows one us past: not found
    qd = torch.from_numpy(q).cuda()
    e_idx, _, e_sc = gs.search_exact(qd, 5, now_us=now)
    oidx, osc = retrieval_c.search(os_, q, 5, now_us=now)
    assert np.array_equal(e_idx.cpu().numpy(), oidx) and np.array_equal(e_sc.cpu().numpy(), osc)
    kept, keep = gs.compact(now)
    assert kept == n - len(exp[1::3]) and not keep[1::3].any() and keep[0::3].all() and keep[2::3].all()
    gs.close()
