"""pilott_mem_search_exact_noalloc -- the exact search for callers that cannot allocate.

Some callers of this library must not reach the CUDA allocator while they run: PilottAI's
asyncio loop thread (cudaMalloc synchronises the whole device and would freeze every other
coroutine's stream), or a process where PyTorch's caching allocator already holds HBM so a
raw cudaMalloc fails.  For them the library keeps a fixed 16 MiB scratch array in its own
device image and offers this entry point, which only launches kernels.

Checked here, through the C ABI: two calls in a row (the second with other queries, so stale
scratch from the first would show) each return exactly what the allocating entry point
pilott_mem_search_exact returns -- indices equal, fp64 scores bit-identical, the same bar as
tests/test_gpu_retrieval.py -- without the device's free memory moving (the array is loaded
when the store is created, not by the call), and over-sized requests are refused with a clear message.
Collected last (file name) so a `-x` run has already been through every other test.
"""
import numpy as np
import pytest

from pilottai_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _outs(torch, nq, k):
    return (torch.full((nq, k), -7, dtype=torch.int64, device="cuda"),
            torch.full((nq, k), 7.0, dtype=torch.float32, device="cuda"),
            torch.full((nq, k), 7.0, dtype=torch.float64, device="cuda"))


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(torch_cuda):
    torch = torch_cuda
    from pilottai_b200.store import SemanticStore
    n, dim, k = 6000, 200, 12                      # dim not a multiple of 64: padded row stride
    x = synth.corpus(n, dim, seed=31, dup_frac=0.01)
    prio, tag, exp = synth.metadata(n, seed=32)
    st = SemanticStore(dim, 4096, 0)               # ring wraps: 6000 rows into 4096 slots
    st.append(x[:3000], prio[:3000], tag[:3000], exp[:3000])
    st.append(x[3000:], prio[3000:], tag[3000:], exp[3000:])
    flt = dict(min_priority=1, req_tagmask=0b10, now_us=1_000_000)
    for call, (nq, seed, ids) in enumerate([(9, 41, None), (5, 42, [4, 0, 2])]):
        q = torch.from_numpy(synth.queries(nq, dim, seed=seed)).cuda()
        qids = None if ids is None else torch.tensor(ids, dtype=torch.int32, device="cuda")
        want = st.search_exact(q, k, query_ids=qids, out=_outs(torch, nq, k), **flt)
        bufs = _outs(torch, nq, k)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]           # cudaMemGetInfo: the driver's own count
        got = st.search_exact_noalloc(q, k, bufs, query_ids=qids, **flt)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info(0)[0] == free_before, f"call {call} changed the device's free memory"
        for w, g, name in zip(want, got, ("index", "score", "score64")):
            assert torch.equal(w, g), f"call {call}: {name} differs from pilott_mem_search_exact"
        rows = list(range(nq)) if ids is None else ids
        assert (got[0][rows] >= -1).all() and (got[0][rows, 0] >= 0).all()      # really written
    st.close()


def test_noalloc_rejects_requests_beyond_the_fixed_scratch(torch_cuda):
    torch = torch_cuda
    from pilottai_b200 import _lib
    from pilottai_b200.store import SemanticStore
    st = SemanticStore(64, 50_000, 0)
    st.append(synth.corpus(50_000, 64, seed=5))
    nq = 64                                        # 64 queries x 50 000 rows x 8 B = 25.6 MB > 16 MiB
    q = torch.from_numpy(synth.queries(nq, 64, seed=6)).cuda()
    with pytest.raises(_lib.PilottError) as ei:
        st.search_exact_noalloc(q, 5, _outs(torch, nq, 5))
    assert ei.value.code == _lib.EINVAL
    assert str(_lib.NOALLOC_SCRATCH_BYTES) in str(ei.value) and "scratch" in str(ei.value)
    many = torch.zeros(65536, dtype=torch.int32, device="cuda")  # more selected queries than one launch grid takes
    with pytest.raises(_lib.PilottError) as ei:
        st.search_exact_noalloc(q, 5, _outs(torch, nq, 5), query_ids=many)
    assert ei.value.code == _lib.EINVAL and "65535" in str(ei.value)
    ok = torch.arange(8, dtype=torch.int32, device="cuda")      # a subset that fits is served
    got = st.search_exact_noalloc(q, 5, _outs(torch, nq, 5), query_ids=ok)
    want = st.search_exact(q, 5, query_ids=ok, out=_outs(torch, nq, 5))
    torch.cuda.synchronize()
    assert torch.equal(got[0][:8], want[0][:8]) and torch.equal(got[2][:8], want[2][:8])
    st.close()
