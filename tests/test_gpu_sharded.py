"""GPU tests of the sharded path: the merge kernel (K6) on one GPU, and -- when the box has
two or more GPUs -- the full path (one process per GPU) with both exchanges, the NVLink peer-memory
one and the NCCL one, against the CPU oracle."""
import os
import socket

import numpy as np
import pytest

from oracle import retrieval_c
from oracle.retrieval_oracle import OracleStore
from pilottai_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def test_topk_merge_kernel_matches_host_rule(torch_cuda):
    torch = torch_cuda
    from pilottai_b200.store import topk_merge
    rng = np.random.default_rng(0)
    shards, nq, k = 8, 300, 10
    sc = np.round(rng.standard_normal((shards, nq, k)), 1)            # coarse -> many score ties
    idx = rng.permutation(shards * nq * k).reshape(shards, nq, k).astype(np.int64)
    sc = -np.sort(-sc, axis=2)
    pad = rng.random((shards, nq, k)) < 0.15                          # ragged shards
    pad = np.sort(pad, axis=2)
    idx[pad], sc[pad] = -1, -np.inf
    pad[:, 5] = True; idx[:, 5] = -1; sc[:, 5] = -np.inf              # a query with no result anywhere
    gi, gs, gs64 = topk_merge(torch.from_numpy(idx).cuda(), torch.from_numpy(sc).cuda())
    gi, gs64 = gi.cpu().numpy(), gs64.cpu().numpy()
    for q in range(nq):
        ci, cs = idx[:, q].ravel(), sc[:, q].ravel()
        ok = ci >= 0
        order = np.lexsort((ci[ok], -cs[ok]))[:k]
        want_i = np.full(k, -1, np.int64); want_s = np.full(k, -np.inf)
        want_i[: order.size], want_s[: order.size] = ci[ok][order], cs[ok][order]
        assert np.array_equal(gi[q], want_i) and np.array_equal(gs64[q], want_s)


def _worker(rank, world, port, n, dim, nq, k, out, exchange):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from pilottai_b200.sharded import ShardedSemanticStore
    x = synth.corpus(n, dim, seed=3, dup_frac=0.01)
    prio, tag, exp = synth.metadata(n, seed=5)
    q = torch.from_numpy(synth.queries(nq, dim, seed=4)).cuda()
    fallback = exchange == "copy-fallback"               # one rank fails to map its peers: everybody must end up on NCCL
    if fallback:
        os.environ["PILOTT_XCHG_FORCE_FAIL"] = "1"
        exchange = "copy"
    st = ShardedSemanticStore(dim, n, rank, world, rank, exchange=exchange)
    for lo in range(0, n, 7000):
        hi = min(n, lo + 7000)
        st.append(x[lo:hi], prio[lo:hi], tag[lo:hi], exp[lo:hi])
    # several steps in a row: the p2p exchange alternates between its two slots and every merge
    # has to wait for the blocks of ITS step
    for rep in range(5):
        idx, sc, sc64, inexact = st.search(q if rep % 2 == 0 else q.flip(0).contiguous(), k, min_priority=1, now_us=1_000_000)
    torch.cuda.synchronize()
    st.check_exchange()
    assert st.exchange == ("nccl" if fallback else exchange)      # no silent fallback unless a rank really failed
    assert int(inexact.item()) == 0
    # pipelined form (what bench.py runs): begin(i+1) before end(i); with the peer-memory exchange the
    # result rows are stored into the peers by finish_kernel itself
    want_i, want_s = idx.clone(), sc64.clone()
    qf = q.flip(0).contiguous()
    for rep in range(3):
        t0 = st.search_begin(q, k, slot=0, min_priority=1, now_us=1_000_000)
        t1 = st.search_begin(qf, k, slot=1, min_priority=1, now_us=1_000_000)
        i0, _, s0, _ = st.search_end(t0)
        i1, _, s1, _ = st.search_end(t1)
        torch.cuda.synchronize()
        assert torch.equal(i0, want_i) and torch.equal(s0, want_s)
        assert torch.equal(i1.flip(0), want_i) and torch.equal(s1.flip(0), want_s)
    st.check_exchange()
    if exchange == "copy" and not fallback:
        # the one-call C entry point a non-Python host uses: search + publish by copy + merge on one stream
        import ctypes as C
        from pilottai_b200 import _lib
        lib = _lib.load()
        xs = st._exchange_for(nq, k)
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for rep in range(10):                           # more steps than the exchange has slots
            blk = torch.empty((2, nq, k), dtype=torch.int64, device="cuda")
            o_i = torch.empty((nq, k), dtype=torch.int64, device="cuda")
            o_s = torch.empty((nq, k), dtype=torch.float32, device="cuda")
            o_s64 = torch.empty((nq, k), dtype=torch.float64, device="cuda")
            cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
            _lib.check(lib.pilott_mem_search_sharded(st.local._h, xs[0], xs[1], q.data_ptr(), nq, k, 1, 0, 1_000_000,
                                                     o_i.data_ptr(), o_s.data_ptr(), o_s64.data_ptr(), cnt.data_ptr(),
                                                     None, blk.data_ptr(), stream))
            xs[1] += 1
            torch.cuda.synchronize()
            assert int(cnt.item()) == 0 and torch.equal(o_i, want_i) and torch.equal(o_s64, want_s)
        st.check_exchange()
    # k > 32: every shard answers through the large-k path, the exchange and the merge are the same
    i64, _, s64, _ = st.search(q, 64, min_priority=1, now_us=1_000_000)
    torch.cuda.synchronize()
    if rank == 0:
        np.save(out + "_i.npy", want_i.cpu().numpy()); np.save(out + "_s.npy", want_s.cpu().numpy())
        np.save(out + "_i64.npy", i64.cpu().numpy()); np.save(out + "_s64.npy", s64.cpu().numpy())
    st.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("exchange", ["copy", "p2p", "nccl", "copy-fallback"])
def test_sharded_search_equals_single_store(torch_cuda, tmp_path, exchange):
    torch = torch_cuda
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs two or more GPUs")
    import torch.multiprocessing as mp
    n, dim, nq, k = 30_000, 256, 64, 10
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "merged")
    mp.spawn(_worker, args=(world, port, n, dim, nq, k, out, exchange), nprocs=world, join=True)
    whole = OracleStore(dim, n)
    prio, tag, exp = synth.metadata(n, seed=5)
    whole.append(synth.corpus(n, dim, seed=3, dup_frac=0.01), prio, tag, exp)
    wi, ws = retrieval_c.search(whole, synth.queries(nq, dim, seed=4), k, min_priority=1, now_us=1_000_000)
    assert np.array_equal(np.load(out + "_i.npy"), wi)
    assert np.array_equal(np.load(out + "_s.npy"), ws)
    wi, ws = retrieval_c.search(whole, synth.queries(nq, dim, seed=4), 64, min_priority=1, now_us=1_000_000)
    assert np.array_equal(np.load(out + "_i64.npy"), wi)
    assert np.array_equal(np.load(out + "_s64.npy"), ws)
