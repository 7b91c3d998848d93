"""GPU parity of the second-chance pass (pilott_mem_search_refine) against the CPU oracle.

The fused search proves its own answers with a gap check; what it cannot prove it flags.  These
corpora are built to be flagged: runs of near-duplicate rows stored side by side (a query's whole
top-k sits in one candidate list), exact duplicates in greater number than a list holds (ties
decided by insertion order alone), and more near-ties than the refine pass collects (brute force).
Bar as everywhere: indices and fp64 scores bit-identical to the oracle.
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


def _bursty(n, dim, burst, seed, noise=0.05):
    rng = np.random.default_rng(seed)
    topics = rng.standard_normal(((n + burst - 1) // burst, dim)).astype(np.float32)
    x = np.repeat(topics, burst, axis=0)[:n] + noise * rng.standard_normal((n, dim)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x.astype(np.float32), topics


def _stores(x, capacity, prio=None, tag=None):
    from pilottai_b200.store import SemanticStore
    gs, os_ = SemanticStore(x.shape[1], capacity, 0), OracleStore(x.shape[1], capacity)
    for lo in range(0, x.shape[0], capacity):
        sl = slice(lo, min(lo + capacity, x.shape[0]))
        gs.append(x[sl], None if prio is None else prio[sl], None if tag is None else tag[sl])
        os_.append(x[sl], None if prio is None else prio[sl], None if tag is None else tag[sl])
    return gs, os_


def _device_search_with_refine(torch, gs, q, k, **flt):
    """Fused search -> flags -> refine, all through the device entry points."""
    qd = torch.from_numpy(q).cuda()
    idx, sc, sc64, inexact, flags = gs.search(qd, k, want_score64=True, **flt)
    n_flag = int(inexact.item())
    n_brute = 0
    if n_flag:
        n_brute = gs.refine(qd, k, flags.nonzero(), (idx, sc, sc64), **flt)
    torch.cuda.synchronize()
    return idx.cpu().numpy(), sc.cpu().numpy(), sc64.cpu().numpy(), n_flag, n_brute


def _same(idx, sc, sc64, os_, q, k, **flt):
    oidx, osc = retrieval_c.search(os_, q, k, **flt)
    assert np.array_equal(idx, oidx), f"indices differ in {np.sum(np.any(idx != oidx, axis=1))} queries"
    assert np.array_equal(sc64, osc)
    assert np.array_equal(sc, osc.astype(np.float32))


@pytest.mark.parametrize("burst,list_len", [(12, 8), (24, 16), (40, 32), (40, 0)])
def test_bursty_corpus_is_flagged_and_repaired(torch_cuda, burst, list_len):
    n, dim, k = 30000, 256, 10
    # noise 0.02: the burst's scores differ by ~2e-5, well inside the gap check's 1e-4 margin
    x, topics = _bursty(n, dim, burst, seed=burst, noise=0.02)
    gs, os_ = _stores(x, 32768)
    gs.set_list_len(list_len)
    rng = np.random.default_rng(1)
    q = topics[rng.integers(0, topics.shape[0], 200)] + 0.02 * rng.standard_normal((200, dim)).astype(np.float32)
    q = q.astype(np.float32)
    idx, sc, sc64, n_flag, n_brute = _device_search_with_refine(torch_cuda, gs, q, k)
    assert n_flag > 0, "corpus was meant to defeat the candidate lists"
    assert n_brute == 0, "the collect pass should have settled every flagged query"
    _same(idx, sc, sc64, os_, q, k)
    # the host entry point runs the same repair inside the library
    hidx, hsc, hsc64 = gs.search(q, k, want_score64=True)
    assert gs.last_fallbacks > 0
    _same(hidx, hsc, hsc64, os_, q, k)
    gs.close()


def test_exact_duplicates_wider_than_a_list(torch_cuda):
    """300 identical rows: every score ties, the answer is the k OLDEST copies."""
    n, dim, k = 20000, 128, 10
    x, _ = _bursty(n, dim, 1, seed=5)
    x[7000:7300] = x[7000]
    x[15000:15040] = x[7000]
    gs, os_ = _stores(x, n)
    q = np.stack([x[7000], x[100], x[7000] + 0.01 * x[3]]).astype(np.float32)
    idx, sc, sc64, n_flag, n_brute = _device_search_with_refine(torch_cuda, gs, q, k)
    assert n_flag >= 1 and n_brute == 0
    assert list(idx[0]) == list(range(7000, 7010))
    _same(idx, sc, sc64, os_, q, k)
    gs.close()


def test_more_near_ties_than_the_collect_pass_holds(torch_cuda):
    """1500 identical rows > 1024 collected candidates: that query goes to brute force, the rest do not."""
    n, dim, k = 20000, 128, 5
    x, topics = _bursty(n, dim, 20, seed=6)
    x[2000:3500] = x[2000]
    gs, os_ = _stores(x, n)
    q = np.concatenate([x[2000:2001], topics[500:520]]).astype(np.float32)
    idx, sc, sc64, n_flag, n_brute = _device_search_with_refine(torch_cuda, gs, q, k)
    assert n_brute == 1
    assert list(idx[0]) == list(range(2000, 2005))
    _same(idx, sc, sc64, os_, q, k)
    gs.close()


def test_refine_respects_filters_and_ring_wrap(torch_cuda):
    n, dim, cap, k = 50000, 192, 16384, 8
    x, topics = _bursty(n, dim, 16, seed=8)
    rng = np.random.default_rng(9)
    prio = rng.integers(0, 4, n).astype(np.int32)
    tag = np.zeros(n, np.uint64)
    for b in range(6):
        tag |= (rng.random(n) < 0.5).astype(np.uint64) << np.uint64(b)
    gs, os_ = _stores(x, cap, prio, tag)              # ring wrapped three times
    gs.set_list_len(8)
    live = topics[(n - cap) // 16 + 5: (n - cap) // 16 + 205]
    q = (live + 0.02 * rng.standard_normal(live.shape)).astype(np.float32)
    for flt in ({}, {"min_priority": 2}, {"req_tagmask": 0b101}, {"min_priority": 1, "req_tagmask": 0b10}):
        idx, sc, sc64, n_flag, n_brute = _device_search_with_refine(torch_cuda, gs, q, k, **flt)
        _same(idx, sc, sc64, os_, q, k, **flt)
    assert n_flag >= 0
    gs.close()


def test_list_length_floor_rises_after_flagged_batches(torch_cuda):
    """Automatic list length: a batch that needed the repair step makes the next one use longer
    lists (fewer flags); results stay exact either way."""
    n, dim, k = 300000, 64, 10        # big enough for >= 16 candidate lists per query (8-entry lists apply)
    x, topics = _bursty(n, dim, 12, seed=11)
    gs, os_ = _stores(x, n)
    rng = np.random.default_rng(12)
    q = (topics[rng.integers(0, topics.shape[0], 128)] + 0.02 * rng.standard_normal((128, dim))).astype(np.float32)
    flagged = []
    for _ in range(3):
        idx, sc, sc64 = gs.search(q, k, want_score64=True)
        flagged.append(gs.last_fallbacks)
    _same(idx, sc, sc64, os_, q, k)
    assert flagged[0] > 0 and flagged[-1] < flagged[0]
    gs.close()


def test_pipelined_host_search_matches_oracle_and_repairs(torch_cuda):
    """pilott_mem_search_host_begin / _end: three batches in flight at once (pageable numpy inputs),
    one of them built to be flagged; slots are reusable and a busy slot is refused."""
    from pilottai_b200._lib import PilottError
    n, dim, k = 30000, 256, 10
    x, topics = _bursty(n, dim, 12, seed=21, noise=0.02)
    gs, os_ = _stores(x, 32768)
    gs.set_list_len(8)
    rng = np.random.default_rng(22)
    near = lambda m: (topics[rng.integers(0, topics.shape[0], m)] + 0.02 * rng.standard_normal((m, dim))).astype(np.float32)
    batches = [near(100), rng.standard_normal((64, dim)).astype(np.float32), near(33)]
    for rep in range(2):                                   # second round reuses the slots
        for s, q in enumerate(batches):
            gs.search_host_begin(s, q, k)
        outs, fallbacks = [], []
        for s in range(3):
            outs.append(gs.search_host_end(s, want_score64=True))
            fallbacks.append(gs.last_fallbacks)
        for q, (idx, sc, sc64) in zip(batches, outs):
            _same(idx, sc, sc64, os_, q, k)
        if rep == 0:
            assert fallbacks[0] > 0 and fallbacks[2] > 0
    gs.search_host_begin(0, batches[1], k)
    with pytest.raises(PilottError):
        gs.search_host_begin(0, batches[1], k)
    with pytest.raises(ValueError):
        gs.search_host_end(1)
    idx, sc = gs.search_host_end(0)
    sidx, ssc = gs.search(batches[1], k)
    assert np.array_equal(idx, sidx) and np.array_equal(sc, ssc)
    gs.close()
