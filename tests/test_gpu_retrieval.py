"""GPU parity: the CUDA store/search path (through the C ABI) vs the CPU oracle.

Bar: top-k indices bit-exact (same ranking rule: score desc, insertion sequence asc),
scores within 1e-3 relative as BASELINE.json states -- in fact the exact re-score makes the
fp64 scores bit-identical and the fp32 ones correctly rounded, which is what is asserted.
"""
import numpy as np
import pytest

from oracle import retrieval_c
from oracle.retrieval_oracle import OracleStore, bf16_round_bits
from pilottai_b200 import synth

pytestmark = pytest.mark.gpu
KERNELS = ["generic", "tc", "tc_pair"]


@pytest.fixture(scope="module")
def torch_cuda():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _mk(dim, capacity):
    from pilottai_b200.store import SemanticStore
    return SemanticStore(dim, capacity, 0)


def _fill(n, dim, capacity, seed=0, dup=0.01, with_meta=True, chunks=3):
    """Same rows into the GPU store and the oracle store, in `chunks` appends."""
    x = synth.corpus(n, dim, seed=seed, dup_frac=dup)
    prio = tag = exp = None
    if with_meta:
        prio, tag, exp = synth.metadata(n, seed=seed + 1)
    gs, os_ = _mk(dim, capacity), OracleStore(dim, capacity)
    bounds = np.linspace(0, n, chunks + 1).astype(int)
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        if hi - lo > capacity:      # a batch may not exceed the ring
            for l2 in range(lo, hi, capacity):
                h2 = min(hi, l2 + capacity)
                sl = slice(l2, h2)
                a = gs.append(x[sl], None if prio is None else prio[sl], None if tag is None else tag[sl],
                              None if exp is None else exp[sl])
                b = os_.append(x[sl], None if prio is None else prio[sl], None if tag is None else tag[sl],
                               None if exp is None else exp[sl])
                assert a == b
            continue
        sl = slice(lo, hi)
        a = gs.append(x[sl], None if prio is None else prio[sl], None if tag is None else tag[sl],
                      None if exp is None else exp[sl])
        b = os_.append(x[sl], None if prio is None else prio[sl], None if tag is None else tag[sl],
                       None if exp is None else exp[sl])
        assert a == b
    return gs, os_, x


def _check(gs, os_, q, k, **flt):
    idx, sc, sc64 = gs.search(q, k, want_score64=True, **flt)
    oidx, osc = retrieval_c.search(os_, q, k, **flt)
    assert np.array_equal(idx, oidx), f"indices differ in {np.sum(np.any(idx != oidx, axis=1))} queries"
    assert np.array_equal(sc64, osc)                      # fp64 scores bit-identical
    assert np.array_equal(sc, osc.astype(np.float32))     # fp32 = correctly rounded fp64
    fin = np.isfinite(osc)
    assert np.all(np.abs(sc[fin] - osc[fin]) <= 1e-3 * np.abs(osc[fin]) + 1e-12)   # the stated tolerance


def test_append_matches_oracle_rounding(torch_cuda):
    gs, os_, x = _fill(3000, 768, 4096, seed=3)
    ex = gs.export()
    assert np.array_equal(ex["vec_bits"], bf16_round_bits(x))
    assert np.array_equal(ex["prio"], os_.prio[:3000])
    assert np.array_equal(ex["tagmask"], os_.tagmask[:3000])
    assert np.array_equal(ex["expires_us"], os_.expires_us[:3000])
    xb = (ex["vec_bits"].astype(np.uint32) << 16).view(np.float32).astype(np.float64)
    want = 1.0 / np.sqrt((xb * xb).sum(axis=1))
    assert np.allclose(ex["inv_norm"], want, rtol=1e-5)
    gs.close()


def test_append_ring_wrap_and_odd_dim(torch_cuda):
    gs, os_, x = _fill(2500, 100, 1000, seed=4, chunks=7)       # dim not a multiple of 64
    assert gs.info() == {"total": 2500, "count": 1000, "capacity": 1000, "dim": 100, "row_stride": 128}
    ex = gs.export()                                            # age order == last 1000 rows
    assert np.array_equal(ex["vec_bits"], bf16_round_bits(x[1500:]))
    gs.close()


def test_device_append_equals_host_append(torch_cuda):
    torch = torch_cuda
    x = synth.corpus(1000, 256, seed=9)
    prio, tag, exp = synth.metadata(1000, seed=10)
    a, b = _mk(256, 1500), _mk(256, 1500)
    a.append(x, prio, tag, exp)
    b.append(torch.from_numpy(x).cuda(), torch.from_numpy(prio).cuda(),
             torch.from_numpy(tag.view(np.int64)).cuda(), torch.from_numpy(exp).cuda())
    torch.cuda.synchronize()
    ea, eb = a.export(), b.export()
    for key in ea:
        assert np.array_equal(ea[key], eb[key]), key
    a.close(); b.close()


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("dim,n,cap,nq", [(768, 5000, 8192, 37), (100, 3000, 3000, 5), (1024, 2600, 2048, 130)])
def test_search_small_matches_oracle(torch_cuda, kernel, dim, n, cap, nq):
    gs, os_, x = _fill(n, dim, cap, seed=dim)
    gs.set_kernel(kernel)
    q = synth.queries(nq, dim, seed=7)
    q[1] = x[n - 5]                         # exact hit (and its duplicates) -> score ties
    _check(gs, os_, q, 10)
    _check(gs, os_, q, 1)
    _check(gs, os_, q, 32)
    _check(gs, os_, q, 10, min_priority=2)
    _check(gs, os_, q, 10, req_tagmask=0b101, now_us=1_000_000)
    _check(gs, os_, q, 10, min_priority=3, req_tagmask=0xFFFF, now_us=1_000_000)    # nearly all masked
    _check(gs, os_, q, 10, min_priority=99)                                          # all masked
    gs.close()


@pytest.mark.parametrize("kernel", KERNELS)
def test_duplicate_rows_tie_break_by_age(torch_cuda, kernel):
    """40 identical rows straddling the ring seam: more exact ties than the fused path keeps,
    so the gap check must hand those queries to the exact path; older rows win."""
    dim, cap = 128, 1024
    x = synth.corpus(1500, dim, seed=21)
    x[480:520] = x[480]          # after wrap these sit on both sides of physical slot 0/1023
    x[1490:1500] = x[480]
    gs, os_ = _mk(dim, cap), OracleStore(dim, cap)
    for lo in range(0, 1500, 500):
        gs.append(x[lo:lo + 500]); os_.append(x[lo:lo + 500])
    gs.set_kernel(kernel)
    q = np.stack([x[480], x[700], -x[480]])
    _check(gs, os_, q, 10)
    _check(gs, os_, q, 32)
    assert gs.last_fallbacks >= 1
    gs.close()


@pytest.mark.parametrize("kernel", KERNELS)
def test_empty_and_tiny_stores(torch_cuda, kernel):
    gs, os_ = _mk(64, 100), OracleStore(64, 100)
    gs.set_kernel(kernel)
    q = synth.queries(3, 64)
    idx, sc = gs.search(q, 5)
    assert np.all(idx == -1) and np.all(np.isneginf(sc))       # empty store: empty result, no crash
    x = synth.corpus(3, 64, seed=1)
    gs.append(x); os_.append(x)
    _check(gs, os_, q, 5)                                       # k > rows held -> -1 padding
    z = np.zeros((2, 64), np.float32)
    gs.append(z); os_.append(z)                                 # zero vectors score 0.0
    _check(gs, os_, np.vstack([q, z[:1]]), 5)
    gs.close()


def test_exact_path_large_k(torch_cuda):
    gs, os_, x = _fill(4000, 256, 4096, seed=5)
    q = synth.queries(9, 256, seed=6)
    _check(gs, os_, q, 50)                       # 32 < k <= 1024 -> threshold-collect path (tests/test_gpu_largek.py)
    _check(gs, os_, q, 100, min_priority=1, now_us=1_000_000)
    assert gs.last_fallbacks == 0                # nobody needed the brute-force kernels
    _check(gs, os_, q, 1500)                     # k > 1024 -> brute-force fp64 path
    assert gs.last_fallbacks == 9
    torch = torch_cuda                           # and the brute-force entry point on its own
    idx, sc, sc64 = gs.search_exact(torch.from_numpy(q).cuda(), 50)
    oidx, osc = retrieval_c.search(os_, q, 50)
    assert np.array_equal(idx.cpu().numpy(), oidx) and np.array_equal(sc64.cpu().numpy(), osc)
    gs.close()


def test_compact_drops_expired(torch_cuda):
    gs, os_, x = _fill(2000, 128, 2048, seed=8)
    now = 1_000_000
    kept, keep = gs.compact(now)
    want_keep = ~((os_.expires_us[:2000] != 0) & (now > os_.expires_us[:2000]))
    assert np.array_equal(keep, want_keep) and kept == want_keep.sum()
    ex = gs.export()
    assert np.array_equal(ex["vec_bits"], bf16_round_bits(x)[want_keep])
    # searching after compaction == searching the oracle with the expiry filter, modulo renumbering
    q = synth.queries(4, 128, seed=2)
    idx, sc, sc64 = gs.search(q, 10, want_score64=True)
    oidx, osc = retrieval_c.search(os_, q, 10, now_us=now)
    renum = np.full(2000, -1, np.int64)
    renum[want_keep] = gs.total - kept + np.arange(kept)
    assert np.array_equal(idx, renum[oidx]) and np.array_equal(sc64, osc)
    gs.close()


@pytest.mark.parametrize("kernel", KERNELS)
def test_search_medium_matches_oracle(torch_cuda, kernel):
    gs, os_, x = _fill(60_000, 768, 65_536, seed=11, dup=0.001)
    gs.set_kernel(kernel)
    q = synth.queries(64, 768, seed=12)
    _check(gs, os_, q, 10)
    _check(gs, os_, q, 10, min_priority=1, req_tagmask=0b10, now_us=1_000_000)
    assert gs.last_fallbacks == 0
    gs.close()


def test_full_size_properties(torch_cuda):
    """BASELINE.json config 2 shape (1M x 768, 1024 queries, top-10): the tcgen05 kernels
    must agree with the generic CUDA-core kernel on every query, a quarter of the queries (256)
    must match the CPU oracle bit for bit, and planted duplicates of the queries must come back
    first with score 1.0."""
    torch = torch_cuda
    n, dim, nq = 1_000_000, 768, 1024
    x = synth.corpus(n, dim, seed=1234)
    q = synth.queries(nq, dim, seed=5678)
    plant = np.arange(0, 64) * 15_000 + 7
    x[plant] = q[:64]
    gs, os_ = _mk(dim, n), OracleStore(dim, n)
    for lo in range(0, n, 250_000):
        gs.append(x[lo:lo + 250_000])
    os_.append(x)
    res = {}
    for kern in KERNELS:
        gs.set_kernel(kern)
        res[kern] = gs.search(q, 10, want_score64=True)
        assert gs.last_fallbacks == 0
    for kern in KERNELS[1:]:
        assert np.array_equal(res[kern][0], res["generic"][0]), kern
        assert np.array_equal(res[kern][2], res["generic"][2]), kern
    idx, sc, sc64 = res["tc"]
    assert np.array_equal(idx[:64, 0], plant) and np.all(sc[:64, 0] == 1.0)
    assert np.all(np.diff(sc64, axis=1) <= 0)                      # sorted
    # every fourth query (256 of the 1024) against the oracle over all 1M rows, bit for bit
    # (row-range form of the C twin: one pass over the corpus scores all sampled queries)
    sample = np.arange(0, nq, 4)
    oidx, osc = retrieval_c.search_rows(os_, q[sample], 10)
    assert np.array_equal(idx[sample], oidx) and np.array_equal(sc64[sample], osc)
    gs.close()


def test_snapshot_round_trip(torch_cuda, tmp_path):
    from pilottai_b200.store import load_snapshot, save_snapshot
    gs, os_, x = _fill(3000, 200, 2048, seed=17, chunks=5)            # ring wrapped: 3000 rows into 2048 slots
    q = synth.queries(11, 200, seed=18)
    want = gs.search(q, 10, want_score64=True, min_priority=1, now_us=1_000_000)
    path = str(tmp_path / "snap")
    save_snapshot(gs, path)
    gs2 = load_snapshot(path)
    assert gs2.info() == gs.info()
    got = gs2.search(q, 10, want_score64=True, min_priority=1, now_us=1_000_000)
    for w, g in zip(want, got):
        assert np.array_equal(w, g)                                   # same sequence numbers, same bits
    ea, eb = gs.export(), gs2.export()
    for key in ea:
        assert np.array_equal(ea[key], eb[key]), key
    gs.close(); gs2.close()


@pytest.mark.parametrize("kernel", ["tc", "tc_pair"])
def test_many_queries_multi_wave_grid(torch_cuda, kernel):
    """More query groups than SMs: the scan grid runs in several waves, the lists of one query
    are not all resident together, and the cold-start exchange must neither stall nor change
    the result."""
    gs, os_, x = _fill(20_000, 64, 20_000, seed=23, dup=0.002)
    gs.set_kernel(kernel)
    q = synth.queries(3000, 64, seed=24)
    _check(gs, os_, q, 10)
    _check(gs, os_, q[:700], 10, min_priority=2, req_tagmask=0b1, now_us=1_000_000)
    gs.close()


@pytest.mark.parametrize("kernel", KERNELS)
def test_wide_vectors_unstaged_rescore(torch_cuda, kernel):
    """dim 1536: 24 k-blocks per tile, and the exact re-score reads the rows straight from HBM
    (33 staged rows would not fit the finish kernel's shared memory)."""
    gs, os_, x = _fill(3000, 1536, 4096, seed=25)
    gs.set_kernel(kernel)
    q = synth.queries(20, 1536, seed=26)
    q[0] = x[1234]
    _check(gs, os_, q, 10)
    _check(gs, os_, q, 24, min_priority=1)
    gs.close()


def test_both_finish_forms_give_the_same_answers(torch_cuda):
    """Option 9: the default finish step re-scores only the survivors that can reach the top-k (and
    lets the gap check prove the cut); form 1 is the original kernel that re-scores all 32 -- same
    indices, same fp64 scores, both against the oracle."""
    gs, os_, x = _fill(9000, 768, 8192, seed=21)          # ring wrapped once
    q = synth.queries(300, 768, seed=22)
    for form in (1, 0, 1):
        gs.set_finish_form(form)
        _check(gs, os_, q, 10)
        _check(gs, os_, q, 32, min_priority=1, now_us=1_000_000)
        _check(gs, os_, q[:3], 1, req_tagmask=0b11)
    gs.close()
