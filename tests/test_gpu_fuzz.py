"""Seeded differential fuzz of the retrieval path against the CPU oracle: random shapes (dimension,
rows, ring capacity smaller or larger than the row count, batch size, k on both sides of 32),
random filters, duplicated and zero rows, appends interleaved with searches.  Every case must give
the oracle's indices and fp64 scores bit for bit, whichever of the fused / second-chance / large-k /
brute-force routes its queries take."""
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


def _case(rng):
    dim = int(rng.choice([8, 33, 64, 100, 128, 200, 256, 320]))
    cap = int(rng.integers(40, 5000))
    n = int(rng.integers(1, 2 * cap))
    nq = int(rng.choice([1, 2, 7, 31, 32, 33, 100, 129, 257]))
    k = int(rng.choice([1, 2, 5, 10, 16, 31, 32, 33, 40, 64, 200]))
    return dim, cap, n, nq, k


@pytest.mark.parametrize("seed", range(48))
def test_random_case_matches_oracle(torch_cuda, seed):
    from pilottai_b200.store import SemanticStore
    rng = np.random.default_rng(1000 + seed)
    dim, cap, n, nq, k = _case(rng)
    x = rng.standard_normal((n, dim)).astype(np.float32)
    if rng.random() < 0.5:                                   # clustered rows: near-ties and exact ties
        centers = rng.standard_normal((max(1, n // 37), dim)).astype(np.float32)
        x = centers[rng.integers(0, centers.shape[0], n)] + 0.01 * x
    dup = rng.random(n) < 0.05
    x[dup] = x[rng.integers(0, n, int(dup.sum()))]
    x[rng.random(n) < 0.01] = 0.0                            # zero rows score 0.0 (den == 0)
    prio = rng.integers(0, 4, n).astype(np.int32)
    tag = np.zeros(n, np.uint64)
    for b in range(5):
        tag |= (rng.random(n) < 0.4).astype(np.uint64) << np.uint64(b)
    exp = np.where(rng.random(n) < 0.1, rng.integers(1, 2_000_000, n), 0).astype(np.int64)
    gs, os_ = SemanticStore(dim, cap, 0), OracleStore(dim, cap)
    q = rng.standard_normal((nq, dim)).astype(np.float32)
    if n > 3:
        q[: nq // 2] = x[rng.integers(0, n, nq // 2)] + 0.01 * q[: nq // 2]      # on-topic queries
    done = 0
    while done < n:                                          # appends in ragged batches, searches in between
        m = int(min(n - done, cap, rng.integers(1, cap + 1)))
        sl = slice(done, done + m)
        assert gs.append(x[sl], prio[sl], tag[sl], exp[sl]) == os_.append(x[sl], prio[sl], tag[sl], exp[sl])
        done += m
        if rng.random() < 0.5 or done == n:
            flt = {}
            if rng.random() < 0.5:
                flt["min_priority"] = int(rng.integers(0, 4))
            if rng.random() < 0.5:
                flt["req_tagmask"] = int(rng.integers(0, 32))
            if rng.random() < 0.5:
                flt["now_us"] = int(rng.integers(0, 2_000_000))
            idx, sc, sc64 = gs.search(q, k, want_score64=True, **flt)
            oidx, osc = retrieval_c.search(os_, q, k, **flt)
            ctx = f"seed {seed}: dim {dim} cap {cap} rows {done}/{n} nq {nq} k {k} filters {flt}"
            assert np.array_equal(idx, oidx), ctx
            assert np.array_equal(sc64, osc), ctx
            assert np.array_equal(sc, osc.astype(np.float32)), ctx
    gs.close()
