"""The C twin of the retrieval oracle against the definitional numpy statement (CPU only).

Every GPU parity test checks against oracle/retrieval_oracle.c (fast enough for long corpora);
this file holds both of its entry points -- the per-query scan and the row-range scan bench.py
uses for its parity sample -- to oracle/retrieval_oracle.py bit for bit: indices and fp64 scores,
with filters, an expiry exactly at `now`, duplicated rows, a wrapped ring and odd dimensions."""
import numpy as np
import pytest

from oracle import retrieval_c
from oracle.retrieval_oracle import OracleStore, bf16_round_bits, exact_scores
from pilottai_b200 import synth


@pytest.mark.parametrize("n,dim,cap,seed", [(700, 96, 512, 1), (300, 200, 300, 2), (1500, 768, 2048, 3), (64, 7, 64, 4)])
def test_c_twin_equals_numpy_oracle(n, dim, cap, seed):
    x = synth.corpus(n, dim, seed=seed, dup_frac=0.02)
    prio, tag, exp = synth.metadata(n, seed=seed + 10)
    now = 1_000_000
    exp[: n // 7] = now                          # expires exactly now: strict '>' keeps the row (enhanced_memory.py:21)
    exp[n // 7: n // 5] = now - 1                # one microsecond earlier: gone
    st = OracleStore(dim, cap)
    st.append(x, prio, tag, exp)
    q = synth.queries(9, dim, seed=seed + 20)
    q[0] = x[n - 1]
    for k in (1, 10, 33):
        for flt in ({}, {"min_priority": 1, "req_tagmask": 0b10, "now_us": now}, {"now_us": now}):
            want_idx, want_sc = st.search(q, k, **flt)
            got_idx, got_sc = retrieval_c.search(st, q, k, threads=3, **flt)
            assert np.array_equal(got_idx, want_idx) and np.array_equal(got_sc, want_sc)
            for thr in (1, 4):
                r_idx, r_sc = retrieval_c.search_rows(st, q, k, threads=thr, **flt)
                assert np.array_equal(r_idx, want_idx) and np.array_equal(r_sc, want_sc)


def test_exact_scores_twin():
    rng = np.random.default_rng(5)
    m = bf16_round_bits((rng.standard_normal((257, 130)) * np.exp(rng.uniform(-20, 20, (257, 1)))).astype(np.float32))
    q = bf16_round_bits(rng.standard_normal((3, 130)).astype(np.float32))
    m[5] = 0                                     # zero row: score defined as 0
    assert np.array_equal(retrieval_c.exact_scores(q, m), exact_scores(q, m))


def test_expiry_boundary_is_kept():
    """now_us == expires_us must be kept (strict '>' in MemoryItem.is_expired, enhanced_memory.py:20-21)."""
    st = OracleStore(8, 4)
    x = np.eye(4, 8, dtype=np.float32)
    st.append(x, expires_us=np.array([100, 99, 0, 101], np.int64))
    idx, _ = retrieval_c.search_rows(st, x[:1] + 0.1, 4, now_us=100)
    assert sorted(idx[0].tolist()) == [-1, 0, 2, 3]
    assert st.valid_mask(now_us=100).tolist() == [True, False, True, True]
