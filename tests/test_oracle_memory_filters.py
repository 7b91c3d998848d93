"""The retrieval oracle's candidate rules (priority floor, tag AND, expiry, older-first order)
against lists the UNMODIFIED reference returned (tests/golden/memory_filters.json, written by
oracle/gen_golden_memory.py).  The reference has no vector score to pin, but it does define which
rows a search may return and in what order; every text in the fixture matches the query, so those
lists are exactly that definition."""
import json
import os

import numpy as np
import pytest

from oracle.retrieval_oracle import OracleStore


@pytest.fixture(scope="module")
def fixture(golden_dir):
    with open(os.path.join(golden_dir, "memory_filters.json")) as f:
        return json.load(f)


def _store(items):
    bits = {}
    st = OracleStore(8, 10000)
    n = len(items)
    mask = np.zeros(n, np.uint64)
    for i, it in enumerate(items):
        for t in it["tags"]:
            mask[i] |= np.uint64(1) << np.uint64(bits.setdefault(t, len(bits)))
    prio = np.array([it["priority"] for it in items], np.int32)
    # ttl 0 -> expired before any search (expires at 1 us, searches run at 2 us); ttl 3600 -> far future
    exp = np.array([0 if it["ttl"] is None else (1 if it["ttl"] == 0 else 10**15) for it in items], np.int64)
    st.append(np.ones((n, 8), np.float32), prio, mask, exp)
    return st, bits


@pytest.mark.parametrize("name", ["small", "mid"])
def test_oracle_candidate_rules_match_reference_lists(fixture, name):
    sc = fixture[name]
    items = sc["items"]
    st, bits = _store(items)
    for q in sc["queries"]:
        assert "expect" in q, q
        if any(t not in bits for t in q["tags"]):
            assert q["expect"] == []                    # a tag nobody stored: the reference returns nothing
            continue
        req = 0
        for t in q["tags"]:
            req |= 1 << bits[t]
        ok = st.valid_mask(q["min_priority"], req, now_us=2)
        seqs = np.nonzero(ok)[0]                        # store never wrapped: slot == insertion sequence
        # reference order (:114): priority descending, then timestamp (= insertion order) ascending
        order = sorted(seqs, key=lambda s: (-items[s]["priority"], s))
        assert [items[s]["text"] for s in order] == q["expect"], q


def test_all_equal_scores_rank_by_insertion_order(fixture):
    """Identical vectors: every score ties and the oracle's rule (score desc, sequence asc) must
    give the oldest eligible rows -- the reference's older-first order among equals."""
    sc = fixture["small"]
    st, _ = _store(sc["items"])
    idx, _ = st.search(np.ones((1, 8), np.float32), 5, min_priority=1, now_us=2)
    elig = [i for i, it in enumerate(sc["items"]) if it["priority"] >= 1 and it["ttl"] != 0]
    assert list(idx[0]) == elig[:5]
