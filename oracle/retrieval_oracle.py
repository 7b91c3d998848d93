"""Retrieval oracle -- CPU restatement of the semantic-memory store/search path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Parity status: **parity unpinned by the reference** for the *scores*.  The
reference (appyjain21/pilottai) has no vector arithmetic at all: its
``EnhancedMemory.semantic_search`` is a case-insensitive substring test
(/root/reference/pilott/memory/enhanced_memory.py:110).  Everything the
reference DOES specify is followed here and cited line by line:

* bounded ring, oldest row silently evicted  -> enhanced_memory.py:27,86
* empty text / empty query guards            -> enhanced_memory.py:68-69,100-101
* candidate filter: priority >= min_priority AND every requested tag present
                                             -> enhanced_memory.py:118-131
* expiry test ``now > expires_at`` (strict)  -> enhanced_memory.py:20-21,108
* older-first ordering among equals (stable sort on timestamp)
                                             -> enhanced_memory.py:114

Those candidate rules ARE pinned on the reference itself: oracle/gen_golden_memory.py runs the
unmodified reference on texts that all match the query and commits the lists it returns
(tests/golden/memory_filters.json); tests/test_oracle_memory_filters.py holds `valid_mask` and the
tie order of this file to them.  Only the scoring rule below has nothing in the reference to be
checked against.

THE STATED SCORING RULE (this build's definition, the GPU must reproduce it):

1. Vectors are rounded fp32 -> bf16, round-to-nearest-even (NaN -> 0x7FFF).
2. dot(q, m) is accumulated in IEEE fp64 in EIGHT INTERLEAVED PARTIAL SUMS: partial j takes
   the terms q_d*m_d for d = j, j+8, j+16, ... in ascending order; the total is
   ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)).  (Each bf16*bf16 product is exact in fp64, so
   FMA == mul+add; eight short chains instead of one long one let the device keep the
   same fixed order at an eighth of the dependent latency.)
   nq = sum q_d*q_d, nm = sum m_d*m_d, same scheme.
3. den = sqrt(nq) * sqrt(nm)  (fp64, correctly rounded sqrt, then one multiply)
   score = dot / den  if den > 0 else 0.0          (fp64 divide)
4. Rank candidates by (score descending, insertion sequence ascending); take k.
   Missing results are padded with index -1 and score -inf.

Indices returned are global insertion sequence numbers (0 for the first row
ever stored), which is what a host needs to find the MemoryItem again.
"""
from __future__ import annotations

import numpy as np

__all__ = [
    "bf16_round_bits", "bf16_bits_to_f32", "exact_scores", "topk_rows",
    "OracleStore", "fast_fp32_search",
]


def bf16_round_bits(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit patterns (uint16), round-to-nearest-even, NaN -> 0x7FFF."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    lsb = (u >> np.uint32(16)) & np.uint32(1)
    rounded = ((u + np.uint32(0x7FFF) + lsb) >> np.uint32(16)).astype(np.uint16)
    return np.where(np.isnan(x), np.uint16(0x7FFF), rounded)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(b, dtype=np.uint16).astype(np.uint32) << np.uint32(16)).view(np.float32)


def _tree8(p):
    return ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]))


def _seq_dot(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """[Q,D] x [N,D] -> [Q,N] fp64: eight interleaved partial sums, each sequential (rule 2)."""
    acc = np.zeros((8, a.shape[0], b.shape[0]), dtype=np.float64)
    for d in range(a.shape[1]):
        acc[d & 7] += a[:, d, None] * b[None, :, d]
    return _tree8(acc)


def _seq_sqnorm(a: np.ndarray) -> np.ndarray:
    acc = np.zeros((8, a.shape[0]), dtype=np.float64)
    for d in range(a.shape[1]):
        acc[d & 7] += a[:, d] * a[:, d]
    return _tree8(acc)


def exact_scores(q_bits: np.ndarray, m_bits: np.ndarray) -> np.ndarray:
    """Rule 2+3 on bf16 bit patterns.  Returns fp64 [Q, N]."""
    q = bf16_bits_to_f32(q_bits).astype(np.float64)
    m = bf16_bits_to_f32(m_bits).astype(np.float64)
    dot = _seq_dot(q, m)
    den = np.sqrt(_seq_sqnorm(q))[:, None] * np.sqrt(_seq_sqnorm(m))[None, :]
    with np.errstate(divide="ignore", invalid="ignore"):
        s = dot / den
    return np.where(den > 0, s, 0.0)


def topk_rows(scores: np.ndarray, seq: np.ndarray, valid: np.ndarray, k: int):
    """Rule 4.  scores [Q,N] fp64, seq [N] int64, valid [N] bool."""
    Q = scores.shape[0]
    out_idx = np.full((Q, k), -1, dtype=np.int64)
    out_sc = np.full((Q, k), -np.inf, dtype=np.float64)
    cand = np.nonzero(valid)[0]
    if cand.size == 0:
        return out_idx, out_sc
    cseq = seq[cand]
    for qi in range(Q):
        s = scores[qi, cand]
        order = np.lexsort((cseq, -s))[:k]          # primary -s, secondary seq
        out_idx[qi, : order.size] = cseq[order]
        out_sc[qi, : order.size] = s[order]
    return out_idx, out_sc


class OracleStore:
    """Ring-buffer store with the reference's capacity / filter / expiry rules."""

    def __init__(self, dim: int, capacity: int):
        self.dim, self.capacity = int(dim), int(capacity)
        self.vec = np.zeros((capacity, dim), dtype=np.uint16)     # bf16 bits, by slot
        self.prio = np.zeros(capacity, dtype=np.int32)
        self.tagmask = np.zeros(capacity, dtype=np.uint64)
        self.expires_us = np.zeros(capacity, dtype=np.int64)      # 0 = never
        self.total = 0                                            # rows ever appended

    # deque(maxlen) semantics: enhanced_memory.py:27,86
    def append(self, x, prio=None, tagmask=None, expires_us=None) -> int:
        x = np.atleast_2d(np.asarray(x, dtype=np.float32))
        n = x.shape[0]
        prio = np.zeros(n, np.int32) if prio is None else np.asarray(prio, np.int32)
        tagmask = np.zeros(n, np.uint64) if tagmask is None else np.asarray(tagmask, np.uint64)
        expires_us = np.zeros(n, np.int64) if expires_us is None else np.asarray(expires_us, np.int64)
        first = self.total
        bits = bf16_round_bits(x)
        for i in range(n):
            s = (self.total + i) % self.capacity
            self.vec[s], self.prio[s] = bits[i], prio[i]
            self.tagmask[s], self.expires_us[s] = tagmask[i], expires_us[i]
        self.total += n
        return first

    @property
    def count(self) -> int:
        return min(self.total, self.capacity)

    def slot_seq(self) -> np.ndarray:
        """Insertion sequence held by every slot (-1 = never written)."""
        s = np.arange(self.capacity, dtype=np.int64)
        if self.total <= self.capacity:
            return np.where(s < self.total, s, -1)
        head = self.total % self.capacity
        base = self.total - self.capacity
        return base + ((s - head) % self.capacity)

    def valid_mask(self, min_priority=0, req_tagmask=0, now_us=0) -> np.ndarray:
        seq = self.slot_seq()
        req = np.uint64(req_tagmask)
        ok = seq >= 0
        ok &= self.prio >= min_priority                             # :124-125
        ok &= (self.tagmask & req) == req                           # :126-130
        ok &= ~((self.expires_us != 0) & (now_us > self.expires_us))  # :20-21
        return ok

    def search(self, queries, k, min_priority=0, req_tagmask=0, now_us=0, scorer=None):
        q_bits = bf16_round_bits(np.atleast_2d(np.asarray(queries, dtype=np.float32)))
        scorer = scorer or exact_scores
        scores = scorer(q_bits, self.vec)
        return topk_rows(scores, self.slot_seq(), self.valid_mask(min_priority, req_tagmask, now_us), k)


def fast_fp32_search(corpus_f32: np.ndarray, inv_norm: np.ndarray, queries_f32: np.ndarray, k: int):
    """The TIMED CPU baseline: BLAS fp32 q @ M.T, argpartition,
    sort.  Not bit-defined; never used as a parity checker."""
    s = (queries_f32 @ corpus_f32.T) * inv_norm[None, :]
    part = np.argpartition(-s, k - 1, axis=1)[:, :k]
    ps = np.take_along_axis(s, part, axis=1)
    order = np.argsort(-ps, axis=1, kind="stable")
    return np.take_along_axis(part, order, axis=1), np.take_along_axis(ps, order, axis=1)
