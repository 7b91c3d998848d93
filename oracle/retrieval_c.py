"""ctypes binding of oracle/retrieval_oracle.c -- TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .retrieval_oracle import OracleStore, bf16_round_bits

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libretrieval_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "retrieval_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libretrieval_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.oracle_exact_scores.restype = None
        _lib.oracle_search.restype = None
        _lib.oracle_search_rows.restype = None
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def exact_scores(q_bits: np.ndarray, m_bits: np.ndarray) -> np.ndarray:
    q_bits = np.ascontiguousarray(q_bits, np.uint16)
    m_bits = np.ascontiguousarray(m_bits, np.uint16)
    Q, D = q_bits.shape
    N = m_bits.shape[0]
    out = np.empty((Q, N), np.float64)
    lib().oracle_exact_scores(_p(q_bits), C.c_int64(Q), _p(m_bits), C.c_int64(N), C.c_int(D),
                              C.c_int64(D), C.c_int64(m_bits.shape[1]), _p(out))
    return out


def search(store: OracleStore, queries, k, min_priority=0, req_tagmask=0, now_us=0, threads=None):
    q_bits = bf16_round_bits(np.atleast_2d(np.asarray(queries, np.float32)))
    Q, D = q_bits.shape
    seq = np.ascontiguousarray(store.slot_seq(), np.int64)
    out_idx = np.empty((Q, k), np.int64)
    out_sc = np.empty((Q, k), np.float64)
    L = lib()

    def run(lo, hi):      # ctypes drops the GIL: query chunks run on separate host threads
        L.oracle_search(_p(q_bits[lo:hi]), C.c_int64(hi - lo), _p(store.vec),
                        C.c_int64(store.capacity), C.c_int(D), C.c_int64(D),
                        C.c_int64(store.dim), _p(seq), _p(store.prio), _p(store.tagmask),
                        _p(store.expires_us), C.c_int32(min_priority), C.c_uint64(req_tagmask),
                        C.c_int64(now_us), C.c_int(k), _p(out_idx[lo:hi]), _p(out_sc[lo:hi]))

    nthr = max(1, min(threads or (os.cpu_count() or 1), Q))
    bounds = np.linspace(0, Q, nthr + 1).astype(int)
    if nthr == 1:
        run(0, Q)
    else:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(nthr) as ex:
            list(ex.map(lambda i: run(bounds[i], bounds[i + 1]), range(nthr)))
    return out_idx, out_sc


def search_rows(store, queries, k, min_priority=0, req_tagmask=0, now_us=0, threads=None, q_bits=None):
    """Same answers as search(), organised for long corpora: the slot range is split across host
    threads, each thread scores ALL queries against its rows in one pass (oracle_search_rows), and
    the per-range top-k lists are merged by (score desc, sequence asc).  `store` needs the
    OracleStore attributes (vec [capacity, dim] uint16, prio, tagmask, expires_us, capacity, dim,
    slot_seq()); q_bits passes queries that are already bf16 bit patterns."""
    if q_bits is None:
        q_bits = bf16_round_bits(np.atleast_2d(np.asarray(queries, np.float32)))
    q_bits = np.ascontiguousarray(q_bits, np.uint16)
    Q, D = q_bits.shape
    seq = np.ascontiguousarray(store.slot_seq(), np.int64)
    N = int(store.capacity)
    nthr = max(1, min(threads or (os.cpu_count() or 1), max(1, N // 1024)))
    bounds = np.linspace(0, N, nthr + 1).astype(np.int64)
    part_idx = np.empty((nthr, Q, k), np.int64)
    part_sc = np.empty((nthr, Q, k), np.float64)
    L = lib()

    def run(t):
        L.oracle_search_rows(_p(q_bits), C.c_int64(Q), _p(store.vec), C.c_int64(int(bounds[t])),
                             C.c_int64(int(bounds[t + 1])), C.c_int(D), C.c_int64(D),
                             C.c_int64(store.vec.shape[1]), _p(seq), _p(store.prio), _p(store.tagmask),
                             _p(store.expires_us), C.c_int32(min_priority), C.c_uint64(req_tagmask),
                             C.c_int64(now_us), C.c_int(k), _p(part_idx[t]), _p(part_sc[t]))

    if nthr == 1:
        run(0)
    else:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(nthr) as ex:
            list(ex.map(run, range(nthr)))
    out_idx = np.full((Q, k), -1, np.int64)
    out_sc = np.full((Q, k), -np.inf, np.float64)
    for qi in range(Q):
        ci, cs = part_idx[:, qi].ravel(), part_sc[:, qi].ravel()
        ok = ci >= 0
        ci, cs = ci[ok], cs[ok]
        order = np.lexsort((ci, -cs))[:k]
        out_idx[qi, :order.size] = ci[order]
        out_sc[qi, :order.size] = cs[order]
    return out_idx, out_sc
