"""SemanticStore -- thin Python handle over the device-resident ring in libpilott_b200.so.

Mirrors the C ABI one to one (include/pilott_b200.h); torch is used only to own device
buffers and to name the current CUDA stream.  numpy arrays go through the *_host entry
points (pinned staging + copies inside the call), CUDA tensors through the device ones.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _lib
from ._lib import check, ptr

KERNELS = {"auto": 0, "generic": 1, "tc": 2, "tc_pair": 3}


def _torch():
    import torch
    return torch


class SemanticStore:
    def __init__(self, dim: int, capacity: int, device: int = 0):
        self._lib = _lib.load()
        h = C.c_void_p()
        check(self._lib.pilott_mem_create(int(dim), int(capacity), int(device), C.byref(h)))
        self._h = h
        self.dim, self.capacity, self.device = int(dim), int(capacity), int(device)

    # -- lifetime -------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.pilott_mem_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- bookkeeping ----------------------------------------------------------------------
    def info(self) -> dict:
        total, count, cap = C.c_int64(), C.c_int64(), C.c_int64()
        dim, ld = C.c_int(), C.c_int()
        check(self._lib.pilott_mem_info(self._h, C.byref(total), C.byref(count), C.byref(cap),
                                        C.byref(dim), C.byref(ld)))
        return {"total": total.value, "count": count.value, "capacity": cap.value,
                "dim": dim.value, "row_stride": ld.value}

    @property
    def count(self) -> int:
        return self.info()["count"]

    @property
    def total(self) -> int:
        return self.info()["total"]

    def set_kernel(self, name: str) -> None:
        check(self._lib.pilott_mem_set_option(self._h, 0, KERNELS[name]))

    def set_ctas(self, n: int) -> None:
        check(self._lib.pilott_mem_set_option(self._h, 1, int(n)))

    def set_list_len(self, n: int) -> None:
        check(self._lib.pilott_mem_set_option(self._h, 3, int(n)))

    def set_finish_form(self, form: int) -> None:
        """0: selective re-score (default); 1: the original kernel that re-scores all 32 survivors."""
        check(self._lib.pilott_mem_set_option(self._h, 9, int(form)))

    def set_profile(self, on: bool) -> None:
        check(self._lib.pilott_mem_set_option(self._h, 4, 1 if on else 0))

    def profile_read(self):
        """(summed scan-kernel milliseconds, number of scans) since the last read."""
        ms, n = C.c_double(), C.c_int64()
        check(self._lib.pilott_mem_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def set_sequence_map(self, stride: int, offset: int) -> None:
        check(self._lib.pilott_mem_set_option(self._h, 5, int(stride)))
        check(self._lib.pilott_mem_set_option(self._h, 6, int(offset)))

    def set_debug(self, flags: int) -> None:
        """Developer switches of the scan kernel: only in a -DPILOTT_DEBUG build (PILOTT_DEBUG=1
        python -m pilottai_b200.build --force); the production library refuses anything but 0."""
        check(self._lib.pilott_mem_set_option(self._h, 2, int(flags)))

    def launch_count(self) -> int:
        return int(self._lib.pilott_mem_launch_count(self._h))

    def flagged_total(self, reset: bool = False) -> int:
        """Queries the fused path flagged as not provably exact since creation / the last reset
        (pilott_mem_flagged_total; synchronises the device)."""
        n = C.c_int64()
        check(self._lib.pilott_mem_flagged_total(self._h, C.byref(n), 1 if reset else 0))
        return n.value

    def clear(self) -> None:
        check(self._lib.pilott_mem_clear(self._h, None))

    def _stream(self):
        torch = _torch()
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- append ---------------------------------------------------------------------------
    def append(self, x, prio=None, tagmask=None, expires_us=None) -> int:
        """Append rows.  x: [n, dim] fp32 (numpy -> host path, CUDA tensor -> device path).
        Returns the insertion sequence number of the first row."""
        first = C.c_int64()
        if isinstance(x, np.ndarray):
            x = np.ascontiguousarray(np.atleast_2d(x), dtype=np.float32)
            n = x.shape[0]
            if x.shape[1] != self.dim:
                raise ValueError(f"expected vectors of dim {self.dim}, got {x.shape[1]}")
            p = None if prio is None else np.ascontiguousarray(prio, dtype=np.int32)
            t = None if tagmask is None else np.ascontiguousarray(tagmask, dtype=np.uint64)
            e = None if expires_us is None else np.ascontiguousarray(expires_us, dtype=np.int64)
            for a in (p, t, e):
                if a is not None and a.shape != (n,):
                    raise ValueError("metadata columns must have one entry per row")
            check(self._lib.pilott_mem_append_host(self._h, ptr(x), n, ptr(p), ptr(t), ptr(e), C.byref(first)))
            return first.value
        torch = _torch()
        if x.dim() == 1:
            x = x[None]
        if not x.is_cuda or x.dtype != torch.float32 or x.shape[1] != self.dim:
            raise ValueError("device append needs a CUDA fp32 tensor [n, dim]")
        x = x.contiguous()
        n = x.shape[0]

        def col(a, dt):
            if a is None:
                return None
            a = a.to(device=x.device, dtype=dt).contiguous()
            if a.shape != (n,):
                raise ValueError("metadata columns must have one entry per row")
            return a
        p, e = col(prio, torch.int32), col(expires_us, torch.int64)
        t = None
        if tagmask is not None:                    # uint64 masks travel as int64 bit patterns
            t = tagmask.to(device=x.device).contiguous()
            if t.dtype not in (torch.int64, torch.uint64) or t.shape != (n,):
                raise ValueError("tagmask must be int64/uint64 [n]")
        check(self._lib.pilott_mem_append(self._h, ptr(x), n, ptr(p), ptr(t), ptr(e), C.byref(first),
                                          self._stream()))
        return first.value

    # -- search ---------------------------------------------------------------------------
    def search(self, q, k: int, min_priority: int = 0, req_tagmask: int = 0, now_us: int = 0,
               want_score64: bool = False):
        """Top-k per query.  numpy in -> numpy out (host path, exact fallback included);
        CUDA tensor in -> CUDA tensors out plus the device-side inexact counter."""
        if isinstance(q, np.ndarray):
            q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
            nq = q.shape[0]
            if q.shape[1] != self.dim:
                raise ValueError(f"expected queries of dim {self.dim}, got {q.shape[1]}")
            idx = np.empty((nq, k), np.int64)
            sc = np.empty((nq, k), np.float32)
            sc64 = np.empty((nq, k), np.float64) if want_score64 else None
            nfb = C.c_int32()
            check(self._lib.pilott_mem_search_host(self._h, ptr(q), nq, int(k), int(min_priority),
                                                   int(req_tagmask), int(now_us), ptr(idx), ptr(sc),
                                                   ptr(sc64), C.byref(nfb)))
            self.last_fallbacks = nfb.value
            return (idx, sc, sc64) if want_score64 else (idx, sc)
        torch = _torch()
        if q.dim() == 1:
            q = q[None]
        if not q.is_cuda or q.dtype != torch.float32 or q.shape[1] != self.dim:
            raise ValueError("device search needs a CUDA fp32 tensor [nq, dim]")
        q = q.contiguous()
        nq = q.shape[0]
        idx = torch.empty((nq, k), dtype=torch.int64, device=q.device)
        sc = torch.empty((nq, k), dtype=torch.float32, device=q.device)
        sc64 = torch.empty((nq, k), dtype=torch.float64, device=q.device) if want_score64 else None
        inexact = torch.zeros(1, dtype=torch.int32, device=q.device)
        flags = torch.empty(nq, dtype=torch.uint8, device=q.device)
        check(self._lib.pilott_mem_search(self._h, ptr(q), nq, int(k), int(min_priority), int(req_tagmask),
                                          int(now_us), ptr(idx), ptr(sc), ptr(sc64), ptr(inexact), ptr(flags),
                                          self._stream()))
        return idx, sc, sc64, inexact, flags

    def search_host_begin(self, slot: int, q: np.ndarray, k: int, min_priority: int = 0, req_tagmask: int = 0,
                          now_us: int = 0) -> None:
        """First half of the pipelined host path (pilott_mem_search_host_begin): returns at once.
        q: fp32 [nq, dim] numpy array; keep it alive (and unchanged) until search_host_end(slot)."""
        q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        if q.shape[1] != self.dim:
            raise ValueError(f"expected queries of dim {self.dim}, got {q.shape[1]}")
        check(self._lib.pilott_mem_search_host_begin(self._h, int(slot), ptr(q), q.shape[0], int(k), int(min_priority),
                                                     int(req_tagmask), int(now_us)))
        self._host_inflight = getattr(self, "_host_inflight", {})
        self._host_inflight[slot] = (q, k)

    def search_host_end(self, slot: int, want_score64: bool = False):
        """Second half: waits for the slot, repairs flagged queries, returns numpy (idx, score[, score64])."""
        if slot not in getattr(self, "_host_inflight", {}):
            raise ValueError(f"slot {slot} is not in flight")
        q, k = self._host_inflight.pop(slot)
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int64)
        sc = np.empty((nq, k), np.float32)
        sc64 = np.empty((nq, k), np.float64) if want_score64 else None
        nfb = C.c_int32()
        check(self._lib.pilott_mem_search_host_end(self._h, int(slot), ptr(idx), ptr(sc), ptr(sc64), C.byref(nfb)))
        self.last_fallbacks = nfb.value
        return (idx, sc, sc64) if want_score64 else (idx, sc)

    def search_into(self, q, k, idx, sc, sc64, inexact, flags, min_priority=0, req_tagmask=0, now_us=0):
        """Device search writing into caller-owned tensors (no allocation in the hot loop)."""
        check(self._lib.pilott_mem_search(self._h, ptr(q), q.shape[0], int(k), int(min_priority),
                                          int(req_tagmask), int(now_us), ptr(idx), ptr(sc), ptr(sc64),
                                          ptr(inexact), ptr(flags), self._stream()))

    def search_publish_into(self, xchg, step: int, q, k, idx, sc, sc64, inexact, flags, min_priority=0,
                            req_tagmask=0, now_us=0):
        """search_into for a row-sharded store: finish_kernel also stores the result rows into every
        rank's exchange buffer and releases the step's flags (pilott_mem_search_publish)."""
        check(self._lib.pilott_mem_search_publish(self._h, xchg, int(step), ptr(q), q.shape[0], int(k),
                                                  int(min_priority), int(req_tagmask), int(now_us), ptr(idx), ptr(sc),
                                                  ptr(sc64), ptr(inexact), ptr(flags), self._stream()))

    def search_large(self, q, k: int, min_priority: int = 0, req_tagmask: int = 0, now_us: int = 0, out=None):
        """Top-k for 32 < k <= 1024 on the device (pilott_mem_search_large): CUDA fp32 [nq, dim] in,
        (idx, score, score64) CUDA tensors out (or written into `out`); self.last_fallbacks = queries
        that needed brute force."""
        torch = _torch()
        q = q.contiguous()
        nq = q.shape[0]
        if out is not None:
            idx, sc, sc64 = out
        else:
            idx = torch.empty((nq, k), dtype=torch.int64, device=q.device)
            sc = torch.empty((nq, k), dtype=torch.float32, device=q.device)
            sc64 = torch.empty((nq, k), dtype=torch.float64, device=q.device)
        nb = C.c_int32()
        check(self._lib.pilott_mem_search_large(self._h, ptr(q), nq, int(k), int(min_priority), int(req_tagmask),
                                                int(now_us), ptr(idx), ptr(sc), ptr(sc64), C.byref(nb), self._stream()))
        self.last_fallbacks = nb.value
        return idx, sc, sc64

    def search_exact(self, q, k: int, query_ids=None, out=None, min_priority: int = 0, req_tagmask: int = 0,
                     now_us: int = 0):
        """Brute-force fp64 search on the device (any k).  q: CUDA fp32 [nq, dim]."""
        torch = _torch()
        q = q.contiguous()
        nq = q.shape[0]
        if out is None:
            idx = torch.full((nq, k), -1, dtype=torch.int64, device=q.device)
            sc = torch.full((nq, k), float("-inf"), dtype=torch.float32, device=q.device)
            sc64 = torch.full((nq, k), float("-inf"), dtype=torch.float64, device=q.device)
        else:
            idx, sc, sc64 = out
        ids = None
        nsel = nq
        if query_ids is not None:
            ids = query_ids.to(device=q.device, dtype=torch.int32).contiguous()
            nsel = ids.numel()
        check(self._lib.pilott_mem_search_exact(self._h, ptr(q), nq, ptr(ids), nsel, int(k), int(min_priority),
                                                int(req_tagmask), int(now_us), ptr(idx), ptr(sc), ptr(sc64),
                                                self._stream()))
        return idx, sc, sc64

    def refine(self, q, k: int, query_ids, out, min_priority: int = 0, req_tagmask: int = 0, now_us: int = 0) -> int:
        """Second-chance pass for the queries a fused search flagged (pilott_mem_search_refine):
        `out` = (idx, score, score64) tensors still holding that search's results, `query_ids` the
        flagged rows (e.g. flags.nonzero()).  Overwrites those rows of `out` with the exact answer and
        returns how many queries needed the brute-force path.  Synchronises the stream."""
        torch = _torch()
        idx, sc, sc64 = out
        if sc64 is None:
            raise ValueError("refine needs the fp64 scores of the fused search (pass sc64)")
        ids = query_ids.to(device=q.device, dtype=torch.int32).contiguous().view(-1)
        nb = C.c_int32()
        check(self._lib.pilott_mem_search_refine(self._h, ptr(q.contiguous()), q.shape[0], ptr(ids), ids.numel(), int(k),
                                                 int(min_priority), int(req_tagmask), int(now_us), ptr(idx), ptr(sc),
                                                 ptr(sc64), C.byref(nb), self._stream()))
        return nb.value

    def search_resolved(self, q, k: int, min_priority: int = 0, req_tagmask: int = 0, now_us: int = 0):
        """Device search with the repair step included: CUDA fp32 [nq, dim] in, (idx, score, score64)
        CUDA tensors out, every row provably exact.  One host sync (reads the inexact counter)."""
        idx, sc, sc64, inexact, flags = self.search(q, k, min_priority, req_tagmask, now_us, want_score64=True)
        self.last_fallbacks = int(inexact.item())
        if self.last_fallbacks:
            self.refine(q, k, flags.nonzero(), (idx, sc, sc64), min_priority, req_tagmask, now_us)
        return idx, sc, sc64

    def search_exact_noalloc(self, q, k: int, out, query_ids=None, min_priority: int = 0,
                             req_tagmask: int = 0, now_us: int = 0):
        """Exact fp64 search that allocates nothing: for callers that cannot touch the CUDA
        allocator while they run (the asyncio loop thread, where cudaMalloc's device-wide
        sync would stall everything else; processes whose framework allocator already owns
        HBM).  Same results as search_exact.  The caller passes every buffer: q (CUDA fp32
        [nq, dim], contiguous), out = (idx int64, score fp32, score64 fp64) CUDA tensors
        [nq, k], query_ids an int32 CUDA tensor or None.  Scratch is the library's fixed
        16 MiB device array, so nq * row_stride * 2 + n_selected * rows_held * 8 bytes must
        fit in it (PilottError otherwise); calls that overlap on the GPU share that array and
        race.  Its one-time set-up per device runs when a store is created, not in this call."""
        idx, sc, sc64 = out
        nsel = q.shape[0] if query_ids is None else query_ids.numel()
        check(self._lib.pilott_mem_search_exact_noalloc(
            self._h, ptr(q), q.shape[0], ptr(query_ids), nsel, int(k), int(min_priority), int(req_tagmask),
            int(now_us), ptr(idx), ptr(sc), ptr(sc64), self._stream()))
        return idx, sc, sc64

    # -- maintenance ----------------------------------------------------------------------
    def compact(self, now_us: int) -> Tuple[int, np.ndarray]:
        n = self.count
        keep = np.ones(n, dtype=np.uint8)
        kept = C.c_int64()
        check(self._lib.pilott_mem_compact(self._h, int(now_us), C.byref(kept), ptr(keep) if n else None))
        return kept.value, keep.astype(bool)

    def export(self) -> dict:
        n = self.count
        out = {"vec_bits": np.empty((n, self.dim), np.uint16), "inv_norm": np.empty(n, np.float32),
               "prio": np.empty(n, np.int32), "tagmask": np.empty(n, np.uint64),
               "expires_us": np.empty(n, np.int64)}
        if n:
            check(self._lib.pilott_mem_export(self._h, ptr(out["vec_bits"]), ptr(out["inv_norm"]),
                                              ptr(out["prio"]), ptr(out["tagmask"]), ptr(out["expires_us"])))
        return out


def save_snapshot(store: "SemanticStore", path: str) -> None:
    """Corpus snapshot (the reference has no persistence): the held rows in
    age order -- bf16 bit patterns and metadata columns -- plus the ring bookkeeping."""
    ex = store.export()
    info = store.info()
    np.savez(path, vec_bits=ex["vec_bits"], prio=ex["prio"], tagmask=ex["tagmask"], expires_us=ex["expires_us"],
             meta=np.array([info["dim"], info["capacity"], info["total"]], dtype=np.int64))


def load_snapshot(path: str, device: int = 0, capacity: Optional[int] = None) -> "SemanticStore":
    """Rebuilds a store from save_snapshot().  bf16 -> fp32 is exact and fp32 -> bf16 (RNE) of an
    exactly representable value gives the same bits back, so rows, inverse norms and therefore
    every search result are bit-identical to the original store's."""
    z = np.load(path if path.endswith(".npz") else path + ".npz")
    dim, cap, total = (int(v) for v in z["meta"])
    st = SemanticStore(dim, capacity or cap, device)
    x = (z["vec_bits"].astype(np.uint32) << 16).view(np.float32)
    step = max(1, min(st.capacity, 1 << 18))
    for lo in range(0, x.shape[0], step):
        hi = min(x.shape[0], lo + step)
        st.append(x[lo:hi], z["prio"][lo:hi], z["tagmask"][lo:hi], z["expires_us"][lo:hi])
    check(st._lib.pilott_mem_set_option(st._h, 7, max(total, st.count)))
    return st


def topk_merge(idx, score64, k: Optional[int] = None):
    """[shards, nq, k] per-shard results -> global top-k (device tensors)."""
    torch = _torch()
    lib = _lib.load()
    shards, nq, kk = idx.shape
    k = kk if k is None else k
    assert k == kk
    idx, score64 = idx.contiguous(), score64.contiguous()
    o_idx = torch.empty((nq, k), dtype=torch.int64, device=idx.device)
    o_sc = torch.empty((nq, k), dtype=torch.float32, device=idx.device)
    o_sc64 = torch.empty((nq, k), dtype=torch.float64, device=idx.device)
    st = C.c_void_p(torch.cuda.current_stream(idx.device).cuda_stream)
    check(lib.pilott_topk_merge(ptr(idx), ptr(score64), shards, nq, k, ptr(o_idx), ptr(o_sc), ptr(o_sc64), st))
    return o_idx, o_sc, o_sc64


def topk_merge_packed(packed, out=None):
    """[shards, 2, nq, k] int64 words (plane 0 indices, plane 1 fp64 score bits) -> global top-k."""
    torch = _torch()
    lib = _lib.load()
    shards, two, nq, k = packed.shape
    assert two == 2 and packed.dtype == torch.int64 and packed.is_contiguous()
    if out is None:
        out = (torch.empty((nq, k), dtype=torch.int64, device=packed.device),
               torch.empty((nq, k), dtype=torch.float32, device=packed.device),
               torch.empty((nq, k), dtype=torch.float64, device=packed.device))
    st = C.c_void_p(torch.cuda.current_stream(packed.device).cuda_stream)
    check(lib.pilott_topk_merge_packed(ptr(packed), shards, nq, k, ptr(out[0]), ptr(out[1]), ptr(out[2]), st))
    return out
