#!/usr/bin/env python
"""bench.py -- headline benchmark: semantic-memory queries/sec (1M x 768, top-10).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one 1024-query batch searched against the whole corpus (BASELINE.json
configs[1]: 1M entries x 768-d synthetic fp32 -> bf16, top-10).  For N > 1 the same corpus is
row-sharded over the N GPUs (launch with torch.distributed.run): local fused search, NCCL
all-gather of the per-GPU top-k, merge kernel -- strong scaling.

Prints ONE JSON line on rank 0.  Headline keys: `value` is device-resident throughput (queries
already in HBM), `e2e` the same metric through the host-buffer entry point with the H2D/D2H copies
inside the timed region, `roofline` the tcgen05 scan kernel against the measured bf16 peaks
(MEASURED_PEAKS.json: burst and sustained, `frac` taken against the one the timed region's length
puts it in), `cpu_baseline` the CPU oracle port (numpy BLAS fp32 + argpartition) on a bounded
sample plus the reference's OWN CPU path on BASELINE configs[0] (`reference_path`).

Also on the line, measured in the same run at the same N:
  parity     -- rank 0 re-scores a fixed sample of one more (untimed) batch, planted exact
                duplicates included, with the CPU oracle over the FULL corpus and compares indices
                and fp64 scores of the merged result bit for bit; a mismatch exits non-zero;
  sustained  -- (N = 1) >= 3 s of back-to-back batches, so the power-capped regime is on record;
  secondary  -- BASELINE configs[2] (10M x 1024 row-sharded: value, ms_per_step, kernel_ms,
                roofline, its own parity sample) and configs[3] (router scoring, 1M dict tasks x
                1024 agents, with a parity sample against the routing oracle).
`--impl reference` times the CPU port alone (rank 0; BLAS threads set explicitly).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NQ, TOPK = 1024, 10
WORKLOADS = {
    "1m": {"rows": 1_000_000, "dim": 768,
           "metric": "semantic-memory queries/sec (1M x 768, top-10)",
           "workload": "semantic retrieval 1M entries x 768-d synthetic fp32->bf16, 1024-query batches, top-10"},
    "10m": {"rows": 10_000_000, "dim": 1024,
            "metric": "semantic-memory queries/sec (10M x 1024, top-10)",
            "workload": "semantic retrieval 10M entries x 1024-d bf16 row-sharded, 1024-query batches, top-10"},
}
Q_POOL = 8                     # distinct query batches cycled through the timed loop
FILL_CHUNK = 131072
N_PLANTED = 64                 # corpus rows 0..63 are copied to 64 other positions (exact score ties)
BURST_SECONDS = 1.0            # a timed region shorter than this runs at burst clocks (no power capping yet)


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    fb = {"burst": 1400.0, "sustained": 1400.0, "hbm": 6650.0, "src": "fallback (B200_PROFILING.md)"}
    try:
        with open(path) as f:
            p = json.load(f)
        burst = float(p.get("bf16_tflops") or p.get("bf16_tflops_burst") or fb["burst"])
        return {"burst": burst, "sustained": float(p.get("bf16_tflops_sustained") or burst),
                "hbm": float(p.get("hbm_gbs") or fb["hbm"]), "src": "MEASURED_PEAKS.json"}
    except (OSError, ValueError, TypeError):
        return fb


def _traffic(rows, dim, world):
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if world == 1 and rows == 1_000_000 and dim == 768 and os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("scan_tc_dram_bytes_per_launch")
    return None


def _host_threads() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


class _blas_threads:
    """numpy's BLAS on every host core this process may use, whatever OMP_NUM_THREADS says
    (torch.distributed.run sets it to 1 before numpy loads)."""

    def __enter__(self):
        self.n = _host_threads()
        self._ctx = None
        try:
            from threadpoolctl import threadpool_limits
            self._ctx = threadpool_limits(limits=self.n, user_api="blas")
            self._ctx.__enter__()
        except Exception:
            self._ctx = None
        return self.n

    def __exit__(self, *a):
        if self._ctx is not None:
            self._ctx.__exit__(*a)
        return False


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled through NVML while the timed region runs
    (same fields as the nvidia-smi clocks line in the profiling recipe)."""

    def __init__(self, index: int, interval: float = 0.001):
        super().__init__(daemon=True)
        self.index, self.sm, self.reasons, self.max_sm, self.interval = index, [], 0, None, interval
        self._stop_evt = threading.Event()
        self._ready = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_sm = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            self._nv = nv
            self._ready.set()
            while not self._stop_evt.is_set():
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                self.reasons |= nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                self._stop_evt.wait(self.interval)
        except Exception:
            self._nv = None
            self._ready.set()

    def begin(self):
        self.start()
        self._ready.wait(5)          # NVML is initialised before the timed region starts
        return self

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=5)
        if not self.sm or self._nv is None:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": []}
        nv = self._nv
        names = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_sm,
                "reasons": [k for k, bit in names.items() if self.reasons & bit], "samples": len(sm)}


def _config(wl, world: int, exchange: str):
    """The `config` object of the JSON line -- the same for both arms (the reference arm times the
    CPU port on this configuration; how it samples it is said in cpu_baseline.sample)."""
    rows, dim = wl["rows"], wl["dim"]
    return {"workload": wl["workload"], "rows": rows, "dim": dim, "queries_per_step": NQ, "k": TOPK,
            "sharding": f"rows round-robin over {world} GPU(s), exchange {exchange}" if world > 1 else "single GPU",
            "l2": f"corpus ({rows * dim * 2 / 1e9:.2f} GB bf16) is larger than L2; 8 distinct query batches cycled"}


def _planted_positions(rows: int):
    """Row j (j < N_PLANTED) is copied to position dst[j]: spread over the corpus and over the ranks."""
    import numpy as np
    n = min(N_PLANTED, rows // 4)
    stride = rows // (n + 1) if n else 0
    return np.array([(j + 1) * stride + 7 + (j % 5) for j in range(n)], dtype=np.int64)


# ------------------------------------------------------------------------------------------
def _cpu_port(corpus, inv, qs, nsamp_target_s: float):
    """numpy BLAS fp32 q @ M.T + argpartition + sort (oracle port) on a bounded sample."""
    from oracle.retrieval_oracle import fast_fp32_search
    with _blas_threads() as nthr:
        fast_fp32_search(corpus, inv, qs[:8], TOPK)              # warm BLAS
        t0 = time.perf_counter()
        fast_fp32_search(corpus, inv, qs[:16], TOPK)
        probe = time.perf_counter() - t0
        nsamp = int(max(16, min(qs.shape[0], 16 * (nsamp_target_s / max(probe, 1e-3)))))
        t0 = time.perf_counter()
        for lo in range(0, nsamp, 64):
            fast_fp32_search(corpus, inv, qs[lo:min(nsamp, lo + 64)], TOPK)
        cdt = time.perf_counter() - t0
    return nsamp / cdt, nsamp, cdt, nthr


def run_reference(args, wl):
    """CPU arm: the oracle port of the path (numpy BLAS fp32 q @ M.T, argpartition, sort; the
    reference itself has no vector search to time -- SURVEY.md section 0) on the host cores, plus
    the reference's own CPU code on BASELINE configs[0] (cpu_baseline.reference_path)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from oracle.retrieval_oracle import fast_fp32_search
    from oracle import reference_path
    from pilottai_b200 import synth
    rows, dim = wl["rows"], wl["dim"]
    sample_q = 32
    corpus = synth.corpus(rows, dim, seed=1234)
    inv = (1.0 / np.linalg.norm(corpus, axis=1)).astype(np.float32)
    q = synth.queries(NQ, dim, seed=5678)
    with _blas_threads() as nthr:
        for w in range(args.warmup):
            fast_fp32_search(corpus, inv, q[:sample_q], TOPK)
        t0 = time.perf_counter()
        for s in range(args.steps):
            lo = (s * sample_q) % (NQ - sample_q + 1)
            fast_fp32_search(corpus, inv, q[lo:lo + sample_q], TOPK)
        dt = time.perf_counter() - t0
    v = sample_q * args.steps / dt
    line = {
        "impl": "reference", "metric": wl["metric"], "value": v, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": _config(wl, int(os.environ.get("WORLD_SIZE", "1")), os.environ.get("PILOTT_EXCHANGE", "nccl")),
        "cpu_baseline": {"value": v, "unit": "queries/s", "cores": nthr, "blas_threads": nthr, "kind": "port",
                         "sample": f"each step = a {sample_q}-query sample of a {NQ}-query batch against all {rows} rows on the host cores "
                                   f"of rank 0 (numpy BLAS fp32 q@M.T + argpartition + sort), {args.steps} steps",
                         "reference_path": reference_path.run()},
        "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
class _HostCorpus:
    """What oracle.retrieval_c.search_rows needs: the full corpus as bf16 bit patterns in slot order."""

    def __init__(self, rows, dim):
        import numpy as np
        self.capacity, self.dim = rows, dim
        self.vec = np.empty((rows, dim), np.uint16)
        self.prio = np.zeros(rows, np.int32)
        self.tagmask = np.zeros(rows, np.uint64)
        self.expires_us = np.zeros(rows, np.int64)

    def slot_seq(self):
        import numpy as np
        return np.arange(self.capacity, dtype=np.int64)


def run_workload(name, args, ctx, full: bool):
    """Builds the store for one workload, times it, checks a sample against the oracle.
    full = headline block (e2e, sustained, cpu baseline); else the secondary one."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from pilottai_b200.sharded import ShardedSemanticStore

    wl = WORKLOADS[name]
    rows, dim = wl["rows"], wl["dim"]
    world, rank, local, dev = ctx["world"], ctx["rank"], ctx["local"], ctx["dev"]
    peaks = _peaks()

    # ---- corpus + queries (synthetic) -------------------------------
    want_oracle = rank == 0 and not args.no_parity
    host = None
    if want_oracle:
        try:
            import psutil
            avail = psutil.virtual_memory().available
        except Exception:
            avail = 1 << 62
        need = rows * dim * 2 * 1.3 + (8 << 30)
        if avail > need:
            host = _HostCorpus(rows, dim)
        else:
            ctx["notes"].append(f"{name}: parity sample skipped, host memory {avail >> 30} GiB < {int(need) >> 30} GiB")
    store = ShardedSemanticStore(dim, rows, rank, world, local)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)
    dst = _planted_positions(rows)
    dst_t = torch.from_numpy(dst).to(dev)
    planted = None
    for lo in range(0, rows, FILL_CHUNK):
        m = min(FILL_CHUNK, rows - lo)
        x = torch.randn(m, dim, generator=gen, device=dev)
        x = x / x.norm(dim=1, keepdim=True)
        if lo == 0:
            planted = x[:dst.size].clone()
        sel = ((dst_t >= lo) & (dst_t < lo + m)).nonzero().view(-1)
        if sel.numel():
            x[dst_t[sel] - lo] = planted[sel]                   # exact copies: score ties the index must break
        store.append(x)
        if host is not None:
            host.vec[lo:lo + m] = x.to(torch.bfloat16).view(torch.int16).cpu().numpy().view(np.uint16)
    gen.manual_seed(5678)
    qpool = torch.randn(Q_POOL, NQ, dim, generator=gen, device=dev)
    qpool = qpool / qpool.norm(dim=2, keepdim=True)
    torch.cuda.synchronize()
    loc = store.local
    n_local = loc.count

    idx = torch.empty((NQ, TOPK), dtype=torch.int64, device=dev)
    sc = torch.empty((NQ, TOPK), dtype=torch.float32, device=dev)
    sc64 = torch.empty((NQ, TOPK), dtype=torch.float64, device=dev)
    flags = torch.empty(NQ, dtype=torch.uint8, device=dev)
    inexact = torch.zeros(1, dtype=torch.int32, device=dev)
    pending = []
    depth = 1 if store.exchange == "p2p" else 3      # batches in flight behind the one being searched

    def step(i: int, q=None):
        q = qpool[i % Q_POOL] if q is None else q
        if world == 1:
            loc.search_into(q, TOPK, idx, sc, sc64, inexact, flags)
            return
        # sharded: local search, one packed all-gather (async), merge of an EARLIER batch -- the
        # exchange of batch i overlaps the local searches of the batches behind it; the merge kernel
        # runs on the store's side stream (wait=False: nothing on this stream consumes the result)
        pending.append(store.search_begin(q, TOPK, slot=i % (depth + 1)))
        if len(pending) > depth:
            store.search_end(pending.pop(0), wait=False)

    def drain():
        out = None
        while pending:
            out = store.search_end(pending.pop(0))       # (waits for the merge: the caller reads the result)
        return out

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def flagged_everywhere() -> int:
        """Queries the gap check could not prove, summed over every search since the last call and
        over all ranks (the library counts them per handle; nothing is lost between steps)."""
        n = torch.tensor([loc.flagged_total(reset=True)], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(n)
        return int(n.item())

    def timed(nsteps: int, first: int, sample_clocks, profile: bool):
        # sample_clocks: False, True (a sampler for this pass alone) or a running ClockSampler the caller stops
        # profile=True records CUDA events around every scan kernel, on its stream; an event between
        # two kernels also serialises them, which turns off the programmatic dependent launches the
        # step otherwise uses -- so throughput is taken from an un-profiled pass and the kernel time
        # from a profiled pass of the same steps.
        loc.set_profile(profile)
        launches0 = loc.launch_count()
        sampler = ClockSampler(local).begin() if (sample_clocks is True and rank == 0) else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for i in range(nsteps):
            step(first + i)
        drain()
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        scan_ms, n_scans = loc.profile_read() if profile else (0.0, 0)
        loc.set_profile(False)
        # + the merge kernel of the exchange (the peer-memory publish is part of the finish kernel)
        launches = loc.launch_count() - launches0 + (nsteps if world > 1 else 0)
        return float(ms.item()), scan_ms / max(1, n_scans), int(launches), clocks

    def roofline_of(kernel_ms: float, region_s: float, clk=None):
        flops = 2.0 * n_local * dim * NQ                        # algorithmic: 2*N*D per query
        achieved = flops / (kernel_ms * 1e-3) / 1e12
        alg_bytes = 2.0 * n_local * dim + 4.0 * n_local + 2.0 * NQ * dim + 12.0 * NQ * TOPK
        # burst = short region at full clocks; a region of a second or more, or one the clock samples
        # show power-capped well below the maximum clock (16 ms launches back to back get there in a
        # few steps), is measured against the sustained peak
        capped = bool(clk and clk.get("sm_mhz") and clk.get("sm_max_mhz") and "sw_power_cap" in clk.get("reasons", [])
                      and clk["sm_mhz"] < 0.85 * clk["sm_max_mhz"])
        regime = "sustained" if (region_s >= BURST_SECONDS or capped) else "burst"
        return {"bound": "tensor", "achieved": achieved, "peak": peaks[regime], "unit": "TFLOP/s",
                "frac": achieved / peaks[regime], "regime": regime,
                "frac_burst": achieved / peaks["burst"], "frac_sustained": achieved / peaks["sustained"],
                "peak_burst": peaks["burst"], "peak_sustained": peaks["sustained"], "peak_source": peaks["src"],
                "traffic": _traffic(rows, dim, world), "kernel": "scan_tc_kernel<2,8>", "kernel_ms": kernel_ms,
                "kernel_ms_source": "CUDA events around each launch on its stream, profiled pass of the same steps",
                "algorithmic_flops_per_launch": flops, "algorithmic_bytes_per_launch": alg_bytes,
                "hbm_gbs": alg_bytes / (kernel_ms * 1e-3) / 1e9,
                "hbm_frac": alg_bytes / (kernel_ms * 1e-3) / 1e9 / peaks["hbm"]}

    # ---- device-resident throughput -------------------------------------------------------
    for i in range(args.warmup):
        step(i)
    drain()
    barrier()
    flagged_everywhere()
    # one clock sampler over both passes (the same K steps twice): NVML answers in 1-20 ms on these
    # boxes, so a 20 ms region alone can end up with a single sample
    both = ClockSampler(local).begin() if rank == 0 else None
    total_ms, _, launches, _ = timed(args.steps, args.warmup, both, False)
    prof_ms, kernel_ms, _, _ = timed(args.steps, args.warmup, both, True)
    clocks = both.stop() if both else None
    if clocks:
        clocks["covers"] = "the timed pass and the profiled pass of the same steps"
    if world > 1:
        store.check_exchange()
    n_flagged = flagged_everywhere()
    if n_flagged:
        raise SystemExit(f"{name}: {n_flagged} queries failed the exactness gap check on synthetic data: result invalid")
    block = {
        "metric": wl["metric"], "value": NQ * args.steps / (total_ms * 1e-3), "unit": "queries/s",
        "ms_per_step": total_ms / args.steps, "kernel_ms": kernel_ms, "ms_per_step_profiled_pass": prof_ms / args.steps,
        "gpu_launches": launches, "clocks": clocks,
        "roofline": roofline_of(kernel_ms, total_ms * 1e-3, clocks),
        "config": _config(wl, world, store.exchange),
    }

    # ---- parity sample: the timed path on one more batch, merged result vs the oracle over every row ----
    if not args.no_parity:
        nsample = 64 if rows <= 2_000_000 else 16
        qcheck = qpool[0].clone()
        sample = np.linspace(0, NQ - 1, nsample).astype(np.int64)
        nplant = min(nsample // 4, dst.size)
        qcheck[torch.from_numpy(sample[:nplant]).to(dev)] = planted[:nplant]   # queries that ARE corpus rows: two exact hits each
        step(0, qcheck)
        res = drain()
        got_idx, got_sc64 = (idx, sc64) if world == 1 else (res[0], res[2])
        repaired = flagged_everywhere()
        if repaired:                                             # the library's own repair step, then compare
            r = loc.search_resolved(qcheck, TOPK) if world == 1 else store.search(qcheck, TOPK)
            got_idx, got_sc64 = r[0], r[2]
            flagged_everywhere()
        torch.cuda.synchronize()
        par = {"queries": int(nsample), "planted_duplicates": int(nplant), "repaired": int(repaired), "ok": None}
        if rank == 0 and host is not None:
            from oracle import retrieval_c
            from oracle.retrieval_oracle import bf16_round_bits
            gi = got_idx[torch.from_numpy(sample).to(dev)].cpu().numpy()
            gs = got_sc64[torch.from_numpy(sample).to(dev)].cpu().numpy()
            t0 = time.perf_counter()
            qbits = bf16_round_bits(qcheck[torch.from_numpy(sample).to(dev)].cpu().numpy())
            want_idx, want_sc = retrieval_c.search_rows(host, None, TOPK, q_bits=qbits, threads=_host_threads())
            par["oracle_s"] = time.perf_counter() - t0
            par["ok"] = bool(np.array_equal(gi, want_idx) and np.array_equal(gs, want_sc))
            par["index_mismatches"] = int((gi != want_idx).sum())
            par["score_mismatches"] = int((gs != want_sc).sum())
            par["tie_pairs_seen"] = int(sum(1 for j in range(nplant) if set(gi[j, :2].tolist()) == {j, int(dst[j])}))
            par["oracle"] = "oracle/retrieval_oracle.c (search_rows) over all %d rows" % rows
        elif rank == 0:
            par["skipped"] = "not enough host memory for the full corpus"
        ok = torch.tensor([0 if par["ok"] is False else 1], dtype=torch.int32, device=dev)
        if world > 1:
            dist.broadcast(ok, 0)
        block["parity"] = par
        ctx["parity_failed"] = ctx.get("parity_failed", False) or int(ok.item()) == 0

    if not full:
        host = None
        store.close()
        del store
        torch.cuda.empty_cache()
        return block

    # ---- end to end through the host-buffer entry point --------------------------------------
    qhost = torch.empty((Q_POOL, NQ, dim), dtype=torch.float32).pin_memory()
    qhost.copy_(qpool.cpu())
    qnp = qhost.numpy()
    h2d, d2h = NQ * dim * 4, NQ * TOPK * (8 + 4)
    # Both forms keep several batches in flight (two on one GPU, four in the sharded form): the H2D
    # copy of batch i+1 and the D2H copy of batch i-1 overlap the search of batch i.  Every step's inputs come from pinned host memory and every
    # step's results are read back to the host inside the timed region.
    if world == 1:
        d2h = NQ * TOPK * (8 + 8 + 4) + NQ + 8          # idx, fp64 + fp32 scores, flags, counter

        def e2e_run(n):
            for i in range(n):
                loc.search_host_begin(i % 2, qnp[i % Q_POOL], TOPK)      # pilott_mem_search_host_begin
                if i > 0:
                    loc.search_host_end((i - 1) % 2)                    # ... _end: waits, repairs, copies out
            loc.search_host_end((n - 1) % 2)
    else:
        NSLOT = 4          # batches in flight: the host runs up to three steps ahead of the results it reads
        h2d_s, d2h_s = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        qdev = [torch.empty((NQ, dim), dtype=torch.float32, device=dev) for _ in range(NSLOT)]
        pack_bytes = NQ * TOPK * 12 + 8                  # merged idx int64 | score fp32 | inexact counter (sharded.py)
        out_p = [torch.empty(pack_bytes, dtype=torch.uint8).pin_memory() for _ in range(NSLOT)]
        ev_h2d = [torch.cuda.Event() for _ in range(NSLOT)]
        ev_used = [torch.cuda.Event() for _ in range(NSLOT)]
        ev_done = [torch.cuda.Event() for _ in range(NSLOT)]
        ev_d2h = [torch.cuda.Event() for _ in range(NSLOT)]
        d2h = pack_bytes

        def e2e_run(n):
            main = torch.cuda.current_stream(dev)
            tickets = [None] * NSLOT

            def finish(j):
                s = j % NSLOT
                store.search_end(tickets[s], wait=False)             # merge (side stream) into the slot's packed result block
                d2h_s.wait_event(tickets[s][0]["merged"])
                with torch.cuda.stream(d2h_s):
                    out_p[s].copy_(tickets[s][0]["pack"], non_blocking=True)    # ONE D2H: indices, scores, counter
                ev_d2h[s].record(d2h_s)

            for i in range(n):
                s = i % NSLOT
                if i >= NSLOT:
                    ev_d2h[s].synchronize()                          # results of batch i-NSLOT are on the host
                    h2d_s.wait_event(ev_used[s])                     # its queries were consumed
                with torch.cuda.stream(h2d_s):
                    qdev[s].copy_(qhost[i % Q_POOL], non_blocking=True)
                ev_h2d[s].record(h2d_s)
                main.wait_event(ev_h2d[s])
                tickets[s] = store.search_begin(qdev[s], TOPK, slot=s)
                ev_used[s].record(main)
                if i > 0:
                    finish(i - 1)
            finish(n - 1)
            for s in range(min(NSLOT, n)):
                ev_d2h[s].synchronize()
    e2e_run(max(4, min(8, args.warmup)))
    barrier()
    flagged_everywhere()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    n_flagged = flagged_everywhere() if world > 1 else 0            # (one GPU: _end repairs flagged queries itself)
    if n_flagged:
        raise SystemExit(f"{n_flagged} queries failed the exactness gap check on synthetic data: result invalid")
    block["e2e"] = {"value": NQ * args.steps / float(dt.item()), "unit": "queries/s",
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}

    # ---- sustained regime: >= 3 s of back-to-back batches (one GPU) -------------------------------
    if world == 1 and args.sustained_s > 0:
        nsus = int(args.sustained_s * 1e3 / block["ms_per_step"]) + 1
        s_ms, s_kernel_ms, _, s_clocks = timed(nsus, 0, True, True)
        if flagged_everywhere():
            raise SystemExit("sustained run: queries failed the exactness gap check: result invalid")
        rf = roofline_of(s_kernel_ms, s_ms * 1e-3, s_clocks)
        block["sustained"] = {"seconds": s_ms * 1e-3, "steps": nsus, "value": NQ * nsus / (s_ms * 1e-3), "unit": "queries/s",
                              "ms_per_step": s_ms / nsus, "kernel_ms": s_kernel_ms, "regime": rf["regime"],
                              "achieved_tflops": rf["achieved"], "frac_sustained": rf["frac_sustained"],
                              "frac_burst": rf["frac_burst"], "clocks": s_clocks}

    # ---- CPU baselines, rank 0, N = 1 only ----------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import reference_path
        ex = loc.export()
        corpus = (ex["vec_bits"].astype(np.uint32) << 16).view(np.float32)
        v, nsamp, cdt, nthr = _cpu_port(corpus, ex["inv_norm"], qnp[0], 12.0)
        cpu = {"value": v, "unit": "queries/s", "cores": nthr, "blas_threads": nthr, "kind": "port",
               "sample": f"{nsamp} of the {NQ} queries of one batch against all {rows} rows "
                         f"(numpy BLAS fp32 q@M.T + argpartition + sort, {cdt:.1f} s)",
               "reference_path": reference_path.run()}
        del corpus, ex
    block["cpu_baseline"] = cpu
    host = None
    store.close()
    del store
    torch.cuda.empty_cache()
    return block


def run_router(args, ctx):
    """BASELINE configs[3]: 1M pending task DICTS x 1024 agents, load threshold 0.8 -- dicts in,
    assignments out (encode on the host, H2D, route_kernel, D2H), with the kernel timed alone
    beside it and a 400-task sample checked against the routing oracle (hex-exact scores)."""
    import numpy as np
    import torch
    from pilottai_b200 import synth
    from pilottai_b200.routing import RoutingEngine
    from oracle import routing_oracle as ro
    dev = ctx["dev"]
    nt, na = args.router_tasks, 1024
    agents = synth.agent_table(na, seed=42)
    tasks = synth.task_table_fast(nt, seed=43)
    eng = RoutingEngine(agents, ctx["local"])
    t0 = time.perf_counter()
    enc = eng.encode(tasks)
    t_enc = time.perf_counter() - t0
    dev_enc = {k: eng.to_device(enc[k]) for k in ("req", "tags", "type", "flags")}
    for _ in range(3):
        a, s = eng.route_encoded(dev_enc, "router", load_threshold=0.8)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 10
    e0.record()
    for _ in range(iters):
        a, s = eng.route_encoded(dev_enc, "router", load_threshold=0.8)
    e1.record()
    torch.cuda.synchronize()
    k_ms = e0.elapsed_time(e1) / iters
    t0 = time.perf_counter()
    ga, gs = eng.route(tasks, mode="router", load_threshold=0.8, as_numpy=True)      # dicts in -> assignments out
    t_e2e = time.perf_counter() - t0
    sample = np.linspace(0, nt - 1, 400).astype(np.int64)
    bad = 0
    for i in sample:
        want, wscore = ro.router_route(agents, tasks[int(i)], 0.8)
        got = agents[int(ga[i])]["id"] if ga[i] >= 0 else None
        if got != want or (want is not None and float(gs[i]).hex() != float(wscore).hex()):
            bad += 1
    ctx["parity_failed"] = ctx.get("parity_failed", False) or bad > 0
    return {"workload": f"router scoring: {nt} pending task dicts x {na} agents, load_threshold 0.8",
            "kernel": "route_kernel", "kernel_ms": k_ms, "kernel_routes_per_s": nt / (k_ms * 1e-3),
            "pair_evals_per_s": nt * float(na) / (k_ms * 1e-3),
            "encode_s": t_enc, "encode_tasks_per_s": nt / t_enc,
            "e2e_s": t_e2e, "e2e_routes_per_s": nt / t_e2e,
            "assigned_frac": float((ga >= 0).mean()),
            "parity": {"tasks": int(sample.size), "ok": bad == 0, "mismatches": bad,
                       "oracle": "oracle/routing_oracle.py (pinned on reference fixtures)"}}


def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: pilottai_b200 has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # stdout carries exactly one JSON line: anything libraries print there (NCCL's version banner)
    # is sent to stderr for the whole run; the line itself goes to the saved descriptor.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.gpus != world:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    ctx = {"world": world, "rank": rank, "local": local, "dev": dev, "notes": []}

    head = run_workload(args.workload, args, ctx, full=True)
    secondary = {}
    if not args.no_secondary:
        other = "10m" if args.workload == "1m" else "1m"
        t0 = time.perf_counter()
        try:
            sec = run_workload(other, args, ctx, full=False)
            sec["wall_s"] = time.perf_counter() - t0
        except SystemExit:
            raise
        except Exception as e:                                      # secondary: never costs the headline line
            sec = {"error": f"{type(e).__name__}: {e}"}
        secondary["retrieval_" + other] = sec
        if rank == 0 and world == 1:
            try:
                secondary["router"] = run_router(args, ctx)
            except Exception as e:                                  # secondary: never costs the headline line
                secondary["router"] = {"error": f"{type(e).__name__}: {e}"}

    if rank == 0:
        line = {
            "metric": head["metric"], "value": head["value"], "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": head["config"], "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
            "roofline": head["roofline"], "cpu_baseline": head["cpu_baseline"], "clocks": head["clocks"],
            "parity": head.get("parity"), "sustained": head.get("sustained"), "secondary": secondary,
            "notes": ctx["notes"],
        }
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    if ctx.get("parity_failed"):
        raise SystemExit("parity sample differs from the oracle: see the `parity` blocks of the line above")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check of a sample (developer runs)")
    ap.add_argument("--no-secondary", action="store_true", help="headline workload only (developer runs)")
    ap.add_argument("--sustained-s", type=float, default=3.0, help="length of the sustained sub-measurement (N = 1); 0 = off")
    ap.add_argument("--router-tasks", type=int, default=1_000_000)
    ap.add_argument("--workload", default="1m", choices=["1m", "10m"],
                    help="headline workload: 1m = BASELINE configs[1] (default); 10m = configs[2]; the other one runs as `secondary`")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args, WORKLOADS[args.workload])
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
