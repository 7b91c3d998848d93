#!/usr/bin/env python
"""bench.py -- headline benchmark: semantic-memory queries/sec (1M x 768, top-10).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one 1024-query batch searched against the whole corpus (BASELINE.json
configs[1]: 1M entries x 768-d synthetic fp32 -> bf16, top-10).  For N > 1 the same corpus is
row-sharded over the N GPUs (launch with torch.distributed.run): local fused search, NCCL
all-gather of the per-GPU top-k, merge kernel -- strong scaling.

Prints ONE JSON line on rank 0.  `value` is device-resident throughput (queries already in
HBM), `e2e` the same metric through the host-buffer entry point with the H2D/D2H copies
inside the timed region, `roofline` the tcgen05 scan kernel against the measured bf16 peak
(MEASURED_PEAKS.json), `cpu_baseline` the CPU oracle port (numpy BLAS fp32 + argpartition)
on a bounded sample.  `--impl reference` times that CPU port alone.
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

N_ROWS, DIM, NQ, TOPK = 1_000_000, 768, 1024, 10
WORKLOAD = "semantic retrieval 1M entries x 768-d synthetic fp32->bf16, 1024-query batches, top-10"
METRIC = "semantic-memory queries/sec (1M x 768, top-10)"
Q_POOL = 8                     # distinct query batches cycled through the timed loop
FILL_CHUNK = 131072


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"tflops": float(p["bf16_tflops_sustained"]), "hbm": float(p["hbm_gbs"]), "src": "measured (sustained)"}
    return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback"}


def _traffic():
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f).get("scan_tc_dram_bytes_per_launch")
    return None


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled through NVML while the timed region runs
    (same fields as the nvidia-smi clocks line in the profiling recipe)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.sm, self.reasons, self.max_sm = index, [], 0, None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_sm = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            while not self._stop_evt.is_set():
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                self.reasons |= nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                self._stop_evt.wait(0.005)
            self._nv = nv
        except Exception:
            self._nv = None

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=5)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": []}
        nv = self._nv
        names = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_sm,
                "reasons": [k for k, bit in names.items() if self.reasons & bit], "samples": len(sm)}


# ------------------------------------------------------------------------------------------
def run_reference(args):
    """CPU arm: the oracle port of the path (numpy BLAS fp32 q @ M.T, argpartition, sort; the
    reference itself has no vector search to time -- SURVEY.md section 0) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from oracle.retrieval_oracle import fast_fp32_search
    from pilottai_b200 import synth
    sample_q = 32
    corpus = synth.corpus(N_ROWS, DIM, seed=1234)
    inv = (1.0 / np.linalg.norm(corpus, axis=1)).astype(np.float32)
    q = synth.queries(NQ, DIM, seed=5678)
    for w in range(args.warmup):
        fast_fp32_search(corpus, inv, q[:sample_q], TOPK)
    t0 = time.perf_counter()
    for s in range(args.steps):
        lo = (s * sample_q) % (NQ - sample_q + 1)
        fast_fp32_search(corpus, inv, q[lo:lo + sample_q], TOPK)
    dt = time.perf_counter() - t0
    v = sample_q * args.steps / dt
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "rows": N_ROWS, "dim": DIM, "queries_per_step": NQ, "k": TOPK,
                   "sharding": "host cores (rank 0 only)",
                   "step": f"{sample_q}-query sample of a {NQ}-query batch against the full corpus"},
        "cpu_baseline": {"value": v, "unit": "queries/s", "cores": cores, "kind": "port",
                         "sample": f"{sample_q} queries/step x {args.steps} steps against all {N_ROWS} rows (numpy BLAS fp32 + argpartition)"},
        "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from pilottai_b200.sharded import ShardedSemanticStore

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

    # ---- corpus + queries (synthetic) -------------------------------
    store = ShardedSemanticStore(DIM, N_ROWS, rank, world, local)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)
    for lo in range(0, N_ROWS, FILL_CHUNK):
        m = min(FILL_CHUNK, N_ROWS - lo)
        x = torch.randn(m, DIM, generator=gen, device=dev)
        x = x / x.norm(dim=1, keepdim=True)
        store.append(x)
    gen.manual_seed(5678)
    qpool = torch.randn(Q_POOL, NQ, DIM, generator=gen, device=dev)
    qpool = qpool / qpool.norm(dim=2, keepdim=True)
    torch.cuda.synchronize()
    loc = store.local
    n_local = loc.count

    idx = torch.empty((NQ, TOPK), dtype=torch.int64, device=dev)
    sc = torch.empty((NQ, TOPK), dtype=torch.float32, device=dev)
    sc64 = torch.empty((NQ, TOPK), dtype=torch.float64, device=dev)
    flags = torch.empty(NQ, dtype=torch.uint8, device=dev)
    inexact = torch.zeros(args.steps + args.warmup + 1, dtype=torch.int32, device=dev)
    pending = []

    def step(i: int):
        if world == 1:
            loc.search_into(qpool[i % Q_POOL], TOPK, idx, sc, sc64, inexact[i:i + 1], flags)
            return
        # sharded: local search, one packed all-gather (async), merge of the PREVIOUS batch -- the
        # exchange of batch i overlaps the local search of batch i+1
        pending.append(store.search_begin(qpool[i % Q_POOL], TOPK, slot=i % 2))
        if len(pending) > 1:
            store.search_end(pending.pop(0))

    def drain():
        while pending:
            store.search_end(pending.pop(0))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput -------------------------------------------------------
    for i in range(args.warmup):
        step(i)
    drain()
    barrier()
    loc.set_profile(True)
    launches0 = loc.launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    drain()
    e1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    scan_ms, n_scans = loc.profile_read()
    loc.set_profile(False)
    # + the merge kernel of the exchange (the peer-memory publish is part of finish_kernel)
    launches = loc.launch_count() - launches0 + (args.steps if world > 1 else 0)
    if world > 1:
        store.check_exchange()
    n_inexact = int(inexact.sum().item()) + sum(int(b["inexact"].item()) for b in store._bufs.values())
    if n_inexact:
        raise SystemExit(f"{n_inexact} queries failed the exactness gap check on synthetic data: result invalid")
    value = NQ * args.steps / (total_ms * 1e-3)

    # ---- end to end through the host-buffer entry point --------------------------------------
    qhost = torch.empty((Q_POOL, NQ, DIM), dtype=torch.float32).pin_memory()
    qhost.copy_(qpool.cpu())
    qnp = qhost.numpy()
    h2d, d2h = NQ * DIM * 4, NQ * TOPK * (8 + 4)
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
        qdev = [torch.empty((NQ, DIM), dtype=torch.float32, device=dev) for _ in range(NSLOT)]
        pack_bytes = NQ * TOPK * 12 + 8                  # merged idx int64 | score fp32 | inexact counter (sharded.py)
        out_p = [torch.empty(pack_bytes, dtype=torch.uint8).pin_memory() for _ in range(NSLOT)]
        out_c = [p[NQ * TOPK * 12: NQ * TOPK * 12 + 4].view(torch.int32) for p in out_p]
        ev_h2d = [torch.cuda.Event() for _ in range(NSLOT)]
        ev_used = [torch.cuda.Event() for _ in range(NSLOT)]
        ev_done = [torch.cuda.Event() for _ in range(NSLOT)]
        ev_d2h = [torch.cuda.Event() for _ in range(NSLOT)]
        d2h = pack_bytes

        def e2e_run(n):
            main = torch.cuda.current_stream(dev)
            tickets = [None] * NSLOT
            flagged = 0

            def finish(j):
                s = j % NSLOT
                store.search_end(tickets[s])                         # merge into the slot's packed result block
                ev_done[s].record(main)
                d2h_s.wait_event(ev_done[s])
                with torch.cuda.stream(d2h_s):
                    out_p[s].copy_(tickets[s][0]["pack"], non_blocking=True)    # ONE D2H: indices, scores, counter
                ev_d2h[s].record(d2h_s)

            for i in range(n):
                s = i % NSLOT
                if i >= NSLOT:
                    ev_d2h[s].synchronize()                          # results of batch i-NSLOT are on the host
                    flagged += int(out_c[s].item())
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
                flagged += int(out_c[s].item())
            if flagged:
                raise SystemExit(f"{flagged} queries failed the exactness gap check on synthetic data: result invalid")
    e2e_run(max(4, min(8, args.warmup)))
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = NQ * args.steps / float(dt.item())

    # ---- roofline of the dominant kernel (tcgen05 scan) ------------------------------------
    peaks = _peaks()
    avg_scan_ms = scan_ms / max(1, n_scans)
    flops = 2.0 * n_local * DIM * NQ                         # algorithmic: 2*N*D per query
    achieved = flops / (avg_scan_ms * 1e-3) / 1e12
    alg_bytes = 2.0 * n_local * DIM + 4.0 * n_local + 2.0 * NQ * DIM + 12.0 * NQ * TOPK
    roofline = {"bound": "tensor", "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                "frac": achieved / peaks["tflops"], "traffic": _traffic() if (world == 1 and N_ROWS == 1_000_000 and DIM == 768) else None,
                "kernel": "scan_tc_kernel<2,8>", "kernel_ms": avg_scan_ms, "peak_source": peaks["src"],
                "hbm_gbs": alg_bytes / (avg_scan_ms * 1e-3) / 1e9, "hbm_frac": alg_bytes / (avg_scan_ms * 1e-3) / 1e9 / peaks["hbm"]}

    # ---- CPU baseline (oracle port) on a bounded sample, rank 0, N = 1 only ----------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.retrieval_oracle import fast_fp32_search
        ex = loc.export()
        corpus = (ex["vec_bits"].astype(np.uint32) << 16).view(np.float32)
        inv = ex["inv_norm"]
        qs = qnp[0]
        fast_fp32_search(corpus, inv, qs[:8], TOPK)              # warm BLAS
        t0 = time.perf_counter()
        fast_fp32_search(corpus, inv, qs[:16], TOPK)
        probe = time.perf_counter() - t0
        nsamp = int(max(16, min(NQ, 16 * (12.0 / max(probe, 1e-3)))))
        t0 = time.perf_counter()
        for lo in range(0, nsamp, 64):
            fast_fp32_search(corpus, inv, qs[lo:min(nsamp, lo + 64)], TOPK)
        cdt = time.perf_counter() - t0
        cpu = {"value": nsamp / cdt, "unit": "queries/s", "cores": os.cpu_count() or 1, "kind": "port",
               "sample": f"{nsamp} of the {NQ} queries of one batch against all {N_ROWS} rows "
                         f"(numpy BLAS fp32 q@M.T + argpartition + sort, {cdt:.1f} s)"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows": N_ROWS, "dim": DIM, "queries_per_step": NQ, "k": TOPK,
                       "sharding": f"rows round-robin over {world} GPU(s), exchange {store.exchange}" if world > 1 else "single GPU",
                       "l2": f"corpus ({N_ROWS * DIM * 2 / 1e9:.2f} GB bf16) is larger than L2; 8 distinct query batches cycled"},
            "e2e": {"value": e2e_value, "unit": "queries/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
        }
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    store.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="1m", choices=["1m", "10m"],
                    help="1m = BASELINE configs[1] (default, the headline); 10m = configs[2]: 10M x 1024 row-sharded")
    args = ap.parse_args()
    if args.workload == "10m":
        global N_ROWS, DIM, WORKLOAD, METRIC
        N_ROWS, DIM = 10_000_000, 1024
        METRIC = "semantic-memory queries/sec (10M x 1024, top-10)"
        WORKLOAD = "semantic retrieval 10M entries x 1024-d bf16 row-sharded, 1024-query batches, top-10"
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
