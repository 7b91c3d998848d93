"""CPU (gloo, world_size 2) test of the multi-GPU exchange logic: round-robin row ownership,
global index mapping, all_gather of per-shard top-k and the merge rule.  The per-shard
searcher here is the CPU oracle (the CUDA search needs a GPU); the GPU twin of this test is
tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.retrieval_oracle import OracleStore
from pilottai_b200 import synth
from pilottai_b200.sharded import shard_rows

N, DIM, NQ, K = 600, 32, 9, 7


def merge_gathered_cpu(idx, score, k):
    """Host restatement of pilott_topk_merge: [shards, Q, k] -> [Q, k] by (score desc, index asc)."""
    shards, nq, kk = idx.shape
    fi = np.transpose(idx, (1, 0, 2)).reshape(nq, shards * kk)
    fs = np.transpose(score, (1, 0, 2)).reshape(nq, shards * kk)
    out_i, out_s = np.full((nq, k), -1, np.int64), np.full((nq, k), -np.inf)
    for q in range(nq):
        ok = fi[q] >= 0
        ci, cs = fi[q][ok], fs[q][ok]
        order = np.lexsort((ci, -cs))[:k]
        out_i[q, : order.size], out_s[q, : order.size] = ci[order], cs[order]
    return out_i, out_s


def test_shard_rows_partitions_every_batch():
    for world in (1, 2, 3, 8):
        for first, n in ((0, 10), (7, 1), (13, 40), (5, 0)):
            owned = [shard_rows(first, n, r, world) for r in range(world)]
            allpos = np.sort(np.concatenate(owned)) if n else np.array([], np.int64)
            assert np.array_equal(allpos, np.arange(n))
            for r, rows in enumerate(owned):
                assert np.all((first + rows) % world == r)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x = synth.corpus(N, DIM, seed=3, dup_frac=0.05)
    q = synth.queries(NQ, DIM, seed=4)
    prio, tag, exp = synth.metadata(N, seed=5)
    local = OracleStore(DIM, N)
    total = 0
    for lo in range(0, N, 170):                       # ragged batches
        hi = min(N, lo + 170)
        rows = lo + shard_rows(total, hi - lo, rank, world)
        local.append(x[rows], prio[rows], tag[rows], exp[rows])
        total += hi - lo
    li, ls = local.search(q, K, min_priority=1, now_us=1_000_000)
    gidx = np.where(li >= 0, li * world + rank, -1)    # sequence map: stride = world, offset = rank
    gi = [torch.zeros(NQ, K, dtype=torch.int64) for _ in range(world)]
    gs = [torch.zeros(NQ, K, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gi, torch.from_numpy(gidx))
    dist.all_gather(gs, torch.from_numpy(ls))
    mi, ms = merge_gathered_cpu(torch.stack(gi).numpy(), torch.stack(gs).numpy(), K)
    if rank == 0:
        np.save(out + "_i.npy", mi); np.save(out + "_s.npy", ms)
    dist.destroy_process_group()


def test_two_rank_exchange_equals_single_store(tmp_path):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "merged")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    x = synth.corpus(N, DIM, seed=3, dup_frac=0.05)
    prio, tag, exp = synth.metadata(N, seed=5)
    whole = OracleStore(DIM, N)
    whole.append(x, prio, tag, exp)
    wi, ws = whole.search(synth.queries(NQ, DIM, seed=4), K, min_priority=1, now_us=1_000_000)
    assert np.array_equal(np.load(out + "_i.npy"), wi)
    assert np.array_equal(np.load(out + "_s.npy"), ws)
