"""world_size-2 gloo test of the multi-rank host logic bench.py uses: batch sharding with no data-path
collective, barrier + max-over-ranks timing reduction."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = bench.shard_batch(7, world, rank)
    w = bench.make_workload(hi - lo, 16, 2, 16, 2, 2, "encoder", "cpu", seed=100 + rank)
    bench.barrier(world)
    slowest = bench.max_over_ranks(10.0 + rank, world, "cpu")
    # every rank's units together cover the global batch exactly once
    counts = [torch.zeros(1, dtype=torch.long) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([w["value"].shape[0]]))
    out[rank] = (lo, hi, slowest, int(sum(c.item() for c in counts)), float(w["value"].sum()))
    dist.destroy_process_group()


def test_two_rank_sharding_and_timing_reduction():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert out[0][:2] == (0, 4) and out[1][:2] == (4, 7)
    assert out[0][2] == out[1][2] == 11.0          # max over ranks
    assert out[0][3] == out[1][3] == 7             # shards cover the batch
    assert out[0][4] != out[1][4]                  # ranks hold different videos (rank-offset seeds)
