"""Host-side multi-GPU logic on CPU: gloo, world size 2 (no data-path collective exists to test; what is covered is
the partition, the barrier and the max-over-ranks reduction bench.py uses)."""
import os
import socket

import torch.distributed as dist
import torch.multiprocessing as mp

from openvm_b200 import scheduler


def test_partition_covers_all_jobs_once():
    for n in (0, 1, 5, 16, 17):
        for world in (1, 2, 3, 8):
            parts = [scheduler.partition(n, world, r) for r in range(world)]
            assert sorted(j for p in parts for j in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def test_aggregation_layers():
    assert scheduler.aggregation_layers(16) == [16, 6, 2, 1]
    assert scheduler.aggregation_layers(1) == [1]
    assert scheduler.aggregation_layers(3) == [3, 1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = scheduler.partition(7, world, rank)
    dist.barrier()
    slow = scheduler.max_over_ranks(10.0 + rank)
    counts = scheduler.gather_counts(len(mine))
    q.put((rank, mine, slow, counts))
    dist.destroy_process_group()


def test_gloo_world_size_2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs: p.join(timeout=60)
    assert res[0][1] == [0, 2, 4, 6] and res[1][1] == [1, 3, 5]
    assert res[0][2] == res[1][2] == 11.0
    assert res[0][3] == res[1][3] == [4, 3]
