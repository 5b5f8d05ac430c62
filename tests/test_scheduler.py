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


def test_prover_pool_runs_the_aggregation_dag_in_dependency_order():
    import threading
    import time
    started = []
    lock = threading.Lock()

    def make_worker(k):
        return {"id": k}

    def seg(w, i):
        time.sleep(0.002 * (7 - i % 7))            # finish out of order
        with lock:
            started.append(("seg", i))
        return ("seg", i, w["id"])

    def leaf(w, i, proof):
        assert proof[:2] == ("seg", i)
        with lock:
            started.append(("leaf", i))
        return ("leaf", [i])

    def internal(w, i, *kids):
        assert 1 <= len(kids) <= 3 and all(k[0] in ("leaf", "node") for k in kids)
        return ("node", sorted(x for k in kids for x in k[1]))

    pool = scheduler.ProverPool(make_worker, engines=4)
    try:
        root, layers = scheduler.submit_aggregation_tree(pool, 16, seg, leaf, internal)
        assert [len(l) for l in layers] == [16] + scheduler.aggregation_layers(16)
        assert pool.result(root, timeout=30) == ("node", list(range(16)))
        used = {r[2] for r in pool.results(layers[0])}
        assert len(used) > 1                        # more than one engine took segments
        for i in range(16):                         # a leaf never starts before its segment is done
            assert started.index(("seg", i)) < started.index(("leaf", i))
    finally:
        pool.close()


def test_prover_pool_failures_propagate_and_do_not_wedge():
    def boom(w):
        raise ValueError("bad trace")

    pool = scheduler.ProverPool(lambda k: k, engines=2)
    try:
        a = pool.submit(boom)
        b = pool.submit(lambda w, x: x, deps=[a])
        c = pool.submit(lambda w: 7)
        assert pool.result(c, timeout=10) == 7
        import pytest
        with pytest.raises(ValueError, match="bad trace"):
            pool.result(a, timeout=10)
        with pytest.raises(ValueError, match="bad trace"):
            pool.result(b, timeout=10)
        with pytest.raises(ValueError):
            pool.submit(lambda w: 0, deps=[999])
    finally:
        pool.close()
    bad = scheduler.ProverPool(lambda k: (_ for _ in ()).throw(RuntimeError("no device")), engines=1)
    try:
        j = bad.submit(lambda w: 1)
        import pytest
        with pytest.raises(RuntimeError, match="failed to start"):
            bad.result(j, timeout=10)
    finally:
        bad.close()


def _dag_worker(rank, world, port, q):
    import threading
    import time
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    store = dist.distributed_c10d._get_default_store()
    kinds, deps = scheduler.aggregation_dag(7)
    fifo = scheduler.DagFifo(store, deps, tag="t")
    logs = []

    def fn(jid, args):
        # a job sees exactly the results of its dependencies
        assert [int(a.decode().split(":")[0]) for a in args] == deps[jid]
        time.sleep(0.003 * (1 + jid % 3))
        return f"{jid}:{rank}".encode()

    def run():
        logs.extend(fifo.work(fn, clock=time.time))

    ths = [threading.Thread(target=run) for _ in range(2)]
    for t in ths: t.start()
    for t in ths: t.join()
    dist.barrier()
    root = fifo.result(len(deps) - 1).decode()
    q.put((rank, logs, root))
    dist.barrier()
    dist.destroy_process_group()


def test_dag_fifo_shared_by_two_ranks_over_gloo():
    """world_size 2, two engine threads per rank: every job of the 7-segment aggregation DAG runs exactly once somewhere, never
    before its dependencies finished, and both ranks take part (one FIFO, job -> first free engine)."""
    kinds, deps = scheduler.aggregation_dag(7)
    assert [k for k, _, _ in kinds] == ["segment"] * 7 + ["leaf"] * 7 + ["internal"] * 3 + ["internal"]
    assert deps[14] == [7, 8, 9] and deps[16] == [13] and deps[17] == [14, 15, 16]
    assert scheduler.critical_path(deps, [1.0] * len(deps)) == 4.0
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_dag_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs: p.join(timeout=60)
    done_at, ran_on = {}, {}
    for rank, logs, root in res:
        assert root.split(":")[0] == str(len(deps) - 1)
        for jid, t0, t1, t2 in logs:
            assert jid not in done_at, "job ran twice"
            done_at[jid], ran_on[jid] = (t1, t2), rank
    assert sorted(done_at) == list(range(len(deps)))
    for j, d in enumerate(deps):
        for k in d:
            assert done_at[k][1] <= done_at[j][0] + 1e-3, (j, k)
    assert set(ran_on.values()) == {0, 1}


def test_dag_fifo_in_one_process_with_a_hash_store():
    """world size 1 (bench.py at N = 1): the same FIFO over a torch HashStore, four engine threads"""
    import threading
    import time
    kinds, deps = scheduler.aggregation_dag(5)
    fifo = scheduler.DagFifo(dist.HashStore(), deps, tag="solo")
    logs, lock = [], threading.Lock()

    def fn(jid, args):
        assert [int(a) for a in args] == deps[jid]
        time.sleep(0.001)
        return str(jid).encode()

    def run():
        out = fifo.work(fn, clock=time.time)
        with lock:
            logs.extend(out)

    ths = [threading.Thread(target=run) for _ in range(4)]
    for t in ths: t.start()
    for t in ths: t.join()
    assert sorted(j for j, *_ in logs) == list(range(len(deps)))
    done = {j: t2 for j, _, _, t2 in logs}
    ready = {j: t1 for j, _, t1, _ in logs}
    assert all(done[k] <= ready[j] + 1e-3 for j, d in enumerate(deps) for k in d)
    assert fifo.result(len(deps) - 1) == str(len(deps) - 1).encode()
