"""One commitment across ranks (openvm_b200/sharded.py). On CPU (gloo, world size 2) the oracle stands in for the device
pieces so that the host logic -- column shards, the row re-shard through all_to_all_single with ragged split sizes, the
sub-tree / top-of-tree composition and the opening paths -- is checked against the oracle's single commitment. The GPU
version of the same check (NCCL, two B200s) is tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from openvm_b200 import sharded

P = 2013265921


def test_column_shards():
    assert sharded.column_shards(8, 2) == [(0, 4), (4, 8)]
    assert sharded.column_shards(7, 4) == [(0, 2), (2, 4), (4, 6), (6, 7)]
    with pytest.raises(ValueError):
        sharded.column_shards(3, 4)


def test_combine_and_top_path():
    import oracle_binding
    orc = oracle_binding.load()
    rng = np.random.default_rng(3)
    leaves = [rng.integers(0, P, size=8, dtype=np.uint64).astype(np.uint32) for _ in range(8)]
    root, levels = sharded.combine_subroots(leaves, orc.compress)
    assert [len(l) for l in levels] == [8, 4, 2, 1]
    for g in range(8):
        node, i = leaves[g], g
        for sib in sharded.top_path(levels, g):
            node = orc.compress(node, sib) if i % 2 == 0 else orc.compress(sib, node)
            i >>= 1
        assert np.array_equal(node, root)
    with pytest.raises(ValueError):
        sharded.combine_subroots(leaves[:3], orc.compress)


def _worker(rank, world, port, q, height, width, b):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import oracle_binding
    orc = oracle_binding.load()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    trace = np.random.default_rng(5).integers(0, P, size=(height, width), dtype=np.uint64).astype(np.uint32)
    lo, hi = sharded.column_shards(width, world)[rank]
    lde = orc.coset_lde(trace[:, lo:hi], b)                        # (rows, w_local), canonical
    flat = torch.from_numpy(np.ascontiguousarray(lde.T).astype(np.int64).astype(np.int32).reshape(-1))
    recv = sharded.reshard_rows(flat, width)                        # (width, r)
    block = np.ascontiguousarray(recv.numpy().astype(np.uint32).T)  # (r, width)
    subroot = orc.mmcs_commit([block])
    gathered = [None] * world
    dist.all_gather_object(gathered, [int(x) for x in subroot])
    root, levels = sharded.combine_subroots([np.array(g, dtype=np.uint32) for g in gathered], orc.compress)
    q.put((rank, [int(x) for x in root], block[:2].tolist(), [[int(x) for x in s] for s in sharded.top_path(levels, rank)]))
    dist.destroy_process_group()


@pytest.mark.parametrize("height,width,b", [(64, 6, 1), (32, 5, 2)])
def test_gloo_sharded_commit_equals_single(height, width, b):
    import oracle_binding
    orc = oracle_binding.load()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, height, width, b)) for r in range(2)]
    for p in procs: p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs: p.join(timeout=60)
    trace = np.random.default_rng(5).integers(0, P, size=(height, width), dtype=np.uint64).astype(np.uint32)
    want_root, want_ldes = orc.commit([trace], b, want_lde=True)
    rows = height << b
    for rank, root, head, top in res:
        assert root == [int(x) for x in want_root]
        assert head == want_ldes[0][rank * rows // 2: rank * rows // 2 + 2].tolist()   # rank g holds rows [g R, (g+1) R)
        assert len(top) == 1
