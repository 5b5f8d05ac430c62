"""GPU parity of the sharded commitment (openvm_b200/sharded.py): two ranks produce the root, the rows and the opening
paths of the single-GPU b2s_commit of the same trace -- through the NCCL all-to-all baseline and through the fused form
(last NTT pass storing into the peers' buffers over NVLink). Needs two GPUs; skipped on a one-GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
P = 2013265921


def _worker(rank, world, port, q, height, width, b, fused, hash_kind=0):
    try:
        import torch.distributed as dist
        from openvm_b200 import B2Engine, FriParameters, sharded
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        eng = B2Engine.new(FriParameters(log_blowup=b), rank)
        eng.set_hash(hash_kind)
        dev = eng.device()
        trace = np.random.default_rng(9).integers(0, P, size=(height, width), dtype=np.uint64).astype(np.uint32)
        lo, hi = sharded.column_shards(width, world)[rank]
        local = dev.transport_matrix_to_device(np.ascontiguousarray(trace[:, lo:hi]))
        if fused:
            committer = sharded.ShardedCommitter(eng, height, width, b)
            root0, _ = committer.commit(local)          # a second commit on the same buffers must give the same tree
            root, spcs = committer.commit(local)
            assert [int(x) for x in root0] == [int(x) for x in root]
        else:
            root, spcs = sharded.sharded_commit(eng, local, width, b)
        torch.cuda.synchronize()
        rows = height << b
        indices = [0, 1, rows // 2 - 1, rows // 2, rows - 1, 12345 % rows]
        opened = {i: spcs.open_batch(i) for i in indices}
        out = {"rank": rank, "root": [int(x) for x in root],
               "opened": {i: (v[0][0].tolist(), np.asarray(v[1]).tolist()) for i, v in opened.items() if v is not None}}
        if rank == 0:
            want_root, pcs = dev.commit([dev.transport_matrix_to_device(trace)])
            out["want_root"] = [int(x) for x in want_root]
            out["want_opened"] = {}
            for i in indices:
                r_, p_ = pcs.open_batch(i)
                out["want_opened"][i] = (r_[0].tolist(), np.asarray(p_).tolist())
        dist.barrier()
        q.put(out)
        if fused:
            del spcs
            committer.close()
        eng.close()
        dist.destroy_process_group()
    except Exception as ex:  # surface the failure in the parent
        import traceback
        q.put({"rank": rank, "error": traceback.format_exc() + repr(ex)})


@pytest.mark.parametrize("fused", [False, True], ids=["nccl_all_to_all", "fused_peer_stores"])
@pytest.mark.parametrize("height,width,b,hash_kind", [(1 << 12, 19, 1, 0), (1 << 18, 64, 1, 0), (1 << 14, 9, 2, 0), (4, 2, 1, 0),
                                                      (1 << 12, 19, 1, 1)], ids=["h12", "h18", "h14b2", "h2", "h12_root_hash"])
def test_sharded_commit_two_gpus(height, width, b, hash_kind, fused):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, height, width, b, fused, hash_kind)) for r in range(2)]
    for p in procs: p.start()
    res = [q.get(timeout=600) for _ in range(2)]
    for p in procs: p.join(timeout=120)
    for r in res:
        assert "error" not in r, r.get("error")
    res.sort(key=lambda r: r["rank"])
    want = res[0]
    for r in res:
        assert r["root"] == want["want_root"]
    seen = 0
    for r in res:
        for i, (row, path) in r["opened"].items():
            wrow, wpath = want["want_opened"][i]
            assert row == wrow and path == wpath, (r["rank"], i)
            seen += 1
    assert seen == len(want["want_opened"])      # every index was answered by exactly one rank


@pytest.mark.parametrize("height,width,b,peers", [(1 << 10, 7, 1, 1), (1 << 12, 19, 1, 4), (1 << 15, 5, 2, 8), (1 << 14, 33, 1, 2), (4, 3, 1, 2)])
def test_lde_scatter_layout_one_gpu(engine, engine_b2, height, width, b, peers):
    """b2s_coset_lde_scatter on ONE GPU: the "peers" are buffers of this process (b2s_shared_alloc), so the row-block layout
    the last NTT pass writes -- rank g gets LDE rows [g R, (g+1) R) of every column, column-major R x total_width, columns at
    their global index -- is checked against b2s_coset_lde without a second GPU. Two column shards are scattered one after
    the other, as two ranks would."""
    import ctypes as C
    e = engine if b == 1 else engine_b2
    dev = e.device()
    rng = np.random.default_rng(21)
    trace = rng.integers(0, P, size=(height, width), dtype=np.uint64).astype(np.uint32)
    want = dev.coset_lde(dev.transport_matrix_to_device(trace), b).to_host()       # (rows, width) canonical
    rows = height << b
    r = rows // peers
    bufs = []
    for _ in range(peers):
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        e._check(e._lib.b2s_shared_alloc(e._h, r * width * 4, C.byref(ptr), handle))
        bufs.append(ptr.value)
    try:
        arr = (C.c_void_p * peers)(*bufs)
        split = max(1, width // 3)
        for lo, hi in ((0, split), (split, width)):
            if hi > lo:
                shard = dev.transport_matrix_to_device(np.ascontiguousarray(trace[:, lo:hi]))
                e._check(e._lib.b2s_coset_lde_scatter(e._h, shard._h, b, 1, arr, peers, width, lo))
        e.synchronize()
        for g in range(peers):
            block = dev.adopt_device_matrix(bufs[g], r, width).to_host()
            assert np.array_equal(block, want[g * r:(g + 1) * r]), g
    finally:
        e.synchronize()
        for pbuf in bufs:
            e._check(e._lib.b2s_shared_free(e._h, pbuf))


def test_lde_scatter_rejects_bad_arguments(engine):
    import ctypes as C
    from openvm_b200 import B2StarkError
    e = engine
    dev = e.device()
    m = dev.transport_matrix_to_device(np.ones((8, 2), dtype=np.uint32))
    ptr, handle = C.c_void_p(), C.create_string_buffer(64)
    e._check(e._lib.b2s_shared_alloc(e._h, 4096, C.byref(ptr), handle))
    try:
        one = (C.c_void_p * 3)(ptr.value, ptr.value, ptr.value)
        with pytest.raises(B2StarkError):   # three peers: not a power of two
            e._check(e._lib.b2s_coset_lde_scatter(e._h, m._h, 1, 1, one, 3, 2, 0))
        with pytest.raises(B2StarkError):   # shard outside the trace
            e._check(e._lib.b2s_coset_lde_scatter(e._h, m._h, 1, 1, one, 1, 2, 1))
        with pytest.raises(B2StarkError):   # fewer than 4 rows per peer
            e._check(e._lib.b2s_coset_lde_scatter(e._h, m._h, 1, 1, (C.c_void_p * 8)(*([ptr.value] * 8)), 8, 2, 0))
    finally:
        e._check(e._lib.b2s_shared_free(e._h, ptr.value))
