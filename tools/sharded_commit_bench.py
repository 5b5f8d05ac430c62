"""Strong scaling of ONE commitment across the GPUs of a node (openvm_b200/sharded.py):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/sharded_commit_bench.py [log_height] [width] [log_blowup] [steps]
Every rank fills its column shard on the device, then times `steps` sharded commits (CUDA events on the stream the work
is queued on, max over ranks) and the phases of one more. Prints one JSON line on rank 0."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from openvm_b200 import B2Engine, FriParameters, sharded

log_h = int(sys.argv[1]) if len(sys.argv) > 1 else 24
width = int(sys.argv[2]) if len(sys.argv) > 2 else 256
b = int(sys.argv[3]) if len(sys.argv) > 3 else 1
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local_rank = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
eng = B2Engine.new(FriParameters(log_blowup=b), local_rank)
dev = eng.device()
lo, hi = sharded.column_shards(width, world)[rank]
local = dev.alloc_matrix(1 << log_h, hi - lo)
dev.fill_synthetic(local, 0xB200000000000001 + rank)
stream = torch.cuda.Stream()
roots = []
with torch.cuda.stream(stream):
    for it in range(2):  # warm-up: twiddle tables, NCCL channels, allocator
        root, spcs = sharded.sharded_commit(eng, local, width, b)
        del spcs
    torch.cuda.synchronize()
    dist.barrier()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream)
    for it in range(steps):
        root, spcs = sharded.sharded_commit(eng, local, width, b)
        roots.append([int(x) for x in root])
        del spcs
    t1.record(stream)
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / steps
    # the fused form: last NTT pass stores into the peers' buffers, no send buffer, no collective
    committer = sharded.ShardedCommitter(eng, 1 << log_h, width, b)
    for it in range(2):
        froot, fp = committer.commit(local)
        del fp
    torch.cuda.synchronize()
    dist.barrier()
    import time
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.perf_counter()
    f0.record(stream)
    for it in range(steps):
        froot, fp = committer.commit(local)
        del fp
    f1.record(stream)
    torch.cuda.synchronize()
    fused_ms = f0.elapsed_time(f1) / steps
    fused_wall_ms = (time.perf_counter() - w0) * 1e3 / steps
    fused_same = [int(x) for x in froot] == roots[0]
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier()
    g0.record(stream)
    eng._check(eng._lib.b2s_coset_lde_scatter(eng._h, local._h, b, 1, committer._peer_arr, world, width, lo))
    g1.record(stream)
    torch.cuda.synchronize()
    fused_lde_ms = g0.elapsed_time(g1)
    dist.barrier()
    # phases of one more commit
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    rows = (1 << log_h) << b
    ev[0].record(stream)
    lde = dev.coset_lde(local, b)
    ev[1].record(stream)
    lde_t = torch.as_tensor(sharded._DevBuf(lde.device_ptr, (hi - lo) * rows), device=torch.device("cuda", local_rank))
    recv = sharded.reshard_rows(lde_t, width)
    ev[2].record(stream)
    block = dev.adopt_device_matrix(recv.data_ptr(), rows // world, width)
    sub, pcs = dev.mmcs_commit([block])
    ev[3].record(stream)
    torch.cuda.synchronize()
    phases = [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
t = torch.tensor([ms] + phases + [fused_ms, fused_wall_ms, fused_lde_ms], dtype=torch.float64, device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    ms, lde_ms, a2a_ms, hash_ms, fused_ms, fused_wall_ms, fused_lde_ms = [float(x) for x in t.cpu()]
    bytes_alg = (1 + (1 << b)) * (1 << log_h) * width * 4 + 2 * rows * 32
    print(json.dumps({"what": "one commitment sharded across GPUs (strong scaling)", "n_gpus": world, "log_height": log_h,
                      "width": width, "log_blowup": b,
                      "fused_peer_stores": {"ms_per_commit": fused_ms, "wall_ms_per_commit": fused_wall_ms,
                                            "GBps_algorithmic": bytes_alg / fused_ms / 1e6,
                                            "lde_with_scatter_ms": fused_lde_ms, "root_equals_nccl_path": fused_same},
                      "nccl_all_to_all": {"ms_per_commit": ms, "GBps_algorithmic": bytes_alg / ms / 1e6},
                      "phase_ms_max_over_ranks": {"lde_column_shard": lde_ms, "regroup_plus_all_to_all": a2a_ms,
                                                  "leaf_hash_plus_subtree": hash_ms},
                      "all_to_all_bytes_sent_per_rank": (hi - lo) * rows * 4 * (world - 1) // world,
                      "roots_equal_across_steps": all(r == roots[0] for r in roots), "root": roots[0]}))
committer.close()
eng.close()
dist.destroy_process_group()
