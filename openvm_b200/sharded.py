"""One commitment across the GPUs of a node (SURVEY.md 8e "if a single proof must span GPUs", 8f-4).

Segments are independent proofs and normally go one per GPU (scheduler.py). This module is for the other case: ONE trace
too tall to wait for on a single GPU. The reference has no such path (TraceCommitter::commit runs on one device,
crates/vm/src/arch/vm.rs:735); the result here is bit-identical to `B2Device.commit` of the whole trace on one GPU.

Plan for G ranks, trace N x W, blowup 2^b (R = 2^b N / G rows per rank afterwards):
  1. every rank holds a COLUMN shard of the trace (N x W_g): the NTT is per column, so the coset LDE needs no exchange
     (b2s_coset_lde on the shard);
  2. the leaf sponge runs along a row, so the LDE is re-sharded by ROWS: rank g receives rows [g R, (g+1) R) of every
     column -- one all-to-all over NVLink / NVSwitch (torch.distributed all_to_all_single on the NCCL backend; the send
     buffer is the column shard regrouped by destination, a single strided copy);
  3. rows are stored bit-reversed, so a contiguous block of R = 2^k rows is exactly one sub-tree of the Merkle tree:
     rank g hashes its rows and builds that sub-tree (b2s_mmcs_commit on the adopted row shard);
  4. the G sub-roots are all-gathered (32 bytes each) and the top log2(G) levels are compressed on every rank.
Openings: the rank owning the row answers with the row and its sub-tree path; the top-level siblings come from the
gathered sub-roots (`ShardedPcsData.open_batch`).

All device work is queued on torch's current stream (the engine is switched to it), so the collective and the kernels
are ordered without host synchronisation.

`ShardedCommitter` is the fused form of steps 1-2 (the one to use): the rows never pass through a send buffer or a
collective. Each rank allocates its row-block buffer once, the ranks exchange CUDA IPC handles and map each other's
buffers, and the LAST NTT PASS of the LDE stores every output row straight into the buffer of the rank that will hash
it, over NVLink (b2s_coset_lde_scatter: ntt_contig_kernel with a peer table). What is left of the exchange is two barriers.
`sharded_commit` (NCCL all-to-all) is kept as the baseline it is measured against.
"""
import ctypes as C

import numpy as np

from .engine import B2Engine, DeviceMatrix  # noqa: F401  (types for callers)


def column_shards(width, world_size):
    """Column ranges [lo, hi) per rank: contiguous, sizes differing by at most one (the order of columns in a row is
    the order of absorption into the leaf sponge, so shards must be contiguous and in rank order)."""
    if world_size < 1 or width < world_size:
        raise ValueError("need at least one column per rank")
    base, extra = divmod(width, world_size)
    out, lo = [], 0
    for g in range(world_size):
        hi = lo + base + (1 if g < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def _compress_of(engine):
    """The 2-to-1 compression of the engine's configuration: TruncatedPermutation over Poseidon2-BabyBear-16 (digests of
    8 words), or over Poseidon2-BN254-3 on [left, right, 0] for the root configuration (digests of one scalar, 8 limbs)."""
    dev = engine.device()
    if engine.hash() != 0:
        zero = np.zeros(8, dtype=np.uint32)
        return lambda left, right: dev.poseidon2_bn254_permute(np.concatenate([left, right, zero]))[0, :8].copy()
    return lambda left, right: dev.poseidon2_permute(np.concatenate([left, right]))[0, :8].copy()


def combine_subroots(subroots, compress):
    """Top of the tree over G = 2^k sub-roots (in rank order). `compress(left8, right8) -> 8 words` is the 2-to-1
    Poseidon2 compression. Returns (root, levels) with levels[0] = the sub-roots, levels[-1] = [root]."""
    g = len(subroots)
    if g & (g - 1):
        raise ValueError("world size must be a power of two")
    levels = [[np.asarray(r, dtype=np.uint32) for r in subroots]]
    while len(levels[-1]) > 1:
        cur = levels[-1]
        levels.append([compress(cur[2 * i], cur[2 * i + 1]) for i in range(len(cur) // 2)])
    return levels[-1][0], levels


def top_path(levels, rank):
    """Sibling digests above sub-tree `rank`, bottom-up."""
    path, i = [], rank
    for lvl in levels[:-1]:
        path.append(lvl[i ^ 1])
        i >>= 1
    return path


def reshard_rows(lde_t, width, group=None):
    """Column shard -> row shard. `lde_t`: this rank's LDE columns as one flat int32 tensor ([column][row], rows =
    full LDE height), on the device of the process group's backend. Returns the (width, rows / G) tensor holding rows
    [g R, (g+1) R) of ALL columns ([column][row], i.e. a column-major R x width matrix). The send buffer is the shard
    regrouped [column][dest][row] -> [dest][column][row] (one strided copy); ragged column shards go through the
    split sizes of all_to_all_single."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    shards = column_shards(width, world)
    w_local = shards[rank][1] - shards[rank][0]
    rows = lde_t.numel() // w_local
    r = rows // world
    send = lde_t.view(w_local, world, r).permute(1, 0, 2).contiguous().view(world * w_local, r)
    recv = torch.empty((width, r), dtype=lde_t.dtype, device=lde_t.device)
    dist.all_to_all_single(recv, send, output_split_sizes=[hi - lo for lo, hi in shards],
                           input_split_sizes=[w_local] * world, group=group)
    return recv


class _DevBuf:
    """A device pointer as a CUDA-array-interface object (torch.as_tensor shares the memory)."""

    def __init__(self, ptr, nwords):
        self.__cuda_array_interface__ = {"shape": (int(nwords),), "typestr": "<i4", "data": (int(ptr), False), "version": 2}


class ShardedPcsData:
    def __init__(self, rank, world, log_height, width, local_pcs, levels, keep):
        self.rank, self.world = rank, world
        self.log_height, self.width = log_height, width   # of the whole LDE
        self.local = local_pcs                             # PcsData of this rank's row block (a sub-tree)
        self.levels = levels
        self._keep = keep

    def owner(self, index):
        return index >> (self.log_height - (self.world.bit_length() - 1))

    def open_batch(self, index):
        """On the owning rank: (row, path) exactly as PcsData.open_batch of the single-GPU commitment; None elsewhere."""
        if self.owner(index) != self.rank:
            return None
        rows_per_rank = 1 << (self.log_height - (self.world.bit_length() - 1))
        row, path = self.local.open_batch(index & (rows_per_rank - 1))
        top = top_path(self.levels, self.rank)
        if top:
            path = np.concatenate([np.asarray(path, dtype=np.uint32).reshape(-1, 8), np.stack(top)])
        return row, path


def sharded_commit(engine, local_columns, width, log_blowup, group=None):
    """Commit the N x `width` trace whose columns [lo_g, hi_g) = column_shards(width, G)[g] live on rank g as the
    DeviceMatrix `local_columns` (N x W_g, column-major). Returns (root as 8 canonical words, ShardedPcsData); every rank
    gets the same root, equal to B2Device.commit([whole trace]) on one GPU."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = engine.device()
    cuda_dev = torch.device("cuda", torch.cuda.current_device())
    cur = torch.cuda.current_stream().cuda_stream
    if engine.stream != cur:
        engine.set_stream(cur)
    shards = column_shards(width, world)
    w_local = shards[rank][1] - shards[rank][0]
    if local_columns.width != w_local:
        raise ValueError(f"rank {rank} must hold {w_local} columns, has {local_columns.width}")
    n = local_columns.height
    rows = n << log_blowup
    if rows % world or (world & (world - 1)):
        raise ValueError("world size must be a power of two dividing the LDE height")
    r = rows // world

    # 1. coset LDE of the column shard
    lde = dev.coset_lde(local_columns, log_blowup)
    lde_t = torch.as_tensor(_DevBuf(lde.device_ptr, w_local * rows), device=cuda_dev)
    # 2. regroup by destination and exchange
    recv = reshard_rows(lde_t, width, group)
    lde.free()
    del lde_t
    # 3. this rank's sub-tree over its R rows of all columns
    block = dev.adopt_device_matrix(recv.data_ptr(), r, width)
    subroot, local_pcs = dev.mmcs_commit([block])
    # 4. top levels from the gathered sub-roots
    mine = torch.from_numpy(subroot.astype(np.int64)).to(cuda_dev)
    allroots = torch.empty((world, 8), dtype=torch.int64, device=cuda_dev)
    dist.all_gather_into_tensor(allroots, mine, group=group)
    subroots = allroots.cpu().numpy().astype(np.uint32)

    root, levels = combine_subroots(list(subroots), _compress_of(engine))
    log_height = rows.bit_length() - 1
    return root, ShardedPcsData(rank, world, log_height, width, local_pcs, levels, keep=(recv, block))


class ShardedCommitter:
    """Fused sharded commitment for traces of one shape (height x width, log_blowup), reusable across commits.

    The row-block buffers are allocated once (cudaMalloc, exported with CUDA IPC, mapped by every peer). `commit`
    returns a ShardedPcsData whose sub-tree BORROWS this rank's buffer: it is valid until the next `commit` on the same
    committer (use two committers to keep two commitments alive)."""

    def __init__(self, engine, height, width, log_blowup, group=None):
        import torch.distributed as dist
        self.engine, self.group = engine, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.height, self.width, self.b = height, width, log_blowup
        self.rows = height << log_blowup
        if self.world & (self.world - 1) or self.world > 16 or self.rows // self.world < 4:
            raise ValueError("world size must be a power of two (<= 16) with at least 4 LDE rows per rank")
        self.r = self.rows // self.world
        self.shards = column_shards(width, self.world)
        e = engine
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        e._check(e._lib.b2s_shared_alloc(e._h, self.r * width * 4, C.byref(ptr), handle))
        self._own = ptr.value
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=group)
        self._peers = []
        for g, h in enumerate(handles):
            if g == self.rank:
                self._peers.append(self._own)
            else:
                q = C.c_void_p()
                e._check(e._lib.b2s_shared_open(e._h, h, C.byref(q)))
                self._peers.append(q.value)
        self._peer_arr = (C.c_void_p * self.world)(*self._peers)
        self._block = engine.device().adopt_device_matrix(self._own, self.r, width)

    def commit(self, local_columns):
        """local_columns: this rank's DeviceMatrix of columns column_shards(width, G)[rank] (height x W_g)."""
        import torch
        import torch.distributed as dist
        e, dev = self.engine, self.engine.device()
        lo, hi = self.shards[self.rank]
        if local_columns.width != hi - lo or local_columns.height != self.height:
            raise ValueError(f"rank {self.rank} must hold a {self.height} x {hi - lo} column shard")
        cur = torch.cuda.current_stream().cuda_stream
        if e.stream != cur:
            e.set_stream(cur)
        # nobody may still be hashing the previous contents of a buffer we are about to overwrite
        torch.cuda.current_stream().synchronize()
        dist.barrier(group=self.group)
        # 1 + 2. coset LDE whose last pass writes each row into its owner's buffer over NVLink
        e._check(e._lib.b2s_coset_lde_scatter(e._h, local_columns._h, self.b, 1, self._peer_arr, self.world, self.width, lo))
        torch.cuda.current_stream().synchronize()
        dist.barrier(group=self.group)          # every peer's rows have landed
        # 3. this rank's sub-tree
        subroot, local_pcs = dev.mmcs_commit([self._block])
        # 4. top levels
        cuda_dev = torch.device("cuda", torch.cuda.current_device())
        mine = torch.from_numpy(subroot.astype(np.int64)).to(cuda_dev)
        allroots = torch.empty((self.world, 8), dtype=torch.int64, device=cuda_dev)
        dist.all_gather_into_tensor(allroots, mine, group=self.group)
        subroots = allroots.cpu().numpy().astype(np.uint32)

        root, levels = combine_subroots(list(subroots), _compress_of(self.engine))
        return root, ShardedPcsData(self.rank, self.world, self.rows.bit_length() - 1, self.width, local_pcs, levels,
                                    keep=(self,))

    def close(self):
        import torch
        import torch.distributed as dist
        e = self.engine
        if self._own is None:
            return
        torch.cuda.synchronize()
        dist.barrier(group=self.group)          # no peer is still writing into a buffer that is about to go away
        for g, q in enumerate(self._peers):
            if g != self.rank:
                e._check(e._lib.b2s_shared_close(e._h, q))
        dist.barrier(group=self.group)
        self._block.free()
        e._check(e._lib.b2s_shared_free(e._h, self._own))
        self._own = None
