"""Host-side partitioning of independent proving jobs across GPUs.

Continuation segments and aggregation-tree nodes are independent STARK proofs (the reference proves them in a
sequential loop on one engine, crates/vm/src/arch/vm.rs:979-1007, and documents the distributed model in
docs/vocs/docs/pages/specs/architecture/distributed-proving.mdx). One process per GPU, one engine per process; the
only cross-rank traffic is a barrier, the max-over-ranks time and the gather of proof sizes -- no data-path collective.
"""
import torch.distributed as dist


def partition(n_jobs, world_size, rank):
    """Job indices of `rank`: round-robin, so layer k of an arity-3 aggregation tree (ceil(S / 3^k) nodes) stays balanced."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    return list(range(rank, n_jobs, world_size))


def aggregation_layers(n_segments, num_children_leaf=1, num_children_internal=3):
    """Number of independent jobs per layer: leaves (1:1 with segments by default, crates/sdk/src/config/mod.rs:26), then
    internal layers of arity 3 (:27) until one proof is left."""
    layers = [-(-n_segments // num_children_leaf)]
    while layers[-1] > 1:
        layers.append(-(-layers[-1] // num_children_internal))
    return layers


def max_over_ranks(value, device=None):
    """Max of a python float over all ranks (time-like quantities are reported as the slowest rank's)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_counts(value):
    """All ranks' integer `value` (e.g. segments proven), on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [int(value)]
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, int(value))
    return out
