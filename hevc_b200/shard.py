"""Sharding of independent frames across GPUs: "replicas only" (DESIGN.md section 8).

Frames (or independent streams) are the unit: frame i belongs to rank i mod world, every
rank owns its context, DPB and reference pictures, and nothing is exchanged on the data
path.  torch.distributed only carries the start/stop barrier and the max-over-ranks
reduction of the timing.
"""


def frames_of_rank(n_frames, rank, world):
    """Indices of the frames rank `rank` reconstructs."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, n_frames, world))


def seed_of_frame(base, frame):
    """Synthetic frame `frame` has the same content whichever rank builds it."""
    return base + 17 * frame


def max_over_ranks(values, device=None):
    """Element-wise maximum of a list of floats over all ranks (identity without a group)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()
