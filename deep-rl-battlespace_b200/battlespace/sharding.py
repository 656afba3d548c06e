"""Multi-GPU plumbing: envs shard by index with no collective on the step path; one small
all-reduce per report interval sums the episode statistics (main.py:231-253)."""
from __future__ import annotations

import numpy as np
import torch


def shard_range(n_envs_total: int, rank: int, world: int):
    """Contiguous env-index range of ``rank``: (offset, count).  Philox streams are keyed by the
    GLOBAL env index, so env g evolves identically for any world size."""
    if not (0 <= rank < world) or n_envs_total < world:
        raise ValueError("need 0 <= rank < world <= n_envs_total")
    base, rem = divmod(n_envs_total, world)
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def allreduce_stats(local, group=None):
    """Sum the int64[8] counter vector over all ranks.  ``local`` is a CUDA tensor (NCCL over
    NVLink, no host hop), a CPU tensor or a numpy array (gloo).  Returns the same kind."""
    import torch.distributed as dist
    is_np = isinstance(local, np.ndarray)
    t = torch.from_numpy(local.copy()) if is_np else local.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.numpy() if is_np else t


def broadcast_actors(actors, src=0, group=None):
    """Config 5: after the learner rank's update, every rank's ``TeamActors`` takes
    rank ``src``'s parameters -- ONE broadcast of the ~5.3k floats per plane packed into a flat
    buffer (NCCL over NVLink for CUDA tensors, gloo for CPU ones).  The device pointers the kernels
    hold stay valid: the parameters are overwritten in place."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return actors
    flat = torch.cat([t.reshape(-1) for row in actors.tensors for t in row])
    dist.broadcast(flat, src=src, group=group)
    off = 0
    for row in actors.tensors:
        for t in row:
            t.copy_(flat[off:off + t.numel()].view_as(t))
            off += t.numel()
    return actors


def summarize(stats):
    """Report-interval numbers of main.py:231-253 from the 8 counters."""
    s = [int(x) for x in stats]
    games = max(s[3], 1)
    return {"red_wins": s[0], "blue_wins": s[1], "ties": s[2], "games": s[3],
            "red_winrate": s[0] / games, "blue_winrate": s[1] / games, "tie_rate": s[2] / games,
            "red_avg_score": s[4] / 1048576.0 / games, "blue_avg_score": s[5] / 1048576.0 / games,
            "env_ticks": s[6], "agent_steps": s[7]}
