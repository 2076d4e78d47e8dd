"""Sharding of a Monte-Carlo batch over the GPUs of one box: filters are independent, so rank r owns a
contiguous block of global instance ids and runs it with `BatchVIEKF.set_instance_offset(first)`; the per-instance noise is
keyed by the GLOBAL id (include/viekf_noise.h), which makes the sharded run reproduce the unsharded one bit for bit.
The only collective of the path is the sum of the per-rank statistics blocks (NCCL on device pointers, gloo in the CPU tests)."""
import numpy as np


def shard_range(total, rank, world):
    """Contiguous block [first, first + count) of `total` instances owned by `rank`; earlier ranks take the remainder."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    base, rem = divmod(int(total), int(world))
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def owner_of(instance, total, world):
    """Rank that owns global instance id `instance` under shard_range."""
    base, rem = divmod(int(total), int(world))
    cut = rem * (base + 1)
    return instance // (base + 1) if instance < cut else rem + (instance - cut) // max(base, 1)


def allreduce_stats(stats, group=None):
    """Sum of the [buckets][STAT_WIDTH] statistics blocks over the ranks (torch.distributed, any backend).  `stats` is a
    numpy array (host copy: gloo / CPU tests) or a torch tensor (device-resident: NCCL); returns the same kind."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return stats
    if isinstance(stats, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(stats, dtype=np.float64).copy())
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return t.numpy()
    dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def rmse_nees(stats):
    """Per-bucket RMSE of the 16 base error states and mean NEES (16-dof, 6-dof pose) from summed statistics blocks."""
    stats = np.asarray(stats)
    n = np.maximum(stats[:, 0:1], 1.0)
    return {"count": stats[:, 0], "rmse": np.sqrt(stats[:, 17:33] / n), "mean_err": stats[:, 1:17] / n,
            "nees16": stats[:, 33] / n[:, 0], "nees6": stats[:, 34] / n[:, 0]}
