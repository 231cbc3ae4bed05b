"""How the path shards over GPUs: environments are independent, so rank r owns a contiguous range of GLOBAL env
ids and every kernel is handed ``env_id_offset`` (the RNG contract of csrc/pf_rng.cuh keys on the global id, so the
union of the shards is the same simulation as one big launch).  The only exchanges are the PPO reductions
below and the broadcast of rank 0's initial parameters (BASELINE.json north_star: "NCCL ... only for the PPO gradient allreduce and the reward/episode statistics");
the env kernels need no collective.  Pure host logic: works on any torch.distributed backend (nccl on the GPUs,
gloo in tests/test_shard_gloo.py)."""
import torch


def dist_or_none():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 else None


def shard_range(total_envs, rank, world_size):
    """(env_id_offset, num_envs) of ``rank``: contiguous, sizes differ by at most one, earlier ranks get the extra."""
    if not (0 <= rank < world_size) or total_envs < 0:
        raise ValueError(f"bad shard request rank={rank} world_size={world_size} total_envs={total_envs}")
    base, extra = divmod(total_envs, world_size)
    return rank * base + min(rank, extra), base + (1 if rank < extra else 0)


def allreduce_sum_(t):
    d = dist_or_none()
    if d:
        d.all_reduce(t)
    return t


def allreduce_mean_(t):
    """Gradient averaging of PPO.update: the minibatch loss is a mean over the rank's samples and shards are equal
    sized, so the global-minibatch gradient is the mean of the rank gradients."""
    d = dist_or_none()
    if d:
        d.all_reduce(t)
        t.div_(d.get_world_size())
    return t


def advantage_moments(stats):
    """(mean, unbiased std) from the all-reduced float64 triple [sum, sum of squares, count] that pf_rl_gae emits;
    the same arithmetic pf_norm_adv_kernel does on the device (rollout_storage.py compute_returns)."""
    s, ss, n = (float(x) for x in stats)
    mean = s / n
    var = max((ss - n * mean * mean) / (n - 1.0), 0.0)
    return mean, var ** 0.5


def merge_episode_stats(stats, counts):
    """extras['episode'] means over several shards: per-shard means [W, R] weighted by per-shard reset counts [W]."""
    counts = counts.to(stats.dtype)
    tot = counts.sum().clamp_min(1.0)
    return (stats * counts[:, None]).sum(0) / tot
