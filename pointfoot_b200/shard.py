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
    sized, so the global-minibatch gradient is the mean of the rank gradients.  ONE collective (NCCL's AVG) where the
    backend has it; gloo (the CPU tests) sums and divides."""
    d = dist_or_none()
    if d:
        if d.get_backend() == "nccl":
            d.all_reduce(t, op=d.ReduceOp.AVG)
        else:
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


def allreduce_episode_infos(ep_infos, num_envs):
    """``extras["episode"]`` of the steps of one rollout, merged over all shards (BASELINE.json north_star: NCCL "for the
    PPO gradient allreduce and the reward/episode statistics").  ``ep_infos``: one dict per step, the env's 0-d tensors
    plus ``"_count"`` = the number of envs of THIS shard that reset on that step.  A ``rew_*`` entry is the mean over the
    shard's reset envs (point_foot.py:276-280), so the job-wide value of a step is the count-weighted mean over shards;
    ``terrain_level`` is a mean over all envs of the shard (point_foot.py:283): weighted by ``num_envs``.  Steps on which
    no env of any shard reset carry no new information and are left out; the result is the plain mean over the other
    steps, like the runner's mean over ``ep_infos`` upstream.  One all-reduce of (S x K) x 2 doubles.  Collective: every
    rank calls it, with the same number of steps.  Returns {key: float}; non-tensor entries are passed through."""
    import torch
    if not ep_infos:
        return {}
    keys = [k for k, v in ep_infos[0].items() if k != "_count" and torch.is_tensor(v)]
    out = {k: v for k, v in ep_infos[-1].items() if k != "_count" and not torch.is_tensor(v)}
    if not keys:
        return out
    dev = ep_infos[0][keys[0]].device
    S, K = len(ep_infos), len(keys)
    vals = torch.stack([torch.stack([ep[k].reshape(()).to(torch.float64) for k in keys]) for ep in ep_infos])
    cnt = torch.stack([torch.as_tensor(ep["_count"], device=dev).reshape(()).to(torch.float64) for ep in ep_infos])
    w = cnt[:, None].expand(S, K).clone()
    for j, k in enumerate(keys):
        if k == "terrain_level":
            w[:, j] = float(num_envs)
    buf = torch.cat(((vals * w).flatten(), w.flatten()))
    allreduce_sum_(buf)
    num, den = buf[:S * K].view(S, K), buf[S * K:].view(S, K)
    valid = (den > 0).to(torch.float64)
    mean = (num / den.clamp_min(1.0) * valid).sum(0) / valid.sum(0).clamp_min(1.0)
    mean = mean.cpu().tolist()
    out.update({k: mean[j] for j, k in enumerate(keys)})
    return out


def shutdown(grace_s=20.0):
    """``destroy_process_group()`` that cannot hang the job: callers close their PPO objects first (PPO.close drops the
    CUDA graph that holds captured NCCL kernels); if the communicator's teardown still has not returned after ``grace_s``
    seconds, the process exits with status 0 -- everything that had to be printed or saved was, before this call."""
    import gc
    import os
    import sys
    import threading
    d = dist_or_none()
    if d is None:
        return
    gc.collect()
    torch.cuda.synchronize()
    sys.stdout.flush(); sys.stderr.flush()

    def bail():
        sys.stderr.write("pointfoot_b200.shard.shutdown: destroy_process_group() did not return; exiting\n")
        sys.stderr.flush()
        os._exit(0)
    t = threading.Timer(grace_s, bail)
    t.daemon = True
    t.start()
    try:
        d.barrier()
        d.destroy_process_group()
    finally:
        t.cancel()
