"""N>1 path on CPU: two gloo ranks (127.0.0.1 rendezvous).  Covers what multi-GPU adds to the hot path:
  * env sharding by global env id (pointfoot_b200/shard.py + the env_id_offset RNG contract): the union of two
    shards stepped separately is bit-identical to one unsharded run (host emulation of the kernel core);
  * the three PPO reductions (gradient mean, loss statistics, advantage moments) against the oracle on the
    concatenated data.
No compute call reaches the CUDA library here."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for _p in (ROOT, HERE):                      # spawned workers do not run conftest.py
    if _p not in sys.path:
        sys.path.insert(0, _p)

from pointfoot_b200 import shard, synth   # noqa: E402

PER_ENV_ROWS = {"dof_state": synth.NUM_DOF, "contact_forces": synth.NUM_BODIES, "rigid_body_states": synth.NUM_BODIES}
SHARED = ("height_samples", "terrain_origins", "common_step_counter")


def slice_state(state, off, n, N):
    """rows [off, off+n) of every per-env tensor of a synthetic state; terrain tables are shared."""
    out = {}
    for k, v in state.items():
        if k in SHARED or not torch.is_tensor(v):
            out[k] = v
        elif k == "episode_sums":
            out[k] = v[:, off:off + n].clone()
        elif v.shape[0] == N * PER_ENV_ROWS.get(k, 1):
            m = PER_ENV_ROWS.get(k, 1)
            out[k] = v[off * m:(off + n) * m].clone()
        else:
            out[k] = v
    return out


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _init(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)


FIELDS = ("obs_buf", "priv_obs_buf", "rew_buf", "reset_buf", "time_out_buf", "torques", "episode_length_buf", "commands",
          "root_states", "dof_state", "feet_air_time", "last_actions")


def _env_worker(rank, world, port, task, N, steps, q):
    try:
        _init(rank, world, port)
        import parity
        state = synth.make_state(N, task, seed=1, adversarial=True)
        if task == "pointfoot_flat":               # one global spawn grid (PF:988-997), sliced like every per-env tensor
            from oracle.pointfoot_ref import grid_env_origins
            state["env_origins"] = grid_env_origins(N, 3.0)
        off, n = shard.shard_range(N, rank, world)
        env = parity.EnvUnderTest(task, slice_state(state, off, n, N), "cpu", backend="host", seed=1, env_id_offset=off)
        for i in range(steps):
            a = synth.make_state(N, task, seed=101 + i)["actions"][off:off + n]
            env.step(a.clone())
        mine = {k: env.t[k].reshape(n, -1).to(torch.float64) for k in FIELDS if env.t.get(k) is not None}
        ok = True
        if rank == 0:
            whole = parity.EnvUnderTest(task, state, "cpu", backend="host", seed=1)
            for i in range(steps):
                whole.step(synth.make_state(N, task, seed=101 + i)["actions"].clone())
        resets = torch.tensor([int(env.t["reset_count"].item())], dtype=torch.int64)
        dist.all_reduce(resets)
        for k, v in sorted(mine.items()):
            parts = [None] * world
            dist.all_gather_object(parts, v)            # shard sizes may differ by one
            if rank == 0:
                got = torch.cat(parts)
                want = whole.t[k].reshape(N, -1).to(torch.float64)
                if not torch.equal(got, want):
                    ok = False
                    q.put(("mismatch", k, int((got != want).sum())))
        if rank == 0:
            if int(resets.item()) != int(whole.t["reset_count"].item()) or int(resets.item()) == 0:
                ok = False
                q.put(("reset_count", int(resets.item()), int(whole.t["reset_count"].item())))
            q.put(("ok", ok))
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:      # surface the failure in the parent
        q.put(("error", rank, repr(e)))
        raise


def _run_two_ranks(worker, extra, timeout):
    """Two spawned gloo ranks on a fresh local port; the port is only known free at probe time, so ONE retry on a new port if the
    ranks did not both exit cleanly (a lost rendezvous race must not look like a parity failure)."""
    ctx = mp.get_context("spawn")
    msgs, procs = [], []
    for attempt in range(2):
        q = ctx.Queue()
        port = free_port()
        procs = [ctx.Process(target=worker, args=(r, 2, port) + tuple(extra) + (q,)) for r in range(2)]
        [p.start() for p in procs]
        [p.join(timeout) for p in procs]
        for p in procs:
            if p.is_alive():
                p.kill()
        msgs = []
        while not q.empty():
            msgs.append(q.get())
        if all(p.exitcode == 0 for p in procs) and ("ok", True) in msgs:
            break
    assert all(p.exitcode == 0 for p in procs), msgs
    assert ("ok", True) in msgs, msgs


@pytest.mark.parametrize("task,N", [("pointfoot_flat", 64), ("pointfoot_rough", 51)])
def test_two_shards_equal_one_unsharded_run(task, N):
    _run_two_ranks(_env_worker, (task, N, 3), 180)


def _ppo_worker(rank, world, port, q):
    try:
        _init(rank, world, port)
        g = torch.Generator().manual_seed(5)
        grads_all = torch.randn(world, 1000, generator=g)
        adv_all = torch.randn(world, 24 * 32, generator=g) * 3 + 1
        kl_all = torch.rand(world, 2, generator=g, dtype=torch.float64)
        grads = grads_all[rank].clone()
        shard.allreduce_mean_(grads)
        a = adv_all[rank].double()
        stats = torch.tensor([a.sum(), (a * a).sum(), float(a.numel())], dtype=torch.float64)
        shard.allreduce_sum_(stats)
        loss = kl_all[rank].clone()
        shard.allreduce_sum_(loss)
        mean, std = shard.advantage_moments(stats)
        flat = adv_all.flatten()
        ok = (torch.allclose(grads, grads_all.mean(0), atol=1e-6)
              and abs(mean - float(flat.mean())) < 1e-5 and abs(std - float(flat.std())) < 1e-5
              and torch.allclose(loss, kl_all.sum(0)))
        merged = shard.merge_episode_stats(torch.tensor([[1.0, 2.0], [3.0, 6.0]]), torch.tensor([1, 3]))
        ok = ok and torch.allclose(merged, torch.tensor([2.5, 5.0]))
        # extras["episode"] of a 3-step rollout merged over the two shards (what OnPolicyRunner.learn logs when sharded):
        # step 0: rank 0 had 1 reset (mean 2.0), rank 1 had 3 (mean 6.0) -> 5.0; step 1: nobody reset -> left out;
        # step 2: only rank 1 reset (mean 1.0; rank 0's entry is a stale 9.0 with count 0) -> 1.0.  Mean over steps: 3.0.
        # terrain_level is a mean over ALL envs of a shard: weighted by num_envs (equal shards -> plain mean over ranks).
        rew = ([2.0, 7.0, 9.0], [6.0, 6.0, 1.0])[rank]
        cnt = ([1, 0, 0], [3, 0, 1])[rank]
        lvl = ([1.0, 1.0, 2.0], [3.0, 3.0, 4.0])[rank]
        infos = [{"rew_torques": torch.tensor(rew[i]), "terrain_level": torch.tensor(lvl[i]), "max_command_x": 1.0,
                  "_count": torch.tensor(cnt[i], dtype=torch.int32)} for i in range(3)]
        m = shard.allreduce_episode_infos(infos, num_envs=8)
        ok = ok and abs(m["rew_torques"] - 3.0) < 1e-12 and abs(m["terrain_level"] - (2.0 + 2.0 + 3.0) / 3) < 1e-12
        ok = ok and m["max_command_x"] == 1.0 and "_count" not in m
        if rank == 0:
            q.put(("ok", bool(ok)))
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:
        q.put(("error", rank, repr(e)))
        raise


def test_ppo_reductions_match_oracle_on_concatenated_shards():
    _run_two_ranks(_ppo_worker, (), 120)


def test_shard_range_partitions_every_env_once():
    for total, world in [(4096, 8), (50, 3), (7, 8), (0, 2)]:
        ranges = [shard.shard_range(total, r, world) for r in range(world)]
        ids = np.concatenate([np.arange(o, o + n) for o, n in ranges]) if total else np.array([])
        assert np.array_equal(ids, np.arange(total))
        assert max(n for _, n in ranges) - min(n for _, n in ranges) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(8, 2, 2)
    assert shard.dist_or_none() is None            # single process: reductions are no-ops
    t = torch.ones(3)
    assert shard.allreduce_mean_(t) is t and torch.equal(t, torch.ones(3))
