"""legged_gym/scripts/play.py:42-114 over the B200 path (headless: no viewer, no frame recording, no matplotlib Logger):
load (or randomly initialise) a policy, run it closed loop, export the actor as TorchScript, print episode rewards.

    python scripts/play.py --task pointfoot_flat --num_envs 50 [--load_run RUN --checkpoint N] [--steps 200]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pointfoot_b200.helpers import get_args, export_policy_as_jit   # noqa: E402
from pointfoot_b200.task_registry import task_registry             # noqa: E402

EXPORT_POLICY = True


def play(args, steps=None, export_dir=None):
    env_cfg, train_cfg = task_registry.get_cfgs(name=args.task)
    # play.py:44-51 overrides
    env_cfg.env.num_envs = min(env_cfg.env.num_envs, 50) if getattr(args, "num_envs", None) is None else args.num_envs
    env_cfg.terrain.num_rows = 5
    env_cfg.terrain.num_cols = 5
    env_cfg.terrain.curriculum = False
    env_cfg.noise.add_noise = False
    env_cfg.domain_rand.randomize_friction = False
    env_cfg.domain_rand.push_robots = False

    env, _ = task_registry.make_env(name=args.task, args=args, env_cfg=env_cfg)
    obs = env.get_observations()
    train_cfg.runner.resume = bool(getattr(args, "resume", False) or getattr(args, "load_run", None))
    ppo_runner, train_cfg = task_registry.make_alg_runner(env=env, name=args.task, args=args, train_cfg=train_cfg,
                                                          log_root="default" if train_cfg.runner.resume else None)
    policy = ppo_runner.get_inference_policy(device=env.device)

    exported = None
    if EXPORT_POLICY:
        path = export_dir or os.path.join(ROOT, "logs", train_cfg.runner.experiment_name, "exported", "policies")
        exported = export_policy_as_jit(ppo_runner.alg.actor_critic, path)
        print("Exported policy as jit script to: ", exported)

    n = steps if steps is not None else 10 * int(env.max_episode_length)
    rew_sums, episodes = {}, 0
    for i in range(n):
        actions = policy(obs.detach())
        obs, _, rews, dones, infos = env.step(actions.detach())
        if infos.get("episode"):                                   # play.py:103-107 (Logger.log_rewards)
            k = int(torch.sum(env.reset_buf).item())
            if k > 0:
                episodes += k
                for name, v in infos["episode"].items():
                    if "rew" in name:
                        rew_sums[name] = rew_sums.get(name, 0.0) + float(v) * k
    if episodes:                                                   # Logger.print_rewards
        print("Average rewards per second:")
        for name, v in rew_sums.items():
            print(f" - {name}: {v / episodes}")
        print(f"Total number of episodes: {episodes}")
    return exported


if __name__ == "__main__":
    a = get_args()
    play(a, steps=int(os.environ.get("PF_PLAY_STEPS", "0")) or None)
