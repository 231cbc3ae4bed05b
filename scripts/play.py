"""legged_gym/scripts/play.py:42-114 over the B200 path (headless: no viewer, no frame recording; Logger = pointfoot_b200/logger.py):
load (or randomly initialise) a policy, run it closed loop, export the actor as TorchScript, print episode rewards.

    python scripts/play.py --task pointfoot_flat --num_envs 50 [--load_run RUN --checkpoint N] [--steps 200]
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pointfoot_b200.helpers import get_args, export_policy_as_jit   # noqa: E402
from pointfoot_b200.logger import Logger                           # noqa: E402
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
    logger = Logger(env.dt)
    robot_index, joint_index = 0, 1                                # play.py:64-65: which robot / joint is recorded
    stop_state_log = min(100, n)                                   # play.py:66: states of the first steps, rewards of all of them
    for i in range(n):
        actions = policy(obs.detach())
        obs, _, rews, dones, infos = env.step(actions.detach())
        if i < stop_state_log:                                     # play.py:88-102
            logger.log_states({
                'dof_pos_target': actions[robot_index, joint_index].item() * env.cfg.control.action_scale,
                'dof_pos': env.dof_pos[robot_index, joint_index].item(),
                'dof_vel': env.dof_vel[robot_index, joint_index].item(),
                'dof_torque': env.torques[robot_index, joint_index].item(),
                'command_x': env.commands[robot_index, 0].item(),
                'command_y': env.commands[robot_index, 1].item(),
                'command_yaw': env.commands[robot_index, 2].item(),
                'base_vel_x': env.base_lin_vel[robot_index, 0].item(),
                'base_vel_y': env.base_lin_vel[robot_index, 1].item(),
                'base_vel_z': env.base_lin_vel[robot_index, 2].item(),
                'base_vel_yaw': env.base_ang_vel[robot_index, 2].item(),
                'contact_forces_z': env.contact_forces[robot_index, env.feet_indices, 2].cpu().numpy()})
        elif i == stop_state_log and os.environ.get("PF_PLAY_PLOT"):
            print("Logged states written to:", logger.plot_states(os.environ["PF_PLAY_PLOT"]))
        if infos.get("episode"):                                   # play.py:103-107
            k = int(torch.sum(env.reset_buf).item())
            if k > 0:
                logger.log_rewards(infos["episode"], k)
    if logger.num_episodes:
        logger.print_rewards()
    play.logger = logger
    return exported


if __name__ == "__main__":
    a = get_args()
    play(a, steps=int(os.environ.get("PF_PLAY_STEPS", "0")) or None)
