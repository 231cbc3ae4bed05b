"""legged_gym/scripts/train.py:41-49 over the B200 path: python scripts/train.py --task pointfoot_flat --num_envs 4096"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointfoot_b200.helpers import get_args              # noqa: E402
from pointfoot_b200.task_registry import task_registry   # noqa: E402


def train(args):
    env, env_cfg = task_registry.make_env(name=args.task, args=args)
    ppo_runner, train_cfg = task_registry.make_alg_runner(env=env, name=args.task, args=args)
    ppo_runner.learn(num_learning_iterations=train_cfg.runner.max_iterations, init_at_random_ep_len=True)


if __name__ == "__main__":
    train(get_args())
