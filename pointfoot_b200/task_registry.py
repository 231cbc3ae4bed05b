"""``task_registry`` with the reference's API (legged_gym/utils/task_registry.py:44-155): register / get_task_class /
get_cfgs / make_env / make_alg_runner, and the two PointFoot tasks registered under the reference's names
(legged_gym/envs/__init__.py:51-52)."""
import os
from datetime import datetime

from . import config, helpers
from .env import PointFoot
from .rl import OnPolicyRunner

ROOT_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class TaskRegistry:
    def __init__(self):
        self.task_classes, self.env_cfgs, self.train_cfgs = {}, {}, {}

    def register(self, name, task_class, env_cfg, train_cfg):
        self.task_classes[name], self.env_cfgs[name], self.train_cfgs[name] = task_class, env_cfg, train_cfg

    def get_task_class(self, name):
        return self.task_classes[name]

    def get_cfgs(self, name):
        env_cfg, train_cfg = self.env_cfgs[name], self.train_cfgs[name]
        env_cfg.seed = train_cfg.seed               # task_registry.py:61
        return env_cfg, train_cfg

    def make_env(self, name, args=None, env_cfg=None, **env_kwargs):
        args = args if args is not None else helpers.get_args([])
        if name not in self.task_classes:
            raise ValueError(f"Task with name: {name} was not registered")
        if env_cfg is None:
            env_cfg, _ = self.get_cfgs(name)
        env_cfg, _ = helpers.update_cfg_from_args(env_cfg, None, args)
        seed = helpers.set_seed(env_cfg.seed)
        sim_params = helpers.parse_sim_params(args, {"sim": config.class_to_dict(env_cfg.sim)})
        env = self.task_classes[name](cfg=env_cfg, sim_params=sim_params, physics_engine=args.physics_engine,
                                      sim_device=args.sim_device, headless=args.headless, seed=seed, **env_kwargs)
        return env, env_cfg

    def make_alg_runner(self, env, name=None, args=None, train_cfg=None, log_root="default"):
        args = args if args is not None else helpers.get_args([])
        if train_cfg is None:
            if name is None:
                raise ValueError("Either 'name' or 'train_cfg' must be not None")
            _, train_cfg = self.get_cfgs(name)
        elif name is not None:
            print(f"'train_cfg' provided -> Ignoring 'name={name}'")
        _, train_cfg = helpers.update_cfg_from_args(None, train_cfg, args)
        stamp = datetime.now().strftime("%b%d_%H-%M-%S") + "_" + train_cfg.runner.run_name
        if log_root == "default":
            log_root = os.path.join(ROOT_DIR, "logs", train_cfg.runner.experiment_name)
        log_dir = None if log_root is None else os.path.join(log_root, stamp)
        if log_dir is not None:
            os.makedirs(log_dir, exist_ok=True)
        runner = OnPolicyRunner(env, config.class_to_dict(train_cfg), log_dir, device=args.rl_device)
        if train_cfg.runner.resume:
            path = helpers.get_load_path(log_root, load_run=train_cfg.runner.load_run, checkpoint=train_cfg.runner.checkpoint)
            print(f"Loading model from: {path}")
            runner.load(path)
        return runner, train_cfg


task_registry = TaskRegistry()
task_registry.register("pointfoot_rough", PointFoot, config.PointFootRoughCfg(), config.PointFootRoughCfgPPO())
task_registry.register("pointfoot_flat", PointFoot, config.PointFootFlatCfg(), config.PointFootFlatCfgPPO())
