"""Command-line flags, cfg overrides, seeding and checkpoint lookup: the part of legged_gym/utils/helpers.py the hot
path's callers use (get_args helpers.py:152-178, update_cfg_from_args :127-150, set_seed :67-77, get_load_path
:103-125).  Isaac Gym's ``gymutil.parse_arguments`` is replaced by argparse with the same flag names."""
import argparse
import os
import random
import re

import numpy as np
import torch

FLAGS = [
    ("--task", dict(type=str, default="pointfoot_rough")),
    ("--resume", dict(action="store_true", default=False)),
    ("--experiment_name", dict(type=str)),
    ("--run_name", dict(type=str)),
    ("--load_run", dict(type=str)),
    ("--checkpoint", dict(type=int)),
    ("--headless", dict(action="store_true", default=False)),
    ("--horovod", dict(action="store_true", default=False)),
    ("--rl_device", dict(type=str, default="cuda:0")),
    ("--num_envs", dict(type=int)),
    ("--seed", dict(type=int)),
    ("--max_iterations", dict(type=int)),
    ("--sim_device", dict(type=str, default="cuda:0")),      # Isaac Gym's own flags, kept for command-line compatibility
    ("--pipeline", dict(type=str, default="gpu")),
    ("--physics_engine", dict(type=str, default="physx")),
    ("--num_threads", dict(type=int, default=0)),
    ("--subscenes", dict(type=int, default=0)),
]


def get_args(argv=None):
    p = argparse.ArgumentParser(description="RL Policy")
    for name, kw in FLAGS:
        p.add_argument(name, **kw)
    args = p.parse_args(argv)
    args.sim_device_id = int(args.sim_device.split(":")[1]) if ":" in args.sim_device else 0
    args.sim_device_type = args.sim_device.split(":")[0]
    return args


def update_cfg_from_args(env_cfg, cfg_train, args):
    if env_cfg is not None and getattr(args, "num_envs", None) is not None:
        env_cfg.env.num_envs = args.num_envs
    if cfg_train is not None:
        if getattr(args, "seed", None) is not None:
            cfg_train.seed = args.seed
        r = cfg_train.runner
        for flag in ("max_iterations", "experiment_name", "run_name", "load_run", "checkpoint"):
            v = getattr(args, flag, None)
            if v is not None:
                setattr(r, flag, v)
        if getattr(args, "resume", False):
            r.resume = True
    return env_cfg, cfg_train


def set_seed(seed):
    """helpers.py:67-77 (which leaves torch's CPU generator unseeded; the kernels draw from the counter-based
    contract keyed by cfg seed instead, so that is harmless here)."""
    if seed == -1:
        seed = np.random.randint(0, 10000)
    print("Setting seed: {}".format(seed))
    random.seed(seed)
    np.random.seed(seed)
    os.environ["PYTHONHASHSEED"] = str(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed(seed)
        torch.cuda.manual_seed_all(seed)
    return seed


def get_load_path(root, load_run=-1, checkpoint=-1):
    """Latest run directory (lexicographic) and highest-numbered model file unless named explicitly."""
    try:
        runs = sorted(r for r in os.listdir(root) if r != "exported")
        newest = runs[-1]
    except (OSError, IndexError):
        raise ValueError("No runs in this directory: " + root)
    run_dir = os.path.join(root, newest if load_run == -1 else load_run)
    if checkpoint == -1:
        models = [f for f in os.listdir(run_dir) if "model" in f]
        models.sort(key=lambda m: int(re.sub(r"\D", "", m) or 0))
        name = models[-1]
    else:
        name = "model_{}.pt".format(checkpoint)
    return os.path.join(run_dir, name)


def parse_sim_params(args, cfg):
    """The reference turns this dict into a gymapi.SimParams (helpers.py:79-101); without Isaac Gym the dict itself is
    what the env constructor receives."""
    out = dict(cfg.get("sim", {}))
    out["use_gpu_pipeline"] = getattr(args, "pipeline", "gpu") == "gpu"
    if getattr(args, "num_threads", 0):
        out.setdefault("physx", {})["num_threads"] = args.num_threads
    return out
