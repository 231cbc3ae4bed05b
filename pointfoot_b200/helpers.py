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


# ---------------------------------------------------------------------------------------------------- export
def actor_module_from_state_dict(state_dict, activation="elu"):
    """A plain ``torch.nn.Sequential`` (Linear, ELU, ..., Linear) holding the ``actor.*`` entries of an rsl_rl
    ``ActorCritic`` state dict -- the module the reference deep-copies out of ``actor_critic.actor`` (helpers.py:180-190,
    export_policy_as_onnx.py:13-46).  Works for checkpoints of either stack (same keys: ``actor.{0,2,4,..}.{weight,bias}``)."""
    import torch
    if activation != "elu":
        raise NotImplementedError("only the 'elu' activation of the pointfoot configs is implemented")
    idx = sorted({int(k.split(".")[1]) for k in state_dict if k.startswith("actor.") and k.endswith(".weight")})
    if not idx:
        raise KeyError("state dict has no actor.*.weight entries")
    layers = []
    for n, i in enumerate(idx):
        w = state_dict[f"actor.{i}.weight"].detach().to("cpu", torch.float)
        b = state_dict[f"actor.{i}.bias"].detach().to("cpu", torch.float)
        lin = torch.nn.Linear(w.shape[1], w.shape[0])
        with torch.no_grad():
            lin.weight.copy_(w); lin.bias.copy_(b)
        layers.append(lin)
        if n + 1 < len(idx):
            layers.append(torch.nn.ELU())
    return torch.nn.Sequential(*layers)


def export_policy_as_jit(actor_critic, path):
    """helpers.py:180-190 of the reference: TorchScript file ``<path>/policy_1.pt`` of the actor on the CPU.  The policy
    deployed on the robot is this file, so it is plain torch (no dependency on the CUDA library)."""
    import os
    import torch
    if getattr(actor_critic, "is_recurrent", False) or hasattr(actor_critic, "memory_a"):
        raise NotImplementedError("recurrent policies are not used by the pointfoot configs (DESIGN.md section 6)")
    os.makedirs(path, exist_ok=True)
    file = os.path.join(path, "policy_1.pt")
    model = actor_module_from_state_dict(actor_critic.state_dict())
    torch.jit.script(model).save(file)
    return file


def export_policy_as_onnx(actor_critic, path, num_obs=None, file_name="policy.onnx"):
    """export_policy_as_onnx.py:13-46 of the reference (``torch.onnx.export`` of the actor with a [1, num_obs] dummy input).
    Needs the ``onnx`` package at call time; raises ImportError otherwise (it is not in this image)."""
    import os
    import torch
    model = actor_module_from_state_dict(actor_critic.state_dict()).eval()
    n = num_obs if num_obs is not None else model[0].in_features
    os.makedirs(path, exist_ok=True)
    file = os.path.join(path, file_name)
    torch.onnx.export(model, torch.zeros(1, n), file, input_names=["obs"], output_names=["actions"], opset_version=13)
    return file
