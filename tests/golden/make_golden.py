"""Generate tests/golden/*.npz by running the UNMODIFIED reference point_foot.py (oracle/ref_harness.py) on seeded
synthetic states.  Only runs where /root/reference exists; the .npz files are committed and travel to the GPU box.

    python tests/golden/make_golden.py

Each file holds the inputs (every state tensor except the 5.5 MB heightfield, which is regenerated from its seed
and pinned by a checksum) and, for each of STEPS chained steps, the outputs and the state the reference left.
"""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import common                                     # noqa: E402
from pointfoot_b200 import synth                  # noqa: E402
from oracle import ref_harness                    # noqa: E402

STEPS = 3
N = 64
FIELDS = ("root_states", "dof_state", "commands", "episode_length_buf", "last_actions", "last_dof_vel",
          "last_root_vel", "feet_air_time", "last_feet_air_time", "current_max_feet_height", "last_max_feet_height",
          "contact_filt", "first_contact", "last_contacts", "base_lin_vel", "base_ang_vel", "projected_gravity",
          "torques", "terminate", "time_out_buf")


def actions_for(task, seed, i):
    return synth.make_state(N, task, seed=seed + 100 + i)["actions"] * (30.0 if i == 1 else 1.0)


def hf_checksum(hf):
    return hashlib.sha256(hf.numpy().tobytes()).hexdigest()


def generate(fname, task, mutate=None, start_step=0, seed=1):
    # (19 rows of episode sums unless the cfg mutation enables more terms; kept so the older files' inputs stay as they were)
    state = synth.make_state(N, task, seed=seed, adversarial=True,
                             num_reward_terms=24 if mutate is common.all_reward_terms else 19)
    state["common_step_counter"] = start_step
    out = {"start_step": np.int64(start_step)}
    for k, v in state.items():
        if k == "height_samples":
            out["hf_sha256"] = np.frombuffer(hf_checksum(v).encode(), dtype=np.uint8)
        elif torch.is_tensor(v):
            out["in_" + k] = v.numpy().copy()
    env = ref_harness.build_reference_env(task, common.clone_state(state), seed=seed, mutate_cfg=mutate)
    names = list(env.episode_sums.keys())
    out["reward_names"] = np.array(names)
    for i in range(STEPS):
        a = actions_for(task, seed, i)
        obs, priv, rew, reset, extras = env.step(a.clone())
        p = f"s{i}_"
        out[p + "obs"], out[p + "rew"], out[p + "reset"] = obs.numpy().copy(), rew.numpy().copy(), reset.numpy().copy()
        if priv is not None:
            out[p + "priv"] = priv.numpy().copy()
        for k in FIELDS:
            out[p + k] = getattr(env, k).numpy().copy()
        out[p + "episode_sums"] = torch.stack([env.episode_sums[n] for n in names]).numpy().copy()
        out[p + "reset_ids"] = reset.nonzero().flatten().to(torch.int32).numpy()
        if task == "pointfoot_rough":
            out[p + "terrain_levels"] = env.terrain_levels.numpy().copy()
            out[p + "env_origins"] = env.env_origins.numpy().copy()
            out[p + "measured_heights"] = env.measured_heights.numpy().copy()
        ep = extras.get("episode", {})
        out[p + "extras_episode"] = np.array([float(ep["rew_" + n]) for n in names] if ep else [], np.float32)
        out[p + "time_outs"] = extras["time_outs"].numpy().copy()
        out[p + "push_forces"] = env.rigid_body_external_forces.numpy().copy()      # PF:590-611 (zeros until a push step)
    path = os.path.join(HERE, fname)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


ACTUATOR_FILE = "anymal_actuator_n64.npz"


def generate_actuator(steps=6, seed=3):
    """Anymal._compute_torques with the actuator network (anymal.py:62-78) over ``steps`` chained sub-steps, the recurrent
    state of a few envs zeroed before step 3 as Anymal.reset_idx does (anymal.py:52-56)."""
    env = ref_harness.build_reference_anymal_torques(N)
    net = env.actuator_network
    sd = net.state_dict()
    keys = [k for k in sd if k not in ("in_scale", "out_scale")]
    out = {"weights": torch.cat([sd[k].flatten() for k in keys]).numpy().copy(), "weight_keys": np.array(keys),
           "in_scale": sd["in_scale"].flatten().numpy().copy(), "out_scale": sd["out_scale"].flatten().numpy().copy(),
           "action_scale": np.float32(env.cfg.control.action_scale), "default_dof_pos": env.default_dof_pos.flatten().numpy().copy(),
           "reset_ids": np.array([0, 5, 17, 63], np.int32), "reset_before_step": np.int64(3)}
    g = torch.Generator().manual_seed(seed)
    for i in range(steps):
        if i == 3:
            ids = torch.from_numpy(out["reset_ids"]).long()
            env.sea_hidden_state_per_env[:, ids] = 0.          # anymal.py:55-56
            env.sea_cell_state_per_env[:, ids] = 0.
        actions = torch.randn(N, 12, generator=g) * (3.0 if i == 4 else 1.0)
        env.dof_state[:, 0] = torch.randn(N * 12, generator=g) * 0.5
        env.dof_state[:, 1] = torch.randn(N * 12, generator=g) * 4.0
        t = env._compute_torques(actions)
        p = f"s{i}_"
        out[p + "actions"], out[p + "dof_state"] = actions.numpy().copy(), env.dof_state.numpy().copy()
        out[p + "torques"] = t.numpy().copy().reshape(N, 12)
        out[p + "hidden"], out[p + "cell"] = env.sea_hidden_state.numpy().copy(), env.sea_cell_state.numpy().copy()
    path = os.path.join(HERE, ACTUATOR_FILE)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


LR_FIELDS = ("root_states", "dof_state", "commands", "episode_length_buf", "last_actions", "last_dof_vel", "last_root_vel",
             "feet_air_time", "last_contacts", "base_lin_vel", "base_ang_vel", "projected_gravity", "torques", "time_out_buf")


def generate_lr(fname, mutate=None, start_step=0, seed=1):
    """The UNMODIFIED LeggedRobot (legged_gym/envs/base/legged_robot.py) with the a1 config, STEPS chained step() calls."""
    import parity_lr
    state = synth.make_lr_state(N, seed=seed, adversarial=True)
    state["common_step_counter"] = start_step
    out = {"start_step": np.int64(start_step)}
    for k, v in state.items():
        if k == "height_samples":
            out["hf_sha256"] = np.frombuffer(hf_checksum(v).encode(), dtype=np.uint8)
        elif torch.is_tensor(v):
            out["in_" + k] = v.numpy().copy()
    env = ref_harness.build_reference_legged_robot(common.clone_state(state), seed=seed, mutate_cfg=mutate)
    names = list(env.episode_sums.keys())
    out["reward_names"] = np.array(names)
    out["num_obs"] = np.int64(env.num_obs)
    cover = {"resets": 0, "resampled": 0, "pushes": 0, "clipped_rew": 0, "positive_rew": 0}
    for i in range(STEPS):
        a = parity_lr.actions_for(N, seed, i)
        before = env.commands.clone()
        obs, priv, rew, reset, extras = env.step(a.clone())
        assert priv is None
        p = f"s{i}_"
        out[p + "obs"], out[p + "rew"], out[p + "reset"] = obs.numpy().copy(), rew.numpy().copy(), reset.numpy().copy()
        for k in LR_FIELDS:
            out[p + k] = getattr(env, k).numpy().copy()
        out[p + "episode_sums"] = torch.stack([env.episode_sums[n] for n in names]).numpy().copy()
        out[p + "reset_ids"] = reset.nonzero().flatten().to(torch.int32).numpy()
        if env.custom_origins:
            out[p + "terrain_levels"] = env.terrain_levels.numpy().copy()
            out[p + "env_origins"] = env.env_origins.numpy().copy()
        if torch.is_tensor(env.measured_heights):
            out[p + "measured_heights"] = env.measured_heights.numpy().copy()
        ep = extras.get("episode", {})
        out[p + "extras_episode"] = np.array([float(ep["rew_" + n]) for n in names] if ep else [], np.float32)
        if "terrain_level" in ep:
            out[p + "extras_terrain_level"] = np.float32(float(ep["terrain_level"]))
        out[p + "time_outs"] = extras["time_outs"].numpy().copy()
        cover["resets"] += int(reset.sum())
        cover["resampled"] += int(((env.commands[:, :2] != before[:, :2]).any(1) & ~reset.bool()).sum())
        cover["pushes"] += int(env.common_step_counter % env.cfg.domain_rand.push_interval == 0 and env.cfg.domain_rand.push_robots)
        cover["clipped_rew"] += int((rew == 0).sum())
        cover["positive_rew"] += int((rew > 0).sum())
    path = os.path.join(HERE, fname)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes", cover)


def generate_terrain(fname, mutate, keep_field):
    """The UNMODIFIED legged_gym/utils/terrain.py::Terrain over oracle/terrain_ref.py (restated isaacgym.terrain_utils)."""
    from oracle import terrain_ref
    import parity_terrain
    t = ref_harness.build_reference_terrain(mutate, seed=1)
    hf = t.height_field_raw
    out = {"shape": np.array(hf.shape), "geometry": np.array([t.tot_rows, t.tot_cols, t.border, t.width_per_env_pixels]),
           "hf_sha256": np.frombuffer(parity_terrain.sha(hf).encode(), np.uint8), "env_origins": t.env_origins.copy()}
    if keep_field:
        out["height_field"] = hf.copy()
    else:
        r0 = t.border + 3 * t.length_per_env_pixels - 40           # two rows of cells' edges, every column type
        out["slab_rows"], out["slab"] = np.array([r0, r0 + 160]), hf[r0:r0 + 160].copy()
    assert t.type == "trimesh"
    moved = terrain_ref.convert_heightfield_to_trimesh(hf, t.cfg.horizontal_scale, t.cfg.vertical_scale, None)[0]
    out["vertices_sha256"] = np.frombuffer(parity_terrain.sha(t.vertices).encode(), np.uint8)
    out["triangles_sha256"] = np.frombuffer(parity_terrain.sha(t.triangles).encode(), np.uint8)
    out["moved_vertices"] = np.int64((moved[:, :2] != t.vertices[:, :2]).any(1).sum())
    path = os.path.join(HERE, fname)
    np.savez_compressed(path, **out)
    kinds = {}
    print("wrote", path, os.path.getsize(path), "bytes; min/max", hf.min(), hf.max(), "moved vertices", int(out["moved_vertices"]))


if __name__ == "__main__":
    if not ref_harness.reference_available():
        raise SystemExit("reference not available here; golden files can only be regenerated in the build container")
    import parity
    only = sys.argv[1:]
    for fname, (task, mutate, start_step) in parity.GOLDEN.items():
        if only and fname not in only:
            continue
        generate(fname, task, mutate, start_step)
    if not only or ACTUATOR_FILE in only:
        generate_actuator()
    import parity_terrain
    for fname, (mutate, keep) in parity_terrain.TERRAIN_GOLDEN.items():
        if not only or fname in only:
            generate_terrain(fname, mutate, keep)
    import parity_lr
    for fname, (mutate, start_step) in parity_lr.LR_GOLDEN.items():
        if not only or fname in only:
            generate_lr(fname, mutate, start_step)
