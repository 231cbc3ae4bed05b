"""ctypes binding of include/pointfoot_b200.h (the C-ABI of the CUDA library).

There is no CPU fallback: if ``libpointfoot_b200.so`` is missing or the process has no CUDA device, the
product path raises.  Build with ``python -c "import __graft_entry__ as g; g.build()"``.
"""
import ctypes as C
import os

PF_ABI_VERSION = 2
MAX_DOF, MAX_BODIES, NUM_FEET, MAX_REWARD_TERMS, MAX_HEIGHT_POINTS = 6, 9, 2, 24, 256

REWARD_IDS = {name: i for i, name in enumerate((
    "lin_vel_z", "ang_vel_xy", "orientation", "base_height", "torques", "dof_vel", "dof_acc", "action_rate",
    "collision", "termination", "dof_pos_limits", "dof_vel_limits", "torque_limits", "tracking_lin_vel",
    "tracking_ang_vel", "feet_air_time", "no_fly", "unbalance_feet_air_time", "unbalance_feet_height",
    "stumble", "stand_still", "feet_contact_forces", "feet_distance", "survival"))}
CONTROL_TYPES = {"P": 0, "V": 1, "T": 2}
TERRAIN_PLANE, TERRAIN_HEIGHTFIELD = 0, 1
DIV_TRUE, DIV_RECIPROCAL = 0, 1

_f, _i, _u = C.c_float, C.c_int32, C.c_uint32


class PfParams(C.Structure):
    _fields_ = [
        ("abi_version", _i), ("num_envs", _i), ("env_id_offset", C.c_int64),
        ("num_dof", _i), ("num_bodies", _i), ("num_commands", _i), ("num_obs", _i), ("num_priv_obs", _i),
        ("seed", _u),
        ("control_type", _i), ("decimation", _i),
        ("action_scale", _f), ("sim_dt", _f), ("dt", _f), ("clip_actions", _f), ("clip_obs", _f),
        ("p_gains", _f * MAX_DOF), ("d_gains", _f * MAX_DOF), ("torque_limits", _f * MAX_DOF),
        ("default_dof_pos", _f * MAX_DOF),
        ("dof_pos_limits", (_f * 2) * MAX_DOF), ("dof_vel_limits", _f * MAX_DOF),
        ("feet_indices", _i * NUM_FEET),
        ("num_termination", _i), ("termination_indices", _i * MAX_BODIES),
        ("num_penalised", _i), ("penalised_indices", _i * MAX_BODIES),
        ("max_episode_length", _f), ("max_episode_length_s", _f),
        ("resample_period", _i), ("heading_command", _i), ("resample_commands", _i),
        ("cmd_range", (_f * 2) * 4),
        ("obs_scale_ang_vel", _f), ("obs_scale_dof_pos", _f), ("obs_scale_dof_vel", _f), ("obs_scale_height", _f),
        ("commands_scale", _f * 3),
        ("add_noise", _i), ("measure_heights_actor", _i), ("measure_heights_critic", _i),
        ("terrain_kind", _i), ("index_div_mode", _i), ("hf_rows", _i), ("hf_cols", _i),
        ("border_size", _f), ("horizontal_scale", _f), ("vertical_scale", _f),
        ("num_height_points", _i),
        ("custom_origins", _i), ("terrain_curriculum", _i), ("max_terrain_level", _i), ("num_terrain_types", _i),
        ("terrain_env_length", _f), ("base_init_state", _f * 13), ("send_timeouts", _i),
        ("num_reward_terms", _i), ("reward_ids", _i * MAX_REWARD_TERMS), ("reward_scales", _f * MAX_REWARD_TERMS),
        ("has_termination_term", _i), ("termination_scale", _f),
        ("only_positive_rewards", _i), ("include_feet_height_rewards", _i),
        ("base_height_target", _f), ("soft_dof_vel_limit", _f), ("soft_torque_limit", _f),
        ("tracking_sigma", _f), ("max_contact_force", _f),
        ("min_feet_air_time", _f), ("max_feet_air_time", _f), ("min_feet_distance", _f),
        ("max_push_vel_xy", _f),
    ]


_P = C.c_void_p
ENV_BUFFER_FIELDS = (
    "root_states", "dof_state", "contact_forces", "rigid_body_states",
    "actions", "torques", "commands", "episode_length_buf",
    "base_lin_vel", "base_ang_vel", "projected_gravity",
    "last_actions", "last_dof_vel", "last_root_vel",
    "feet_air_time", "last_feet_air_time",
    "current_max_feet_height", "last_max_feet_height", "feet_height",
    "contact_filt", "first_contact", "last_contacts",
    "episode_sums", "rew_buf", "reset_buf", "time_out_buf", "terminate",
    "obs_buf", "priv_obs_buf", "measured_heights",
    "height_samples", "terrain_origins", "terrain_levels", "terrain_types", "env_origins",
    "reset_ids", "reset_count", "episode_stats", "extras_time_outs", "step_counter",
)


class PfEnvBuffers(C.Structure):
    _fields_ = [(name, _P) for name in ENV_BUFFER_FIELDS]


def lib_path():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_C", "libpointfoot_b200.so")


_lib = None


def load():
    """Load the CUDA library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.isfile(path):
        raise RuntimeError(f"pointfoot_b200: CUDA library not built ({path}); run __graft_entry__.build(). "
                           "There is no CPU fallback.")
    lib = C.CDLL(path)
    lib.pf_last_error.restype = C.c_char_p
    lib.pf_abi_version.restype = C.c_int
    lib.pf_debug_launch_count.restype = C.c_ulonglong
    if lib.pf_abi_version() != PF_ABI_VERSION:
        raise RuntimeError("pointfoot_b200: ABI version mismatch between capi.py and the built library")
    if lib.pf_sizeof_params() != C.sizeof(PfParams) or lib.pf_sizeof_env_buffers() != C.sizeof(PfEnvBuffers):
        raise RuntimeError("pointfoot_b200: struct layout mismatch between capi.py and include/pointfoot_b200.h")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().pf_last_error()
        raise RuntimeError(f"pointfoot_b200 {what} failed ({rc}): {msg.decode() if msg else ''}")


def ptr(t):
    """Raw device (or host) address of a torch tensor, NULL for None."""
    return C.c_void_p(None if t is None else t.data_ptr())


class PfHostState(C.Structure):
    _fields_ = [(n, _P) for n in ("root_states", "dof_state", "contact_forces", "rigid_body_states")]


# ------------------------------------------------------------------------------------------------ rsl_rl side
MAX_MLP_LAYERS = 6
RL_FP32, RL_BF16_TC = 0, 1


class PfRlParams(C.Structure):
    _fields_ = [
        ("abi_version", _i), ("num_envs", _i), ("env_id_offset", C.c_int64), ("num_steps", _i),
        ("num_obs", _i), ("num_critic_obs", _i), ("num_actions", _i),
        ("actor_layers", _i), ("actor_dims", _i * (MAX_MLP_LAYERS + 1)),
        ("critic_layers", _i), ("critic_dims", _i * (MAX_MLP_LAYERS + 1)),
        ("seed", _u),
        ("gamma", _f), ("lam", _f), ("clip_param", _f), ("value_loss_coef", _f), ("entropy_coef", _f),
        ("max_grad_norm", _f),
        ("use_clipped_value_loss", _i), ("adaptive_schedule", _i),
        ("num_mini_batches", _i), ("num_learning_epochs", _i), ("precision", _i),
        ("desired_kl", C.c_double), ("adam_beta1", C.c_double), ("adam_beta2", C.c_double), ("adam_eps", C.c_double),
    ]


RL_STORAGE_FIELDS = ("observations", "privileged_observations", "actions", "rewards", "dones", "values",
                     "actions_log_prob", "mu", "sigma", "returns", "advantages")
RL_LEARNER_FIELDS = ("params", "grads", "adam_m", "adam_v", "lr", "adam_step", "loss_stats")


class PfRlStorage(C.Structure):
    _fields_ = [(n, _P) for n in RL_STORAGE_FIELDS]


class PfRlLearner(C.Structure):
    _fields_ = [(n, _P) for n in RL_LEARNER_FIELDS]


def load_rl():
    """The same library, with the rsl_rl-side prototypes set and the struct layout checked."""
    lib = load()
    if getattr(lib, "_pf_rl_ready", False):
        return lib
    if lib.pf_sizeof_rl_params() != C.sizeof(PfRlParams):
        raise RuntimeError("pointfoot_b200: pf_rl_params layout mismatch between capi.py and the header")
    lib.pf_rl_param_count.restype = C.c_int64
    lib.pf_rl_param_offset.restype = C.c_int64
    lib.pf_rl_param_count.argtypes = [C.c_void_p]
    lib.pf_rl_param_offset.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    lib._pf_rl_ready = True
    return lib


# ------------------------------------------------------------------------------------------------ actuator network
ACTUATOR_MAX_DOF, ACTUATOR_NET_WEIGHTS = 12, 969


class PfActuatorNetParams(C.Structure):
    _fields_ = [("num_envs", _i), ("num_dof", _i), ("action_scale", _f), ("in_scale", _f * 2), ("out_scale", _f),
                ("default_dof_pos", _f * ACTUATOR_MAX_DOF)]


# ------------------------------------------------------------------------------------------------ LeggedRobot (N2)
LR_MAX_DOF, LR_MAX_BODIES, LR_MAX_FEET = 12, 24, 4


class PfLrParams(C.Structure):
    _fields_ = [
        ("abi_version", _i), ("num_envs", _i), ("env_id_offset", C.c_int64),
        ("num_dof", _i), ("num_bodies", _i), ("num_feet", _i), ("num_commands", _i), ("num_obs", _i),
        ("seed", _u),
        ("control_type", _i), ("decimation", _i),
        ("action_scale", _f), ("sim_dt", _f), ("dt", _f), ("clip_actions", _f), ("clip_obs", _f),
        ("p_gains", _f * LR_MAX_DOF), ("d_gains", _f * LR_MAX_DOF), ("torque_limits", _f * LR_MAX_DOF),
        ("default_dof_pos", _f * LR_MAX_DOF),
        ("dof_pos_limits", (_f * 2) * LR_MAX_DOF), ("dof_vel_limits", _f * LR_MAX_DOF),
        ("feet_indices", _i * LR_MAX_FEET),
        ("num_termination", _i), ("termination_indices", _i * LR_MAX_BODIES),
        ("num_penalised", _i), ("penalised_indices", _i * LR_MAX_BODIES),
        ("max_episode_length", _f), ("max_episode_length_s", _f),
        ("resample_period", _i), ("heading_command", _i),
        ("cmd_range", (_f * 2) * 4),
        ("obs_scale_lin_vel", _f), ("obs_scale_ang_vel", _f), ("obs_scale_dof_pos", _f), ("obs_scale_dof_vel", _f),
        ("obs_scale_height", _f),
        ("commands_scale", _f * 3),
        ("add_noise", _i), ("measure_heights", _i),
        ("terrain_kind", _i), ("index_div_mode", _i), ("hf_rows", _i), ("hf_cols", _i),
        ("border_size", _f), ("horizontal_scale", _f), ("vertical_scale", _f),
        ("num_height_points", _i),
        ("custom_origins", _i), ("terrain_curriculum", _i), ("max_terrain_level", _i), ("num_terrain_types", _i),
        ("terrain_env_length", _f), ("base_init_state", _f * 13), ("send_timeouts", _i),
        ("num_reward_terms", _i), ("reward_ids", _i * MAX_REWARD_TERMS), ("reward_scales", _f * MAX_REWARD_TERMS),
        ("has_termination_term", _i), ("termination_scale", _f),
        ("only_positive_rewards", _i),
        ("base_height_target", _f), ("soft_dof_vel_limit", _f), ("soft_torque_limit", _f),
        ("tracking_sigma", _f), ("max_contact_force", _f),
        ("max_push_vel_xy", _f),
    ]


LR_BUFFER_FIELDS = (
    "root_states", "dof_state", "contact_forces", "actions", "torques", "commands", "episode_length_buf",
    "base_lin_vel", "base_ang_vel", "projected_gravity", "last_actions", "last_dof_vel", "last_root_vel",
    "feet_air_time", "last_contacts", "episode_sums", "rew_buf", "reset_buf", "time_out_buf", "obs_buf",
    "measured_heights", "height_samples", "terrain_origins", "terrain_levels", "terrain_types", "env_origins",
    "reset_ids", "reset_count", "episode_stats", "extras_time_outs", "noise_scale_vec", "height_points_xy",
)


class PfLrBuffers(C.Structure):
    _fields_ = [(name, _P) for name in LR_BUFFER_FIELDS]


def load_lr():
    """The same library with the LeggedRobot struct layouts checked."""
    lib = load()
    if getattr(lib, "_pf_lr_ready", False):
        return lib
    if lib.pf_lr_sizeof_params() != C.sizeof(PfLrParams) or lib.pf_lr_sizeof_buffers() != C.sizeof(PfLrBuffers):
        raise RuntimeError("pointfoot_b200: pf_lr_params / pf_lr_buffers layout mismatch between capi.py and the header")
    lib.pf_lr_post_physics_step.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]
    lib.pf_lr_pd_torques.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib._pf_lr_ready = True
    return lib


# ------------------------------------------------------------------------------------------------ terrain generation (N1)
class PfTerrainParams(C.Structure):
    _fields_ = [("abi_version", _i), ("num_rows", _i), ("num_cols", _i), ("cell_px", _i), ("border_px", _i), ("tot_rows", _i),
                ("tot_cols", _i), ("seed", _u), ("origin_patch", _i * 4), ("env_length", C.c_double), ("env_width", C.c_double),
                ("vertical_scale", C.c_double)]


def load_terrain():
    lib = load()
    if getattr(lib, "_pf_terrain_ready", False):
        return lib
    if lib.pf_terrain_sizeof_params() != C.sizeof(PfTerrainParams):
        raise RuntimeError("pointfoot_b200: pf_terrain_params layout mismatch between capi.py and the header")
    lib._pf_terrain_ready = True
    return lib
