"""Actuator-network torques of legged_gym/envs/anymal_c/anymal.py: what ``Anymal`` adds to
``LeggedRobot`` -- ``_compute_torques`` through the TorchScript SEA network (anymal.py:62-72), its recurrent state
buffers (anymal.py:43-50) and their reset in ``reset_idx`` (anymal.py:52-56) -- as launches into libpointfoot_b200.so
(``pf_actuator_net_torques`` / ``pf_actuator_net_reset``).  No CPU fallback.
"""
import ctypes as C

import torch

from . import capi

WEIGHT_KEYS = ("lstm.weight_ih_l0", "lstm.weight_hh_l0", "lstm.bias_ih_l0", "lstm.bias_hh_l0", "lstm.weight_ih_l1",
               "lstm.weight_hh_l1", "lstm.bias_ih_l1", "lstm.bias_hh_l1", "linear.weight", "linear.bias")


class ActuatorNetwork:
    """``sea_*`` buffers + ``compute_torques`` for ``num_envs`` x ``num_dof`` actuators on ``device``."""

    def __init__(self, num_envs, num_dof, action_scale, default_dof_pos, device="cuda:0"):
        self.device = torch.device(device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("pointfoot_b200.anymal needs a CUDA device (no CPU fallback)")
        if not 0 < num_dof <= capi.ACTUATOR_MAX_DOF:
            raise ValueError(f"num_dof must be in 1..{capi.ACTUATOR_MAX_DOF}")
        self._lib = capi.load()
        self.num_envs, self.num_dof = num_envs, num_dof
        P = capi.PfActuatorNetParams()
        P.num_envs, P.num_dof, P.action_scale = num_envs, num_dof, action_scale
        for i, v in enumerate(torch.as_tensor(default_dof_pos, dtype=torch.float).flatten().tolist()):
            P.default_dof_pos[i] = v
        self._P = P
        self.weights = None
        # anymal.py:47-50
        self.sea_hidden_state = torch.zeros(2, num_envs * num_dof, 8, device=self.device)
        self.sea_cell_state = torch.zeros(2, num_envs * num_dof, 8, device=self.device)
        self.sea_hidden_state_per_env = self.sea_hidden_state.view(2, num_envs, num_dof, 8)
        self.sea_cell_state_per_env = self.sea_cell_state.view(2, num_envs, num_dof, 8)
        self.torques = torch.zeros(num_envs, num_dof, device=self.device)

    def load_state_dict(self, sd):
        """state dict of the reference's module (``torch.jit.load(actuator_net_file).state_dict()``, anymal.py:39-40)"""
        flat = torch.cat([torch.as_tensor(sd[k], dtype=torch.float).flatten() for k in WEIGHT_KEYS])
        if flat.numel() != capi.ACTUATOR_NET_WEIGHTS:
            raise RuntimeError("actuator network is not the 2-layer LSTM(2 -> 8 -> 8) + Linear(8 -> 1) of anydrive_v3_lstm.pt")
        self.weights = flat.to(self.device).contiguous()
        s_in = torch.as_tensor(sd["in_scale"], dtype=torch.float).flatten().tolist()
        self._P.in_scale[0], self._P.in_scale[1] = s_in
        self._P.out_scale = float(torch.as_tensor(sd["out_scale"]).flatten()[0])
        return self

    def load(self, path):
        return self.load_state_dict(torch.jit.load(path, map_location="cpu").state_dict())

    def _st(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def compute_torques(self, actions, dof_state):
        """anymal.py:62-72; ``dof_state`` is the PhysX tensor [N * A, 2] (pos, vel)."""
        if self.weights is None:
            raise RuntimeError("actuator network weights not loaded")
        n = self.num_envs * self.num_dof
        for name, t, cnt in (("actions", actions, n), ("dof_state", dof_state, 2 * n)):
            if t.device != self.device or t.dtype != torch.float32 or t.numel() != cnt or not t.is_contiguous():
                raise RuntimeError(f"compute_torques({name}): expected a contiguous fp32 tensor of {cnt} elements on {self.device}")
        p = lambda t: C.c_void_p(t.data_ptr())
        capi.check(self._lib.pf_actuator_net_torques(C.byref(self._P), p(self.weights), p(actions), p(dof_state),
                                                     p(self.sea_hidden_state), p(self.sea_cell_state), p(self.torques),
                                                     self._st()), "pf_actuator_net_torques")
        return self.torques

    def reset(self, reset_ids, reset_count):
        """anymal.py:52-56; ``reset_ids`` int32[N] / ``reset_count`` int32[1] on the device, as the env step leaves them."""
        p = lambda t: C.c_void_p(t.data_ptr())
        capi.check(self._lib.pf_actuator_net_reset(C.c_int32(self.num_envs), C.c_int32(self.num_dof), p(reset_ids), p(reset_count),
                                                   p(self.sea_hidden_state), p(self.sea_cell_state), self._st()),
                   "pf_actuator_net_reset")
