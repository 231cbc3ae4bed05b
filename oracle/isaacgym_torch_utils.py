"""Restatement of the ``isaacgym.torch_utils`` helpers the hot path calls.  ORACLE / TEST CODE.

``isaacgym`` (Preview 3) is a closed binary that is not under /root/reference; these follow the
published op order recorded in SURVEY.md Appendix A.  Quaternions are xyzw.  Each statement is one
eager torch op so that fp32 rounding matches what the reference executes.
Call sites in the reference: legged_gym/envs/pointfoot/point_foot.py:209-211, 488, 557, 577, 583,
598-603, 730-732; legged_gym/utils/math.py:34-42.
"""
import torch


def to_torch(x, dtype=torch.float, device='cpu', requires_grad=False):
    return torch.tensor(x, dtype=dtype, device=device, requires_grad=requires_grad)


def get_axis_params(value, axis_idx, x_value=0., dtype=float, n_dims=3):
    zs = [0.0] * n_dims
    zs[axis_idx] = 1.0
    params = [z * value for z in zs]
    params[0] = x_value if axis_idx != 0 else params[0]
    return [dtype(p) for p in params]


def normalize(x, eps: float = 1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def quat_apply(a, b):
    shape = b.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 3)
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + a[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def quat_rotate(q, v):
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a + b + c


def quat_rotate_inverse(q, v):
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def torch_rand_float(lower, upper, shape, device):
    """Upstream: ``(upper-lower)*torch.rand(*shape)+lower``.  Here ``torch.rand`` is replaced by the
    counter-based stream when a draw context is active (see ``oracle.rng.DrawContext``)."""
    from . import rng
    r = rng.active_uniform(shape, device)
    if r is None:
        r = torch.rand(*shape, device=device)
    return (upper - lower) * r + lower
