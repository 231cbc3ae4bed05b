"""Anymal's actuator-network torques (legged_gym/envs/anymal_c/anymal.py:62-78, 52-56) through the C-ABI against the
outputs of the UNMODIFIED reference method run with the reference's own TorchScript module
(tests/golden/anymal_actuator_n64.npz, made by tests/golden/make_golden.py): torques and recurrent state after each of six
chained sub-steps, a state reset in between.  fp32 with expf / tanhf on both sides: the recurrent state (unit scale) within
the suite's rtol 1e-5 / atol 2e-6 (parity.RTOL / ATOL); the torque is that unit-scale output times out_scale = 20, so its
absolute bar is out_scale x ATOL = 4e-5 (measured worst case on the B200: 1.05e-5), same rtol."""
import os

import numpy as np
import pytest
import torch

import parity

GOLD = os.path.join(parity.GOLDEN_DIR, "anymal_actuator_n64.npz")


def _net(z, n, device="cuda:0"):
    from pointfoot_b200.anymal import ActuatorNetwork, WEIGHT_KEYS
    net = ActuatorNetwork(n, 12, float(z["action_scale"]), torch.from_numpy(z["default_dof_pos"]), device)
    sd, off = {"in_scale": torch.from_numpy(z["in_scale"]), "out_scale": torch.from_numpy(z["out_scale"])}, 0
    shapes = {"lstm.weight_ih_l0": (32, 2), "lstm.weight_hh_l0": (32, 8), "lstm.weight_ih_l1": (32, 8), "lstm.weight_hh_l1": (32, 8),
              "linear.weight": (1, 8), "linear.bias": (1,)}
    assert [str(k) for k in z["weight_keys"]] == list(WEIGHT_KEYS)
    for k in WEIGHT_KEYS:
        shp = shapes.get(k, (32,))
        cnt = int(np.prod(shp))
        sd[k] = torch.from_numpy(z["weights"][off:off + cnt].reshape(shp).copy()); off += cnt
    return net.load_state_dict(sd)


@pytest.mark.gpu
def test_actuator_network_matches_reference_golden():
    z = np.load(GOLD)
    n = z["s0_actions"].shape[0]
    net = _net(z, n)
    ids = torch.zeros(n, dtype=torch.int32, device="cuda")
    ids[:len(z["reset_ids"])] = torch.from_numpy(z["reset_ids"]).cuda()
    cnt = torch.tensor([len(z["reset_ids"])], dtype=torch.int32, device="cuda")
    for i in range(6):
        if i == int(z["reset_before_step"]):
            net.reset(ids, cnt)
        t = net.compute_torques(torch.from_numpy(z[f"s{i}_actions"]).cuda(), torch.from_numpy(z[f"s{i}_dof_state"]).cuda())
        torch.cuda.synchronize()
        for name, got in (("torques", t), ("hidden", net.sea_hidden_state), ("cell", net.sea_cell_state)):
            atol = parity.ATOL * (float(z["out_scale"][0]) if name == "torques" else 1.0)
            torch.testing.assert_close(got.cpu(), torch.from_numpy(z[f"s{i}_{name}"]), rtol=parity.RTOL, atol=atol,
                                       msg=lambda m, nm=name, i=i: f"step {i} {nm}: {m}")


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 333, 20000])
def test_actuator_network_matches_torch_lstm_at_other_sizes(n):
    """ragged / large env counts against a plain torch.nn.LSTM + Linear carrying the golden file's weights (fp32, CPU)"""
    z = np.load(GOLD)
    net = _net(z, n)
    lstm, lin = torch.nn.LSTM(2, 8, 2, batch_first=True), torch.nn.Linear(8, 1)
    off = 0
    with torch.no_grad():
        for k in [str(k) for k in z["weight_keys"]]:
            prm = getattr(lstm, k[5:]) if k.startswith("lstm.") else getattr(lin, k[7:])
            prm.copy_(torch.from_numpy(z["weights"][off:off + prm.numel()].copy()).view(prm.shape)); off += prm.numel()
    g = torch.Generator().manual_seed(n)
    h, c = torch.zeros(2, n * 12, 8), torch.zeros(2, n * 12, 8)
    default = torch.from_numpy(z["default_dof_pos"])
    for i in range(3):
        a, ds = torch.randn(n, 12, generator=g), torch.randn(n * 12, 2, generator=g) * torch.tensor([0.5, 4.0])
        x = torch.stack(((a * float(z["action_scale"]) + default - ds[:, 0].view(n, 12)).flatten(), ds[:, 1]), 1).unsqueeze(1)
        with torch.no_grad():
            y, (h, c) = lstm(x * torch.from_numpy(z["in_scale"]), (h, c))
            want = float(z["out_scale"][0]) * lin(y).squeeze()
        got = net.compute_torques(a.cuda(), ds.cuda().contiguous())
        torch.testing.assert_close(got.cpu().flatten(), want.flatten(), rtol=parity.RTOL, atol=parity.ATOL * float(z["out_scale"][0]))
    torch.testing.assert_close(net.sea_cell_state.cpu(), c, rtol=1e-5, atol=2e-6)


def test_actuator_golden_matches_live_reference():
    """where the reference tree exists: re-run the unmodified Anymal._compute_torques and compare with the committed file"""
    from oracle import ref_harness
    if not os.path.isfile("/root/reference/resources/actuator_nets/anydrive_v3_lstm.pt"):
        pytest.skip("reference tree not present (GPU box)")
    z = np.load(GOLD)
    n = z["s0_actions"].shape[0]
    env = ref_harness.build_reference_anymal_torques(n)
    for i in range(6):
        if i == int(z["reset_before_step"]):
            ids = torch.from_numpy(z["reset_ids"]).long()
            env.sea_hidden_state_per_env[:, ids] = 0.
            env.sea_cell_state_per_env[:, ids] = 0.
        env.dof_state.copy_(torch.from_numpy(z[f"s{i}_dof_state"]))
        t = env._compute_torques(torch.from_numpy(z[f"s{i}_actions"]))
        assert torch.equal(t.reshape(n, 12), torch.from_numpy(z[f"s{i}_torques"]))
        assert torch.equal(env.sea_cell_state, torch.from_numpy(z[f"s{i}_cell"]))
