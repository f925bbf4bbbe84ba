"""Actor forward + action sampling: ``go1_policy_act`` / ``k_policy`` against a plain fp32
reference of the same op (SB3 ``ActorCriticPolicy``: 48-64-64-12 tanh MLP, ``a = clip(mu + exp(log_std) eps)``),
with the weights of a SHIPPED policy (``tests/golden/policy_pos_0428.npz``, exported from the reference's
``models/2024-04-28.../final_model.zip`` by ``tests/golden/make_returns.py``)."""
import os

import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.lib import Policy, Sim

TOL = 2e-5     # fp32 MLP, different summation order; actions are O(1)


def _weights():
    return dict(np.load(os.path.join(pc.GOLDEN, "policy_pos_0428.npz")))


def _ref(w, obs, lo, hi):
    h = np.tanh(obs @ w["w0"].T + w["b0"])
    h = np.tanh(h @ w["w1"].T + w["b1"])
    return np.clip(h @ w["w2"].T + w["b2"], lo, hi)


def _obs(n, seed=0):
    rng = np.random.default_rng(seed)
    o = rng.normal(size=(n, 48)).astype(np.float32)
    o[: n // 8] *= 20.0                      # some saturated rows (tanh tails, clipping at the action bounds)
    return o


def test_policy_act_emu(emu_lib_f32):
    """CPU: the same CUDA source through the SIMT emulation (device pointers are host pointers there)."""
    w = _weights()
    n = 37                                   # ragged: not a multiple of the block size
    sim = Sim(pc.model_blob("position"), n, lib=emu_lib_f32, seed=4)
    lo, hi = sim.action_bounds()
    keep = [np.ascontiguousarray(w[k], dtype=np.float32) for k in ("w0", "b0", "w1", "b1", "w2", "b2", "log_std")]
    obs = _obs(n); act = np.zeros((n, 12), np.float32)
    det = Policy(*[k.ctypes.data for k in keep[:6]], None)
    sim.policy_act(det, obs, act)
    assert np.abs(act - _ref(w, obs, lo, hi)).max() < TOL
    sim.policy_act(det, obs, act, n=0)       # empty input is a no-op
    # stochastic: noise = exp(log_std) * N(0,1), fresh every call, clipped to the Box
    sto = Policy(*[k.ctypes.data for k in keep])
    a1 = np.zeros((n, 12), np.float32); a2 = np.zeros((n, 12), np.float32)
    sim.policy_act(sto, obs, a1); sim.policy_act(sto, obs, a2)
    assert not np.array_equal(a1, a2)
    assert np.all(a1 >= lo - 1e-6) and np.all(a1 <= hi + 1e-6)
    sim.close()


@pytest.mark.gpu
def test_policy_act_gpu(cuda_lib_f32):
    import torch
    w = _weights()
    n = 16384 + 5
    sim = Sim(pc.model_blob("position"), 64, lib=cuda_lib_f32, seed=4)
    lo, hi = sim.action_bounds()
    dev = [torch.from_numpy(np.ascontiguousarray(w[k], dtype=np.float32)).cuda() for k in ("w0", "b0", "w1", "b1", "w2", "b2", "log_std")]
    obs = torch.from_numpy(_obs(n)).cuda(); act = torch.zeros(n, 12, device="cuda")
    det = Policy(*[d.data_ptr() for d in dev[:6]], None)
    sim.policy_act(det, obs, act, n=n); torch.cuda.synchronize()
    # plain PyTorch fp32 reference of the same op
    torch.backends.cuda.matmul.allow_tf32 = False
    h = torch.tanh(obs @ dev[0].T + dev[1]); h = torch.tanh(h @ dev[2].T + dev[3])
    ref = torch.minimum(torch.maximum(h @ dev[4].T + dev[5], torch.from_numpy(lo).cuda()), torch.from_numpy(hi).cuda())
    assert (act - ref).abs().max().item() < TOL
    # sampling statistics: wide bounds so clipping does not bias the moments
    big = Sim(pc.model_blob("torque"), 64, lib=cuda_lib_f32, seed=9)
    z = [torch.zeros_like(d) for d in dev[:6]]
    ls = torch.full((12,), -1.0, device="cuda")
    sto = Policy(*[t.data_ptr() for t in z], ls.data_ptr())
    o0 = torch.zeros(n, 48, device="cuda"); a = torch.zeros(n, 12, device="cuda")
    big.policy_act(sto, o0, a, n=n); torch.cuda.synchronize()
    s = float(np.exp(-1.0))
    assert abs(a.mean().item()) < 0.01 and abs(a.std().item() - s) < 0.01
    c = torch.corrcoef(a.T)
    assert (c - torch.eye(12, device="cuda")).abs().max().item() < 0.05      # independent across action dims
    big.close(); sim.close()
