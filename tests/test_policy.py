"""Actor forward + action sampling: ``go1_policy_act`` / ``k_policy`` against a plain fp32
reference of the same op (SB3 ``ActorCriticPolicy``: 48-64-64-12 tanh MLP, ``a = clip(mu + exp(log_std) eps)``),
with the weights of a SHIPPED policy (``tests/golden/policy_pos_0428.npz``, exported from the reference's
``models/2024-04-28.../final_model.zip`` by ``tests/golden/make_returns.py``)."""
import os

import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.lib import Policy, Sim

TOL = 2e-5     # fp32 scalar kernel (CPU emulation build, go1_policy_act_scalar): different summation order; actions are O(1)
# tensor-core kernel (go1_policy_act / go1_policy_forward on the GPU): tf32 operands (10-bit mantissa, relative rounding 4.9e-4 per
# operand), fp32 accumulation.  Three layers with O(1) tanh activations and weight rows of norm ~1..5: the error of the action means
# is 4e-4 on average and 1.4e-2 at worst on N(0,1) observations (the shipped policy has weight rows of norm up to 6.7, which
# amplify the 5e-4 operand rounding layer by layer; tools/policy_tc_debug.py prints the anatomy).  Declared bar: 2e-2 absolute.
TOL_TF32 = 2e-2


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
    # plain PyTorch fp32 reference of the same op
    torch.backends.cuda.matmul.allow_tf32 = False
    h = torch.tanh(obs @ dev[0].T + dev[1]); h = torch.tanh(h @ dev[2].T + dev[3])
    ref = torch.minimum(torch.maximum(h @ dev[4].T + dev[5], torch.from_numpy(lo).cuda()), torch.from_numpy(hi).cuda())
    torch.cuda.synchronize()            # the library runs on the handle's own stream: torch's fills / copies above must have landed
    # (1) the fp32 scalar kernel
    sim.policy_act_scalar(det, obs, act, n=n); torch.cuda.synchronize()
    assert (act - ref).abs().max().item() < TOL
    # (2) the tensor-core kernel (the product path: go1_policy_act, go1_rollout(GO1_ACTION_POLICY))
    act_tc = torch.zeros(n, 12, device="cuda"); torch.cuda.synchronize()
    sim.policy_act(det, obs, act_tc, n=n); torch.cuda.synchronize()
    err = (act_tc - ref).abs()
    print("tcgen05 tf32 policy: max |action - fp32 reference| = %.2e, mean %.2e" % (err.max().item(), err.mean().item()))
    # rows [0, n/8) are the 20x-scaled observations (|obs| up to ~80): tf32 rounds the INPUT to 3 digits, so their bar scales with it
    assert err[n // 8:].max().item() < TOL_TF32 and err[: n // 8].max().item() < 20 * TOL_TF32 and err.mean().item() < 1e-3
    # sampling statistics: wide bounds so clipping does not bias the moments
    big = Sim(pc.model_blob("torque"), 64, lib=cuda_lib_f32, seed=9)
    z = [torch.zeros_like(d) for d in dev[:6]]
    ls = torch.full((12,), -1.0, device="cuda")
    sto = Policy(*[t.data_ptr() for t in z], ls.data_ptr())
    o0 = torch.zeros(n, 48, device="cuda"); a = torch.zeros(n, 12, device="cuda")
    torch.cuda.synchronize()
    big.policy_act(sto, o0, a, n=n); torch.cuda.synchronize()
    s = float(np.exp(-1.0))
    assert abs(a.mean().item()) < 0.01 and abs(a.std().item() - s) < 0.01
    c = torch.corrcoef(a.T)
    assert (c - torch.eye(12, device="cuda")).abs().max().item() < 0.05      # independent across action dims
    # the noise stream does not depend on which kernel evaluates the network: same seed / counter -> same sample up to the tf32 mean
    big2 = Sim(pc.model_blob("torque"), 64, lib=cuda_lib_f32, seed=9)
    a2 = torch.zeros(n, 12, device="cuda")
    torch.cuda.synchronize()
    big2.policy_act_scalar(sto, o0, a2, n=n); torch.cuda.synchronize()
    assert (a - a2).abs().max().item() < 1e-6
    big.close(); big2.close(); sim.close()


@pytest.mark.gpu
def test_policy_forward_actor_critic_gpu(cuda_lib_f32):
    """ActorCriticPolicy.forward (train.py:122-128): actions, values and log-probs of actor 48-64-64-12 + critic 48-64-64-1 in ONE
    tcgen05 kernel, against plain fp32 PyTorch.  Ragged n, random (SB3-style orthogonal) critic, shipped actor."""
    import torch
    from quadruped_rl_locomotion_b200.lib import Critic
    w = _weights()
    n = 1000 + 3
    sim = Sim(pc.model_blob("torque"), 64, lib=cuda_lib_f32, seed=11)       # bounds +-1
    lo, hi = sim.action_bounds()
    dev = [torch.from_numpy(np.ascontiguousarray(w[k], dtype=np.float32)).cuda() for k in ("w0", "b0", "w1", "b1", "w2", "b2", "log_std")]
    dev[4] = (dev[4] * 0.05).contiguous(); dev[5] = torch.zeros_like(dev[5]); dev[6] = torch.full_like(dev[6], -2.0)   # small means, sigma 0.135: samples stay inside the bounds
    g = torch.Generator().manual_seed(3)
    def ortho(o, i, gain):
        x = torch.empty(o, i); torch.nn.init.orthogonal_(x, gain=gain, generator=g); return x.cuda().contiguous()
    cw = [ortho(64, 48, 2 ** 0.5), torch.randn(64, generator=g).cuda() * 0.1, ortho(64, 64, 2 ** 0.5), torch.randn(64, generator=g).cuda() * 0.1,
          ortho(1, 64, 1.0), torch.randn(1, generator=g).cuda()]
    pol = Policy(*[d.data_ptr() for d in dev])
    cri = Critic(*[t.data_ptr() for t in cw])
    obs = torch.from_numpy(_obs(n, seed=5)[n // 8:]).cuda().contiguous(); n = obs.shape[0]
    act = torch.zeros(n, 12, device="cuda"); val = torch.zeros(n, device="cuda"); lp = torch.zeros(n, device="cuda")
    torch.cuda.synchronize()
    sim.policy_forward(pol, cri, obs, act, val, lp, n=n); torch.cuda.synchronize()
    torch.backends.cuda.matmul.allow_tf32 = False
    h = torch.tanh(obs @ dev[0].T + dev[1]); h = torch.tanh(h @ dev[2].T + dev[3]); mu = h @ dev[4].T + dev[5]
    hc = torch.tanh(obs @ cw[0].T + cw[1]); hc = torch.tanh(hc @ cw[2].T + cw[3]); vref = (hc @ cw[4].T + cw[5])[:, 0]
    assert (val - vref).abs().max().item() < TOL_TF32, (val - vref).abs().max().item()
    # unclipped sample = mu + sigma * eps where the action was not clipped; log-prob as SB3's DiagGaussianDistribution
    sig = torch.exp(dev[6])
    lo_t, hi_t = torch.from_numpy(lo).cuda(), torch.from_numpy(hi).cuda()
    inside = (act > lo_t + 1e-6) & (act < hi_t - 1e-6)
    rows = inside.all(dim=1)
    assert rows.float().mean().item() > 0.2
    eps = (act[rows] - mu[rows]) / sig
    lp_ref = (-0.5 * eps ** 2 - dev[6] - 0.5 * np.log(2 * np.pi)).sum(dim=1)
    # eps is recovered through the tf32 mean: |d eps| = |d mu| / sigma, so the log-prob agrees to ~ |eps| * 5e-3 / sigma per dim
    assert (lp[rows] - lp_ref).abs().max().item() < 0.5 and (lp[rows] - lp_ref).abs().mean().item() < 0.05
    assert abs(eps.mean().item()) < 0.05 and abs(eps.std().item() - 1.0) < 0.05
    # critic == NULL leaves the value buffer alone; deterministic actor returns the mean
    val2 = torch.full((n,), 7.0, device="cuda"); act2 = torch.zeros(n, 12, device="cuda")
    det = Policy(*[d.data_ptr() for d in dev[:6]], None)
    torch.cuda.synchronize()
    sim.policy_forward(det, None, obs, act2, val2, None, n=n); torch.cuda.synchronize()
    assert torch.all(val2 == 7.0)
    assert (act2 - torch.minimum(torch.maximum(mu, lo_t), hi_t)).abs().max().item() < TOL_TF32
    sim.close()
