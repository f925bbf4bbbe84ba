"""The device-resident PPO loop (caller side of reference train.py:112-138): its arithmetic on CPU tensors
against plain numpy restatements of SB3 2.3.0's RolloutBuffer / PPO.train, the flat gradient bucket over two gloo ranks, and the
stable-baselines3 zip layout (a shipped final_model.zip loads; a zip written here holds SB3's tensor names)."""
import io
import os
import socket
import zipfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity_common as pc
from quadruped_rl_locomotion_b200 import ppo

SB3_KEYS = {"log_std", "mlp_extractor.policy_net.0.weight", "mlp_extractor.policy_net.0.bias", "mlp_extractor.policy_net.2.weight",
            "mlp_extractor.policy_net.2.bias", "mlp_extractor.value_net.0.weight", "mlp_extractor.value_net.0.bias",
            "mlp_extractor.value_net.2.weight", "mlp_extractor.value_net.2.bias", "action_net.weight", "action_net.bias",
            "value_net.weight", "value_net.bias"}


def test_network_has_sb3_names_and_defaults():
    torch.manual_seed(0)
    net = ppo.ActorCritic()
    assert set(net.state_dict()) == SB3_KEYS
    assert sum(p.numel() for p in net.parameters()) == 15449          # 8 088 actor + 7 361 critic: the gradient bucket
    assert torch.all(net.log_std == 0) and net.action_net.weight.abs().max() < 0.05     # log_std_init = 0, action gain 0.01
    w = net.mlp_extractor.policy_net[0].weight
    assert torch.allclose(w.T @ w, 2.0 * torch.eye(48), atol=1e-4)    # orthogonal columns, gain sqrt 2


def _gae_numpy(rew, val, done, last_val, gamma, lam):
    """stable_baselines3.common.buffers.RolloutBuffer.compute_returns_and_advantage, written out."""
    T = len(rew)
    adv = np.zeros_like(rew)
    last = 0.0
    for step in reversed(range(T)):
        if step == T - 1:
            next_non_terminal = 1.0 - done[step]; next_values = last_val
        else:
            next_non_terminal = 1.0 - done[step]; next_values = val[step + 1]
        delta = rew[step] + gamma * next_values * next_non_terminal - val[step]
        last = delta + gamma * lam * next_non_terminal * last
        adv[step] = last
    return adv, adv + val


def test_gae_matches_sb3_restatement():
    rng = np.random.default_rng(0)
    T, N = 17, 5
    rew = rng.normal(size=(T, N)); val = rng.normal(size=(T, N)); last = rng.normal(size=N)
    done = (rng.random((T, N)) < 0.2).astype(np.float64)
    a, r = ppo.compute_gae(torch.tensor(rew), torch.tensor(val), torch.tensor(done), torch.tensor(last), 0.99, 0.95)
    a_np, r_np = _gae_numpy(rew, val, done, last, 0.99, 0.95)
    assert np.allclose(a.numpy(), a_np, atol=1e-12) and np.allclose(r.numpy(), r_np, atol=1e-12)
    # an episode end cuts the recursion: the advantage at a done step does not see later rewards
    rew2 = rew.copy(); rew2[10:, 0] += 100.0
    done2 = done.copy(); done2[9, 0] = 1.0
    a2, _ = ppo.compute_gae(torch.tensor(rew2), torch.tensor(val), torch.tensor(done2), torch.tensor(last), 0.99, 0.95)
    a1, _ = ppo.compute_gae(torch.tensor(rew), torch.tensor(val), torch.tensor(done2), torch.tensor(last), 0.99, 0.95)
    assert np.allclose(a2[:10, 0].numpy(), a1[:10, 0].numpy())


def test_truncation_bootstrap():
    """collect_rollouts: reward += gamma * V(terminal_observation) where TimeLimit.truncated, nowhere else."""
    rew = torch.tensor([1.0, 2.0, 3.0]); trunc = torch.tensor([0, 1, 0], dtype=torch.uint8); tv = torch.tensor([10.0, 20.0, 30.0])
    out = ppo.bootstrap_truncated(rew, trunc.bool(), tv, 0.99)
    assert torch.allclose(out, torch.tensor([1.0, 2.0 + 0.99 * 20.0, 3.0]))


def test_ppo_loss_clipping_and_terms():
    torch.manual_seed(1)
    net = ppo.ActorCritic().double()
    B = 64
    obs = torch.randn(B, 48, dtype=torch.float64); raw = torch.randn(B, 12, dtype=torch.float64)
    with torch.no_grad():
        logp = net.log_prob(obs, raw)
        # reference log-prob: torch.distributions (what SB3's DiagGaussianDistribution wraps)
        d = torch.distributions.Normal(net.mean(obs), net.log_std.exp())
        assert torch.allclose(logp, d.log_prob(raw).sum(-1), atol=1e-10)
    adv = torch.randn(B, dtype=torch.float64); ret = torch.randn(B, dtype=torch.float64)
    old = logp + torch.linspace(-0.6, 0.6, B, dtype=torch.float64)         # ratios from e^-0.6 to e^0.6: both clip sides are active
    loss, diag = ppo.ppo_loss(net, obs, raw, old, adv, ret, clip_range=0.2, vf_coef=0.5, ent_coef=0.0)
    with torch.no_grad():
        a = (adv - adv.mean()) / (adv.std() + 1e-8)
        ratio = (logp - old).exp()
        pl = -torch.min(ratio * a, ratio.clamp(0.8, 1.2) * a).mean()
        vl = ((net.value(obs) - ret) ** 2).mean()
    assert torch.allclose(loss, pl + 0.5 * vl, atol=1e-12)
    assert 0.3 < float(diag["clip_fraction"]) < 0.9
    # a clipped sample contributes no policy gradient: with every ratio far outside the clip range and the sign of the advantage
    # such that the clipped branch is the minimum, d(policy loss)/d(actor) is zero
    old_far = logp - 2.0                                                     # ratio = e^2 > 1.2
    apos = torch.ones(B, dtype=torch.float64)
    loss2, _ = ppo.ppo_loss(net, obs, raw, old_far, apos, ret, normalize_advantage=False, vf_coef=0.0)
    net.zero_grad(); loss2.backward()
    assert net.action_net.weight.grad.abs().max() == 0 and net.log_std.grad.abs().max() == 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_update(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)                          # identical initial weights on every rank, no broadcast
    net = ppo.ActorCritic()
    params = list(net.parameters())
    opt = torch.optim.Adam(params, lr=3e-4, eps=1e-5)
    g = torch.Generator().manual_seed(100 + rank)  # DIFFERENT data per rank (each rank owns its slab of environments)
    for _ in range(3):
        obs = torch.randn(32, 48, generator=g); raw = torch.randn(32, 12, generator=g)
        adv = torch.randn(32, generator=g); ret = torch.randn(32, generator=g)
        with torch.no_grad():
            old = net.log_prob(obs, raw)
        loss, _ = ppo.ppo_loss(net, obs, raw, old, adv, ret)
        opt.zero_grad(set_to_none=False); loss.backward()
        nflat = ppo.allreduce_grads(params, world)
        torch.nn.utils.clip_grad_norm_(params, 0.5)
        opt.step()
    torch.save({"sd": net.state_dict(), "nflat": nflat}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier(); dist.destroy_process_group()


def test_gradient_bucket_keeps_two_ranks_identical(tmp_path):
    mp.spawn(_rank_update, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    a = torch.load(tmp_path / "rank0.pt"); b = torch.load(tmp_path / "rank1.pt")
    assert a["nflat"] == b["nflat"] == 15449
    for k in a["sd"]:
        assert torch.equal(a["sd"][k], b["sd"][k]), k
    torch.manual_seed(0)
    init = ppo.ActorCritic().state_dict()
    assert any(not torch.equal(init[k], a["sd"][k]) for k in init)      # the weights did move


def test_sb3_zip_round_trip(tmp_path):
    torch.manual_seed(3)
    net = ppo.ActorCritic()
    opt = torch.optim.Adam(net.parameters(), lr=1e-3, eps=1e-5)
    net.value(torch.randn(4, 48)).sum().backward(); opt.step()
    path = ppo.save_sb3_zip(str(tmp_path / "model"), net, opt, dict(num_timesteps=1234, n_envs=16384, learning_rate=1e-3))
    with zipfile.ZipFile(path) as zf:
        assert {"data", "policy.pth", "policy.optimizer.pth", "pytorch_variables.pth", "_stable_baselines3_version", "system_info.txt"} <= set(zf.namelist())
        sd_raw = torch.load(io.BytesIO(zf.read("policy.pth")), weights_only=True)
    assert set(sd_raw) == SB3_KEYS
    sd, data, ostate = ppo.load_sb3_zip(path)
    assert data["num_timesteps"] == 1234 and data["n_envs"] == 16384 and ostate is not None
    net2 = ppo.actor_critic_from_state_dict(sd)
    for k, v in net.state_dict().items():
        assert torch.equal(v, net2.state_dict()[k]), k
    bad = dict(sd); bad["mlp_extractor.policy_net.0.weight"] = torch.zeros(64, 45)      # the reference's 45-d zips
    with pytest.raises(ValueError):
        ppo.actor_critic_from_state_dict(bad)


def test_shipped_final_model_zip_loads():
    """``PPO.load(path)`` of the reference (train.py:112-119) on a shipped zip: the tensors are the ones exported to
    tests/golden/policy_pos_0428.npz bit for bit, and the actor reproduces the fixture's actions."""
    ref = os.environ.get("GO1_REFERENCE_ROOT", "/root/reference")
    z = os.path.join(ref, "models/2024-04-28_22-10-59=1_pos_ctrl_20mil_iter_walking_with_normal_steps/final_model.zip")
    if not os.path.exists(z):
        pytest.skip("the reference's models/ are not on this machine")
    sd, data, _ = ppo.load_sb3_zip(z)
    net = ppo.actor_critic_from_state_dict(sd)
    w = dict(np.load(os.path.join(pc.GOLDEN, "policy_pos_0428.npz")))
    pn = net.mlp_extractor.policy_net
    for t, k in ((pn[0].weight, "w0"), (pn[0].bias, "b0"), (pn[2].weight, "w1"), (pn[2].bias, "b1"), (net.action_net.weight, "w2"),
                 (net.action_net.bias, "b2"), (net.log_std, "log_std")):
        assert np.array_equal(t.detach().numpy(), w[k]), k
    obs = np.random.default_rng(0).normal(size=(50, 48)).astype(np.float32)
    h = np.tanh(obs @ w["w0"].T + w["b0"]); h = np.tanh(h @ w["w1"].T + w["b1"]); mu = h @ w["w2"].T + w["b2"]
    with torch.no_grad():
        assert np.abs(net.mean(torch.from_numpy(obs)).numpy() - mu).max() < 2e-6
    assert isinstance(data, dict) and int(data.get("num_timesteps", 0)) > 1_000_000


@pytest.mark.gpu
def test_rollout_buffers_belong_to_their_observations(cuda_lib_f32):
    """What collect_rollouts stores at step t (values, unclipped samples) must be what the kernels computed FROM the observations of
    step t.  (The library reads a NULL stream as "the handle's own stream"; a rollout that passes torch's default stream stores the
    outputs of the PREVIOUS step, and training silently degenerates.)"""
    from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv
    env = Go1CudaVecEnv(4096, ctrl_type="position", library=cuda_lib_f32, seed=3)
    m = ppo.DevicePPO(env, n_steps=6, seed=0)
    with torch.no_grad():
        m.net.action_net.weight.mul_(60.0)                  # means that move with the observation (the SB3 init has gain 0.01)
        m.net.log_std.fill_(-1.0)
    batch, _ = m.collect_rollouts()
    torch.cuda.synchronize()
    B = m.buf
    with torch.no_grad():
        for t in range(6):
            v = m.net.value(B["obs"][t])
            assert (B["val"][t] - v).abs().max().item() < 2e-2, t                          # tf32 forward vs fp32
            eps = (B["raw"][t] - m.net.mean(B["obs"][t])) / m.net.log_std.exp()
            assert abs(eps.std().item() - 1.0) < 0.05 and abs(eps.mean().item()) < 0.02, (t, eps.std().item())
        # two consecutive steps see different observations, so a one-step lag would show in both checks
        assert (m.net.value(B["obs"][1]) - m.net.value(B["obs"][0])).abs().mean().item() > 0.05
        assert (m.net.mean(B["obs"][1]) - m.net.mean(B["obs"][0])).abs().mean().item() > 0.05
    env.close()


@pytest.mark.gpu
def test_device_ppo_learns_and_round_trips(cuda_lib_f32, tmp_path):
    """A few iterations of the whole loop on the GPU (tcgen05 policy forward in the rollout), then save / load / predict."""
    from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv
    env = Go1CudaVecEnv(2048, ctrl_type="position", library=cuda_lib_f32, seed=1, desired_velocity_min=(0.5, 0, 0), desired_velocity_max=(0.5, 0, 0))
    m = ppo.DevicePPO(env, n_steps=16, n_minibatches=4, n_epochs=4, seed=0)
    m.learn(3 * 16 * 2048, log=False)
    h = m.history
    assert len(h) == 3 and h[-1]["timesteps"] == 3 * 16 * 2048
    assert all(np.isfinite(r["policy_loss"]) and np.isfinite(r["value_loss"]) for r in h)
    assert 0 <= h[0]["approx_kl"] < 0.1 and h[0]["clip_fraction"] < 0.5
    # every update lowers the critic's error on the batch it was computed from (the returns themselves grow from iteration to
    # iteration while the episodes get longer, so losses of different iterations are not comparable)
    assert all(r["value_loss_after"] < r["value_loss_before"] for r in h)
    path = m.save(str(tmp_path / "ppo_b200"))
    m2 = ppo.DevicePPO.load(path, env)
    obs = np.random.default_rng(0).normal(size=(8, 48))
    a1, _ = m.predict(obs); a2, _ = m2.predict(obs)
    assert np.array_equal(a1, a2) and m2.num_timesteps == m.num_timesteps
    # the tensor-core rollout and the plain PyTorch rollout drive the same update rule: both leave finite, moved weights
    m3 = ppo.DevicePPO(env, n_steps=8, n_minibatches=2, n_epochs=2, seed=0, use_tensor_core_policy=False)
    m3._started = False
    m3.learn(8 * 2048, log=False)
    assert np.isfinite(m3.history[-1]["policy_loss"])
    env.close()
