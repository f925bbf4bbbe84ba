"""Device-resident PPO for the CUDA Go1 environments: the caller side of the hot path.

Replaces what the reference runs at ``train.py:112-138``::

    model = PPO("MlpPolicy", vec_env, learning_rate=..., ...)      # or PPO.load(path, env=vec_env)
    model.learn(total_timesteps=..., callback=...)
    model.save(path)

with the rollout on the GPU: observations, actions, rewards and the rollout buffer never leave device memory, the policy forward
of every control step (actions, values, log-probs for all environments) is ONE tcgen05 kernel (``go1_policy_forward``), the
optimisation is plain PyTorch on the same device.  Kept from SB3 2.3.0's PPO defaults (the reference passes only
``learning_rate``; the shipped zips' ``data`` confirm them): MlpPolicy = separate 64-64 tanh actor / critic, state-independent
``log_std`` (init 0), orthogonal init (gains sqrt 2 / 0.01 / 1), gamma 0.99, GAE lambda 0.95, clip 0.2, 10 epochs, advantage
normalisation per minibatch, vf_coef 0.5, ent_coef 0, max_grad_norm 0.5, Adam eps 1e-5, value bootstrap through the terminal
observation on time-limit truncation.  With 16k environments ``n_steps`` is 16-32 instead of 2048 and minibatches are large.

Multi-GPU: one process + one slab per GPU; NCCL is used for exactly two things: the flat gradient bucket
(15 449 floats) every optimizer step and the rollout statistics.

Files: :func:`save_sb3_zip` / :func:`load_sb3_zip` write and read stable-baselines3's zip layout (``policy.pth`` with SB3's
parameter names, ``policy.optimizer.pth``, ``pytorch_variables.pth``, ``data`` JSON, ``_stable_baselines3_version``,
``system_info.txt``), so a shipped ``final_model.zip`` loads here and a zip written here has the tensors where SB3 looks for them.
"""
from __future__ import annotations

import io
import json
import os
import time
import zipfile
from typing import Callable, Optional

import numpy as np
import torch
import torch.nn as nn

OBS_DIM, ACT_DIM, HID = 48, 12, 64


# ------------------------------------------------------------------------------------------------ network (SB3 MlpPolicy names)
class _Extractor(nn.Module):
    def __init__(self):
        super().__init__()
        self.policy_net = nn.Sequential(nn.Linear(OBS_DIM, HID), nn.Tanh(), nn.Linear(HID, HID), nn.Tanh())
        self.value_net = nn.Sequential(nn.Linear(OBS_DIM, HID), nn.Tanh(), nn.Linear(HID, HID), nn.Tanh())


class ActorCritic(nn.Module):
    """SB3 ``ActorCriticPolicy`` with ``net_arch=dict(pi=[64, 64], vf=[64, 64])``, tanh: the ``state_dict`` keys are SB3's
    (``mlp_extractor.policy_net.0.weight`` ... ``action_net.weight``, ``value_net.weight``, ``log_std``)."""

    def __init__(self, log_std_init: float = 0.0):
        super().__init__()
        self.mlp_extractor = _Extractor()
        self.action_net = nn.Linear(HID, ACT_DIM)
        self.value_net = nn.Linear(HID, 1)
        self.log_std = nn.Parameter(torch.full((ACT_DIM,), float(log_std_init)))
        for seq in (self.mlp_extractor.policy_net, self.mlp_extractor.value_net):
            for m in seq:
                if isinstance(m, nn.Linear):
                    nn.init.orthogonal_(m.weight, gain=2 ** 0.5); nn.init.zeros_(m.bias)
        nn.init.orthogonal_(self.action_net.weight, gain=0.01); nn.init.zeros_(self.action_net.bias)
        nn.init.orthogonal_(self.value_net.weight, gain=1.0); nn.init.zeros_(self.value_net.bias)

    def mean(self, obs):
        return self.action_net(self.mlp_extractor.policy_net(obs))

    def value(self, obs):
        return self.value_net(self.mlp_extractor.value_net(obs)).squeeze(-1)

    def log_prob(self, obs, raw_actions):
        """log N(raw_action; mean(obs), exp(log_std)) summed over the action dims (SB3 DiagGaussianDistribution.log_prob)."""
        mu = self.mean(obs)
        return (-0.5 * ((raw_actions - mu) / self.log_std.exp()) ** 2 - self.log_std - 0.5 * np.log(2 * np.pi)).sum(-1)

    def kernel_pointers(self):
        """(Policy, Critic) structs of device pointers for ``go1_policy_forward`` (weights must be contiguous fp32 CUDA tensors)."""
        from .lib import Critic, Policy
        pn, vn = self.mlp_extractor.policy_net, self.mlp_extractor.value_net
        ts = [pn[0].weight, pn[0].bias, pn[2].weight, pn[2].bias, self.action_net.weight, self.action_net.bias, self.log_std,
              vn[0].weight, vn[0].bias, vn[2].weight, vn[2].bias, self.value_net.weight, self.value_net.bias]
        for t in ts:
            assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
        p = [t.data_ptr() for t in ts]
        return Policy(*p[:7]), Critic(*p[7:])


# ------------------------------------------------------------------------------------------------ pure functions (CPU-testable)
def bootstrap_truncated(rew, trunc, terminal_values, gamma: float):
    """SB3 ``collect_rollouts``: a time-limit truncation is not a failure, the reward gets ``gamma * V(terminal_observation)``."""
    return rew + gamma * terminal_values * trunc.to(rew.dtype)


def compute_gae(rew, val, done, last_val, gamma: float = 0.99, lam: float = 0.95):
    """SB3 ``RolloutBuffer.compute_returns_and_advantage`` on [T, N] tensors; ``done[t]`` = the episode ended AT step t (the
    observation of step t + 1 belongs to a new episode).  Returns (advantages, returns)."""
    T = rew.shape[0]
    adv = torch.zeros_like(rew)
    gae = torch.zeros_like(rew[0])
    for t in reversed(range(T)):
        nv = last_val if t == T - 1 else val[t + 1]
        nonterm = 1.0 - done[t]
        delta = rew[t] + gamma * nv * nonterm - val[t]
        gae = delta + gamma * lam * nonterm * gae
        adv[t] = gae
    return adv, adv + val


def ppo_loss(net: ActorCritic, obs, raw_act, old_logp, adv, ret, clip_range: float = 0.2, vf_coef: float = 0.5, ent_coef: float = 0.0,
             normalize_advantage: bool = True):
    """SB3 ``PPO.train`` inner loss for one minibatch.  Returns (loss, dict of detached diagnostics)."""
    if normalize_advantage and adv.numel() > 1:
        adv = (adv - adv.mean()) / (adv.std() + 1e-8)
    logp = net.log_prob(obs, raw_act)
    log_ratio = logp - old_logp
    ratio = log_ratio.clamp(-20.0, 20.0).exp()
    pl = -torch.min(ratio * adv, ratio.clamp(1 - clip_range, 1 + clip_range) * adv).mean()
    vl = (net.value(obs) - ret).pow(2).mean()
    ent = (0.5 + 0.5 * np.log(2 * np.pi) + net.log_std).sum()
    loss = pl + vf_coef * vl - ent_coef * ent
    with torch.no_grad():
        diag = dict(policy_loss=pl.detach(), value_loss=vl.detach(), approx_kl=((ratio - 1) - log_ratio).mean(),
                    clip_fraction=((ratio - 1).abs() > clip_range).float().mean())
    return loss, diag


def allreduce_grads(params, world: int):
    """One flat bucket (15 449 floats = 61.8 KB for this policy), SUM then / world: every rank applies the same update, so
    identical initial weights stay identical without broadcasts."""
    import torch.distributed as dist
    flat = torch.cat([p.grad.reshape(-1) for p in params])
    dist.all_reduce(flat)
    flat /= world
    o = 0
    for p in params:
        n = p.numel(); p.grad.copy_(flat[o:o + n].view_as(p)); o += n
    return flat.numel()


# ------------------------------------------------------------------------------------------------ SB3 zip layout
SB3_VERSION = "2.3.0"


def save_sb3_zip(path: str, net: ActorCritic, optimizer: Optional[torch.optim.Optimizer] = None, data: Optional[dict] = None) -> str:
    """Writes stable-baselines3's model zip (``BaseAlgorithm.save`` -> ``save_to_zip_file``): ``data`` (JSON), ``policy.pth``,
    ``policy.optimizer.pth``, ``pytorch_variables.pth``, ``_stable_baselines3_version``, ``system_info.txt``.  ``data`` holds the
    plain hyper-parameters; SB3 additionally stores cloudpickled objects (spaces, schedules) there, which need SB3 itself to
    produce -- ``load_sb3_zip`` reads both kinds."""
    if not path.endswith(".zip"):
        path += ".zip"
    d = dict(policy_class="ActorCriticPolicy", observation_space=dict(shape=[OBS_DIM], dtype="float64", low=None, high=None),
             action_space=dict(shape=[ACT_DIM], dtype="float32"), n_envs=None, num_timesteps=0, gamma=0.99, gae_lambda=0.95, n_steps=None,
             batch_size=None, n_epochs=10, ent_coef=0.0, vf_coef=0.5, max_grad_norm=0.5, clip_range=0.2, learning_rate=None,
             policy_kwargs={}, writer="quadruped_rl_locomotion_b200.ppo")
    d.update(data or {})
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with zipfile.ZipFile(path, "w") as zf:
        zf.writestr("data", json.dumps(d, indent=4, default=lambda o: np.asarray(o).tolist()))
        buf = io.BytesIO(); torch.save({k: v.detach().cpu() for k, v in net.state_dict().items()}, buf); zf.writestr("policy.pth", buf.getvalue())
        buf = io.BytesIO(); torch.save(optimizer.state_dict() if optimizer is not None else {}, buf); zf.writestr("policy.optimizer.pth", buf.getvalue())
        buf = io.BytesIO(); torch.save({}, buf); zf.writestr("pytorch_variables.pth", buf.getvalue())
        zf.writestr("_stable_baselines3_version", SB3_VERSION)
        zf.writestr("system_info.txt", f"- writer: quadruped_rl_locomotion_b200.ppo\n- PyTorch: {torch.__version__}\n- Numpy: {np.__version__}\n")
    return path


def load_sb3_zip(path: str, device: str = "cpu"):
    """Reads a model zip written by stable-baselines3 (e.g. the reference's ``models/**/final_model.zip``) or by
    :func:`save_sb3_zip`.  Returns (state_dict, data dict, optimizer state or None).  SB3's pickled ``data`` entries (dicts with a
    ``":serialized:"`` key) are left as they are: only the plain values are used here."""
    with zipfile.ZipFile(path) as zf:
        names = set(zf.namelist())
        sd = torch.load(io.BytesIO(zf.read("policy.pth")), map_location=device, weights_only=True)
        data = json.loads(zf.read("data").decode()) if "data" in names else {}
        opt = None
        if "policy.optimizer.pth" in names:
            try:
                opt = torch.load(io.BytesIO(zf.read("policy.optimizer.pth")), map_location=device, weights_only=True)
            except Exception:  # noqa: BLE001 - an optimizer state is optional for evaluation
                opt = None
    return sd, data, opt


def actor_critic_from_state_dict(sd: dict, device="cpu") -> ActorCritic:
    net = ActorCritic()
    missing = set(net.state_dict()) - set(sd)
    if missing:
        raise KeyError(f"policy.pth lacks {sorted(missing)}: not a 48-64-64 MlpPolicy")
    shapes = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    for k, shp in shapes.items():
        if tuple(sd[k].shape) != shp:
            raise ValueError(f"{k}: shape {tuple(sd[k].shape)} != {shp} (HEAD's environment has 48-d observations; the reference's "
                             f"45-d zips cannot drive it, SURVEY appendix E)")
    net.load_state_dict({k: sd[k] for k in shapes})
    return net.to(device)


# ------------------------------------------------------------------------------------------------ the algorithm
class DevicePPO:
    """``PPO("MlpPolicy", env, learning_rate=...)`` / ``.learn`` / ``.save`` / ``.load`` / ``.predict`` of the reference's train.py
    on a :class:`~quadruped_rl_locomotion_b200.vec_env.Go1CudaVecEnv` (its ``sim`` is stepped through device buffers)."""

    def __init__(self, env, learning_rate: float = 3e-4, n_steps: int = 24, n_minibatches: int = 4, n_epochs: int = 10, gamma: float = 0.99,
                 gae_lambda: float = 0.95, clip_range: float = 0.2, vf_coef: float = 0.5, ent_coef: float = 0.0, max_grad_norm: float = 0.5,
                 seed: int = 0, device: Optional[torch.device] = None, use_tensor_core_policy: bool = True):
        self.env, self.sim = env, env.sim
        self.n = env.num_envs
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.hp = dict(learning_rate=learning_rate, n_steps=n_steps, n_minibatches=n_minibatches, n_epochs=n_epochs, gamma=gamma,
                       gae_lambda=gae_lambda, clip_range=clip_range, vf_coef=vf_coef, ent_coef=ent_coef, max_grad_norm=max_grad_norm)
        torch.manual_seed(seed)                       # same initial weights on every rank (no broadcast needed)
        self.net = ActorCritic().to(self.device)
        self.params = list(self.net.parameters())
        self.opt = torch.optim.Adam(self.params, lr=learning_rate, eps=1e-5)
        self.num_timesteps = 0
        self.use_tc = use_tensor_core_policy
        self.history = []
        import torch.distributed as dist
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        n, T, f = self.n, n_steps, dict(device=self.device, dtype=torch.float32)
        self.obs = torch.zeros(n, OBS_DIM, **f); self.nobs = torch.zeros(n, OBS_DIM, **f); self.tobs = torch.zeros(n, OBS_DIM, **f)
        self.rew = torch.zeros(n, **f); self.info = torch.zeros(n, 20, **f)
        self.done = torch.zeros(n, dtype=torch.uint8, device=self.device); self.trunc = torch.zeros(n, dtype=torch.uint8, device=self.device)
        self.act = torch.zeros(n, ACT_DIM, **f); self.raw = torch.zeros(n, ACT_DIM, **f); self.val = torch.zeros(n, **f)
        self.buf = dict(obs=torch.zeros(T, n, OBS_DIM, **f), raw=torch.zeros(T, n, ACT_DIM, **f), rew=torch.zeros(T, n, **f),
                        val=torch.zeros(T, n, **f), done=torch.zeros(T, n, **f))
        lo, hi = self.sim.action_bounds()
        self.lo, self.hi = torch.from_numpy(lo).to(self.device), torch.from_numpy(hi).to(self.device)
        self._started = False
        # ONE stream for the whole rollout, the library's kernels and torch's ops alike.  It must not be torch's default stream: the
        # C ABI reads a NULL stream as "the handle's own stream" (non-blocking), on which the policy / step kernels would run
        # unordered with torch's copies of their outputs.
        self.stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None

    # -- policy forward for all environments: one tcgen05 kernel (actions, raw samples, values)
    def _act(self):
        if self.use_tc:
            pol, cri = self.net.kernel_pointers()
            self.sim.policy_forward(pol, cri, self.obs, self.act, self.val, None, stream=self.stream.cuda_stream, raw_act=self.raw)
        else:
            mu = self.net.mean(self.obs)
            self.raw.copy_(mu + self.net.log_std.exp() * torch.randn_like(mu))
            self.act.copy_(torch.minimum(torch.maximum(self.raw, self.lo), self.hi))
            self.val.copy_(self.net.value(self.obs))

    def collect_rollouts(self):
        self.stream.wait_stream(torch.cuda.current_stream(self.device))      # the optimiser's last update is visible to the rollout
        with torch.cuda.stream(self.stream):
            out = self._collect()
        torch.cuda.current_stream(self.device).wait_stream(self.stream)      # ... and the batch to the update
        return out

    def _collect(self):
        hp, B = self.hp, self.buf
        stream = self.stream.cuda_stream
        assert stream != 0
        if not self._started:
            self.sim.reset(None, self.obs, stream)
            self._started = True
        ep_ret = torch.zeros((), device=self.device); ep_cnt = torch.zeros((), device=self.device)
        with torch.no_grad():
            for t in range(hp["n_steps"]):
                self._act()
                B["obs"][t] = self.obs; B["raw"][t] = self.raw; B["val"][t] = self.val
                self.sim.step(self.act, self.nobs, self.rew, self.done, self.trunc, self.tobs, self.info, stream)
                tr = self.trunc.bool()
                r = self.rew
                if tr.any():
                    r = bootstrap_truncated(self.rew, tr, self.net.value(self.tobs) * tr, hp["gamma"])
                dn = self.done.bool()
                ep_ret += self.info[dn, 10].sum(); ep_cnt += dn.sum()
                B["rew"][t] = r; B["done"][t] = dn.float()
                self.obs, self.nobs = self.nobs, self.obs
            last_val = self.net.value(self.obs)
            adv, ret = compute_gae(B["rew"], B["val"], B["done"], last_val, hp["gamma"], hp["gae_lambda"])
            T, n = hp["n_steps"], self.n
            fobs, fraw = B["obs"].reshape(T * n, OBS_DIM), B["raw"].reshape(T * n, ACT_DIM)
            old_logp = self.net.log_prob(fobs, fraw)          # fp32, the same arithmetic as the update: ratio == 1 at the first epoch
        return dict(obs=fobs, raw=fraw, old_logp=old_logp, adv=adv.reshape(-1), ret=ret.reshape(-1)), (ep_ret, ep_cnt, B["rew"].sum())

    def train(self, batch):
        hp = self.hp
        N = batch["obs"].shape[0]
        mb = N // hp["n_minibatches"]
        diag = None
        with torch.no_grad():
            vl_before = ((self.net.value(batch["obs"]) - batch["ret"]) ** 2).mean()
        for _ in range(hp["n_epochs"]):
            perm = torch.randperm(N, device=self.device)
            for k in range(hp["n_minibatches"]):
                idx = perm[k * mb:(k + 1) * mb]
                loss, diag = ppo_loss(self.net, batch["obs"][idx], batch["raw"][idx], batch["old_logp"][idx], batch["adv"][idx], batch["ret"][idx],
                                      hp["clip_range"], hp["vf_coef"], hp["ent_coef"])
                self.opt.zero_grad(set_to_none=False)
                loss.backward()
                if self.world > 1:
                    allreduce_grads(self.params, self.world)
                nn.utils.clip_grad_norm_(self.params, hp["max_grad_norm"])
                self.opt.step()
        with torch.no_grad():       # SB3 logs train/explained_variance the same way (ppo.py:train)
            res = batch["ret"] - self.net.value(batch["obs"])
            diag["value_loss_before"], diag["value_loss_after"] = vl_before, (res ** 2).mean()
            diag["explained_variance"] = 1 - res.var() / batch["ret"].var().clamp_min(1e-12)
        return diag

    def learn(self, total_timesteps: int, callback: Optional[Callable[[dict], bool]] = None, log: bool = True):
        """``callback(record) -> bool`` is called once per iteration (return False to stop), like SB3's ``BaseCallback._on_step``
        at rollout granularity; ``record`` carries the fields the reference's callbacks log (timesteps, mean return, losses)."""
        import torch.distributed as dist
        t0, start = time.time(), self.num_timesteps
        while self.num_timesteps - start < total_timesteps:
            batch, (ep_ret, ep_cnt, rsum) = self.collect_rollouts()
            diag = self.train(batch)
            stats = torch.stack([ep_ret, ep_cnt, rsum, torch.tensor(float(batch["obs"].shape[0]), device=self.device)])
            if self.world > 1:
                dist.all_reduce(stats)
            self.num_timesteps += self.hp["n_steps"] * self.n * self.world
            s = stats.tolist()
            rec = dict(iteration=len(self.history) + 1, timesteps=self.num_timesteps, mean_step_reward=s[2] / s[3], episodes=int(s[1]),
                       mean_episode_return=(s[0] / s[1]) if s[1] > 0 else None, fps=(self.num_timesteps - start) / (time.time() - t0),
                       policy_loss=float(diag["policy_loss"]), value_loss=float(diag["value_loss"]), approx_kl=float(diag["approx_kl"]),
                       clip_fraction=float(diag["clip_fraction"]), std=float(self.net.log_std.detach().exp().mean()),
                       value_loss_before=float(diag["value_loss_before"]), value_loss_after=float(diag["value_loss_after"]),
                       explained_variance=float(diag["explained_variance"]))
            self.history.append(rec)
            if log and self.rank == 0:
                print(rec, flush=True)
            if callback is not None and callback(rec) is False:
                break
        return self

    def predict(self, obs, deterministic: bool = True):
        """SB3 ``predict``: numpy observations in, clipped numpy actions out (the mean when deterministic)."""
        with torch.no_grad():
            o = torch.as_tensor(np.asarray(obs), dtype=torch.float32, device=self.device)
            a = self.net.mean(o)
            if not deterministic:
                a = a + self.net.log_std.exp() * torch.randn_like(a)
            return torch.minimum(torch.maximum(a, self.lo), self.hi).cpu().numpy(), None

    def save(self, path: str) -> str:
        data = dict(self.hp, n_envs=self.n * self.world, num_timesteps=self.num_timesteps, batch_size=self.hp["n_steps"] * self.n // self.hp["n_minibatches"],
                    ctrl_type=getattr(self.env, "ctrl_type", None), biped=getattr(self.env, "biped", None))
        return save_sb3_zip(path, self.net, self.opt, data)

    @classmethod
    def load(cls, path: str, env, **kw) -> "DevicePPO":
        sd, data, opt = load_sb3_zip(path)
        m = cls(env, **{k: data[k] for k in ("learning_rate",) if isinstance(data.get(k), (int, float))}, **kw)
        m.net.load_state_dict({k: v for k, v in sd.items() if k in m.net.state_dict()})
        m.num_timesteps = int(data.get("num_timesteps") or 0)
        if opt:
            try:
                m.opt.load_state_dict(opt)
            except Exception:  # noqa: BLE001 - optimizer states of other layouts (SB3 param groups) are optional
                pass
        return m
