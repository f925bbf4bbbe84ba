"""PPO on the CUDA Go1 environments, device-resident: the loop of the reference's ``train.py:110-136``
(``PPO("MlpPolicy", vec_env, learning_rate=...).learn(total_timesteps)``) with the rollout on the GPU.

    python examples/train_ppo.py --num_parallel_envs 4096 --ctrl_type position --total_timesteps 5000000
    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_ppo.py --num_parallel_envs 16384 ...

What is kept from SB3 2.3.0's PPO defaults (the reference passes only ``learning_rate``): MlpPolicy = separate
48-64-64 tanh actor / critic, state-independent ``log_std`` (init 0), orthogonal init (gains sqrt2 / 0.01 / 1),
gamma 0.99, GAE lambda 0.95, clip 0.2, 10 epochs, advantage normalisation per minibatch, vf_coef 0.5,
max_grad_norm 0.5, value bootstrap through ``terminal_observation`` on time-limit truncation.  What changes with
16k environments: ``n_steps`` is 16-32 instead of 2048 and the minibatch is large; both are flags.

Environments are sharded one slab per rank (no data-path collective).  NCCL is used for exactly two things
(SURVEY 8(e)): the flat gradient bucket (15 449 floats) every optimizer step and the rollout statistics."""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from quadruped_rl_locomotion_b200.lib import Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402


def mlp(i, o, gain):
    layers = [nn.Linear(i, 64), nn.Tanh(), nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, o)]
    for k, l in enumerate(x for x in layers if isinstance(x, nn.Linear)):
        nn.init.orthogonal_(l.weight, gain=gain if k == 2 else 2 ** 0.5)
        nn.init.zeros_(l.bias)
    return nn.Sequential(*layers)


class ActorCritic(nn.Module):
    def __init__(self):
        super().__init__()
        self.pi, self.vf = mlp(48, 12, 0.01), mlp(48, 1, 1.0)
        self.log_std = nn.Parameter(torch.zeros(12))

    def dist(self, obs):
        return torch.distributions.Normal(self.pi(obs), self.log_std.exp())


def allreduce_grads(params, world):
    """One flat bucket, SUM then / world (what DDP would do; 61.8 KB, latency bound on NVLink)."""
    flat = torch.cat([p.grad.reshape(-1) for p in params])
    dist.all_reduce(flat)
    flat /= world
    o = 0
    for p in params:
        n = p.numel(); p.grad.copy_(flat[o:o + n].view_as(p)); o += n


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--num_parallel_envs", type=int, default=4096)          # per rank (train.py: --num_parallel_envs)
    ap.add_argument("--ctrl_type", default="position", choices=["torque", "position"])
    ap.add_argument("--total_timesteps", type=int, default=2_000_000)
    ap.add_argument("--learning_rate", type=float, default=3e-4)
    ap.add_argument("--n_steps", type=int, default=24)
    ap.add_argument("--n_minibatches", type=int, default=4)
    ap.add_argument("--n_epochs", type=int, default=10)
    ap.add_argument("--rand_power", type=float, default=0.0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--fixed_velocity", type=float, default=0.5)           # the "x=1" target (scaled obs slot = 1)
    ap.add_argument("--save", default=None)
    ap.add_argument("--biped", action="store_true")                          # train.py:192-196
    args = ap.parse_args(argv)

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(args.seed)        # same initial weights on every rank (no broadcast needed)
    n = args.num_parallel_envs
    lib = Go1Lib()
    cfg = lib.default_config()
    cfg.rand_power = args.rand_power
    cfg.biped = 1 if args.biped else 0
    cfg.desired_vel_min[0] = cfg.desired_vel_max[0] = args.fixed_velocity
    sim = Sim(load_model_blob(args.ctrl_type), n, lib=lib, cfg=cfg, device=local, seed=args.seed, env_id_offset=rank * n)
    lo, hi = (torch.from_numpy(x).to(dev) for x in sim.action_bounds())
    net = ActorCritic().to(dev)
    params = list(net.parameters())
    opt = torch.optim.Adam(params, lr=args.learning_rate, eps=1e-5)
    T = args.n_steps
    f = dict(device=dev, dtype=torch.float32)
    obs = torch.zeros(n, 48, **f); nobs = torch.zeros(n, 48, **f); tobs = torch.zeros(n, 48, **f)
    rew = torch.zeros(n, **f); info = torch.zeros(n, 16, **f)
    done = torch.zeros(n, dtype=torch.uint8, device=dev); trunc = torch.zeros(n, dtype=torch.uint8, device=dev)
    B = dict(obs=torch.zeros(T, n, 48, **f), act=torch.zeros(T, n, 12, **f), logp=torch.zeros(T, n, **f),
             rew=torch.zeros(T, n, **f), val=torch.zeros(T, n, **f), done=torch.zeros(T, n, **f))
    stream = torch.cuda.current_stream().cuda_stream
    sim.reset(None, obs, stream)
    steps_done, it, t0 = 0, 0, time.time()
    ep_ret_sum = torch.zeros((), **f); ep_cnt = torch.zeros((), **f)
    history = []
    while steps_done < args.total_timesteps:
        with torch.no_grad():
            for t in range(T):
                d = net.dist(obs)
                a = d.sample()
                B["obs"][t] = obs; B["act"][t] = a; B["logp"][t] = d.log_prob(a).sum(-1); B["val"][t] = net.vf(obs).squeeze(-1)
                sim.step(torch.minimum(torch.maximum(a, lo), hi).contiguous(), nobs, rew, done, trunc, tobs, info, stream)
                r = rew.clone()
                tr = trunc.bool()
                if tr.any():      # SB3: bootstrap with the value of the terminal observation on time-limit truncation
                    r[tr] += 0.99 * net.vf(tobs[tr]).squeeze(-1)
                dn = done.bool()
                ep_ret_sum += info[dn, 10].sum(); ep_cnt += dn.sum()
                B["rew"][t] = r; B["done"][t] = dn.float()
                obs, nobs = nobs, obs
            last_val = net.vf(obs).squeeze(-1)
            adv = torch.zeros(T, n, **f); gae = torch.zeros(n, **f)
            for t in reversed(range(T)):
                nv = last_val if t == T - 1 else B["val"][t + 1]
                nonterm = 1.0 - B["done"][t]
                delta = B["rew"][t] + 0.99 * nv * nonterm - B["val"][t]
                gae = delta + 0.99 * 0.95 * nonterm * gae
                adv[t] = gae
            ret = adv + B["val"]
        flat = {k: v.reshape(T * n, *v.shape[2:]) for k, v in B.items()}
        fadv, fret = adv.reshape(-1), ret.reshape(-1)
        mb = T * n // args.n_minibatches
        for _ in range(args.n_epochs):
            perm = torch.randperm(T * n, device=dev)
            for k in range(args.n_minibatches):
                idx = perm[k * mb:(k + 1) * mb]
                d = net.dist(flat["obs"][idx])
                logp = d.log_prob(flat["act"][idx]).sum(-1)
                a_ = fadv[idx]; a_ = (a_ - a_.mean()) / (a_.std() + 1e-8)
                log_ratio = logp - flat["logp"][idx]
                ratio = log_ratio.clamp(-20.0, 20.0).exp()
                pl = -torch.min(ratio * a_, ratio.clamp(0.8, 1.2) * a_).mean()
                vl = (net.vf(flat["obs"][idx]).squeeze(-1) - fret[idx]).pow(2).mean()
                opt.zero_grad(set_to_none=False)
                (pl + 0.5 * vl).backward()
                last = (pl.detach(), vl.detach(), (-log_ratio.detach()).mean())
                if world > 1:
                    allreduce_grads(params, world)
                nn.utils.clip_grad_norm_(params, 0.5)
                opt.step()
        steps_done += T * n * world
        it += 1
        stats = torch.stack([ep_ret_sum, ep_cnt, B["rew"].sum(), torch.tensor(float(T * n), **f)])
        if world > 1:
            dist.all_reduce(stats)
        ep_ret_sum.zero_(); ep_cnt.zero_()
        s = stats.tolist()
        rec = dict(iteration=it, timesteps=steps_done, mean_step_reward=s[2] / s[3], episodes=int(s[1]),
                   mean_episode_return=(s[0] / s[1]) if s[1] > 0 else None, fps=steps_done / (time.time() - t0),
                   policy_loss=float(last[0]), value_loss=float(last[1]), approx_kl=float(last[2]), std=float(net.log_std.exp().mean()))
        history.append(rec)
        if rank == 0:
            print(rec, flush=True)
    if rank == 0 and args.save:
        torch.save({"policy": net.state_dict(), "history": history}, args.save)
    sim.close()
    if world > 1:
        dist.destroy_process_group()
    return history, net


if __name__ == "__main__":
    main()
