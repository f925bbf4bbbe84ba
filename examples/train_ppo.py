"""Command-line front end of ``quadruped_rl_locomotion_b200.ppo.DevicePPO``: the reference's ``train.py --run train``
(``train.py:110-138``) with the rollout on the GPU.

    python examples/train_ppo.py --num_parallel_envs 4096 --ctrl_type position --total_timesteps 5000000 --save runs/ppo
    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 examples/train_ppo.py --num_parallel_envs 16384 ...

Flags follow the reference's (``--num_parallel_envs``, ``--ctrl_type``, ``--total_timesteps``, ``--learning_rate``, ``--biped``,
``--model_path`` to continue from a zip).  ``--save`` writes stable-baselines3's zip layout plus the reference's
``training_history.json`` (``quadruped_rl_locomotion_b200.callbacks``)."""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from quadruped_rl_locomotion_b200.ppo import DevicePPO  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv  # noqa: E402


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
    ap.add_argument("--model_path", default=None)                            # train.py:112-119: continue from a zip
    ap.add_argument("--biped", action="store_true")                          # train.py:192-196
    ap.add_argument("--no_tensor_core_policy", action="store_true")
    args = ap.parse_args(argv)

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.num_parallel_envs
    env = Go1CudaVecEnv(n, ctrl_type=args.ctrl_type, biped=args.biped, rand_power=args.rand_power, device=local, seed=args.seed,
                        env_id_offset=rank * n, desired_velocity_min=(args.fixed_velocity, 0, 0), desired_velocity_max=(args.fixed_velocity, 0, 0))
    kw = dict(n_steps=args.n_steps, n_minibatches=args.n_minibatches, n_epochs=args.n_epochs, seed=args.seed,
              use_tensor_core_policy=not args.no_tensor_core_policy)
    model = DevicePPO.load(args.model_path, env, **kw) if args.model_path else DevicePPO(env, learning_rate=args.learning_rate, **kw)
    model.learn(args.total_timesteps)
    if rank == 0 and args.save:
        from quadruped_rl_locomotion_b200.callbacks import TrainingHistory
        path = model.save(os.path.join(args.save, "final_model"))
        hist = TrainingHistory(args.save)
        for r in model.history:
            hist.append(r["timesteps"], r["mean_episode_return"] if r["mean_episode_return"] is not None else float("nan"))
        hist.save()
        print("saved", path)
    env.close()
    if world > 1:
        dist.destroy_process_group()
    return model.history, model.net


if __name__ == "__main__":
    main()
