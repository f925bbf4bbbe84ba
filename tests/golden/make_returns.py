"""Generates ``tests/golden/returns_<tag>.npz``: episode returns / lengths of the reference's SHIPPED 48-d
position-control policies (``models/**/final_model.zip``) in the reference's own environment code
(unmodified ``go1_mujoco_env.py``, imported from /root/reference) on the fp64 oracle physics.

    python tests/golden/make_returns.py [n_episodes=1000] [n_procs=8] [max_steps=1500] [tag,tag,... = all] [first_seed=1000]

With a first_seed other than 1000 the new episodes are APPENDED to the existing fixture (disjoint seeds), which is how the
4000-episode files were extended to 12 000 for the policies whose first comparison sat near the 2 % bar.

north_star's third correctness criterion: the CUDA simulator must reproduce the MEAN episode return of these
policies within 2 % (trajectories are chaotic, so only the statistics are comparable).  Episodes are capped at
``max_steps`` control steps (30 s of simulated time at the default) to bound the CPU cost of the reference arm;
the GPU test applies the same cap.  Evaluation protocol: deterministic actions (the
Gaussian mean), reset joint noise ``rand_power = 1`` (the value the reference's curriculum callback ramps to), clipped to the action space like SB3's ``predict``; desired velocity fixed to (0.5, 0, 0) (the
"x=1" target: scaled observation slot [9:12] = (1, 0, 0)); the 20 % fallen-start branch stays on.

Also exports the policy weights to ``tests/golden/policy_<tag>.npz`` (plain arrays, so the GPU box needs
neither the zip nor stable-baselines3).
"""
from __future__ import annotations

import io
import multiprocessing as mp
import os
import sys
import zipfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")

POLICIES = {
    "pos_0428": "models/2024-04-28_22-10-59=1_pos_ctrl_20mil_iter_walking_with_normal_steps/final_model.zip",
    "pos_0421": "models/2024-04-21_20-04-20=1_pso_ctrl_30mil_iter_walking_with_oriented_rear_legs/final_model.zip",
    "pos_0427": "models/2024-04-27_18-04-12=1_pos_ctrl_20mil_iter_walking_with_fast_steps/final_model.zip",
    "pos_0428_best": "models/2024-04-28_22-10-59=1_pos_ctrl_20mil_iter_walking_with_normal_steps/best_model.zip",
    # not a reference artefact: trained HERE by examples/train_ppo.py on the CUDA environments (200 M steps, 16384 envs, one B200,
    # position control, rand_power 1); policy_ppo_b200.npz is committed.  Evaluated in the reference env like the shipped ones
    # ("eval of the trained policy in both simulators").
    "ppo_b200": None,
    # the same in round 2: DevicePPO, 240 M steps, position control, rand_power 0, on the simulator WITH robot-robot contacts
    # (profiles/r2_ppo_700M_steps_16384envs.log); policy_ppo_b200_r2.npz is committed
    "ppo_b200_r2": None,
}
V_DES = (0.5, 0.0, 0.0)
RAND_POWER = 1.0     # reset joint noise N(0, 0.1 rad) (go1_mujoco_env.py:974-982): without it every upright episode of a
                     # deterministic policy is the SAME trajectory and the "mean" would be one sample of a chaotic system


def load_policy(zip_path: str) -> dict:
    import torch
    with zipfile.ZipFile(zip_path) as zf:
        sd = torch.load(io.BytesIO(zf.read("policy.pth")), map_location="cpu", weights_only=True)
    g = lambda k: sd[k].numpy().astype(np.float32)
    return dict(w0=g("mlp_extractor.policy_net.0.weight"), b0=g("mlp_extractor.policy_net.0.bias"),
                w1=g("mlp_extractor.policy_net.2.weight"), b1=g("mlp_extractor.policy_net.2.bias"),
                w2=g("action_net.weight"), b2=g("action_net.bias"), log_std=g("log_std"))


def act(pol: dict, obs: np.ndarray, lo: np.ndarray, hi: np.ndarray) -> np.ndarray:
    """SB3 ActorCriticPolicy.predict(deterministic=True): fp32 MLP, tanh, mean action, clip to the Box."""
    x = obs.astype(np.float32)
    h = np.tanh(pol["w0"] @ x + pol["b0"])
    h = np.tanh(pol["w1"] @ h + pol["b1"])
    return np.clip(pol["w2"] @ h + pol["b2"], lo, hi).astype(np.float32)


def _worker(args):
    tag, seeds, max_steps = args
    from oracle.ref_env import make_reference_env
    pol = dict(np.load(os.path.join(OUT, f"policy_{tag}.npz")))
    env = make_reference_env("position", rand_power=RAND_POWER)
    env._desired_velocity_min = np.array(V_DES); env._desired_velocity_max = np.array(V_DES)
    lo, hi = env.action_space.low, env.action_space.high
    out = []
    for s in seeds:
        np.random.seed(s)                       # reset_model draws from the global numpy RNG (go1_mujoco_env.py:953-959)
        obs, _ = env.reset()
        fallen = float(abs(env.data.qpos[2] - 0.1) < 1e-9)
        ret, n, term = 0.0, 0, False
        for _ in range(max_steps):
            obs, r, term, trunc, info = env.step(act(pol, obs, lo, hi))
            ret += r; n += 1
            if term or trunc:
                break
        out.append((s, ret, n, fallen, float(term), float(env.data.qpos[0])))
    return out


def main():
    n_ep = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    procs = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    max_steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1500
    only = sys.argv[4].split(",") if len(sys.argv) > 4 else list(POLICIES)
    seed0 = int(sys.argv[5]) if len(sys.argv) > 5 else 1000
    for tag, rel in POLICIES.items():
        if tag not in only or rel is None:
            continue
        pol = load_policy(os.path.join("/root/reference", rel))
        assert pol["w0"].shape == (64, 48) and pol["w2"].shape == (12, 64)
        np.savez(os.path.join(OUT, f"policy_{tag}.npz"), **pol)
    for tag in only:
        seeds = np.arange(n_ep) + seed0
        chunks = [(tag, list(map(int, c)), max_steps) for c in np.array_split(seeds, procs * 4)]
        with mp.Pool(procs) as pool:
            res = [r for part in pool.map(_worker, chunks) for r in part]
        a = np.array(sorted(res))
        path = os.path.join(OUT, f"returns_{tag}.npz")
        if seed0 != 1000 and os.path.exists(path):
            old = np.load(path)
            assert int(old["max_steps"]) == max_steps and not set(old["seed"].astype(int)) & set(seeds.tolist())
            a = np.concatenate([np.stack([old[k] for k in ("seed", "ret", "length", "fallen", "terminated", "x_final")], axis=1), a])
        np.savez(os.path.join(OUT, f"returns_{tag}.npz"), seed=a[:, 0], ret=a[:, 1], length=a[:, 2], fallen=a[:, 3],
                 terminated=a[:, 4], x_final=a[:, 5], max_steps=max_steps, v_des=np.array(V_DES), rand_power=RAND_POWER)
        print(f"{tag}: {len(a)} episodes, mean return {a[:, 1].mean():.3f} +- {a[:, 1].std() / np.sqrt(len(a)):.3f} (s.e.), "
              f"mean length {a[:, 2].mean():.1f}, fallen starts {a[:, 3].mean():.3f}, mean x_final {a[:, 5].mean():.2f}", flush=True)


if __name__ == "__main__":
    main()
