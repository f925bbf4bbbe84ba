"""north_star correctness part 3: the reference's SHIPPED policies reach the same mean episode return in the CUDA
simulator as in the reference environment (most episodes are chaotic, so only statistics are comparable).

Reference arm = ``tests/golden/returns_<tag>.npz``: 12 000 seeded episodes of the unmodified ``go1_mujoco_env.py``
on the fp64 oracle physics (``tests/golden/make_returns.py``; episodes capped at ``max_steps`` control steps,
deterministic actions, desired velocity (0.5, 0, 0), 20 % fallen starts left on, reset joint noise rand_power = 1
so that the upright starts form an ensemble instead of one deterministic trajectory).  Two CUDA arms:

* ``test_shipped_policy_mean_return``: the same protocol through ``Go1CudaVecEnv.step`` (host policy, host buffers) from the
  kernel's OWN resets, 16 384 episodes.  Bar: |mean_ours - mean_ref| <= max(2 % of mean_ref, 3 standard errors of the
  difference of the two means); the two start strata are also compared separately.  The reset draws of either side are a noisy
  sample (two seeds of 8192 CUDA episodes differ by 2.6 %), which is why the bar is not a bare 2 %.
* ``test_mean_return_from_the_reference_reset_states``: common random numbers: the CUDA simulator starts from the reference env's
  own reset states (``reset_states_pos.npz``), so only dynamics and env logic can differ (measured over all 12 000 seeds:
  +0.15 % +- 0.17 % for ``ppo_b200``; profiles/r2_experiments.md section 10)."""
import os

import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv

# + "ppo_b200": a policy trained HERE on the CUDA environments (examples/train_ppo.py, 200 M steps) and evaluated in the
# reference env code afterwards: the transfer direction SURVEY 8(d) config 5 asks for
TAGS = ("pos_0428", "pos_0421", "pos_0427", "pos_0428_best", "ppo_b200", "ppo_b200_r2")     # the three 48-d position final_model.zip + the 04-28 best_model.zip


def _policy(tag):
    w = dict(np.load(os.path.join(pc.GOLDEN, f"policy_{tag}.npz")))

    def act(obs, lo, hi):
        x = obs.astype(np.float32)
        h = np.tanh(x @ w["w0"].T + w["b0"]); h = np.tanh(h @ w["w1"].T + w["b1"])
        return np.clip(h @ w["w2"].T + w["b2"], lo, hi).astype(np.float32)
    return act


def run_episodes(lib, tag, n_envs, max_steps, seed=0, rand_power=1.0):
    """First episode of every env, capped at max_steps: returns (ret, length, fallen_start)."""
    env = Go1CudaVecEnv(n_envs, ctrl_type="position", rand_power=rand_power, library=lib, seed=seed, desired_velocity_min=(0.5, 0, 0),
                        desired_velocity_max=(0.5, 0, 0))
    act = _policy(tag)
    lo, hi = env.action_space.low, env.action_space.high
    obs = env.reset()
    assert np.allclose(obs[:, 9:12], [1.0, 0, 0])                     # the "x=1" target in the scaled observation
    fallen = np.array([abs(env.sim.get_state(e)["qpos"][2] - 0.1) < 1e-6 for e in range(n_envs)])
    ret = np.zeros(n_envs); length = np.zeros(n_envs, int); alive = np.ones(n_envs, bool)
    for _ in range(max_steps):
        obs, rew, dones, _ = env.step(act(obs, lo, hi))
        ret += np.where(alive, rew, 0.0); length += alive
        alive &= ~dones
        if not alive.any():
            break
    env.close()
    return ret, length, fallen


def test_protocol_smoke_emu(emu_lib_f32):
    """CPU: the evaluation harness itself (few envs, few steps) on the emulation build."""
    ret, length, fallen = run_episodes(emu_lib_f32, "pos_0428", 6, 12)
    assert np.all(length == 12) and np.all(ret >= 0) and np.isfinite(ret).all()
    ref = np.load(os.path.join(pc.GOLDEN, "returns_pos_0428.npz"))
    assert len(ref["ret"]) >= 1000 and abs(ref["fallen"].mean() - 0.2) < 0.04
    assert np.allclose(ref["v_des"], [0.5, 0, 0])


@pytest.mark.gpu
@pytest.mark.parametrize("tag", TAGS)
def test_shipped_policy_mean_return(cuda_lib_f32, tag):
    ref = np.load(os.path.join(pc.GOLDEN, f"returns_{tag}.npz"))
    max_steps = int(ref["max_steps"])
    ret, length, fallen = run_episodes(cuda_lib_f32, tag, 16384, max_steps, seed=2024, rand_power=float(ref["rand_power"]))
    r_ref, f_ref = ref["ret"], ref["fallen"] > 0.5
    se = r_ref.std() / np.sqrt(len(r_ref))
    se_d = float(np.hypot(se, ret.std() / np.sqrt(len(ret))))        # standard error of the difference of the two sample means
    tol = max(0.02 * r_ref.mean(), 3 * se_d)
    print(f"{tag}: ours {ret.mean():.2f} (n={len(ret)}, fallen {fallen.mean():.3f}, len {length.mean():.0f}) | reference "
          f"{r_ref.mean():.2f} +- {se:.2f} (n={len(r_ref)}, fallen {f_ref.mean():.3f}, len {ref['length'].mean():.0f}) | "
          f"rel diff {(ret.mean() - r_ref.mean()) / r_ref.mean():+.4f}, tol {tol / r_ref.mean():.4f}")
    for name, mo, mr in (("upright", ~fallen, ~f_ref), ("fallen", fallen, f_ref)):
        se_s = float(np.hypot(r_ref[mr].std() / np.sqrt(mr.sum()), ret[mo].std() / np.sqrt(mo.sum())))
        d = ret[mo].mean() - r_ref[mr].mean()
        print(f"   {name:8s} ours {ret[mo].mean():.2f} | reference {r_ref[mr].mean():.2f} +- {se_s:.2f} | rel diff {d / r_ref[mr].mean():+.4f}")
        assert abs(d) <= max(0.02 * abs(r_ref[mr].mean()), 3.5 * se_s), name
    assert abs(fallen.mean() - 0.2) < 0.02
    # stratified mean with the nominal 20/80 start mix removes the binomial noise of the fallen fraction
    strat_o = 0.2 * ret[fallen].mean() + 0.8 * ret[~fallen].mean()
    strat_r = 0.2 * r_ref[f_ref].mean() + 0.8 * r_ref[~f_ref].mean()
    assert abs(strat_o - strat_r) <= tol
    assert abs(ret.mean() - r_ref.mean()) <= tol


PAIRED_TAGS = ("ppo_b200", "ppo_b200_r2", "pos_0421", "pos_0428")


@pytest.mark.gpu
@pytest.mark.parametrize("tag", PAIRED_TAGS)
def test_mean_return_from_the_reference_reset_states(cuda_lib_f32, tag):
    """Common random numbers: the CUDA simulator starts from the reference environment's OWN reset states (``reset_states_pos.npz``,
    the first 4096 seeds of the fixtures), so the reset draws no longer add sampling noise and the comparison is episode by episode.
    ``ppo_b200`` is a contraction (93 % of its episodes end within 1 % of the reference's return); the shipped policies fall and are
    chaotic, for them the pairing only removes the reset noise.  Bar: |mean of the paired differences| <= max(1 % of the reference
    mean, 3 standard errors of the paired differences)."""
    st = np.load(os.path.join(pc.GOLDEN, "reset_states_pos.npz"))
    ref = np.load(os.path.join(pc.GOLDEN, f"returns_{tag}.npz"))
    idx = {int(s): i for i, s in enumerate(ref["seed"])}
    sel = np.array([int(s) in idx for s in st["seed"]])          # the 4000-episode fixture of ppo_b200_r2 lacks the last 96 seeds
    seeds, qpos0, obs0 = st["seed"][sel], st["qpos"][sel], st["obs0"][sel]
    n = len(seeds)
    assert n >= 4000
    env = Go1CudaVecEnv(n, ctrl_type="position", rand_power=float(st["rand_power"]), library=cuda_lib_f32, seed=1,
                        desired_velocity_min=tuple(st["v_des"]), desired_velocity_max=tuple(st["v_des"]))
    env.reset()
    for e in range(n):
        env.sim.set_state(e, qpos=qpos0[e], qvel=np.zeros(18), qacc_warmstart=np.zeros(18))
    act = _policy(tag)
    lo, hi = env.action_space.low, env.action_space.high
    obs = obs0.copy()
    ret = np.zeros(n); alive = np.ones(n, bool)
    for _ in range(int(ref["max_steps"])):
        obs, rew, dones, _ = env.step(act(obs, lo, hi))
        ret += np.where(alive, rew, 0.0)
        alive &= ~dones
        if not alive.any():
            break
    env.close()
    r_ref = np.array([ref["ret"][idx[int(s)]] for s in seeds])
    d = ret - r_ref
    se = d.std() / np.sqrt(n)
    close = np.abs(d) <= 0.01 * np.maximum(1.0, np.abs(r_ref))
    print(f"{tag}: {n} paired episodes | ours {ret.mean():.2f} reference {r_ref.mean():.2f} | paired difference {d.mean():+.2f} +- {se:.2f} "
          f"({d.mean() / r_ref.mean():+.4f}) | episodes within 1 %: {close.mean():.3f}")
    assert abs(d.mean()) <= max(0.01 * abs(r_ref.mean()), 3 * se)
    if tag == "ppo_b200":      # this policy is a contraction: the episodes pair up one by one (the round-2 policy: 22 %)
        assert close.mean() > 0.85

