"""Diagnostic: is the small negative offset of the CUDA mean returns (tests/test_policy_returns.py) sampling noise of the reset
draws or a real difference?  A walking policy is a contraction, so from IDENTICAL initial states the two simulators should follow
the same episodes; this tool runs the CUDA simulator from the reference env's own reset states (same seeds as the fixture) and
compares the returns episode by episode.

    python tests/golden/make_reset_states.py 12000 tools/scratch/reset_states.npz   # here (needs /root/reference); optional: all fixture seeds
    python tools/paired_returns.py run <tag> [f32|f64]                              # on the GPU box

Without tools/scratch/reset_states.npz the 4096 committed states of tests/golden/reset_states_pos.npz are used."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
OUT = os.path.join(ROOT, "tools", "scratch", "reset_states.npz")
if not os.path.exists(OUT):
    OUT = os.path.join(ROOT, "tests", "golden", "reset_states_pos.npz")


def run(tag, prec):
    import test_policy_returns as tpr
    from quadruped_rl_locomotion_b200.lib import Go1Lib
    from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv
    st0 = np.load(OUT)
    ref = np.load(os.path.join(ROOT, "tests", "golden", f"returns_{tag}.npz"))
    idx = {int(s): i for i, s in enumerate(ref["seed"])}
    keep = np.array([int(s) in idx for s in st0["seed"]])      # a fixture may hold fewer seeds than the reset-state file
    st = {k: st0[k][keep] for k in ("seed", "qpos", "obs0")}
    n = len(st["seed"])
    env = Go1CudaVecEnv(n, ctrl_type="position", rand_power=float(ref["rand_power"]), library=Go1Lib(precision=prec), seed=1,
                        desired_velocity_min=(0.5, 0, 0), desired_velocity_max=(0.5, 0, 0))
    env.reset()
    for e in range(n):
        env.sim.set_state(e, qpos=st["qpos"][e], qvel=np.zeros(18), qacc_warmstart=np.zeros(18))
    act = tpr._policy(tag)
    lo, hi = env.action_space.low, env.action_space.high
    obs = st["obs0"].astype(np.float64)
    ret = np.zeros(n); length = np.zeros(n, int); alive = np.ones(n, bool)
    for _ in range(int(ref["max_steps"])):
        obs, rew, dones, _ = env.step(act(obs, lo, hi))
        ret += np.where(alive, rew, 0.0); length += alive
        alive &= ~dones
        if not alive.any():
            break
    env.close()
    r_ref = np.array([ref["ret"][idx[int(s)]] for s in st["seed"]]); l_ref = np.array([ref["length"][idx[int(s)]] for s in st["seed"]])
    d = ret - r_ref
    same_len = length == l_ref
    close = np.abs(d) <= 0.01 * np.maximum(1.0, np.abs(r_ref))
    print(f"{tag} {prec}: n {n} | mean ours {ret.mean():.2f} reference {r_ref.mean():.2f} | paired mean difference {d.mean():+.3f} +- {d.std() / np.sqrt(n):.3f} "
          f"({d.mean() / r_ref.mean():+.4%}) | episodes within 1 %: {close.mean():.3f}, same length: {same_len.mean():.3f} | "
          f"median |d| {np.median(np.abs(d)):.4f}, mean of d over the others {d[~close].mean() if (~close).any() else 0:+.2f} (n {int((~close).sum())})")
    worst = np.argsort(-np.abs(d))[:5]
    for e in worst:
        print(f"   seed {int(st['seed'][e])}: ours {ret[e]:.1f} / {length[e]} steps, reference {r_ref[e]:.1f} / {int(l_ref[e])} steps")


if __name__ == "__main__":
    run(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "f32")
