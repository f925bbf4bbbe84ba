"""Generates ``tests/golden/reset_states_pos.npz``: the reference environment's own reset states (qpos and first observation) for
the first 4096 seeds of the return fixtures (``make_returns.py``: ``np.random.seed(s); env.reset()``; the reset does not depend on
the policy).  ``tests/test_policy_returns.py`` starts the CUDA simulator from exactly these states, which removes the sampling
noise of the reset draws from the comparison of the mean returns.

    python tests/golden/make_reset_states.py [n=4096] [output.npz]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


def main():
    import make_returns as mr
    from oracle.ref_env import make_reference_env
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    seeds = np.load(os.path.join(mr.OUT, "returns_ppo_b200.npz"))["seed"].astype(int)[:n]
    env = make_reference_env("position", rand_power=mr.RAND_POWER)
    env._desired_velocity_min = np.array(mr.V_DES); env._desired_velocity_max = np.array(mr.V_DES)
    qpos, obs0 = [], []
    for s in seeds:
        np.random.seed(int(s))
        o, _ = env.reset()
        qpos.append(env.data.qpos.copy()); obs0.append(o.copy())
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(mr.OUT, "reset_states_pos.npz")
    np.savez_compressed(out, seed=seeds, qpos=np.array(qpos, np.float64), obs0=np.array(obs0, np.float64),
                        rand_power=mr.RAND_POWER, v_des=np.array(mr.V_DES))
    print("wrote", out, len(seeds))


if __name__ == "__main__":
    main()
