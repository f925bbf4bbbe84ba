"""Where the time of one Go1CudaVecEnv.step goes (16384 envs, random actions).  Variants are interleaved step by
step so that the drift of the kernel time along the rollout cancels.

    python tools/e2e_breakdown.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv  # noqa: E402


def main():
    n = 16384
    env = Go1CudaVecEnv(n, seed=1234)
    env.reset()
    s = env.sim
    rng = np.random.default_rng(0)
    lo, hi = env.action_space.low, env.action_space.high
    acts = [rng.uniform(lo, hi, (n, 12)).astype(np.float32) for _ in range(8)]
    s.rollout(300, ACTION_RANDOM, None); s.stagger_phases(); s.rollout(100, ACTION_RANDOM, None)     # bench.py's burn-in: episodes end every step
    for i in range(10):
        env.step(acts[i % 8])
    obs_pageable = np.zeros((n, 48), np.float32)

    def v_env(a):
        env.step(a)

    def v_host(a):
        np.copyto(env._actions, a)
        s.step_host(env._actions, env._obs, env._rew, env._done, env._trunc, env._tobs, env._info)

    def v_host_min(a):
        np.copyto(env._actions, a)
        s.step_host(env._actions, env._obs, env._rew, env._done, env._trunc)

    def v_host_pageable(a):
        s.step_host(a, obs_pageable, env._rew, env._done, env._trunc, env._tobs, env._info)

    dev = []

    def v_dev(a):
        dev.append(s.rollout(1, ACTION_RANDOM, None)["device_ms"])

    variants = [("Go1CudaVecEnv.step", v_env), ("go1_step_host, registered buffers", v_host), ("same without terminal_obs / info", v_host_min),
                ("go1_step_host, pageable obs + actions", v_host_pageable), ("go1_rollout (device actions), wall", v_dev)]
    tot = {k: 0.0 for k, _ in variants}
    rounds = 40
    for r in range(rounds):
        for k, f in variants:
            a = acts[(r * 5 + len(k)) % 8]
            t0 = time.perf_counter(); f(a); tot[k] += time.perf_counter() - t0
    for k, _ in variants:
        print(f"{k:45s} {tot[k] / rounds * 1e3:7.3f} ms")
    print(f"{'k_step + k_order on the device (CUDA events)':45s} {np.mean(dev):7.3f} ms")
    env.close()


if __name__ == "__main__":
    main()
