import os, sys, time, subprocess
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM, Go1Lib, Sim
from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv, load_model_blob

def clk():
    return subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()

n = 16384
def dev(s, k=50):
    return np.mean([s.rollout(1, ACTION_RANDOM, None)["device_ms"] for _ in range(k)])
sim = Sim(load_model_blob("torque"), n, lib=Go1Lib(), seed=1234)
sim.rollout(300, ACTION_RANDOM, None); sim.stagger_phases(); sim.rollout(100, ACTION_RANDOM, None)
print("plain Sim, device rollouts:", dev(sim), clk())
sim.close()
env = Go1CudaVecEnv(n, seed=1234)
env.reset()
s = env.sim
s.rollout(300, ACTION_RANDOM, None); s.stagger_phases(); s.rollout(100, ACTION_RANDOM, None)
print("VecEnv sim, device rollouts:", dev(s), clk())
rng = np.random.default_rng(0)
lo, hi = env.action_space.low, env.action_space.high
acts = [rng.uniform(lo, hi, (n, 12)).astype(np.float32) for _ in range(8)]
t0 = time.perf_counter()
for i in range(50): env.step(acts[i % 8])
print("VecEnv.step wall (8 recycled action sets):", (time.perf_counter() - t0) / 50 * 1e3, clk())
print("device rollouts after:", dev(s), clk())
t0 = time.perf_counter()
for i in range(50): env.step(rng.uniform(lo, hi, (n, 12)).astype(np.float32))
print("VecEnv.step wall (fresh actions, incl. rng):", (time.perf_counter() - t0) / 50 * 1e3, clk())
print("device rollouts after:", dev(s), clk())
a = rng.uniform(lo, hi, (50, n, 12)).astype(np.float32)
t0 = time.perf_counter()
for i in range(50): env.step(a[i])
print("VecEnv.step wall (50 pre-drawn action sets):", (time.perf_counter() - t0) / 50 * 1e3, clk())
print("device rollouts after:", dev(s), clk())
env.close()
