"""Diagnostic: distribution of Newton iterations / active contacts per control step in a random rollout.

    python tools/iter_hist.py [n_envs] [n_steps] [ctrl]

Uses the C ABI (go1_step_host with the info block: info[14] = solver iterations summed over the
frame_skip substeps, info[15] = active contacts of the last substep)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from quadruped_rl_locomotion_b200.lib import Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    ctrl = sys.argv[3] if len(sys.argv) > 3 else "torque"
    lib = Go1Lib(os.environ["GO1_LIB"]) if os.environ.get("GO1_LIB") else Go1Lib()      # alternative build (tools/build_variant.sh)
    sim = Sim(load_model_blob(ctrl), n, lib=lib, seed=1234)
    lo, hi = sim.action_bounds()
    rng = np.random.default_rng(0)
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32)
    done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    tobs = np.zeros((n, 48), np.float32); info = np.zeros((n, 20), np.float32)
    dump = os.environ.get("GO1_ITER_DUMP")
    hist_it, hist_nc = [], []
    for t in range(steps):
        a = rng.uniform(lo, hi, size=(n, 12)).astype(np.float32)
        t0 = time.perf_counter()
        sim.step_host(a, obs, rew, done, trunc, tobs, info)
        dt = time.perf_counter() - t0
        if dump:
            hist_it.append(info[:, 14].copy()); hist_nc.append(info[:, 15].copy())
        if t % 10 == 0 or t == steps - 1:
            it = info[:, 14]; nc = info[:, 15]
            q = np.percentile(it, [0, 50, 90, 99, 100])
            # block-level cost: the 12 envs of a block run max-over-12 rounds (slab order here, the kernel's queue is work-sorted)
            wmax = it.reshape(-1, 12).max(axis=1) if n % 12 == 0 else it[: n // 12 * 12].reshape(-1, 12).max(axis=1)
            print(f"step {t:4d} {dt*1e3:7.2f} ms | iters/ctrl-step min/med/p90/p99/max {q} mean {it.mean():.1f} warp-max mean {wmax.mean():.1f}"
                  f" | ncon med {np.median(nc):.0f} max {nc.max():.0f} mean {nc.mean():.2f} | healthy {info[:,13].mean():.3f} done {done.sum()}")
    if dump:
        np.savez_compressed(dump, iters=np.array(hist_it, np.int16), ncon=np.array(hist_nc, np.int16))
    sim.close()


if __name__ == "__main__":
    main()
