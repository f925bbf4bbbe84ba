"""Diagnostic: Newton iterations PER SUBSTEP (frame_skip = 1 stepping with each random action held for 10 substeps)
and how well the previous substep predicts them.  Question behind it: blocks advance in lockstep, so a block costs
max over its 12 environments; would re-sorting the queue between substeps balance the blocks?

    python tools/substep_iters.py [n_envs] [n_control_steps] [dump.npz]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from quadruped_rl_locomotion_b200.lib import Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 120
    dump = sys.argv[3] if len(sys.argv) > 3 else None
    lib = Go1Lib()
    cfg = lib.default_config()
    cfg.frame_skip = 1; cfg.max_unhealthy_steps = 7500; cfg.max_episode_steps = 60000
    sim = Sim(load_model_blob("torque"), n, lib=lib, cfg=cfg, seed=1234)
    lo, hi = sim.action_bounds()
    rng = np.random.default_rng(0)
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    tobs = np.zeros((n, 48), np.float32); info = np.zeros((n, 20), np.float32)
    its, ncs = [], []
    for t in range(steps):
        a = rng.uniform(lo, hi, size=(n, 12)).astype(np.float32)
        for s in range(10):
            sim.step_host(a, obs, rew, done, trunc, tobs, info)
            if t >= steps - 40:
                its.append(info[:, 14].copy()); ncs.append(info[:, 15].copy())
    it = np.array(its).reshape(-1, 10, n); nc = np.array(ncs).reshape(-1, 10, n)      # [control step][substep][env]
    if dump:
        np.savez_compressed(dump, iters=it.astype(np.int16), ncon=nc.astype(np.int16))
    m = n - n % 12
    tot = it.sum(axis=1)
    print(f"iterations per substep mean {it.mean():.2f}, per control step mean {tot.mean():.1f}")
    print(f"corr(substep s, s+1) = {np.mean([np.corrcoef(it[c, s], it[c, s + 1])[0, 1] for c in range(len(it)) for s in range(9)]):.3f}; "
          f"corr(control-step totals t, t+1) = {np.mean([np.corrcoef(tot[c], tot[c + 1])[0, 1] for c in range(len(it) - 1)]):.3f}")
    rnd = np.random.default_rng(1).permutation(n)[:m]

    def rounds(order_of):            # mean over blocks of (sum over substeps of max over the block's 12 envs)
        r = []
        for c in range(1, len(it)):
            for s in range(10):
                o = order_of(c, s)[:m]
                r.append(it[c, s][o].reshape(-1, 12).max(axis=1).mean())
        return 10 * np.mean(r)
    print(f"block rounds per control step: fixed random groups {rounds(lambda c, s: rnd):.1f} | "
          f"re-sorted every substep by the previous substep's iterations {rounds(lambda c, s: np.argsort(-(it[c, s - 1] if s else it[c - 1, 9]), kind='stable')):.1f} | "
          f"by previous substep's iterations*8+contacts {rounds(lambda c, s: np.argsort(-((it[c, s - 1] if s else it[c - 1, 9]) * 64 + (nc[c, s - 1] if s else nc[c - 1, 9])), kind='stable')):.1f} | "
          f"perfect {rounds(lambda c, s: np.argsort(-it[c, s], kind='stable')):.1f}")
    sim.close()


if __name__ == "__main__":
    main()
