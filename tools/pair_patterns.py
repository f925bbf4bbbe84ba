"""Diagnostic: which Hessian path (go1_warp.cuh pair_hessian) the steady-state population of the bench takes, from the contact lists
of the last substep: no pair contact / pair contacts inside the block-arrow / one coupled leg pair / several.

    python tools/pair_patterns.py [n_envs=4096]"""
import os
import sys
from collections import Counter

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM, Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.mjcf import Model  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    blob = load_model_blob("torque")
    gb = np.asarray(Model.from_blob(blob).geom_bodyid)
    sim = Sim(blob, n, lib=Go1Lib(), seed=1234)
    sim.rollout(300, ACTION_RANDOM, None); sim.stagger_phases(); sim.rollout(100, ACTION_RANDOM, None)
    sim.enable_diag(True)
    sim.rollout(1, ACTION_RANDOM, None)
    pat, kinds = Counter(), Counter()
    leg = lambda b: -1 if b < 2 else (b - 2) // 3          # MuJoCo body ids: 0 world, 1 trunk, 2.. = 3 per leg
    for e in range(n):
        pairs, inarrow = set(), 0
        for c in sim.get_contacts(e):
            if c["geom1"] <= 0 or not np.any(np.asarray(c["force"]) != 0):
                continue
            l1, l2 = leg(int(gb[c["geom1"]])), leg(int(gb[c["geom"]]))
            if l1 >= 0 and l2 >= 0 and l1 != l2:
                pairs.add((min(l1, l2), max(l1, l2)))
                kinds[(min(l1, l2), max(l1, l2))] += 1
            else:
                inarrow += 1
        if not pairs:
            pat["arrow only" if inarrow else "no active pair contact"] += 1
        elif len(pairs) == 1:
            pat["one coupled leg pair"] += 1
        else:
            legs = sorted(pairs)
            disjoint = len(pairs) == 2 and not (set(legs[0]) & set(legs[1]))
            pat["two disjoint leg pairs" if disjoint else "several connected leg pairs"] += 1
    for k, v in pat.most_common():
        print(f"{k:32s} {v / n:.3f}")
    print("coupled pairs by legs (0 FR, 1 FL, 2 RR, 3 RL):", {k: round(v / n, 3) for k, v in kinds.most_common()})
    sim.close()


if __name__ == "__main__":
    main()
