"""Is the small negative offset of the CUDA returns (tests/test_policy_returns.py) an fp32 effect?  Same protocol, same
seeds, fp32 product build vs fp64 check build, against the reference-env fixtures.

    python tools/returns_f64_vs_f32.py [tag ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_policy_returns as tpr  # noqa: E402
from quadruped_rl_locomotion_b200.lib import Go1Lib  # noqa: E402


def main():
    tags = sys.argv[1:] or ["pos_0421", "ppo_b200"]
    for tag in tags:
        ref = np.load(os.path.join(ROOT, "tests", "golden", f"returns_{tag}.npz"))
        r_ref, f_ref = ref["ret"], ref["fallen"] > 0.5
        for prec in ("f32", "f64"):
            for seed in (2024, 7):
                ret, length, fallen = tpr.run_episodes(Go1Lib(precision=prec), tag, 8192, int(ref["max_steps"]), seed=seed, rand_power=float(ref["rand_power"]))
                up_o, up_r = ret[~fallen].mean(), r_ref[~f_ref].mean()
                se_o = ret[~fallen].std() / np.sqrt((~fallen).sum()); se_r = r_ref[~f_ref].std() / np.sqrt((~f_ref).sum())
                print(f"{tag} {prec} seed {seed}: upright ours {up_o:.1f} +- {se_o:.1f} | reference {up_r:.1f} +- {se_r:.1f} | rel {(up_o - up_r) / up_r:+.4f}; "
                      f"fallen ours {ret[fallen].mean():.1f} | reference {r_ref[f_ref].mean():.1f}", flush=True)


if __name__ == "__main__":
    main()
