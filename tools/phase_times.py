"""Where a warp's time goes inside k_step, by Hessian path of its environment (diagnostic build -DGO1_PHASE_TIMING).

    tools/build_variant.sh pt -DGO1_PHASE_TIMING && GO1_LIB=tools/variants/libgo1sim_pt.so python tools/phase_times.py [n_envs] [steps]

Per class of environment (no pair rows / pair rows never active / active pair rows inside the block-arrow / one coupled leg pair /
dense 12 x 12): share of the warp's control step spent WORKING in each phase and WAITING at the barrier that ends it."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM, Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402

PH = ["kinematics..floor collision (->F1)", "pair collision (->F2)", "rows + evaluate + gradient (->B1)", "Hessian (->B2)", "factor + solve (->B3)",
      "line search (->round barrier)", "Euler + tail (->substep barrier)"]
CLS = ["no pair rows", "pair rows, never active", "active pair rows, block-arrow", "one coupled leg pair", "dense 12x12"]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    lib = Go1Lib(os.environ["GO1_LIB"])
    sim = Sim(load_model_blob("torque"), n, lib=lib, seed=1234)
    sim.rollout(300, ACTION_RANDOM, None); sim.stagger_phases(); sim.rollout(100, ACTION_RANDOM, None)
    buf = (C.c_ulonglong * 80)()
    lib.L.go1_debug_phase_times.argtypes = [C.c_void_p]
    lib.L.go1_debug_phase_times(buf)                       # clear
    st = sim.rollout(steps, ACTION_RANDOM, None)
    lib.L.go1_debug_phase_times(buf)
    a = np.array(buf[:], dtype=np.float64).reshape(5, 16)
    tot_all = a[:, :7].sum() + a[:, 8:15].sum()
    print(f"{st['device_ms'] / steps:.3f} ms per control step; warp-time shares of the whole launch (work / wait):")
    w, q = a[:, :7].sum(0), a[:, 8:15].sum(0)
    for k in range(7):
        print(f"  {PH[k]:42s} work {w[k] / tot_all:6.3f}  wait {q[k] / tot_all:6.3f}")
    print(f"  {'total':42s} work {w.sum() / tot_all:6.3f}  wait {q.sum() / tot_all:6.3f}")
    for c in range(5):
        cnt = a[c, 7]
        if cnt == 0:
            continue
        tot = a[c, :7].sum() + a[c, 8:15].sum()
        print(f"\n{CLS[c]}: {cnt / (n * steps):.3f} of the env-steps, {a[c, 15] / cnt:.1f} Newton rounds with a Hessian per control step, "
              f"{tot / cnt / 1e3:.0f} k cycles per env-step")
        for k in range(7):
            print(f"  {PH[k]:42s} work {a[c, k] / cnt / 1e3:8.1f} k  wait {a[c, 8 + k] / cnt / 1e3:8.1f} k")
    sim.close()


if __name__ == "__main__":
    main()
