"""Profiling driver: W warm-up + K timed device-resident control steps of the fused step kernel.

    python tools/prof_step.py [n_envs] [warmup] [steps] [ctrl] [stagger]

With a fifth argument "stagger" the warm-up is bench.py's burn-in: `warmup` steps, go1_stagger_phases, 100 more steps (the
steady-state mix of episode phases; k_step launches before the timed ones: warmup + 100).

Prints ms per control step (CUDA events inside go1_rollout) and the rollout statistics.  Short enough to
wrap in `ncu --set full -k regex:k_step -s <warmup+2> -c 1` (B200_PROFILING.md)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM, Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    warm = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    ctrl = sys.argv[4] if len(sys.argv) > 4 else "torque"
    libpath = os.environ.get("GO1_LIB")          # alternative build of the same source (tuning experiments)
    sim = Sim(load_model_blob(ctrl), n, lib=Go1Lib(libpath) if libpath else Go1Lib(), seed=1234)
    for _ in range(warm):
        sim.rollout(1, ACTION_RANDOM, None)
    if len(sys.argv) > 5 and sys.argv[5] == "stagger":
        sim.stagger_phases()
        for _ in range(100):
            sim.rollout(1, ACTION_RANDOM, None)
    tot = 0.0
    for _ in range(steps):
        st = sim.rollout(1, ACTION_RANDOM, None)
        tot += st["device_ms"]
    print(f"n_envs {n} ctrl {ctrl}: {tot / steps:.3f} ms / control step -> {n * steps / (tot * 1e-3):.4g} env-steps/s; last stats {st}")
    sim.close()


if __name__ == "__main__":
    main()
