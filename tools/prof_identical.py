"""Diagnostic: all environments identical (no fallen starts, zero actions) -> all warps execute the same
path at about the same time.  Compared with the random-action run this separates instruction-fetch
sharing effects from per-instruction costs (IPC under ncu)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from quadruped_rl_locomotion_b200.lib import ACTION_ZERO, Go1Lib, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
lib = Go1Lib()
cfg = lib.default_config()
cfg.fallen_start_prob = 0.0
cfg.desired_vel_min[0] = 0.5; cfg.desired_vel_max[0] = 0.5
sim = Sim(load_model_blob("torque"), n, lib=lib, cfg=cfg, seed=1)
for _ in range(20):
    sim.rollout(1, ACTION_ZERO, None)
tot = 0.0
for _ in range(5):
    st = sim.rollout(1, ACTION_ZERO, None)
    tot += st["device_ms"]
print(f"identical envs: {tot / 5:.3f} ms / control step")
sim.enable_diag(True)
sim.rollout(1, ACTION_ZERO, None)
d = sim.get_diag(0)
print("iters_total", d["solver_iter_total"], "ncon", d["ncon"])
sim.close()
