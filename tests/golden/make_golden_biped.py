"""Generates ``tests/golden/starts_biped_{torque,position}.npz``: (pre, post) pairs of the reference's UNMODIFIED
``Go1MujocoEnv(biped=True)`` (imported from /root/reference) on the fp64 oracle physics.  Run HERE:

    python tests/golden/make_golden_biped.py

Same layout as ``starts_*.npz`` (``make_golden.py``) plus the biped extras: the ten biped reward_info terms
(go1_mujoco_env.py:893-902), ``bipedal_success`` (:520-521), the front-feet-touched latch (:66, :477-479) and the
kinematic quantities the terms read from mjData.  Chunks: the biped reset pose (:967-972) under random actions,
fallen / tilted / airborne / joint-limit starts, and single steps at the episode time limit with and without an
earlier front-foot touch.  Physics half = oracle (PARITY UNPINNED vs MuJoCo); env half = reference code.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import make_golden as mg  # noqa: E402
from oracle.env_logic import BIPED_INFO_KEYS  # noqa: E402
from oracle.ref_env import make_reference_env  # noqa: E402

BODY = {"FR_hip": 2, "FR_calf": 4, "FL_hip": 5, "FL_calf": 7, "RR_hip": 8, "RL_hip": 11}   # MuJoCo body ids (world = 0)
SITE = {"FR": 2, "FL": 3}                                                                    # sites: head, imu, FR, FL, RR, RL


def kin_vector(d):
    """Order of the kernel's snapshot (go1_warp.cuh BK_*): hip z FR FL RR RL, rear hip y RR RL, front calf z, front foot z, trunk x-axis z."""
    return np.array([d.xpos[BODY["FR_hip"], 2], d.xpos[BODY["FL_hip"], 2], d.xpos[BODY["RR_hip"], 2], d.xpos[BODY["RL_hip"], 2],
                     d.xpos[BODY["RR_hip"], 1], d.xpos[BODY["RL_hip"], 1], d.xpos[BODY["FR_calf"], 2], d.xpos[BODY["FL_calf"], 2],
                     d.site_xpos[SITE["FR"], 2], d.site_xpos[SITE["FL"], 2], d.xmat[1].reshape(3, 3)[2, 0]])


def front_forces(d):
    return np.linalg.norm(d.cfrc_ext[[BODY["FR_calf"], BODY["FL_calf"]]], axis=1)


def record(env, actions, label):
    """``make_golden.record_trace`` plus the biped extras, one row per step."""
    ex = {k: [] for k in ("pre_fft", "post_fft", "binfo", "bipedal_success", "kin", "front_forces")}
    chunks = []
    for a in actions:
        pre_fft = bool(env._front_feet_touched) or bool(np.any(front_forces(env.data) > 1.0))    # what :477-479 will latch
        tr = mg.record_trace(env, [a], label)
        info = env._last_info
        ex["pre_fft"].append(pre_fft)
        ex["post_fft"].append(bool(env._front_feet_touched) or bool(np.any(front_forces(env.data) > 1.0)))
        ex["binfo"].append([info[k] for k in BIPED_INFO_KEYS]); ex["bipedal_success"].append(bool(info["bipedal_success"]))
        ex["kin"].append(kin_vector(env.data)); ex["front_forces"].append(front_forces(env.data))
        chunks.append(tr)
        if tr["terminated"][-1] or tr["truncated"][-1]:
            break
    return chunks, {k: np.array(v) for k, v in ex.items()}


def main():
    import builtins
    quiet = lambda *a, **k: None
    for ctrl in ("torque", "position"):
        env = make_reference_env(ctrl, biped=True)
        step0 = env.step

        def step(a, _s=step0, _e=env):          # keep the info dict of the last step (record_trace drops the biped keys)
            out = _s(a)
            _e._last_info = out[4]
            return out
        env.step = step
        all_chunks, extras = [], []
        plan = [("biped", 11, 120), ("biped", 12, 60), ("fallen", 13, 40), ("tilted", 14, 40), ("airborne", 15, 40), ("limits", 16, 30)]
        for kind, seed, n in plan:
            rng = np.random.default_rng(seed)
            if kind == "biped":
                np.random.seed(seed)
                while True:
                    env.reset()
                    if env.data.qpos[2] == 0.65:
                        break
            else:
                mg.random_start(env, rng, kind)
            acts = mg.sample_actions(env, rng, n)
            if kind == "biped" and seed == 12:      # small actions: the robot stays on its rear legs longer
                if ctrl == "torque":
                    acts = [a * 0.2 for a in acts]
                else:
                    ready = np.array(env.BIPEDAL_READY_JOINTS, dtype=np.float32)
                    acts = [np.clip(ready + 0.1 * rng.normal(size=12).astype(np.float32), env.action_space.low, env.action_space.high) for _ in acts]
            real_print = builtins.print
            builtins.print = quiet
            try:
                ch, ex = record(env, acts, kind)
            finally:
                builtins.print = real_print
            print(f"[{ctrl}] {kind} seed {seed}: {len(ch)} steps, mean reward {np.mean([c['reward'][0] for c in ch]):.3f}, "
                  f"front touched at end {bool(ex['post_fft'][-1])}, mean upright {ex['binfo'][:, 0].mean():.2f}")
            all_chunks += ch; extras.append(ex)
            # one step at the time limit (:498, :520-521), with the latch as the episode left it
            env._step = 5999
            builtins.print = quiet
            try:
                ch, ex = record(env, mg.sample_actions(env, rng, 1), "limit")
            finally:
                builtins.print = real_print
            assert ch[0]["truncated"][0]
            print(f"[{ctrl}]   time-limit step: bipedal_success {bool(ex['bipedal_success'][0])}")
            all_chunks += ch; extras.append(ex)
        pre_keys = ("qpos", "qvel", "warm", "ep_step", "ep_last_action", "ep_feet_air_time", "ep_last_contacts",
                    "ep_desired_velocity", "ep_time_unhealthy", "ep_last_dev")
        merged = {}
        for k in all_chunks[0]:
            if k in pre_keys:
                merged["pre_" + k] = np.concatenate([c[k][:-1] for c in all_chunks])
                merged["post_" + k] = np.concatenate([c[k][1:] for c in all_chunks])
            else:
                merged[k] = np.concatenate([c[k] for c in all_chunks])
        for k in extras[0]:
            merged[("" if k in ("binfo", "bipedal_success", "kin", "front_forces") else "x_") + k] = np.concatenate([e[k] for e in extras])
        merged["pre_fft"] = merged.pop("x_pre_fft"); merged["post_fft"] = merged.pop("x_post_fft")
        out = os.path.join(mg.OUT, f"starts_biped_{ctrl}.npz")
        np.savez_compressed(out, **merged)
        print(out, os.path.getsize(out) // 1024, "KiB,", len(merged["reward"]), "pairs, bipedal_success count", int(merged["bipedal_success"].sum()))


if __name__ == "__main__":
    main()
