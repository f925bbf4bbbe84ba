"""Generates the committed golden fixtures.  Run HERE (needs /root/reference):

    python tests/golden/make_golden.py

Produces
  quadruped_rl_locomotion_b200/models/go1_{torque,position}.go1m   compiled model blobs (product assets)
  tests/golden/trace_{torque,position}.npz      BASELINE config 1 traces (reference env code, oracle physics)
  tests/golden/starts_{torque,position}.npz     randomised-start traces (fallen / airborne / joint-limit cases)
  tests/golden/substeps_torque.npz              per-substep (nstep=1) states for the first 50 control steps

How: the reference's UNMODIFIED ``go1_mujoco_env.py`` is imported from /root/reference and run on the
fp64 oracle physics (``oracle/ref_env.py``); MuJoCo itself is not installable in this image, so the
physics half of every vector is the oracle's (PARITY UNPINNED vs MuJoCo), while the
observation / reward / termination half is produced by the reference's literal code.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.env_logic import INFO_KEYS  # noqa: E402
from oracle.ref_env import make_reference_env  # noqa: E402
from quadruped_rl_locomotion_b200.mjcf import compile_mjcf  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
MAXC = 48


def _snapshot_episode(env):
    return dict(
        ep_step=np.int64(env._step), ep_last_action=np.array(env._last_action, dtype=np.float64),
        ep_feet_air_time=np.array(env._feet_air_time, dtype=np.float64),
        ep_last_contacts=np.array(env._last_contacts, dtype=np.float64),
        ep_desired_velocity=np.array(env._desired_velocity, dtype=np.float64),
        ep_time_unhealthy=np.float64(env._time_in_unhealthy_state),
        ep_last_dev=np.array([env._last_health_deviation[k] for k in ("z", "roll", "pitch")], dtype=np.float64),
    )


def _snapshot_physics(env):
    d = env.data
    return dict(qpos=d.qpos.copy(), qvel=d.qvel.copy(), warm=d.qacc_warmstart.copy())


def record_trace(env, actions, label):
    """Runs env.step over `actions`; returns dict of stacked arrays (index 0 = initial state)."""
    rec = {k: [v] for k, v in {**_snapshot_physics(env), **_snapshot_episode(env)}.items()}
    post = {k: [] for k in ("action", "qacc", "qfrc_actuator", "cfrc_ext", "trunk_z", "ncon", "con_geom", "con_geom1", "con_dim",
                            "con_dist", "con_fn", "con_force", "con_pos", "obs", "reward", "info", "healthy", "terminated", "truncated", "solver_iter")}
    for a in actions:
        obs, reward, terminated, truncated, info = env.step(a)
        d = env.data
        for k, v in {**_snapshot_physics(env), **_snapshot_episode(env)}.items():
            rec[k].append(v)
        post["action"].append(np.array(a, dtype=np.float64))
        post["qacc"].append(d.qacc.copy()); post["qfrc_actuator"].append(d.qfrc_actuator.copy())
        post["cfrc_ext"].append(d.cfrc_ext.copy()); post["trunk_z"].append(d.xpos[1, 2])
        cons = [c for c in d.contact if c.exclude == 0]
        g = -np.ones(MAXC, np.int32); g1 = -np.ones(MAXC, np.int32); dim = np.zeros(MAXC, np.int32)
        dist = np.zeros(MAXC); fn = np.zeros(MAXC); frc = np.zeros((MAXC, 6)); cpos = np.zeros((MAXC, 3))
        f = d.efc_force
        for i, c in enumerate(cons[:MAXC]):
            g[i] = c.geom2; g1[i] = c.geom1; dim[i] = c.dim; dist[i] = c.dist; fn[i] = f[c.efc_address]
            frc[i, :c.dim] = f[c.efc_address:c.efc_address + c.dim]; cpos[i] = c.pos
        post["ncon"].append(len(cons)); post["con_geom"].append(g); post["con_dist"].append(dist); post["con_fn"].append(fn)
        post["con_geom1"].append(g1); post["con_dim"].append(dim); post["con_force"].append(frc); post["con_pos"].append(cpos)
        post["obs"].append(obs.copy()); post["reward"].append(reward)
        post["info"].append(np.array([info.get(k, 0.0) for k in INFO_KEYS], dtype=np.float64))   # biped mode drops two keys
        post["healthy"].append(info["termination_reason"] == "not_terminated")
        post["terminated"].append(terminated); post["truncated"].append(truncated)
        post["solver_iter"].append(d.solver_iter)
        if terminated or truncated:
            break
    out = {k: np.array(v) for k, v in rec.items()}
    out.update({k: np.array(v) for k, v in post.items()})
    nself = sum(int(np.sum(g1 > 0)) for g1 in post["con_geom1"])
    print(f"  {label}: {len(post['reward'])} steps, mean reward {np.mean(post['reward']):.4f}, "
          f"max ncon {max(post['ncon'])}, robot-robot contacts {nself}, mean iters {np.mean(post['solver_iter']):.2f}, "
          f"healthy frac {np.mean(post['healthy']):.2f}")
    return out


def sample_actions(env, rng, n):
    lo, hi = env.action_space.low.astype(np.float64), env.action_space.high.astype(np.float64)
    return [rng.uniform(lo, hi).astype(np.float32) for _ in range(n)]


def home_reset(env, v_des=(0.5, 0.0, 0.0)):
    """BASELINE config 1 start: mj_resetData, qpos <- key_qpos[0], fallen branch disabled, v_des fixed."""
    np.random.seed(12345)
    while True:
        env.reset()
        if abs(env.data.qpos[2] - 0.30) < 1e-12 and env.data.qpos[3] == 1.0:  # not the fallen branch
            break
    env._desired_velocity = np.array(v_des, dtype=np.float64)


def random_start(env, rng, kind):
    np.random.seed(int(rng.integers(1 << 31)))
    env.reset()
    d = env.data
    jr = env.model.jnt_range[1:]
    if kind == "fallen":
        ang = rng.uniform(np.pi / 2.1, np.pi / 1.5)
        if rng.random() < 0.5:
            d.qpos[3:7] = [np.cos(ang / 2), np.sin(ang / 2), 0, 0]
        else:
            d.qpos[3:7] = [np.cos(ang / 2), 0, np.sin(ang / 2), 0]
        d.qpos[2] = 0.1
    elif kind == "airborne":
        d.qpos[2] = rng.uniform(0.45, 0.6)
        q = rng.normal(size=4); d.qpos[3:7] = q / np.linalg.norm(q)
        d.qpos[7:] = rng.uniform(jr[:, 0], jr[:, 1])
        d.qvel[:] = rng.normal(size=18)
    elif kind == "limits":
        d.qpos[2] = rng.uniform(0.25, 0.4)
        d.qpos[7:] = np.where(rng.random(12) < 0.5, jr[:, 0] + 0.01, jr[:, 1] - 0.01)
        d.qvel[6:] = rng.normal(size=12) * 3
    elif kind == "folded":       # knees fully folded, hips abducted inwards, legs crossing: robot-robot contacts from the first step
        d.qpos[2] = rng.uniform(0.12, 0.3)
        e = rng.uniform(-0.8, 0.8, size=3)
        q = np.array([1.0, *(0.5 * e)]); d.qpos[3:7] = q / np.linalg.norm(q)
        lo, hi = jr[:, 0], jr[:, 1]
        u = rng.random(12)
        d.qpos[7:] = lo + (hi - lo) * np.where(np.arange(12) % 3 == 2, 0.06 * u, u)      # knees near their lower limit
        d.qpos[7::3] = np.array([1, -1, 1, -1]) * rng.uniform(0.4, 0.86, size=4)           # abduction towards the body
        d.qvel[:] = rng.normal(size=18) * 0.5
    elif kind == "tilted":
        d.qpos[2] = rng.uniform(0.25, 0.35)
        e = rng.uniform(-0.6, 0.6, size=3)
        q = np.array([1.0, *(0.5 * e)]); d.qpos[3:7] = q / np.linalg.norm(q)
        d.qvel[:6] = rng.normal(size=6) * 0.5
    env._desired_velocity = np.array([rng.uniform(-0.5, 0.5), 0.0, 0.0])


def record_substeps(env, actions):
    """nstep=1 stepping (same physics as nstep=10) to expose every substep state."""
    d = env.data
    out = {k: [] for k in ("qpos", "qvel", "warm", "qacc", "ctrl", "ncon", "solver_iter", "cfrc_ext", "con_geom", "con_geom1", "con_dim", "con_force")}
    for a in actions:
        d.ctrl[:] = a
        for _ in range(env.frame_skip):
            out["qpos"].append(d.qpos.copy()); out["qvel"].append(d.qvel.copy()); out["warm"].append(d.qacc_warmstart.copy())
            out["ctrl"].append(np.array(a, dtype=np.float64))
            env.sim.step(1)
            env.sim.rne_post_constraint()      # cfrc_ext of THIS mj_step (does not change the state)
            cons = [c for c in d.contact if c.exclude == 0]
            g = -np.ones(MAXC, np.int32); g1 = -np.ones(MAXC, np.int32); dim = np.zeros(MAXC, np.int32); frc = np.zeros((MAXC, 6))
            for i, c in enumerate(cons[:MAXC]):
                g[i] = c.geom2; g1[i] = c.geom1; dim[i] = c.dim; frc[i, :c.dim] = d.efc_force[c.efc_address:c.efc_address + c.dim]
            out["qacc"].append(d.qacc.copy()); out["ncon"].append(len(cons)); out["solver_iter"].append(d.solver_iter)
            out["cfrc_ext"].append(d.cfrc_ext.copy()); out["con_geom"].append(g); out["con_geom1"].append(g1); out["con_dim"].append(dim)
            out["con_force"].append(frc)
    out["qpos_final"] = d.qpos.copy(); out["qvel_final"] = d.qvel.copy()
    return {k: np.array(v) for k, v in out.items()}


def main():
    os.makedirs(OUT, exist_ok=True)
    mdir = os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models")
    os.makedirs(mdir, exist_ok=True)
    for ctrl in ("torque", "position"):
        model = compile_mjcf(f"/root/reference/unitree_go1/scene_{ctrl}.xml")
        model.save(os.path.join(mdir, f"go1_{ctrl}.go1m"))
        print(f"[{ctrl}] model blob written")
        env = make_reference_env(ctrl)
        # --- BASELINE config 1: home start, seeded random actions, 1000 control steps
        home_reset(env)
        rng = np.random.default_rng(0)
        n = 1000 if ctrl == "torque" else 400
        tr = record_trace(env, sample_actions(env, rng, n), "config-1 trace")
        np.savez_compressed(os.path.join(OUT, f"trace_{ctrl}.npz"), **tr)
        # --- randomised starts
        chunks = []
        for seed, kind in enumerate(("fallen", "airborne", "limits", "tilted", "fallen", "airborne", "folded", "folded"), start=1):
            rng = np.random.default_rng(seed)
            random_start(env, rng, kind)
            chunks.append(record_trace(env, sample_actions(env, rng, 120), f"start seed {seed} ({kind})"))
        # concatenate as independent (pre, post) pairs
        pre_keys = ("qpos", "qvel", "warm", "ep_step", "ep_last_action", "ep_feet_air_time", "ep_last_contacts",
                    "ep_desired_velocity", "ep_time_unhealthy", "ep_last_dev")
        merged = {}
        for k in chunks[0]:
            if k in pre_keys:
                merged["pre_" + k] = np.concatenate([c[k][:-1] for c in chunks])
                merged["post_" + k] = np.concatenate([c[k][1:] for c in chunks])
            else:
                merged[k] = np.concatenate([c[k] for c in chunks])
        np.savez_compressed(os.path.join(OUT, f"starts_{ctrl}.npz"), **merged)
        if ctrl == "torque":
            home_reset(env)
            rng = np.random.default_rng(0)
            sub = record_substeps(env, sample_actions(env, rng, 50))
            np.savez_compressed(os.path.join(OUT, "substeps_torque.npz"), **sub)
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)) // 1024, "KiB")


if __name__ == "__main__":
    main()
