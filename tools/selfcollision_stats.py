"""How often would the robot-robot contacts that this round leaves out (DESIGN.md "deliberate scope limit") occur?

    python tools/selfcollision_stats.py sample out.npz     # GPU: sample qpos from policy and random-action rollouts
    python tools/selfcollision_stats.py analyse out.npz    # CPU: geometric overlap test of the non-adjacent geom pairs

The reference model gives every collision geom contype = conaffinity = 1, so MuJoCo also collides robot geoms of
bodies that are not parent and child.  The overlap test here is geometric only (capsules and spheres exact,
cylinders as the inscribed capsule, boxes through their signed distance at 7 points of the other geom's axis):
a LOWER bound of how often MuJoCo would have generated such a contact."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def sample(out):
    from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv
    import test_policy_returns as tpr
    n = 2048
    res = {}
    # shipped walking policy, the protocol of tests/test_policy_returns.py
    env = Go1CudaVecEnv(n, ctrl_type="position", rand_power=1.0, seed=11, desired_velocity_min=(0.5, 0, 0), desired_velocity_max=(0.5, 0, 0))
    act = tpr._policy("pos_0428")
    lo, hi = env.action_space.low, env.action_space.high
    obs = env.reset()
    qs = []
    for t in range(500):
        obs, _, _, _ = env.step(act(obs, lo, hi))
        if t % 25 == 24:
            qs.append(np.array([env.sim.get_state(e)["qpos"] for e in range(0, n, 8)]))
    res["policy"] = np.concatenate(qs)
    env.close()
    for ctrl in ("torque", "position"):
        env = Go1CudaVecEnv(n, ctrl_type=ctrl, seed=12)
        lo, hi = env.action_space.low, env.action_space.high
        env.reset()
        rng = np.random.default_rng(0)
        qs = []
        for t in range(200):
            env.step(rng.uniform(lo, hi, size=(n, 12)).astype(np.float32))
            if t % 20 == 19:
                qs.append(np.array([env.sim.get_state(e)["qpos"] for e in range(0, n, 8)]))
        res["random_" + ctrl] = np.concatenate(qs)
        env.close()
    np.savez_compressed(out, **res)
    print({k: v.shape for k, v in res.items()})


def _seg_seg(p1, d1, p2, d2):
    """min distance between segments p1 + s d1, p2 + t d2 (s, t in [-1, 1])."""
    best = np.inf
    for _ in range(2):          # alternate projections (converges for segments; 2 sweeps from 3 starts is plenty here)
        pass
    r = p1 - p2
    a, e, f = d1 @ d1, d2 @ d2, d2 @ r
    b, c = d1 @ d2, d1 @ r
    den = a * e - b * b
    cand = []
    s = np.clip((b * f - c * e) / den, -1, 1) if den > 1e-12 else 0.0
    for s0 in (s, -1.0, 1.0):
        t = np.clip((b * s0 + f) / e, -1, 1) if e > 1e-12 else 0.0
        s1 = np.clip((b * t - c) / a, -1, 1) if a > 1e-12 else 0.0
        cand.append(np.linalg.norm(r + s1 * d1 - t * d2))
    return min(cand)


def _box_sdf(p, half):
    q = np.abs(p) - half
    return np.linalg.norm(np.maximum(q, 0)) + min(max(q[0], q[1], q[2]), 0.0)


def analyse(path):
    from quadruped_rl_locomotion_b200.mjcf import Model, kinematics, quat_to_mat
    data = np.load(path)
    for key in data.files:
        m = Model.load(os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models",
                                    "go1_torque.go1m" if key == "random_torque" else "go1_position.go1m"))
        geoms = [g for g in range(1, m.ngeom) if m.geom_contype[g] != 0]
        par = m.body_parentid
        pairs = [(a, b) for i, a in enumerate(geoms) for b in geoms[i + 1:]
                 if m.geom_bodyid[a] != m.geom_bodyid[b] and par[m.geom_bodyid[a]] != m.geom_bodyid[b] and par[m.geom_bodyid[b]] != m.geom_bodyid[a]]
        Q = data[key]
        hits = np.zeros(len(Q), bool); by_pair = {}
        for i, q in enumerate(Q):
            k = kinematics(m, q)
            G = {}
            for g in geoms:
                b = m.geom_bodyid[g]
                R = k["xmat"][b].reshape(3, 3) @ quat_to_mat(m.geom_quat[g]).reshape(3, 3)
                pos = k["xpos"][b] + k["xmat"][b].reshape(3, 3) @ m.geom_pos[g]
                t, sz = int(m.geom_type[g]), m.geom_size[g]
                if t == 6:
                    G[g] = ("box", pos, R, sz.copy())
                else:
                    rad = sz[0]
                    hl = 0.0 if t == 2 else (sz[1] if t == 3 else max(sz[1] - sz[0], 0.0))      # cylinder -> inscribed capsule
                    G[g] = ("cap", pos, R[:, 2] * hl, rad)
            for a, b in pairs:
                ga, gb = G[a], G[b]
                if ga[0] == "box" and gb[0] == "box":
                    continue
                if ga[0] == "box" or gb[0] == "box":
                    bx, cp = (ga, gb) if ga[0] == "box" else (gb, ga)
                    d = min(_box_sdf(bx[2].T @ (cp[1] + s * cp[2] - bx[1]), bx[3]) for s in np.linspace(-1, 1, 7)) - cp[3]
                else:
                    d = _seg_seg(ga[1], ga[2], gb[1], gb[2]) - ga[3] - gb[3]
                if d < 0:
                    hits[i] = True
                    nm = tuple(sorted((m.names["body"][m.geom_bodyid[a]].split("_")[-1], m.names["body"][m.geom_bodyid[b]].split("_")[-1])))
                    by_pair[nm] = by_pair.get(nm, 0) + 1
        print(f"{key}: {len(Q)} sampled states, {hits.mean() * 100:.1f} % have at least one overlapping non-adjacent robot geom pair; "
              f"by body kinds: {dict(sorted(by_pair.items(), key=lambda kv: -kv[1])[:6])}")


if __name__ == "__main__":
    {"sample": sample, "analyse": analyse}[sys.argv[1]](sys.argv[2])
