"""Shared parity machinery: drives a ``libgo1sim`` build (real CUDA or the CPU SIMT emulation of the same
source) through the C ABI, teacher-forced from golden states, and compares with the golden outputs
(reference env code + oracle physics, ``tests/golden/make_golden.py``)."""
from __future__ import annotations

import os

import numpy as np

from quadruped_rl_locomotion_b200.lib import Sim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
MODELS = os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models")


def model_blob(ctrl: str) -> bytes:
    with open(os.path.join(MODELS, f"go1_{ctrl}.go1m"), "rb") as f:
        return f.read()


def load_trace(ctrl: str):
    """Config-1 trace as (pre, post) dicts of equal length arrays."""
    tr = np.load(os.path.join(GOLDEN, f"trace_{ctrl}.npz"))
    n = len(tr["reward"])
    pre = {k: tr[k][:n] for k in ("qpos", "qvel", "warm", "ep_step", "ep_last_action", "ep_feet_air_time",
                                  "ep_last_contacts", "ep_desired_velocity", "ep_time_unhealthy", "ep_last_dev")}
    post = {k: tr[k][1:n + 1] for k in ("qpos", "qvel", "warm", "ep_feet_air_time", "ep_last_contacts",
                                        "ep_time_unhealthy", "ep_last_dev")}
    for k in ("action", "qacc", "qfrc_actuator", "cfrc_ext", "trunk_z", "ncon", "con_geom", "con_geom1", "con_dim", "con_dist", "con_fn",
              "con_force", "con_pos", "obs", "reward", "info", "healthy", "terminated", "truncated", "solver_iter"):
        post[k] = tr[k][:n]
    return pre, post


def load_starts(ctrl: str, biped: bool = False):
    """``biped=True``: tests/golden/make_golden_biped.py pairs (reference env with biped=True)."""
    st = np.load(os.path.join(GOLDEN, f"starts_biped_{ctrl}.npz" if biped else f"starts_{ctrl}.npz"))
    pre = {k[4:]: st[k] for k in st.files if k.startswith("pre_")}
    post = {k[5:]: st[k] for k in st.files if k.startswith("post_")}
    for k in st.files:
        if not k.startswith(("pre_", "post_")):
            post[k] = st[k]
    return pre, post


def run_teacher_forced(lib, ctrl: str, pre, post, idx, capture_contacts: bool = False, biped: bool = False):
    """Loads golden state idx[i] into env i, steps all envs ONCE with the golden actions; returns results."""
    idx = np.asarray(idx)
    n = len(idx)
    cfg = lib.default_config()
    cfg.biped = 1 if biped else 0
    sim = Sim(model_blob(ctrl), n, lib=lib, cfg=cfg, seed=7)
    sim.enable_diag(True)
    for e, t in enumerate(idx):
        if biped:
            sim.set_state(e, front_feet_touched=int(pre["fft"][t]))
        sim.set_state(
            e, qpos=pre["qpos"][t], qvel=pre["qvel"][t], qacc_warmstart=pre["warm"][t],
            last_action=pre["ep_last_action"][t], feet_air_time=pre["ep_feet_air_time"][t],
            last_health_deviation=pre["ep_last_dev"][t], desired_velocity=pre["ep_desired_velocity"][t],
            last_contacts=(pre["ep_last_contacts"][t] > 0).astype(int), step=int(pre["ep_step"][t]),
            unhealthy_steps=int(round(float(pre["ep_time_unhealthy"][t]) / 0.02)), episode_return=0.0, bad_state=0)
    actions = np.ascontiguousarray(post["action"][idx].astype(np.float32))
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32)
    done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    tobs = np.zeros((n, 48), np.float32); info = np.zeros((n, 20), np.float32)
    sim.step_host(actions, obs, rew, done, trunc, tobs, info)
    out = dict(obs=obs, rew=rew, done=done, trunc=trunc, tobs=tobs, info=info, state=[], diag=[], contacts=[])
    if biped:
        out["binfo"] = sim.biped_info_host()
    for e in range(n):
        # a done env was auto-reset: its physics state is the reset state; parity on state is skipped for it
        out["state"].append(sim.get_state(e))
        out["diag"].append(sim.get_diag(e))
        if capture_contacts:
            out["contacts"].append(sim.get_contacts(e))
    sim.close()
    return out


def rel_err(a, b, floor=1.0):
    """max |a-b| / max(floor, max|b|): relative to the vector's scale, absolute below `floor`."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(floor, np.abs(b).max()))


def compare(out, pre, post, idx):
    """Returns dict of per-quantity worst errors over the batch + flag mismatches."""
    idx = np.asarray(idx)
    w = dict(qpos=0.0, qvel=0.0, qacc=0.0, qfrc_actuator=0.0, cfrc_ext=0.0, obs=0.0, reward=0.0, info=0.0,
             trunk_z=0.0, air_time=0.0)
    mism = dict(ncon=0, done=0, trunc=0, healthy=0, last_contacts=0, unhealthy_steps=0)
    per = dict(qpos=[], qvel=[], qacc=[], cfrc_ext=[])
    for e, t in enumerate(idx):
        s, d = out["state"][e], out["diag"][e]
        gold_done = bool(post["terminated"][t] or post["truncated"][t])
        mism["done"] += int(bool(out["done"][e]) != gold_done)
        mism["trunc"] += int(bool(out["trunc"][e]) != bool(post["truncated"][t] and not post["terminated"][t]))
        mism["ncon"] += int(d["ncon"] != int(post["ncon"][t]))
        mism["healthy"] += int(bool(out["info"][e, 13] > 0.5) != bool(post["healthy"][t]))
        if not gold_done:
            per["qpos"].append(rel_err(s["qpos"], post["qpos"][t])); per["qvel"].append(rel_err(s["qvel"], post["qvel"][t]))
            w["qpos"] = max(w["qpos"], per["qpos"][-1])
            w["qvel"] = max(w["qvel"], per["qvel"][-1])
            w["air_time"] = max(w["air_time"], rel_err(s["feet_air_time"], post["ep_feet_air_time"][t]))
            mism["last_contacts"] += int(np.any(np.asarray(s["last_contacts"]) != (post["ep_last_contacts"][t] > 0)))
            mism["unhealthy_steps"] += int(s["unhealthy_steps"] != int(round(float(post["ep_time_unhealthy"][t]) / 0.02)))
            w["obs"] = max(w["obs"], float(np.abs(out["obs"][e] - post["obs"][t]).max()))
        else:
            w["obs"] = max(w["obs"], float(np.abs(out["tobs"][e] - post["obs"][t]).max()))
        per["qacc"].append(rel_err(d["qacc"], post["qacc"][t]))
        w["qacc"] = max(w["qacc"], per["qacc"][-1])
        w["qfrc_actuator"] = max(w["qfrc_actuator"], rel_err(d["qfrc_actuator"], post["qfrc_actuator"][t]))
        per["cfrc_ext"].append(rel_err(d["cfrc_ext"], post["cfrc_ext"][t]))
        w["cfrc_ext"] = max(w["cfrc_ext"], per["cfrc_ext"][-1])
        w["trunk_z"] = max(w["trunk_z"], abs(d["trunk_xpos_z"] - post["trunk_z"][t]))
        # reward = max(0, sum of 6 rewards - 11 costs): the error is measured against the largest term that entered the sum
        # (info rows hold the 8 terms the reference exports), never below 1
        rscale = max(1.0, abs(float(post["reward"][t])), float(np.abs(post["info"][t]).max()))
        w["reward"] = max(w["reward"], abs(float(out["rew"][e]) - float(post["reward"][t])) / rscale)
        w["info"] = max(w["info"], float(np.abs(out["info"][e, 2:10] - post["info"][t]).max() /
                                         max(1.0, np.abs(post["info"][t]).max())))
    # 99th percentile over the batch of the three state quantities (check_state): a capsule lying ALONG a cylinder side or a
    # box face is a line contact whose contact point is ill-conditioned in any precision (MuJoCo's MPR point jumps there too),
    # so the bar applies to 99 % of the steps and 10x the bar to all of them
    w["q99"] = {k: (float(np.quantile(v, 0.99)) if len(v) else 0.0) for k, v in per.items()}
    w["q95"] = {k: (float(np.quantile(v, 0.95)) if len(v) else 0.0) for k, v in per.items()}     # for the ~50-step samples of the CPU suite
    return w, mism


def check_state(w, tol, slack=10.0, what=(), q="q99"):
    """tol[k] on the 99th (q="q95": 95th) percentile of qpos / qvel / qacc / cfrc_ext and slack * tol[k] on the worst step; tol[k] on
    everything else.  cfrc_ext in fp32: the friction force of a SLIDING contact points along the tangential part of J qacc - aref,
    a difference of O(100) accelerations; where that slip term is small its direction, and with it the body wrench, carries the 1e-5
    error of qacc amplified 100-1000x (1.3e-2 on one step of the random-torque trace), hence a quantile bar here as well."""
    for k, t in tol.items():
        if k in ("qpos", "qvel", "qacc") or (k == "cfrc_ext" and k in w.get(q, {})):
            assert w[q][k] <= t, (*what, k, q, w[q][k])
            assert w[k] <= slack * t, (*what, k, "worst", w[k])
        else:
            assert w[k] <= t, (*what, k, w[k])


MAX_CONTACTS_PER_ENV = 40     # GO1_CW: contacts kept per environment; more are dropped and COUNTED (Go1Diag.contact_overflow)


def drop_overflowed(out, idx, max_dropped=2):
    """Removes the environments in which a substep had more than GO1_CW contacts (reported by the kernel itself).
    The biped fixture holds two consecutive synthetic states (rear-leg stance joints with the trunk laid on its side
    10 cm above the floor: 42 contacts) beyond that documented capacity; nothing else may be dropped."""
    idx = np.asarray(idx)
    keep = np.array([d["contact_overflow"] == 0 for d in out["diag"]])
    assert (~keep).sum() <= max_dropped, idx[~keep]
    sel = np.flatnonzero(keep)
    res = {}
    for k, v in out.items():
        if isinstance(v, list):
            res[k] = [v[i] for i in sel] if len(v) == len(keep) else v
        else:
            res[k] = v[sel]
    return res, idx[sel]


def compare_biped(out, post, idx):
    """Worst error of the ten biped reward_info terms (relative to max(1, |term|)) + mismatches of the two flags.
    ``out['binfo'][:, 11]`` is the latch INCLUDING this step's front-foot forces (golden ``post_fft``)."""
    idx = np.asarray(idx)
    worst = np.zeros(10); flags = dict(bipedal_success=0, front_feet_touched=0)
    for e, t in enumerate(idx):
        g = post["binfo"][t]
        worst = np.maximum(worst, np.abs(out["binfo"][e, :10] - g) / np.maximum(1.0, np.abs(g)))
        flags["bipedal_success"] += int(bool(out["binfo"][e, 10] > 0.5) != bool(post["bipedal_success"][t]))
        # a front-foot force within 1e-3 of the 1 N threshold may legitimately fall on either side in fp32
        near = np.any(np.abs(post["front_forces"][t] - 1.0) < 1e-3)
        if not near:
            flags["front_feet_touched"] += int(bool(out["binfo"][e, 11] > 0.5) != bool(post["fft"][t]))
    return worst, flags


def contact_sets_equal(contacts, post, t, tie_eps=1e-6, margin=0.001):
    """Contact-set equality ((geom1, geom2) pairs with multiplicity; geom1 = 0 is the floor) excluding contacts within
    tie_eps of the margin."""
    gold = [(int(g1), int(g), float(dd)) for g1, g, dd in zip(post["con_geom1"][t], post["con_geom"][t], post["con_dist"][t]) if g >= 0]
    ours = [(c["geom1"], c["geom"], c["dist"]) for c in contacts]
    near = lambda dd: abs(dd - margin) < tie_eps
    a = sorted((g1, g) for g1, g, dd in gold if not near(dd))
    b = sorted((g1, g) for g1, g, dd in ours if not near(dd))
    excluded = sum(near(dd) for _, _, dd in gold) + sum(near(dd) for _, _, dd in ours)
    return a == b, excluded


def _force_invariants(f):
    """normal, |tangential|, torsional, |rolling| of a contact-frame force: independent of the choice of the two tangent axes
    (MuJoCo aligns the first tangent of a plane-capsule contact with the capsule axis, the kernel keeps the floor's constant
    frame; the cone is isotropic in the tangent plane, so the physics are the same)."""
    f = np.asarray(f, dtype=np.float64)
    return np.array([f[0], np.hypot(f[1], f[2]), f[3], np.hypot(f[4], f[5])])


def compare_contacts(contacts, post, t):
    """Per-contact comparison with the golden mjContact / efc_force records of step t (same MuJoCo order on both sides:
    floor contacts in geom order, then robot-robot contacts in body-pair order).  Returns the worst errors of
    dist (absolute), pos (absolute) and of the contact force RELATIVE TO THE LARGEST FORCE ROW OF THE STEP (north_star:
    "contact forces match mj_step within 1e-5 relative"): `force` = frame-independent invariants of every contact,
    `force_rows` = all rows of the robot-robot contacts (same frame on both sides).  None when the sets differ."""
    n = int(post["ncon"][t])
    if n != len(contacts) or n > post["con_geom"].shape[1]:
        return None
    fscale = max(1.0, float(np.abs(post["con_force"][t][:n]).max())) if n else 1.0
    w = dict(dist=0.0, pos=0.0, force=0.0, force_rows=0.0)
    for i, c in enumerate(contacts):
        if c["geom"] != int(post["con_geom"][t][i]) or c["geom1"] != int(post["con_geom1"][t][i]) or c["dim"] != int(post["con_dim"][t][i]):
            return None
        w["dist"] = max(w["dist"], abs(c["dist"] - float(post["con_dist"][t][i])))
        w["pos"] = max(w["pos"], float(np.abs(c["pos"] - post["con_pos"][t][i]).max()))
        w["force"] = max(w["force"], float(np.abs(_force_invariants(c["force"]) - _force_invariants(post["con_force"][t][i])).max()) / fscale)
        if c["geom1"] > 0:
            w["force_rows"] = max(w["force_rows"], float(np.abs(c["force"] - post["con_force"][t][i]).max()) / fscale)
    return w


def self_contact_categories(post, geom_bodyid):
    """Golden step indices by the kind of robot-robot contact they hold: cross (two different legs: dense leg block),
    same (one leg: inside its 3x3 block), trunk (trunk vs leg), foot (condim-6 foot contact with a robot geom)."""
    gb = np.asarray(geom_bodyid)
    cats = dict(cross=[], same=[], trunk=[], foot=[], none=[])
    for t in range(len(post["con_geom1"])):
        sel = post["con_geom1"][t] > 0
        if not sel.any():
            cats["none"].append(t)
            continue
        b1 = gb[post["con_geom1"][t][sel]]; b2 = gb[post["con_geom"][t][sel]]
        l1 = np.where(b1 >= 2, (b1 - 2) // 3, -1); l2 = np.where(b2 >= 2, (b2 - 2) // 3, -1)
        if ((l1 >= 0) & (l2 >= 0) & (l1 != l2)).any():
            cats["cross"].append(t)
        if ((l1 >= 0) & (l1 == l2)).any():
            cats["same"].append(t)
        if ((l1 < 0) | (l2 < 0)).any():
            cats["trunk"].append(t)
        if (post["con_dim"][t][sel] == 6).any():
            cats["foot"].append(t)
    return {k: np.array(v, dtype=int) for k, v in cats.items()}
