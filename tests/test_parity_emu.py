"""CPU-side parity: the CUDA source compiled against the SIMT emulation shim (tests/emu) is driven through the
same C ABI and compared with the golden vectors.  Small sample sizes keep the CPU suite at a few minutes;
the full-size versions of these checks run on the GPU (tests/test_parity_gpu.py)."""
import numpy as np
import pytest

import parity_common as pc

# tolerances (relative to the vector scale, see parity_common.rel_err); one control step = 10 mj_steps
F64 = dict(qpos=1e-10, qvel=1e-10, qacc=1e-9, qfrc_actuator=1e-10, cfrc_ext=1e-9, trunk_z=1e-12)
F32 = dict(qpos=5e-6, qvel=2e-5, qacc=1e-4, qfrc_actuator=1e-5, cfrc_ext=2e-3, trunk_z=1e-6, obs=2e-5, reward=2e-4, info=1e-5)     # see tests/test_parity_gpu.py for the reasons
F32_FORCE = 1e-2     # contact-frame force rows relative to the largest row of the step (see DESIGN.md, fp32 floors)


def _sample(n_total, k, seed=0):
    rng = np.random.default_rng(seed)
    head = np.arange(min(8, n_total))
    rest = rng.choice(np.arange(8, n_total), size=min(k, n_total - 8), replace=False) if n_total > 8 else []
    return np.unique(np.concatenate([head, rest, [n_total - 1]])).astype(int)


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_trace_f64(emu_lib_f64, ctrl):
    pre, post = pc.load_trace(ctrl)
    idx = _sample(len(post["reward"]), 40)
    out = pc.run_teacher_forced(emu_lib_f64, ctrl, pre, post, idx)
    w, mism = pc.compare(out, pre, post, idx)
    pc.check_state(w, F64, slack=100.0, q="q95")
    assert w["obs"] <= 2e-6 and w["reward"] <= 1e-5 and w["info"] <= 1e-6   # f32 output buffers
    assert all(v == 0 for v in mism.values()), mism


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_trace_f32(emu_lib_f32, ctrl):
    pre, post = pc.load_trace(ctrl)
    idx = _sample(len(post["reward"]), 40, seed=1)
    out = pc.run_teacher_forced(emu_lib_f32, ctrl, pre, post, idx)
    w, mism = pc.compare(out, pre, post, idx)
    pc.check_state(w, F32, q="q95")
    assert all(v == 0 for v in mism.values()), mism


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_random_starts_f64(emu_lib_f64, ctrl):
    """fallen / airborne / joint-limit / tilted starts: many body-floor contacts, limit rows, cylinders, boxes."""
    pre, post = pc.load_starts(ctrl)
    idx = _sample(len(post["reward"]), 60, seed=2)
    out = pc.run_teacher_forced(emu_lib_f64, ctrl, pre, post, idx, capture_contacts=True)
    w, mism = pc.compare(out, pre, post, idx)
    pc.check_state(w, F64, slack=100.0, q="q95")
    assert all(v == 0 for v in mism.values()), mism
    excluded = 0
    for e, t in enumerate(idx):
        ok, ex = pc.contact_sets_equal(out["contacts"][e], post, t)
        excluded += ex
        assert ok, f"contact set differs at golden index {t}"
        assert out["diag"][e]["contact_overflow"] == 0
    assert excluded <= 2


def test_single_substep_f32(emu_lib_f32):
    """The north_star's fp32 bar is per mj_step: one substep from identical state, 1e-5 relative."""
    from quadruped_rl_locomotion_b200.lib import Sim
    import os
    sub = np.load(os.path.join(pc.GOLDEN, "substeps_torque.npz"))
    cfg = emu_lib_f32.default_config(); cfg.frame_skip = 1
    idx = np.arange(0, len(sub["qacc"]), 7)
    sim = Sim(pc.model_blob("torque"), len(idx), lib=emu_lib_f32, cfg=cfg)
    sim.enable_diag(True)
    for e, t in enumerate(idx):
        sim.set_state(e, qpos=sub["qpos"][t], qvel=sub["qvel"][t], qacc_warmstart=sub["warm"][t], step=0, unhealthy_steps=0)
    n = len(idx)
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    sim.step_host(np.ascontiguousarray(sub["ctrl"][idx].astype(np.float32)), obs, rew, done, trunc)
    e_acc, e_pos, e_vel = [], [], []
    for e, t in enumerate(idx):
        d = sim.get_diag(e); s = sim.get_state(e)
        nq = sub["qpos"][t + 1] if t + 1 < len(sub["qacc"]) else sub["qpos_final"]
        nv = sub["qvel"][t + 1] if t + 1 < len(sub["qacc"]) else sub["qvel_final"]
        e_acc.append(pc.rel_err(d["qacc"], sub["qacc"][t])); e_pos.append(pc.rel_err(s["qpos"], nq)); e_vel.append(pc.rel_err(s["qvel"], nv))
    assert np.median(e_acc) < 2e-6 and max(e_acc) < 5e-5, (np.median(e_acc), max(e_acc))
    assert max(e_pos) < 1e-5 and max(e_vel) < 1e-5, (max(e_pos), max(e_vel))


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_biped_mode_f64(emu_lib_f64, ctrl):
    """cfg.biped = 1 against the reference's Go1MujocoEnv(biped=True): physics as before, plus the ten biped
    reward_info terms (:869-904), the alternating-rear-feet reward (:646-649), bipedal_success (:520-521) and the
    front-feet-touched latch (:477-479).  The fixture holds steps at the time limit with and without a touch."""
    pre, post = pc.load_starts(ctrl, biped=True)
    n = len(post["reward"])
    lim = np.flatnonzero(post["truncated"])
    idx = np.unique(np.concatenate([_sample(n, 50, seed=3), lim, [182, 183]])).astype(int)
    out = pc.run_teacher_forced(emu_lib_f64, ctrl, pre, post, idx, biped=True)
    out, idx = pc.drop_overflowed(out, idx)
    w, mism = pc.compare(out, pre, post, idx)
    pc.check_state(w, F64, slack=100.0, q="q95")
    assert w["obs"] <= 2e-6 and w["reward"] <= 1e-5 and w["info"] <= 1e-6
    assert all(v == 0 for v in mism.values()), mism
    worst, flags = pc.compare_biped(out, post, idx)
    assert worst.max() <= 1e-6, worst          # f32 output buffer
    assert all(v == 0 for v in flags.values()), flags
    assert post["bipedal_success"][lim].sum() == 2 and len(lim) == 6 and set(lim) <= set(idx)


def test_biped_mode_f32(emu_lib_f32):
    pre, post = pc.load_starts("position", biped=True)
    idx = _sample(len(post["reward"]), 40, seed=4)
    out = pc.run_teacher_forced(emu_lib_f32, "position", pre, post, idx, biped=True)
    out, idx = pc.drop_overflowed(out, idx)
    w, mism = pc.compare(out, pre, post, idx)
    pc.check_state(w, {k: v for k, v in F32.items() if k != "reward"}, q="q95")
    # reward = max(0, 15 upright - 150 |cfrc|^2 - ...): measured against the largest of the ten biped terms of the step
    rerr = max(abs(float(out["rew"][e]) - float(post["reward"][t])) / max(1.0, abs(float(post["reward"][t])), float(np.abs(post["binfo"][t]).max()))
               for e, t in enumerate(idx))
    assert rerr <= 1e-3, rerr
    assert all(v == 0 for v in mism.values()), mism
    worst, flags = pc.compare_biped(out, post, idx)
    # unwanted_contact = 150 * |cfrc|^2 reaches 1e6: relative 1e-4 of the term is the cfrc_ext tolerance squared-through
    assert worst.max() <= 2e-3, worst
    assert all(v == 0 for v in flags.values()), flags


def test_biped_reset_pose(emu_lib_f64):
    """Non-fallen biped resets start from BIPEDAL_READY_JOINTS at z = 0.65 pitched -95 deg (:967-972)."""
    from quadruped_rl_locomotion_b200.lib import Sim
    cfg = emu_lib_f64.default_config(); cfg.biped = 1
    sim = Sim(pc.model_blob("torque"), 24, lib=emu_lib_f64, cfg=cfg, seed=3)
    sim.reset_host()
    ready = np.array([0, 4.0, -2.0, 0, 4.0, -2.0, 0, 2.8, -1.2, 0, 2.8, -1.2])
    kinds = []
    for e in range(24):
        s = sim.get_state(e)
        if abs(s["qpos"][2] - 0.65) < 1e-12:
            kinds.append("biped")
            assert np.allclose(s["qpos"][7:], ready, atol=1e-15)
            a = np.deg2rad(-95.0)
            assert np.allclose(s["qpos"][3:7], [np.cos(a / 2), 0, np.sin(a / 2), 0], atol=1e-15)
        else:
            kinds.append("fallen")
            assert abs(s["qpos"][2] - 0.1) < 1e-12
        assert s["front_feet_touched"] == 0
    assert "biped" in kinds and "fallen" in kinds
    sim.close()


def test_contact_overflow_is_counted(emu_lib_f64):
    """More than GO1_CW = 40 simultaneous contacts: the extra ones (geom order) are dropped, and the drop is reported."""
    pre, post = pc.load_starts("torque", biped=True)
    over = np.flatnonzero(post["ncon"] > pc.MAX_CONTACTS_PER_ENV)
    assert len(over) >= 1
    out = pc.run_teacher_forced(emu_lib_f64, "torque", pre, post, over[:2], biped=True)
    for e in range(len(over[:2])):
        assert out["diag"][e]["contact_overflow"] > 0 and out["diag"][e]["ncon"] == pc.MAX_CONTACTS_PER_ENV
        assert np.isfinite(out["state"][e]["qpos"]).all()


def _self_contact_indices(ctrl, post, per_cat=6, seed=5):
    from quadruped_rl_locomotion_b200.mjcf import Model
    cats = pc.self_contact_categories(post, Model.from_blob(pc.model_blob(ctrl)).geom_bodyid)
    rng = np.random.default_rng(seed)
    pick = [rng.choice(v, size=min(per_cat, len(v)), replace=False) for k, v in cats.items() if k != "none" and len(v)]
    return cats, np.unique(np.concatenate(pick)).astype(int)


@pytest.mark.parametrize("prec", ["f64", "f32"])
def test_robot_robot_contacts(emu_lib_f64, emu_lib_f32, prec):
    """Rows a6/a7: steps whose golden contact list holds robot-robot contacts of every kind -- two different legs (the dense
    12 x 12 leg block of the Hessian), one leg (foot against its own hip motor), trunk vs leg, condim-6 foot contacts.
    State, contact set INCLUDING geom1, per-contact dist / pos / force rows against the oracle's mjContact + efc_force."""
    lib = emu_lib_f64 if prec == "f64" else emu_lib_f32
    pre, post = pc.load_starts("torque")
    cats, idx = _self_contact_indices("torque", post)
    assert all(len(cats[k]) > 0 for k in ("cross", "same", "foot")), {k: len(v) for k, v in cats.items()}
    pre_t, post_t = pc.load_trace("torque")
    cats_t, idx_t = _self_contact_indices("torque", post_t)
    assert len(cats_t["trunk"]) > 0
    for (p0, p1, ii) in ((pre, post, idx), (pre_t, post_t, np.unique(np.concatenate([idx_t, cats_t["trunk"]])))):
        out = pc.run_teacher_forced(lib, "torque", p0, p1, ii, capture_contacts=True)
        w, mism = pc.compare(out, p0, p1, ii)
        pc.check_state(w, F64 if prec == "f64" else F32, slack=100.0 if prec == "f64" else 10.0, q="q95")
        assert all(v == 0 for v in mism.values()), mism
        worst = dict(dist=0.0, pos=0.0, force=0.0, force_rows=0.0)
        for e, t in enumerate(ii):
            assert out["diag"][e]["contact_overflow"] == 0
            ok, _ = pc.contact_sets_equal(out["contacts"][e], p1, t)
            assert ok, f"contact set differs at golden index {t}"
            cw = pc.compare_contacts(out["contacts"][e], p1, t)
            if cw is not None:
                for k in worst:
                    worst[k] = max(worst[k], cw[k])
        lim = dict(dist=1e-12, pos=1e-11, force=1e-8, force_rows=1e-8) if prec == "f64" else dict(dist=2e-7, pos=2e-6, force=F32_FORCE, force_rows=F32_FORCE)
        for k in worst:
            assert worst[k] <= lim[k], (k, worst)


def test_biped_self_collision_cost(emu_lib_f64):
    """go1_mujoco_env.py:269-312: 25 x the number of contacts between the front-right and front-left limbs (or the rear pair)."""
    pre, post = pc.load_starts("torque", biped=True)
    hit = np.flatnonzero(post["binfo"][:, 9] != 0)
    assert len(hit) >= 1
    idx = np.unique(np.concatenate([hit, hit - 1, [0, 1]])).astype(int)
    out = pc.run_teacher_forced(emu_lib_f64, "torque", pre, post, idx, biped=True)
    worst, flags = pc.compare_biped(out, post, idx)
    assert worst.max() <= 1e-6, worst
    for e, t in enumerate(idx):
        assert abs(out["binfo"][e, 9] - post["binfo"][t, 9]) < 1e-6
