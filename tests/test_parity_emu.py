"""CPU-side parity: the CUDA source compiled against the SIMT emulation shim (tests/emu) is driven through the
same C ABI and compared with the golden vectors.  Small sample sizes keep the CPU suite at a few minutes;
the full-size versions of these checks run on the GPU (tests/test_parity_gpu.py)."""
import numpy as np
import pytest

import parity_common as pc

# tolerances (relative to the vector scale, see parity_common.rel_err); one control step = 10 mj_steps
F64 = dict(qpos=1e-10, qvel=1e-10, qacc=1e-9, qfrc_actuator=1e-10, cfrc_ext=1e-9, trunk_z=1e-12)
F32 = dict(qpos=2e-6, qvel=1e-5, qacc=5e-5, qfrc_actuator=1e-5, cfrc_ext=5e-3, trunk_z=1e-6, obs=1e-5, reward=1e-3, info=1e-5)


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
    for k, tol in F64.items():
        assert w[k] <= tol, (k, w[k])
    assert w["obs"] <= 2e-6 and w["reward"] <= 1e-5 and w["info"] <= 1e-6   # f32 output buffers
    assert all(v == 0 for v in mism.values()), mism


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_trace_f32(emu_lib_f32, ctrl):
    pre, post = pc.load_trace(ctrl)
    idx = _sample(len(post["reward"]), 40, seed=1)
    out = pc.run_teacher_forced(emu_lib_f32, ctrl, pre, post, idx)
    w, mism = pc.compare(out, pre, post, idx)
    for k, tol in F32.items():
        assert w[k] <= tol, (k, w[k])
    assert all(v == 0 for v in mism.values()), mism


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_random_starts_f64(emu_lib_f64, ctrl):
    """fallen / airborne / joint-limit / tilted starts: many body-floor contacts, limit rows, cylinders, boxes."""
    pre, post = pc.load_starts(ctrl)
    idx = _sample(len(post["reward"]), 60, seed=2)
    out = pc.run_teacher_forced(emu_lib_f64, ctrl, pre, post, idx, capture_contacts=True)
    w, mism = pc.compare(out, pre, post, idx)
    for k, tol in F64.items():
        assert w[k] <= tol, (k, w[k])
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
