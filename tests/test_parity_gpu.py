"""GPU parity (the parity tests proper): every golden vector, through the C ABI of the nvcc-built library.

Each golden step t becomes environment t of ONE batched launch (teacher-forced from the golden pre-state),
so the whole BASELINE config-1 trace (769 control steps) and all randomised-start sets are checked at once."""
import os

import numpy as np
import pytest

import parity_common as pc

pytestmark = pytest.mark.gpu

# one control step = frame_skip(10) x mj_step; tolerances relative to the vector scale (parity_common.rel_err)
F64 = dict(qpos=1e-10, qvel=1e-10, qacc=1e-9, qfrc_actuator=1e-10, cfrc_ext=1e-9, trunk_z=1e-12)
F32 = dict(qpos=2e-6, qvel=1e-5, qacc=5e-5, qfrc_actuator=1e-5, cfrc_ext=5e-3, trunk_z=1e-6, obs=1e-5, reward=1e-3, info=1e-5)


def _check(lib, ctrl, which, tol, flags_exact=True, contacts=True):
    pre, post = pc.load_trace(ctrl) if which == "trace" else pc.load_starts(ctrl)
    idx = np.arange(len(post["reward"]))
    out = pc.run_teacher_forced(lib, ctrl, pre, post, idx, capture_contacts=contacts)
    w, mism = pc.compare(out, pre, post, idx)
    for k, t in tol.items():
        assert w[k] <= t, (ctrl, which, k, w[k])
    if flags_exact:
        assert all(v == 0 for v in mism.values()), mism
    if contacts:
        excluded = 0
        for e, t in enumerate(idx):
            ok, ex = pc.contact_sets_equal(out["contacts"][e], post, t)
            excluded += ex
            assert ok, f"contact set differs at golden index {t}"
            assert out["diag"][e]["contact_overflow"] == 0
        assert excluded <= 4
    return w


@pytest.mark.parametrize("which", ["trace", "starts"])
@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_f64_check_build(cuda_lib_f64, ctrl, which):
    assert cuda_lib_f64.precision_bits == 64
    _check(cuda_lib_f64, ctrl, which, F64)


@pytest.mark.parametrize("which", ["trace", "starts"])
@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_f32_product_build(cuda_lib_f32, ctrl, which):
    assert cuda_lib_f32.precision_bits == 32
    w = _check(cuda_lib_f32, ctrl, which, F32)
    print(ctrl, which, {k: float("%.2g" % v) for k, v in w.items()})


def test_f32_single_mj_step(cuda_lib_f32):
    """north_star bar: ONE mj_step from identical state, qpos/qvel within 1e-5 relative in fp32."""
    from quadruped_rl_locomotion_b200.lib import Sim
    sub = np.load(os.path.join(pc.GOLDEN, "substeps_torque.npz"))
    cfg = cuda_lib_f32.default_config(); cfg.frame_skip = 1
    n = len(sub["qacc"])
    sim = Sim(pc.model_blob("torque"), n, lib=cuda_lib_f32, cfg=cfg)
    sim.enable_diag(True)
    for t in range(n):
        sim.set_state(t, qpos=sub["qpos"][t], qvel=sub["qvel"][t], qacc_warmstart=sub["warm"][t], step=0, unhealthy_steps=0)
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    sim.step_host(np.ascontiguousarray(sub["ctrl"].astype(np.float32)), obs, rew, done, trunc)
    e_acc, e_pos, e_vel = [], [], []
    for t in range(n):
        d = sim.get_diag(t); s = sim.get_state(t)
        nq = sub["qpos"][t + 1] if t + 1 < n else sub["qpos_final"]
        nv = sub["qvel"][t + 1] if t + 1 < n else sub["qvel_final"]
        e_acc.append(pc.rel_err(d["qacc"], sub["qacc"][t])); e_pos.append(pc.rel_err(s["qpos"], nq)); e_vel.append(pc.rel_err(s["qvel"], nv))
    print("single mj_step fp32: qacc median %.2e max %.2e | qpos max %.2e | qvel max %.2e" % (np.median(e_acc), max(e_acc), max(e_pos), max(e_vel)))
    assert max(e_pos) < 1e-5 and max(e_vel) < 1e-5 and max(e_acc) < 5e-5


def test_determinism_and_env_independence(cuda_lib_f32):
    """Identical inputs -> bit-identical outputs regardless of env index / batch composition."""
    pre, post = pc.load_starts("torque")
    idx = np.arange(0, 64)
    a = pc.run_teacher_forced(cuda_lib_f32, "torque", pre, post, idx)
    b = pc.run_teacher_forced(cuda_lib_f32, "torque", pre, post, idx[::-1].copy())
    for e in range(len(idx)):
        sa, sb = a["state"][e], b["state"][len(idx) - 1 - e]
        assert np.array_equal(sa["qpos"], sb["qpos"]) and np.array_equal(sa["qvel"], sb["qvel"])
    assert np.array_equal(a["rew"], b["rew"][::-1])


def test_nan_injection_is_isolated(cuda_lib_f32):
    """Fault injection: a NaN state in one env resets that env only; neighbours are bit-identical."""
    from quadruped_rl_locomotion_b200.lib import Sim
    pre, post = pc.load_starts("torque")
    idx = np.arange(0, 32)
    ref = pc.run_teacher_forced(cuda_lib_f32, "torque", pre, post, idx)
    pre2 = {k: v.copy() for k, v in pre.items()}
    pre2["qvel"][5, 3] = np.nan
    out = pc.run_teacher_forced(cuda_lib_f32, "torque", pre2, post, idx)
    assert out["done"][5] == 1 and out["state"][5]["bad_state"] == 1
    assert np.isfinite(out["obs"]).all() and np.isfinite(out["rew"]).all()
    for e in range(len(idx)):
        if e != 5:
            assert np.array_equal(out["state"][e]["qpos"], ref["state"][e]["qpos"])


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_biped_mode_f64(cuda_lib_f64, ctrl):
    """cfg.biped = 1 against the reference's Go1MujocoEnv(biped=True) (tests/golden/make_golden_biped.py): every pair."""
    pre, post = pc.load_starts(ctrl, biped=True)
    idx = np.arange(len(post["reward"]))
    out = pc.run_teacher_forced(cuda_lib_f64, ctrl, pre, post, idx, biped=True)
    out, idx = pc.drop_overflowed(out, idx)
    w, mism = pc.compare(out, pre, post, idx)
    for k, tol in F64.items():
        assert w[k] <= tol, (k, w[k])
    assert w["obs"] <= 2e-6 and w["reward"] <= 1e-5 and w["info"] <= 1e-6      # f32 output buffers
    assert all(v == 0 for v in mism.values()), mism
    worst, flags = pc.compare_biped(out, post, idx)
    assert worst.max() <= 1e-6, worst
    assert all(v == 0 for v in flags.values()), flags


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_biped_mode_f32(cuda_lib_f32, ctrl):
    pre, post = pc.load_starts(ctrl, biped=True)
    idx = np.arange(len(post["reward"]))
    out = pc.run_teacher_forced(cuda_lib_f32, ctrl, pre, post, idx, biped=True)
    out, idx = pc.drop_overflowed(out, idx)
    w, mism = pc.compare(out, pre, post, idx)
    print(ctrl, {k: float("%.2g" % v) for k, v in w.items()})
    # qacc of the last substep: the fixture's impacts (40 contacts, 1e3 N wrenches after a 0.65 m drop) are worse conditioned
    # than the walking fixtures; 1e-4 here, qpos / qvel keep the common tolerances
    tol32 = dict(F32, qacc=1e-4)
    for k, tol in tol32.items():
        if k != "reward":          # reward = max(0, 15 - 150 |cfrc|^2 ...): absolute error scales with the 1e6-sized cost terms
            assert w[k] <= tol, (k, w[k])
    assert all(v == 0 for v in mism.values()), mism
    worst, flags = pc.compare_biped(out, post, idx)
    print(ctrl, "biped terms worst rel", worst)
    assert worst.max() <= 5e-3, worst      # unwanted_contact = 150 |cfrc_ext|^2: twice the cfrc_ext tolerance
    assert all(v == 0 for v in flags.values()), flags
    ok = post["reward"][idx] > 0
    if ok.any():
        assert np.abs(out["rew"][ok] - post["reward"][idx][ok]).max() <= 1e-3


def test_biped_rollout_statistics(cuda_lib_f32):
    """4096 biped environments, zero actions for 30 control steps: every non-fallen env starts from the -95 deg
    rear-leg stance, falls from 0.65 m and the upright term decays; no bad states, terms stay finite."""
    from quadruped_rl_locomotion_b200.lib import Sim
    n = 4096
    cfg = cuda_lib_f32.default_config(); cfg.biped = 1
    sim = Sim(pc.model_blob("position"), n, lib=cuda_lib_f32, cfg=cfg, seed=5)
    obs0 = sim.reset_host()
    z0 = np.array([sim.get_state(e)["qpos"][2] for e in range(0, n, 64)])
    assert set(np.round(z0, 3)) <= {0.65, 0.1}
    act = np.tile(np.array([0, 2.8, -1.2] * 4, np.float32), (n, 1))
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    ups = []
    for _ in range(30):
        sim.step_host(act, obs, rew, done, trunc)
        b = sim.biped_info_host()
        assert np.isfinite(b).all() and np.all(b[:, 0] >= 0) and np.all(b[:, 0] <= 15.0 + 1e-4) and np.all(b[:, 1:10] <= 0)
        ups.append(b[:, 0].mean())
    assert ups[0] > 10.0            # 80 % of the slab starts with the trunk x-axis 5 deg from vertical
    assert not done.any() and np.isfinite(obs).all()
    sim.close()
