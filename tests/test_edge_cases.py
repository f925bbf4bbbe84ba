"""Edge cases of the C ABI: masked resets, invalid sizes / configs, null buffers, out-of-range env indices, a zero-length
rollout, long quiet standing (no contact changes), and the action clamp the actuators apply (ctrlrange)."""
import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200 import lib as L
from quadruped_rl_locomotion_b200.lib import ACTION_ZERO, Sim


def test_masked_reset_touches_only_masked_envs(emu_lib_f32):
    n = 7
    sim = Sim(pc.model_blob("torque"), n, lib=emu_lib_f32, seed=2)
    sim.reset_host()
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    rng = np.random.default_rng(0)
    for _ in range(3):
        sim.step_host(rng.uniform(-1, 1, (n, 12)).astype(np.float32), obs, rew, done, trunc)
    before = [sim.get_state(e) for e in range(n)]
    mask = np.array([0, 1, 0, 0, 1, 0, 1], np.uint8)
    robs = np.full((n, 48), 7.0, np.float32)
    sim.reset_host(robs, mask)
    for e in range(n):
        s = sim.get_state(e)
        if mask[e]:
            assert s["step"] == 0 and np.all(s["qvel"] == 0) and s["reset_count"] == before[e]["reset_count"] + 1
            assert np.all(robs[e, 0:6] == 0) and np.all(robs[e, 36:48] == 0)          # reset observation: zero velocities / last action
        else:
            for k in ("qpos", "qvel", "qacc_warmstart", "last_action", "feet_air_time"):
                assert np.array_equal(s[k], before[e][k]), (e, k)
            assert s["step"] == 3 and s["reset_count"] == before[e]["reset_count"]
            assert np.all(robs[e] == 7.0)                                               # rows of unmasked envs are not written
    sim.close()


def test_invalid_arguments_are_refused(emu_lib_f32):
    blob = pc.model_blob("torque")
    with pytest.raises(L.Go1Error):
        Sim(blob, 0, lib=emu_lib_f32)
    cfg = emu_lib_f32.default_config(); cfg.frame_skip = 0
    with pytest.raises(L.Go1Error):
        Sim(blob, 2, lib=emu_lib_f32, cfg=cfg)
    with pytest.raises(L.Go1Error, match="model"):
        Sim(blob[: len(blob) // 2], 2, lib=emu_lib_f32)                                # truncated blob
    sim = Sim(blob, 2, lib=emu_lib_f32)
    for bad in (-1, 2, 1000):
        with pytest.raises(L.Go1Error):
            sim.get_state(bad)
        with pytest.raises(L.Go1Error):
            sim.get_diag(bad)
    obs = np.zeros((2, 48), np.float32)
    rc = emu_lib_f32.L.go1_step_host(sim._h, None, obs.ctypes.data, None, None, None, None, None)
    assert rc == -1                                                                     # GO1_EINVAL, nothing dereferenced
    with pytest.raises(L.Go1Error):
        sim.get_attr("no_such_attribute")
    st = sim.rollout(0, ACTION_ZERO, None)
    assert st["env_steps"] == 0 and st["launches"] == 0
    sim.close()


def test_actions_outside_ctrlrange_act_as_clamped(emu_lib_f64):
    """MuJoCo clamps ctrl to actuator_ctrlrange (ctrllimited): +-1000 must give exactly the physics of the range ends."""
    blob = pc.model_blob("torque")
    a = Sim(blob, 1, lib=emu_lib_f64, seed=1); b = Sim(blob, 1, lib=emu_lib_f64, seed=1)
    lo, hi = a.action_bounds()
    cfg_state = a.get_state(0)
    b.set_state(0, qpos=cfg_state["qpos"], qvel=cfg_state["qvel"])
    obs = np.zeros((1, 48), np.float32); rew = np.zeros(1, np.float32); done = np.zeros(1, np.uint8); trunc = np.zeros(1, np.uint8)
    big = np.where(np.arange(12) % 2 == 0, 1000.0, -1000.0).astype(np.float32)[None]
    clamped = np.clip(big, lo, hi).astype(np.float32)
    a.step_host(big, obs, rew, done, trunc); b.step_host(clamped, obs, rew, done, trunc)
    sa, sb = a.get_state(0), b.get_state(0)
    assert np.array_equal(sa["qpos"], sb["qpos"]) and np.array_equal(sa["qvel"], sb["qvel"])
    assert np.array_equal(sa["last_action"], big[0].astype(np.float64))                 # the env remembers the RAW action (:529)
    a.close(); b.close()


def test_quiet_standing_is_stationary(emu_lib_f64):
    """Position control at the home keyframe: after settling for 8 s, 50 more control steps barely move the robot
    (same four foot contacts, at most two Newton iterations per substep) and the reward stays put."""
    cfg = emu_lib_f64.default_config(); cfg.fallen_start_prob = 0.0
    sim = Sim(pc.model_blob("position"), 1, lib=emu_lib_f64, cfg=cfg, seed=0)
    sim.enable_diag(True)
    sim.reset_host()
    act = np.array([[0, 0.9, -1.8] * 4], np.float32)
    obs = np.zeros((1, 48), np.float32); rew = np.zeros(1, np.float32); done = np.zeros(1, np.uint8); trunc = np.zeros(1, np.uint8)
    for _ in range(400):
        sim.step_host(act, obs, rew, done, trunc)
    q0 = sim.get_state(0)["qpos"].copy(); r0 = float(rew[0])
    iters = []
    for _ in range(50):
        sim.step_host(act, obs, rew, done, trunc)
        iters.append(sim.get_diag(0)["solver_iter_total"])
    s = sim.get_state(0)
    assert np.abs(s["qpos"] - q0).max() < 2e-3 and np.abs(s["qvel"]).max() < 5e-3      # soft contacts + joint friction: slow creep only
    assert sim.get_diag(0)["ncon"] == 4 and max(iters) <= 20            # <= 2 Newton iterations per substep
    assert abs(float(rew[0]) - r0) < 1e-2 and not done[0]
    sim.close()
