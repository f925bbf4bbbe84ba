"""Environment logic: the numpy restatement ``oracle/env_logic.py`` against the committed
golden vectors (which were produced by the reference's UNMODIFIED ``go1_mujoco_env.py`` running on the oracle
physics, ``tests/golden/make_golden.py``), and -- where ``/root/reference`` exists -- against the reference's
literal code live."""
import numpy as np
import pytest

import parity_common as pc
from oracle import env_logic as el
from oracle.ref_env import reference_available
from quadruped_rl_locomotion_b200.mjcf import Model


def _episode(pre, t):
    return el.EpisodeState(
        step=int(pre["ep_step"][t]), last_action=pre["ep_last_action"][t].copy(),
        feet_air_time=pre["ep_feet_air_time"][t].copy(), last_contacts=pre["ep_last_contacts"][t] > 0,
        desired_velocity=pre["ep_desired_velocity"][t].copy(), time_unhealthy=float(pre["ep_time_unhealthy"][t]),
        last_dev=pre["ep_last_dev"][t].copy())


@pytest.mark.parametrize("which", ["trace", "starts"])
@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_restatement_matches_reference_goldens(ctrl, which):
    """obs / reward / 8 info terms / health / termination / episode bookkeeping, every golden step."""
    m = Model.from_blob(pc.model_blob(ctrl))
    logic = el.EnvLogic(m)
    pre, post = pc.load_trace(ctrl) if which == "trace" else pc.load_starts(ctrl)
    n = len(post["reward"])
    for t in range(n):
        ep = _episode(pre, t)
        obs, rew, term, trunc, info = logic.step(ep, post["action"][t], post["qpos"][t], post["qvel"][t], post["qacc"][t],
                                                 post["qfrc_actuator"][t], post["cfrc_ext"][t], post["trunk_z"][t])
        assert np.allclose(obs, post["obs"][t], rtol=0, atol=1e-12), t
        assert rew == pytest.approx(float(post["reward"][t]), rel=1e-8, abs=1e-8), t    # fp64 rounding: summation order, qacc^2 / (|qvel| + 1e-6)
        assert np.allclose([info[k] for k in el.INFO_KEYS], post["info"][t], rtol=0, atol=1e-10), t
        assert bool(info["healthy"]) == bool(post["healthy"][t])
        assert bool(term) == bool(post["terminated"][t]) and bool(trunc) == bool(post["truncated"][t])
        assert np.allclose(ep.feet_air_time, post["ep_feet_air_time"][t], atol=1e-12)
        assert np.array_equal(ep.last_contacts, post["ep_last_contacts"][t] > 0)
        assert ep.time_unhealthy == pytest.approx(float(post["ep_time_unhealthy"][t]), abs=1e-9)
        assert np.allclose(ep.last_dev, post["ep_last_dev"][t], atol=1e-12)


def test_projected_gravity_and_euler_fixture_properties():
    # reference quirk kept: "projected gravity" is (g . euler) * euler normalised (go1_mujoco_env.py:596-608)
    assert np.array_equal(el.projected_gravity(np.array([1.0, 0, 0, 0])), np.zeros(3))     # zero euler -> zeros branch
    q = np.array([0.9, 0.1, -0.3, 0.2]); q /= np.linalg.norm(q)
    pg = el.projected_gravity(q)
    assert np.linalg.norm(pg) == pytest.approx(1.0)
    e = np.array(el.euler_from_quaternion(*q))
    assert np.allclose(np.cross(pg, e), 0, atol=1e-12)            # parallel to the euler vector
    # asin clamp (:1027-1029)
    r, p, y = el.euler_from_quaternion(np.sqrt(0.5), 0, np.sqrt(0.5) + 1e-9, 0)
    assert p == pytest.approx(np.pi / 2)


def test_health_ranges():
    qv = np.zeros(18)
    q = np.array([0, 0, 0.3, 1, 0, 0, 0] + [0] * 12, dtype=float)
    assert el.health(q, qv) == (True, 0)
    q2 = q.copy(); q2[2] = 0.21
    assert el.health(q2, qv) == (False, 2)
    q3 = q.copy(); q3[4] = 1.4
    assert el.health(q3, qv) == (False, 3)
    q4 = q.copy(); q4[5] = 0.01                 # healthy "pitch" range is [-pi, 0] on the raw quaternion y (:119)
    assert el.health(q4, qv) == (False, 4)
    q5 = q.copy(); q5[0] = np.nan
    assert el.health(q5, qv) == (False, 1)


def test_termination_needs_751_unhealthy_steps():
    # terminated = time_unhealthy > 15.0 with 0.02 increments (:487-496): exact count under fp64 accumulation
    t, n = 0.0, 0
    while not t > el.MAX_UNHEALTHY_TIME:
        t += el.DT; n += 1
    assert n == 751
    assert el.MAX_EPISODE_STEPS == 6000


@pytest.mark.skipif(not reference_available(), reason="reference sources not on this machine")
def test_live_reference_env_equals_restatement():
    """The reference's own class, imported unmodified on shim modules, stepped beside the restatement."""
    from oracle.ref_env import make_reference_env
    env = make_reference_env("torque")
    np.random.seed(5)
    env.reset()
    logic = el.EnvLogic(env.sim.model if hasattr(env.sim, "model") else Model.from_blob(pc.model_blob("torque")))
    ep = el.EpisodeState(desired_velocity=np.array(env._desired_velocity, dtype=float))
    rng = np.random.default_rng(9)
    for _ in range(25):
        a = rng.uniform(-1, 1, 12).astype(np.float32)
        obs, rew, term, trunc, info = env.step(a)
        d = env.data
        o2, r2, t2, tr2, i2 = logic.step(ep, a, d.qpos.copy(), d.qvel.copy(), d.qacc.copy(), d.qfrc_actuator.copy(),
                                         d.cfrc_ext.copy(), d.xpos[1, 2])
        assert np.allclose(obs, o2, atol=1e-12) and rew == pytest.approx(r2, abs=1e-10)
        assert (term, trunc) == (t2, tr2)
        for k in el.INFO_KEYS:
            assert info[k] == pytest.approx(i2[k], abs=1e-10)


def biped_kin(d):
    """The kinematic quantities the biped terms read (go1_mujoco_env.py:322,340,360-361,379,409-410), MuJoCo body ids."""
    ids = {"FR_hip": 2, "FR_calf": 4, "FL_hip": 5, "FL_calf": 7, "RR_hip": 8, "RL_hip": 11}
    return dict(trunk_xaxis_z=d.xmat[1].reshape(3, 3)[2, 0],
                front_calf_z=d.xpos[[ids["FR_calf"], ids["FL_calf"]], 2], front_feet_z=d.site_xpos[[2, 3], 2],
                rear_hip_y=d.xpos[[ids["RR_hip"], ids["RL_hip"]], 1], rear_hip_z=d.xpos[[ids["RR_hip"], ids["RL_hip"]], 2],
                front_hip_z=d.xpos[[ids["FR_hip"], ids["FL_hip"]], 2])


@pytest.mark.skipif(not reference_available(), reason="reference sources not on this machine")
@pytest.mark.parametrize("seed", [5, 6])       # seed 6 draws a fallen start
def test_live_reference_env_biped_equals_restatement(seed):
    """``biped=True`` (go1_mujoco_env.py:91-101, :869-904): the extra terms, the alternating-rear-feet air-time
    reward, the front-feet-touched latch and the biped reset pose."""
    from oracle.ref_env import make_reference_env
    env = make_reference_env("torque", biped=True)
    m = env.sim.model
    assert [m.names["body"][i] for i in el.UNWANTED_CONTACT_BODIES] == ["trunk", "FR_thigh", "FL_thigh", "RR_thigh", "RL_thigh", "FR_calf", "FL_calf"]
    assert [m.names["body"][i] for i in (2, 4, 5, 7, 8, 11)] == ["FR_hip", "FR_calf", "FL_hip", "FL_calf", "RR_hip", "RL_hip"]
    assert list(env._front_feet_site_ids) == [2, 3]      # sites: head, imu, FR, FL, RR, RL
    np.random.seed(seed)
    env.reset()
    if abs(env.data.qpos[2] - 0.1) > 1e-9:
        assert np.allclose(env.data.qpos[7:], el.BIPEDAL_READY_JOINTS) and env.data.qpos[2] == 0.65
    logic = el.EnvLogic(m, biped=True)
    ep = el.EpisodeState(desired_velocity=np.array(env._desired_velocity, dtype=float))
    rng = np.random.default_rng(9)
    prev_ff = np.zeros(2)
    for _ in range(40):
        a = rng.uniform(-1, 1, 12).astype(np.float32)
        obs, rew, term, trunc, info = env.step(a)
        d = env.data
        o2, r2, t2, tr2, i2 = logic.step(ep, a, d.qpos.copy(), d.qvel.copy(), d.qacc.copy(), d.qfrc_actuator.copy(),
                                         d.cfrc_ext.copy(), d.xpos[1, 2], kin=biped_kin(d), prev_front_feet_forces=prev_ff,
                                         n_self_collisions=el.count_self_collisions(d.contact, env.model.geom_bodyid))
        prev_ff = np.linalg.norm(d.cfrc_ext[[4, 7]], axis=1)
        assert np.allclose(obs, o2, atol=1e-12) and rew == pytest.approx(r2, rel=1e-9, abs=1e-9)
        assert (term, trunc) == (t2, tr2)
        assert env._front_feet_touched == ep.front_feet_touched
        for k in el.BIPED_INFO_KEYS[:-1] + ("linear_vel_tracking_reward", "reward_ctrl", "get_up_reward"):
            assert info[k] == pytest.approx(i2[k], rel=1e-9, abs=1e-9), k
        assert info["bipedal_success"] == i2["bipedal_success"]
