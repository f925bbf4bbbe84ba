"""``Go1CudaVecEnv`` against the SB3 2.3.0 ``VecEnv`` contract the reference's PPO loop relies on (SURVEY 8(b);
reference call sites ``train.py:44-50``, ``test_max_envs.py:35-65``, ``training_callback.py:64``).

Runs on the CPU box through the SIMT emulation build of the SAME CUDA source (tests/emu); the GPU version of
the same checks is ``test_vec_env_gpu`` at the bottom (marked gpu, real library)."""
import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200 import lib as L
from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv, make_go1_vec_env


def _contract(env, n):
    assert env.num_envs == n
    assert env.observation_space.shape == (48,) and env.observation_space.dtype == np.float64
    assert env.action_space.shape == (12,) and env.action_space.dtype == np.float32
    obs = env.reset()
    assert obs.shape == (n, 48) and obs.dtype == np.float64 and np.isfinite(obs).all()
    # obs layout (go1_mujoco_env.py:915-947): [9:12] = desired velocity * 2 in x only; [36:48] = last action = 0 after reset
    assert np.all(np.abs(obs[:, 9]) <= 1.0) and np.all(obs[:, 10:12] == 0) and np.all(obs[:, 36:48] == 0)
    # qvel = 0 at reset -> first 6 and joint-velocity block are exactly zero
    assert np.all(obs[:, 0:6] == 0) and np.all(obs[:, 24:36] == 0)
    a = np.stack([env.action_space.sample() for _ in range(n)])
    assert all(env.action_space.contains(x) for x in a)
    obs, rew, dones, infos = env.step(a)
    assert obs.shape == (n, 48) and rew.shape == (n,) and rew.dtype == np.float32
    assert dones.shape == (n,) and dones.dtype == bool and len(infos) == n
    assert np.all(rew >= 0)                                   # reward = max(0, .) (go1_mujoco_env.py:911)
    # the obs is built BEFORE _last_action is updated (go1_mujoco_env.py:483 vs :529): slot [36:48] lags one step
    assert np.all(obs[:, 36:48] == 0)
    obs2, _, dones2, _ = env.step(a)
    keep = ~(dones | dones2)
    assert np.allclose(obs2[keep, 36:48], a[keep], atol=1e-6)
    with pytest.raises(ValueError):
        env.step_async(np.zeros((n, 11), np.float32))         # reference: "Action dimension mismatch"
    assert env.get_attr("render_mode") == [None] * n
    assert env.env_is_wrapped(object) == [False] * n
    assert env.render() is None and env.get_images() == [None] * n
    env.set_attr("_rand_power", 0.25)                         # CurriculumCallback (training_callback.py:64)
    assert env.get_attr("_rand_power") == [0.25] * n
    assert env.seed(3) == [3] * n


def _auto_reset(env, n):
    env.reset()
    sim = env.sim
    # env 1: one step from the 6000-step truncation; env 2: one step from the 751-step unhealthy termination
    s = sim.get_state(1)
    sim.set_state(1, step=5999)
    q = sim.get_state(2)["qpos"].copy(); q[2] = 0.1            # z below healthy_z_range -> unhealthy
    sim.set_state(2, qpos=q, unhealthy_steps=750)
    env.full_info = True
    obs, rew, dones, infos = env.step(np.zeros((n, 12), np.float32))
    st1 = sim.get_state(1)          # state right after the auto-reset
    assert dones[1] and dones[2] and not dones[0]
    assert infos[1]["TimeLimit.truncated"] is True and infos[2]["TimeLimit.truncated"] is False
    for i in (1, 2):
        assert infos[i]["terminal_observation"].shape == (48,)
        assert set(infos[i]["episode"]) == {"r", "l", "t"}
        # returned obs of a finished env is the RESET observation: zero velocities, zero last action
        assert np.all(obs[i, 0:6] == 0) and np.all(obs[i, 36:48] == 0)
        assert not np.array_equal(obs[i], infos[i]["terminal_observation"])
    assert infos[1]["episode"]["l"] == 6000 and infos[2]["termination_reason"] == "prolonged_unhealthy_z"
    # termination_details (go1_mujoco_env.py:444-447, :513): the offending value and the healthy range, then the persistence note
    z2 = float(infos[2]["termination_details"].split("Z-position: ")[1].split(",")[0])
    assert infos[2]["termination_details"].endswith("Healthy Range: [0.22, 1.80] (persisted for > 15.00s)") and z2 < 0.22
    assert infos[0]["termination_details"] in ("No termination",) or infos[0]["termination_reason"] == "unhealthy_state_active"
    # every step hands out its OWN info objects: a wrapper that mutates them cannot leak into the next step or another env
    infos[0]["leak"] = 1
    _, _, _, infos_next = env.step(np.zeros((n, 12), np.float32))
    assert "leak" not in infos_next[0] and infos[0]["leak"] == 1 and infos_next[0] is not infos[0]
    assert len(infos_next) == n and isinstance(infos_next, list) and len(infos_next[:2]) == 2
    # the float64 observation arrays alternate: the previous step's array is intact while the next step is returned
    o1, _, _, _ = env.step(np.zeros((n, 12), np.float32)); keep = o1.copy()
    o2, _, _, _ = env.step(np.zeros((n, 12), np.float32))
    assert o1 is not o2 and np.array_equal(o1, keep) and o1.dtype == np.float64
    with pytest.raises(ValueError):
        env.set_attr("_rand_power", 0.5, indices=[0])          # one value per slab: per-index values are refused, not ignored
    assert env.env_method("get_wrapper_attr", "_rand_power") == env.get_attr("_rand_power")
    assert "terminal_observation" not in infos[0]
    for k in ("x_position", "y_position", "distance_from_origin", "linear_vel_tracking_reward", "reward_ctrl",
              "reward_survive", "recovery_reward", "get_up_reward", "unhealthy_state_cost", "orientation_cost",
              "default_joint_position_cost", "termination_reason", "bipedal_success"):
        assert k in infos[0], k                                # go1_mujoco_env.py:500-517, 860-909
    assert st1["step"] == 0 and st1["reset_count"] >= 1 and np.all(st1["last_action"] == 0)


def test_make_go1_vec_env_signature(emu_lib_f32):
    """``make_vec_env(Go1MujocoEnv, env_kwargs=..., n_envs=..., seed=..., vec_env_cls=SubprocVecEnv)`` (train.py:44-50) with the
    env class positional and SB3's keyword set."""
    from quadruped_rl_locomotion_b200.vec_env import make_go1_vec_env
    env = make_go1_vec_env(object, env_kwargs={"ctrl_type": "position"}, n_envs=3, seed=5, vec_env_cls=object, library=emu_lib_f32)
    assert env.num_envs == 3 and env.ctrl_type == "position"
    env.close()
    with pytest.raises(ValueError):
        make_go1_vec_env(object, n_envs=2, wrapper_class=object, library=emu_lib_f32)


def test_contract_emu(emu_lib_f32):
    env = Go1CudaVecEnv(6, ctrl_type="torque", library=emu_lib_f32, seed=1)
    _contract(env, 6)
    env.close()


def test_position_ctrl_spaces_emu(emu_lib_f32):
    env = make_go1_vec_env(n_envs=4, env_kwargs={"ctrl_type": "position"}, seed=2, library=emu_lib_f32)
    lo, hi = env.action_space.low, env.action_space.high
    assert np.allclose(lo.reshape(4, 3), [[-0.863, -0.686, -2.818]] * 4) and np.allclose(hi.reshape(4, 3), [[0.863, 4.501, -0.888]] * 4)
    _contract(env, 4)
    env.close()


def test_auto_reset_emu(emu_lib_f32):
    env = Go1CudaVecEnv(4, library=emu_lib_f32, seed=3)
    _auto_reset(env, 4)
    env.close()


def test_biped_vec_env_emu(emu_lib_f32):
    """Go1MujocoEnv(biped=True) through the VecEnv: the biped reward_info keys replace orientation / default joint
    position (go1_mujoco_env.py:893-909); bipedal_success is reported on a time-limit truncation without an earlier
    front-foot touch (:520-521) and not after one."""
    from quadruped_rl_locomotion_b200.vec_env import BIPED_INFO_KEYS
    n = 4
    env = Go1CudaVecEnv(n, biped=True, library=emu_lib_f32, seed=3, full_info=True)
    assert env.get_attr("biped") == [True] * n
    env.reset()
    sim = env.sim
    sim.set_state(1, step=5999, front_feet_touched=0)
    sim.set_state(2, step=5999, front_feet_touched=1)
    obs, rew, dones, infos = env.step(np.zeros((n, 12), np.float32))
    assert dones[1] and dones[2] and infos[1]["TimeLimit.truncated"] and infos[2]["TimeLimit.truncated"]
    assert infos[1]["bipedal_success"] is True and infos[2]["bipedal_success"] is False and infos[0]["bipedal_success"] is False
    for k in BIPED_INFO_KEYS:
        assert k in infos[0], k
    assert "orientation_cost" not in infos[0] and "default_joint_position_cost" not in infos[0]
    assert 0.0 <= infos[0]["biped_upright_reward"] <= 15.0 and infos[0]["self_collision_cost"] == 0.0
    assert sim.get_state(1)["front_feet_touched"] == 0 and sim.get_state(1)["step"] == 0      # auto-reset clears the latch
    env.close()
    env = Go1CudaVecEnv(2, library=emu_lib_f32)
    with pytest.raises(Exception, match="biped"):
        env.sim.biped_info_host()
    env.close()


def test_reset_distribution_emu(emu_lib_f32):
    """reset_model (go1_mujoco_env.py:949-1001): 20 % fallen starts (z = 0.1, roll or pitch 50/50, angle in
    [pi/2.1, pi/1.5]), else the home keyframe; desired velocity U[-0.5, 0.5] in x only (:1011-1015)."""
    n = 2000
    sim = L.Sim(pc.model_blob("torque"), n, lib=emu_lib_f32, seed=11)
    sim.reset_host()
    q = np.array([sim.get_state(e)["qpos"] for e in range(n)])
    v = np.array([sim.get_state(e)["desired_velocity"] for e in range(n)])
    fallen = np.abs(q[:, 2] - 0.1) < 1e-6
    assert abs(fallen.mean() - 0.2) < 0.03
    assert np.allclose(q[~fallen, 2], 0.30) and np.allclose(q[~fallen, 3:7], [1, 0, 0, 0])
    ang = 2 * np.arccos(np.clip(q[fallen, 3], -1, 1))
    assert ang.min() >= np.pi / 2.1 - 1e-5 and ang.max() <= np.pi / 1.5 + 1e-5
    roll = np.abs(q[fallen, 4]) > 1e-6
    assert np.all(roll ^ (np.abs(q[fallen, 5]) > 1e-6)) and abs(roll.mean() - 0.5) < 0.08
    assert np.all(q[fallen, 6] == 0)
    assert v[:, 0].min() >= -0.5 and v[:, 0].max() <= 0.5 and np.all(v[:, 1:] == 0)
    assert abs(v[:, 0].mean()) < 0.03 and abs(v[:, 0].std() - 1 / np.sqrt(12)) < 0.02
    assert np.allclose(q[:, 7:], np.tile([0, 0.9, -1.8], 4))          # rand_power = 0: no joint noise
    sim.close()


def test_slab_offset_invariance_emu(emu_lib_f32):
    """RNG stream = GLOBAL env id: env k of a slab created with offset o behaves exactly like env
    o + k of one big slab -- the result does not depend on how environments are sharded over GPUs."""
    blob = pc.model_blob("torque")
    big = L.Sim(blob, 8, lib=emu_lib_f32, seed=5)
    lo = L.Sim(blob, 4, lib=emu_lib_f32, seed=5, env_id_offset=0)
    hi = L.Sim(blob, 4, lib=emu_lib_f32, seed=5, env_id_offset=4)
    for s in (big, lo, hi):
        s.reset_host()
    for _ in range(3):
        sb, sl, sh = big.rollout(1), lo.rollout(1), hi.rollout(1)
    for k in range(4):
        for key in ("qpos", "qvel", "desired_velocity"):
            assert np.array_equal(big.get_state(k)[key], lo.get_state(k)[key])
            assert np.array_equal(big.get_state(4 + k)[key], hi.get_state(k)[key])
    assert sb["sum_reward"] == pytest.approx(sl["sum_reward"] + sh["sum_reward"], rel=1e-6)
    for s in (big, lo, hi):
        s.close()


@pytest.mark.gpu
def test_vec_env_gpu(cuda_lib_f32):
    env = Go1CudaVecEnv(64, library=cuda_lib_f32, seed=1)
    _contract(env, 64)
    env.close()
    env = Go1CudaVecEnv(8, library=cuda_lib_f32, seed=3)
    _auto_reset(env, 8)
    env.close()


def test_reset_joint_noise_emu(emu_lib_f32):
    """rand_power > 0 (go1_mujoco_env.py:974-982): joint noise N(0, 0.1 * rand_power) on the keyframe joints, clipped
    to jnt_range; the curriculum callback changes it through set_attr (training_callback.py:64)."""
    n = 3000
    cfg = emu_lib_f32.default_config(); cfg.rand_power = 1.0; cfg.fallen_start_prob = 0.0
    sim = L.Sim(pc.model_blob("position"), n, lib=emu_lib_f32, cfg=cfg, seed=21)
    sim.reset_host()
    q = np.array([sim.get_state(e)["qpos"][7:] for e in range(n)])
    home = np.tile([0, 0.9, -1.8], 4)
    lo, hi = sim.action_bounds()                      # position model: ctrlrange == jnt_range
    d = q - home
    assert np.all(q >= lo - 1e-6) and np.all(q <= hi + 1e-6)
    free = [0, 3, 6, 9, 1, 4, 7, 10]                  # abduction / hip joints: the range is far from the keyframe value
    assert abs(d[:, free].mean()) < 0.01 and abs(d[:, free].std() - 0.1) < 0.01
    assert abs(np.corrcoef(d[:, 0], d[:, 1])[0, 1]) < 0.08
    sim.set_attr("rand_power", 0.5)
    sim.reset_host()
    q2 = np.array([sim.get_state(e)["qpos"][7:] for e in range(n)])
    assert abs((q2 - home)[:, free].std() - 0.05) < 0.006
    sim.close()


@pytest.mark.gpu
def test_registered_buffers_gpu(cuda_lib_f32):
    """Real cudaHostRegister path: VecEnv (page-locked arrays, direct DMA) == Sim.step_host on pageable arrays."""
    _registered_vs_pageable(cuda_lib_f32, 1000)


def test_registered_and_pageable_buffers_give_identical_steps(emu_lib_f32):
    _registered_vs_pageable(emu_lib_f32, 5)


def _registered_vs_pageable(emu_lib_f32, n):
    """go1_step_host with caller buffers registered through go1_host_register (the VecEnv's own arrays) and with
    plain pageable arrays (Sim.step_host from a test) must return the same bytes; partial overlap with a registered
    range falls back to staging."""
    from quadruped_rl_locomotion_b200.lib import Sim
    env = Go1CudaVecEnv(n, library=emu_lib_f32, seed=4)
    assert len(env._pinned) == 9          # actions, obs, terminal obs, reward, done, trunc, info + the two float64 observation batches
    sim2 = Sim(pc.model_blob("torque"), n, lib=emu_lib_f32, seed=4)
    env.reset(); sim2.reset_host()
    rng = np.random.default_rng(0)
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    for _ in range(3):
        a = rng.uniform(-1, 1, (n, 12)).astype(np.float32)
        o1, r1, d1, _ = env.step(a)
        sim2.step_host(a, obs, rew, done, trunc)
        assert np.array_equal(o1.astype(np.float32), obs) and np.array_equal(r1, rew) and np.array_equal(d1, done.astype(bool))
    # go1_step_host_f64 (the VecEnv's path, registered float64 arrays above) through a pageable float64 array: staged, same values
    o64 = np.full((n, 48), np.nan)
    env.sim.step_host_f64(a, o64, rew, done, trunc); sim2.step_host(a, obs, rew, done, trunc)
    assert o64.dtype == np.float64 and np.array_equal(o64, obs.astype(np.float64))
    big = np.zeros((n + 1, 48), np.float32)
    env.sim.host_register(big)
    env.sim.step_host(a, big[1:], rew, done, trunc)          # inside the registered range, not at its start
    sim2.step_host(a, obs, rew, done, trunc)
    assert np.array_equal(big[1:], obs)
    env.sim.host_unregister(big)
    with pytest.raises(Exception):
        env.sim.host_unregister(big)
    env.close(); sim2.close()
