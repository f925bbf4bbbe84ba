"""BASELINE full sizes on the GPU (config 2: position control, 4096 envs; config 3: torque control, 16384 envs)
through size-independent properties: the oracle cannot step 16k environments in seconds, so at these sizes the
checks are invariants every correct mj_step / env step satisfies, determinism, independence of the batch
composition, consistency of the device-side statistics with the per-env outputs, and a sampled comparison with
the fp64 oracle (64 of the 16384 environments, free-running for 3 control steps)."""
import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM, Sim

pytestmark = pytest.mark.gpu


def _step_n(sim, n_steps, seed):
    n = sim.n_envs
    lo, hi = sim.action_bounds()
    rng = np.random.default_rng(seed)
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32)
    done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    tobs = np.zeros((n, 48), np.float32); info = np.zeros((n, 20), np.float32)
    hist = []
    for _ in range(n_steps):
        a = rng.uniform(lo, hi, size=(n, 12)).astype(np.float32)
        sim.step_host(a, obs, rew, done, trunc, tobs, info)
        hist.append((a, obs.copy(), rew.copy(), done.copy(), info.copy()))
    return hist


@pytest.mark.parametrize("ctrl,n", [("position", 4096), ("torque", 16384)])
def test_invariants_at_full_size(cuda_lib_f32, ctrl, n):
    sim = Sim(pc.model_blob(ctrl), n, lib=cuda_lib_f32, seed=42)
    sim.reset_host()
    hist = _step_n(sim, 25, seed=1)
    for a, obs, rew, done, info in hist:
        assert np.isfinite(obs).all() and np.isfinite(rew).all()
        assert np.all(np.abs(obs) <= 100.0) and np.all(rew >= 0)          # obs clip (:945), reward = max(0, .) (:911)
        assert np.all(np.abs(obs[:, 9]) <= 1.0 + 1e-6) and np.all(obs[:, 10:12] == 0)   # desired velocity * 2, x only
    idx = np.linspace(0, n - 1, 97).astype(int)
    for e in idx:
        s = sim.get_state(int(e))
        assert abs(np.linalg.norm(s["qpos"][3:7]) - 1.0) < 1e-5          # quaternion renormalised every substep
        assert s["step"] == 25 and s["bad_state"] == 0
        assert np.all(np.abs(s["qvel"]) < 1e3)
    # last_action slot of the next observation == action applied one step earlier (:483 vs :529)
    a_prev = hist[-2][0]; obs_last = hist[-1][1]; done_any = hist[-1][3].astype(bool) | hist[-2][3].astype(bool)
    assert np.allclose(obs_last[~done_any, 36:48], a_prev[~done_any], atol=1e-6)
    sim.close()


def test_determinism_and_batch_independence_16k(cuda_lib_f32):
    """Same seed -> bit-identical slab; env k of a 16384 slab == env k of a 512 slab (same global id, same inputs)."""
    blob = pc.model_blob("torque")
    a = Sim(blob, 16384, lib=cuda_lib_f32, seed=7); b = Sim(blob, 16384, lib=cuda_lib_f32, seed=7)
    c = Sim(blob, 512, lib=cuda_lib_f32, seed=7)
    for s in (a, b, c):
        s.reset_host()
        for _ in range(6):
            s.rollout(1, ACTION_RANDOM, None)                           # device-side Philox actions keyed by global env id
    for e in (0, 1, 37, 255, 511):
        sa, sb, sc = a.get_state(e), b.get_state(e), c.get_state(e)
        for k in ("qpos", "qvel", "qacc_warmstart", "last_action"):
            assert np.array_equal(sa[k], sb[k]) and np.array_equal(sa[k], sc[k]), (e, k)
    for e in (4096, 16383):
        assert np.array_equal(a.get_state(e)["qpos"], b.get_state(e)["qpos"])
    for s in (a, b, c):
        s.close()


def test_device_statistics_match_outputs(cuda_lib_f32):
    """go1_rollout's on-device reduction (the 8-vector the ranks all-reduce) == sums of the per-env outputs."""
    blob = pc.model_blob("torque")
    n = 16384
    a = Sim(blob, n, lib=cuda_lib_f32, seed=3); b = Sim(blob, n, lib=cuda_lib_f32, seed=3)
    a.reset_host(); b.reset_host()
    st = a.rollout(4, ACTION_RANDOM, None)
    assert st["env_steps"] == 4 * n and st["launches"] == 8 and st["n_bad_state"] == 0
    tot = 0.0
    for _ in range(4):
        s1 = b.rollout(1, ACTION_RANDOM, None)
        tot += s1["sum_reward"]
    assert st["sum_reward"] == pytest.approx(tot, rel=1e-9)
    a.close(); b.close()


def test_sampled_envs_follow_the_oracle(cuda_lib_f32):
    """64 environments of a 16384 slab, free-running (no teacher forcing) for 3 control steps from the home keyframe,
    against the fp64 oracle stepping the same actions: chaos has not set in yet, positions agree to 1e-4."""
    from oracle.oracle import OracleSim
    from quadruped_rl_locomotion_b200.mjcf import Model
    blob = pc.model_blob("torque")
    model = Model.from_blob(blob)
    n = 16384
    cfg = cuda_lib_f32.default_config(); cfg.fallen_start_prob = 0.0
    sim = Sim(blob, n, lib=cuda_lib_f32, cfg=cfg, seed=5)
    sim.reset_host()
    rng = np.random.default_rng(0)
    pick = rng.choice(n, size=64, replace=False)
    oracles = {int(e): OracleSim(model) for e in pick}
    for o in oracles.values():
        o.data.qpos[:] = model.key_qpos[0]
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    for _ in range(3):
        a = rng.uniform(-1, 1, size=(n, 12)).astype(np.float32)
        sim.step_host(a, obs, rew, done, trunc)
        for e, o in oracles.items():
            o.data.ctrl[:] = a[e]
            o.step(10)
    worst = 0.0
    for e, o in oracles.items():
        s = sim.get_state(e)
        worst = max(worst, float(np.abs(s["qpos"] - o.data.qpos).max()))
    assert worst < 1e-4, worst
    sim.close()


@pytest.mark.parametrize("n", [1, 13, 3553])
def test_ragged_slab_sizes(cuda_lib_f32, n):
    """Slab sizes that are not a multiple of the 12 environments a block pulls from the queue (1 env, one env more
    than a block, one env more than a full wave of 296 blocks): env k equals env k of a 16-env slab, bit for bit."""
    blob = pc.model_blob("torque")
    a = Sim(blob, n, lib=cuda_lib_f32, seed=9); b = Sim(blob, 16, lib=cuda_lib_f32, seed=9)
    for s in (a, b):
        s.reset_host()
        for _ in range(5):
            st = s.rollout(1, ACTION_RANDOM, None)
        assert st["env_steps"] == s.n_envs and st["n_bad_state"] == 0
    for e in range(min(n, 16)):
        sa, sb = a.get_state(e), b.get_state(e)
        assert np.array_equal(sa["qpos"], sb["qpos"]) and np.array_equal(sa["qvel"], sb["qvel"]) and sa["step"] == 5
    last = a.get_state(n - 1)
    assert last["step"] == 5 and np.isfinite(last["qpos"]).all() and abs(np.linalg.norm(last["qpos"][3:7]) - 1) < 1e-5
    a.close(); b.close()
