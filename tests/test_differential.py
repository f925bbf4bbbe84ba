"""Differential test on RANDOM states: the CUDA source (fp64 build) against the fp64 oracle, one control step from
states no fixture holds.  Seeded draws over the whole state space the environment can reach and some it cannot
(arbitrary trunk orientation, heights from deep penetration to airborne, joints anywhere inside their limits, velocities up
to 10 rad/s, arbitrary warm start), both control modes.  The fixtures pin the oracle; this pins the kernel to the oracle
where the fixtures do not look.  CPU: SIMT emulation of the same source, 48 states; GPU: 1024 states per mode."""
import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.lib import Sim
from quadruped_rl_locomotion_b200.mjcf import Model


def random_states(model, n, seed):
    rng = np.random.default_rng(seed)
    jr = np.array(model.jnt_range)[1:]
    lo, hi = np.array(model.actuator_ctrlrange).T
    out = []
    for i in range(n):
        q = np.zeros(19); v = np.zeros(18)
        kind = i % 4
        q[:2] = rng.uniform(-3, 3, 2)
        quat = rng.normal(size=4); q[3:7] = quat / np.linalg.norm(quat)
        q[7:] = rng.uniform(jr[:, 0], jr[:, 1])
        if kind == 0:      # near the home pose, small velocities (the walking regime)
            q[:] = model.key_qpos[0]; q[7:] += rng.normal(size=12) * 0.2; q[7:] = np.clip(q[7:], jr[:, 0], jr[:, 1])
            q[2] = rng.uniform(0.24, 0.36); v[:] = rng.normal(size=18) * 0.5
        elif kind == 1:    # on / in the floor in an arbitrary orientation (many contacts, some penetrating)
            q[2] = rng.uniform(0.08, 0.25); v[:] = rng.normal(size=18) * 2.0
        elif kind == 2:    # airborne, fast
            q[2] = rng.uniform(0.5, 1.0); v[:6] = rng.normal(size=6) * 3.0; v[6:] = rng.uniform(-10, 10, 12)
        else:              # joints at their limits
            q[7:] = np.where(rng.random(12) < 0.5, jr[:, 0] - 0.02, jr[:, 1] + 0.02)
            q[2] = rng.uniform(0.2, 0.5); v[6:] = rng.normal(size=12) * 5.0
        out.append(dict(qpos=q, qvel=v, warm=rng.normal(size=18) * (0.0 if i % 3 == 0 else 5.0), action=rng.uniform(lo, hi)))
    return out


def _run(lib, ctrl, n, seed, tol, quantile=None):
    from oracle.oracle import OracleSim
    blob = pc.model_blob(ctrl)
    model = Model.from_blob(blob)
    states = random_states(model, n, seed)
    sim = Sim(blob, n, lib=lib, seed=1)
    sim.enable_diag(True)
    for e, s in enumerate(states):
        sim.set_state(e, qpos=s["qpos"], qvel=s["qvel"], qacc_warmstart=s["warm"], step=0, unhealthy_steps=0, last_action=np.zeros(12))
    act = np.ascontiguousarray(np.array([s["action"] for s in states], np.float32))
    obs = np.zeros((n, 48), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); trunc = np.zeros(n, np.uint8)
    sim.step_host(act, obs, rew, done, trunc)
    o = OracleSim(model)
    worst = dict(qpos=0.0, qvel=0.0, qacc=0.0, cfrc_ext=0.0)
    per_state = {k: [] for k in worst}
    ncon_mismatch = overflow = 0
    for e, s in enumerate(states):
        o.reset_data()
        o.data.qpos[:] = s["qpos"]; o.data.qvel[:] = s["qvel"]; o.data.qacc_warmstart[:] = s["warm"]
        o.data.ctrl[:] = act[e].astype(np.float64)
        o.step(10); o.rne_post_constraint()
        d = sim.get_diag(e); st = sim.get_state(e)
        if d["contact_overflow"]:
            overflow += 1
            continue
        ncon_mismatch += int(d["ncon"] != o.data.ncon)
        err = dict(qpos=pc.rel_err(st["qpos"], o.data.qpos), qvel=pc.rel_err(st["qvel"], o.data.qvel),
                   qacc=pc.rel_err(d["qacc"], o.data.qacc), cfrc_ext=pc.rel_err(d["cfrc_ext"], o.data.cfrc_ext))
        for k, v in err.items():
            per_state[k].append(v); worst[k] = max(worst[k], v)
    sim.close()
    assert overflow <= n // 16, overflow           # more than 40 contacts (deep penetration): counted by the kernel, outside its capacity
    if quantile is None:
        assert ncon_mismatch == 0
        for k, t in tol.items():
            assert worst[k] <= t, (ctrl, k, worst[k])
        return worst
    # fp32: a contact within rounding of the 1 mm margin may be on either side, and ten substeps of a violent state amplify
    # the last bits; the bar is on the `quantile` of the states, the worst case is reported
    assert ncon_mismatch <= max(2, n // 100), ncon_mismatch
    q = {k: float(np.quantile(v, quantile)) for k, v in per_state.items()}
    for k, t in tol.items():
        assert q[k] <= t, (ctrl, k, q[k])
    return dict(worst=worst, quantile=q, ncon_mismatch=ncon_mismatch, overflow=overflow)


F64 = dict(qpos=1e-10, qvel=1e-10, qacc=1e-9, cfrc_ext=1e-9)      # measured: 2e-13, 2e-12, 4e-12, 7e-12
F32 = dict(qpos=2e-6, qvel=2e-5, qacc=1e-4, cfrc_ext=5e-3)         # on the 95 % quantile of the states (ten mj_steps)


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_random_states_emu_f64(emu_lib_f64, ctrl):
    _run(emu_lib_f64, ctrl, 48, seed=11, tol=F64)


@pytest.mark.gpu
@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_random_states_gpu_f64(cuda_lib_f64, ctrl):
    w = _run(cuda_lib_f64, ctrl, 1024, seed=12, tol=F64)
    print(ctrl, {k: float("%.2g" % v) for k, v in w.items()})


@pytest.mark.gpu
@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_random_states_gpu_f32(cuda_lib_f32, ctrl):
    r = _run(cuda_lib_f32, ctrl, 1024, seed=13, tol=F32, quantile=0.95)
    print(ctrl, r)
    # worst case over the 1024 states after TEN mj_steps, measured 3.7e-7 / 5.0e-6 (north_star: 1e-5 per mj_step)
    assert r["worst"]["qpos"] <= 2e-6 and r["worst"]["qvel"] <= 2e-5


def test_random_states_emu_f32(emu_lib_f32):
    r = _run(emu_lib_f32, "torque", 48, seed=14, tol=F32, quantile=0.9)
    print(r)
