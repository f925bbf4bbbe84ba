"""The fp64 C oracle (oracle/go1_oracle.c): oracle-free physics invariants and
reproduction of the committed golden vectors.

PARITY UNPINNED vs MuJoCo itself: mujoco==2.3.0 is not installable in this image and the
reference ships no golden vectors for this path, so the restatement is pinned by invariants every correct
mj_step must satisfy and by the notebook's model constants (tests/test_mjcf.py)."""
import os

import numpy as np
import pytest

import parity_common as pc
from oracle.oracle import OracleSim
from quadruped_rl_locomotion_b200.mjcf import Model, dense_mass_matrix

MASS = 12.743448


@pytest.fixture(scope="module")
def model():
    return Model.from_blob(pc.model_blob("torque"))


def _random_pose(model, rng, z=0.8):
    q = np.array(model.key_qpos[0], dtype=np.float64)
    q[2] = z
    quat = rng.normal(size=4); q[3:7] = quat / np.linalg.norm(quat)
    q[7:] = rng.uniform(model.jnt_range[1:, 0], model.jnt_range[1:, 1])
    return q


def test_mass_matrix_matches_jacobian_form(model):
    """mj_crb (sparse qM) == sum_b m J'J + J' I J computed independently in numpy; M[0:3,0:3] = m I3."""
    sim = OracleSim(model)
    rng = np.random.default_rng(0)
    for _ in range(4):
        q = _random_pose(model, rng)
        sim.reset_data(); sim.data.qpos[:] = q
        sim.forward()
        M = sim.full_M()
        Md, _ = dense_mass_matrix(model, q)
        assert np.allclose(M, Md, atol=1e-10)
        assert np.allclose(M[:3, :3], MASS * np.eye(3), atol=1e-5)
        assert np.linalg.eigvalsh(M).min() > 0


def test_bias_force_at_rest_is_weight(model):
    sim = OracleSim(model)
    sim.data.qpos[:] = model.key_qpos[0]; sim.data.qpos[2] = 1.0
    sim.forward()
    assert np.allclose(sim.data.qfrc_bias[:3], [0, 0, MASS * 9.81], atol=1e-4)     # = -m g  (125.013 N)
    assert sim.data.ncon == 0


def test_airborne_com_falls_with_g(model):
    """No contacts, zero control: joint damping / frictionloss are internal forces, the COM accelerates with g and
    the linear momentum changes by m g h per substep."""
    sim = OracleSim(model)
    rng = np.random.default_rng(1)
    sim.data.qpos[:] = _random_pose(model, rng, z=1.5)
    sim.data.qvel[:] = rng.normal(size=18) * 0.5
    sim.forward()
    mass = np.asarray(model.body_mass)

    def com_and_momentum():
        sim.forward()
        com = (mass[:, None] * sim.data.xipos).sum(0) / mass.sum()
        return com

    h = model.opt_timestep
    c0 = com_and_momentum(); sim.step(1); c1 = com_and_momentum(); sim.step(1); c2 = com_and_momentum()
    acc = (c2 - 2 * c1 + c0) / (h * h)
    assert np.allclose(acc, [0, 0, -9.81], atol=2e-2)      # semi-implicit Euler: O(h) discretisation error
    assert sim.data.ncon == 0
    assert abs(np.linalg.norm(sim.data.qpos[3:7]) - 1.0) < 1e-12


def test_standing_contact_forces_balance_weight(model):
    """Settle on the floor with zero torque control off the home pose held by a PD (position model): the sum of
    the foot normal forces equals the weight and every contact force lies inside its elliptic cone."""
    mp = Model.from_blob(pc.model_blob("position"))
    sim = OracleSim(mp)
    sim.data.qpos[:] = mp.key_qpos[0]
    sim.data.ctrl[:] = mp.key_ctrl[0]
    sim.step(4000)
    sim.forward()
    assert np.abs(sim.data.qvel).max() < 5e-3
    cons = sim.data.contact
    assert len(cons) >= 4
    f = sim.data.efc_force
    fz = 0.0
    for c in cons:
        fn = f[c.efc_address]
        ft = f[c.efc_address + 1:c.efc_address + 3]
        assert fn >= 0
        assert np.hypot(*(ft / c.friction[:2])) <= fn * (1 + 1e-6) + 1e-9       # elliptic cone (tangential part)
        fz += fn
    assert fz == pytest.approx(MASS * 9.81, rel=2e-3)
    # KKT: M (a - a_smooth) = J' f
    J = sim.data.efc_J
    M = sim.full_M()
    res = M @ (sim.data.qacc - sim.data.qacc_smooth) - J.T @ f
    assert np.abs(res).max() < 1e-6


def test_energy_does_not_grow_without_control(model):
    sim = OracleSim(model)
    sim.data.qpos[:] = model.key_qpos[0]; sim.data.qpos[2] = 0.5
    mass = np.asarray(model.body_mass)

    def energy():
        sim.forward()
        M = sim.full_M()
        v = sim.data.qvel
        return 0.5 * v @ M @ v + 9.81 * (mass * sim.data.xipos[:, 2]).sum()

    e = [energy()]
    for _ in range(60):
        sim.step(10)
        e.append(energy())
    assert max(np.diff(e)) < 1e-3         # dissipative: drops on impact, never climbs (up to integrator error)
    assert np.isfinite(e).all()


def test_determinism(model):
    a, b = OracleSim(model), OracleSim(model)
    rng = np.random.default_rng(3)
    for s in (a, b):
        s.data.qpos[:] = model.key_qpos[0]
    for _ in range(20):
        u = rng.uniform(-1, 1, 12)
        a.data.ctrl[:] = u; b.data.ctrl[:] = u
        a.step(10); b.step(10)
    assert np.array_equal(a.data.qpos, b.data.qpos) and np.array_equal(a.data.qvel, b.data.qvel)


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_oracle_reproduces_golden_trace(ctrl):
    """The committed vectors are what this oracle produces (guards against the oracle drifting from the goldens)."""
    m = Model.from_blob(pc.model_blob(ctrl))
    pre, post = pc.load_trace(ctrl)
    sim = OracleSim(m)
    for t in list(range(0, 12)) + list(range(40, len(post["reward"]), 97)):
        sim.reset_data()
        sim.data.qpos[:] = pre["qpos"][t]; sim.data.qvel[:] = pre["qvel"][t]
        sim.data.qacc_warmstart[:] = pre["warm"][t]
        sim.data.ctrl[:] = post["action"][t]
        sim.step(10); sim.rne_post_constraint()
        assert np.allclose(sim.data.qpos, post["qpos"][t], rtol=0, atol=1e-12), t
        assert np.allclose(sim.data.qvel, post["qvel"][t], rtol=0, atol=1e-11), t
        assert np.allclose(sim.data.cfrc_ext, post["cfrc_ext"][t], rtol=1e-9, atol=1e-9), t
        assert sim.data.ncon == post["ncon"][t]


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_every_robot_pair_of_the_golden_states_is_collided(ctrl):
    """The oracle COUNTS a geom pair it cannot collide (warn_unsupported_pair: a cylinder pair that is not coplanar, cylinder-box,
    box-box) instead of silently skipping it: no state of the golden sets holds one, and the random-torque sets do exercise the one
    cylinder-cylinder case this robot has (left vs right rear hip motor)."""
    m = Model.from_blob(pc.model_blob(ctrl))
    gt = np.asarray(m.geom_type)
    sim = OracleSim(m)
    n_states = n_cylcyl = 0
    for pre, post in (pc.load_trace(ctrl), pc.load_starts(ctrl)):
        for t in range(len(post["reward"])):
            sim.reset_data()
            sim.data.qpos[:] = pre["qpos"][t]; sim.data.qvel[:] = pre["qvel"][t]
            sim.data.ctrl[:] = post["action"][t]
            sim.step(1)
            n_states += 1
            n_cylcyl += sum(1 for c in sim.data.contact if c.geom1 > 0 and gt[c.geom1] == 5 and gt[c.geom2] == 5)
    assert n_states > 800 and sim.data.warn_unsupported_pair == 0
    assert n_cylcyl > 0 or ctrl == "position"


def test_substep_golden_is_consistent_with_trace():
    sub = np.load(os.path.join(pc.GOLDEN, "substeps_torque.npz"))
    pre, post = pc.load_trace("torque")
    # state after 10 substeps == trace state after 1 control step
    assert np.allclose(sub["qpos"][10], post["qpos"][0], atol=1e-12)
    assert np.allclose(sub["qvel"][10], post["qvel"][0], atol=1e-12)
