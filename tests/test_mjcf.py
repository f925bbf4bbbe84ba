"""MJCF compiler against the only known-answer values the reference holds: the saved outputs
of ``test_scripts/go1_mujoco_test_env.ipynb`` (notebook lines cited per assert) and the
MJCF sources themselves (``unitree_go1/go1_torque.xml`` / ``go1_position.xml``).

The committed blobs in ``quadruped_rl_locomotion_b200/models`` are what travels to the GPU box; when
``/root/reference`` is present (this container) the compiler is also re-run on the XML and must reproduce
the committed blob bit for bit."""
import os

import numpy as np
import pytest

from quadruped_rl_locomotion_b200.mjcf import Model, MjcfError, compile_mjcf, dense_mass_matrix, kinematics

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODELS = os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models")
REF = "/root/reference/unitree_go1"


@pytest.fixture(scope="module", params=["torque", "position"])
def model(request):
    m = Model.load(os.path.join(MODELS, f"go1_{request.param}.go1m"))
    m.ctrl = request.param
    return m


def test_sizes(model):
    # notebook :158-179: nv 18, nq 19, nbody 14, nsite 6; timestep 0.002
    assert (model.nq, model.nv, model.nu, model.nbody, model.nsite) == (19, 18, 12, 14, 6)
    assert model.ngeom == 56 and model.njnt == 13
    assert model.opt_timestep == pytest.approx(0.002)
    assert model.qpos0[2] == pytest.approx(0.445)
    assert np.allclose(model.qpos0[3:7], [1, 0, 0, 0])


def test_foot_geom_FL(model):
    # notebook :112-133 prints model.geom("FL")
    g = model.name2id("geom", "FL")
    assert g == 31
    assert model.geom_bodyid[g] == 7 and model.geom_condim[g] == 6 and model.geom_priority[g] == 1
    assert np.allclose(model.geom_friction[g], [2, 0.02, 0.01])
    assert model.geom_margin[g] == pytest.approx(0.001)
    assert model.geom_rbound[g] == pytest.approx(0.023)
    assert np.allclose(model.geom_solimp[g], [0.015, 1, 0.031, 0.5, 2])
    assert np.allclose(model.geom_solref[g], [0.02, 1])


def test_joint_and_ctrl_ranges(model):
    # notebook :253-290 (jnt_range, actuator_ctrlrange); go1_torque.xml:7,216-227 / go1_position.xml:15-26,216-230
    jr = model.jnt_range[1:]
    assert np.allclose(jr[0::3], [[-0.863, 0.863]] * 4)
    assert np.allclose(jr[1::3], [[-0.686, 4.501]] * 4)
    assert np.allclose(jr[2::3], [[-2.818, -0.888]] * 4)
    cr = model.actuator_ctrlrange
    if model.ctrl == "torque":
        assert np.allclose(cr, [[-1, 1]] * 12)
        assert np.allclose(np.asarray(model.actuator_gear).reshape(4, 3), [[23, 23, 35]] * 4)     # gear 23|23|35
    else:
        assert np.allclose(cr, jr)
        assert np.allclose(model.actuator_gainprm[:, 0], 100.0) and np.allclose(model.actuator_biasprm[:, 1], -100.0)
        assert np.allclose(model.actuator_forcerange.reshape(4, 3, 2)[:, :, 1], [[23.7, 23.7, 35.55]] * 4)


def test_dof_parameters(model):
    # go1_torque.xml:10,14 / go1_position.xml:14,17: armature 0.01, frictionloss 0.2, damping 2 (abduction class: 1)
    assert np.allclose(model.dof_armature[6:], 0.01) and np.allclose(model.dof_armature[:6], 0)
    assert np.allclose(model.dof_frictionloss[6:], 0.2)
    assert np.allclose(model.dof_damping[6:].reshape(4, 3), [[1, 2, 2]] * 4) and np.allclose(model.dof_damping[:6], 0)


def test_keyframe_home(model):
    # go1_torque.xml:231: z = 0.30 (torque) / keyframe joints (0, 0.9, -1.8) x 4
    k = model.key_qpos[0]
    assert np.allclose(k[7:], [0, 0.9, -1.8] * 4)
    assert np.allclose(k[3:7], [1, 0, 0, 0])
    # key ctrl: go1_torque.xml:232 zeros, go1_position.xml:236 the home joint angles
    assert np.allclose(model.key_ctrl[0], [0] * 12 if model.ctrl == "torque" else [0, 0.9, -1.8] * 4)
    assert k[2] == pytest.approx(0.30)


def test_total_mass_and_mass_matrix(model):
    # SURVEY section 4 invariants: total mass 12.743448 kg; M[0:3,0:3] = m*I3 for any pose; M symmetric PD
    assert model.body_mass.sum() == pytest.approx(12.743448, abs=2e-6)
    assert model.body_subtreemass[1] == pytest.approx(model.body_mass.sum())
    rng = np.random.default_rng(0)
    for _ in range(5):
        q = model.qpos0.copy()
        quat = rng.normal(size=4); q[3:7] = quat / np.linalg.norm(quat)
        q[7:] = rng.uniform(model.jnt_range[1:, 0], model.jnt_range[1:, 1])
        M, _ = dense_mass_matrix(model, q)
        assert np.allclose(M, M.T, atol=1e-12)
        assert np.allclose(M[:3, :3], model.body_mass.sum() * np.eye(3), atol=1e-9)
        assert np.linalg.eigvalsh(M).min() > 0


def test_kinematics_home_pose(model):
    # feet (sites FR, FL, RR, RL, fixed to the calves) at the home keyframe: mirror-symmetric, just above the floor
    k = kinematics(model, model.key_qpos[0])
    names = model.names["site"]
    feet = {}
    for n in ("FR", "FL", "RR", "RL"):
        s = names.index(n)
        b = model.site_bodyid[s]
        feet[n] = k["xpos"][b] + k["xmat"][b].reshape(3, 3) @ model.site_pos[s]
    assert np.allclose(feet["FR"][[0, 2]], feet["FL"][[0, 2]], atol=1e-9) and feet["FR"][1] == pytest.approx(-feet["FL"][1])
    assert np.allclose(feet["RR"][[0, 2]], feet["RL"][[0, 2]], atol=1e-9)
    assert all(0.0 < f[2] < 0.05 for f in feet.values())      # foot centres (radius 2.3 cm) at trunk z = 0.30


def test_collision_geometry_census(model):
    # floor plane vs 4 spheres, 24 capsules, 12 cylinders, 2 boxes (43 colliding geoms)
    col = (model.geom_contype != 0) | (model.geom_conaffinity != 0)
    t = model.geom_type[col]
    # mjtGeom: 0 plane, 2 sphere, 3 capsule, 5 cylinder, 6 box
    assert col.sum() == 43
    assert [(t == c).sum() for c in (0, 2, 3, 5, 6)] == [1, 4, 24, 12, 2]


def test_blob_round_trip(model):
    again = Model.from_blob(model.to_blob())
    assert sorted(again.arrays) == sorted(model.arrays)
    for k, v in model.arrays.items():
        assert np.array_equal(again.arrays[k], v), k
    assert again.names == model.names


def test_bad_blob_rejected():
    with pytest.raises(MjcfError):
        Model.from_blob(b"NOTAGO1M" + b"\0" * 100)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference MJCF sources not on this machine")
@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_compiler_reproduces_committed_blob(ctrl):
    fresh = compile_mjcf(os.path.join(REF, f"scene_{ctrl}.xml"))
    shipped = Model.load(os.path.join(MODELS, f"go1_{ctrl}.go1m"))
    assert sorted(fresh.arrays) == sorted(shipped.arrays)
    for k in fresh.arrays:
        a, b = np.asarray(fresh.arrays[k]), np.asarray(shipped.arrays[k])
        assert a.shape == b.shape and np.array_equal(a.astype(b.dtype), b), k
