"""The pin of the physics oracle to MuJoCo itself (SURVEY 8(c) plan A).

``tests/golden/mujoco_*.npz`` are produced by ``tests/golden/make_mujoco_golden.py`` from REAL ``mujoco.mj_step`` on a machine
that has the ``mujoco`` wheel.  Neither this image nor the GPU box has it (no wheel, no network), so the files are absent and
these tests SKIP with the explicit status "parity vs MuJoCo unpinned": the oracle is then pinned only by the notebook's model
constants (tests/test_mjcf.py) and by physics invariants (tests/test_oracle.py).  The day the fixtures exist, the same tests
turn that status green without any other change.

Known, documented deviations of the oracle from MuJoCo 2.3.0 are excluded from the 1e-10 comparison, not hidden: contacts of a
cylinder with a sphere / capsule (MuJoCo: libccd MPR, tolerance 1e-6) and capsule-box contacts (the second-contact rule of
mjc_CapsuleBox is not reproduced): steps whose MuJoCo contact list holds such a pair are compared on the contact SET only."""
import os

import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.mjcf import Model

UNPINNED = ("parity vs MuJoCo unpinned: tests/golden/mujoco_{kind}_{ctrl}.npz is absent (needs `import mujoco`; run "
            "tests/golden/make_mujoco_golden.py where the wheel exists)")


def _load(kind, ctrl):
    p = os.path.join(pc.GOLDEN, f"mujoco_{kind}_{ctrl}.npz")
    if not os.path.exists(p):
        pytest.skip(UNPINNED.format(kind=kind, ctrl=ctrl))
    return np.load(p)


def _replay(ctrl, g):
    from oracle.oracle import OracleSim
    model = Model.from_blob(pc.model_blob(ctrl))
    gt = np.asarray(model.geom_type)
    o = OracleSim(model)
    worst = dict(qpos=0.0, qvel=0.0, qacc=0.0, cfrc_ext=0.0, force=0.0)
    n_exact = n_set_only = set_mismatch = 0
    for t in range(len(g["action"])):
        o.reset_data()
        o.data.qpos[:] = g["pre_qpos"][t]; o.data.qvel[:] = g["pre_qvel"][t]; o.data.qacc_warmstart[:] = g["pre_warm"][t]
        o.data.ctrl[:] = g["action"][t]
        o.step(10); o.rne_post_constraint()
        n = int(g["ncon"][t])
        gold = sorted((int(a), int(b)) for a, b in zip(g["con_geom1"][t][:n], g["con_geom"][t][:n]))
        ours = sorted((c.geom1, c.geom2) for c in o.data.contact if c.exclude == 0)
        approx = any(a > 0 and (5 in (gt[a], gt[b]) or {int(gt[a]), int(gt[b])} == {3, 6}) for a, b in gold)
        if approx:      # cylinder pair (MPR in MuJoCo) or capsule-box: documented deviation, contact set only
            n_set_only += 1
            set_mismatch += int({p for p in gold} != {p for p in ours})
            continue
        assert gold == ours, (t, gold, ours)
        n_exact += 1
        worst["qpos"] = max(worst["qpos"], pc.rel_err(o.data.qpos, g["qpos"][t]))
        worst["qvel"] = max(worst["qvel"], pc.rel_err(o.data.qvel, g["qvel"][t]))
        worst["qacc"] = max(worst["qacc"], pc.rel_err(o.data.qacc, g["qacc"][t]))
        worst["cfrc_ext"] = max(worst["cfrc_ext"], pc.rel_err(o.data.cfrc_ext, g["cfrc_ext"][t]))
    return worst, n_exact, n_set_only, set_mismatch


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_oracle_matches_mujoco_trace(ctrl):
    g = _load("trace", ctrl)
    worst, n_exact, n_set_only, set_mismatch = _replay(ctrl, g)
    print("mujoco", str(g["mujoco_version"]), ctrl, worst, n_exact, n_set_only, set_mismatch)
    assert n_exact > 0
    # MuJoCo's Newton tolerance is 1e-8 on the scaled improvement: two correct implementations agree to ~1e-9, not to rounding
    assert worst["qpos"] <= 1e-9 and worst["qvel"] <= 1e-8 and worst["qacc"] <= 1e-7 and worst["cfrc_ext"] <= 1e-7, worst
    assert set_mismatch <= max(1, n_set_only // 10)


@pytest.mark.parametrize("ctrl", ["torque", "position"])
def test_oracle_matches_mujoco_random_states(ctrl):
    g = _load("random", ctrl)
    worst, n_exact, n_set_only, set_mismatch = _replay(ctrl, g)
    print("mujoco", str(g["mujoco_version"]), ctrl, worst, n_exact, n_set_only, set_mismatch)
    assert n_exact > 100
    assert worst["qpos"] <= 1e-8 and worst["qvel"] <= 1e-7, worst


def test_pin_status_is_reported():
    """Always runs: states, in the test log, whether the oracle is pinned to MuJoCo on this machine."""
    have = [f for f in os.listdir(pc.GOLDEN) if f.startswith("mujoco_") and f.endswith(".npz")]
    print("MuJoCo pin fixtures:", have if have else "none -> parity vs MuJoCo UNPINNED (DESIGN.md section 1)")
    import importlib.machinery
    # on disk, not in sys.modules: oracle/ref_env.py injects a stand-in module named `mujoco` for the reference's imports
    have_mujoco = importlib.machinery.PathFinder.find_spec("mujoco") is not None
    assert have or not have_mujoco, "mujoco is importable here: run tests/golden/make_mujoco_golden.py and commit the fixtures"
