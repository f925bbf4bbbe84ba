"""The MuJoCo pin of the physics oracle (SURVEY 8(c) plan A).  Runs ONLY where ``import mujoco`` works -- it does not in this
image nor on the GPU box (no wheel, no network), so ``tests/golden/mujoco_*.npz`` are absent and ``tests/test_mujoco_pin.py``
skips with "parity vs MuJoCo unpinned".  On a machine that has the wheel (the reference pins mujoco==2.3.0,
requirements.txt:81):

    GO1_REFERENCE_ROOT=/path/to/quadruped-rl-locomotion python tests/golden/make_mujoco_golden.py

writes, from REAL ``mujoco.mj_step``,
  tests/golden/mujoco_trace_<ctrl>.npz    BASELINE config 1: home keyframe, actions rng(0).uniform(ctrlrange) as float32,
                                          per control step (10 x mj_step + mj_rnePostConstraint, what gymnasium's
                                          MujocoEnv.do_simulation does for go1_mujoco_env.py:481): pre-state, qpos, qvel, qacc,
                                          qacc_warmstart, cfrc_ext, contacts (geom1, geom2, dim, dist, pos, frame, efc_force rows)
  tests/golden/mujoco_random_<ctrl>.npz   the random states of tests/test_differential.py stepped once the same way
and records ``mujoco.__version__``.  The reference's MJCF references ``assets/trunk.stl`` which the repository does not ship
(.MISSING_LARGE_BLOBS): a one-triangle placeholder is written into a temporary copy; the mesh is visual only (contype 0).
"""
from __future__ import annotations

import os
import shutil
import struct
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
OUT = os.path.join(ROOT, "tests", "golden")
MAXC = 64


def _placeholder_stl(path):
    with open(path, "wb") as f:
        f.write(b"placeholder".ljust(80, b" "))
        f.write(struct.pack("<I", 1))
        f.write(struct.pack("<12fH", 0, 0, 1, 0, 0, 0, 0.01, 0, 0, 0, 0.01, 0, 0))


def load_model(mujoco, ctrl):
    ref = os.environ.get("GO1_REFERENCE_ROOT", "/root/reference")
    tmp = tempfile.mkdtemp(prefix="go1_mjcf_")
    dst = os.path.join(tmp, "unitree_go1")
    shutil.copytree(os.path.join(ref, "unitree_go1"), dst)
    assets = os.path.join(dst, "assets")
    os.makedirs(assets, exist_ok=True)
    import re
    for fn in os.listdir(dst):
        if fn.endswith(".xml"):
            for mesh in re.findall(r'file="([^"]+\.stl)"', open(os.path.join(dst, fn)).read(), flags=re.I):
                p = os.path.join(assets, os.path.basename(mesh))
                if not os.path.exists(p):
                    _placeholder_stl(p)
    return mujoco.MjModel.from_xml_path(os.path.join(dst, f"scene_{ctrl}.xml"))


def do_simulation(mujoco, m, d, ctrl, frame_skip=10):
    d.ctrl[:] = ctrl
    mujoco.mj_step(m, d, nstep=frame_skip)
    mujoco.mj_rnePostConstraint(m, d)


def snapshot(m, d):
    g1 = -np.ones(MAXC, np.int32); g2 = -np.ones(MAXC, np.int32); dim = np.zeros(MAXC, np.int32)
    dist = np.zeros(MAXC); pos = np.zeros((MAXC, 3)); frame = np.zeros((MAXC, 9)); frc = np.zeros((MAXC, 6))
    k = 0
    for i in range(d.ncon):
        c = d.contact[i]
        if c.exclude != 0 or c.efc_address < 0 or k >= MAXC:
            continue
        g1[k], g2[k], dim[k], dist[k] = c.geom1, c.geom2, c.dim, c.dist
        pos[k] = c.pos; frame[k] = c.frame
        frc[k, :c.dim] = d.efc_force[c.efc_address:c.efc_address + c.dim]
        k += 1
    return dict(qpos=d.qpos.copy(), qvel=d.qvel.copy(), qacc=d.qacc.copy(), warm=d.qacc_warmstart.copy(), cfrc_ext=d.cfrc_ext.copy(),
                ncon=k, con_geom1=g1, con_geom=g2, con_dim=dim, con_dist=dist, con_pos=pos, con_frame=frame, con_force=frc,
                solver_iter=int(d.solver_iter[0]) if np.ndim(d.solver_iter) else int(d.solver_iter))


def main():
    import mujoco
    for ctrl in ("torque", "position"):
        m = load_model(mujoco, ctrl)
        d = mujoco.MjData(m)
        lo, hi = m.actuator_ctrlrange[:, 0].copy(), m.actuator_ctrlrange[:, 1].copy()
        # ---- config-1 trace
        mujoco.mj_resetData(m, d)
        d.qpos[:] = m.key_qpos[0]
        rng = np.random.default_rng(0)
        rec = {}
        for t in range(1000 if ctrl == "torque" else 400):
            pre = dict(pre_qpos=d.qpos.copy(), pre_qvel=d.qvel.copy(), pre_warm=d.qacc_warmstart.copy())
            a = rng.uniform(lo, hi).astype(np.float32)
            do_simulation(mujoco, m, d, a.astype(np.float64))
            row = {**pre, "action": a.astype(np.float64), **snapshot(m, d)}
            for k, v in row.items():
                rec.setdefault(k, []).append(v)
        np.savez_compressed(os.path.join(OUT, f"mujoco_trace_{ctrl}.npz"), mujoco_version=mujoco.__version__,
                            **{k: np.array(v) for k, v in rec.items()})
        # ---- random states (the generator of tests/test_differential.py, so the oracle test replays exactly these)
        from quadruped_rl_locomotion_b200.mjcf import Model
        from test_differential import random_states
        model = Model.load(os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models", f"go1_{ctrl}.go1m"))
        rec = {}
        for s in random_states(model, 1000, 12):
            mujoco.mj_resetData(m, d)
            d.qpos[:] = s["qpos"]; d.qvel[:] = s["qvel"]; d.qacc_warmstart[:] = s["warm"]
            a = np.float32(s["action"]).astype(np.float64)
            do_simulation(mujoco, m, d, a)
            row = {"pre_qpos": s["qpos"], "pre_qvel": s["qvel"], "pre_warm": s["warm"], "action": a, **snapshot(m, d)}
            for k, v in row.items():
                rec.setdefault(k, []).append(v)
        np.savez_compressed(os.path.join(OUT, f"mujoco_random_{ctrl}.npz"), mujoco_version=mujoco.__version__,
                            **{k: np.array(v) for k, v in rec.items()})
        print(ctrl, "written; mujoco", mujoco.__version__)


if __name__ == "__main__":
    main()
