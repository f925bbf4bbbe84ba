"""Run the reference's UNMODIFIED ``go1_mujoco_env.py`` on top of the oracle physics.

TEST INFRASTRUCTURE ONLY, and only usable where ``/root/reference`` exists (this container):
it is how ``tests/golden/make_golden.py`` produces the env-logic golden vectors.  The
reference imports ``gymnasium`` / ``mujoco`` (``go1_mujoco_env.py:1-4``), neither of which is
installed here, so minimal stand-in modules are injected into ``sys.modules``:

* ``gymnasium.spaces.Box``            -- shape/dtype/low/high container + ``sample``
* ``gymnasium.envs.mujoco.MujocoEnv`` -- restates gymnasium 0.29.1 ``MujocoEnv``
  (``__init__`` compiles the MJCF, ``do_simulation`` = ``ctrl <- a; mj_step(nstep); mj_rnePostConstraint``,
  ``reset`` = ``mj_resetData`` then ``reset_model()``, ``state_vector``, ``dt``), backed by
  :class:`oracle.oracle.OracleSim` instead of the MuJoCo C library
* ``mujoco.mj_name2id`` / ``mujoco.mjtObj``

Nothing of the reference is copied: the file is imported from where it lies.
"""
from __future__ import annotations

import enum
import importlib.util
import os
import sys
import types

import numpy as np

from quadruped_rl_locomotion_b200.mjcf import compile_mjcf
from .oracle import OracleSim

REFERENCE_ROOT = os.environ.get("GO1_REFERENCE_ROOT", "/root/reference")


class Box:
    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        if shape is None:
            shape = np.asarray(low).shape
        self.shape = tuple(shape)
        self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), self.shape).copy()
        self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), self.shape).copy()
        self._rng = np.random.default_rng()

    def seed(self, seed=None):
        self._rng = np.random.default_rng(seed)
        return [seed]

    def sample(self):
        return self._rng.uniform(self.low, self.high).astype(self.dtype)


class _Opt:
    def __init__(self, m):
        self.gravity = np.array(m.opt_gravity)
        self.timestep = float(m.opt_timestep)


class _ModelView:
    """The few ``mjModel`` attributes the env reads (go1_mujoco_env.py:104-105,139-143,282,921,950,981)."""

    def __init__(self, m):
        self._m = m
        self.opt = _Opt(m)
        self.key_qpos = m.key_qpos
        self.key_ctrl = m.key_ctrl
        self.actuator_ctrlrange = m.actuator_ctrlrange
        self.jnt_range = m.jnt_range
        self.geom_bodyid = m.geom_bodyid
        self.nq, self.nv, self.nu = m.nq, m.nv, m.nu


class MujocoEnv:
    """gymnasium 0.29.1 ``MujocoEnv`` semantics on the oracle (see module docstring)."""

    def __init__(self, model_path, frame_skip, observation_space, render_mode=None, width=480, height=480,
                 camera_id=None, camera_name=None, default_camera_config=None, **kwargs):
        self.fullpath = model_path
        self._compiled = compile_mjcf(model_path)
        self.sim = OracleSim(self._compiled)
        self.model = _ModelView(self._compiled)
        self.data = self.sim.data
        self.init_qpos = self.data.qpos.ravel().copy()
        self.init_qvel = self.data.qvel.ravel().copy()
        self.frame_skip = frame_skip
        self.render_mode = render_mode
        self.observation_space = observation_space
        cr = self._compiled.actuator_ctrlrange.astype(np.float32)
        self.action_space = Box(low=cr[:, 0], high=cr[:, 1], dtype=np.float32)
        self.np_random = np.random.default_rng()

    @property
    def dt(self):
        return self.model.opt.timestep * self.frame_skip

    def do_simulation(self, ctrl, n_frames):
        if np.array(ctrl).shape != (self.model.nu,):
            raise ValueError(f"Action dimension mismatch. Expected {(self.model.nu,)}, found {np.array(ctrl).shape}")
        self.data.ctrl[:] = ctrl
        self.sim.step(n_frames)
        self.sim.rne_post_constraint()

    def state_vector(self):
        return np.concatenate([self.data.qpos.flat, self.data.qvel.flat])

    def reset(self, *, seed=None, options=None):
        if seed is not None:
            self.np_random = np.random.default_rng(seed)
        self.sim.reset_data()
        ob = self.reset_model()
        return ob, {}

    def render(self):
        return None

    def close(self):
        pass


class _mjtObj(enum.Enum):
    mjOBJ_BODY = 1
    mjOBJ_JOINT = 3
    mjOBJ_GEOM = 5
    mjOBJ_SITE = 6
    mjOBJ_ACTUATOR = 19


def _mj_name2id(model, objtype, name):
    kind = {1: "body", 3: "joint", 5: "geom", 6: "site", 19: "actuator"}[int(objtype)]
    return model._m.name2id(kind, name)


def _install_shims():
    if "gymnasium" in sys.modules and not getattr(sys.modules["gymnasium"], "_go1_shim", False):
        raise RuntimeError("a real gymnasium is importable; the shim is only for boxes without it")
    gym = types.ModuleType("gymnasium"); gym._go1_shim = True
    spaces = types.ModuleType("gymnasium.spaces"); spaces.Box = Box
    envs = types.ModuleType("gymnasium.envs")
    mj_envs = types.ModuleType("gymnasium.envs.mujoco"); mj_envs.MujocoEnv = MujocoEnv
    gym.spaces = spaces; gym.envs = envs; envs.mujoco = mj_envs
    mujoco = types.ModuleType("mujoco"); mujoco._go1_shim = True
    mujoco.mj_name2id = _mj_name2id
    mujoco.mjtObj = _mjtObj
    sys.modules.update({"gymnasium": gym, "gymnasium.spaces": spaces, "gymnasium.envs": envs,
                        "gymnasium.envs.mujoco": mj_envs, "mujoco": mujoco})


def reference_available() -> bool:
    return os.path.exists(os.path.join(REFERENCE_ROOT, "go1_mujoco_env.py"))


def load_reference_env_class():
    """Import the reference's Go1MujocoEnv class, unmodified, from ``/root/reference``."""
    if not reference_available():
        raise FileNotFoundError(f"{REFERENCE_ROOT}/go1_mujoco_env.py not present on this machine")
    _install_shims()
    spec = importlib.util.spec_from_file_location("_reference_go1_mujoco_env", os.path.join(REFERENCE_ROOT, "go1_mujoco_env.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.Go1MujocoEnv


def make_reference_env(ctrl_type="torque", biped=False, rand_power=0.0):
    cls = load_reference_env_class()
    cwd = os.getcwd()
    os.chdir(REFERENCE_ROOT)  # go1_mujoco_env.py:42 resolves ./unitree_go1/scene_<ctrl>.xml from the cwd
    try:
        env = cls(ctrl_type=ctrl_type, biped=biped, rand_power=rand_power)
    finally:
        os.chdir(cwd)
    return env
