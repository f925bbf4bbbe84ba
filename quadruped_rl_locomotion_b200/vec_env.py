"""``Go1CudaVecEnv``: the reference's vectorised environment surface on top of the CUDA library.

Drop-in for ``make_vec_env(Go1MujocoEnv, ..., vec_env_cls=SubprocVecEnv)`` (reference ``train.py:44-50``,
``test_max_envs.py:35-41``): same constructor keywords as ``Go1MujocoEnv`` (``go1_mujoco_env.py:41``),
same spaces (``Box(-inf, inf, (48,), float64)`` observations, ``Box(ctrlrange, float32)`` actions), and
the SB3 2.3.0 ``VecEnv`` semantics the PPO loop relies on: auto-reset inside ``step``, the returned
observation of a finished env is its reset observation, ``infos[i]["terminal_observation"]``,
``infos[i]["TimeLimit.truncated"]``, and the ``Monitor`` record ``infos[i]["episode"] = {"r","l","t"}``.

Host code only: every number comes from the CUDA kernels through ``go1_step_host`` (pinned staging,
H2D actions -> one kernel -> D2H results).  If ``stable_baselines3`` / ``gymnasium`` are importable the
class derives from SB3's ``VecEnv`` and uses gymnasium spaces; otherwise (this image) minimal
stand-ins with the same attributes are used so the class is still importable and testable.
"""
from __future__ import annotations

import os
import time
from typing import Any, List, Optional, Sequence

import numpy as np

from . import lib as _lib
from .mjcf import Model, compile_mjcf

try:  # pragma: no cover - not installed in the build image
    from stable_baselines3.common.vec_env.base_vec_env import VecEnv as _SB3VecEnv
    _HAVE_SB3 = True
except Exception:  # noqa: BLE001
    _SB3VecEnv = object
    _HAVE_SB3 = False

try:  # pragma: no cover
    from gymnasium import spaces as _spaces
    Box = _spaces.Box
except Exception:  # noqa: BLE001
    class Box:  # minimal gymnasium.spaces.Box stand-in
        def __init__(self, low, high, shape=None, dtype=np.float32):
            self.dtype = np.dtype(dtype)
            self.shape = tuple(shape) if shape is not None else np.asarray(low).shape
            self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), self.shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), self.shape).copy()
            self._rng = np.random.default_rng()

        def seed(self, seed=None):
            self._rng = np.random.default_rng(seed)
            return [seed]

        def sample(self):
            return self._rng.uniform(self.low, self.high).astype(self.dtype)

        def contains(self, x):
            x = np.asarray(x)
            return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

        def __eq__(self, other):
            return (isinstance(other, Box) and self.shape == other.shape and self.dtype == other.dtype
                    and np.array_equal(self.low, other.low) and np.array_equal(self.high, other.high))

        def __repr__(self):
            return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"

_MODEL_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "models")
INFO_KEYS = ("linear_vel_tracking_reward", "reward_ctrl", "reward_survive", "recovery_reward", "get_up_reward",
             "unhealthy_state_cost", "orientation_cost", "default_joint_position_cost")
BIPED_INFO_KEYS = ("biped_upright_reward", "biped_front_contact_cost", "biped_rear_feet_airborne_cost",
                   "biped_front_foot_height_cost", "biped_crossed_legs_cost", "biped_low_rear_hips_cost",
                   "biped_front_feet_below_hips_cost", "biped_abduction_joints_cost", "biped_unwanted_contact_cost",
                   "self_collision_cost")      # go1_mujoco_env.py:893-902
_REASONS = {0: "not_terminated", 1: "state_not_finite", 2: "unhealthy_z", 3: "unhealthy_roll", 4: "unhealthy_pitch"}


def load_model_blob(ctrl_type: str = "torque", model_dir: Optional[str] = None) -> bytes:
    """Compiled model for ``scene_<ctrl_type>.xml``.

    ``model_dir`` = a directory with the reference's MJCF files (``unitree_go1/``): compiled on the fly with
    :func:`compile_mjcf`; otherwise the blob shipped in ``models/`` (compiled from the reference's MJCF by
    ``tests/golden/make_golden.py``) is used.
    """
    if ctrl_type not in ("torque", "position"):
        raise ValueError(f"ctrl_type must be 'torque' or 'position', got {ctrl_type!r}")
    if model_dir is not None:
        return compile_mjcf(os.path.join(model_dir, f"scene_{ctrl_type}.xml")).to_blob()
    path = os.path.join(_MODEL_DIR, f"go1_{ctrl_type}.go1m")
    with open(path, "rb") as f:
        return f.read()


class StepInfos(list):
    """``infos`` of one ``step``: a list of N per-environment dicts with every key of ``Go1MujocoEnv.step``'s info
    (go1_mujoco_env.py:500-521, 860-909) plus SB3's ``terminal_observation`` / ``TimeLimit.truncated`` / ``episode`` for
    finished environments.  Building 16 384 dicts per step would cost more than the simulation, so a dict is built the
    first time its index is read and then kept (mutations by wrappers stay visible, and stay inside this step's object).
    The object owns copies of the step's info rows: it remains valid after later steps."""

    def __init__(self, env: "Go1CudaVecEnv", dones: np.ndarray):
        super().__init__([None] * env.num_envs)
        self._env = env
        self._rows = env._info.copy()
        self._brows = env._binfo.copy() if env.biped else None
        self._trunc = env._trunc.copy()
        self._dones = dones
        self._t = round(time.time() - env._t_start, 6)
        self._tobs = {int(i): env._tobs[i].astype(np.float64) for i in np.flatnonzero(dones)}

    def _build(self, i: int) -> dict:
        env = self._env
        done = bool(self._dones[i])
        d = env._info_dict(i, self._rows[i], None if self._brows is None else self._brows[i], done, bool(self._trunc[i]))
        if done:                                   # SB3 VecEnv / Monitor contract
            d["terminal_observation"] = self._tobs[i]
            d["TimeLimit.truncated"] = bool(self._trunc[i])
            d["episode"] = {"r": float(self._rows[i, 10]), "l": int(self._rows[i, 11]), "t": self._t}
        return d

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        v = list.__getitem__(self, i)
        if v is None:
            j = i if i >= 0 else i + len(self)
            v = self._build(j)
            list.__setitem__(self, j, v)
        return v

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    def copy(self):
        return [self[i] for i in range(len(self))]


class Go1CudaVecEnv(_SB3VecEnv):
    """N Go1 environments stepped by one CUDA kernel per control step."""

    metadata = {"render_modes": ["rgb_array"]}

    def __init__(self, num_envs: int, ctrl_type: str = "torque", biped: bool = False, rand_power: float = 0.0, *,
                 device: int = 0, seed: int = 0, env_id_offset: int = 0, desired_velocity_min=(-0.5, 0.0, 0.0),
                 desired_velocity_max=(0.5, 0.0, 0.0), model_dir: Optional[str] = None, precision: str = "f32",
                 library: Optional[_lib.Go1Lib] = None, full_info: bool = True, render_mode: Optional[str] = None,
                 camera_name: str = "tracking", width: int = 1024, height: int = 768):
        self._lib = library or _lib.Go1Lib(precision=precision)
        blob = load_model_blob(ctrl_type, model_dir)
        self.model = Model.from_blob(blob)
        cfg = self._lib.default_config()
        cfg.rand_power = float(rand_power)
        cfg.biped = 1 if biped else 0           # go1_mujoco_env.py:43
        for i in range(3):
            cfg.desired_vel_min[i] = float(desired_velocity_min[i])
            cfg.desired_vel_max[i] = float(desired_velocity_max[i])
        self.sim = _lib.Sim(blob, num_envs, lib=self._lib, cfg=cfg, device=device, seed=seed, env_id_offset=env_id_offset)
        lo, hi = self.sim.action_bounds()
        observation_space = Box(low=-np.inf, high=np.inf, shape=(_lib.OBS_DIM,), dtype=np.float64)
        action_space = Box(low=lo, high=hi, dtype=np.float32)
        self.ctrl_type, self.biped, self._rand_power = ctrl_type, bool(biped), float(rand_power)
        self.full_info = bool(full_info)
        if render_mode not in (None, "rgb_array"):
            raise ValueError(f"render_mode {render_mode!r}: only None and 'rgb_array' (headless CUDA ray-caster) are available on this box")
        if camera_name != "tracking":
            raise ValueError("only the model's `tracking` camera (go1_torque.xml:84) is implemented")
        self.render_mode, self.render_width, self.render_height = render_mode, int(width), int(height)
        n = int(num_envs)
        self._actions = np.zeros((n, _lib.ACT_DIM), np.float32)
        self._obs = np.zeros((n, _lib.OBS_DIM), np.float32)
        self._tobs = np.zeros((n, _lib.OBS_DIM), np.float32)
        self._rew = np.zeros(n, np.float32)
        self._done = np.zeros(n, np.uint8)
        self._trunc = np.zeros(n, np.uint8)
        self._info = np.zeros((n, _lib.INFO_DIM), np.float32)
        self._binfo = np.zeros((n, _lib.BIPED_INFO_DIM), np.float32) if biped else None
        # the step buffers live as long as the env: page-lock them once so every step is DMA to / from these arrays
        self._pinned = []
        for a in (self._actions, self._obs, self._tobs, self._rew, self._done, self._trunc, self._info):
            try:
                self.sim.host_register(a)
                self._pinned.append(a)
            except _lib.Go1Error:      # e.g. locked-memory limit: the library then stages through its own pinned buffers
                pass
        self._t_start = time.time()
        # float64 observation batches (observation_space.dtype): two arrays that alternate from step to step, written by the
        # library itself (go1_step_host_f64 widens on the device): no 6 MB allocation, no host-side conversion pass per step
        self._obs_f64 = [np.zeros((n, _lib.OBS_DIM), np.float64), np.zeros((n, _lib.OBS_DIM), np.float64)]
        self._obs_flip = 0
        for a in self._obs_f64:
            try:
                self.sim.host_register(a)
                self._pinned.append(a)
            except _lib.Go1Error:
                pass
        self._healthy_ranges = ((float(cfg.healthy_z[0]), float(cfg.healthy_z[1])), (float(cfg.healthy_roll[0]), float(cfg.healthy_roll[1])),
                                (float(cfg.healthy_pitch[0]), float(cfg.healthy_pitch[1])))
        if _HAVE_SB3:
            super().__init__(n, observation_space, action_space)
        else:
            self.num_envs = n
            self.observation_space = observation_space
            self.action_space = action_space
            self.reset_infos = [{} for _ in range(n)]

    # ------------------------------------------------------------------ VecEnv API (SB3 2.3.0 base_vec_env.py)
    def _next_obs64(self) -> np.ndarray:
        """The array this step's float64 observations go to: the two preallocated batches alternate, so the array of the previous
        step (SB3 keeps it as ``_last_obs`` until the rollout buffer has copied it) stays intact for one more step."""
        self._obs_flip ^= 1
        return self._obs_f64[self._obs_flip]

    def reset(self) -> np.ndarray:
        self.sim.reset_host(self._obs)
        self.reset_infos = [{} for _ in range(self.num_envs)]
        out = self._next_obs64()
        np.copyto(out, self._obs)
        return out

    def step_async(self, actions: np.ndarray) -> None:
        a = np.asarray(actions, dtype=np.float32)
        if a.shape != (self.num_envs, _lib.ACT_DIM):
            # the reference raises from MujocoEnv.do_simulation on a shape mismatch
            raise ValueError(f"Action dimension mismatch. Expected {(self.num_envs, _lib.ACT_DIM)}, found {a.shape}")
        np.copyto(self._actions, a)

    def step_wait(self):
        obs = self._next_obs64()
        self.sim.step_host_f64(self._actions, obs, self._rew, self._done, self._trunc, self._tobs, self._info)
        dones = self._done.astype(bool)
        if self.biped:
            self.sim.biped_info_host(self._binfo)
        infos = StepInfos(self, dones)
        return obs, self._rew.copy(), dones, infos

    def step(self, actions: np.ndarray):
        self.step_async(actions)
        return self.step_wait()

    def _info_dict(self, i: int, r: np.ndarray, b: Optional[np.ndarray], done: bool, trunc: bool) -> dict:
        """The info dict of ``Go1MujocoEnv.step`` (go1_mujoco_env.py:500-521) for environment i from its info row."""
        d = {"x_position": float(r[0]), "y_position": float(r[1]), "distance_from_origin": float(np.hypot(r[0], r[1])),
             "termination_reason": "not_terminated", "termination_details": "No termination", "bipedal_success": False}
        healthy = r[13] > 0.5
        code = int(r[12])
        if code == 1:
            details = "State values are not finite"                       # the reference prints the whole state vector here (:439)
        elif code in (2, 3, 4):
            name, unit, rng = {2: ("Z-position", "", self._healthy_ranges[0]), 3: ("Roll", " rad", self._healthy_ranges[1]),
                               4: ("Pitch", " rad", self._healthy_ranges[2])}[code]
            details = f"{name}: {float(r[16]):.3f}{unit}, Healthy Range: [{rng[0]:.2f}, {rng[1]:.2f}]{unit}"      # :444-459
        else:
            details = "No termination"
        reason = _REASONS.get(code, "not_terminated")
        if done and not trunc:                                                                                       # :511-514
            d["termination_reason"] = f"prolonged_{reason}"
            d["termination_details"] = f"{details} (persisted for > {15.0:.2f}s)"
        elif not healthy:                                                                                            # :515-518
            d["termination_reason"] = "unhealthy_state_active"
            d["termination_details"] = details
        for k, key in enumerate(INFO_KEYS[:6] if self.biped else INFO_KEYS):      # biped mode has no orientation / default-joint terms (:905-909)
            d[key] = float(r[2 + k])
        if self.biped:
            for k, key in enumerate(BIPED_INFO_KEYS):
                d[key] = float(b[k])
            d["bipedal_success"] = bool(b[10] > 0.5)                               # :520-521
        return d

    def __del__(self):
        try:
            self.close()       # unregister the page-locked arrays before numpy frees them
        except Exception:
            pass

    def close(self) -> None:
        for a in getattr(self, "_pinned", []):
            try:
                self.sim.host_unregister(a)
            except _lib.Go1Error:
                pass
        self._pinned = []
        self.sim.close()

    def _shared(self, indices, what):
        """All environments of the slab share one value of the simulator-level attributes (kernel parameters)."""
        idx = self._indices(indices)
        if sorted(set(idx)) != list(range(self.num_envs)):
            raise ValueError(f"{what}: the attribute is shared by all {self.num_envs} environments of the slab; per-index values are "
                             f"not supported (got indices={indices!r})")

    def get_attr(self, attr_name: str, indices=None) -> List[Any]:
        n = len(self._indices(indices))
        if attr_name in ("_rand_power", "rand_power"):
            return [self.sim.get_attr("rand_power")] * n
        if attr_name == "_desired_velocity_min":
            return [np.array([self.sim.get_attr(f"desired_vel_min_{c}") for c in "xyz"])] * n
        if attr_name == "_desired_velocity_max":
            return [np.array([self.sim.get_attr(f"desired_vel_max_{c}") for c in "xyz"])] * n
        if hasattr(self, attr_name):
            return [getattr(self, attr_name)] * n
        raise AttributeError(attr_name)

    def set_attr(self, attr_name: str, value: Any, indices=None) -> None:
        """``training_env.set_attr("_rand_power", v)`` (training_callback.py:64) really reaches the simulator
        (all envs share one value; per-index values are not supported)."""
        if attr_name in ("_rand_power", "rand_power", "_desired_velocity_min", "_desired_velocity_max"):
            self._shared(indices, f"set_attr({attr_name!r})")
        if attr_name in ("_rand_power", "rand_power"):
            self._rand_power = float(value)
            self.sim.set_attr("rand_power", float(value))
        elif attr_name in ("_desired_velocity_min", "_desired_velocity_max"):
            tag = "min" if attr_name.endswith("min") else "max"
            for c, v in zip("xyz", np.asarray(value, dtype=np.float64)):
                self.sim.set_attr(f"desired_vel_{tag}_{c}", float(v))
        else:
            setattr(self, attr_name, value)

    def env_method(self, method_name: str, *method_args, indices=None, **method_kwargs) -> List[Any]:
        """There are no per-environment Python objects behind this VecEnv; the methods callers of the reference use through
        ``env_method`` are answered here, anything else raises."""
        idx = self._indices(indices)
        if method_name in ("get_wrapper_attr", "__getattr__") and method_args:
            return self.get_attr(method_args[0], indices)
        if method_name == "set_wrapper_attr" and len(method_args) >= 2:
            self.set_attr(method_args[0], method_args[1], indices)
            return [None] * len(idx)
        if method_name == "seed":
            return list(self.seed(*method_args))[: len(idx)]
        raise NotImplementedError(f"env_method({method_name!r}): there are no per-env Python objects behind this VecEnv")

    def env_is_wrapped(self, wrapper_class, indices=None) -> List[bool]:
        return [False] * len(self._indices(indices))

    def seed(self, seed: Optional[int] = None) -> Sequence[Optional[int]]:
        # the reference's resets ignore the gymnasium seed (unseeded np.random, go1_mujoco_env.py:953-959,1012);
        # here the Philox key is fixed at construction.  action_space sampling is seeded like make_vec_env does.
        self.action_space.seed(seed)
        return [seed] * self.num_envs

    def get_images(self, max_envs: int = 16):
        """SB3 ``VecEnv.get_images``: one frame per environment (the first ``max_envs``; the rest are None -- 16 384 frames of
        1024 x 768 would be 38 GB).  None everywhere unless ``render_mode == "rgb_array"``."""
        if self.render_mode != "rgb_array":
            return [None] * self.num_envs
        k = min(self.num_envs, int(max_envs))
        frames = self.sim.render(range(k), self.render_width, self.render_height)
        return [frames[i] for i in range(k)] + [None] * (self.num_envs - k)

    def render(self, mode: Optional[str] = None):
        """``Go1MujocoEnv(render_mode="rgb_array", camera_name="tracking", width=1024, height=768).render()`` (train.py:53-61,
        training_callback.py:711) for environment 0: an [height, width, 3] uint8 frame from the CUDA ray-caster (the reference's
        OpenGL renderer is unavailable on this box).  None when the env was not created with ``render_mode="rgb_array"``."""
        if (mode or self.render_mode) != "rgb_array":
            return None
        return self.sim.render((0,), self.render_width, self.render_height)[0]

    def _indices(self, indices):
        if indices is None:
            return list(range(self.num_envs))
        if isinstance(indices, int):
            return [indices]
        return list(indices)


def make_go1_vec_env(env_id=None, n_envs: int = 12, seed: int = 0, start_index: int = 0, monitor_dir=None, wrapper_class=None,
                     env_kwargs: Optional[dict] = None, vec_env_cls=None, vec_env_kwargs: Optional[dict] = None,
                     monitor_kwargs=None, wrapper_kwargs=None, **kw) -> Go1CudaVecEnv:
    """Replacement for the call at reference ``train.py:44-50`` with SB3 ``make_vec_env``'s signature:
    ``make_vec_env(Go1MujocoEnv, env_kwargs=..., n_envs=..., seed=..., vec_env_cls=SubprocVecEnv)``.  ``env_id`` (the env class),
    ``vec_env_cls`` / ``vec_env_kwargs`` and the Monitor arguments are accepted and ignored: the slab IS the vectorised,
    monitored environment (episode records come from the kernel).  ``wrapper_class`` is not supported (no per-env objects)."""
    if wrapper_class is not None:
        raise ValueError("make_go1_vec_env: wrapper_class is not supported (there are no per-env Python objects to wrap)")
    if isinstance(env_id, int) and n_envs == 12:      # make_go1_vec_env(4096): the round-1 positional form
        n_envs, env_id = env_id, None
    env_kwargs = dict(env_kwargs or {})
    kw.setdefault("env_id_offset", int(start_index))
    env = Go1CudaVecEnv(n_envs, seed=seed, **env_kwargs, **kw)
    env.seed(seed)
    return env
