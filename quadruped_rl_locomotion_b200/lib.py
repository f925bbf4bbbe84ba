"""ctypes binding of the C ABI declared in ``include/go1sim.h``.

The product path has exactly one implementation: the nvcc-built ``libgo1sim.so`` (fp32) or
``libgo1sim_f64.so`` (fp64 check build) next to this file.  If it is missing, loading fails loudly
-- there is no CPU fallback.  ``Go1Lib(path=...)`` accepts an explicit path so that the test-suite
can point the same binding at other builds of the same source.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

OBS_DIM, ACT_DIM, NQ, NV, NBODY, INFO_DIM, MAX_CONTACTS = 48, 12, 19, 18, 14, 20, 40
BIPED_INFO_DIM = 12

ACTION_RANDOM, ACTION_POLICY, ACTION_ZERO = 0, 1, 2


class Go1Error(RuntimeError):
    pass


class EnvConfig(C.Structure):
    _fields_ = [
        ("biped", C.c_int32), ("frame_skip", C.c_int32), ("max_episode_steps", C.c_int32),
        ("max_unhealthy_steps", C.c_int32), ("rand_power", C.c_float), ("fallen_start_prob", C.c_float),
        ("desired_vel_min", C.c_float * 3), ("desired_vel_max", C.c_float * 3), ("reward_w", C.c_float * 6),
        ("cost_w", C.c_float * 11), ("healthy_z", C.c_float * 2), ("healthy_roll", C.c_float * 2),
        ("healthy_pitch", C.c_float * 2), ("tracking_sigma", C.c_float), ("obs_clip", C.c_float),
        ("solver_tolerance", C.c_float), ("solver_iterations", C.c_int32), ("biped_w", C.c_float * 10),
        ("biped_ready_joints", C.c_float * 12), ("biped_reset_z", C.c_float), ("biped_reset_pitch_deg", C.c_float),
    ]


class State(C.Structure):
    _fields_ = [
        ("qpos", C.c_double * NQ), ("qvel", C.c_double * NV), ("qacc_warmstart", C.c_double * NV),
        ("last_action", C.c_double * ACT_DIM), ("feet_air_time", C.c_double * 4),
        ("last_health_deviation", C.c_double * 3), ("desired_velocity", C.c_double * 3),
        ("episode_return", C.c_double), ("last_contacts", C.c_int32 * 4), ("step", C.c_int32),
        ("unhealthy_steps", C.c_int32), ("reset_count", C.c_int32), ("bad_state", C.c_int32),
        ("front_feet_touched", C.c_int32), ("pad_", C.c_int32),
    ]


class Diag(C.Structure):
    _fields_ = [
        ("qacc", C.c_double * NV), ("qfrc_actuator", C.c_double * NV), ("qfrc_constraint", C.c_double * NV),
        ("cfrc_ext", (C.c_double * 6) * NBODY), ("trunk_xpos_z", C.c_double), ("subtree_com", C.c_double * 3),
        ("ncon", C.c_int32), ("solver_iter", C.c_int32), ("solver_iter_total", C.c_int32),
        ("contact_overflow", C.c_int32),
    ]


class Contact(C.Structure):
    _fields_ = [("geom", C.c_int32), ("dim", C.c_int32), ("dist", C.c_double), ("pos", C.c_double * 3),
                ("force", C.c_double * 6), ("geom1", C.c_int32), ("pad_", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("sum_reward", "n_episodes", "sum_episode_length", "sum_episode_return",
                                          "n_terminated", "n_truncated", "n_bad_state", "env_steps",
                                          "n_contact_overflow", "sum_contacts", "sum_solver_iterations", "sum_self_contacts")]


class Policy(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("w0", "b0", "w1", "b1", "w2", "b2", "log_std")]


class Camera(C.Structure):
    """``Go1Camera``: a trackcom camera (position in the trunk frame at qpos0, xyaxes, vertical field of view)."""
    _fields_ = [("pos", C.c_float * 3), ("xaxis", C.c_float * 3), ("yaxis", C.c_float * 3), ("fovy_deg", C.c_float)]


class Critic(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("w0", "b0", "w1", "b1", "w2", "b2")]


EXPORTS = (
    "go1_default_config", "go1_precision_bits", "go1_version", "go1_create", "go1_destroy", "go1_last_error",
    "go1_num_envs", "go1_action_bounds", "go1_reset", "go1_step", "go1_step_host", "go1_step_host_f64", "go1_reset_host", "go1_rollout", "go1_policy_act",
    "go1_last_rollout_timing", "go1_set_state", "go1_get_state", "go1_enable_diag", "go1_get_diag",
    "go1_get_contacts", "go1_set_attr_f32", "go1_get_attr_f32", "go1_biped_info_host", "go1_biped_info_dev",
    "go1_host_register", "go1_host_unregister", "go1_stagger_phases", "go1_step_noalloc", "go1_policy_forward", "go1_policy_act_scalar", "go1_render",
)


def library_path(precision: str = "f32") -> str:
    return os.path.join(_HERE, "libgo1sim.so" if precision == "f32" else "libgo1sim_f64.so")


class Go1Lib:
    """Loaded shared library with typed entry points."""

    def __init__(self, path: str | None = None, precision: str = "f32"):
        path = path or library_path(precision)
        if not os.path.exists(path):
            raise Go1Error(
                f"{path} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()'). "
                "There is no CPU fallback.")
        self.path = path
        L = C.CDLL(path)
        for name in EXPORTS:
            if not hasattr(L, name):
                raise Go1Error(f"{path} does not export {name}")
        vp, i32, u64, f32p, u8p = C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p
        L.go1_default_config.argtypes = [C.POINTER(EnvConfig)]; L.go1_default_config.restype = None
        L.go1_precision_bits.restype = i32
        L.go1_version.restype = C.c_char_p
        L.go1_create.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(EnvConfig), i32, i32, u64, u64, C.POINTER(vp)]
        L.go1_destroy.argtypes = [vp]
        L.go1_last_error.argtypes = [vp]; L.go1_last_error.restype = C.c_char_p
        L.go1_num_envs.argtypes = [vp]
        L.go1_action_bounds.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.go1_reset.argtypes = [vp, u8p, f32p, vp]
        L.go1_step.argtypes = [vp, f32p, f32p, f32p, u8p, u8p, f32p, f32p, vp]
        L.go1_step_noalloc.argtypes = [vp, f32p, f32p, f32p, u8p, u8p, f32p, f32p, vp]
        L.go1_step_host.argtypes = [vp, f32p, f32p, f32p, u8p, u8p, f32p, f32p]
        L.go1_step_host_f64.argtypes = [vp, f32p, C.c_void_p, f32p, u8p, u8p, f32p, f32p]
        L.go1_reset_host.argtypes = [vp, u8p, f32p]
        L.go1_rollout.argtypes = [vp, i32, i32, C.POINTER(Policy), C.POINTER(Stats), vp]
        L.go1_policy_act.argtypes = [vp, C.POINTER(Policy), f32p, f32p, i32, vp]
        L.go1_last_rollout_timing.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        L.go1_set_state.argtypes = [vp, i32, C.POINTER(State)]
        L.go1_get_state.argtypes = [vp, i32, C.POINTER(State)]
        L.go1_enable_diag.argtypes = [vp, i32]
        L.go1_get_diag.argtypes = [vp, i32, C.POINTER(Diag)]
        L.go1_get_contacts.argtypes = [vp, i32, C.POINTER(Contact), i32, C.POINTER(C.c_int)]
        L.go1_set_attr_f32.argtypes = [vp, C.c_char_p, C.c_float]
        L.go1_get_attr_f32.argtypes = [vp, C.c_char_p, C.POINTER(C.c_float)]
        L.go1_biped_info_host.argtypes = [vp, f32p]
        L.go1_host_register.argtypes = [vp, vp, C.c_size_t]
        L.go1_host_unregister.argtypes = [vp, vp]
        L.go1_biped_info_dev.argtypes = [vp, C.POINTER(C.c_void_p)]
        L.go1_stagger_phases.argtypes = [vp]
        L.go1_render.argtypes = [vp, C.POINTER(Camera), C.POINTER(C.c_int32), i32, i32, i32, vp, i32]
        L.go1_policy_forward.argtypes = [vp, C.POINTER(Policy), C.POINTER(Critic), vp, vp, vp, vp, vp, i32, vp]
        L.go1_policy_act_scalar.argtypes = [vp, C.POINTER(Policy), vp, vp, i32, vp]
        self.L = L

    def default_config(self) -> EnvConfig:
        cfg = EnvConfig()
        self.L.go1_default_config(C.byref(cfg))
        return cfg

    @property
    def precision_bits(self) -> int:
        return int(self.L.go1_precision_bits())


def _ptr(a):
    """Device pointer (int), torch tensor, numpy array or None -> c_void_p value."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    raise TypeError(type(a))


class Sim:
    """One handle = one slab of environments on one GPU (go1_create ... go1_destroy)."""

    def __init__(self, model_blob: bytes, n_envs: int, *, lib: Go1Lib | None = None, cfg: EnvConfig | None = None,
                 device: int = 0, seed: int = 0, env_id_offset: int = 0):
        self.lib = lib or Go1Lib()
        self.n_envs = int(n_envs)
        self._h = C.c_void_p()
        cfg = cfg if cfg is not None else self.lib.default_config()
        self.cfg = cfg
        rc = self.lib.L.go1_create(model_blob, len(model_blob), C.byref(cfg), self.n_envs, int(device), int(seed),
                                   int(env_id_offset), C.byref(self._h))
        if rc != 0:
            msg = self.lib.L.go1_last_error(None)
            self._h = C.c_void_p()
            raise Go1Error(f"go1_create failed ({rc}): {msg.decode() if msg else ''}")

    # -- helpers
    def _check(self, rc, what):
        if rc != 0:
            msg = self.lib.L.go1_last_error(self._h)
            raise Go1Error(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.L.go1_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def action_bounds(self):
        lo = (C.c_float * 12)(); hi = (C.c_float * 12)()
        self._check(self.lib.L.go1_action_bounds(self._h, lo, hi), "go1_action_bounds")
        return np.array(lo[:], np.float32), np.array(hi[:], np.float32)

    # -- device-pointer API
    def reset(self, mask=None, obs=None, stream=None):
        self._check(self.lib.L.go1_reset(self._h, _ptr(mask), _ptr(obs), stream), "go1_reset")

    def step(self, actions, obs, rew, done, trunc, terminal_obs=None, info=None, stream=None):
        self._check(self.lib.L.go1_step(self._h, _ptr(actions), _ptr(obs), _ptr(rew), _ptr(done), _ptr(trunc),
                                        _ptr(terminal_obs), _ptr(info), stream), "go1_step")

    def step_noalloc(self, actions, obs, rew, done, trunc, terminal_obs=None, info=None, stream=None):
        """``go1_step`` with the kernel's scratch in the library's static device array (``go1_step_noalloc``): no allocation on the
        call path, at most ``NOALLOC_MAX_ENVS`` environments, calls on one device must not overlap."""
        self._check(self.lib.L.go1_step_noalloc(self._h, _ptr(actions), _ptr(obs), _ptr(rew), _ptr(done), _ptr(trunc),
                                                _ptr(terminal_obs), _ptr(info), stream), "go1_step_noalloc")

    # -- host-buffer API (numpy, C-contiguous)
    def step_host(self, actions, obs, rew, done, trunc, terminal_obs=None, info=None):
        assert actions.dtype == np.float32 and actions.flags.c_contiguous and actions.shape == (self.n_envs, 12)
        self._check(self.lib.L.go1_step_host(self._h, _ptr(actions), _ptr(obs), _ptr(rew), _ptr(done), _ptr(trunc),
                                             _ptr(terminal_obs), _ptr(info)), "go1_step_host")

    def step_host_f64(self, actions, obs64, rew, done, trunc, terminal_obs=None, info=None):
        """``step_host`` with the observations delivered as float64 (widened on the device)."""
        assert actions.dtype == np.float32 and actions.flags.c_contiguous and actions.shape == (self.n_envs, 12)
        assert obs64.dtype == np.float64 and obs64.flags.c_contiguous and obs64.shape == (self.n_envs, 48)
        self._check(self.lib.L.go1_step_host_f64(self._h, _ptr(actions), _ptr(obs64), _ptr(rew), _ptr(done), _ptr(trunc),
                                                 _ptr(terminal_obs), _ptr(info)), "go1_step_host_f64")

    def reset_host(self, obs=None, mask=None):
        self._check(self.lib.L.go1_reset_host(self._h, _ptr(mask), _ptr(obs)), "go1_reset_host")

    def rollout(self, n_steps: int, mode: int = ACTION_RANDOM, policy: Policy | None = None, stream=None):
        st = Stats()
        self._check(self.lib.L.go1_rollout(self._h, int(n_steps), int(mode), C.byref(policy) if policy is not None else None,
                                           C.byref(st), stream), "go1_rollout")
        ms = C.c_float(); launches = C.c_int()
        self.lib.L.go1_last_rollout_timing(self._h, C.byref(ms), C.byref(launches))
        out = {k: getattr(st, k) for k, _ in Stats._fields_}
        out["device_ms"] = float(ms.value); out["launches"] = int(launches.value)
        return out

    def policy_act(self, policy: Policy, obs, act, n: int | None = None, stream=None):
        """Actor forward + sampling on device buffers (``go1_policy_act``)."""
        self._check(self.lib.L.go1_policy_act(self._h, C.byref(policy), _ptr(obs), _ptr(act),
                                              int(self.n_envs if n is None else n), stream), "go1_policy_act")

    def policy_act_scalar(self, policy: Policy, obs, act, n: int | None = None, stream=None):
        """The fp32 scalar reference kernel of the same forward (``go1_policy_act_scalar``)."""
        self._check(self.lib.L.go1_policy_act_scalar(self._h, C.byref(policy), _ptr(obs), _ptr(act),
                                                     int(self.n_envs if n is None else n), stream), "go1_policy_act_scalar")

    def policy_forward(self, policy: Policy, critic, obs, act, values=None, log_probs=None, n: int | None = None, stream=None, raw_act=None):
        """ActorCriticPolicy.forward on the tensor cores (``go1_policy_forward``): actions, values, log-probs; device buffers."""
        self._check(self.lib.L.go1_policy_forward(self._h, C.byref(policy), C.byref(critic) if critic is not None else None, _ptr(obs), _ptr(act),
                                                  _ptr(values) if values is not None else None, _ptr(log_probs) if log_probs is not None else None,
                                                  _ptr(raw_act) if raw_act is not None else None,
                                                  int(self.n_envs if n is None else n), stream), "go1_policy_forward")

    # -- state / diagnostics
    def set_state(self, env: int, **kw):
        s = self.get_state_struct(env)
        for k, v in kw.items():
            f = getattr(s, k)
            if hasattr(f, "__len__"):
                v = np.asarray(v).ravel()
                assert len(v) == len(f), k
                for i in range(len(f)):
                    f[i] = type(f[0])(v[i]) if not isinstance(f[0], float) else float(v[i])
            else:
                setattr(s, k, v)
        self._check(self.lib.L.go1_set_state(self._h, int(env), C.byref(s)), "go1_set_state")

    def get_state_struct(self, env: int) -> State:
        s = State()
        self._check(self.lib.L.go1_get_state(self._h, int(env), C.byref(s)), "go1_get_state")
        return s

    def get_state(self, env: int) -> dict:
        s = self.get_state_struct(env)
        out = {}
        for k, _ in State._fields_:
            v = getattr(s, k)
            out[k] = np.array(v[:]) if hasattr(v, "__len__") else v
        return out

    def host_register(self, arr: np.ndarray) -> None:
        """Page-lock a persistent numpy buffer so ``step_host`` DMAs straight into it (keep ``arr`` alive until
        ``host_unregister`` / ``close``)."""
        assert arr.flags.c_contiguous
        self._check(self.lib.L.go1_host_register(self._h, arr.ctypes.data, arr.nbytes), "go1_host_register")

    def host_unregister(self, arr: np.ndarray) -> None:
        self._check(self.lib.L.go1_host_unregister(self._h, arr.ctypes.data), "go1_host_unregister")

    def biped_info_host(self, out: np.ndarray | None = None) -> np.ndarray:
        """[N][BIPED_INFO_DIM] extra reward_info terms of the last control step (biped mode only)."""
        if out is None:
            out = np.zeros((self.n_envs, BIPED_INFO_DIM), np.float32)
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.size == self.n_envs * BIPED_INFO_DIM
        self._check(self.lib.L.go1_biped_info_host(self._h, _ptr(out)), "go1_biped_info_host")
        return out

    def biped_info_dev(self) -> int:
        p = C.c_void_p()
        self._check(self.lib.L.go1_biped_info_dev(self._h, C.byref(p)), "go1_biped_info_dev")
        return int(p.value)

    def enable_diag(self, on: bool = True):
        self._check(self.lib.L.go1_enable_diag(self._h, 1 if on else 0), "go1_enable_diag")

    def get_diag(self, env: int) -> dict:
        d = Diag()
        self._check(self.lib.L.go1_get_diag(self._h, int(env), C.byref(d)), "go1_get_diag")
        out = {}
        for k, _ in Diag._fields_:
            v = getattr(d, k)
            out[k] = np.array(v) if hasattr(v, "__len__") else v
        out["cfrc_ext"] = np.array(out["cfrc_ext"]).reshape(NBODY, 6)
        return out

    def get_contacts(self, env: int):
        buf = (Contact * MAX_CONTACTS)()
        n = C.c_int()
        self._check(self.lib.L.go1_get_contacts(self._h, int(env), buf, MAX_CONTACTS, C.byref(n)), "go1_get_contacts")
        return [dict(geom=b.geom, geom1=b.geom1, dim=b.dim, dist=b.dist, pos=np.array(b.pos[:]), force=np.array(b.force[:]))
                for b in buf[: n.value]]

    def render(self, env_ids=(0,), width: int = 1024, height: int = 768, camera: "Camera | None" = None, out=None):
        """RGB frames [n][height][width][3] uint8 of the listed environments from the tracking camera (``go1_render``: CUDA
        ray-caster, no graphics API).  ``out``: numpy array (host) or torch CUDA tensor (device) to write into."""
        ids = (C.c_int32 * len(env_ids))(*[int(e) for e in env_ids])
        if out is None:
            out = np.zeros((len(env_ids), int(height), int(width), 3), np.uint8)
        host = isinstance(out, np.ndarray)
        self._check(self.lib.L.go1_render(self._h, C.byref(camera) if camera is not None else None, ids, len(env_ids), int(width), int(height),
                                          _ptr(out), 1 if host else 0), "go1_render")
        return out

    def stagger_phases(self):
        """Throughput runs: spread the episode phases of a slab that was reset all at once (go1_stagger_phases)."""
        self._check(self.lib.L.go1_stagger_phases(self._h), "go1_stagger_phases")

    def set_attr(self, name: str, value: float):
        self._check(self.lib.L.go1_set_attr_f32(self._h, name.encode(), float(value)), "go1_set_attr_f32")

    def get_attr(self, name: str) -> float:
        v = C.c_float()
        self._check(self.lib.L.go1_get_attr_f32(self._h, name.encode(), C.byref(v)), "go1_get_attr_f32")
        return float(v.value)
