"""ctypes wrapper around ``oracle/libgo1oracle.so`` -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu-baseline / reference arm may
import this module.  It exposes a ``mujoco``-flavoured surface (``OracleSim.data.qpos`` ...) so
that the reference's unmodified ``go1_mujoco_env.py`` can run on top of it (``oracle/ref_env.py``)
and so that parity tests read like MuJoCo code:

    sim = OracleSim(model)            # model: quadruped_rl_locomotion_b200.mjcf.Model
    sim.reset_data()                  # mj_resetData
    sim.data.qpos[:] = model.key_qpos[0]
    sim.step(10); sim.rne_post_constraint()   # gymnasium MujocoEnv.do_simulation (go1_mujoco_env.py:481)
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc). Building the checker is not using it."""
    so = os.path.join(_HERE, "libgo1oracle.so")
    src = os.path.join(_HERE, "go1_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libgo1oracle.so"])
    return so


def _lib():
    global _LIB
    if _LIB is None:
        so = build()
        L = ctypes.CDLL(so)
        L.o1_model_load.restype = ctypes.c_void_p
        L.o1_model_load.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
        L.o1_model_free.argtypes = [ctypes.c_void_p]
        L.o1_data_new.restype = ctypes.c_void_p
        L.o1_data_new.argtypes = [ctypes.c_void_p]
        L.o1_data_free.argtypes = [ctypes.c_void_p]
        for f in ("o1_reset_data", "o1_forward", "o1_step", "o1_rne_post_constraint"):
            getattr(L, f).argtypes = [ctypes.c_void_p]
            getattr(L, f).restype = None
        L.o1_step_n.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.o1_step_n.restype = None
        L.o1_field.restype = ctypes.POINTER(ctypes.c_double)
        L.o1_field.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
        L.o1_ifield.restype = ctypes.POINTER(ctypes.c_int)
        L.o1_ifield.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
        L.o1_contact.restype = ctypes.c_int
        L.o1_contact.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
        _LIB = L
    return _LIB


_SHAPES = {
    "xpos": 3, "xquat": 4, "xmat": 9, "xipos": 3, "ximat": 9, "xanchor": 3, "xaxis": 3, "geom_xpos": 3,
    "geom_xmat": 9, "site_xpos": 3, "site_xmat": 9, "subtree_com": 3, "cinert": 10, "crb": 10, "cdof": 6,
    "cdof_dot": 6, "cvel": 6, "cacc": 6, "cfrc_ext": 6, "efc_KBIP": 4,
}
_CONTACT_FIELDS = ("dist", "pos", "frame", "includemargin", "friction", "solref", "solimp", "mu", "dim",
                   "geom1", "geom2", "exclude", "efc_address")


class OracleContact:
    __slots__ = _CONTACT_FIELDS

    def __init__(self, rec):
        self.dist = rec[0]; self.pos = rec[1:4].copy(); self.frame = rec[4:13].copy()
        self.includemargin = rec[13]; self.friction = rec[14:19].copy(); self.solref = rec[19:21].copy()
        self.solimp = rec[21:26].copy(); self.mu = rec[26]; self.dim = int(rec[27]); self.geom1 = int(rec[28])
        self.geom2 = int(rec[29]); self.exclude = int(rec[30]); self.efc_address = int(rec[31])


class _Data:
    """Live numpy views into the C arrays (like ``mujoco.MjData``)."""

    def __init__(self, sim):
        object.__setattr__(self, "_sim", sim)

    def _view(self, name):
        sim = self._sim
        n = ctypes.c_int(0)
        p = _lib().o1_field(sim._d, name.encode(), ctypes.byref(n))
        if not p:
            return None
        a = np.ctypeslib.as_array(p, shape=(max(n.value, 0),)) if n.value > 0 else np.zeros(0)
        if name in _SHAPES and n.value > 0:
            a = a.reshape(-1, _SHAPES[name])
        if name == "efc_J" and n.value > 0:
            a = a.reshape(-1, sim.model.nv)
        return a

    def __getattr__(self, name):
        sim = self._sim
        if name in ("ncon", "nefc", "nf", "solver_iter", "ls_total", "warn_bad", "warn_contact_full", "warn_unsupported_pair",
                    "self_collision"):
            n = ctypes.c_int(0)
            return int(_lib().o1_ifield(sim._d, name.encode(), ctypes.byref(n))[0])
        if name in ("efc_type", "efc_id", "efc_state"):
            n = ctypes.c_int(0)
            p = _lib().o1_ifield(sim._d, name.encode(), ctypes.byref(n))
            return np.ctypeslib.as_array(p, shape=(n.value,)).copy() if n.value else np.zeros(0, np.int32)
        if name in ("time", "solver_improvement", "solver_gradient"):
            return float(self._view(name)[0])
        if name == "contact":
            out = []
            rec = (ctypes.c_double * 32)()
            for i in range(self.ncon):
                _lib().o1_contact(sim._d, i, rec)
                out.append(OracleContact(np.array(rec[:])))
            return out
        v = self._view(name)
        if v is None:
            raise AttributeError(name)
        return v

    def __setattr__(self, name, value):
        if name == "time":
            self._view("time")[0] = value
            return
        v = self._view(name)
        if v is None:
            raise AttributeError(name)
        v[...] = value


class OracleSim:
    def __init__(self, model):
        self.model = model
        blob = model.to_blob()
        self._m = _lib().o1_model_load(blob, len(blob))
        if not self._m:
            raise RuntimeError("oracle: model blob rejected")
        self._d = _lib().o1_data_new(self._m)
        self.data = _Data(self)
        self.reset_data()

    def __del__(self):
        try:
            if getattr(self, "_d", None):
                _lib().o1_data_free(self._d)
            if getattr(self, "_m", None):
                _lib().o1_model_free(self._m)
        except Exception:
            pass

    def reset_data(self):
        _lib().o1_reset_data(self._d)

    def forward(self):
        _lib().o1_forward(self._d)

    def step(self, nstep: int = 1):
        _lib().o1_step_n(self._d, int(nstep))

    def rne_post_constraint(self):
        _lib().o1_rne_post_constraint(self._d)

    def set_self_collision(self, on: bool):
        """Robot-robot pairs on (default, MuJoCo's behaviour for this MJCF) or off (floor contacts only)."""
        n = ctypes.c_int(0)
        _lib().o1_ifield(self._d, b"self_collision", ctypes.byref(n))[0] = 1 if on else 0

    def set_tolerance(self, tol: float):
        """Test knob: override the Newton tolerance (negative -> model value)."""
        self.data._view("tolerance_override")[0] = tol

    # convenience -------------------------------------------------------------------------
    def full_M(self):
        m = self.model
        qM = self.data.qM
        M = np.zeros((m.nv, m.nv))
        for i in range(m.nv):
            a = m.dof_Madr[i]
            j = i
            while j >= 0:
                M[i, j] = M[j, i] = qM[a]
                a += 1
                j = m.dof_parentid[j]
        return M

    def get_state(self):
        d = self.data
        return dict(qpos=d.qpos.copy(), qvel=d.qvel.copy(), ctrl=d.ctrl.copy(),
                    qacc_warmstart=d.qacc_warmstart.copy(), time=d.time)

    def set_state(self, s):
        d = self.data
        d.qpos[:] = s["qpos"]; d.qvel[:] = s["qvel"]; d.ctrl[:] = s["ctrl"]
        d.qacc_warmstart[:] = s["qacc_warmstart"]; d.time = s.get("time", 0.0)
