"""MJCF compiler: flattens ``scene_torque.xml`` / ``scene_position.xml`` into a flat model.

This replaces the MuJoCo MJCF compiler call the reference makes through
``MujocoEnv.__init__`` (``go1_mujoco_env.py:46-53`` -> ``MjModel.from_xml_path``) for the
subset of MJCF those files use (``unitree_go1/go1_torque.xml:1-234``,
``go1_position.xml:1-251``, ``scene_*.xml:1-23``):

* ``<include>``, ``<compiler angle autolimits>``, ``<option>``
* nested ``<default class>`` trees, ``childclass`` on bodies, per-element ``class``
* bodies with explicit ``<inertial>``, ``<freejoint/>``, hinge ``<joint>``s
* geoms: plane / sphere / capsule / cylinder / box (+ ``fromto``), visual mesh geoms are
  kept as non-colliding placeholders so that geom ids equal MuJoCo's (feet = 20/31/43/55,
  notebook ``test_scripts/go1_mujoco_test_env.ipynb`` cell 4: geom "FL" has id 31)
* ``<motor>`` / ``<position>`` / ``<general>`` actuators on joints, ``<keyframe>``
* the ``mj_setConst`` quantities the constraint solver needs
  (``dof_invweight0``, ``body_invweight0``, ``stat.meaninertia``) computed at ``qpos0``.

The result is a :class:`Model` (named numpy arrays, field names follow ``mjModel``) that
serialises to a self-describing binary blob consumed by the CUDA library
(``go1_create``), and by the CPU oracle under ``oracle/`` (tests only).

Nothing here needs MuJoCo; the known-answer values recorded in the reference notebook are
checked in ``tests/test_mjcf.py``.
"""
from __future__ import annotations

import json
import math
import os
import struct
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

# mjtGeom
GEOM_PLANE, GEOM_HFIELD, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID, GEOM_CYLINDER, GEOM_BOX, GEOM_MESH = range(8)
_GEOM_TYPES = {
    "plane": GEOM_PLANE, "hfield": GEOM_HFIELD, "sphere": GEOM_SPHERE, "capsule": GEOM_CAPSULE,
    "ellipsoid": GEOM_ELLIPSOID, "cylinder": GEOM_CYLINDER, "box": GEOM_BOX, "mesh": GEOM_MESH,
}
# mjtJoint
JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE = range(4)
_JNT_TYPES = {"free": JNT_FREE, "ball": JNT_BALL, "slide": JNT_SLIDE, "hinge": JNT_HINGE}

MJ_MINVAL = 1e-15


class MjcfError(ValueError):
    pass


# ----------------------------------------------------------------------------- math helpers
def _normalize(v):
    v = np.asarray(v, dtype=np.float64)
    n = np.linalg.norm(v)
    if n < MJ_MINVAL:
        raise MjcfError("zero-length vector cannot be normalised")
    return v / n


def quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([
        aw * bw - ax * bx - ay * by - az * bz,
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
    ])


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
    ])


def axis_angle_quat(axis, angle):
    if angle == 0.0:
        return np.array([1.0, 0.0, 0.0, 0.0])
    s = math.sin(angle * 0.5)
    return np.array([math.cos(angle * 0.5), axis[0] * s, axis[1] * s, axis[2] * s])


def z_to_quat(vec):
    """Quaternion rotating the z axis onto ``vec`` (the MJCF ``fromto`` / ``zaxis`` rule)."""
    vec = _normalize(vec)
    axis = np.cross([0.0, 0.0, 1.0], vec)
    s = np.linalg.norm(axis)
    if s < 1e-10:
        axis = np.array([1.0, 0.0, 0.0])
    else:
        axis = axis / s
    ang = math.atan2(s, vec[2])
    return axis_angle_quat(axis, ang)


def _mat_to_quat(m):
    # robust conversion (largest-component branch)
    tr = m[0, 0] + m[1, 1] + m[2, 2]
    if tr > 0:
        s = math.sqrt(tr + 1.0) * 2
        q = [0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s]
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        s = math.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
        q = [(m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s]
    elif m[1, 1] > m[2, 2]:
        s = math.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
        q = [(m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s]
    else:
        s = math.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
        q = [(m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s]
    return _normalize(q)


# ----------------------------------------------------------------------------- defaults
# Built-in MuJoCo defaults for the attributes we read (XML reference, "default" values).
_BUILTIN = {
    "geom": {
        "type": "sphere", "contype": "1", "conaffinity": "1", "condim": "3", "group": "0",
        "priority": "0", "size": "0 0 0", "friction": "1 0.005 0.0001", "solmix": "1",
        "solref": "0.02 1", "solimp": "0.9 0.95 0.001 0.5 2", "margin": "0", "gap": "0",
        "pos": "0 0 0", "quat": "1 0 0 0",
    },
    "joint": {
        "type": "hinge", "pos": "0 0 0", "axis": "0 0 1", "limited": "auto", "range": "0 0",
        "damping": "0", "armature": "0", "frictionloss": "0", "stiffness": "0", "margin": "0",
        "solreflimit": "0.02 1", "solimplimit": "0.9 0.95 0.001 0.5 2",
        "solreffriction": "0.02 1", "solimpfriction": "0.9 0.95 0.001 0.5 2", "ref": "0",
        "springref": "0",
    },
    "site": {"pos": "0 0 0", "quat": "1 0 0 0", "size": "0.005 0.005 0.005", "type": "sphere"},
    "actuator": {
        "ctrllimited": "auto", "forcelimited": "auto", "ctrlrange": "0 0", "forcerange": "0 0",
        "gear": "1 0 0 0 0 0", "gainprm": "1 0 0", "biasprm": "0 0 0", "gaintype": "fixed",
        "biastype": "none", "dyntype": "none",
    },
}
# array attributes whose XML value may be shorter than the full array (partial overwrite)
_ARRAY_LEN = {
    ("geom", "size"): 3, ("geom", "friction"): 3, ("geom", "solimp"): 5, ("geom", "solref"): 2,
    ("joint", "solimplimit"): 5, ("joint", "solreflimit"): 2, ("joint", "solimpfriction"): 5,
    ("joint", "solreffriction"): 2, ("site", "size"): 3, ("actuator", "gear"): 6,
    ("actuator", "gainprm"): 3, ("actuator", "biasprm"): 3,
}
_ACTUATOR_TAGS = ("general", "motor", "position", "velocity")


def _floats(s):
    return [float(x) for x in str(s).split()]


def _overlay(kind: str, base: Dict[str, str], attrs: Dict[str, str]) -> Dict[str, str]:
    """Overlay XML attributes on a default dict; short arrays overwrite a prefix only."""
    out = dict(base)
    for k, v in attrs.items():
        n = _ARRAY_LEN.get((kind, k))
        if n is not None and k in out:
            new = _floats(v)
            old = _floats(out[k])
            old = old + [0.0] * (n - len(old))
            old[: len(new)] = new
            out[k] = " ".join(repr(x) for x in old[: max(n, len(new))])
        else:
            out[k] = v
    return out


def _actuator_shortcut(tag: str, attrs: Dict[str, str]) -> Dict[str, str]:
    """Expand <motor>/<position>/<velocity> shortcuts to <general> attributes."""
    a = dict(attrs)
    if tag == "motor":
        a.setdefault("gaintype", "fixed")
        a["biastype"] = "none"
        a["dyntype"] = "none"
        a["gainprm"] = "1 0 0"
        a["biasprm"] = "0 0 0"
    elif tag == "position":
        if "kp" in a:
            kp = float(a.pop("kp"))
            a["gainprm"] = f"{kp!r} 0 0"
            a["biasprm"] = f"0 {-kp!r} 0"
        a["gaintype"] = "fixed"
        a["biastype"] = "affine"
        a["dyntype"] = "none"
    elif tag == "velocity":
        if "kv" in a:
            kv = float(a.pop("kv"))
            a["gainprm"] = f"{kv!r} 0 0"
            a["biasprm"] = f"0 0 {-kv!r}"
        a["gaintype"] = "fixed"
        a["biastype"] = "affine"
        a["dyntype"] = "none"
    return a


class _DefaultClass:
    def __init__(self, name: str, parent: Optional["_DefaultClass"]):
        self.name = name
        self.parent = parent
        if parent is None:
            self.vals = {k: dict(v) for k, v in _BUILTIN.items()}
        else:
            self.vals = {k: dict(v) for k, v in parent.vals.items()}

    def apply(self, elem: ET.Element):
        tag = elem.tag
        if tag in _ACTUATOR_TAGS:
            attrs = _actuator_shortcut(tag, elem.attrib)
            self.vals["actuator"] = _overlay("actuator", self.vals["actuator"], attrs)
        elif tag in ("geom", "joint", "site"):
            self.vals[tag] = _overlay(tag, self.vals[tag], elem.attrib)
        # mesh / material / camera / light / ... defaults do not affect the physics we build


# ----------------------------------------------------------------------------- model container
_DTYPE_CODES = {np.dtype("float64"): 0, np.dtype("int32"): 1, np.dtype("uint8"): 2}
_CODE_DTYPES = {v: k for k, v in _DTYPE_CODES.items()}
BLOB_MAGIC = b"GO1MDL01"


@dataclass
class Model:
    """Flat compiled model. ``arrays`` maps mjModel-style field names to numpy arrays."""
    arrays: Dict[str, np.ndarray] = field(default_factory=dict)
    names: Dict[str, List[str]] = field(default_factory=dict)

    def __getattr__(self, key):
        arrays = object.__getattribute__(self, "arrays")
        if key in arrays:
            a = arrays[key]
            return a.item() if a.shape == () or (a.ndim == 1 and a.size == 1 and key.startswith(("n", "opt_", "stat_"))) else a
        raise AttributeError(key)

    def name2id(self, kind: str, name: str) -> int:
        """Equivalent of ``mujoco.mj_name2id`` (used at go1_mujoco_env.py:161-167)."""
        try:
            return self.names[kind].index(name)
        except (KeyError, ValueError):
            return -1

    # --- serialisation -----------------------------------------------------
    def to_blob(self) -> bytes:
        """Self-describing little-endian blob: magic, n, table(name[48],dtype,ndim,shape[4],off,nbytes), data."""
        arrays = dict(self.arrays)
        arrays["names_json"] = np.frombuffer(json.dumps(self.names).encode("utf-8"), dtype=np.uint8).copy()
        entries = []
        payload = bytearray()
        for name in sorted(arrays):
            a = np.ascontiguousarray(arrays[name])
            if a.dtype == np.int64:
                a = a.astype(np.int32)
            if a.dtype not in _DTYPE_CODES:
                a = a.astype(np.float64)
            while len(payload) % 8:
                payload.append(0)
            shape = list(a.shape) + [0] * (4 - a.ndim)
            entries.append((name.encode("ascii"), _DTYPE_CODES[a.dtype], a.ndim, shape, len(payload), a.nbytes))
            payload += a.tobytes()
        header = bytearray(BLOB_MAGIC)
        header += struct.pack("<II", len(entries), 0)
        table_size = len(entries) * (48 + 4 + 4 + 16 + 8 + 8)
        data_off = len(header) + table_size
        while data_off % 8:
            data_off += 1
        for (nm, code, ndim, shape, off, nbytes) in entries:
            if len(nm) > 47:
                raise MjcfError(f"field name too long: {nm!r}")
            header += nm.ljust(48, b"\0")
            header += struct.pack("<ii4iqq", code, ndim, *shape, data_off + off, nbytes)
        while len(header) < data_off:
            header.append(0)
        return bytes(header) + bytes(payload)

    @staticmethod
    def from_blob(blob: bytes) -> "Model":
        if blob[:8] != BLOB_MAGIC:
            raise MjcfError("not a go1 model blob")
        n, _ = struct.unpack_from("<II", blob, 8)
        pos = 16
        arrays = {}
        for _ in range(n):
            nm = blob[pos:pos + 48].split(b"\0", 1)[0].decode("ascii")
            code, ndim, s0, s1, s2, s3, off, nbytes = struct.unpack_from("<ii4iqq", blob, pos + 48)
            pos += 48 + 4 + 4 + 16 + 8 + 8
            shape = (s0, s1, s2, s3)[:ndim]
            a = np.frombuffer(blob, dtype=_CODE_DTYPES[code], count=nbytes // _CODE_DTYPES[code].itemsize, offset=off)
            arrays[nm] = a.reshape(shape).copy()
        names = json.loads(bytes(arrays.pop("names_json")).decode("utf-8"))
        return Model(arrays=arrays, names=names)

    def save(self, path: str):
        with open(path, "wb") as f:
            f.write(self.to_blob())

    @staticmethod
    def load(path: str) -> "Model":
        with open(path, "rb") as f:
            return Model.from_blob(f.read())


# ----------------------------------------------------------------------------- parser
class _Body:
    def __init__(self, name, parent, pos, quat, childclass):
        self.name = name
        self.parent = parent  # _Body or None
        self.pos = pos
        self.quat = quat
        self.childclass = childclass
        self.inertial = None  # (ipos, iquat, mass, diaginertia)
        self.joints = []
        self.geoms = []
        self.sites = []
        self.children = []
        self.id = -1


def _read_orientation(attrs: Dict[str, str], angle_deg: bool, default_quat="1 0 0 0"):
    if "quat" in attrs:
        return _normalize(_floats(attrs["quat"]))
    if "axisangle" in attrs:
        v = _floats(attrs["axisangle"])
        ang = math.radians(v[3]) if angle_deg else v[3]
        return axis_angle_quat(_normalize(v[:3]), ang)
    if "euler" in attrs:
        e = _floats(attrs["euler"])
        if angle_deg:
            e = [math.radians(x) for x in e]
        q = np.array([1.0, 0, 0, 0])
        for i, ax in enumerate(np.eye(3)):  # default eulerseq "xyz", intrinsic
            q = quat_mul(q, axis_angle_quat(ax, e[i]))
        return _normalize(q)
    if "xyaxes" in attrs:
        v = _floats(attrs["xyaxes"])
        x = _normalize(v[:3])
        y = np.array(v[3:6]) - x * np.dot(x, v[3:6])
        y = _normalize(y)
        z = np.cross(x, y)
        return _mat_to_quat(np.stack([x, y, z], axis=1))
    if "zaxis" in attrs:
        return z_to_quat(_floats(attrs["zaxis"]))
    return _normalize(_floats(default_quat))


class MjcfCompiler:
    def __init__(self, path: str):
        self.path = os.path.abspath(path)
        self.root = self._load_with_includes(self.path)
        self.angle_deg = True
        self.autolimits = False
        self.opt = {
            "timestep": 0.002, "gravity": [0.0, 0.0, -9.81], "impratio": 1.0, "tolerance": 1e-8,
            "ls_tolerance": 0.01, "iterations": 100, "ls_iterations": 50, "cone": "pyramidal",
            "integrator": "Euler", "solver": "Newton", "noslip_iterations": 0,
        }
        self.classes: Dict[str, _DefaultClass] = {}
        self.world = _Body("world", None, np.zeros(3), np.array([1.0, 0, 0, 0]), None)
        self.actuators = []
        self.keys = []

    # --- xml loading -------------------------------------------------------
    def _load_with_includes(self, path: str) -> ET.Element:
        root = ET.parse(path).getroot()
        if root.tag != "mujoco":
            raise MjcfError(f"{path}: root element must be <mujoco>")
        base = os.path.dirname(path)

        def expand(parent: ET.Element):
            i = 0
            while i < len(parent):
                ch = parent[i]
                if ch.tag == "include":
                    inc = self._load_with_includes(os.path.join(base, ch.attrib["file"]))
                    parent.remove(ch)
                    for k, sub in enumerate(list(inc)):
                        parent.insert(i + k, sub)
                    i += len(inc)
                else:
                    expand(ch)
                    i += 1
        expand(root)
        return root

    # --- sections ----------------------------------------------------------
    def _parse_compiler_option(self):
        for c in self.root.findall("compiler"):
            if "angle" in c.attrib:
                self.angle_deg = c.attrib["angle"] == "degree"
            if "autolimits" in c.attrib:
                self.autolimits = c.attrib["autolimits"] == "true"
        for o in self.root.findall("option"):
            for k, v in o.attrib.items():
                if k in ("timestep", "impratio", "tolerance", "ls_tolerance"):
                    self.opt[k] = float(v)
                elif k in ("iterations", "ls_iterations", "noslip_iterations"):
                    self.opt[k] = int(v)
                elif k == "gravity":
                    self.opt[k] = _floats(v)
                elif k in ("cone", "integrator", "solver"):
                    self.opt[k] = v
            if o.find("flag") is not None:
                for k, v in o.find("flag").attrib.items():
                    if v != {"warmstart": "enable", "contact": "enable", "frictionloss": "enable",
                             "limit": "enable", "eulerdamp": "enable", "refsafe": "enable",
                             "gravity": "enable", "actuation": "enable", "constraint": "enable",
                             "passive": "enable", "clampctrl": "enable", "filterparent": "enable"}.get(k, v):
                        raise MjcfError(f"<flag {k}={v}> is not supported")

    def _parse_defaults(self):
        main = _DefaultClass("main", None)
        self.classes["main"] = main

        # apply own elements before descending so that children inherit them regardless of order
        def rec_ordered(elem: ET.Element, cls: _DefaultClass):
            for ch in elem:
                if ch.tag != "default":
                    cls.apply(ch)
            for ch in elem:
                if ch.tag == "default":
                    name = ch.attrib.get("class")
                    if name is None:
                        raise MjcfError("nested <default> needs a class name")
                    sub = _DefaultClass(name, cls)
                    self.classes[name] = sub
                    rec_ordered(ch, sub)
        for d in self.root.findall("default"):
            rec_ordered(d, main)

    def _class_for(self, elem: ET.Element, body: Optional[_Body]) -> _DefaultClass:
        name = elem.attrib.get("class")
        if name is None:
            b = body
            while b is not None and name is None:
                name = b.childclass
                b = b.parent
        if name is None:
            name = "main"
        if name not in self.classes:
            raise MjcfError(f"unknown default class {name!r}")
        return self.classes[name]

    def _parse_body(self, elem: ET.Element, body: _Body):
        for ch in elem:
            t = ch.tag
            if t == "body":
                pos = np.array(_floats(ch.attrib.get("pos", "0 0 0")))
                quat = _read_orientation(ch.attrib, self.angle_deg)
                cc = ch.attrib.get("childclass")
                sub = _Body(ch.attrib.get("name", ""), body, pos, quat, cc)
                body.children.append(sub)
                self._parse_body(ch, sub)
            elif t == "inertial":
                if "fullinertia" in ch.attrib:
                    raise MjcfError("fullinertia is not supported")
                body.inertial = (
                    np.array(_floats(ch.attrib.get("pos", "0 0 0"))),
                    _read_orientation(ch.attrib, self.angle_deg),
                    float(ch.attrib["mass"]),
                    np.array(_floats(ch.attrib["diaginertia"])),
                )
            elif t == "freejoint":
                body.joints.append({"type": "free", "name": ch.attrib.get("name", ""), "_free": True})
            elif t == "joint":
                cls = self._class_for(ch, body)
                a = _overlay("joint", cls.vals["joint"], {k: v for k, v in ch.attrib.items() if k != "class"})
                body.joints.append(a)
            elif t == "geom":
                cls = self._class_for(ch, body)
                a = _overlay("geom", cls.vals["geom"], {k: v for k, v in ch.attrib.items() if k != "class"})
                body.geoms.append(a)
            elif t == "site":
                cls = self._class_for(ch, body)
                a = _overlay("site", cls.vals["site"], {k: v for k, v in ch.attrib.items() if k != "class"})
                body.sites.append(a)
            elif t in ("camera", "light"):
                pass
            else:
                raise MjcfError(f"unsupported element <{t}> inside <body>")

    def _parse_actuators(self):
        for sec in self.root.findall("actuator"):
            for ch in sec:
                if ch.tag not in _ACTUATOR_TAGS:
                    raise MjcfError(f"unsupported actuator <{ch.tag}>")
                cls = self._class_for(ch, None)
                attrs = _actuator_shortcut(ch.tag, {k: v for k, v in ch.attrib.items() if k != "class"})
                base = dict(cls.vals["actuator"])
                if ch.tag == "motor":
                    # a <motor> never inherits gain/bias set by a <position> default of the same class
                    pass
                a = _overlay("actuator", base, attrs)
                if "joint" not in a:
                    raise MjcfError("only joint transmissions are supported")
                self.actuators.append(a)

    def _parse_keys(self):
        for sec in self.root.findall("keyframe"):
            for k in sec.findall("key"):
                self.keys.append(dict(k.attrib))

    # --- compile -----------------------------------------------------------
    def compile(self) -> Model:
        self._parse_compiler_option()
        self._parse_defaults()
        for wb in self.root.findall("worldbody"):
            self._parse_body(wb, self.world)
        self._parse_actuators()
        self._parse_keys()

        # depth-first body order (MuJoCo's)
        bodies: List[_Body] = []

        def order(b: _Body):
            b.id = len(bodies)
            bodies.append(b)
            for c in b.children:
                order(c)
        order(self.world)
        nbody = len(bodies)

        A: Dict[str, np.ndarray] = {}
        names = {"body": [b.name for b in bodies], "joint": [], "geom": [], "site": [], "actuator": [], "key": []}

        body_parentid = np.zeros(nbody, np.int32)
        body_rootid = np.zeros(nbody, np.int32)
        body_pos = np.zeros((nbody, 3)); body_quat = np.tile([1.0, 0, 0, 0], (nbody, 1))
        body_ipos = np.zeros((nbody, 3)); body_iquat = np.tile([1.0, 0, 0, 0], (nbody, 1))
        body_mass = np.zeros(nbody); body_inertia = np.zeros((nbody, 3))
        body_jntnum = np.zeros(nbody, np.int32); body_jntadr = -np.ones(nbody, np.int32)
        body_dofnum = np.zeros(nbody, np.int32); body_dofadr = -np.ones(nbody, np.int32)
        body_geomnum = np.zeros(nbody, np.int32); body_geomadr = -np.ones(nbody, np.int32)

        jnt = {k: [] for k in ("type", "qposadr", "dofadr", "bodyid", "pos", "axis", "limited", "range",
                               "solref", "solimp", "margin")}
        dof = {k: [] for k in ("bodyid", "jntid", "parentid", "armature", "damping", "frictionloss",
                               "solref", "solimp")}
        qpos0: List[float] = []
        geom = {k: [] for k in ("type", "contype", "conaffinity", "condim", "bodyid", "priority", "size", "pos",
                                "quat", "friction", "margin", "gap", "solmix", "solref", "solimp", "rbound", "group")}
        site = {k: [] for k in ("bodyid", "pos", "quat")}

        last_dof_of_body = -np.ones(nbody, np.int32)

        for b in bodies:
            i = b.id
            if b.parent is not None:
                body_parentid[i] = b.parent.id
                body_rootid[i] = i if b.parent.id == 0 else body_rootid[b.parent.id]
                body_pos[i] = b.pos
                body_quat[i] = b.quat
                if b.inertial is None:
                    raise MjcfError(f"body {b.name!r}: explicit <inertial> required (inertiafromgeom is not implemented)")
                body_ipos[i], body_iquat[i], body_mass[i], body_inertia[i] = b.inertial
            # joints / dofs
            parent_dof = -1
            if b.parent is not None:
                p = b.parent
                while p is not None and last_dof_of_body[p.id] < 0:
                    p = p.parent
                parent_dof = int(last_dof_of_body[p.id]) if p is not None else -1
            body_jntnum[i] = len(b.joints)
            if b.joints:
                body_jntadr[i] = len(jnt["type"])
                body_dofadr[i] = len(dof["bodyid"])
            for j in b.joints:
                jid = len(jnt["type"])
                names["joint"].append(j.get("name", ""))
                if j.get("_free"):
                    jt = JNT_FREE
                    jnt["pos"].append(np.zeros(3)); jnt["axis"].append(np.array([0.0, 0, 1]))
                    jnt["limited"].append(0); jnt["range"].append([0.0, 0.0]); jnt["margin"].append(0.0)
                    jnt["solref"].append([0.02, 1.0]); jnt["solimp"].append([0.9, 0.95, 0.001, 0.5, 2.0])
                    ndof, nqp = 6, 7
                    q0 = list(bodies[i].pos) + list(bodies[i].quat)
                    damping = armature = frictionloss = 0.0
                    dsolref, dsolimp = [0.02, 1.0], [0.9, 0.95, 0.001, 0.5, 2.0]
                else:
                    jt = _JNT_TYPES[j["type"]]
                    if jt not in (JNT_HINGE, JNT_SLIDE):
                        raise MjcfError("only free / hinge / slide joints are supported")
                    jnt["pos"].append(np.array(_floats(j["pos"])))
                    jnt["axis"].append(_normalize(_floats(j["axis"])))
                    rng = _floats(j["range"])
                    if jt == JNT_HINGE and self.angle_deg:
                        rng = [math.radians(x) for x in rng]
                    lim = j["limited"]
                    if lim == "auto":
                        if not self.autolimits and "range" in j and (rng[0] != 0 or rng[1] != 0):
                            # MuJoCo 2.3.0 errors here ("range specified without limited"); keep strict
                            raise MjcfError("joint range given but limited unspecified and autolimits is off")
                        lim = "true" if (rng[0] != 0 or rng[1] != 0) else "false"
                    jnt["limited"].append(1 if lim == "true" else 0)
                    jnt["range"].append(rng); jnt["margin"].append(float(j["margin"]))
                    jnt["solref"].append(_floats(j["solreflimit"])); jnt["solimp"].append(_floats(j["solimplimit"]))
                    ndof, nqp = 1, 1
                    q0 = [float(j["ref"])]
                    damping, armature, frictionloss = float(j["damping"]), float(j["armature"]), float(j["frictionloss"])
                    dsolref, dsolimp = _floats(j["solreffriction"]), _floats(j["solimpfriction"])
                    if float(j["stiffness"]) != 0.0:
                        raise MjcfError("joint stiffness is not supported")
                jnt["type"].append(jt); jnt["qposadr"].append(len(qpos0)); jnt["dofadr"].append(len(dof["bodyid"]))
                jnt["bodyid"].append(i)
                qpos0 += q0
                for k in range(ndof):
                    d = len(dof["bodyid"])
                    dof["bodyid"].append(i); dof["jntid"].append(jid)
                    dof["parentid"].append(parent_dof)
                    parent_dof = d
                    dof["armature"].append(armature); dof["damping"].append(damping)
                    dof["frictionloss"].append(frictionloss)
                    dof["solref"].append(dsolref); dof["solimp"].append(dsolimp)
                last_dof_of_body[i] = parent_dof
            body_dofnum[i] = sum(6 if j.get("_free") else 1 for j in b.joints)

        # geoms are numbered body by body (world body's geoms first)
        for b in bodies:
            i = b.id
            body_geomnum[i] = len(b.geoms)
            if b.geoms:
                body_geomadr[i] = len(geom["type"])
            for g in b.geoms:
                names["geom"].append(g.get("name", ""))
                gt = _GEOM_TYPES[g["type"]]
                size = _floats(g["size"]) + [0.0, 0.0, 0.0]
                size = size[:3]
                pos = np.array(_floats(g["pos"]))
                quat = _read_orientation(g, self.angle_deg)
                if "fromto" in g:
                    ft = _floats(g["fromto"])
                    if gt not in (GEOM_CAPSULE, GEOM_CYLINDER, GEOM_BOX, GEOM_ELLIPSOID):
                        raise MjcfError("fromto needs capsule/cylinder/box/ellipsoid")
                    vec = np.array(ft[0:3]) - np.array(ft[3:6])
                    ln = np.linalg.norm(vec)
                    if ln < MJ_MINVAL:
                        raise MjcfError("fromto points too close")
                    if gt in (GEOM_CAPSULE, GEOM_CYLINDER):
                        size[1] = ln / 2
                    else:
                        size[2] = ln / 2
                    pos = 0.5 * (np.array(ft[0:3]) + np.array(ft[3:6]))
                    quat = z_to_quat(vec)
                if gt == GEOM_SPHERE:
                    rb = size[0]
                elif gt == GEOM_CAPSULE:
                    rb = size[0] + size[1]
                elif gt == GEOM_CYLINDER:
                    rb = math.sqrt(size[0] ** 2 + size[1] ** 2)
                elif gt == GEOM_BOX:
                    rb = math.sqrt(size[0] ** 2 + size[1] ** 2 + size[2] ** 2)
                elif gt == GEOM_ELLIPSOID:
                    rb = max(size)
                else:  # plane, mesh (visual placeholder), hfield
                    rb = 0.0
                contype, conaff = int(g["contype"]), int(g["conaffinity"])
                if gt in (GEOM_MESH, GEOM_HFIELD) and (contype or conaff):
                    raise MjcfError("colliding mesh/hfield geoms are not supported")
                geom["type"].append(gt); geom["contype"].append(contype); geom["conaffinity"].append(conaff)
                geom["condim"].append(int(g["condim"])); geom["bodyid"].append(i)
                geom["priority"].append(int(g["priority"])); geom["size"].append(size)
                geom["pos"].append(pos); geom["quat"].append(quat)
                geom["friction"].append(_floats(g["friction"])[:3]); geom["margin"].append(float(g["margin"]))
                geom["gap"].append(float(g["gap"])); geom["solmix"].append(float(g["solmix"]))
                geom["solref"].append(_floats(g["solref"])[:2]); geom["solimp"].append(_floats(g["solimp"])[:5])
                geom["rbound"].append(rb); geom["group"].append(int(g["group"]))
            for s in b.sites:
                names["site"].append(s.get("name", ""))
                site["bodyid"].append(i); site["pos"].append(_floats(s["pos"]))
                site["quat"].append(_read_orientation(s, self.angle_deg))

        nq, nv = len(qpos0), len(dof["bodyid"])
        njnt, ngeom, nsite = len(jnt["type"]), len(geom["type"]), len(site["bodyid"])

        # sparse-M addressing (dof_Madr): row i holds M[i, ancestors(i) incl. i], MuJoCo layout
        dof_parentid = np.array(dof["parentid"], np.int32)
        dof_Madr = np.zeros(nv + 1, np.int32)
        for i in range(nv):
            cnt, j = 0, i
            while j >= 0:
                cnt += 1
                j = dof_parentid[j]
            dof_Madr[i + 1] = dof_Madr[i] + cnt

        # actuators
        nu = len(self.actuators)
        act = {k: [] for k in ("trnid", "gear", "gainprm", "biasprm", "ctrllimited", "ctrlrange",
                               "forcelimited", "forcerange", "biastype")}
        for a in self.actuators:
            names["actuator"].append(a.get("name", ""))
            if a["joint"] not in names["joint"]:
                raise MjcfError(f"actuator joint {a['joint']!r} not found")
            jid = names["joint"].index(a["joint"])
            if jnt["type"][jid] not in (JNT_HINGE, JNT_SLIDE):
                raise MjcfError("actuators must act on hinge/slide joints")
            if a.get("dyntype", "none") != "none" or a.get("gaintype", "fixed") != "fixed":
                raise MjcfError("only stateless fixed-gain actuators are supported")
            cr, fr = _floats(a["ctrlrange"]), _floats(a["forcerange"])

            def lim(flag, r):
                if flag == "auto":
                    if not self.autolimits and (r[0] != 0 or r[1] != 0):
                        raise MjcfError("range given but limited unspecified and autolimits is off")
                    return 1 if (r[0] != 0 or r[1] != 0) else 0
                return 1 if flag == "true" else 0
            act["trnid"].append(jid); act["gear"].append(_floats(a["gear"])[0])
            act["gainprm"].append(_floats(a["gainprm"])[:3]); act["biasprm"].append(_floats(a["biasprm"])[:3])
            act["ctrllimited"].append(lim(a["ctrllimited"], cr)); act["ctrlrange"].append(cr[:2])
            act["forcelimited"].append(lim(a["forcelimited"], fr)); act["forcerange"].append(fr[:2])
            act["biastype"].append({"none": 0, "affine": 1}[a.get("biastype", "none")])

        # keyframes
        nkey = len(self.keys)
        key_qpos = np.tile(np.array(qpos0), (max(nkey, 1), 1))[:nkey]
        key_qvel = np.zeros((nkey, nv)); key_ctrl = np.zeros((nkey, nu))
        for k, kd in enumerate(self.keys):
            names["key"].append(kd.get("name", ""))
            if "qpos" in kd:
                v = _floats(kd["qpos"])
                if len(v) != nq:
                    raise MjcfError("keyframe qpos has wrong size")
                key_qpos[k] = v
            if "qvel" in kd:
                key_qvel[k] = _floats(kd["qvel"])
            if "ctrl" in kd:
                v = _floats(kd["ctrl"])
                if len(v) != nu:
                    raise MjcfError("keyframe ctrl has wrong size")
                key_ctrl[k] = v

        f64 = lambda x, shape=None: np.array(x, np.float64).reshape(shape) if shape else np.array(x, np.float64)
        i32 = lambda x: np.array(x, np.int32)
        A.update(
            nq=i32([nq]), nv=i32([nv]), nu=i32([nu]), nbody=i32([nbody]), njnt=i32([njnt]), ngeom=i32([ngeom]),
            nsite=i32([nsite]), nkey=i32([nkey]), nM=i32([int(dof_Madr[nv])]),
            opt_timestep=f64([self.opt["timestep"]]), opt_gravity=f64(self.opt["gravity"]),
            opt_impratio=f64([self.opt["impratio"]]), opt_tolerance=f64([self.opt["tolerance"]]),
            opt_ls_tolerance=f64([self.opt["ls_tolerance"]]), opt_iterations=i32([self.opt["iterations"]]),
            opt_ls_iterations=i32([self.opt["ls_iterations"]]),
            opt_cone=i32([{"pyramidal": 0, "elliptic": 1}[self.opt["cone"]]]),
            opt_integrator=i32([{"Euler": 0, "RK4": 1, "implicit": 2}[self.opt["integrator"]]]),
            opt_solver=i32([{"PGS": 0, "CG": 1, "Newton": 2}[self.opt["solver"]]]),
            body_parentid=body_parentid, body_rootid=body_rootid, body_pos=body_pos, body_quat=body_quat,
            body_ipos=body_ipos, body_iquat=body_iquat, body_mass=body_mass, body_inertia=body_inertia,
            body_jntnum=body_jntnum, body_jntadr=body_jntadr, body_dofnum=body_dofnum, body_dofadr=body_dofadr,
            body_geomnum=body_geomnum, body_geomadr=body_geomadr,
            jnt_type=i32(jnt["type"]), jnt_qposadr=i32(jnt["qposadr"]), jnt_dofadr=i32(jnt["dofadr"]),
            jnt_bodyid=i32(jnt["bodyid"]), jnt_pos=f64(jnt["pos"], (njnt, 3)), jnt_axis=f64(jnt["axis"], (njnt, 3)),
            jnt_limited=i32(jnt["limited"]), jnt_range=f64(jnt["range"], (njnt, 2)),
            jnt_solref=f64(jnt["solref"], (njnt, 2)), jnt_solimp=f64(jnt["solimp"], (njnt, 5)),
            jnt_margin=f64(jnt["margin"]),
            dof_bodyid=i32(dof["bodyid"]), dof_jntid=i32(dof["jntid"]), dof_parentid=dof_parentid,
            dof_Madr=dof_Madr, dof_armature=f64(dof["armature"]), dof_damping=f64(dof["damping"]),
            dof_frictionloss=f64(dof["frictionloss"]), dof_solref=f64(dof["solref"], (nv, 2)),
            dof_solimp=f64(dof["solimp"], (nv, 5)),
            qpos0=f64(qpos0),
            geom_type=i32(geom["type"]), geom_contype=i32(geom["contype"]), geom_conaffinity=i32(geom["conaffinity"]),
            geom_condim=i32(geom["condim"]), geom_bodyid=i32(geom["bodyid"]), geom_priority=i32(geom["priority"]),
            geom_size=f64(geom["size"], (ngeom, 3)), geom_pos=f64(geom["pos"], (ngeom, 3)),
            geom_quat=f64(geom["quat"], (ngeom, 4)), geom_friction=f64(geom["friction"], (ngeom, 3)),
            geom_margin=f64(geom["margin"]), geom_gap=f64(geom["gap"]), geom_solmix=f64(geom["solmix"]),
            geom_solref=f64(geom["solref"], (ngeom, 2)), geom_solimp=f64(geom["solimp"], (ngeom, 5)),
            geom_rbound=f64(geom["rbound"]), geom_group=i32(geom["group"]),
            site_bodyid=i32(site["bodyid"]), site_pos=f64(site["pos"], (nsite, 3)),
            site_quat=f64(site["quat"], (nsite, 4)),
            actuator_trnid=i32(act["trnid"]), actuator_gear=f64(act["gear"]),
            actuator_gainprm=f64(act["gainprm"], (nu, 3)), actuator_biasprm=f64(act["biasprm"], (nu, 3)),
            actuator_biastype=i32(act["biastype"]),
            actuator_ctrllimited=i32(act["ctrllimited"]), actuator_ctrlrange=f64(act["ctrlrange"], (nu, 2)),
            actuator_forcelimited=i32(act["forcelimited"]), actuator_forcerange=f64(act["forcerange"], (nu, 2)),
            key_qpos=key_qpos, key_qvel=key_qvel, key_ctrl=key_ctrl,
        )
        model = Model(arrays=A, names=names)
        _set_const(model)
        return model


# ----------------------------------------------------------------------------- mj_setConst quantities
def kinematics(m: Model, qpos: np.ndarray):
    """Reference forward kinematics in numpy (compile-time only). Returns a dict of world-frame arrays."""
    nbody, njnt = m.nbody, m.njnt
    xpos = np.zeros((nbody, 3)); xquat = np.tile([1.0, 0, 0, 0], (nbody, 1)); xmat = np.tile(np.eye(3), (nbody, 1, 1))
    xipos = np.zeros((nbody, 3)); ximat = np.tile(np.eye(3), (nbody, 1, 1))
    xanchor = np.zeros((njnt, 3)); xaxis = np.zeros((njnt, 3))
    for i in range(1, nbody):
        p = m.body_parentid[i]
        jn, ja = m.body_jntnum[i], m.body_jntadr[i]
        if jn == 1 and m.jnt_type[ja] == JNT_FREE:
            qa = m.jnt_qposadr[ja]
            pos = qpos[qa:qa + 3].copy()
            quat = _normalize(qpos[qa + 3:qa + 7])
            xanchor[ja] = pos; xaxis[ja] = [0, 0, 1]
        else:
            pos = xpos[p] + xmat[p] @ m.body_pos[i]
            quat = quat_mul(xquat[p], m.body_quat[i])
            for j in range(ja, ja + jn):
                qa = m.jnt_qposadr[j]
                R = quat_to_mat(quat)
                xanchor[j] = pos + R @ m.jnt_pos[j]
                xaxis[j] = R @ m.jnt_axis[j]
                if m.jnt_type[j] == JNT_HINGE:
                    qloc = axis_angle_quat(m.jnt_axis[j], qpos[qa] - m.qpos0[qa])
                    quat = quat_mul(quat, qloc)
                    pos = xanchor[j] - quat_to_mat(quat) @ m.jnt_pos[j]
                elif m.jnt_type[j] == JNT_SLIDE:
                    pos = pos + xaxis[j] * (qpos[qa] - m.qpos0[qa])
        quat = _normalize(quat)
        xpos[i], xquat[i], xmat[i] = pos, quat, quat_to_mat(quat)
        xipos[i] = pos + xmat[i] @ m.body_ipos[i]
        ximat[i] = quat_to_mat(quat_mul(quat, m.body_iquat[i]))
    return dict(xpos=xpos, xquat=xquat, xmat=xmat, xipos=xipos, ximat=ximat, xanchor=xanchor, xaxis=xaxis)


def dense_mass_matrix(m: Model, qpos: np.ndarray):
    """Dense joint-space inertia from body Jacobians (independent of the CRBA used at run time)."""
    k = kinematics(m, qpos)
    nv = m.nv
    M = np.zeros((nv, nv))
    jacs = {}
    for b in range(1, m.nbody):
        jp, jr = body_jacobian(m, k, b, k["xipos"][b])
        jacs[b] = (jp, jr)
        Iw = k["ximat"][b] @ np.diag(m.body_inertia[b]) @ k["ximat"][b].T
        M += m.body_mass[b] * jp.T @ jp + jr.T @ Iw @ jr
    M += np.diag(m.dof_armature)
    return M, k


def body_jacobian(m: Model, k, body: int, point: np.ndarray):
    """Translational / rotational Jacobian (3 x nv each) of a point attached to ``body``."""
    nv = m.nv
    jp, jr = np.zeros((3, nv)), np.zeros((3, nv))
    b = body
    while b > 0:
        for j in range(m.body_jntadr[b], m.body_jntadr[b] + m.body_jntnum[b]):
            d = m.jnt_dofadr[j]
            t = m.jnt_type[j]
            if t == JNT_FREE:
                jp[:, d:d + 3] = np.eye(3)
                R = k["xmat"][b]
                for a in range(3):
                    ax = R[:, a]
                    jr[:, d + 3 + a] = ax
                    jp[:, d + 3 + a] = np.cross(ax, point - k["xpos"][b])
            elif t == JNT_HINGE:
                jr[:, d] = k["xaxis"][j]
                jp[:, d] = np.cross(k["xaxis"][j], point - k["xanchor"][j])
            elif t == JNT_SLIDE:
                jp[:, d] = k["xaxis"][j]
        b = m.body_parentid[b]
    return jp, jr


def _set_const(m: Model):
    """dof_invweight0 / body_invweight0 / meaninertia / subtree mass at qpos0 (engine_setconst.c semantics)."""
    nv, nbody = m.nv, m.nbody
    M, k = dense_mass_matrix(m, m.qpos0)
    Minv = np.linalg.inv(M)
    m.arrays["stat_meaninertia"] = np.array([np.trace(M) / max(nv, 1)])
    biw = np.zeros((nbody, 2))
    for b in range(1, nbody):
        jp, jr = body_jacobian(m, k, b, k["xipos"][b])
        J = np.vstack([jp, jr])
        Ainv = J @ Minv @ J.T
        biw[b, 0] = np.trace(Ainv[:3, :3]) / 3.0
        biw[b, 1] = np.trace(Ainv[3:, 3:]) / 3.0
    diw = np.zeros(nv)
    for j in range(m.njnt):
        d = m.jnt_dofadr[j]
        if m.jnt_type[j] == JNT_FREE:
            diw[d:d + 3] = np.trace(Minv[d:d + 3, d:d + 3]) / 3.0
            diw[d + 3:d + 6] = np.trace(Minv[d + 3:d + 6, d + 3:d + 6]) / 3.0
        else:
            diw[d] = Minv[d, d]
    sub = m.body_mass.copy()
    for b in range(nbody - 1, 0, -1):
        sub[m.body_parentid[b]] += sub[b]
    m.arrays["body_invweight0"] = biw
    m.arrays["dof_invweight0"] = diw
    m.arrays["body_subtreemass"] = sub


def compile_mjcf(path: str) -> Model:
    """Compile an MJCF scene file into a flat :class:`Model`."""
    return MjcfCompiler(path).compile()
