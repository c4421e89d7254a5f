"""Minimal MJCF reader for the Jaco model family (offline model compiler, stage 1).

Reads the subset of MJCF that /root/reference/env_script/assets/jaco2/jaco2_curtain_torque.xml uses
(compiler angle=radian, bodies, joints, inertials, box/sphere/plane/mesh geoms, sites, actuators,
touch sensors) into plain Python objects.  No MuJoCo involved; semantics restated from the MJCF
documentation.  This is build-time tooling: nothing here runs on the hot path.
"""
from __future__ import annotations

import os
import struct
from dataclasses import dataclass, field
from xml.etree import ElementTree

import numpy as np

# ----------------------------------------------------------------------------- quaternion helpers (w,x,y,z)


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
    axis = np.asarray(axis, dtype=np.float64)
    s = np.sin(angle / 2)
    return np.array([np.cos(angle / 2), axis[0] * s, axis[1] * s, axis[2] * s])


def euler_xyz_quat(e):
    """MJCF default eulerseq="xyz": intrinsic rotations about x, then y', then z''  (R = Rx Ry Rz)."""
    q = np.array([1.0, 0, 0, 0])
    for ax, ang in zip(np.eye(3), e):
        q = quat_mul(q, axis_angle_quat(ax, ang))
    return q


def normalize(v):
    v = np.asarray(v, dtype=np.float64)
    n = np.linalg.norm(v)
    return v / n if n > 0 else v


# ----------------------------------------------------------------------------- data classes


@dataclass
class Geom:
    name: str
    type: str
    size: np.ndarray
    pos: np.ndarray
    quat: np.ndarray
    density: float = 1000.0
    mesh: str | None = None
    contype: int = 1
    conaffinity: int = 1
    condim: int = 3
    friction: np.ndarray = field(default_factory=lambda: np.array([1.0, 0.005, 0.0001]))
    solref: np.ndarray = field(default_factory=lambda: np.array([0.02, 1.0]))
    solimp: np.ndarray = field(default_factory=lambda: np.array([0.9, 0.95, 0.001, 0.5, 2.0]))
    margin: float = 0.0


@dataclass
class Site:
    name: str
    type: str
    size: np.ndarray
    pos: np.ndarray
    quat: np.ndarray


@dataclass
class Joint:
    name: str
    type: str  # hinge | free
    axis: np.ndarray
    pos: np.ndarray
    ref: float = 0.0
    limited: bool = False
    range: tuple = (0.0, 0.0)
    damping: float = 0.0


@dataclass
class Body:
    name: str
    parent: int
    pos: np.ndarray
    quat: np.ndarray
    mocap: bool = False
    joints: list = field(default_factory=list)
    geoms: list = field(default_factory=list)
    sites: list = field(default_factory=list)
    inertial: dict | None = None  # explicit <inertial>
    children: list = field(default_factory=list)


def _vec(s, n=None, default=None):
    if s is None:
        return None if default is None else np.array(default, dtype=np.float64)
    v = np.array([float(x) for x in s.split()], dtype=np.float64)
    if n is not None and len(v) < n and default is not None:
        out = np.array(default, dtype=np.float64)
        out[: len(v)] = v
        return out
    return v


def _frame(el):
    pos = _vec(el.get("pos"), default=[0, 0, 0])
    if el.get("quat") is not None:
        quat = normalize(_vec(el.get("quat")))
    elif el.get("euler") is not None:
        quat = euler_xyz_quat(_vec(el.get("euler")))
    else:
        quat = np.array([1.0, 0, 0, 0])
    return pos, quat


class MJCF:
    """Parsed model: bodies[0] is the world."""

    def __init__(self, xml_path):
        self.xml_path = xml_path
        root = ElementTree.parse(xml_path).getroot()
        comp = root.find("compiler")
        assert comp is not None and comp.get("angle") == "radian"
        self.meshdir = os.path.join(os.path.dirname(xml_path), comp.get("meshdir", ""))
        self.balanceinertia = comp.get("balanceinertia", "false") == "true"
        self.mesh_files = {}
        for m in root.findall("asset/mesh"):
            fn = m.get("file")
            name = m.get("name") or os.path.splitext(os.path.basename(fn))[0]
            self.mesh_files[name] = (fn, _vec(m.get("scale"), default=[1, 1, 1]))
        self.bodies: list[Body] = [Body("world", -1, np.zeros(3), np.array([1.0, 0, 0, 0]))]
        self._auto = 0
        wb = root.find("worldbody")
        self._read_children(wb, 0)
        # actuators
        self.actuators = []
        for a in root.find("actuator"):
            self.actuators.append(dict(
                kind=a.tag, name=a.get("name"), joint=a.get("joint"),
                kp=float(a.get("kp", "1")),
                ctrllimited=a.get("ctrllimited", "false") == "true",
                ctrlrange=_vec(a.get("ctrlrange"), default=[0, 0]),
                forcelimited=a.get("forcelimited", "false") == "true",
                forcerange=_vec(a.get("forcerange"), default=[0, 0]),
            ))
        self.sensors = [dict(kind=s.tag, name=s.get("name"), site=s.get("site")) for s in root.find("sensor")]

    # -- tree
    def _read_children(self, el, bid):
        body = self.bodies[bid]
        for ch in el:
            if ch.tag == "geom":
                body.geoms.append(self._geom(ch))
            elif ch.tag == "site":
                pos, quat = _frame(ch)
                body.sites.append(Site(ch.get("name"), ch.get("type", "sphere"),
                                       _vec(ch.get("size"), default=[0.005]), pos, quat))
            elif ch.tag == "joint" or ch.tag == "freejoint":
                if ch.tag == "freejoint":
                    body.joints.append(Joint(ch.get("name") or f"free_{body.name}", "free",
                                             np.zeros(3), np.zeros(3)))
                else:
                    assert ch.get("type", "hinge") == "hinge"
                    rng = _vec(ch.get("range"), default=[0, 0])
                    body.joints.append(Joint(ch.get("name"), "hinge", normalize(_vec(ch.get("axis"))),
                                             _vec(ch.get("pos"), default=[0, 0, 0]), float(ch.get("ref", "0")),
                                             ch.get("limited", "false") == "true", (rng[0], rng[1]),
                                             float(ch.get("damping", "0"))))
            elif ch.tag == "inertial":
                body.inertial = dict(pos=_vec(ch.get("pos")), mass=float(ch.get("mass")),
                                     diaginertia=_vec(ch.get("diaginertia")))
            elif ch.tag == "body":
                pos, quat = _frame(ch)
                nb = Body(ch.get("name"), bid, pos, quat, mocap=ch.get("mocap", "false") == "true")
                self.bodies.append(nb)
                nid = len(self.bodies) - 1
                body.children.append(nid)
                self._read_children(ch, nid)

    def _geom(self, el):
        pos, quat = _frame(el)
        gtype = el.get("type", "sphere")
        name = el.get("name")
        if name is None:
            name = f"_geom{self._auto}"
            self._auto += 1
        g = Geom(name, gtype, _vec(el.get("size"), default=[0, 0, 0]), pos, quat,
                 density=float(el.get("density", "1000")), mesh=el.get("mesh"),
                 contype=int(el.get("contype", "1")), conaffinity=int(el.get("conaffinity", "1")),
                 condim=int(el.get("condim", "3")), margin=float(el.get("margin", "0")))
        if el.get("friction") is not None:
            g.friction = _vec(el.get("friction"), 3, [1.0, 0.005, 0.0001])
        if el.get("solref") is not None:
            g.solref = _vec(el.get("solref"), 2, [0.02, 1.0])
        if el.get("solimp") is not None:
            g.solimp = _vec(el.get("solimp"), 5, [0.9, 0.95, 0.001, 0.5, 2.0])
        return g

    def body_id(self, name):
        for i, b in enumerate(self.bodies):
            if b.name == name:
                return i
        raise KeyError(name)

    # -- meshes
    def load_mesh(self, name):
        fn, scale = self.mesh_files[name]
        return load_stl(os.path.join(self.meshdir, fn)) * scale


def load_stl(path):
    """Binary STL -> float64 array [ntri, 3 verts, 3]."""
    raw = open(path, "rb").read()
    n = struct.unpack("<I", raw[80:84])[0]
    assert len(raw) == 84 + 50 * n, f"{path}: not a binary STL"
    rec = np.frombuffer(raw[84:], dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]))
    return rec["v"].astype(np.float64)


def mesh_mass_props(tris):
    """Volume, centre of mass and inertia tensor about the COM (unit density) of a closed triangle
    mesh by signed tetrahedra against the origin."""
    a, b, c = tris[:, 0], tris[:, 1], tris[:, 2]
    vol6 = np.einsum("ij,ij->i", a, np.cross(b, c))
    vol = vol6.sum() / 6.0
    com = ((a + b + c) / 4.0 * (vol6 / 6.0)[:, None]).sum(0) / vol
    # second moments: integral of x_i x_j over each tetra (origin,a,b,c) = vol/20 * (sum_k v_k^i v_k^j + s^i s^j), s = a+b+c
    s = a + b + c
    C = np.zeros((3, 3))
    for v in (a, b, c, s):
        C += np.einsum("n,ni,nj->ij", vol6 / 6.0, v, v)
    C /= 20.0
    C -= vol * np.outer(com, com)  # shift to COM
    I = np.trace(C) * np.eye(3) - C
    return vol, com, I
