"""ctypes binding of the CPU ORACLE (oracle/jaco_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; the product (mujoco_jaco_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
TASKS = {"reaching": 0, "grasping": 1, "picking": 2, "placing": 3, "pickAndplace": 4, "releasing": 5, "carrying": 6}
VARIANTS = {"base": 0, "cis": 1, "trans": 2}
NQ, NV, NU, NOBS, NAUX = 23, 21, 9, 26, 64
_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def load_blob():
    z = np.load(os.path.join(REPO, "mujoco_jaco_b200", "model", "jaco2_curtain_torque.npz"))
    return np.ascontiguousarray(z["blob"], dtype=np.float64)


_libs = {}


def build_fast():
    """-O3 -march=native build for the CPU baseline; built on the machine that times it.  Returns True on success."""
    try:
        subprocess.check_call(["make", "-s", "-C", HERE, "fast"])
        return os.path.exists(os.path.join(HERE, "libjaco_oracle_fast.so"))
    except Exception:
        return False


def lib(f32=False, fast=False):
    name = "libjaco_oracle_fast.so" if fast else ("libjaco_oracle_f32.so" if f32 else "libjaco_oracle.so")
    if name not in _libs:
        path = os.path.join(HERE, name)
        if not os.path.exists(path):
            build_fast() if fast else build()
        L = C.CDLL(path)
        L.jo_create.restype = C.c_void_p
        L.jo_create.argtypes = [_dp, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int]
        L.jo_destroy.argtypes = [C.c_void_p]
        L.jo_reset.argtypes = [C.c_void_p, _fp]
        L.jo_step.argtypes = [C.c_void_p, _dp, _fp, _dp, _ip, _ip, _dp]
        L.jo_get_state.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.jo_set_state.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.jo_mj_step.argtypes = [C.c_void_p, _dp, _ip, _dp]
        L.jo_forward.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _ip, _dp]
        L.jo_get_contacts.argtypes = [C.c_void_p, _dp, C.c_int]
        L.jo_get_contacts.restype = C.c_int
        L.jo_get_efc.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp, C.c_int]
        L.jo_get_efc.restype = C.c_int
        L.jo_render_prims.argtypes = [C.c_void_p, _dp, C.c_int]
        L.jo_render_prims.restype = C.c_int
        L.jo_osc.argtypes = [C.c_void_p] + [_dp] * 8
        L.jo_solver_stats.argtypes = [C.c_void_p, _ip, _dp, _ip]
        L.jo_touch_knife_edge.argtypes = [C.c_void_p]
        L.jo_set_probe.argtypes = [C.c_void_p, C.c_int]
        L.jo_osc_knife_edge.argtypes = []
        L.jo_osc_knife_edge_clear.argtypes = []
        L.jo_philox.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32)]
        _libs[name] = L
    return _libs[name]


def _d(a):
    return a.ctypes.data_as(_dp)


class OracleEnv:
    """One reference-semantics environment on the CPU (JacoMujocoEnv restated)."""

    def __init__(self, task="picking", variant="base", obs_mode=1, seed=0, env_id=0, f32=False, blob=None, fast=False, binary=False):
        self.L = lib(f32, fast)
        self.blob = load_blob() if blob is None else blob
        v = VARIANTS[variant] | (obs_mode << 4) | (int(bool(binary)) << 8)
        self.h = self.L.jo_create(_d(self.blob), len(self.blob), TASKS[task], v, seed, env_id)
        if not self.h:
            raise RuntimeError("jo_create failed (blob size mismatch?)")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.jo_destroy(self.h)
            self.h = None

    def reset(self):
        obs = np.zeros(NOBS, np.float32)
        self.L.jo_reset(self.h, obs.ctypes.data_as(_fp))
        return obs

    def step(self, action):
        a = np.zeros(7)
        a[: len(action)] = action
        obs = np.zeros(NOBS, np.float32)
        r, wb = C.c_double(), C.c_double()
        d, s = C.c_int(), C.c_int()
        self.L.jo_step(self.h, _d(a), obs.ctypes.data_as(_fp), C.byref(r), C.byref(d), C.byref(s), C.byref(wb))
        return obs, r.value, bool(d.value), s.value, wb.value

    def get_state(self):
        q, v, aux = np.zeros(NQ), np.zeros(NV), np.zeros(NAUX)
        self.L.jo_get_state(self.h, _d(q), _d(v), _d(aux))
        return q, v, aux

    def set_state(self, qpos, qvel, aux=None):
        q = np.ascontiguousarray(qpos, np.float64)
        v = np.ascontiguousarray(qvel, np.float64)
        a = None if aux is None else np.ascontiguousarray(aux, np.float64)
        self.L.jo_set_state(self.h, _d(q), _d(v), None if a is None else _d(a))

    def mj_step(self, ctrl):
        c = np.ascontiguousarray(ctrl, np.float64)
        ncon = C.c_int()
        sd = np.zeros(20)
        self.L.jo_mj_step(self.h, _d(c), C.byref(ncon), _d(sd))
        return ncon.value, sd

    def forward(self, ctrl=None):
        c = None if ctrl is None else np.ascontiguousarray(ctrl, np.float64)
        M, b, J = np.zeros((NV, NV)), np.zeros(NV), np.zeros((6, 6))
        p, q, a, sd = np.zeros(3), np.zeros(4), np.zeros(NV), np.zeros(20)
        ncon = C.c_int()
        self.L.jo_forward(self.h, None if c is None else _d(c), _d(M), _d(b), _d(J), _d(p), _d(q), _d(a),
                          C.byref(ncon), _d(sd))
        return dict(M=M, bias=b, J=J, ee_pos=p, ee_quat=q, qacc=a, ncon=ncon.value, sensordata=sd)

    def contacts(self, maxcon=96):
        out = np.zeros((maxcon, 10))
        n = self.L.jo_get_contacts(self.h, _d(out), maxcon)
        return out[:n]

    def efc(self, maxrows=1024):
        """constraint problem of the last forward: (J [nefc, 21], D, aref, qacc_smooth)"""
        J, D, aref, qs = np.zeros((maxrows, NV)), np.zeros(maxrows), np.zeros(maxrows), np.zeros(NV)
        n = self.L.jo_get_efc(self.h, _d(J), _d(D), _d(aref), _d(qs), maxrows)
        assert n <= maxrows
        return J[:n], D[:n], aref[:n], qs

    def render_prims(self, maxprim=64):
        """world primitives of the headless camera at the current qpos: float64 [n, 16] = pos 3, R 9 (row-major), size 3, type"""
        out = np.zeros((maxprim, 16))
        n = self.L.jo_render_prims(self.h, _d(out), maxprim)
        return out[:n]

    def osc(self, J, M, bias, ee_pos, ee_quat, dq, target):
        u = np.zeros(6)
        args = [np.ascontiguousarray(x, np.float64) for x in (J, M, bias, ee_pos, ee_quat, dq, target)]
        self.L.jo_osc(self.h, *[_d(x) for x in args], _d(u))
        return u

    def set_probe(self, on=True):
        """enable the contact-margin probe (every forward also collides with margins +-2 um)"""
        self.L.jo_set_probe(self.h, int(on))

    def osc_knife_edge(self, clear=True):
        """True when an OSC call since the last clear sat on the pinv branch threshold or the rcond cut-off (process-wide flag)"""
        r = bool(self.L.jo_osc_knife_edge())
        if clear:
            self.L.jo_osc_knife_edge_clear()
        return r

    def touch_knife_edge(self):
        """True when a touch-sensor attribution of the last forward flips if the sensor site is resized by +-2 um, or (with
        set_probe) a geom pair is within 2 um of making / breaking contact"""
        return bool(self.L.jo_touch_knife_edge(self.h))

    def solver_stats(self):
        it, n = C.c_int(), C.c_int()
        g = C.c_double()
        self.L.jo_solver_stats(self.h, C.byref(it), C.byref(g), C.byref(n))
        return it.value, g.value, n.value


def philox(seed, c0, c1, c2, c3):
    out = (C.c_uint32 * 4)()
    lib().jo_philox(seed, c0, c1, c2, c3, out)
    return np.array(list(out), dtype=np.uint32)
