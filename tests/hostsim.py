"""ctypes wrapper of tests/harness/libjaco_hostsim.so: the per-env device code compiled for the CPU.

TEST INFRASTRUCTURE: lets the CPU-only suite compare the kernel logic (fp32, same source as the CUDA kernels)
with the oracle.  Mirrors the array interface of mujoco_jaco_b200.env.JacoVecEnv (numpy instead of torch).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
NQ, NV, NU, NOBS, NAUX = 23, 21, 9, 26, 72
NSTATE = NQ + NV + NAUX
TASKS = {"reaching": 0, "grasping": 1, "picking": 2, "placing": 3, "pickAndplace": 4, "releasing": 5, "carrying": 6}
VARIANTS = {"base": 0, "cis": 1, "trans": 2}
_f = C.POINTER(C.c_float)
_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "harness")])
        _lib = C.CDLL(os.path.join(HERE, "harness", "libjaco_hostsim.so"))
        blob = np.load(os.path.join(REPO, "mujoco_jaco_b200", "model", "jaco2_curtain_torque.npz"))["blob"]
        blob = np.ascontiguousarray(blob, np.float64)
        _lib.hh_init(blob.ctypes.data_as(C.POINTER(C.c_double)))
    return _lib


def set_blob(blob=None):
    """(re)initialise the harness' model from `blob` (None = the shipped model).  The model is a global of the harness library."""
    L = lib()
    if blob is None:
        blob = np.load(os.path.join(REPO, "mujoco_jaco_b200", "model", "jaco2_curtain_torque.npz"))["blob"]
    blob = np.ascontiguousarray(blob, np.float64)
    L.hh_init(blob.ctypes.data_as(C.POINTER(C.c_double)))


def layout_offsets():
    """{name: offset} parsed from include/jaco_model_layout.h"""
    import re
    txt = open(os.path.join(REPO, "include", "jaco_model_layout.h")).read()
    return {m.group(1): int(m.group(2)) for m in re.finditer(r"#define JM_([A-Z0-9_]+) (\d+)", txt)}


def fp(a):
    return a.ctypes.data_as(_f)


class HostSim:
    def __init__(self, n, task="picking", variant="base", obs_mode=1, seed=0, env_offset=0, binary=False):
        self.L = lib()
        self.n, self.task, self.seed, self.env_offset = n, TASKS[task], seed, env_offset
        self.variant = VARIANTS[variant] | (obs_mode << 4) | (int(bool(binary)) << 8)
        self.state = np.zeros((NSTATE, n), np.float32)
        self.stats = np.zeros(16, np.float32)

    def reset(self):
        obs = np.zeros((self.n, NOBS), np.float32)
        self.L.hh_reset(self.task, self.variant, C.c_ulonglong(self.seed), self.env_offset, fp(self.state), self.n, fp(obs))
        return obs

    def step(self, action, auto_reset=False):
        a = np.zeros((self.n, 7), np.float32)
        a[:, : action.shape[1]] = action
        obs = np.zeros((self.n, NOBS), np.float32)
        rew, wb = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        done, succ = np.zeros(self.n, np.uint8), np.zeros(self.n, np.uint8)
        self.L.hh_step(self.task, self.variant, C.c_ulonglong(self.seed), self.env_offset, fp(self.state), self.n, fp(a), fp(obs),
                       fp(rew), done.ctypes.data_as(C.POINTER(C.c_ubyte)), succ.ctypes.data_as(C.POINTER(C.c_ubyte)),
                       fp(wb), int(auto_reset), fp(self.stats))
        return obs, rew, done, succ, wb

    def step_fast(self, action, auto_reset=False, allow_fallback=True):
        """Fast path (jaco_fast.cuh) with general-path fall-back; returns (..., why) with the bail reason bits."""
        a = np.zeros((self.n, 7), np.float32)
        a[:, : action.shape[1]] = action
        obs = np.zeros((self.n, NOBS), np.float32)
        rew, wb = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        done, succ = np.zeros(self.n, np.uint8), np.zeros(self.n, np.uint8)
        why = np.zeros(self.n, np.int32)
        self.L.hh_step_fast(self.task, self.variant, C.c_ulonglong(self.seed), self.env_offset, fp(self.state), self.n, fp(a), fp(obs),
                            fp(rew), done.ctypes.data_as(C.POINTER(C.c_ubyte)), succ.ctypes.data_as(C.POINTER(C.c_ubyte)),
                            fp(wb), int(auto_reset), fp(self.stats), why.ctypes.data_as(C.POINTER(C.c_int)), int(allow_fallback))
        return obs, rew, done, succ, wb, why

    def step_hybrid(self, action, auto_reset=False):
        """The production env step (jaco_kernel_coop.cu) emulated sequentially; returns (..., info) with per-env path bits
        (1 = served by the cooperative path, 2 = general-path fall-back, 4 = general reset)."""
        a = np.zeros((self.n, 7), np.float32)
        a[:, : action.shape[1]] = action
        obs = np.zeros((self.n, NOBS), np.float32)
        rew, wb = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        done, succ = np.zeros(self.n, np.uint8), np.zeros(self.n, np.uint8)
        info = np.zeros(self.n, np.int32)
        self.L.hh_step_hybrid(self.task, self.variant, C.c_ulonglong(self.seed), self.env_offset, fp(self.state), self.n, fp(a), fp(obs),
                            fp(rew), done.ctypes.data_as(C.POINTER(C.c_ubyte)), succ.ctypes.data_as(C.POINTER(C.c_ubyte)),
                            fp(wb), int(auto_reset), fp(self.stats), info.ctypes.data_as(C.POINTER(C.c_int)))
        return obs, rew, done, succ, wb, info

    def render(self, i, camera19, palette, W, H):
        """headless camera of env i through the kernel's per-pixel code (no tile culling): (rgb, depth, prims [35, 16])"""
        cam, pal = np.ascontiguousarray(camera19, np.float32), np.ascontiguousarray(palette, np.float32)
        rgb, depth, prims = np.zeros((H, W, 3), np.uint8), np.zeros((H, W), np.float32), np.zeros((len(pal), 16), np.float32)
        self.L.hh_render(fp(self.state), self.n, i, fp(cam), fp(pal), W, H, rgb.ctypes.data_as(C.POINTER(C.c_ubyte)), fp(depth), fp(prims))
        return rgb, depth, prims

    def get_state(self):
        s = self.state
        return s[:NQ].T.copy(), s[NQ:NQ + NV].T.copy(), s[NQ + NV:].T.copy()

    def set_state(self, qpos=None, qvel=None, aux=None):
        if qpos is not None:
            self.state[:NQ] = np.asarray(qpos, np.float32).T
        if qvel is not None:
            self.state[NQ:NQ + NV] = np.asarray(qvel, np.float32).T
        if aux is not None:
            a = np.asarray(aux, np.float32)
            self.state[NQ + NV:NQ + NV + a.shape[1]] = a.T

    def substep(self, ctrl):
        c = np.ascontiguousarray(ctrl, np.float32)
        ncon = np.zeros(self.n, np.int32)
        sd = np.zeros((self.n, 20), np.float32)
        qacc = np.zeros((self.n, NV), np.float32)
        self.L.hh_substep(fp(self.state), self.n, fp(c), ncon.ctypes.data_as(C.POINTER(C.c_int)), fp(sd), fp(qacc))
        return ncon, sd, qacc

    def forward(self, i, ctrl):
        c = np.ascontiguousarray(ctrl, np.float32)
        Mc, bias, J = np.zeros(45, np.float32), np.zeros(21, np.float32), np.zeros((6, 6), np.float32)
        p, q, a = np.zeros(3, np.float32), np.zeros(4, np.float32), np.zeros(21, np.float32)
        ncon, it = C.c_int(), C.c_int()
        con = np.zeros((48, 8), np.float32)
        self.L.hh_forward(fp(self.state), self.n, i, fp(c), fp(Mc), fp(bias), fp(J), fp(p), fp(q), fp(a),
                          C.byref(ncon), C.byref(it), fp(con))
        M = np.zeros((9, 9), np.float32)
        for r in range(9):
            for cc in range(r + 1):
                M[r, cc] = M[cc, r] = Mc[r * (r + 1) // 2 + cc]
        return dict(M=M, bias=bias, J=J, ee_pos=p, ee_quat=q, qacc=a, ncon=ncon.value, iters=it.value,
                    contacts=con[:ncon.value])
