"""Host-side mirror of the reference's environment surface on top of the C ABI.

* JacoVecEnv   -- n_envs environments resident on one B200; torch tensors in / out, stream ordered.
* JacoMujocoEnv -- the reference's single-env class (env_script/env_mujoco.py:9-167) with the same constructor
  kwargs, attributes and return types (numpy obs f32[26], float reward, bool done, {0: 0}), so that
  main.py:48,254,260 and the stable-baselines HPC loop can drive it unchanged.
PyTorch is used for device memory and streams only; all arithmetic happens in libjaco_b200.so.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib

HERE = os.path.dirname(os.path.abspath(__file__))
NQ, NV, NU, NOBS, NACT, NAUX, NSTATS = 23, 21, 9, 26, 7, 72, 16
TASKS = {"reaching": 0, "grasping": 1, "picking": 2, "placing": 3, "pickAndplace": 4, "releasing": 5, "carrying": 6}
VARIANTS = {"base": 0, "cis": 1, "trans": 2}
STAT_NAMES = ["episodes", "successes", "return_sum", "timeouts", "drops", "singular_exits", "nan_resets",
              "env_steps", "length_sum"]


# Camera 0 of the reference's model: <camera fovy="60"/> default class (jaco2_curtain_torque.xml:9) and
# <camera name="visual_camera" pos="0 0.2 0.5" euler="1.3 0 0"> (xml:286; compiler angle="radian", default eulerseq "xyz").
CAMERA0 = dict(pos=(0.0, 0.2, 0.5), euler=(1.3, 0.0, 0.0), fovy=60.0)
RENDER_PRIMS = 35      # include/jaco_b200.h JACO_RENDER_PRIMS: the 33 compiled geoms + 2 upper-arm "bone" capsules


def camera_vector(pos=CAMERA0["pos"], euler=CAMERA0["euler"], fovy=CAMERA0["fovy"], extent=1.0, znear=0.01, zfar=50.0, ambient=0.35):
    """float32[19] for jaco_render: pos, R = Rx Ry Rz (MuJoCo's intrinsic xyz euler sequence; columns = camera right / up / backward),
    fovy (degrees), znear * extent, zfar * extent (MuJoCo: vis.map.znear / zfar scaled by stat.extent), ambient."""
    a, b, c = euler
    Rx = np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
    Ry = np.array([[np.cos(b), 0, np.sin(b)], [0, 1, 0], [-np.sin(b), 0, np.cos(b)]])
    Rz = np.array([[np.cos(c), -np.sin(c), 0], [np.sin(c), np.cos(c), 0], [0, 0, 1]])
    return np.concatenate([np.asarray(pos, np.float64), (Rx @ Ry @ Rz).reshape(-1), [fovy, znear * extent, zfar * extent, ambient, 0, 0, 0]]).astype(np.float32)


def render_palette(geom_names=None):
    """Flat colours per primitive (float32[35, 3]): the reference's materials live in its meshes / GL path, unavailable here."""
    rgb = np.tile(np.array([0.75, 0.75, 0.75], np.float32), (RENDER_PRIMS, 1))
    rgb[0] = (0.55, 0.55, 0.55)        # floor
    rgb[1] = (0.50, 0.35, 0.20)        # object holder
    rgb[2] = (0.00, 0.00, 0.50)        # object (xml: rgba="0 0 0.5 0.8")
    rgb[3] = (0.20, 0.50, 0.20)        # destination box
    rgb[4:23] = (0.90, 0.80, 0.10)     # touch pads
    rgb[23:29] = (0.70, 0.70, 0.72)    # finger capsules
    rgb[29] = (0.30, 0.30, 0.32)       # hand
    rgb[30:35] = (0.22, 0.22, 0.25)    # fore-arm links and upper-arm bones
    return rgb


def load_model_blob(robot_file="jaco2_curtain_torque"):
    path = os.path.join(HERE, "model", robot_file + ".npz")
    if not os.path.exists(path):
        raise NotImplementedError(f"no compiled model for robot_file={robot_file!r} ({path}); run "
                                  "python -m mujoco_jaco_b200.model.compiler <xml>")
    return np.ascontiguousarray(np.load(path)["blob"], dtype=np.float64)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class JacoVecEnv:
    """Batched Jaco environment on one GPU (one independent shard of the global env range)."""

    def __init__(self, n_envs, task="picking", variant="base", obs_mode=1, seed=0, device=0, env_offset=0,
                 robot_file="jaco2_curtain_torque", auto_reset=True, model_blob=None, binary=False):
        if not torch.cuda.is_available():
            raise RuntimeError("JacoVecEnv needs a CUDA device: mujoco_jaco_b200 has no CPU fallback")
        if task not in TASKS:
            raise NotImplementedError(f"task {task!r}")
        self.L = _lib.load()
        self.n_envs, self.task, self.auto_reset = int(n_envs), task, auto_reset
        self.device = torch.device("cuda", device)
        # model_blob: a compiled model other than the shipped one (layout of include/jaco_model_layout.h), e.g. another base pose
        blob = load_model_blob(robot_file) if model_blob is None else np.ascontiguousarray(model_blob, np.float64)
        h = C.c_void_p()
        rc = self.L.jaco_create(blob.ctypes.data_as(C.c_void_p), blob.nbytes, self.n_envs, TASKS[task],
                                VARIANTS[variant] | (obs_mode << 4) | (int(bool(binary)) << 8), device, seed, env_offset, C.byref(h))
        if rc != 0:
            msg = self.L.jaco_last_error(h).decode() if h else ""
            raise RuntimeError(f"jaco_create failed ({rc}) {msg}")
        self.h = h
        n, dev = self.n_envs, self.device
        self.obs = torch.zeros(n, NOBS, device=dev)
        self.reward = torch.zeros(n, device=dev)
        self.done = torch.zeros(n, dtype=torch.uint8, device=dev)
        self.succ = torch.zeros(n, dtype=torch.uint8, device=dev)
        self.wb = torch.zeros(n, device=dev)
        self._stats = torch.zeros(NSTATS, device=dev)

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(f"libjaco_b200 call failed ({rc}): {self.L.jaco_last_error(self.h).decode()}")

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "h", None):
            self.L.jaco_destroy(self.h)
            self.h = None

    __del__ = close

    def reset(self, mask=None):
        if mask is not None:
            mask = mask.to(self.device, torch.uint8).contiguous()
        self._check(self.L.jaco_reset(self.h, _ptr(mask), _ptr(self.obs), self._stream()))
        return self.obs

    def step(self, action, term_obs=None):
        """action: float32 cuda tensor [n,7] (or [n,6] for reaching).  Returns views of the env's own buffers."""
        a = action
        if a.shape[1] != NACT:
            a = torch.nn.functional.pad(a, (0, NACT - a.shape[1]))
        a = a.to(self.device, torch.float32).contiguous()
        self._check(self.L.jaco_step(self.h, _ptr(a), _ptr(self.obs), _ptr(self.reward), _ptr(self.done), _ptr(self.succ),
                                     _ptr(self.wb), _ptr(term_obs), int(self.auto_reset), self._stream()))
        return self.obs, self.reward, self.done, self.succ

    def step_host(self, action_np, out=None):
        """Reference-facing path: host (numpy / pinned) buffers in and out, copies inside the call."""
        a = np.ascontiguousarray(action_np, np.float32)
        if out is None:
            out = (np.empty((self.n_envs, NOBS), np.float32), np.empty(self.n_envs, np.float32),
                   np.empty(self.n_envs, np.uint8), np.empty(self.n_envs, np.uint8), np.empty(self.n_envs, np.float32))
        obs, rew, done, succ, wb = out
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        self._check(self.L.jaco_step_host(self.h, vp(a), vp(obs), vp(rew), vp(done), vp(succ), vp(wb),
                                          int(self.auto_reset), self._stream()))
        return out

    def get_state(self):
        n, dev = self.n_envs, self.device
        q, v, a = torch.empty(n, NQ, device=dev), torch.empty(n, NV, device=dev), torch.empty(n, NAUX, device=dev)
        self._check(self.L.jaco_get_state(self.h, _ptr(q), _ptr(v), _ptr(a), self._stream()))
        return q, v, a

    def set_state(self, qpos=None, qvel=None, aux=None):
        f = lambda t: None if t is None else t.to(self.device, torch.float32).contiguous()
        q, v, a = f(qpos), f(qvel), f(aux)
        if a is not None and a.shape[1] != NAUX:
            a = torch.nn.functional.pad(a, (0, NAUX - a.shape[1]))
        self._check(self.L.jaco_set_state(self.h, _ptr(q), _ptr(v), _ptr(a), self._stream()))
        torch.cuda.current_stream(self.device).synchronize()  # keep q/v/a alive until consumed

    def render(self, env_begin=0, env_count=None, width=640, height=480, camera=None, palette=None):
        """_get_camera (util:302-307) for a range of envs: (rgb uint8 [k, H, W, 3] with row 0 at the top, depth float32 [k, H, W],
        OpenGL-normalised like MjRenderContextOffscreen.read_pixels(depth=True)).  Headless CUDA ray caster; see csrc/jaco_render.cuh."""
        k = self.n_envs - env_begin if env_count is None else int(env_count)
        cam = np.ascontiguousarray(camera_vector() if camera is None else camera, np.float32)
        pal = np.ascontiguousarray(render_palette() if palette is None else palette, np.float32)
        assert cam.shape == (19,) and pal.shape == (RENDER_PRIMS, 3)
        rgb = torch.empty(k, height, width, 3, dtype=torch.uint8, device=self.device)
        depth = torch.empty(k, height, width, dtype=torch.float32, device=self.device)
        self._check(self.L.jaco_render(self.h, cam.ctypes.data_as(C.c_void_p), pal.ctypes.data_as(C.c_void_p), int(width), int(height),
                                       int(env_begin), k, _ptr(rgb), _ptr(depth), self._stream()))
        return rgb, depth

    def debug_substep(self, ctrl):
        n, dev = self.n_envs, self.device
        c = ctrl.to(dev, torch.float32).contiguous()
        ncon = torch.zeros(n, dtype=torch.int32, device=dev)
        sd, qacc = torch.zeros(n, 20, device=dev), torch.zeros(n, NV, device=dev)
        self._check(self.L.jaco_debug_substep(self.h, _ptr(c), _ptr(ncon), _ptr(sd), _ptr(qacc), self._stream()))
        return ncon, sd, qacc

    def stats(self, clear=True):
        """Device-side episode counters (float[16] on the GPU): what the multi-GPU harness all-reduces."""
        self._check(self.L.jaco_stats(self.h, _ptr(self._stats), int(clear), self._stream()))
        return self._stats

    def debug_counters(self):
        """Host copy of the step work-list counters: [n_bail, n_reset (last step), bail-reason histogram...]."""
        out = np.zeros(16, np.int32)
        self._check(self.L.jaco_debug_counters(self.h, out.ctypes.data_as(C.c_void_p)))
        return out

    def launch_count(self):
        return int(self.L.jaco_launch_count(self.h))


class _Box:
    """Minimal stand-in for gym.spaces.Box (gym is optional at run time)."""

    def __init__(self, low, high, dtype=np.float32):
        self.low, self.high = np.asarray(low, dtype), np.asarray(high, dtype)
        self.shape, self.dtype = self.low.shape, dtype

    def sample(self):
        return np.random.uniform(self.low, self.high).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))


def _box(low, high):
    try:
        from gym import spaces  # type: ignore
        return spaces.Box(low, high, dtype=np.float32)
    except Exception:
        return _Box(low, high)


class JacoMujocoEnv:
    """Drop-in for env_script.env_mujoco.JacoMujocoEnv (n_envs = 1 on the GPU)."""

    def __init__(self, **kwargs):
        self.n_robots = kwargs.get("n_robots", 1)
        if self.n_robots != 1:
            raise NotImplementedError("\n\t\033[91m[ERROR]: xml_file of the given number of robots doesn't exist\033[0m")
        robot_file = kwargs.get("robot_file", "jaco2_curtain_torque") or "jaco2_curtain_torque"
        self.task = kwargs.get("task", None)
        obs_mode = 2 if kwargs.get("subgoal_obs", False) else (1 if kwargs.get("rulebased_subgoal", False) else 0)
        # the reference draws from unseeded np.random (every worker process its own episodes): without an explicit seed take one from the OS
        seed = kwargs.get("seed", None)
        if seed is None:
            seed = int.from_bytes(os.urandom(4), "little")
        self._vec = JacoVecEnv(1, task=self.task, variant=kwargs.get("variant", "base"), obs_mode=obs_mode, seed=int(seed),
                               device=kwargs.get("device", 0), robot_file=robot_file, auto_reset=False)
        self.current_steps, self.max_steps = 0, 2500
        if self.task in ["reaching", "grasping", "carrying", "releasing", "pushing"]:
            self.task_max_steps = 500
        elif self.task in ["picking", "placing"]:
            self.task_max_steps = 700
        else:
            self.task_max_steps = 1200
        self.skip_frames = 50
        logger_path = kwargs.get("logger_path", "./logger_csv")
        os.makedirs(logger_path, exist_ok=True)
        self.file_logger = open(logger_path + "/SR_" + self.task + "_abalation_GA_learned_2.csv", "w")
        obs_max = np.hstack([[3], [1] * 25]).astype(np.float32)
        self.observation_space = _box(-obs_max, obs_max)
        try:
            self.state_shape = kwargs["stateGen"].get_state_shape()
        except Exception:
            self.state_shape = self.observation_space.shape[0]
        self.pose_action_space_max, self.gripper_action_space_max, self.gripper_action_space_min = 1, 1, -1
        nact = 6 if self.task in ["reaching", "pushing"] else 7
        self.act_max, self.act_min = np.ones(nact), -np.ones(nact)
        self.action_space = _box(self.act_min, self.act_max)
        self.wb, self.accum_rew, self.accum_succ, self.metadata, self.log_save = 0, 0, [0] * 25, None, False
        self.verbose = kwargs.get("verbose", True)
        self.reset()

    def reset(self):
        self.current_steps, self.accum_rew = 0, 0
        return self._vec.reset().cpu().numpy()[0].copy()

    def get_state_shape(self):
        return self.state_shape

    def get_num_observation(self):
        return self.observation_space.shape[0]

    def get_num_action(self):
        return self.action_space.shape[0]

    def get_action_bound(self):
        return self.pose_action_space_max

    def step(self, action, weight=None, subgoal=None, id=None):
        action = np.clip(np.asarray(action, np.float32), self.act_min, self.act_max)
        a = np.zeros((1, NACT), np.float32)
        a[0, : len(action)] = action
        obs, rew, done, succ, wb = self._vec.step_host(a)
        self.current_steps += 1
        total_reward, done, self.wb = float(rew[0]), bool(done[0]), float(wb[0])
        self.accum_rew += total_reward
        if done:
            self.accum_succ.pop()
            self.accum_succ.insert(0, int(succ[0]))
            sr = "{0:3.3f}".format(sum(self.accum_succ) / 25 * 100)
            if self.verbose:
                print("sr: ", sr)
                print("Return: ", self.accum_rew)
            self.file_logger.writelines(sr + "\n")
            self.file_logger.flush()
        assert self.state_shape == obs.shape[1]
        return obs[0].copy(), total_reward, done, {0: 0}

    def get_wb(self):
        return self.wb

    def seed(self, seed):
        pass

    def close(self):
        self._vec.close()
        return None

    def _get_camera(self):
        """util:302-307: (image uint8 [480, 640, 3], depth float32 [480, 640]), both top-down.  Headless CUDA ray caster over the
        simulation's primitives instead of the reference's OpenGL off-screen context (unavailable on this box)."""
        rgb, depth = self._vec.render(0, 1, 640, 480)
        return rgb[0].cpu().numpy(), depth[0].cpu().numpy()

    def set_capture_path(self, path):
        raise NotImplementedError("viewer / GL capture: unavailable on this box")
