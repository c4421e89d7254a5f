"""Cycle breakdown of the warp-cooperative path on an aged population (needs a -DJACO_PROF build, e.g.
_lib.build(defines=("JACO_PROF",), out=".../libjaco_b200_prof.so")).  Fresh i.i.d. actions per step (the bench default).
usage: python scripts/prof_coop.py <prof lib.so> [n_envs] [age_steps] [steps]"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mujoco_jaco_b200 import _lib
_lib.load(sys.argv[1])
from mujoco_jaco_b200.env import JacoVecEnv
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
age = int(sys.argv[3]) if len(sys.argv) > 3 else 200
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
env.reset()
env.L.jaco_debug_prof.argtypes = [C.c_void_p]
g = torch.Generator(device="cuda").manual_seed(0)
act = lambda k: torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
def snap():
    out = np.zeros(16, np.uint64)
    env.L.jaco_debug_prof(out.ctypes.data_as(C.c_void_p))  # clock64 sums of the -DJACO_PROF build
    return out.astype(np.float64)
for k in range(age):
    env.step(act(k))
torch.cuda.synchronize()
prev = snap()
for k in range(steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); env.step(act(age + k)); e1.record(); torch.cuda.synchronize()
    cur = snap(); o = cur - prev; prev = cur
    nb, ns = max(o[8], 1), max(o[3], 1)
    print(f"age {age + k}: {e0.elapsed_time(e1):.1f} ms | stage B calls {int(o[8])} ({100 * o[8] / (n * 50):.2f}% of env-substeps) "
          f"publish {o[9] / nb:.0f} stageB {o[2] / nb:.0f} (collision {o[4] / nb:.0f} constraints {o[5] / nb:.0f}) cyc/call | "
          f"solves {int(o[3])} {o[6] / ns:.0f} cyc, iters {o[7] / ns:.2f}: setup {o[10] / ns:.0f} grad {o[11] / ns:.0f} "
          f"dir {o[12] / ns:.0f} ls {o[14] / ns:.0f} upd {o[15] / ns:.0f}", flush=True)
