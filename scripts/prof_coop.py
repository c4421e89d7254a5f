"""Cycle breakdown of the warp-cooperative path (needs a -DJACO_PROF build: JACO_B200_LIB=...)."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from mujoco_jaco_b200.env import JacoVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
env.reset()
g = torch.Generator(device="cuda").manual_seed(0)
for k in range(6):
    a = torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); env.step(a); e1.record(); torch.cuda.synchronize()
    out = np.zeros(16, np.uint64)
    env.L.jaco_debug_prof.argtypes = [C.c_void_p]
    rc = env.L.jaco_debug_prof(out.ctypes.data_as(C.c_void_p))
    o = out.astype(np.float64)
    print(k, "ms", round(e0.elapsed_time(e1), 2), "stageA calls", int(o[0]), "cyc/call %.0f" % (o[1] / max(o[0], 1)), "| stageB calls", int(o[8]), "publish %.0f" % (o[9] / max(o[8], 1)),
          "stageB %.0f" % (o[2] / max(o[8], 1)), "collision %.0f" % (o[4] / max(o[8], 1)), "constraints %.0f" % (o[5] / max(o[8], 1)), "| solved", int(o[3]),
          "solve cyc %.0f" % (o[6] / max(o[3], 1)), "iters %.2f" % (o[7] / max(o[3], 1)),
          "| per solve: setup %.0f grad %.0f hess %.0f chol %.0f ls %.0f upd %.0f" % tuple(o[10 + k] / max(o[3], 1) for k in range(6)), flush=True)
