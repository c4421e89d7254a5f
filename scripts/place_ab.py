"""A/B of the step kernel's env placement on the aged picking roll-out: a JACO_DIAG build reads JACO_SORT_WAVE at
jaco_create (0 = identity placement, w = the sorted envs dealt over the first w blocks, -1 = the library's rule: one block per SM).
  python scripts/place_ab.py <lib.so> [envs] [age] [steps] [waves...]"""
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mujoco_jaco_b200 import _lib  # noqa: E402

_lib.load(sys.argv[1])
from mujoco_jaco_b200.env import JacoVecEnv  # noqa: E402

n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
age = int(sys.argv[3]) if len(sys.argv) > 3 else 200
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 20
waves = [int(x) for x in sys.argv[5:]] or [0, 148]
for w in waves:
    os.environ["JACO_SORT_WAVE"] = str(w)
    env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
    env.reset()
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    trace = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for k in range(age + steps):
        a = torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
        e0.record()
        env.step(a)
        e1.record()
        torch.cuda.synchronize()
        if k in (0, 5, 20, 50, 100, 150) or k >= age:
            trace.append((k, e0.elapsed_time(e1)))
    timed = [t for k, t in trace if k >= age]
    st = env.stats().cpu().numpy()
    print(f"wave {w:4d}: ms/step (age {age}) mean {sum(timed) / len(timed):7.2f} min {min(timed):7.2f} max {max(timed):7.2f} | "
          f"trace {[(k, round(t, 1)) for k, t in trace if k < age]} | envs quiet / busy / solving / top key {env.debug_counters()[8:12].tolist()} | episodes {st[0]:.0f} stageB {st[11] / st[7] / 50:.4f} pinv {st[12] / st[7] / 50:.4f}", flush=True)
    env.close()
