"""Development probe: device time per env step when every env holds the SAME state and gets the SAME action (all
control flow uniform across lanes and warps) against the normal roll-out: measures how much of the step time is
control-flow divergence between warps (lost instruction-stream sharing, barrier waits).
usage: python scripts/uniform_probe.py [lib.so] [age]"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch
from mujoco_jaco_b200 import _lib
if len(sys.argv) > 1 and sys.argv[1] != "main":
    _lib.SO_PATH = os.path.join(REPO, "mujoco_jaco_b200", f"libjaco_b200_{sys.argv[1]}.so"); _lib.needs_build = lambda: False
age = int(sys.argv[2]) if len(sys.argv) > 2 else 10
from mujoco_jaco_b200.env import JacoVecEnv
n = 65536
env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
env.reset()
g = torch.Generator(device="cuda"); g.manual_seed(1)
acts = torch.rand(16, n, 7, device="cuda", generator=g) * 2 - 1
def timeit(actions, k=6):
    for i in range(3): env.step(actions[i])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(k): env.step(actions[(3 + i) % 16])
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k
for i in range(age): env.step(acts[i % 16])
print("roll-out          ms/step %.2f" % timeit(acts))
q, v, a = env.get_state()
for src in (0, 5, 11):
    env.set_state(q[src:src + 1].repeat(n, 1), v[src:src + 1].repeat(n, 1), a[src:src + 1].repeat(n, 1))
    ua = acts[:, src:src + 1, :].repeat(1, n, 1).contiguous()
    print("all envs = env %2d ms/step %.2f" % (src, timeit(ua)))
# group-uniform: every group of G consecutive envs (a warp: 32, a block: 256) holds one state / action, groups differ
for G in (32, 256, 4096):
    idx = (torch.arange(n, device="cuda") // G) * G
    env.set_state(q[idx], v[idx], a[idx])
    ua = acts[:, idx, :].contiguous()
    print("uniform per group of %4d ms/step %.2f" % (G, timeit(ua)))
