"""Development probe: size of the dirty list / fall-back list / general-reset list per env step as the population ages."""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch
from mujoco_jaco_b200.env import JacoVecEnv
n = 65536
env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
env.reset()
g = torch.Generator(device="cuda"); g.manual_seed(1)
acts = torch.rand(16, n, 7, device="cuda", generator=g) * 2 - 1
for k in range(260):
    env.step(acts[k % 16])
    if k % 20 == 0 or k in (1, 2, 3, 5):
        c = env.debug_counters()
        st = env.stats(clear=True).cpu().numpy()
        print(f"age {k:3d} fallback {c[0]:5d} genreset {c[1]:5d} dirty {c[8]:6d} ({100 * c[8] / n:.1f}%) reasons {c[2:8]}  stageB/substep {st[11] / max(st[7],1) / 50 * 100:.2f}% pinv {st[12] / max(st[7],1) / 50 * 100:.2f}% fingerobj {st[13] / max(st[7],1) / 50 * 100:.3f}% episodes {st[0]:.0f}", flush=True)
