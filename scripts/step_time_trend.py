"""Per-step device time of the picking roll-out over a long horizon (workload drift as episodes age).
usage: python scripts/step_time_trend.py [n_envs] [steps] [action_bank]   (action_bank 0 = fresh i.i.d. actions per step, the bench default;
32 = the rounds 1-2 protocol)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mujoco_jaco_b200.env import JacoVecEnv
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 400
bank = int(sys.argv[3]) if len(sys.argv) > 3 else 0
env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
env.reset()
g = torch.Generator(device="cuda").manual_seed(0)
acts = torch.rand(max(bank, 1), n, 7, device="cuda", generator=g) * 2 - 1
acc = 0.0
for k in range(steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a = acts[k % bank] if bank else torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
    e0.record(); env.step(a); e1.record(); torch.cuda.synchronize()
    acc += e0.elapsed_time(e1)
    if k % 25 == 24:
        st = env.stats(clear=True).cpu().tolist()
        print(f"steps {k - 24:4d}-{k:4d}: {acc / 25:7.2f} ms/step  episodes {st[0]:.0f} successes {st[1]:.0f} timeouts {st[3]:.0f} drops {st[4]:.0f} singular {st[5]:.0f} stageB {st[11] / max(st[7], 1) / 50:.4f} pinv {st[12] / max(st[7], 1) / 50:.4f}", flush=True)
        acc = 0.0
