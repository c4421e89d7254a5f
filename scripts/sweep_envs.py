"""Kernel time of one env step versus batch size (latency/throughput curve of the step kernel)."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mujoco_jaco_b200.env import JacoVecEnv

out = []
for n in [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else "128,1024,4096,16384,65536".split(","))]:
    env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
    env.reset()
    g = torch.Generator(device="cuda").manual_seed(0)
    acts = [torch.rand(n, 7, device="cuda", generator=g) * 2 - 1 for _ in range(8)]
    for a in acts[:3]:
        env.step(a)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for a in acts[3:]:
        env.step(a)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    st = env.stats().cpu().tolist()
    out.append(dict(n=n, ms_per_step=ms, env_steps_per_s=n / ms * 1e3, stats=st[:9]))
    print(json.dumps(out[-1]), flush=True)
    env.close()
