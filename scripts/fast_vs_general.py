"""GPU check: the two-kernel step (fast + general list) against the general kernel alone on identical inputs,
plus timing of both.  Usage: python scripts/fast_vs_general.py [n_envs] [steps]"""
import os, sys, json, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

def run(n, steps, tag):
    from mujoco_jaco_b200.env import JacoVecEnv
    env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
    env.reset()
    rng = np.random.default_rng(0)   # numpy actions: a mismatching env can be replayed on the CPU against the oracle
    acts_np = rng.uniform(-1, 1, (steps, n, 7)).astype(np.float32)
    acts = [torch.from_numpy(acts_np[k]).cuda() for k in range(steps)]
    outs = []
    torch.cuda.synchronize()
    times = []
    for a in acts:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); o, r, d, s = env.step(a); e1.record(); torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
        outs.append((o.cpu().numpy().copy(), r.cpu().numpy().copy(), d.cpu().numpy().copy(), env.debug_counters().copy()))
    q, v, aux = [t.cpu().numpy() for t in env.get_state()]
    np.savez(f"gpurun_out/fvg_{tag}.npz", obs=np.stack([x[0] for x in outs]), rew=np.stack([x[1] for x in outs]),
             done=np.stack([x[2] for x in outs]), cnt=np.stack([x[3] for x in outs]), q=q, v=v, times=np.array(times))
    print(tag, "ms/step", np.round(times, 2).tolist(), "counters(last)", outs[-1][3][:9].tolist(), "stats", env.stats().cpu().tolist()[:9], flush=True)

if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    if len(sys.argv) > 3:
        run(n, steps, sys.argv[3])
    else:
        os.makedirs("gpurun_out", exist_ok=True)
        subprocess.check_call([sys.executable, __file__, str(n), str(steps), "fast"])
        subprocess.check_call([sys.executable, __file__, str(n), str(steps), "general"], env=dict(os.environ, JACO_NO_FAST="1"))
        a, b = np.load("gpurun_out/fvg_fast.npz"), np.load("gpurun_out/fvg_general.npz")
        for k in range(steps):
            same_done = (a["done"][k] == b["done"][k]).mean()
            print(k, "obs maxdiff %.2e  p99 %.2e  rew maxdiff %.2e  done agree %.4f  bail %d reset %d" % (
                np.abs(a["obs"][k] - b["obs"][k]).max(), np.quantile(np.abs(a["obs"][k] - b["obs"][k]).max(1), 0.99),
                np.abs(a["rew"][k] - b["rew"][k]).max(), same_done, a["cnt"][k][0], a["cnt"][k][1]))
