"""Throughput of the headless camera (jaco_render): n envs x 640 x 480 per call, CUDA-event timed, against the HBM-write roofline
(7 bytes per pixel: rgb u8 x 3 + depth f32).  python scripts/render_bench.py [n_envs] [reps]"""
import json
import os
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mujoco_jaco_b200.env import JacoVecEnv  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
W, H = 640, 480
env = JacoVecEnv(n, task="picking", seed=1, auto_reset=True)
env.reset()
for _ in range(3):
    env.step(torch.rand(n, 7, device="cuda") * 2 - 1)
for _ in range(3):
    rgb, depth = env.render(0, n, W, H)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ms = []
for _ in range(reps):
    e0.record()
    rgb, depth = env.render(0, n, W, H)
    e1.record()
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
ms = sorted(ms)[len(ms) // 2]
peak = 6550.7
try:
    peak = json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
pix = n * W * H
print(json.dumps({"metric": "camera_frames_per_sec", "value": n / ms * 1e3, "envs": n, "width": W, "height": H, "ms_per_call": ms,
                  "gpix_per_s": pix / ms / 1e6, "roofline": {"bound": "hbm", "achieved": 7 * pix / ms / 1e6, "peak": peak, "unit": "GB/s",
                                                            "frac": 7 * pix / ms / 1e6 / peak, "bytes_per_pixel": 7},
                  "note": "includes the torch.empty of the two output tensors"}))
