"""Per-block timeline of one step-kernel launch on the aged picking roll-out (JACO_DIAG build: globaltimer at block start / end).
  python scripts/block_times.py <diag lib.so> [envs] [age] [JACO_SORT_WAVE values...]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mujoco_jaco_b200 import _lib  # noqa: E402

L = _lib.load(sys.argv[1])
from mujoco_jaco_b200.env import JacoVecEnv  # noqa: E402

n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
age = int(sys.argv[3]) if len(sys.argv) > 3 else 200
for w in [int(x) for x in sys.argv[4:]] or [-1, 0]:
    os.environ["JACO_SORT_WAVE"] = str(w)
    env = JacoVecEnv(n, task="picking", seed=0, auto_reset=True)
    env.reset()
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    for k in range(age):
        env.step(torch.rand(n, 7, device="cuda", generator=g) * 2 - 1)
    torch.cuda.synchronize()
    cnt = env.debug_counters()
    T = np.zeros((3, 2048), np.uint64)
    L.jaco_diag_block_times(T.ctypes.data_as(C.c_void_p))
    nb = (n + 255) // 256
    t0 = T[0, :nb].min()
    start = (T[0, :nb] - t0).astype(np.float64) / 1e6
    dur = (T[1, :nb] - T[0, :nb]).astype(np.float64) / 1e6
    wave = 0 if w == 0 else (min(nb, 148) if w < 0 else min(w, nb))     # the library's rule: one first-wave block per SM
    first = start < 0.5           # blocks of the first hardware wave
    print(f"JACO_SORT_WAVE {w}: first-wave blocks {wave}, envs quiet / busy / solving / top key {cnt[8:12].tolist()}, kernel span {(T[1, :nb].max() - t0) / 1e6:.2f} ms", flush=True)
    if wave > 0:
        print(f"  first-wave blocks [0, {wave}): duration mean {dur[:wave].mean():.2f} min {dur[:wave].min():.2f} max {dur[:wave].max():.2f} ms")
        if wave < nb:
            print(f"  quiet blocks [{wave}, {nb}): duration mean {dur[wave:].mean():.2f} min {dur[wave:].min():.2f} max {dur[wave:].max():.2f} ms; start mean {start[wave:].mean():.2f} max {start[wave:].max():.2f}")
    else:
        print(f"  all blocks: duration mean {dur.mean():.2f} min {dur.min():.2f} max {dur.max():.2f} ms; first-wave blocks {int(first.sum())}, late start mean {start[~first].mean():.2f}")
    print("  duration deciles:", np.round(np.percentile(dur, [0, 10, 25, 50, 75, 90, 100]), 2).tolist())
    env.close()
