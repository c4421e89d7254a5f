"""Why is one rank slower than another at identical workload fractions?  Reproduces rank r's population of the bench (seed 0,
env_offset r * n, action generator 1234 + r) on one GPU with a JACO_DIAG build and prints the slowest blocks of one step-kernel launch
together with the envs in them that needed cooperative solves (iterations, contacts, cap hits).
  python scripts/rank_tail.py <diag lib.so> <rank> [envs] [steps]"""
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

r = int(sys.argv[2])
n = int(sys.argv[3]) if len(sys.argv) > 3 else 65536
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 210
env = JacoVecEnv(n, task="picking", seed=0, env_offset=r * n, auto_reset=True)
env.reset()
g = torch.Generator(device="cuda"); g.manual_seed(1234 + r)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(steps):
    a = torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
    e0.record(); env.step(a); e1.record()
torch.cuda.synchronize()
T = np.zeros((3, 2048), np.uint64); L.jaco_diag_block_times(T.ctypes.data_as(C.c_void_p))
D = np.zeros((4, 65536), np.int32); L.jaco_diag_env_dbg(D.ctypes.data_as(C.c_void_p))
nb = (n + 255) // 256
dur = (T[1, :nb] - T[0, :nb]).astype(np.float64) / 1e6
print(f"rank {r}: last step {e0.elapsed_time(e1):.2f} ms; block durations: first wave mean {dur[:148].mean():.2f} max {dur[:148].max():.2f}; deciles {np.round(np.percentile(dur[:148], [0, 25, 50, 75, 90, 100]), 2).tolist()}")
blk, it, nc, misc = D[0, :n], D[1, :n], D[2, :n], D[3, :n]
cap, nsolve, npinv = misc % 1000, (misc // 1000) % 100, misc // 100000
q, v, aux = [t.cpu().numpy() for t in env.get_state()]
for b in np.argsort(-dur)[:6]:
    envs = np.nonzero(blk == b)[0]
    s = envs[nsolve[envs] > 0]
    order = s[np.argsort(-it[s])][:6]
    print(f"  block {b}: {dur[b]:.2f} ms, {len(s)} solving envs, iterations (sum over the step) top {it[order].tolist()}, max contacts {nc[order].tolist()}, cap hits {cap[order].tolist()}, "
          f"solves {nsolve[order].tolist()}, pinv substeps max {int(npinv[envs].max())}")
print("  all solving envs: iterations per solve percentiles",
      np.round(np.percentile(it[nsolve > 0] / np.maximum(nsolve[nsolve > 0], 1), [50, 90, 99, 100]), 2).tolist(), "cap hits total", int(cap.sum()))
# blocks ranked by their heaviest env
heavy = np.array([(it[blk == b] / np.maximum(nsolve[blk == b], 1)).max() if (blk == b).any() else 0 for b in range(nb)])
print("  correlation(block duration, heaviest env's iterations per solve) over the first wave:", round(float(np.corrcoef(dur[:148], heavy[:148])[0, 1]), 3))
nc_b = np.array([nc[blk == b].max() if (blk == b).any() else 0 for b in range(nb)])
print("  correlation(block duration, max contacts in block):", round(float(np.corrcoef(dur[:148], nc_b[:148])[0, 1]), 3))
np_b = np.array([npinv[blk == b].sum() if (blk == b).any() else 0 for b in range(nb)])
print("  correlation(block duration, pinv substeps in block):", round(float(np.corrcoef(dur[:148], np_b[:148])[0, 1]), 3))
