"""CPU harness: workload mix of the hybrid step on the picking roll-out distribution (which fraction of env-substeps
need the stage-A proximity test, the stage-B general contact pass, a cooperative Newton solve ...).
Usage: python scripts/workload_mix.py [n_envs] [steps]"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from hostsim import HostSim

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
sim = HostSim(n, task="picking", seed=0)
sim.reset()
rng = np.random.default_rng(0)
out = np.zeros(16, np.int64)
names = ["substeps", "need", "stageB", "solved", "iters", "ncon", "solved_chain", "solved_withcon", "limit", "armprox", "objdest", "armcollide", "narrow", "armcontacts"]
for k in range(steps):
    a = rng.uniform(-1, 1, (n, 7)).astype(np.float32)
    o, r, d, s, wb, why = sim.step_fast(a, auto_reset=True, allow_fallback=2)
    sim.L.hh_wstats(out.ctypes.data_as(C.POINTER(C.c_longlong)), 1)
    tot = max(out[0], 1)
    print(k, "done", int(d.sum()), "bail", int((why != 0).sum()), " ".join(f"{nm}={out[i] / tot:.4f}" for i, nm in enumerate(names) if i not in (0, 4, 5)),
          f"iters/solve={out[4] / max(out[3], 1):.2f} ncon/solve={out[5] / max(out[3], 1):.2f}", flush=True)
pa, pc = np.zeros(64, np.int64), np.zeros(64, np.int64)
sim.L.hh_pairstats(pa.ctypes.data_as(C.POINTER(C.c_longlong)), pc.ctypes.data_as(C.POINTER(C.c_longlong)))
tot = n * steps * 50
print("pair: stageA-survive fraction / contact fraction (of all env-substeps)")
for p in range(60):
    if pa[p] or pc[p]:
        print(p, f"{pa[p] / tot:.4f} {pc[p] / tot:.4f}")
