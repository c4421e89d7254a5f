"""Workload for compute-sanitizer (SURVEY section 5 plan; VERDICT r1 hygiene item): a 1024-env picking population that exercises every
kernel of the library -- reset, the production hybrid step kernel incl. the shared-memory stage-B task queue and the union'd
cooperative workspace (in-contact snapshots injected so that most envs need a cooperative solve), the list kernel (auto-reset),
jaco_debug_substep, get/set_state and the allocation-free entry point.

  compute-sanitizer --tool memcheck|racecheck|initcheck|synccheck python scripts/sanitize_step.py [n_envs] [steps]
"""
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from mujoco_jaco_b200.env import JacoVecEnv  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
torch.manual_seed(0)
env = JacoVecEnv(n, task="picking", seed=7, device=0, auto_reset=True)
env.reset()
for _ in range(steps):                                  # roll-out mix: fast path, pinv, in-place reset
    env.step(torch.rand(n, 7, device="cuda") * 2 - 1)
Z = np.load(os.path.join(REPO, "tests", "golden", "contact_snapshots.npz"))
f = lambda k: torch.tensor(np.resize(np.asarray(Z[k], np.float32), (n,) + Z[k].shape[1:]))
env.set_state(f("qpos"), f("qvel"), f("aux"))
for _ in range(steps):                                  # contact-rich mix: stage-B queue, cooperative solves, touch sensors
    env.step(f("action").cuda())
env.debug_substep(torch.zeros(n, 9, device="cuda"))
qpos, qvel, aux = env.get_state()
torch.cuda.synchronize()
st = env.stats().cpu().numpy()
print("sanitize workload done:", n, "envs,", 2 * steps, "env steps; episodes", st[0], "nan_resets", st[6], "launches", env.launch_count())
env.close()
