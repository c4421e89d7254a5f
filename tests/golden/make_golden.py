"""Generates the fixtures in tests/golden/ (run HERE, where /root/reference exists; the fixtures travel).

1. reference_freefall.npz -- extracted from the reference's own real-MuJoCo recording
   /root/reference/models_baseline/trajectories/reaching.npz: the object box is teleported to z = 0.1898 with
   zero velocity and free-falls onto the floor plane; obs[:, 10] samples its z once per env step (50 substeps
   of 1 ms).  This is the one true MuJoCo known-answer vector in the reference tree: it pins free-body
   integration, the one-substep observation lag, plane-box contact generation, the pyramidal
   soft-contact impedance model and the Newton solver of the oracle against real MuJoCo output.
2. reference_rest_heights.npz -- object rest heights from the grasping recordings (weak invariants).
3. oracle_picking_rollout.npz -- oracle-generated regression vectors (seeded actions -> obs/reward/done) so a
   change of the oracle itself is detected; NOT evidence of agreement with MuJoCo.
4. reference_trajectory_sample.npz -- the first three episodes of the reference's grasping_trajectory_expert5.npz, verbatim keys and
   dtypes: the on-disk format mujoco_jaco_b200/formats.py reads and writes.
(policy_picking.npz: tests/golden/make_policy_golden.py; contact_snapshots.npz: tests/golden/make_contact_snapshots.py.)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, REPO)


def make_reference_fixtures():
    z = np.load(os.path.join(REF, "models_baseline/trajectories/reaching.npz"), allow_pickle=True)
    obs = z["obs"]
    assert z["episode_starts"][0]
    np.savez(os.path.join(HERE, "reference_freefall.npz"), obj_xyz=obs[:12, 8:11].astype(np.float32),
             source="models_baseline/trajectories/reaching.npz obs[0:12, 8:11] (object position per env step)")
    rest = []
    for f in ("grasping_trajectory_expert5", "grasping_trajectory_expert6"):
        g = np.load(os.path.join(REF, f"models_baseline/trajectories/{f}.npz"), allow_pickle=True)
        o, st = g["obs"], np.where(g["episode_starts"])[0]
        rest.append(np.stack([o[i + 5, 10] for i in st if i + 5 < len(o)]))
    np.savez(os.path.join(HERE, "reference_rest_heights.npz"), holder_rest_z=np.concatenate(rest).astype(np.float32),
             floor_rest_z=obs[8:12, 10].astype(np.float32))


def make_trajectory_sample():
    g = np.load(os.path.join(REF, "models_baseline/trajectories/grasping_trajectory_expert5.npz"), allow_pickle=True)
    end = int(np.where(g["episode_starts"])[0][3])
    np.savez_compressed(os.path.join(HERE, "reference_trajectory_sample.npz"), **{k: (g[k][:3] if k == "episode_returns" else g[k][:end]) for k in g.files})


def make_oracle_rollout():
    from oracle.oracle import OracleEnv
    n, steps = 16, 4
    rng = np.random.default_rng(2024)
    acts = rng.uniform(-1, 1, (steps, n, 7)).astype(np.float32)
    envs = [OracleEnv(task="picking", seed=77, env_id=i) for i in range(n)]
    obs0 = np.stack([e.reset() for e in envs])
    O, R, D = [], [], []
    for s in range(steps):
        res = [e.step(acts[s, i].astype(np.float64)) for i, e in enumerate(envs)]
        O.append(np.stack([r[0] for r in res])); R.append([r[1] for r in res]); D.append([r[2] for r in res])
    q = np.stack([e.get_state()[0] for e in envs])
    np.savez(os.path.join(HERE, "oracle_picking_rollout.npz"), seed=77, actions=acts, obs0=obs0, obs=np.stack(O),
             reward=np.array(R), done=np.array(D), qpos_final=q)


if __name__ == "__main__":
    make_reference_fixtures()
    make_trajectory_sample()
    make_oracle_rollout()
    print("golden fixtures written to", HERE)
