"""On-disk formats either side of the env-step path.

* Expert-trajectory `.npz` in the layout of the reference's `models_baseline/trajectories/*.npz` (what the stable-baselines
  behaviour-cloning pre-trainer reads): `actions [T, A] f64`, `obs [T, O] f32`, `rewards [T] f64`, `next_obs [T, O] f32`,
  `episode_returns [E] f64`, `episode_starts [T] bool`; transitions are episode-major (each episode contiguous, its first
  transition flagged in `episode_starts`).  `TrajectoryRecorder` fills it from batched device roll-outs of `JacoVecEnv`:
  every env contributes its completed episodes, unfinished tails are dropped like the reference's generator drops them.
* Success-rate CSV of `env_mujoco.py:25-28,129-135`: one line per finished episode, the success percentage over a 25-deep
  FIFO of episode outcomes formatted `"{0:3.3f}"`.  `SuccessRateLog` applies that rule to the episodes a batched env
  finishes (in env-index order within a step), so a vectorised run writes the file the single-env reference would.
"""
from __future__ import annotations

import os

import numpy as np

TRAJ_KEYS = ("actions", "obs", "rewards", "next_obs", "episode_returns", "episode_starts")


class SuccessRateLog:
    """`logger_csv/SR_<task>_abalation_GA_learned_2.csv` writer for a batched env (env_mujoco.py:25-28,129-135)."""

    def __init__(self, task, logger_path="./logger_csv", depth=25):
        os.makedirs(logger_path, exist_ok=True)
        self.path = os.path.join(logger_path, "SR_" + task + "_abalation_GA_learned_2.csv")
        self.file = open(self.path, "w")
        self.accum_succ = [0] * depth
        self.depth = depth

    def update(self, done, succ):
        """done, succ: per-env arrays of one step (numpy or torch); writes one line per finished episode."""
        done = np.asarray(done.cpu() if hasattr(done, "cpu") else done).astype(bool)
        succ = np.asarray(succ.cpu() if hasattr(succ, "cpu") else succ)
        for i in np.nonzero(done)[0]:
            self.accum_succ.pop()
            self.accum_succ.insert(0, int(succ[i]))
            self.file.writelines("{0:3.3f}".format(sum(self.accum_succ) / self.depth * 100) + "\n")
        if done.any():
            self.file.flush()

    def close(self):
        self.file.close()


class TrajectoryRecorder:
    """Collects (obs, action, reward, next_obs) of a batched roll-out and writes completed episodes as an expert-trajectory npz.
    `next_obs` of a terminal transition is the episode's terminal observation (pass `term_obs` when the env auto-resets)."""

    def __init__(self, n_envs):
        self.n = n_envs
        self.cur = [[] for _ in range(n_envs)]      # per env: transitions of its running episode
        self.episodes = []                           # completed episodes: list of lists of (obs, act, rew, next_obs)

    @staticmethod
    def _np(x):
        return np.asarray(x.detach().cpu() if hasattr(x, "detach") else x)

    def add(self, obs, action, reward, next_obs, done, term_obs=None):
        obs, action, reward, next_obs, done = (self._np(x) for x in (obs, action, reward, next_obs, done))
        term = None if term_obs is None else self._np(term_obs)
        for i in range(self.n):
            nxt = term[i] if (term is not None and done[i]) else next_obs[i]
            self.cur[i].append((obs[i].astype(np.float32), action[i].astype(np.float64), float(reward[i]), nxt.astype(np.float32)))
            if done[i]:
                self.episodes.append(self.cur[i])
                self.cur[i] = []

    def save(self, path, max_episodes=None):
        eps = self.episodes if max_episodes is None else self.episodes[:max_episodes]
        if not eps:
            raise ValueError("no completed episode to save")
        flat = [t for ep in eps for t in ep]
        starts = np.zeros(len(flat), bool)
        k = 0
        for ep in eps:
            starts[k] = True
            k += len(ep)
        np.savez(path, actions=np.stack([t[1] for t in flat]), obs=np.stack([t[0] for t in flat]),
                 rewards=np.array([t[2] for t in flat], np.float64), next_obs=np.stack([t[3] for t in flat]),
                 episode_returns=np.array([sum(t[2] for t in ep) for ep in eps], np.float64), episode_starts=starts)
        return len(eps), len(flat)


def load_trajectories(path):
    """Reads an expert-trajectory npz (the reference's or TrajectoryRecorder's) and checks its invariants.  Older reference
    files lack `next_obs`; it is then rebuilt from `obs` within each episode (the last transition repeats its own obs)."""
    z = np.load(path, allow_pickle=False)
    d = {k: z[k] for k in z.files}
    T = len(d["actions"])
    assert len(d["obs"]) == T and len(d["rewards"]) == T and len(d["episode_starts"]) == T
    starts = d["episode_starts"].astype(bool)
    assert starts[0] and starts.sum() == len(d["episode_returns"])
    if "next_obs" not in d:
        nxt = np.roll(d["obs"], -1, axis=0)
        last = np.r_[starts[1:], True]
        nxt[last] = d["obs"][last]
        d["next_obs"] = nxt
    bounds = np.r_[np.nonzero(starts)[0], T]
    d["episode_slices"] = [slice(int(a), int(b)) for a, b in zip(bounds[:-1], bounds[1:])]
    return d
