"""On-disk formats: expert-trajectory npz and success-rate CSV of mujoco_jaco_b200/formats.py against the
reference's own files (tests/golden/reference_trajectory_sample.npz = first three episodes of grasping_trajectory_expert5.npz)."""
import os

import numpy as np
import pytest

from mujoco_jaco_b200.formats import TRAJ_KEYS, SuccessRateLog, TrajectoryRecorder, load_trajectories

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_reader_accepts_the_references_trajectory_file():
    d = load_trajectories(os.path.join(GOLD, "reference_trajectory_sample.npz"))
    assert set(TRAJ_KEYS) <= set(d) and len(d["episode_slices"]) == 3
    assert d["obs"].dtype == np.float32 and d["actions"].dtype == np.float64 and d["episode_starts"].dtype == bool
    for sl, ret in zip(d["episode_slices"], d["episode_returns"]):
        assert abs(d["rewards"][sl].sum() - ret) < 1e-9                      # episode_returns = sum of the episode's rewards
        assert np.array_equal(d["next_obs"][sl][:-1], d["obs"][sl][1:])       # next_obs chains within an episode


def test_recorder_writes_the_same_layout(tmp_path):
    ref = np.load(os.path.join(GOLD, "reference_trajectory_sample.npz"))
    rng = np.random.default_rng(0)
    n, rec = 3, TrajectoryRecorder(3)
    obs = rng.normal(size=(n, 26)).astype(np.float32)
    lens = [4, 2, 5]
    for t in range(6):
        act, rew = rng.uniform(-1, 1, (n, 7)).astype(np.float32), rng.normal(size=n).astype(np.float32)
        nxt = rng.normal(size=(n, 26)).astype(np.float32)
        done = np.array([(t + 1) % L == 0 for L in lens])
        term = nxt + 100.0                                                    # auto-reset: next_obs of a terminal step = term_obs
        rec.add(obs, act, rew, nxt, done, term_obs=term)
        obs = nxt
    path = str(tmp_path / "traj.npz")
    n_ep, n_tr = rec.save(path)
    assert n_ep == 1 + 3 + 1 and n_tr == 4 + 6 + 5
    z = np.load(path)
    assert {k: z[k].dtype for k in TRAJ_KEYS} == {k: ref[k].dtype for k in TRAJ_KEYS} and set(z.files) == set(ref.files)
    d = load_trajectories(path)
    for sl, ret in zip(d["episode_slices"], d["episode_returns"]):
        assert abs(d["rewards"][sl].sum() - ret) < 1e-9
        assert np.array_equal(d["next_obs"][sl][:-1], d["obs"][sl][1:])
        assert (d["next_obs"][sl][-1] > 50).any()                             # the terminal observation, not the reset one


def test_success_rate_csv_follows_the_25_deep_fifo(tmp_path):
    log = SuccessRateLog("picking", logger_path=str(tmp_path))
    assert os.path.basename(log.path) == "SR_picking_abalation_GA_learned_2.csv"      # env_mujoco.py:25-26
    log.update(np.array([1, 0, 1, 0]), np.array([1, 0, 0, 0]))                          # two episodes end: success, failure
    log.update(np.array([0, 0, 0, 0]), np.array([0, 0, 0, 0]))
    log.update(np.array([0, 1, 0, 0]), np.array([0, 1, 0, 0]))
    log.close()
    assert open(log.path).read().split() == ["4.000", "4.000", "8.000"]                  # "{0:3.3f}".format(sum / 25 * 100)


@pytest.mark.gpu
def test_record_a_device_rollout(tmp_path):
    import torch
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 256
    env = JacoVecEnv(n, task="picking", seed=3, auto_reset=True)
    obs = env.reset().clone()
    rec, log = TrajectoryRecorder(n), SuccessRateLog("picking", logger_path=str(tmp_path))
    term = torch.zeros(n, 26, device="cuda")
    g = torch.Generator(device="cuda"); g.manual_seed(0)
    for _ in range(40):
        a = torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
        nobs, rew, done, succ = env.step(a, term_obs=term)
        rec.add(obs, a, rew, nobs, done, term_obs=term)
        log.update(done, succ)
        obs = nobs.clone()
    if rec.episodes:
        n_ep, n_tr = rec.save(str(tmp_path / "t.npz"))
        d = load_trajectories(str(tmp_path / "t.npz"))
        assert len(d["episode_slices"]) == n_ep and d["obs"].shape == (n_tr, 26)
