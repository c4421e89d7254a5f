"""_cis / _trans variants with a NON-ZERO base pose.  On the shipped model `base_link` sits at
the origin with identity orientation, so "minus base_position" subtracts zero and global_to_local is the identity: the offset /
rotation code is only exercised with a synthetic base pose injected through the model blob.  Checks: kernel logic vs oracle, and both
against an independent numpy restatement of env_mujoco_util_trans.py:848-862 (global_to_local = (Rx Ry Rz)^T (prop - base_position),
applied as written to positions AND Euler triples) and env_mujoco_util_cis.py:293-307 (positions minus base_position)."""
import numpy as np
import pytest

import hostsim
from hostsim import HostSim
from oracle.oracle import OracleEnv, load_blob

T = np.array([0.30, -0.20, 0.10])
EUL = np.array([0.20, -0.30, 0.50])


def _rot(a, b, c):
    Rx = np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])
    Ry = np.array([[np.cos(b), 0, np.sin(b)], [0, 1, 0], [-np.sin(b), 0, np.cos(b)]])
    Rz = np.array([[np.cos(c), -np.sin(c), 0], [np.sin(c), np.cos(c), 0], [0, 0, 1]])
    return Rx @ Ry @ Rz


def _blob(eul=EUL):
    off = hostsim.layout_offsets()
    b = load_blob().copy()
    b[off["BASE_LINK_POSITION"]:off["BASE_LINK_POSITION"] + 3] = T
    b[off["BASE_LINK_EULER"]:off["BASE_LINK_EULER"] + 3] = eul
    return b


def _check_frame_change(obs_base, obs_cis, obs_trans, q_trans, aux_trans, tol):
    R = _rot(*EUL)
    g2l = lambda v: (R.T @ (np.asarray(v, np.float64) - T))
    for i in range(len(obs_base)):
        # _cis: positions minus base_position; the arm pose does not depend on the variant (same Philox draws)
        assert np.allclose(obs_cis[i, 1:4], obs_base[i, 1:4] - T, atol=tol)
        assert np.allclose(obs_cis[i, 4:7], obs_base[i, 4:7], atol=tol)
        # _trans: every triple through global_to_local, Euler triples included (the "Euler minus translation" quirk)
        assert np.allclose(obs_trans[i, 1:4], g2l(obs_base[i, 1:4]), atol=tol)
        assert np.allclose(obs_trans[i, 4:7], g2l(obs_base[i, 4:7] * np.pi) / np.pi, atol=tol)
        assert np.allclose(obs_trans[i, 8:11], g2l(q_trans[i, 9:12]), atol=tol)
        assert np.allclose(obs_trans[i, 11:14], g2l([0, 0, 0]), atol=tol)
        assert np.allclose(obs_trans[i, 14:17], g2l(aux_trans[i, 14:17]), atol=tol)
        assert np.allclose(obs_trans[i, 23:26], g2l([0, np.pi / 2, 0]), atol=tol)
        # goals were sampled in the local frame and mapped out with local_to_global (trans:265-270)
        loc = g2l(aux_trans[i, 14:17])
        assert 0.35 - 1e-5 <= loc[0] <= 0.45 + 1e-5 and 0.25 - 1e-5 <= loc[1] <= 0.35 + 1e-5 and abs(loc[2] - 0.3468) < 1e-5
        loc = g2l(aux_trans[i, 11:14])
        assert -0.1 - 1e-5 <= loc[0] <= 0.1 + 1e-5 and 0.57 - 1e-5 <= loc[1] <= 0.67 + 1e-5 and abs(loc[2] - 0.1898) < 1e-5


def test_cis_trans_with_a_synthetic_base_pose_cpu():
    blob, n = _blob(), 6
    try:
        hostsim.set_blob(blob)
        runs = {}
        for variant in ("base", "cis", "trans"):
            hs = HostSim(n, task="picking", variant=variant, seed=3)
            obs = hs.reset()
            ora = [OracleEnv(task="picking", variant=variant, seed=3, env_id=i, blob=blob) for i in range(n)]
            oobs = np.stack([o.reset() for o in ora])
            assert np.abs(obs - oobs).max() < 5e-6, variant                      # kernel logic vs oracle at reset
            q, v, aux = hs.get_state()
            runs[variant] = (oobs.astype(np.float64), q, aux)
            a = np.random.default_rng(1).uniform(-1, 1, (n, 7)).astype(np.float32)
            o1 = hs.step_hybrid(a)[0]
            ro = np.stack([o.step(a[i].astype(np.float64))[0] for i, o in enumerate(ora)])
            assert np.abs(o1 - ro).max() < 2e-4, variant                          # ... and after one env step
        _check_frame_change(runs["base"][0], runs["cis"][0], runs["trans"][0], runs["trans"][1], runs["trans"][2], 2e-6)
        assert np.abs(runs["trans"][0][:, 1:4] - runs["cis"][0][:, 1:4]).max() > 0.05     # the rotation is not a no-op here
    finally:
        hostsim.set_blob(None)


@pytest.mark.gpu
def test_cis_trans_with_a_synthetic_base_pose_gpu():
    from mujoco_jaco_b200.env import JacoVecEnv
    blob, n = _blob(), 64
    runs = {}
    for variant in ("base", "cis", "trans"):
        env = JacoVecEnv(n, task="picking", variant=variant, seed=3, auto_reset=False, model_blob=blob)
        obs = env.reset().cpu().numpy().astype(np.float64)
        q, v, aux = [t.cpu().numpy() for t in env.get_state()]
        runs[variant] = (obs, q, aux)
        oobs = np.stack([OracleEnv(task="picking", variant=variant, seed=3, env_id=i, blob=blob).reset() for i in range(8)])
        assert np.abs(obs[:8] - oobs).max() < 5e-6, variant
        env.close()
    _check_frame_change(runs["base"][0], runs["cis"][0], runs["trans"][0], runs["trans"][1], runs["trans"][2], 5e-6)


def test_cis_touch_never_reports_the_grasp_class():
    """env_mujoco_util_cis.py:568-578: `if` instead of `elif` overwrites class 3 by 1.  In-contact snapshots whose base-variant step ends
    in touch class 3 must end in class 1 under the cis variant (same physics up to the a/100 action scale: zero action used), in the
    kernel logic and in the oracle, with the matching reward difference (+0.75 vs +2.5, times 0.01)."""
    import os
    Z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "contact_snapshots.npz"))
    idx = np.nonzero((Z["o64_step_obs"][:, 0] == 3) & ~Z["ke_step_d"])[0][:12]
    assert len(idx) == 12
    n_seen = 0
    for i in idx:
        res = {}
        for variant in ("base", "cis"):
            o = OracleEnv(task="picking", variant=variant, seed=0, env_id=int(i)); o.reset()
            o.set_state(Z["qpos"][i].astype(np.float64), Z["qvel"][i].astype(np.float64), Z["aux"][i].astype(np.float64))
            ob, r, d, s, wb = o.step(np.zeros(7))
            hs = HostSim(1, task="picking", variant=variant, seed=0, env_offset=int(i)); hs.reset()
            hs.set_state(Z["qpos"][i:i + 1], Z["qvel"][i:i + 1], Z["aux"][i:i + 1])
            kobs, krew, kdone, ksucc, kwb, info = hs.step_hybrid(np.zeros((1, 7), np.float32))
            assert kobs[0, 0] == ob[0] and bool(kdone[0]) == d and abs(krew[0] - r) < 1e-4 * max(1, abs(r))
            res[variant] = (ob[0], r, d)
        if res["base"][0] == 3:
            n_seen += 1
            assert res["cis"][0] == 1
            if not res["base"][2] and not res["cis"][2]:
                assert abs((res["base"][1] - res["cis"][1]) - 0.01 * (2.5 - 0.75)) < 2e-3
    assert n_seen >= 4
