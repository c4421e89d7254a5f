"""CPU tests: pin the oracle (oracle/jaco_oracle.c) against the reference's own real-MuJoCo recordings, against an
independent numpy restatement of the smooth dynamics, and against its committed regression vectors."""
import os

import numpy as np
import pytest

from oracle.oracle import OracleEnv, load_blob, philox

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _drop_object(env_like_step, get_z, n_env_steps):
    zs = []
    for _ in range(n_env_steps):
        for k in range(50):
            if k == 49:
                zs.append(get_z())          # mjData.xpos of the last substep's forward = qpos before its integration
            env_like_step()
    return np.array(zs)


def test_freefall_and_floor_contact_match_real_mujoco_recording():
    """reference models_baseline/trajectories/reaching.npz: object teleported to z=0.1898, v=0, falls on the floor.
    The recorded per-env-step object height (real MuJoCo 2.0) is reproduced to float32 resolution, including the
    one-substep observation lag (49, 99, 149 integrations), the first impact and the settling sequence."""
    rec = np.load(os.path.join(GOLD, "reference_freefall.npz"))["obj_xyz"]
    e = OracleEnv(task="reaching", seed=0, env_id=0)
    e.reset()
    q, v, aux = e.get_state()
    q[9:12] = [rec[0, 0], -abs(rec[0, 1]), 0.1898]    # away from the holder (the recording's model had none)
    q[12:16] = [1, 0, 0, 0]
    v[9:15] = 0
    e.set_state(q, v, aux)
    ctrl = np.zeros(9); ctrl[6:] = 1.1
    zs = _drop_object(lambda: e.mj_step(ctrl), lambda: e.get_state()[0][11], 11)
    assert rec[0, 2] == np.float32(0.1898)
    np.testing.assert_allclose(zs[:3], rec[1:4, 2], rtol=0, atol=2e-8)      # pure free fall: sum_{k<50j} g h^2 k
    np.testing.assert_allclose(zs, rec[1:12, 2], rtol=0, atol=2e-8)         # impact 0.02579194, settle 0.02998548
    assert abs(zs[-1] - 0.02998548) < 1e-8


def test_rest_heights_weak_invariants():
    g = np.load(os.path.join(GOLD, "reference_rest_heights.npz"))
    assert np.all((g["floor_rest_z"] > 0.02998) & (g["floor_rest_z"] < 0.02999))
    e = OracleEnv(task="picking", seed=1, env_id=0)
    e.reset()
    for _ in range(6):
        obs, *_ = e.step(np.zeros(7))
    # object settles on the holder (top z = 0.17) with a penetration of a few 1e-5 m, as in the recordings
    assert 0.19995 < obs[10] < 0.2000
    assert abs(obs[10] - np.median(g["holder_rest_z"])) < 3e-5


def test_oracle_regression_vectors():
    g = np.load(os.path.join(GOLD, "oracle_picking_rollout.npz"))
    acts = g["actions"]
    n = acts.shape[1]
    envs = [OracleEnv(task="picking", seed=int(g["seed"]), env_id=i) for i in range(n)]
    obs0 = np.stack([e.reset() for e in envs])
    np.testing.assert_allclose(obs0, g["obs0"], atol=1e-7)
    for s in range(acts.shape[0]):
        res = [e.step(acts[s, i].astype(np.float64)) for i, e in enumerate(envs)]
        np.testing.assert_allclose(np.stack([r[0] for r in res]), g["obs"][s], atol=2e-6)
        np.testing.assert_allclose([r[1] for r in res], g["reward"][s], atol=1e-7)
        assert [r[2] for r in res] == list(g["done"][s])


# ---------------------------------------------------------------- independent numpy restatement (smooth dynamics)
def _blob_fields(blob):
    import re
    hdr = open(os.path.join(os.path.dirname(GOLD), "..", "include", "jaco_model_layout.h")).read()
    off = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define JM_(\w+) (\d+)", hdr)}
    f = lambda name, shape: blob[off[name]: off[name] + int(np.prod(shape))].reshape(shape)
    return dict(pos=f("CHAIN_POS", (9, 3)), quat=f("CHAIN_QUAT", (9, 4)), axis=f("CHAIN_AXIS", (9, 3)), ref=f("CHAIN_REF", (9,)),
                com=f("CHAIN_COM", (9, 3)), mass=f("CHAIN_MASS", (9,)), inertia=f("CHAIN_INERTIA", (9, 6)), ee=f("EE_OFFSET", (3,)))


def _q2m(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def _rodrigues(axis, ang):
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def _fk(m, q):
    parent = [-1, 0, 1, 2, 3, 4, 5, 5, 5]
    P, R = [], []
    for k in range(9):
        pp, pr = (np.zeros(3), np.eye(3)) if parent[k] < 0 else (P[parent[k]], R[parent[k]])
        P.append(pp + pr @ m["pos"][k])
        R.append(pr @ _q2m(m["quat"][k]) @ _rodrigues(m["axis"][k], q[k] - m["ref"][k]))
    return P, R


def _kinetic_matrix(m, q, eps=1e-6):
    """M = sum_b m Jc^T Jc + Jw^T (R I R^T) Jw with finite-difference Jacobians of an independent FK."""
    def feats(qq):
        P, R = _fk(m, qq)
        return [P[b] + R[b] @ m["com"][b] for b in range(9)], R
    M = np.zeros((9, 9))
    c0, R0 = feats(q)
    Jc = np.zeros((9, 3, 9)); Jw = np.zeros((9, 3, 9))
    for j in range(9):
        dq = np.zeros(9); dq[j] = eps
        cp, Rp = feats(q + dq); cm, Rm = feats(q - dq)
        for b in range(9):
            Jc[b, :, j] = (cp[b] - cm[b]) / (2 * eps)
            W = (Rp[b] - Rm[b]) / (2 * eps) @ R0[b].T
            Jw[b, :, j] = [W[2, 1], W[0, 2], W[1, 0]]
    for b in range(9):
        I6 = m["inertia"][b]
        I = np.array([[I6[0], I6[3], I6[4]], [I6[3], I6[1], I6[5]], [I6[4], I6[5], I6[2]]])
        M += m["mass"][b] * Jc[b].T @ Jc[b] + Jw[b].T @ (R0[b] @ I @ R0[b].T) @ Jw[b]
    return M, c0


def test_mass_matrix_gravity_and_ee_jacobian_against_independent_numpy():
    blob = load_blob()
    m = _blob_fields(blob)
    e = OracleEnv(task="reaching", seed=3, env_id=0)
    rng = np.random.default_rng(0)
    for _ in range(3):
        e.reset()
        q, v, aux = e.get_state()
        q[:9] = np.r_[rng.uniform(0.5, 5.0, 6), rng.uniform(0.2, 1.4, 3)]
        v[:] = 0
        e.set_state(q, v, aux)
        f = e.forward(np.zeros(9))
        M, _ = _kinetic_matrix(m, q[:9])
        np.testing.assert_allclose(f["M"][:9, :9], M, atol=2e-6)
        # bias at zero velocity = dV/dq with V = sum m g z_com
        V = lambda qq: sum(m["mass"][b] * 9.81 * (_fk(m, qq)[0][b] + _fk(m, qq)[1][b] @ m["com"][b])[2] for b in range(9))
        g = np.array([(V(q[:9] + 1e-6 * np.eye(9)[j]) - V(q[:9] - 1e-6 * np.eye(9)[j])) / 2e-6 for j in range(9)])
        np.testing.assert_allclose(f["bias"][:9], g, atol=2e-6)
        # EE position and translational Jacobian (mujoco_config.py:243-301, body 'EE' = link6 origin + 0.12 z)
        ee = lambda qq: _fk(m, qq)[0][5] + _fk(m, qq)[1][5] @ m["ee"]
        np.testing.assert_allclose(f["ee_pos"], ee(q[:9]), atol=1e-9)
        Jp = np.stack([(ee(q[:9] + 1e-6 * np.eye(9)[j]) - ee(q[:9] - 1e-6 * np.eye(9)[j])) / 2e-6 for j in range(6)], 1)
        np.testing.assert_allclose(f["J"][:3], Jp, atol=1e-7)


def test_coriolis_via_energy_conservation():
    """Contact-free passive motion of the arm: d/dt (kinetic + potential) = -damping power; checks the RNE bias."""
    blob = load_blob(); m = _blob_fields(blob)
    e = OracleEnv(task="reaching", seed=4, env_id=0)
    e.reset()
    q, v, aux = e.get_state()
    q[:9] = [1.2, 3.9, 1.3, 2.0, 1.5, 1.2, 0.9, 0.9, 0.9]
    v[:9] = [0.5, -0.4, 0.6, 0.8, -0.7, 0.9, 0, 0, 0]
    e.set_state(q, v, aux)

    def energy():
        qq, vv, _ = e.get_state()
        M, c = _kinetic_matrix(m, qq[:9])
        return 0.5 * vv[:9] @ M @ vv[:9] + sum(m["mass"][b] * 9.81 * c[b][2] for b in range(9))
    E0 = energy()
    ctrl = np.zeros(9); ctrl[6:] = 0.9           # finger servos hold their pose
    for _ in range(20):
        e.mj_step(ctrl)
    E1 = energy()
    # semi-implicit Euler at h = 1 ms: energy drift over 20 ms is O(h) small relative to the energy scale
    assert abs(E1 - E0) < 2e-3 * max(1.0, abs(E0)), (E0, E1)


def test_euler_rxyz_convention_matches_scipy():
    from scipy.spatial.transform import Rotation
    e = OracleEnv(task="reaching", seed=5, env_id=0)
    e.reset()
    f = e.forward(np.zeros(9))
    qw = f["ee_quat"]
    Rm = Rotation.from_quat([qw[1], qw[2], qw[3], qw[0]])
    obs = e.reset()
    # obs[4:7] = euler_from_quaternion(xquat, 'rxyz') / pi  (util:223-271): intrinsic x-y'-z''
    eul = Rm.as_euler("XYZ")
    f2 = e.forward(np.zeros(9))
    q2 = f2["ee_quat"]
    eul = Rotation.from_quat([q2[1], q2[2], q2[3], q2[0]]).as_euler("XYZ")
    np.testing.assert_allclose(obs[4:7] * np.pi, eul, atol=2e-6)


def test_philox_known_answer():
    # Random123 known-answer test for philox4x32-10 with zero counter and key
    np.testing.assert_array_equal(philox(0, 0, 0, 0, 0), np.array([0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8], np.uint32))
    k = 0xffffffffffffffff
    np.testing.assert_array_equal(philox(k, 0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff),
                                  np.array([0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd], np.uint32))
