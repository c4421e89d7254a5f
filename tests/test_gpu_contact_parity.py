"""GPU parity tests, part 2 (through the C ABI): in-contact one-step parity on the 4096 contact snapshots, BASELINE
config 2 at its stated size (4096 contact-free reaching envs), terminal-logic boundaries and the OSC's pinv branch on
the device.  Bars and the oracle-side knife-edge predicates: tests/parity_util.py (no percentage budgets)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from parity_util import (ATOL, H, REW_TOL, RTOL, assert_obs, block_scale, check_step, check_substep, close, live_answers_and_knife_edges, load_snapshots, obs_close,
                         oracle_answers)

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "contact_snapshots.npz")


def _t(x):
    return torch.tensor(np.asarray(x, np.float32))


def _np(*ts):
    return [t.cpu().numpy() for t in ts]


def test_contact_snapshots_one_substep():
    """one mj_step from 4096 injected in-contact states: ncon and the touch-sensor flags bit-exact, state and
    qacc within the bars, everywhere outside the oracle-side knife-edge sets (counts printed)."""
    from mujoco_jaco_b200.env import JacoVecEnv
    Z, o64, ke = load_snapshots(GOLD)
    n = len(Z["qpos"])
    env = JacoVecEnv(n, task="picking", seed=0, auto_reset=False)
    env.reset()
    env.set_state(_t(Z["qpos"]), _t(Z["qvel"]), _t(Z["aux"]))
    ncon, sd, qacc = _np(*env.debug_substep(_t(Z["ctrl"])))
    q1, v1, _ = _np(*env.get_state())
    check_substep("CUDA general path", dict(qpos=q1, qvel=v1, ncon=ncon, sensordata=sd, qacc=qacc), o64, ke["sub_d"], ke["sub_c"])


def test_contact_snapshots_one_env_step():
    """One JacoMujocoEnv.step (the production hybrid kernel) from the same snapshots: done / succ / touch class bit-exact, reward
    1e-5, observation, outside the knife-edge sets; the finger-object contact counter of jaco_stats must have seen them."""
    from mujoco_jaco_b200.env import JacoVecEnv
    Z, o64, ke = load_snapshots(GOLD)
    n = len(Z["qpos"])
    env = JacoVecEnv(n, task="picking", seed=0, auto_reset=False)
    env.reset()
    env.set_state(_t(Z["qpos"]), _t(Z["qvel"]), _t(Z["aux"]))
    env.stats(clear=True)
    obs, rew, done, succ = _np(*env.step(_t(Z["action"]).cuda()))
    check_step("CUDA hybrid kernel", dict(obs=obs, reward=rew, done=done, succ=succ), o64, ke["step_d"], ke["step_c"])
    st = env.stats(clear=True).cpu().numpy()
    assert st[13] > 0.3 * n * 50, st          # STAT_FINGEROBJ: env-substeps with a hand-object contact
    assert st[1] == succ.sum() and st[0] == done.sum()
    print(f"cap counters on the contact snapshots: contact cap {st[9]:.0f}, iteration cap {st[10]:.0f} of {st[11]:.0f} stage-B solves")


@pytest.mark.parametrize("v0_sigma", [0.0, 0.3])
def test_config2_reaching_4096_one_substep_and_one_env_step(v0_sigma):
    """BASELINE config 2: 4096 contact-free reaching envs, identical (qpos, qvel, aux, ctrl / action) injected into oracle and
    kernel; one substep and one env step; v0 = 0 and v0 ~ N(0, 0.3^2) rad/s.  The strict bars (1e-4 rel / 1e-6 abs state, 1e-5
    reward, done bit-exact) hold on every env the float32 oracle itself keeps inside them."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 4096
    env = JacoVecEnv(n, task="reaching", seed=11, auto_reset=False)
    env.reset()
    # one zero-action env step first: at the reset pose the destination box touches the floor with dist == 0 exactly, a knife
    # edge on which float32 (contact) and float64 (no contact) legitimately differ on every env
    env.step(torch.zeros(n, 7, device="cuda"))
    q, v, aux = [x.astype(np.float64) for x in _np(*env.get_state())]
    rng = np.random.default_rng(5)
    if v0_sigma > 0:
        v[:, :6] = rng.normal(0, v0_sigma, (n, 6))
    q, v, aux = q.astype(np.float32), v.astype(np.float32), aux[:, :64].astype(np.float32)
    ctrl = np.concatenate([rng.uniform(-5, 5, (n, 6)), np.zeros((n, 3))], 1).astype(np.float32)
    act = np.concatenate([rng.uniform(-1, 1, (n, 6)), np.zeros((n, 1))], 1).astype(np.float32)
    o64, (kd, kc), (kd2, kc2) = live_answers_and_knife_edges("reaching", q, v, aux, act=act, ctrl=ctrl, seed=11)
    env.set_state(_t(q), _t(v), _t(aux))
    ncon, sd, qacc = _np(*env.debug_substep(_t(ctrl)))
    q1, v1, _ = _np(*env.get_state())
    print(f"[config 2, v0 sigma {v0_sigma}] one substep: {int(kd.sum())} discrete / {int(kc.sum())} strict exclusions of {n}")
    assert kc.mean() < 0.08
    ok = ~kc
    assert np.array_equal(ncon[~kd], o64["sub_ncon"][~kd])
    sc = block_scale(o64["sub_qacc"])
    assert close(q1, o64["sub_qpos"], RTOL, ATOL)[ok].all()
    assert close(v1, o64["sub_qvel"], RTOL, ATOL, H * sc)[ok].all()
    assert close(qacc, o64["sub_qacc"], 1e-4, 1e-3, sc)[ok].all()
    env.set_state(_t(q), _t(v), _t(aux))
    obs, rew, done, succ = _np(*env.step(_t(act).cuda()))
    print(f"[config 2, v0 sigma {v0_sigma}] one env step: {int(kd2.sum())} discrete / {int(kc2.sum())} strict exclusions of {n}")
    assert kc2.mean() < 0.08
    assert np.array_equal(done[~kd2].astype(bool), o64["step_done"][~kd2].astype(bool)) and np.array_equal(succ[~kd2], o64["step_succ"][~kd2])
    ok = ~kc2
    assert (np.abs(rew - o64["step_reward"]) <= REW_TOL * np.maximum(1, np.abs(o64["step_reward"])))[ok].all()
    assert_obs(f"config 2, v0 sigma {v0_sigma}", obs, o64["step_obs"], ok)


def test_terminal_boundaries_on_device():
    """SURVEY 8d C4 boundary states through the production kernel, against the oracle on the same states: drop (obj_z < 0.1),
    time-out at current_steps 699 -> 700, singular exit q2 in pi +- 0.1 (inside / outside the window), picked flag."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 8
    env = JacoVecEnv(n, task="picking", seed=2, auto_reset=False)
    env.reset()
    q, v, aux = _np(*env.get_state())
    q[0, 11] = 0.05; q[0, 9:11] = [0.5, -0.5]                      # drop
    aux[1, 9] = 699                                                 # time-out on this step
    aux[2, 9] = 698                                                 # one short of it
    for i, off in ((3, 0.0), (4, 0.08), (5, 0.12), (6, -0.08), (7, -0.12)):   # singular window pi +- 0.1
        q[i, 2] = np.pi + off; aux[i, 46 + 2] = np.pi + off
    aux = aux[:, :64]
    act = np.zeros((n, 7), np.float32)
    o64 = oracle_answers("picking", q, v, aux, act=act, seed=2, workers=1)
    env.set_state(_t(q), _t(v), _t(aux))
    obs, rew, done, succ = _np(*env.step(_t(act).cuda()))
    assert np.array_equal(done.astype(bool), o64["step_done"].astype(bool)), (done, o64["step_done"])
    assert np.array_equal(succ, o64["step_succ"])
    assert done[0] == 1 and rew[0] < -19 and done[1] == 1 and -10.2 < rew[1] < -9.8 and done[2] == 0
    assert done[3] == 1 and done[4] == 1 and done[5] == 0 and done[6] == 1 and done[7] == 0
    assert (np.abs(rew - o64["step_reward"]) <= REW_TOL * np.maximum(1, np.abs(o64["step_reward"]))).all()


def test_osc_pinv_branch_on_device():
    """Stretched-arm postures (links 2-3 aligned: |det(J M^-1 J^T)| < 1e-3 -> numpy.linalg.pinv(rcond=1e-4), App. A.2) through one
    env step of the production kernel: jaco_stats must count pinv substeps and the result must agree with the oracle wherever
    the float32 oracle does."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 512
    env = JacoVecEnv(n, task="reaching", seed=4, auto_reset=False)
    env.reset()
    q, v, aux = _np(*env.get_state())
    rng = np.random.default_rng(11)
    q[:, 1] = np.pi + rng.normal(0, 0.02, n); q[:, 2] = np.pi + rng.normal(0, 0.02, n)
    q[:, 0] = rng.uniform(0, 6.28, n); q[:, 3:6] = rng.uniform(0, 6.28, (n, 3))
    v[:, :6] = rng.normal(0, 0.3, (n, 6))
    aux[:, 46:52] = q[:, :6]; aux[:, 55:61] = v[:, :6]             # lagged controller state = injected state
    aux = aux[:, :64]
    act = np.concatenate([rng.uniform(-1, 1, (n, 6)), np.zeros((n, 1))], 1).astype(np.float32)
    o64, _, (kd, kc) = live_answers_and_knife_edges("reaching", q, v, aux, act=act, seed=4)
    env.set_state(_t(q), _t(v), _t(aux))
    env.stats(clear=True)
    obs, rew, done, succ = _np(*env.step(_t(act).cuda()))
    st = env.stats(clear=True).cpu().numpy()
    print(f"[pinv] {st[12]:.0f} pinv substeps of {n * 50}; {int(kd.sum())} discrete / {int(kc.sum())} strict exclusions")
    assert st[12] > 0.2 * n * 50, st[12]
    assert kc.mean() < 0.25
    assert np.array_equal(done[~kd].astype(bool), o64["step_done"][~kd].astype(bool))
    ok = ~kc
    assert (np.abs(rew - o64["step_reward"]) <= REW_TOL * np.maximum(1, np.abs(o64["step_reward"])))[ok].all()
    assert obs_close(obs, o64["step_obs"])[ok].all()
