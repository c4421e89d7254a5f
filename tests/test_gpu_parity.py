"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on identical inputs.

Tolerances (BASELINE.json north_star): one-step state 1e-4 relative / 1e-6 absolute in fp32, reward 1e-5,
contact count / success / done bit-exact.  No percentage budgets: an env is excused from a check only by the explicit
oracle-side knife-edge predicate of tests/parity_util.py (the float32 build of the ORACLE flips the outcome / leaves the
bar on that input); the excluded counts are printed and bounded.  In-contact snapshots, config 2 at 4096 envs, terminal
boundaries and the pinv branch: test_gpu_contact_parity.py.
"""
import numpy as np
import pytest

from parity_util import ATOL, H, RTOL, block_scale, check_step, check_substep, close, live_answers_and_knife_edges

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _oracles(task, seed, n, **kw):
    from oracle.oracle import OracleEnv
    return [OracleEnv(task=task, seed=seed, env_id=i, **kw) for i in range(n)]


def _close(a, b, rtol, atol):
    return np.abs(a - b) <= atol + rtol * np.abs(b)


@pytest.mark.parametrize("task", ["picking", "reaching", "pickAndplace", "grasping"])
def test_reset_matches_oracle(task):
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 64 if task == "grasping" else 256
    env = JacoVecEnv(n, task=task, seed=5, auto_reset=False)
    obs = env.reset().cpu().numpy()
    q, v, aux = [t.cpu().numpy() for t in env.get_state()]
    ora = _oracles(task, 5, n)
    oobs = np.stack([o.reset() for o in ora])
    # grasping: two approach servo loops of ~1000 substeps precede the first observation (util:123-167)
    tol = 5e-5 if task == "grasping" else 5e-6
    assert np.abs(obs - oobs).max() < tol
    for i in (0, 17, n - 1):
        oq, ov, oa = ora[i].get_state()
        assert np.abs(q[i] - oq).max() < (5e-5 if task == "grasping" else 1e-6) and np.abs(v[i] - ov).max() < (1e-3 if task == "grasping" else 1e-6)
        assert np.abs(aux[i, :23] - oa[:23]).max() < (5e-5 if task == "grasping" else 1e-5)
        if task != "grasping":
            assert np.abs(aux[i, :64] - oa).max() < 1e-5
    if task == "grasping":
        assert (np.linalg.norm(obs[:, 1:4] - aux[:, 11:14], axis=1) < 0.15 + 1e-3).all()


def test_one_substep_state_and_contacts_picking():
    """mj_step from identical (qpos, qvel, ctrl): state within 1e-4 rel / 1e-6 abs, ncon / touch flags bit-exact, qacc."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 512
    env = JacoVecEnv(n, task="picking", seed=9, auto_reset=False)
    env.reset()
    # one zero-action env step first: at the reset pose the destination box touches the floor with dist == 0 exactly, a
    # knife edge on which fp32 (contact) and the fp64 oracle (no contact) legitimately differ
    env.step(torch.zeros(n, 7, device="cuda"))
    q, v, aux = [t_.cpu().numpy() for t_ in env.get_state()]
    rng = np.random.default_rng(1)
    v[:, :9] = rng.normal(0, 0.3, (n, 9)).astype(np.float32)
    aux = aux[:, :64]
    ctrl = np.concatenate([rng.uniform(-5, 5, (n, 6)), rng.uniform(0.6, 1.0, (n, 1)).repeat(3, 1)], 1).astype(np.float32)
    o64, (kd, kc), _ = live_answers_and_knife_edges("picking", q, v, aux, ctrl=ctrl, seed=9)
    env.set_state(torch.tensor(q), torch.tensor(v), torch.tensor(aux))
    ncon, sd, qacc = [t_.cpu().numpy() for t_ in env.debug_substep(torch.tensor(ctrl))]
    q1, v1, _ = [t_.cpu().numpy() for t_ in env.get_state()]
    # the object rests on the holder and the destination box on the floor: in-contact bars of check_substep (ncon / touch
    # flags exact, qpos strict, second-tier bound on the velocities of the bodies in contact)
    check_substep("picking after reset + 1 step, random joint velocities", dict(qpos=q1, qvel=v1, ncon=ncon, sensordata=sd, qacc=qacc), o64, kd, kc)
    arm = ~kd    # the arm itself is contact-free here: its 9 dofs meet the strict bar relative to the block scale of qacc
    sc = block_scale(o64["sub_qacc"])
    assert close(v1[:, :9], o64["sub_qvel"][:, :9], RTOL, ATOL, (H * sc)[:, :9])[arm].all()


def test_env_step_reward_done_picking():
    """Three env steps of a random roll-out; before each step the kernel's state is injected into the oracle, so every step is a
    one-step comparison from identical inputs: done / succ / touch bit-exact, reward 1e-5, observation."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 256
    env = JacoVecEnv(n, task="picking", seed=21, auto_reset=False)
    env.reset()
    env.step(torch.zeros(n, 7, device="cuda"))      # leave the dist == 0 knife edge of the reset pose (see above)
    rng = np.random.default_rng(3)
    for step in range(3):
        q, v, aux = [t_.cpu().numpy() for t_ in env.get_state()]
        a = rng.uniform(-1, 1, (n, 7)).astype(np.float32)
        o64, _, (kd, kc) = live_answers_and_knife_edges("picking", q, v, aux[:, :64], act=a, seed=21)
        obs, rew, done, succ = [t_.cpu().numpy() for t_ in env.step(torch.from_numpy(a).cuda())]
        assert kc.mean() < 0.05
        check_step(f"roll-out step {step}", dict(obs=obs, reward=rew, done=done, succ=succ), o64, kd, kc)


def test_rollout_50_steps_reaching_drift():
    """Short-horizon drift bound: 50 env steps (2500 substeps) of contact-free reaching.  The bound is the measured
    divergence of the oracle's own fp32 build from its fp64 build on the same roll-out: the CUDA
    path (fp32) must stay within 3x that divergence (+1e-3 rad), and its median env within 5e-4 rad."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 64
    env = JacoVecEnv(n, task="reaching", seed=2, auto_reset=False)
    env.reset()
    ora = _oracles("reaching", 2, n)
    o32 = _oracles("reaching", 2, n, f32=True)
    for o in ora + o32:
        o.reset()
    rng = np.random.default_rng(5)
    alive = np.ones(n, bool)
    for step in range(50):
        a = rng.uniform(-1, 1, (n, 6)).astype(np.float32)
        obs, rew, done, succ = [t.cpu().numpy() for t in env.step(torch.from_numpy(a).cuda())]
        for i, o in enumerate(ora):
            if alive[i]:
                r = o.step(a[i].astype(np.float64))
                r32 = o32[i].step(a[i].astype(np.float64))
                alive[i] &= not (r[2] or r32[2])
        alive &= ~done.astype(bool)
    q, _, _ = [t.cpu().numpy() for t in env.get_state()]
    oq = np.stack([o.get_state()[0] for o in ora])
    oq32 = np.stack([o.get_state()[0] for o in o32])
    err = np.abs(q[alive, :6] - oq[alive, :6]).max(1)
    err32 = np.abs(oq32[alive, :6] - oq[alive, :6]).max(1)
    assert alive.sum() > n // 2
    assert np.median(err) < 5e-4, np.median(err)
    assert err.max() < 3 * err32.max() + 1e-3, (err.max(), err32.max())


def test_step_host_matches_device_path():
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 128
    e1 = JacoVecEnv(n, task="picking", seed=4, auto_reset=True)
    e2 = JacoVecEnv(n, task="picking", seed=4, auto_reset=True)
    e1.reset(); e2.reset()
    a = np.random.default_rng(0).uniform(-1, 1, (n, 7)).astype(np.float32)
    o1, r1, d1, s1 = [t.cpu().numpy() for t in e1.step(torch.from_numpy(a).cuda())]
    o2, r2, d2, s2, w2 = e2.step_host(a)
    assert np.array_equal(o1, o2) and np.array_equal(r1, r2) and np.array_equal(d1, d2)


def test_single_env_shim_runs_1000_random_steps(tmp_path):
    """BASELINE config 1: the JacoMujocoEnv surface, picking, random actions."""
    from mujoco_jaco_b200.env import JacoMujocoEnv
    env = JacoMujocoEnv(task="picking", robot_file="jaco2_curtain_torque", seed=2346, rulebased_subgoal=True,
                        logger_path=str(tmp_path), verbose=False)
    obs = env.reset()
    assert obs.shape == (26,) and obs.dtype == np.float32
    rng = np.random.default_rng(0)
    ndone = 0
    for _ in range(1000):
        obs, r, d, info = env.step(rng.uniform(-1, 1, 7))
        assert obs.shape == (26,) and np.isfinite(obs).all() and np.isfinite(r) and info == {0: 0}
        if d:
            ndone += 1
            obs = env.reset()
    env.close()


@pytest.mark.parametrize("task,variant", [("placing", "base"), ("pickAndplace", "cis"), ("picking", "trans"), ("placing", "cis")])
def test_config4_tasks_and_variants_one_step(task, variant):
    """BASELINE config 4: placing / pickAndplace and the _cis / _trans variants (base-relative observations, a/100 - a/20
    action scale, 3-tuple terminal logic): reset (placing = 150 in-hand settle substeps) and one env step against the oracle.
    Placing starts with the object held between the fingers: its 150-substep settle is compared on the envs where the oracle's
    own float32 build reproduces its float64 reset; the env step is a one-step comparison from the kernel's injected state."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 64
    env = JacoVecEnv(n, task=task, variant=variant, seed=13, auto_reset=False)
    obs = env.reset().cpu().numpy()
    ora = _oracles(task, 13, n, variant=variant)
    oobs = np.stack([o.reset() for o in ora])
    tol = 2e-3 if task == "placing" else 5e-6
    if task == "placing":
        o32 = np.stack([o.reset() for o in _oracles(task, 13, n, variant=variant, f32=True)])
        ke = np.abs(o32 - oobs).max(1) > tol / 4     # the settle is chaotic there: float32 rounding alone moves it
        print(f"placing reset: {int(ke.sum())} of {n} envs excluded (oracle float32 vs float64 reset differ)")
        assert ke.mean() < 0.3
    else:
        ke = np.zeros(n, bool)
    assert (np.abs(obs - oobs).max(1) <= tol)[~ke].all(), np.abs(obs - oobs).max(1)
    if task != "placing":
        env.step(torch.zeros(n, 7, device="cuda"))  # leave the dist == 0 knife edge of the reset pose
    q, v, aux = [t_.cpu().numpy() for t_ in env.get_state()]
    a = np.random.default_rng(17).uniform(-1, 1, (n, 7)).astype(np.float32)
    o64, _, (kd, kc) = live_answers_and_knife_edges(task, q, v, aux[:, :64], act=a, variant=variant, seed=13)
    o1, r1, d1, s1 = [t_.cpu().numpy() for t_ in env.step(torch.from_numpy(a).cuda())]
    check_step(f"{task}/{variant}", dict(obs=o1, reward=r1, done=d1, succ=s1), o64, kd, kc, max_discrete_frac=0.35 if task == "placing" else 0.1)


@pytest.mark.gpu
@pytest.mark.parametrize("task", ["releasing", "carrying"])
def test_releasing_and_carrying_reset_and_step(task):
    """the releasing / carrying branches (util:106-117 in-hand start, util:186-189 nine init angles, util:123-167 servo loops
    for carrying, util:549-566 terminal logic, reward 0, task_max_steps 500): reset against the oracle where the oracle's own float32
    build reproduces its float64 reset, then one env step from the kernel's injected state (done / succ exact, reward 1e-5)."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n = 64
    env = JacoVecEnv(n, task=task, seed=29, auto_reset=False)
    obs = env.reset().cpu().numpy()
    ora = _oracles(task, 29, n)
    for o in ora:
        o.set_probe(True)
    oobs = np.stack([o.reset() for o in ora])
    tke = np.array([bool(o.touch_knife_edge()) for o in ora])      # a pad contact within 2 um of a site boundary / of making or breaking
    o32 = np.stack([o.reset() for o in _oracles(task, 29, n, f32=True)])
    tol = 2e-3
    ke = np.abs(o32 - oobs).max(1) > tol / 4
    print(f"{task} reset: {int(ke.sum())} of {n} envs excluded (oracle float32 vs float64 reset differ), {int(tke.sum())} touch knife edges")
    assert ke.mean() < 0.3
    assert (np.abs(obs - oobs)[:, 1:].max(1) <= tol)[~ke].all(), np.abs(obs - oobs).max(1)
    # touch class after the 150 in-hand settle substeps: exact outside the probes' knife-edge set; a pad force that crosses the 1e-3
    # sensor threshold on the last substep is not covered by a geometric probe -- counted, printed and bounded by 2 envs
    tmis = (obs[:, 0] != oobs[:, 0]) & ~ke & ~tke
    print(f"{task} reset: touch class differs on {int(tmis.sum())} env(s) outside the knife-edge sets: {np.nonzero(tmis)[0].tolist()}")
    assert tmis.sum() <= 2
    q, v, aux = [t_.cpu().numpy() for t_ in env.get_state()]
    if task == "releasing":
        assert (np.abs(obs[:, 8:11] - obs[:, 1:4]).max(1) < 0.12).all()   # object in the hand
    a = np.random.default_rng(31).uniform(-1, 1, (n, 7)).astype(np.float32)
    o64, _, (kd, kc) = live_answers_and_knife_edges(task, q, v, aux[:, :64], act=a, seed=29)
    o1, r1, d1, s1 = [t_.cpu().numpy() for t_ in env.step(torch.from_numpy(a).cuda())]
    check_step(task, dict(obs=o1, reward=r1, done=d1, succ=s1), o64, kd, kc, max_discrete_frac=0.35)
    if task == "carrying":
        assert d1.all() and not s1.any()       # util:549-550: (True, 0, wb) on every step
    env.close()


@pytest.mark.gpu
def test_env_placement_does_not_change_any_env():
    """The step kernel places envs on threads by last step's workload (csrc/jaco_kernel_coop.cu, DESIGN.md 5.4: busiest envs dealt over
    the first wave of blocks, quiet envs in the remaining blocks).  An env's arithmetic must not depend on where it runs: a 40 960-env
    handle (160 blocks: both branches of the placement map, keys non-trivial after the first step) must reproduce, bit for bit, small
    handles that own the same global env ids (env_offset) and run with a different placement."""
    from mujoco_jaco_b200.env import JacoVecEnv
    n, m = 40960, 256
    big = JacoVecEnv(n, task="picking", seed=77, auto_reset=True)
    offs = [0, 20224, 40000, n - m]
    small = [JacoVecEnv(m, task="picking", seed=77, auto_reset=True, env_offset=o) for o in offs]
    ob = big.reset().clone()
    for o, e in zip(offs, small):
        assert torch.equal(e.reset(), ob[o:o + m])
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    for k in range(4):
        a = torch.rand(n, 7, device="cuda", generator=g) * 2 - 1
        if k == 2:      # push a share of the arms down onto the scene so that the keys differ widely
            a[::3, 2] = -1.0
        ob, rb, db, sb = [t.clone() for t in big.step(a)]
        for o, e in zip(offs, small):
            os_, rs, ds, ss = e.step(a[o:o + m].contiguous())
            assert torch.equal(os_, ob[o:o + m]) and torch.equal(rs, rb[o:o + m]) and torch.equal(ds, db[o:o + m]) and torch.equal(ss, sb[o:o + m]), (k, o)
    qb, vb, ab = big.get_state()
    for o, e in zip(offs, small):
        q, v, a_ = e.get_state()
        assert torch.equal(q, qb[o:o + m]) and torch.equal(v, vb[o:o + m])
        e.close()
    big.close()
