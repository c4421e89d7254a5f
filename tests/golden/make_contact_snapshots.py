"""Generates tests/golden/contact_snapshots.npz: in-contact state snapshots for the one-step parity tests
 and the oracle's answers on them.

Two scripted sources, both driven through the CPU oracle only (oracle/jaco_oracle.c):
  A  in-hand: the reference's own `placing` reset (util:106-117) puts the object between the fingers and settles it
     for 150 substeps; 0..5 further env steps follow with small random arm motion and a scripted gripper (open /
     hold / close).  Pad contacts, touch classes 0 / 1 / 3, object above and below the lift threshold 0.2598.
  B  approach-and-close: `picking` reset, proportional control of the end effector towards the object on the holder
     (action = clip(25 * (goal - ee), -1, 1)) with the gripper closing near the object; a snapshot is taken at every
     env-step boundary at which the hand cluster touches the object or the scene.

Every snapshot is (qpos, qvel, aux) at an env-step boundary, rounded to float32 so that oracle and kernel are fed
bit-identical inputs; it is injected into a *picking* env by the tests.  Stored next to it: a seeded action and raw
ctrl vector, and what the oracle returns for ONE physics substep (state, ncon, sensordata, qacc) and for ONE env step
(obs, reward, done, succ, wb), from the float64 build (the reference semantics) and from the float32 build of the same
oracle, and from the float64 build on inputs perturbed by 1, 4 and 16 float32 ulps (random sign patterns).  These define the
knife-edge predicate of the tests: where float32 rounding alone -- of the arithmetic or of the input -- flips a discrete
outcome of the ORACLE or moves its result past the tolerance, no float32 implementation can be held to it.

usage: python tests/golden/make_contact_snapshots.py [n_snapshots=4096] [workers=os.cpu_count()]
"""
import multiprocessing as mp
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)


def _hand_contact(o):
    """number of contacts of the last forward that involve a hand-cluster / fore-arm geom (geom id >= 4)"""
    con = o.contacts()
    return int(sum(1 for c in con if c[7] >= 4 or c[8] >= 4)), len(con)


def _source_a(wid, count):
    from oracle.oracle import OracleEnv
    rng = np.random.default_rng(1000 + wid)
    out = []
    eid = 0
    while len(out) < count:
        o = OracleEnv(task="placing", seed=7000 + wid, env_id=eid)
        eid += 1
        o.reset()
        grip_dir = rng.choice([-1.0, 0.0, 1.0])
        nsteps = int(rng.integers(0, 6))
        dead = False
        for k in range(nsteps):
            a = np.concatenate([rng.uniform(-0.4, 0.4, 6), [grip_dir]])
            _, _, d, _, _ = o.step(a)
            if d:
                dead = True
                break
        if dead:
            continue
        o.forward()
        nh, nc = _hand_contact(o)
        if nh == 0:
            continue
        q, v, aux = o.get_state()
        out.append((q, v, aux, 0))
    return out


def _source_b(wid, count):
    from oracle.oracle import OracleEnv
    rng = np.random.default_rng(2000 + wid)
    out = []
    eid = 0
    while len(out) < count:
        o = OracleEnv(task="picking", seed=9000 + wid, env_id=eid)
        eid += 1
        obs = o.reset()
        above = rng.uniform(0.0, 0.06)
        taken = 0
        for k in range(60):
            ee, obj = obs[1:4].astype(np.float64), obs[8:11].astype(np.float64)
            goal = obj + np.array([0.0, -0.02, above if k < 25 else -0.01])
            a = np.zeros(7)
            a[:3] = np.clip(25.0 * (goal - ee), -1, 1)
            a[3:6] = rng.uniform(-0.15, 0.15, 3)
            a[6] = 1.0 if np.linalg.norm(goal - ee) < 0.05 else -1.0
            obs, _, d, _, _ = o.step(a)
            if d:
                break
            o.forward()
            nh, nc = _hand_contact(o)
            if nh > 0 and rng.uniform() < 0.5:
                q, v, aux = o.get_state()
                out.append((q, v, aux, 1))
                taken += 1
                if taken >= 4 or len(out) >= count:
                    break
    return out


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    workers = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
    from oracle import oracle as _o
    _o.build()
    na = (n * 5 // 8 + workers - 1) // workers      # 5/8 in-hand, 3/8 approach-and-close
    nb = (n - na * workers + workers - 1) // workers
    with mp.get_context("spawn").Pool(workers) as pool:
        ra = pool.starmap(_source_a, [(w, na) for w in range(workers)])
        rb = pool.starmap(_source_b, [(w, max(nb, 1)) for w in range(workers)])
    snaps = [s for chunk in ra + rb for s in chunk][:n]
    q = np.array([s[0] for s in snaps]).astype(np.float32)
    v = np.array([s[1] for s in snaps]).astype(np.float32)
    aux = np.array([s[2] for s in snaps]).astype(np.float32)
    src = np.array([s[3] for s in snaps], np.uint8)
    rng = np.random.default_rng(4242)
    act = rng.uniform(-1, 1, (len(q), 7)).astype(np.float32)
    ctrl = np.concatenate([rng.uniform(-5, 5, (len(q), 6)), rng.uniform(0.6, 1.0, (len(q), 1)).repeat(3, 1)], 1).astype(np.float32)
    sys.path.insert(0, os.path.join(REPO, "tests"))
    from parity_util import oracle_answers
    out = {}
    # conditioning probes: the float64 oracle from inputs perturbed by one float32 ulp (relative 6e-8, random signs)
    prng = np.random.default_rng(77)
    pert = lambda x, ulps: x.astype(np.float64) * (1.0 + ulps * 6e-8 * prng.choice([-1.0, 1.0], x.shape))
    from parity_util import step_knife_edge, substep_knife_edge
    ans64 = oracle_answers("picking", q, v, aux, act=act, ctrl=ctrl, f32=False, workers=workers)
    for k, x in ans64.items():
        keep64 = k in ("sub_qpos", "sub_qvel", "step_reward", "step_qpos", "step_qvel")
        out[f"o64_{k}"] = x.astype(np.float64 if keep64 else (x.dtype if x.dtype.kind in "iub" else np.float32))
    # knife-edge masks: the probes' full answers are not stored, only where they leave the float64 answer
    z = lambda: np.zeros(len(q), bool)
    ke, causes = dict(sub_d=z(), sub_c=z(), step_d=z(), step_c=z()), []
    # float32 arithmetic accumulates a few ulps through the 9-link kinematic chain before a contact point is tested against a
    # 0.5 mm sensor site or a distance against its margin: the DISCRETE predicate therefore also probes +-4 and +-16 ulp
    # (relative 1e-6, the float32 forward-kinematics error level); the strict continuous bars are only excused by the 1-ulp probes
    probes = [("oracle float32 build", True, (q, v), True)]
    probes += [(f"float64, input +- {u} ulp (pattern {k})", False, (pert(q, u), pert(v, u)), u == 1) for u in (1, 4, 16) for k in range(4)]
    for tag, f32, inp, strict in probes:
        ans = oracle_answers("picking", inp[0], inp[1], aux, act=act, ctrl=ctrl, f32=f32, workers=workers)
        (ad, ac), (bd, bc) = substep_knife_edge(ans64, ans), step_knife_edge(ans64, ans)
        causes.append(f"{tag}: substep {int(ad.sum())} discrete / {int(ac.sum())} strict, env step {int(bd.sum())} / {int(bc.sum())}")
        ke["sub_d"] |= ad; ke["step_d"] |= bd
        if strict:
            ke["sub_c"] |= ac
        ke["step_c"] |= bc                  # 50 chained in-contact substeps amplify, and the CUDA build's 2-ulp division / square root and FMA
                                            # contraction differ from the oracle's arithmetic by more than one ulp: the env step's strict bars yield to every probe
    ke["sub_c"] |= ke["sub_d"]; ke["step_c"] |= ke["step_d"]
    out.update({"ke_" + k: x for k, x in ke.items()})
    print("knife-edge probes:\n  " + "\n  ".join(causes) + "\n  union: " + ", ".join(f"{k} {int(x.sum())}" for k, x in ke.items()))
    out.update(qpos=q, qvel=v, aux=aux, source=src, action=act, ctrl=ctrl)
    path = os.path.join(HERE, "contact_snapshots.npz")
    np.savez_compressed(path, **out)
    touch = out["o64_step_obs"][:, 0]
    print(f"{len(q)} snapshots -> {path} ({os.path.getsize(path) / 1e6:.1f} MB); source A {int((src == 0).sum())} B {int((src == 1).sum())}; "
          f"ncon min/mean/max {out['o64_sub_ncon'].min()}/{out['o64_sub_ncon'].mean():.1f}/{out['o64_sub_ncon'].max()}; "
          f"touch classes after one step {[int((touch == c).sum()) for c in range(4)]}; succ {int(out['o64_step_succ'].sum())} done {int(out['o64_step_done'].sum())}")


if __name__ == "__main__":
    main()
