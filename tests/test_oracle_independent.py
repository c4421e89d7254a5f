"""Checks of the oracle (and of the kernels' per-env code) that do not re-use the oracle's own algorithms (round-1 verdict, "make the
oracle less of a mirror"):
  (a) solver optimality: the returned qacc must satisfy the first-order condition of MuJoCo's convex problem
      min_a 1/2 (a - a_s)' M (a - a_s) + sum_r 1/2 D_r min(0, J_r a - aref_r)^2, evaluated here in numpy from the exported problem data;
  (b) geometry: every contact the oracle reports must lie on / inside both geoms (within the margin) with a unit normal, and its
      distance must agree with a brute-force numpy sampling of the two primitives' surfaces;
  (c) touch sensors: a site is credited exactly when the contact's ray passes through the site box (analytic slab test in numpy)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle.oracle import OracleEnv
from parity_util import load_snapshots

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SEL = np.arange(0, 4096, 23)      # 179 in-contact snapshots across all sources / touch classes


def _grad(M, J, D, aref, qs, a):
    jar = J @ a - aref
    f = -D * np.minimum(jar, 0.0)
    return M @ (a - qs) - J.T @ f, f


def test_solver_optimality_kkt_of_oracle_and_kernel_code():
    from hostsim import HostSim
    Z, o64, ke = load_snapshots(os.path.join(GOLD, "contact_snapshots.npz"))
    q, v, aux, ctrl = [np.asarray(Z[k])[SEL] for k in ("qpos", "qvel", "aux", "ctrl")]
    n = len(SEL)
    hs = HostSim(n, task="picking", seed=0)
    hs.reset()
    _, _, a0 = hs.get_state()
    a0[:, :64] = aux
    hs.set_state(q.astype(np.float32), v.astype(np.float32), a0)
    _, _, qacc_k = hs.substep(ctrl.astype(np.float32))          # kernel code (general path), float32
    worst_o, worst_k, rows = 0.0, 0.0, 0
    for i in range(n):
        o = OracleEnv(task="picking", seed=0, env_id=int(SEL[i]))
        o.set_state(q[i].astype(np.float64), v[i].astype(np.float64), aux[i].astype(np.float64))
        f = o.forward(ctrl[i].astype(np.float64))
        J, D, aref, qs = o.efc()
        rows += len(D)
        g, frc = _grad(f["M"], J, D, aref, qs, f["qacc"])
        scale = np.linalg.norm(f["M"] @ (f["qacc"] - qs)) + np.linalg.norm(J.T @ frc) + 1e-12
        worst_o = max(worst_o, np.linalg.norm(g) / scale)
        gk, fk = _grad(f["M"], J, D, aref, qs, qacc_k[i].astype(np.float64))
        scale_k = np.linalg.norm(f["M"] @ (qacc_k[i] - qs)) + np.linalg.norm(J.T @ fk) + 1e-12
        worst_k = max(worst_k, np.linalg.norm(gk) / scale_k)
    print(f"KKT residual |grad| / (|M (a - a_s)| + |J' f|) over {n} in-contact snapshots ({rows} rows): oracle {worst_o:.2e}, kernel code {worst_k:.2e}")
    assert rows > 20 * n
    assert worst_o < 1e-6            # float64 Newton, tolerance 1e-8
    assert worst_k < 2e-3            # float32 kernel code on the float64 problem data


def _surface_points(P, npts=4000, rng=None):
    """random points on the surface of a box / capsule primitive record (pos 3, R 9, size 3, type), world frame"""
    pos, R, size, typ = P[:3], P[3:12].reshape(3, 3), P[12:15], int(P[15])
    rng = rng or np.random.default_rng(0)
    if typ == 1:
        u = rng.uniform(-1, 1, (npts, 3)) * size
        ax = rng.integers(0, 3, npts)
        u[np.arange(npts), ax] = size[ax] * rng.choice([-1.0, 1.0], npts)
    else:
        d = rng.normal(size=(npts, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
        z = rng.uniform(-size[1], size[1], npts)
        cyl = rng.random(npts) < size[1] / (size[1] + size[0] + 1e-12)
        u = np.where(cyl[:, None], np.c_[size[0] * d[:, 0] / np.hypot(d[:, 0], d[:, 1]), size[0] * d[:, 1] / np.hypot(d[:, 0], d[:, 1]), z],
                     size[0] * d + np.c_[np.zeros(npts), np.zeros(npts), np.sign(d[:, 2]) * size[1]])
    return pos + u @ R.T


def _signed_dist(P, x):
    """signed distance of world points x to a plane / box / capsule primitive (negative inside)"""
    pos, R, size, typ = P[:3], P[3:12].reshape(3, 3), P[12:15], int(P[15])
    l = (x - pos) @ R
    if typ == 0:
        return l[:, 2]
    if typ == 1:
        d = np.abs(l) - size
        return np.linalg.norm(np.maximum(d, 0), axis=1) + np.minimum(d.max(1), 0)
    zc = np.clip(l[:, 2], -size[1], size[1])
    return np.linalg.norm(l - np.c_[np.zeros(len(l)), np.zeros(len(l)), zc], axis=1) - size[0]


def test_contact_geometry_against_brute_force_distance():
    Z, o64, ke = load_snapshots(os.path.join(GOLD, "contact_snapshots.npz"))
    q, v, aux, ctrl = [np.asarray(Z[k])[SEL[::3]] for k in ("qpos", "qvel", "aux", "ctrl")]
    ncon, worst_pen, worst_pos, worst_thin = 0, 0.0, 0.0, 0.0
    for i in range(len(q)):
        o = OracleEnv(task="picking", seed=0, env_id=i)
        o.set_state(q[i].astype(np.float64), v[i].astype(np.float64), aux[i].astype(np.float64))
        o.forward(ctrl[i].astype(np.float64))
        prims = o.render_prims()[:33]                      # world pose + size + type of the 33 compiled geoms
        pairs = {}
        for c in o.contacts():
            p, nrm, dist, g1, g2 = c[:3], c[3:6], c[6], int(c[7]), int(c[8])
            assert abs(np.linalg.norm(nrm) - 1) < 1e-9
            ncon += 1
            # the contact point is the midpoint between the two surfaces along the normal: its true (nearest-point) signed distance to
            # either geom lies between 0 and dist / 2 -- equal to dist / 2 away from edges, closer to 0 next to one
            # (exact for boxes thicker than the penetration and for the plane; the 1 mm touch pads are thinner than some penetrations,
            # so the midpoint can lie beyond their far face, and MuJoCo's capsule-box routine places the point on the capsule's nearest
            # segment point, up to 0.6 mm off the midpoint in deep contacts: those two cases get 1 mm)
            for g in (g1, g2):
                d = _signed_dist(prims[g], p[None])[0]
                dev = max(min(dist / 2, 0.0) - d, d - max(dist / 2, 0.0))
                thin = int(prims[g][15]) == 2 or (int(prims[g][15]) == 1 and prims[g][12:15].min() < 2 * abs(dist))
                if thin:
                    worst_thin = max(worst_thin, dev)
                else:
                    worst_pos = max(worst_pos, dev)
            pairs.setdefault((g1, g2), []).append(dist)
        # brute force per geom pair: no sampled surface point of one geom may be deeper inside the other than the deepest contact the
        # oracle reports for that pair (+ slack for the sampling density)
        for (g1, g2), dists in pairs.items():
            for ga, gb in ((g1, g2), (g2, g1)):
                if int(prims[ga][15]) == 0:
                    continue
                s_ = _surface_points(prims[ga], rng=np.random.default_rng(ncon))
                deepest = _signed_dist(prims[gb], s_).min()
                if min(dists) < 0:
                    worst_pen = max(worst_pen, (-deepest) - (-min(dists)))
    print(f"{ncon} contacts: signed distance of the contact point to either geom leaves [min(dist/2, 0), max(dist/2, 0)] by at most {worst_pos:.2e} (thin pads / capsules {worst_thin:.2e}); brute-force penetration exceeds the reported one by at most {worst_pen:.2e}")
    assert ncon > 300
    assert worst_pos < 5e-6 and worst_thin < 5e-3
    assert worst_pen < 1e-3


def test_touch_sensor_sites_against_analytic_ray_box():
    """sensordata of site k = sum of the normal forces of the contacts of its geom whose ray (contact point, +- normal) passes through
    the site box (mj_sensor touch semantics): recomputed in numpy from the oracle's contacts and the model blob's site table."""
    from hostsim import layout_offsets
    from oracle.oracle import load_blob
    blob = load_blob()
    off = layout_offsets()
    sites = blob[off["SITES"]:off["SITES"] + 20 * 20].reshape(20, 20)
    Z, o64, ke = load_snapshots(os.path.join(GOLD, "contact_snapshots.npz"))
    q, v, aux, ctrl = [np.asarray(Z[k])[SEL[::2]] for k in ("qpos", "qvel", "aux", "ctrl")]
    nsite_hits, checked = 0, 0
    for i in range(len(q)):
        o = OracleEnv(task="picking", seed=0, env_id=i)
        o.set_probe(True)
        o.set_state(q[i].astype(np.float64), v[i].astype(np.float64), aux[i].astype(np.float64))
        f = o.forward(ctrl[i].astype(np.float64))
        if o.touch_knife_edge():
            continue                                        # a contact within 2 um of a site boundary
        prims = o.render_prims()[:33]
        con = o.contacts()
        sd = np.zeros(20)
        for k, S in enumerate(sites):
            typ, body, geom = int(S[0]), int(S[1]), int(S[17])       # blob row: type, body, pos 3, R 9, size 3, geom
            spos_l, sR_l, ssize = S[2:5], S[5:14].reshape(3, 3), S[14:17]
            G = prims[geom]                                  # the site rides on its geom's body: body pose from the geom's world pose
            gb = blob[off["GEOMS"] + 32 * geom: off["GEOMS"] + 32 * (geom + 1)]
            gpos_l, gR_l = gb[2:5], gb[5:14].reshape(3, 3)
            Rb = G[3:12].reshape(3, 3) @ gR_l.T
            pb = G[:3] - Rb @ gpos_l
            spos, sR = pb + Rb @ spos_l, Rb @ sR_l
            for c in con:
                if c[9] <= 0 or (int(c[7]) != geom and int(c[8]) != geom):
                    continue
                ray = c[3:6] * (1.0 if int(c[7]) == geom else -1.0)
                ol, dl = (c[:3] - spos) @ sR, ray @ sR
                t0, t1 = -np.inf, np.inf
                axes = (0, 1, 2) if typ == 0 else (2,)       # box: three slabs; cylinder (axis z): the z slab + the infinite cylinder
                miss = False
                for ax in axes:
                    h = ssize[ax] if typ == 0 else ssize[1]
                    if abs(dl[ax]) < 1e-300:
                        miss |= abs(ol[ax]) > h
                        continue
                    ta, tb = sorted(((-h - ol[ax]) / dl[ax], (h - ol[ax]) / dl[ax]))
                    t0, t1 = max(t0, ta), min(t1, tb)
                if typ == 1:
                    aa, bb, cc = dl[0] ** 2 + dl[1] ** 2, ol[0] * dl[0] + ol[1] * dl[1], ol[0] ** 2 + ol[1] ** 2 - ssize[0] ** 2
                    if aa < 1e-300:
                        miss |= cc > 0
                    else:
                        disc = bb * bb - aa * cc
                        if disc < 0:
                            miss = True
                        else:
                            t0, t1 = max(t0, (-bb - np.sqrt(disc)) / aa), min(t1, (-bb + np.sqrt(disc)) / aa)
                if not miss and t1 >= max(t0, 0.0):
                    sd[k] += c[9]
        checked += 1
        nsite_hits += int((sd > 1e-3).sum())
        assert np.array_equal(sd > 1e-3, f["sensordata"] > 1e-3), (i, sd, f["sensordata"])
        assert np.allclose(sd, f["sensordata"], rtol=1e-9, atol=1e-12)
    print(f"touch sensors: {checked} snapshots, {nsite_hits} active sites reproduced by the analytic ray / box test")
    assert checked > 40 and nsite_hits > 40
