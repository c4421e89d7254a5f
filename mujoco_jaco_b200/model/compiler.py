"""Offline model compiler: MJCF + STL  ->  packed constants blob for the oracle and the CUDA kernels.

SURVEY.md §7 step 1.  Input is the reference asset
/root/reference/env_script/assets/jaco2/jaco2_curtain_torque.xml (+ meshes/*.STL); output is
  * mujoco_jaco_b200/model/jaco2_curtain_torque.npz   (named float64 arrays, committed, travels to the GPU box)
  * include/jaco_model_layout.h                         (offsets of every array inside the flat blob)
MuJoCo is not involved; the MJCF compile rules that matter are restated here:
  - a body's inertia comes from its <inertial> if present, else from its geoms at their density
    (mesh geoms: exact signed-volume integrals of the STL triangles);
  - welded children (no joint) are merged into the nearest jointed ancestor -> 9 rigid groups on the arm
    (link1..link6, thumb, index, pinky) + 2 free bodies (object_body, object_dest);
  - body_invweight0 / dof_invweight0 / meaninertia are evaluated at qpos0 on the *unmerged* tree the way
    mj_setConst does (mean diagonal of J M^-1 J^T per body);
  - contact parameters of a geom pair are mixed with MuJoCo 2.0 rules (condim max, friction max,
    solref/solimp equal-weight mean, margin max).
Collision geometry is the north star's primitive set (BASELINE.json): exact boxes for the 19 touch pads, the
object, the holder and the destination box, the floor plane, and fitted capsules standing in for the finger /
hand / fore-arm convex hulls.
Run:  python -m mujoco_jaco_b200.model.compiler /root/reference/env_script/assets/jaco2/jaco2_curtain_torque.xml
"""
from __future__ import annotations

import os
import sys

import numpy as np

from .mjcf import MJCF, mesh_mass_props, quat_mul, quat_to_mat

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))

# fixed topology of the compiled model -------------------------------------------------------------------
CHAIN_BODIES = ["link1", "link2", "link3", "link4", "link5", "link6",
                "thumb_proximal", "index_proximal", "pinky_proximal"]
CHAIN_PARENT = [-1, 0, 1, 2, 3, 4, 5, 5, 5]
FREE_BODIES = ["object_body", "object_dest"]
# body ids used by geoms/sites in the blob: 0 = world, 1..9 = chain bodies, 10 = object, 11 = dest
NB = 12
GEOM_PLANE, GEOM_BOX, GEOM_CAPSULE = 0, 1, 2
SITE_BOX, SITE_CYL = 0, 1
NGEOM_FIELDS = 32   # per-geom record, see GEOM_REC below
NPAIR_FIELDS = 20
NSITE_FIELDS = 20


def rot_inertia(I, R):
    return R @ I @ R.T


def shift_inertia(I, m, d):
    """Inertia about a point displaced by d from the COM (parallel axis)."""
    return I + m * (np.dot(d, d) * np.eye(3) - np.outer(d, d))


def box_inertia(m, h):
    x, y, z = 2 * h
    return np.diag([m * (y * y + z * z) / 12, m * (x * x + z * z) / 12, m * (x * x + y * y) / 12])


class Compiled:
    pass


def geom_mass_props(model, g):
    """(mass, com, inertia about com) of one geom in its body's frame."""
    R = quat_to_mat(g.quat)
    if g.type == "box":
        m = g.density * 8 * g.size[0] * g.size[1] * g.size[2]
        return m, g.pos.copy(), rot_inertia(box_inertia(m, g.size), R)
    if g.type == "sphere":
        m = g.density * 4 / 3 * np.pi * g.size[0] ** 3
        return m, g.pos.copy(), np.eye(3) * 0.4 * m * g.size[0] ** 2
    if g.type == "mesh":
        vol, com, I = mesh_mass_props(model.load_mesh(g.mesh))
        m = g.density * vol
        return m, g.pos + R @ com, rot_inertia(g.density * I, R)
    raise NotImplementedError(g.type)


def body_inertial(model, b):
    """(mass, com, I_com) of one MJCF body in its own frame (explicit <inertial> wins over geoms)."""
    if b.inertial is not None:
        I = np.diag(b.inertial["diaginertia"])
        return b.inertial["mass"], b.inertial["pos"].copy(), I
    mass, com, parts = 0.0, np.zeros(3), []
    for g in b.geoms:
        if g.type == "plane":
            continue
        m, c, I = geom_mass_props(model, g)
        parts.append((m, c, I))
        mass += m
        com += m * c
    if mass == 0:
        return 0.0, np.zeros(3), np.zeros((3, 3))
    com /= mass
    I = sum(shift_inertia(Ii, mi, ci - com) for mi, ci, Ii in parts)
    if model.balanceinertia:
        ev = np.linalg.eigvalsh(I)
        if ev[0] + ev[1] < ev[2]:
            I = np.eye(3) * ev.mean()
    return mass, com, I


def rel_frame(model, bid, root):
    """Pose of body `bid` expressed in ancestor `root`'s frame (zero joint displacement)."""
    pos, q = np.zeros(3), np.array([1.0, 0, 0, 0])
    while bid != root:
        b = model.bodies[bid]
        pos = b.pos + quat_to_mat(b.quat) @ pos
        q = quat_mul(b.quat, q)
        bid = b.parent
        assert bid >= 0
    return pos, q


def welded_group(model, root):
    out, stack = [], [root]
    while stack:
        bid = stack.pop()
        out.append(bid)
        for c in model.bodies[bid].children:
            if not model.bodies[c].joints:
                stack.append(c)
    return out


def fit_capsule(verts, radius):
    """Capsule (p0, p1, r) along the principal axis of a vertex cloud; spans its extent."""
    c = verts.mean(0)
    w, v = np.linalg.eigh(np.cov((verts - c).T))
    ax = v[:, -1]
    if ax[np.argmax(np.abs(ax))] < 0:
        ax = -ax
    s = (verts - c) @ ax
    lo, hi = s.min() + radius, s.max() - radius
    if hi < lo:
        lo = hi = 0.5 * (lo + hi)
    return c + lo * ax, c + hi * ax, radius


def mix_pair(g1, g2):
    """MuJoCo 2.0 contact-parameter mixing for equal priority / solmix."""
    return dict(
        condim=max(g1["condim"], g2["condim"]),
        friction=np.maximum(g1["friction"], g2["friction"]),
        solref=0.5 * (g1["solref"] + g2["solref"]),
        solimp=0.5 * (g1["solimp"] + g2["solimp"]),
        margin=max(g1["margin"], g2["margin"]),
    )


def compile_model(xml_path):
    model = MJCF(xml_path)
    C = Compiled()
    A = {}  # named arrays of the blob

    chain_ids = [model.body_id(n) for n in CHAIN_BODIES]
    free_ids = [model.body_id(n) for n in FREE_BODIES]
    group_of = {}  # mjcf body id -> blob body id (0 world)
    for k, bid in enumerate(chain_ids):
        for w in welded_group(model, bid):
            group_of[w] = 1 + k
    for k, bid in enumerate(free_ids):
        for w in welded_group(model, bid):
            group_of[w] = 10 + k

    # ---- chain kinematics ------------------------------------------------------------------------------
    pos = np.zeros((9, 3)); quat = np.zeros((9, 4)); axis = np.zeros((9, 3))
    ref = np.zeros(9); rng = np.zeros((9, 2)); limited = np.zeros(9); damping = np.zeros(9)
    for k, bid in enumerate(chain_ids):
        b = model.bodies[bid]
        par = chain_ids[CHAIN_PARENT[k]] if CHAIN_PARENT[k] >= 0 else 0
        p, q = rel_frame(model, bid, par)
        pos[k], quat[k] = p, q
        (j,) = b.joints
        assert j.type == "hinge" and np.allclose(j.pos, 0)
        axis[k], ref[k], damping[k] = j.axis, j.ref, j.damping
        limited[k], rng[k] = float(j.limited), j.range
    A["chain_pos"], A["chain_quat"], A["chain_axis"] = pos, quat, axis
    A["chain_ref"], A["chain_range"], A["chain_limited"], A["chain_damping"] = ref, rng, limited, damping

    # ---- merged inertias ------------------------------------------------------------------------------
    def merged(root):
        parts = []
        for w in welded_group(model, root):
            m, c, I = body_inertial(model, model.bodies[w])
            if m == 0:
                continue
            p, q = rel_frame(model, w, root)
            R = quat_to_mat(q)
            parts.append((m, p + R @ c, rot_inertia(I, R)))
        M = sum(p[0] for p in parts)
        com = sum(p[0] * p[1] for p in parts) / M
        I = sum(shift_inertia(Ii, mi, ci - com) for mi, ci, Ii in parts)
        return M, com, I

    def pack_I(I):
        return np.array([I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2]])

    cm = np.zeros(9); cc = np.zeros((9, 3)); cI = np.zeros((9, 6))
    for k, bid in enumerate(chain_ids):
        m, c, I = merged(bid)
        cm[k], cc[k], cI[k] = m, c, pack_I(I)
    A["chain_mass"], A["chain_com"], A["chain_inertia"] = cm, cc, cI
    fm = np.zeros(2); fc = np.zeros((2, 3)); fI = np.zeros((2, 6)); fq0 = np.zeros((2, 7))
    for k, bid in enumerate(free_ids):
        m, c, I = merged(bid)
        fm[k], fc[k], fI[k] = m, c, pack_I(I)
        fq0[k, :3] = model.bodies[bid].pos
        fq0[k, 3:] = model.bodies[bid].quat
    A["free_mass"], A["free_com"], A["free_inertia"], A["free_qpos0"] = fm, fc, fI, fq0

    # ---- actuators -------------------------------------------------------------------------------------
    act = model.actuators
    assert [a["kind"] for a in act] == ["motor"] * 6 + ["position"] * 3
    A["motor_forcerange"] = np.array([a["forcerange"] for a in act[:6]])
    A["finger_act"] = np.array([[a["kp"], *a["ctrlrange"], *a["forcerange"]] for a in act[6:]])  # kp, clo, chi, flo, fhi

    # ---- frames the env logic reads -------------------------------------------------------------------
    link6 = model.body_id("link6")
    A["ee_offset"] = rel_frame(model, model.body_id("EE"), link6)[0]
    A["ee_obj_offset"] = rel_frame(model, model.body_id("EE_obj"), link6)[0]
    A["base_position"] = rel_frame(model, model.body_id("link1"), 0)[0]          # env_mujoco_util.py:35
    A["base_link_position"] = rel_frame(model, model.body_id("base_link"), 0)[0]  # _cis variant :42
    # _trans variant (env_mujoco_util_trans.py:43): base_link orientation as Gohlke 'rxyz' Euler angles, R = Rx(a) Ry(b) Rz(c)
    Rb = quat_to_mat(rel_frame(model, model.body_id("base_link"), 0)[1])
    A["base_link_euler"] = np.array([np.arctan2(-Rb[1, 2], Rb[2, 2]), np.arcsin(np.clip(Rb[0, 2], -1, 1)), np.arctan2(-Rb[0, 1], Rb[0, 0])])

    # ---- unmerged tree at qpos0: M, invweight0 (mj_setConst semantics) ---------------------------------
    nv = 21
    dof_of = {}  # mjcf body id -> list of (dof index, kind, axis_world, anchor_world)
    xpos, xmat = {0: np.zeros(3)}, {0: np.eye(3)}
    order = []
    def walk(bid):
        b = model.bodies[bid]
        if bid != 0:
            xpos[bid] = xpos[b.parent] + xmat[b.parent] @ b.pos
            xmat[bid] = xmat[b.parent] @ quat_to_mat(b.quat)
        order.append(bid)
        for c in b.children:
            walk(c)
    walk(0)
    dofs = []  # (body, kind, axis_world, anchor_world)
    for bid in order:  # MuJoCo numbers dofs depth-first in XML order
        for j in model.bodies[bid].joints:
            if j.type == "hinge":
                dofs.append((bid, "h", xmat[bid] @ j.axis, xpos[bid]))
            else:
                for e in np.eye(3):
                    dofs.append((bid, "t", e, xpos[bid]))
                for e in np.eye(3):
                    dofs.append((bid, "r", xmat[bid] @ e, xpos[bid]))
    assert len(dofs) == nv
    dof_body = [d[0] for d in dofs]
    assert dof_body[:9] == chain_ids and dof_body[9] == free_ids[0] and dof_body[15] == free_ids[1]

    def ancestors(bid):
        out = set()
        while bid > 0:
            out.add(bid)
            bid = model.bodies[bid].parent
        return out

    def jac(bid, point):
        Jp, Jr = np.zeros((3, nv)), np.zeros((3, nv))
        anc = ancestors(bid)
        for i, (db, kind, ax, anchor) in enumerate(dofs):
            if db not in anc:
                continue
            if kind == "t":
                Jp[:, i] = ax
            else:
                Jr[:, i] = ax
                Jp[:, i] = np.cross(ax, point - anchor)
        return Jp, Jr

    M0 = np.zeros((nv, nv))
    binert = {}
    for bid in order[1:]:
        m, c, I = body_inertial(model, model.bodies[bid])
        binert[bid] = (m, c, I)
        if m == 0 or not ancestors(bid) & set(dof_body):
            continue
        Jp, Jr = jac(bid, xpos[bid] + xmat[bid] @ c)
        M0 += m * Jp.T @ Jp + Jr.T @ rot_inertia(I, xmat[bid]) @ Jr
    Minv = np.linalg.inv(M0)
    C.M0 = M0
    invw = {0: (0.0, 0.0)}
    for bid in order[1:]:
        if not ancestors(bid) & set(dof_body):
            invw[bid] = (0.0, 0.0)
            continue
        m, c, _ = binert[bid]
        Jp, Jr = jac(bid, xpos[bid] + xmat[bid] @ c)
        Ap, Ar = Jp @ Minv @ Jp.T, Jr @ Minv @ Jr.T
        invw[bid] = (np.trace(Ap) / 3, np.trace(Ar) / 3)
    dinv = np.diag(Minv).copy()
    for s in (9, 15):  # free joints: mean over the 3 translational / 3 rotational dofs
        dinv[s:s + 3] = dinv[s:s + 3].mean()
        dinv[s + 3:s + 6] = dinv[s + 3:s + 6].mean()
    A["dof_invweight0"] = dinv
    A["meaninertia"] = np.array([np.trace(M0) / nv])
    C.body_invweight0 = {model.bodies[b].name: v for b, v in invw.items()}

    # ---- collision geoms -------------------------------------------------------------------------------
    geoms = []  # dict records

    def add_geom(name, gtype, body_mjcf, p, R, size, src, rbound):
        invw_t, invw_r = invw[body_mjcf]
        geoms.append(dict(name=name, type=gtype, body=group_of.get(body_mjcf, 0), pos=p, R=R, size=np.asarray(size, float),
                          condim=src.condim, friction=src.friction, solref=src.solref, solimp=src.solimp,
                          margin=src.margin, invweight=(invw_t, invw_r), rbound=rbound, src_body=body_mjcf))
        return len(geoms) - 1

    def group_root(bid):
        g = group_of.get(bid, 0)
        if g == 0:
            return 0
        return chain_ids[g - 1] if g < 10 else free_ids[g - 10]

    def add_mjcf_box(bid, g):
        root = group_root(bid)
        p, q = rel_frame(model, bid, root)
        R = quat_to_mat(q)
        return add_geom(g.name, GEOM_BOX, bid, p + R @ g.pos, R @ quat_to_mat(g.quat), g.size, g,
                        float(np.linalg.norm(g.size)))

    def add_capsule(name, bid, g, radius):
        root = group_root(bid)
        p, q = rel_frame(model, bid, root)
        R = quat_to_mat(q) @ quat_to_mat(g.quat)
        v = model.load_mesh(g.mesh).reshape(-1, 3) @ R.T + (p + quat_to_mat(q) @ g.pos)
        p0, p1, r = fit_capsule(v, radius)
        half = 0.5 * np.linalg.norm(p1 - p0)
        zax = (p1 - p0) / (2 * half)
        xax = np.cross([0, 1, 0], zax) if abs(zax[1]) < 0.9 else np.cross([1, 0, 0], zax)
        xax /= np.linalg.norm(xax)
        Rc = np.stack([xax, np.cross(zax, xax), zax], axis=1)
        return add_geom(name, GEOM_CAPSULE, bid, 0.5 * (p0 + p1), Rc, [r, half, 0], g, half + r)

    world = model.bodies[0]
    floor = add_geom("floor", GEOM_PLANE, 0, world.geoms[0].pos, np.eye(3), [0, 0, 0], world.geoms[0], 0.0)
    hb = model.bodies[model.body_id("object_holder")]
    holder = add_geom("holder", GEOM_BOX, 0, hb.pos + hb.geoms[0].pos, np.eye(3), hb.geoms[0].size, hb.geoms[0],
                      float(np.linalg.norm(hb.geoms[0].size)))
    obj = add_mjcf_box(free_ids[0], model.bodies[free_ids[0]].geoms[0])
    dest = add_mjcf_box(free_ids[1], model.bodies[free_ids[1]].geoms[0])
    geoms[obj]["name"], geoms[dest]["name"] = "object", "dest"

    # touch pads: the box geom that shares a body with site "<k>_touch"
    pad_of_site = {}
    pads = []
    site_recs = [None] * 20
    for bid, b in enumerate(model.bodies):
        for s in b.sites:
            if s.name == "EE_touch":
                k = 19          # env_mujoco_util.py:475 puts EE_touch last
            else:
                k = int(s.name.split("_")[0])
            root = group_root(bid)
            p, q = rel_frame(model, bid, root)
            R = quat_to_mat(q)
            site_recs[k] = dict(name=s.name, body=group_of[bid], src_body=bid,
                                type=SITE_BOX if s.type == "box" else SITE_CYL,
                                pos=p + R @ s.pos, R=R @ quat_to_mat(s.quat), size=np.resize(s.size, 3))
            if s.name != "EE_touch":
                (g,) = b.geoms
                gi = add_mjcf_box(bid, g)
                pads.append(gi)
                pad_of_site[k] = gi
    # capsules standing in for hulls (finger links radius 9 mm keeps the capsule surface at or inside the pads)
    caps_hand = []
    for fname in ("thumb", "index", "pinky"):
        for part in ("proximal", "distal"):
            bid = model.body_id(f"{fname}_{part}")
            (g,) = [g for g in model.bodies[bid].geoms if g.type == "mesh"]
            caps_hand.append(add_capsule(f"{fname}_{part}_cap", bid, g, 0.009))
    hand_g = [g for g in model.bodies[link6].geoms if g.name == "link6"][0]
    # hand hull: radius 45 mm capsule along the wrist axis, tip kept 3.5 mm behind the palm pad face
    Rh = quat_to_mat(hand_g.quat)
    zdir = Rh @ np.array([0, 0, -1.0])  # mesh spans z in [-0.121, 0] -> body +z after the 3.14 flip
    p0, p1 = hand_g.pos + zdir * 0.045, hand_g.pos + zdir * 0.070
    half = 0.5 * np.linalg.norm(p1 - p0); zax = (p1 - p0) / (2 * half)
    xax = np.cross([0, 1, 0], zax); xax /= np.linalg.norm(xax)
    hand_cap = add_geom("hand_cap", GEOM_CAPSULE, link6, 0.5 * (p0 + p1),
                        np.stack([xax, np.cross(zax, xax), zax], axis=1), [0.045, half, 0], hand_g, half + 0.045)
    caps_hand.append(hand_cap)
    # fore-arm hulls (link3, link4, link5) as capsules: only tested against the static scene + dest box
    caps_arm = []
    for lname, rad in (("link3", 0.035), ("link4", 0.03), ("link5", 0.03)):
        bid = model.body_id(lname)
        g = [g for g in model.bodies[bid].geoms if g.name == lname][0]
        caps_arm.append(add_capsule(f"{lname}_cap", bid, g, rad))

    # ---- pair table ---------------------------------------------------------------------------------
    pairs = []

    def add_pair(g1, g2, group):
        mx = mix_pair(geoms[g1], geoms[g2])
        pairs.append(dict(g1=g1, g2=g2, group=group, **mx))

    # group 0: free-body scene pairs, always tested
    add_pair(floor, obj, 0); add_pair(holder, obj, 0); add_pair(dest, obj, 0); add_pair(floor, dest, 0)
    # group 1: hand-cluster geoms vs object (culled together by one hand/object sphere test)
    for gi in pads + caps_hand:
        add_pair(gi, obj, 1)
    # group 2: hand-cluster capsules + pads vs holder ; 3: vs floor ; 4: vs dest
    for grp, other in ((2, holder), (3, floor), (4, dest)):
        for gi in caps_hand:
            add_pair(other, gi, grp)
    # group 5: fore-arm capsules vs static scene and dest
    for gi in caps_arm:
        for other in (holder, floor, dest):
            add_pair(other, gi, 5)

    G = np.zeros((len(geoms), NGEOM_FIELDS))
    for i, g in enumerate(geoms):
        G[i, 0] = g["type"]; G[i, 1] = g["body"]; G[i, 2:5] = g["pos"]; G[i, 5:14] = g["R"].reshape(-1)
        G[i, 14:17] = g["size"]; G[i, 17] = g["rbound"]; G[i, 18:20] = g["invweight"]
        G[i, 20] = g["condim"]; G[i, 21:24] = g["friction"]; G[i, 24:26] = g["solref"]; G[i, 26:31] = g["solimp"]
        G[i, 31] = g["margin"]
    P = np.zeros((len(pairs), NPAIR_FIELDS))
    for i, p in enumerate(pairs):
        P[i, 0] = p["g1"]; P[i, 1] = p["g2"]; P[i, 2] = p["group"]; P[i, 3] = p["condim"]
        P[i, 4:7] = p["friction"]; P[i, 7:9] = p["solref"]; P[i, 9:14] = p["solimp"]; P[i, 14] = p["margin"]
        # diagApprox pieces (mj_diagApprox): translational / rotational inverse weights of the two bodies
        P[i, 15] = geoms[p["g1"]]["invweight"][0] + geoms[p["g2"]]["invweight"][0]
        P[i, 16] = geoms[p["g1"]]["invweight"][1] + geoms[p["g2"]]["invweight"][1]
    S = np.zeros((20, NSITE_FIELDS))
    for k, s in enumerate(site_recs):
        S[k, 0] = s["type"]; S[k, 1] = s["body"]; S[k, 2:5] = s["pos"]; S[k, 5:14] = s["R"].reshape(-1)
        S[k, 14:17] = s["size"]
        # which geom's contacts feed this sensor (contacts involving the site's own MJCF body)
        S[k, 17] = pad_of_site[k] if k in pad_of_site else hand_cap
    A["geoms"], A["pairs"], A["sites"] = G, P, S
    # bounding sphere of the hand cluster in link6 frame (for the group broad phase)
    A["hand_sphere"] = np.array([0.0, 0.0, 0.13, 0.125])
    # options (MuJoCo defaults without <option>; timestep overridden by env_script/mujoco.py:39)
    A["opt"] = np.array([0.001, -9.81, 100, 1e-8])  # timestep, gravity z, solver iterations, tolerance

    C.arrays = A
    C.geoms, C.pairs, C.sites = geoms, pairs, site_recs
    C.model = model
    return C


def pack(arrays):
    names = sorted(arrays)
    off, layout = 0, {}
    for n in names:
        a = np.asarray(arrays[n], dtype=np.float64)
        layout[n] = (off, a.shape)
        off += a.size
    blob = np.concatenate([np.asarray(arrays[n], dtype=np.float64).reshape(-1) for n in names])
    return blob, layout


def write_layout_header(path, layout, total, counts):
    lines = ["/* AUTO-GENERATED by mujoco_jaco_b200/model/compiler.py -- offsets (in elements) of each named array",
             " * inside the flat model blob that jaco_create() and the oracle take.  Do not edit. */",
             "#ifndef JACO_MODEL_LAYOUT_H", "#define JACO_MODEL_LAYOUT_H", ""]
    for n, (off, shape) in layout.items():
        lines.append(f"#define JM_{n.upper()} {off}  /* shape {tuple(shape)} */")
    lines.append(f"#define JM_TOTAL {total}")
    for k, v in counts.items():
        lines.append(f"#define JM_{k} {v}")
    lines += ["", "#endif", ""]
    open(path, "w").write("\n".join(lines))


def main(xml_path):
    C = compile_model(xml_path)
    blob, layout = pack(C.arrays)
    counts = dict(NGEOM=len(C.geoms), NPAIR=len(C.pairs), NSITE=20, NGEOM_FIELDS=NGEOM_FIELDS,
                  NPAIR_FIELDS=NPAIR_FIELDS, NSITE_FIELDS=NSITE_FIELDS)
    out = os.path.join(HERE, os.path.splitext(os.path.basename(xml_path))[0] + ".npz")
    np.savez(out, blob=blob, geom_names=np.array([g["name"] for g in C.geoms]),
             **{f"a_{k}": v for k, v in C.arrays.items()})
    write_layout_header(os.path.join(REPO, "include", "jaco_model_layout.h"), layout, blob.size, counts)
    print(f"wrote {out}: {blob.size} doubles, {len(C.geoms)} geoms, {len(C.pairs)} pairs")
    return C


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else
         "/root/reference/env_script/assets/jaco2/jaco2_curtain_torque.xml")
