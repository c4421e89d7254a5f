"""CPU restatement (numpy, float64) of the headless camera -- TEST INFRASTRUCTURE, never imported by the product path.

The reference's `_get_camera` (env_script/env_mujoco_util.py:302-307) is two OpenGL calls (MjRenderContextOffscreen.render +
read_pixels, flipped top-down) on camera 0 of jaco2_curtain_torque.xml:9,286; mujoco-py / OpenGL are unavailable on this box, so this
oracle restates the ray caster of mujoco_jaco_b200/csrc/jaco_render.cuh instead: **parity unpinned** (no reference image exists in the
tree).  What it pins independently of the kernel: the camera model (MuJoCo's convention: the camera looks along -z of its frame, +y up,
fovy vertical), the top-down row order of the reference's np.flip, the OpenGL depth normalisation of read_pixels, and analytic
ray / plane / box / capsule intersections written in vectorised numpy rather than per-ray C.
"""
import numpy as np

PLANE, BOX, CAPSULE = 0, 1, 2


def rays(cam, W, H, dj=0.0, di=0.0):
    """world ray directions [H, W, 3] through the pixel centres (+ sub-pixel offset), unit component along the optical axis"""
    pos, R, fovy = cam[:3].astype(np.float64), cam[3:12].astype(np.float64).reshape(3, 3), float(cam[12])
    ty = np.tan(np.float32(0.5 * fovy * np.pi / 180.0).astype(np.float64))
    tx = ty * W / H
    j, i = np.meshgrid(np.arange(W) + 0.5 + dj, np.arange(H) + 0.5 + di)
    x, y = (2 * j / W - 1) * tx, (1 - 2 * i / H) * ty
    return x[..., None] * R[:, 0] + y[..., None] * R[:, 1] - R[:, 2]


def _hit(P, o, d, tmin, tmax):
    """nearest hit of rays o + t d with one primitive: (t [H, W] with inf where missed, world normal [H, W, 3])"""
    pos, R, size, typ = P[:3], P[3:12].reshape(3, 3), P[12:15], int(P[15])
    ol, dl = R.T @ (o - pos), d @ R
    t = np.full(d.shape[:2], np.inf)
    nl = np.zeros(d.shape)
    with np.errstate(divide="ignore", invalid="ignore"):
        if typ == PLANE:
            tt = -ol[2] / dl[..., 2]
            ok = (np.abs(dl[..., 2]) > 1e-12) & (tt > tmin) & (tt < tmax)
            t[ok] = tt[ok]
            nl[..., 2] = 1.0 if ol[2] > 0 else -1.0
        elif typ == BOX:
            ta, tb = (-size - ol) / dl, (size - ol) / dl
            lo, hi = np.minimum(ta, tb), np.maximum(ta, tb)
            par = np.abs(dl) < 1e-12
            inside = np.abs(ol) <= size
            lo = np.where(par, np.where(inside, -np.inf, np.inf), lo)
            hi = np.where(par, np.where(inside, np.inf, -np.inf), hi)
            ax = np.argmax(lo, axis=-1)
            t0, t1 = np.max(lo, axis=-1), np.minimum(np.min(hi, axis=-1), tmax)
            ok = (t0 > tmin) & (t0 <= t1)
            t[ok] = t0[ok]
            sg = -np.sign(np.take_along_axis(dl, ax[..., None], -1))[..., 0]
            np.put_along_axis(nl, ax[..., None], sg[..., None], -1)
        else:
            rad, h = size[0], size[1]
            a = dl[..., 0] ** 2 + dl[..., 1] ** 2
            b = ol[0] * dl[..., 0] + ol[1] * dl[..., 1]
            c = ol[0] ** 2 + ol[1] ** 2 - rad ** 2
            disc = b * b - a * c
            tc = (-b - np.sqrt(np.where(disc >= 0, disc, np.nan))) / a
            z = ol[2] + tc * dl[..., 2]
            ok = (a > 1e-12) & (disc >= 0) & (tc > tmin) & (tc < tmax) & (np.abs(z) <= h)
            t[ok] = tc[ok]
            nl[ok, 0] = ((ol[0] + tc * dl[..., 0]) / rad)[ok]
            nl[ok, 1] = ((ol[1] + tc * dl[..., 1]) / rad)[ok]
            dd = a + dl[..., 2] ** 2
            for e in (-1.0, 1.0):
                oz = ol[2] - e * h
                bb = b + oz * dl[..., 2]
                cc = ol[0] ** 2 + ol[1] ** 2 + oz ** 2 - rad ** 2
                disc = bb * bb - dd * cc
                ts = (-bb - np.sqrt(np.where(disc >= 0, disc, np.nan))) / dd
                zz = oz + ts * dl[..., 2]
                ok = (disc >= 0) & (ts > tmin) & (ts < np.minimum(t, tmax)) & (zz * e >= 0)
                t[ok] = ts[ok]
                nl[ok, 0] = ((ol[0] + ts * dl[..., 0]) / rad)[ok]
                nl[ok, 1] = ((ol[1] + ts * dl[..., 1]) / rad)[ok]
                nl[ok, 2] = (zz / rad)[ok]
    return t, nl @ R.T


def render(prims, cam, palette, W, H, dj=0.0, di=0.0):
    """(rgb uint8 [H, W, 3], depth float64 [H, W] OpenGL-normalised, hit id int [H, W] with -1 = background, z [H, W] metres)"""
    cam = np.asarray(cam, np.float32)
    o = cam[:3].astype(np.float64)
    zn, zf, amb = float(cam[13]), float(cam[14]), float(cam[15])
    d = rays(cam, W, H, dj, di)
    best = np.full((H, W), zf)
    bid = np.full((H, W), -1)
    bn = np.zeros((H, W, 3))
    for k, P in enumerate(np.asarray(prims, np.float64)):
        t, n = _hit(P, o, d, zn, zf)
        better = t < best
        best[better], bid[better], bn[better] = t[better], k, n[better]
    lam = -(bn * d).sum(-1) / np.linalg.norm(d, axis=-1)
    lam = amb + (1 - amb) * np.maximum(lam, 0)
    col = np.asarray(palette, np.float64)[np.maximum(bid, 0)] * lam[..., None]
    rgb = np.floor(np.clip(col, 0, 1) * 255 + 0.5).astype(np.uint8)
    rgb[bid < 0] = 0
    depth = (1 / zn - 1 / best) / (1 / zn - 1 / zf)
    depth[bid < 0] = 1.0
    return rgb, depth, bid, best
