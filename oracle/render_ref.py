"""numpy restatement of the ray caster (mujoco_sim_b200/csrc/render.cuh) -- TEST INFRASTRUCTURE ONLY.

Same camera model, intersection formulas and shading, in float64 and vectorised over pixels; the geom world poses come from
the oracle's kinematics (stage_dump) or from the caller.  MuJoCo's own OpenGL renderer (entities/camera.py:94-103) is
unavailable on this box; this restatement pins the CUDA kernel to the stated math, not to MuJoCo's pixels."""
import numpy as np

PLANE, SPHERE, CAPSULE, CYLINDER, BOX = 0, 2, 3, 5, 6
AMBIENT, DIFFUSE, EPS = 0.3, 0.7, 1e-6


def quat_to_mat(q):
    q = np.asarray(q, float); q = q / np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)], [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])


def _hit(gtype, size, o, d, tbest, nrm):
    """o, d: [P, 3] local rays.  Updates tbest [P] / nrm [P, 3] in place; returns the mask of pixels that hit."""
    P = len(o)
    hit = np.zeros(P, bool)

    def take(mask, t, n):
        m = mask & (t > EPS) & (t < tbest)
        tbest[m] = t[m]; nrm[m] = n[m]; hit[m] = True

    with np.errstate(divide="ignore", invalid="ignore"):
        if gtype == PLANE:
            t = -o[:, 2] / d[:, 2]
            x, y = o[:, 0] + t * d[:, 0], o[:, 1] + t * d[:, 1]
            ok = (np.abs(d[:, 2]) > EPS) & ((size[0] <= 0) | (np.abs(x) <= size[0])) & ((size[1] <= 0) | (np.abs(y) <= size[1]))
            take(ok, t, np.tile([0.0, 0.0, 1.0], (P, 1)))
        elif gtype == SPHERE:
            b = (o * d).sum(1); c = (o * o).sum(1) - size[0] ** 2
            disc = b * b - c
            t = -b - np.sqrt(np.maximum(disc, 0))
            take(disc >= 0, t, (o + t[:, None] * d) / size[0])
        elif gtype in (CYLINDER, CAPSULE):
            r, h = size[0], size[1]
            a = d[:, 0] ** 2 + d[:, 1] ** 2; b = o[:, 0] * d[:, 0] + o[:, 1] * d[:, 1]; c = o[:, 0] ** 2 + o[:, 1] ** 2 - r * r
            disc = b * b - a * c
            t = (-b - np.sqrt(np.maximum(disc, 0))) / a
            z = o[:, 2] + t * d[:, 2]
            n = np.stack([(o[:, 0] + t * d[:, 0]) / r, (o[:, 1] + t * d[:, 1]) / r, np.zeros(P)], 1)
            take((a > EPS * EPS) & (disc >= 0) & (np.abs(z) <= h), t, n)
            for s in (-1.0, 1.0):
                if gtype == CYLINDER:
                    t = (s * h - o[:, 2]) / d[:, 2]
                    x, y = o[:, 0] + t * d[:, 0], o[:, 1] + t * d[:, 1]
                    take((np.abs(d[:, 2]) > EPS) & (x * x + y * y <= r * r), t, np.tile([0.0, 0.0, s], (P, 1)))
                else:
                    oc = o - np.array([0, 0, s * h])
                    bb = (oc * d).sum(1); cc = (oc * oc).sum(1) - r * r
                    disc = bb * bb - cc
                    t = -bb - np.sqrt(np.maximum(disc, 0))
                    z = oc[:, 2] + t * d[:, 2]
                    take((disc >= 0) & (z * s >= 0), t, (oc + t[:, None] * d) / r)
        elif gtype == BOX:
            tn = np.full(P, -1e30); tf = np.full(P, 1e30); ax = np.zeros(P, int); sg = np.ones(P)
            for k in range(3):
                par = np.abs(d[:, k]) < EPS
                tn = np.where(par & (np.abs(o[:, k]) > size[k]), 1e30, tn)
                t1 = (-size[k] - o[:, k]) / d[:, k]; t2 = (size[k] - o[:, k]) / d[:, k]
                lo, hi = np.minimum(t1, t2), np.maximum(t1, t2)
                upd = ~par & (lo > tn)
                tn = np.where(upd, lo, tn); ax = np.where(upd, k, ax); sg = np.where(upd, np.where(d[:, k] > 0, -1.0, 1.0), sg)
                tf = np.where(par, tf, np.minimum(tf, hi))
            n = np.zeros((P, 3)); n[np.arange(P), ax] = sg
            take(tn <= tf, tn, n)
    return hit


def render(cam_pos, cam_mat, fovy_deg, width, height, visuals):
    """visuals: list of (type, size[3], pos[3], mat[3,3], rgb[3]) resolved in the world frame.  Returns uint8 [H, W, 3]."""
    th = np.tan(0.5 * np.deg2rad(fovy_deg)); aspect = width / height
    cols, rows = np.meshgrid(np.arange(width), np.arange(height))
    dc = np.stack([(2 * (cols + 0.5) / width - 1) * th * aspect, (1 - 2 * (rows + 0.5) / height) * th, -np.ones_like(cols, float)], -1).reshape(-1, 3)
    dc /= np.linalg.norm(dc, axis=1, keepdims=True)
    d = dc @ np.asarray(cam_mat, float).T
    P = len(d)
    tbest = np.full(P, 1e30); nbest = np.zeros((P, 3)); cbest = np.zeros((P, 3)); any_hit = np.zeros(P, bool)
    for (gtype, size, pos, mat, rgb) in visuals:
        mat = np.asarray(mat, float)
        ol = (np.asarray(cam_pos, float) - np.asarray(pos, float)) @ mat
        dl = d @ mat
        nl = np.zeros((P, 3))
        h = _hit(gtype, np.asarray(size, float), np.tile(ol, (P, 1)), dl, tbest, nl)
        nbest[h] = nl[h] @ mat.T
        cbest[h] = np.minimum(np.asarray(rgb, float)[:3], 1.0)
        any_hit |= h
    shade = AMBIENT + DIFFUSE * np.abs((nbest * d).sum(1))
    c = np.where(any_hit[:, None], cbest * shade[:, None], 0.0)
    return np.floor(255.0 * np.minimum(c, 1.0) + 0.5).astype(np.uint8).reshape(height, width, 3)


def resolve_visuals(bundle, geom_xpos, geom_xmat, target, ovr, active):
    """world-frame visuals of one env from the bundle's MjsVisual list and that env's geom poses / anchors"""
    from mujoco_sim_b200.model.cstruct import MACROS

    out = []
    for v in bundle.visuals:
        pos = np.array(list(v.pos)); mat = quat_to_mat(list(v.quat))
        if v.anchor == MACROS["MJS_ANCHOR_GEOM"]:
            pos, mat = geom_xpos[v.geom], geom_xmat[v.geom].reshape(3, 3)
        elif v.anchor == MACROS["MJS_ANCHOR_TARGET"]:
            pos = pos + np.asarray(target, float)
        elif v.anchor == MACROS["MJS_ANCHOR_OVR"]:
            pos = pos + np.asarray(ovr, float)
        elif v.anchor == MACROS["MJS_ANCHOR_GEOMFRAME"]:
            G = geom_xmat[v.geom].reshape(3, 3)
            pos, mat = geom_xpos[v.geom] + G @ pos, G @ mat
        rgb = list(v.rgba_active)[:3] if (v.use_active and active) else list(v.rgba)[:3]
        out.append((v.type, list(v.size), pos, mat, rgb))
    return out


def resolve_camera(cam, geom_xpos, geom_xmat):
    from mujoco_sim_b200.model.cstruct import MACROS

    pos = np.array(list(cam.pos)); mat = quat_to_mat(list(cam.quat))
    if cam.anchor == MACROS["MJS_ANCHOR_GEOMFRAME"]:
        G = geom_xmat[cam.geom].reshape(3, 3)
        pos, mat = geom_xpos[cam.geom] + G @ pos, G @ mat
    return pos, mat
