"""Model flattener: Scene -> MjsModel (include/mjs_model.h).

Restates, for the feature subset of scene.py, what MuJoCo's model compiler
derives before simulation [UPSTREAM-RECALL, mujoco is not installed here]:
depth-first body numbering, inertia from geoms, bounding radii, static
collision-pair filtering, and the qpos0 statistics the constraint code needs
(dof_invweight0, body_invweight0, meaninertia).
"""
from __future__ import annotations

import numpy as np

from . import scene as S
from .cstruct import MACROS, MjsModel

MINVAL = 1e-15


# ---------------------------------------------------------------- quaternions
def qnorm(q):
    q = np.asarray(q, dtype=np.float64)
    return q / np.linalg.norm(q)


def qmul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array(
        [
            aw * bw - ax * bx - ay * by - az * bz,
            aw * bx + ax * bw + ay * bz - az * by,
            aw * by - ax * bz + ay * bw + az * bx,
            aw * bz + ax * by - ay * bx + az * bw,
        ]
    )


def qmat(q):
    w, x, y, z = q
    return np.array(
        [
            [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
            [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
            [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
        ]
    )


# ------------------------------------------------------------- geom inertias
def geom_volume_inertia(g: S.Geom):
    """(volume, principal inertia per unit mass) of a primitive about its centre."""
    s = np.asarray(g.size, dtype=np.float64)
    if g.type == S.SPHERE:
        r = s[0]
        return 4.0 / 3.0 * np.pi * r**3, np.full(3, 0.4 * r * r)
    if g.type == S.CYLINDER:
        r, h = s[0], s[1]
        return np.pi * r * r * 2 * h, np.array(
            [(3 * r * r + 4 * h * h) / 12.0, (3 * r * r + 4 * h * h) / 12.0, r * r / 2.0]
        )
    if g.type == S.BOX:
        return 8 * s[0] * s[1] * s[2], np.array(
            [(s[1] ** 2 + s[2] ** 2) / 3.0, (s[0] ** 2 + s[2] ** 2) / 3.0, (s[0] ** 2 + s[1] ** 2) / 3.0]
        )
    if g.type == S.CAPSULE:
        r, h = s[0], s[1]
        vc = np.pi * r * r * 2 * h
        vs = 4.0 / 3.0 * np.pi * r**3
        v = vc + vs
        # cylinder + two hemispheres (parallel-axis for the caps)
        izz = (vc * r * r / 2.0 + vs * 0.4 * r * r) / v
        ixx = (vc * (3 * r * r + 4 * h * h) / 12.0 + vs * (0.4 * r * r + h * h + 0.75 * r * h)) / v
        return v, np.array([ixx, ixx, izz])
    if g.type == S.PLANE:
        return 0.0, np.zeros(3)
    raise ValueError(g.type)


def geom_rbound(g: S.Geom):
    s = np.asarray(g.size, dtype=np.float64)
    if g.type == S.SPHERE:
        return s[0]
    if g.type == S.CAPSULE:
        return s[0] + s[1]
    if g.type == S.CYLINDER:
        return float(np.hypot(s[0], s[1]))
    if g.type == S.BOX:
        return float(np.linalg.norm(s[:3]))
    return 0.0  # plane


class Compiled:
    """Result of compile_scene: the ctypes struct plus name -> id maps."""

    def __init__(self):
        self.model = MjsModel()
        self.body_id = {}
        self.jnt_id = {}
        self.geom_id = {}
        self.site_id = {}
        self.jnt_qadr = {}
        self.jnt_dadr = {}
        self.geom_rgba = []  # per compiled geom, for the ray caster (not part of the physics model)


def compile_scene(sc: S.Scene) -> Compiled:
    out = Compiled()
    m = out.model
    bodies, parents = [], []

    def walk(b, parent):
        idx = len(bodies)
        bodies.append(b)
        parents.append(parent)
        for c in b.children:
            walk(c, idx)

    walk(sc.world, 0)
    nb = len(bodies)
    assert nb <= MACROS["MJS_MAXBODY"], nb

    f = {k: [] for k in ()}  # noqa: F841
    body_pos, body_quat = [], []
    body_ipos, body_iquat, body_mass, body_inertia, body_gravcomp = [], [], [], [], []
    body_jntadr, body_jntnum, body_dofadr, body_dofnum, body_mocapid, body_weldid = [], [], [], [], [], []
    jnt = dict(type=[], body=[], qadr=[], dadr=[], limited=[], pos=[], axis=[], range=[], margin=[], solref=[], solimp=[])
    dof = dict(body=[], parent=[], jnt=[], armature=[], damping=[])
    geoms, geom_body = [], []
    sites, site_body = [], []
    qpos0 = []
    nq = nv = nmocap = 0
    last_dof_of_body = {}

    for i, b in enumerate(bodies):
        out.body_id[b.name] = i
        p = parents[i]
        body_pos.append(S._v(b.pos, 3))
        body_quat.append(qnorm(S._v(b.quat, 4)))
        body_gravcomp.append(b.gravcomp)
        body_jntadr.append(len(jnt["type"]) if b.joints else -1)
        body_jntnum.append(len(b.joints))
        body_dofadr.append(nv if b.joints else -1)
        if b.mocap:
            assert p == 0 and not b.joints
            body_mocapid.append(nmocap)
            nmocap += 1
        else:
            body_mocapid.append(-1)
        body_weldid.append(i if (b.joints or i == 0) else body_weldid[p])
        # last dof of the nearest ancestor with dofs
        pdof = last_dof_of_body.get(p, -1) if i > 0 else -1
        nd0 = nv
        for j in b.joints:
            jid = len(jnt["type"])
            out.jnt_id[j.name] = jid
            out.jnt_qadr[j.name] = nq
            out.jnt_dadr[j.name] = nv
            jnt["type"].append(j.type)
            jnt["body"].append(i)
            jnt["qadr"].append(nq)
            jnt["dadr"].append(nv)
            jnt["limited"].append(1 if (j.range is not None and j.type != S.FREE) else 0)
            jnt["pos"].append(S._v(j.pos, 3))
            ax = S._v(j.axis, 3)
            jnt["axis"].append(ax / np.linalg.norm(ax))
            jnt["range"].append(S._v(j.range, 2) if j.range is not None else np.zeros(2))
            jnt["margin"].append(j.margin)
            jnt["solref"].append(S._v(j.solref, 2))
            jnt["solimp"].append(S._v(j.solimp, 5))
            if j.type == S.FREE:
                # qpos0 of a free joint is the body's own pose
                qpos0.extend(list(S._v(b.pos, 3)) + list(qnorm(S._v(b.quat, 4))))
                nq += 7
                nd = 6
            else:
                qpos0.append(0.0)
                nq += 1
                nd = 1
            for _ in range(nd):
                dof["body"].append(i)
                dof["parent"].append(pdof)
                dof["jnt"].append(jid)
                dof["armature"].append(j.armature)
                dof["damping"].append(j.damping)
                pdof = nv
                nv += 1
        body_dofnum.append(nv - nd0)
        last_dof_of_body[i] = pdof
        for g in b.geoms:
            out.geom_id[g.name or f"geom{len(geoms)}"] = len(geoms)
            geoms.append(g)
            geom_body.append(i)
        for s in b.sites:
            out.site_id[s.name] = len(sites)
            sites.append(s)
            site_body.append(i)
        # inertial frame
        if b.inertial is not None:
            body_mass.append(b.inertial.mass)
            body_ipos.append(S._v(b.inertial.pos, 3))
            body_iquat.append(qnorm(S._v(b.inertial.quat, 4)))
            body_inertia.append(S._v(b.inertial.diaginertia, 3))
        else:
            massive = [g for g in b.geoms if g.type != S.PLANE and (g.mass is None or g.mass > 0)]
            if i == 0 or not massive or body_weldid[i] == 0:
                # static bodies: inertia never enters the dynamics
                body_mass.append(0.0)
                body_ipos.append(np.zeros(3))
                body_iquat.append(np.array([1.0, 0, 0, 0]))
                body_inertia.append(np.zeros(3))
            else:
                assert len(massive) == 1, f"body {b.name}: give an explicit <inertial> for multi-geom bodies"
                g = massive[0]
                vol, unit = geom_volume_inertia(g)
                mass = g.mass if g.mass is not None else g.density * vol
                body_mass.append(mass)
                body_ipos.append(S._v(g.pos, 3))
                body_iquat.append(qnorm(S._v(g.quat, 4)))
                body_inertia.append(unit * mass)

    assert nv <= MACROS["MJS_MAXV"] and nq <= MACROS["MJS_MAXQ"], (nv, nq)
    ng = len(geoms)
    out.geom_rgba = [tuple(float(x) for x in g.rgba) for g in geoms]
    assert ng <= MACROS["MJS_MAXGEOM"], ng

    # ------------------------------------------------------------ sizes/options
    m.set("nbody", nb)
    m.set("njnt", len(jnt["type"]))
    m.set("nq", nq)
    m.set("nv", nv)
    m.set("ngeom", ng)
    m.set("nsite", len(sites))
    m.set("neq", len(sc.welds))
    m.set("nu", len(sc.actuators))
    m.set("nmocap", nmocap)
    m.set("integrator", sc.integrator)
    m.set("solver_iterations", sc.iterations)
    m.set("timestep", sc.timestep)
    m.set("gravity", sc.gravity)
    m.set("impratio", sc.impratio)
    m.set("tolerance", sc.tolerance)

    m.set("body_parent", parents)
    m.set("body_jntadr", body_jntadr)
    m.set("body_jntnum", body_jntnum)
    m.set("body_dofadr", body_dofadr)
    m.set("body_dofnum", body_dofnum)
    m.set("body_mocapid", body_mocapid)
    m.set("body_weldid", body_weldid)
    m.set("body_pos", np.array(body_pos))
    m.set("body_quat", np.array(body_quat))
    m.set("body_ipos", np.array(body_ipos))
    m.set("body_iquat", np.array(body_iquat))
    m.set("body_mass", body_mass)
    m.set("body_inertia", np.array(body_inertia))
    m.set("body_gravcomp", body_gravcomp)

    if jnt["type"]:
        m.set("jnt_type", jnt["type"])
        m.set("jnt_body", jnt["body"])
        m.set("jnt_qposadr", jnt["qadr"])
        m.set("jnt_dofadr", jnt["dadr"])
        m.set("jnt_limited", jnt["limited"])
        m.set("jnt_pos", np.array(jnt["pos"]))
        m.set("jnt_axis", np.array(jnt["axis"]))
        m.set("jnt_range", np.array(jnt["range"]))
        m.set("jnt_margin", jnt["margin"])
        m.set("jnt_solref", np.array(jnt["solref"]))
        m.set("jnt_solimp", np.array(jnt["solimp"]))
        m.set("qpos0", qpos0)
        m.set("dof_body", dof["body"])
        m.set("dof_parent", dof["parent"])
        m.set("dof_jnt", dof["jnt"])
        m.set("dof_armature", dof["armature"])
        m.set("dof_damping", dof["damping"])

    if ng:
        m.set("geom_type", [g.type for g in geoms])
        m.set("geom_body", geom_body)
        m.set("geom_condim", [g.condim for g in geoms])
        m.set("geom_priority", [g.priority for g in geoms])
        m.set("geom_pos", np.array([S._v(g.pos, 3) for g in geoms]))
        m.set("geom_quat", np.array([qnorm(S._v(g.quat, 4)) for g in geoms]))
        m.set("geom_size", np.array([list(g.size) + [0.0] * (3 - len(g.size)) for g in geoms]))
        m.set("geom_friction", np.array([S._v(g.friction, 3) for g in geoms]))
        m.set("geom_solref", np.array([S._v(g.solref, 2) for g in geoms]))
        m.set("geom_solimp", np.array([S._v(g.solimp, 5) for g in geoms]))
        m.set("geom_solmix", [g.solmix for g in geoms])
        m.set("geom_margin", [g.margin for g in geoms])
        m.set("geom_gap", [g.gap for g in geoms])
        m.set("geom_rbound", [geom_rbound(g) for g in geoms])
    if sites:
        m.set("site_body", site_body)
        m.set("site_pos", np.array([S._v(s.pos, 3) for s in sites]))
        m.set("site_quat", np.array([qnorm(S._v(s.quat, 4)) for s in sites]))

    # ----------------------------------------------------------- FK at qpos0
    xpos = np.zeros((nb, 3))
    xquat = np.zeros((nb, 4))
    xquat[0] = [1, 0, 0, 0]
    for i in range(1, nb):
        p = parents[i]
        xpos[i] = xpos[p] + qmat(xquat[p]) @ body_pos[i]
        xquat[i] = qnorm(qmul(xquat[p], body_quat[i]))
    xmat = np.array([qmat(q) for q in xquat])
    xipos = np.array([xpos[i] + xmat[i] @ body_ipos[i] for i in range(nb)])

    # ------------------------------------------------ M, invweights at qpos0
    def jac(bid, point):
        jp = np.zeros((3, nv))
        jr = np.zeros((3, nv))
        b = bid
        while b > 0:
            for k in range(body_jntnum[b]):
                jid = body_jntadr[b] + k
                d0 = jnt["dadr"][jid]
                t = jnt["type"][jid]
                if t == S.FREE:
                    jp[:, d0 : d0 + 3] = np.eye(3)
                    R = xmat[b]
                    for a in range(3):
                        jr[:, d0 + 3 + a] = R[:, a]
                        jp[:, d0 + 3 + a] = np.cross(R[:, a], point - xpos[b])
                else:
                    ax = xmat[b] @ jnt["axis"][jid]
                    anchor = xpos[b] + xmat[b] @ jnt["pos"][jid]
                    if t == S.SLIDE:
                        jp[:, d0] = ax
                    else:
                        jr[:, d0] = ax
                        jp[:, d0] = np.cross(ax, point - anchor)
            b = parents[b]
        return jp, jr

    if nv:
        M = np.diag(np.array(dof["armature"], dtype=np.float64))
        for i in range(1, nb):
            if body_mass[i] <= 0:
                continue
            jp, jr = jac(i, xipos[i])
            Ri = qmat(qnorm(qmul(xquat[i], body_iquat[i])))
            Iw = Ri @ np.diag(body_inertia[i]) @ Ri.T
            M += body_mass[i] * jp.T @ jp + jr.T @ Iw @ jr
        Minv = np.linalg.inv(M)
        dinv = np.diag(Minv).copy()
        for jid, t in enumerate(jnt["type"]):
            if t == S.FREE:
                d0 = jnt["dadr"][jid]
                dinv[d0 : d0 + 3] = dinv[d0 : d0 + 3].mean()
                dinv[d0 + 3 : d0 + 6] = dinv[d0 + 3 : d0 + 6].mean()
        m.set("dof_invweight0", np.maximum(dinv, MINVAL))
        m.set("meaninertia", float(np.mean(np.diag(M))))
        biw = np.zeros((nb, 2))
        for i in range(1, nb):
            if body_weldid[i] == 0:
                continue
            jp, jr = jac(i, xipos[i])
            biw[i, 0] = max(np.trace(jp @ Minv @ jp.T) / 3.0, MINVAL)
            biw[i, 1] = max(np.trace(jr @ Minv @ jr.T) / 3.0, MINVAL)
        m.set("body_invweight0", biw)
        out.M0 = M
    else:
        m.set("meaninertia", 1.0)

    # ------------------------------------------------------------ equalities
    for e, w in enumerate(sc.welds):
        b1, b2 = out.body_id[w.body1], out.body_id[w.body2]
        m.eq_body1[e] = b1
        m.eq_body2[e] = b2
        # anchor (on body2) = 0; relpose = pose of body2 in body1 frame at qpos0
        R1 = xmat[b1]
        relpos = R1.T @ (xpos[b2] - xpos[b1])
        q1c = xquat[b1] * np.array([1, -1, -1, -1])
        relquat = qnorm(qmul(q1c, xquat[b2]))
        data = np.concatenate([np.zeros(3), relpos, relquat, [w.torquescale]])
        for k in range(11):
            m.eq_data[e * 11 + k] = data[k]
        for k in range(2):
            m.eq_solref[e * 2 + k] = w.solref[k]
        for k in range(5):
            m.eq_solimp[e * 5 + k] = w.solimp[k]

        def has_rot(b):
            while b > 0:
                for k in range(body_jntnum[b]):
                    if jnt["type"][body_jntadr[b] + k] in (S.FREE, S.HINGE):
                        return True
                b = parents[b]
            return False

        m.eq_rot_rows[e] = 1 if (has_rot(b1) or has_rot(b2)) else 0

    # -------------------------------------------------------------- actuators
    for u, a in enumerate(sc.actuators):
        m.act_dof[u] = out.jnt_dadr[a.joint]
        m.act_gear[u] = a.gear
        m.act_gain[u] = a.kp
        m.act_bias[u * 3 + 0] = 0.0
        m.act_bias[u * 3 + 1] = -a.kp
        m.act_bias[u * 3 + 2] = -a.kv
        m.act_ctrllimited[u] = 1 if a.ctrlrange is not None else 0
        m.act_forcelimited[u] = 1 if a.forcerange is not None else 0
        if a.ctrlrange is not None:
            m.act_ctrlrange[u * 2], m.act_ctrlrange[u * 2 + 1] = a.ctrlrange
        if a.forcerange is not None:
            m.act_forcerange[u * 2], m.act_forcerange[u * 2 + 1] = a.forcerange

    # ------------------------------------------------------- collision pairs
    pairs = []
    for a in range(ng):
        for b in range(a + 1, ng):
            ba, bb = geom_body[a], geom_body[b]
            wa, wb = body_weldid[ba], body_weldid[bb]
            if wa == wb:  # same (welded) body, incl. static-static
                continue
            ga, gb = geoms[a], geoms[b]
            if not ((ga.contype & gb.conaffinity) or (gb.contype & ga.conaffinity)):
                continue
            # filterparent: skip parent-child unless the parent is the world
            if wa != 0 and wb != 0:
                pa = body_weldid[parents[wa]]
                pb = body_weldid[parents[wb]]
                if pa == wb or pb == wa:
                    continue
            g1, g2 = (a, b) if ga.type <= gb.type else (b, a)
            pairs.append((g1, g2))
    assert len(pairs) <= MACROS["MJS_MAXPAIR"], len(pairs)
    m.set("npair", len(pairs))
    if pairs:
        m.set("pair_geom1", [p[0] for p in pairs])
        m.set("pair_geom2", [p[1] for p in pairs])

    m.set("ovr_body", out.body_id[sc.ovr_body] if sc.ovr_body else -1)
    m.set("touch_site", out.site_id[sc.touch_site] if sc.touch_site else -1)
    m.set("touch_size", sc.touch_size)
    return out
