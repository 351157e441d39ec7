"""Independent pins of the oracle's contact path (VERDICT r1 "parity first" item 1c / ADVICE r1 item 1).

Nothing here calls the oracle for the quantity under test: impedance, reference acceleration, pyramidal regularisation
(2 mu^2 R), diagApprox from body_invweight0 and the constrained minimiser are rebuilt in numpy from MuJoCo's documented
formulas (Computation chapter: "Constraint model", "Impedance", "Reference acceleration") and the
model constants of robot_planar_push.py / google_block.py:43-49, then compared with what oracle/mjs_oracle.cpp produces.
These pin the restatement to the published algorithm; they do not replace MuJoCo golden vectors (tests/test_reference_traces.py
takes those the day someone runs tools/dump_reference.py where mujoco + dm_control exist).
"""
import numpy as np

from mujoco_sim_b200.model import tasks as T
from tests.helpers import push_states

G = 9.81
MASS, RAD, HH = 0.1, 0.02, 0.019          # CylinderProp (google_block.py:31,43-49 on a primitive)
MU = np.array([1.0, 1.0, 0.05])           # friction [1, 0.05, 0] -> tangential, tangential, torsional (condim 4)
SOLREF = (0.02, 1.0)                      # MuJoCo defaults (the reference overrides neither)
SOLIMP = (0.9, 0.95, 0.001, 0.5, 2.0)
INVWEIGHT_TRAN = 1.0 / MASS               # body_invweight0[0] of a free body; the static plane contributes 0


def impedance(r, solimp=SOLIMP, margin=0.0):
    d0, dw, width, mid, power = solimp
    x = abs(r - margin) / width
    if x >= 1:
        return dw
    y = x**power / mid ** (power - 1) if x <= mid else 1 - (1 - x) ** power / (1 - mid) ** (power - 1)
    return d0 + y * (dw - d0)


def stiffness_damping(h, solref=SOLREF, solimp=SOLIMP):
    tc, dr = max(solref[0], 2 * h), solref[1]   # refsafe: time constant clamped to 2 h
    dmax = solimp[1]
    return 1.0 / (dmax**2 * tc**2 * dr**2), 2.0 / (dmax * tc)


def pyramid_D(r, impratio=1.0):
    """D = 1/R of every edge row of a condim-4 pyramidal contact at penetration r."""
    imp = impedance(r)
    diag = INVWEIGHT_TRAN * (1 + MU[0] ** 2)             # diagApprox of the first edge: tran + mu^2 tran
    R0 = max(1e-15, (1 - imp) / imp * diag)
    return 1.0 / (2 * MU[0] ** 2 / impratio * R0)


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)], [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])


def _settled(n_settle=2000, seed=0):
    b = T.build_planar_push()
    ora = push_states(b, 1, seed, settle=False)
    ora.finish_reset(n_settle)
    return b, ora


def test_flat_cylinder_rests_at_the_closed_form_penetration():
    """Static equilibrium of a flat cylinder on the plane: 3 contacts x 6 pyramid edges, each carrying D(r) * aref(r) with
    aref = -k d(r) r at rest, must sum to m g.  The root of that scalar equation is the resting penetration."""
    b, ora = _settled()
    h = b.model.timestep
    k, _ = stiffness_damping(h)

    def net(r):  # total normal force minus weight
        return 18 * pyramid_D(r) * k * impedance(r) * (-r) - MASS * G

    lo, hi = -1e-3, -1e-9
    assert net(lo) > 0 > net(hi)
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        lo, hi = (lo, mid) if net(mid) < 0 else (mid, hi)
    r_star = 0.5 * (lo + hi)
    s, t = ora.states[0], b.task
    assert max(abs(s.qvel[i]) for i in range(24)) < 1e-8
    for ob in range(t.n_obj):
        z = s.qpos[t.obj_qadr[ob] + 2]
        assert abs((z - HH) - r_star) < 1e-9, (z - HH, r_star)
    d = ora.stage_dump(0)
    assert d["ncon"] == 9 and d["nefc"] == 54
    np.testing.assert_allclose(d["con_dist"], r_star, atol=1e-9)
    np.testing.assert_allclose(d["efc_D"], pyramid_D(r_star), rtol=1e-6)
    np.testing.assert_allclose(d["efc_aref"], k * impedance(r_star) * (-r_star), rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(d["efc_force"], MASS * G / 18, rtol=1e-6)


def _prop_problem(d, s, qa, da, h, rows_of):
    """Rebuild the prop's constrained problem in numpy from the contact GEOMETRY only (dist, pos, frame of the dump)."""
    x = np.array([s.qpos[qa + i] for i in range(3)])
    R = quat_to_mat([s.qpos[qa + 3 + i] for i in range(4)])
    v = np.array([s.qvel[da + i] for i in range(6)])
    k, bdamp = stiffness_damping(h)
    J, aref, D = [], [], []
    for c in rows_of:
        r, p, F = d["con_dist"][c], d["con_pos"][c], d["con_frame"][c].reshape(3, 3)
        lever = p - x
        Jc = [np.concatenate([F[a], R.T @ np.cross(lever, F[a])]) for a in range(3)]   # normal, tangent 1, tangent 2
        Jc.append(np.concatenate([np.zeros(3), R.T @ F[0]]))                            # torsion about the normal
        imp = impedance(r)
        for a in range(3):
            for sgn in (1.0, -1.0):
                row = Jc[0] + sgn * MU[a] * Jc[1 + a]
                J.append(row); D.append(pyramid_D(r)); aref.append(-bdamp * (row @ v) - k * imp * r)
    return np.array(J), np.array(aref), np.array(D), v, R


def _solve_active_set(M, a0, J, aref, D):
    """argmin 1/2 (a-a0)^T M (a-a0) + sum 1/2 D min(0, J a - aref)^2 by active-set iteration; returns (a, force)."""
    a = a0.copy()
    active = (J @ a - aref) < 0
    for _ in range(100):
        Ja = J[active]
        H = M + Ja.T @ (D[active, None] * Ja)
        a = np.linalg.solve(H, M @ a0 + Ja.T @ (D[active] * aref[active]))
        new = (J @ a - aref) < 0
        if (new == active).all():
            break
        active = new
    else:
        raise AssertionError("active set did not settle")
    jar = J @ a - aref
    f = np.where(jar < 0, -D * jar, 0.0)
    np.testing.assert_allclose(M @ (a - a0), J.T @ f, atol=1e-10)  # stationarity of the numpy solution itself
    return a, f


def test_sliding_spinning_prop_matches_the_numpy_minimiser():
    """A settled prop given a tangential velocity and a spin: friction-pyramid edges switch on and off, the torsional rows
    (mu 0.05) load.  qacc and efc_force of the oracle must equal the minimiser of the problem rebuilt here."""
    b, ora = _settled()
    t, h = b.task, b.model.timestep
    s = ora.states[0]
    qa, da = t.obj_qadr[0], t.obj_dadr[0]
    for vel in ([0.3, 0.0, 0.0, 0.0, 0.0, 2.0], [0.05, -0.2, 0.0, 0.0, 0.0, -6.0], [0.0, 0.0, 0.0, 0.0, 0.0, 0.5], [0.4, 0.4, 0.0, 0.3, -0.2, 0.0]):
        for i in range(6):
            s.qvel[da + i] = vel[i]
        d = ora.stage_dump(0)
        x0 = np.array([s.qpos[qa], s.qpos[qa + 1]])
        mine = [c for c in range(d["ncon"]) if np.hypot(*(d["con_pos"][c][:2] - x0)) < 1.5 * RAD]
        assert len(mine) == 3
        J, aref, D, v, R = _prop_problem(d, s, qa, da, h, mine)
        I = np.array([MASS * (3 * RAD**2 + (2 * HH) ** 2) / 12] * 2 + [MASS * RAD**2 / 2])
        M = np.diag([MASS] * 3 + list(I))
        w = v[3:]
        a0 = np.concatenate([[0, 0, -G], -np.cross(w, I * w) / I])
        a, f = _solve_active_set(M, a0, J, aref, D)
        np.testing.assert_allclose(d["qacc"][da:da + 6], a, rtol=1e-7, atol=1e-8)
        rows = [r for c in mine for r in range(6 * c, 6 * c + 6)]  # the oracle lists contacts in order, 6 edge rows each, no limits active
        np.testing.assert_allclose(d["efc_J"][rows][:, da:da + 6], J, atol=1e-12)
        np.testing.assert_allclose(d["efc_aref"][rows], aref, rtol=1e-9, atol=1e-10)
        np.testing.assert_allclose(d["efc_D"][rows], D, rtol=1e-9)
        np.testing.assert_allclose(d["efc_force"][rows], f, rtol=1e-6, atol=1e-8)
        if abs(vel[0]) + abs(vel[1]) > 0:   # Coulomb sanity: friction opposes the slide and stays inside the cone
            ft = MASS * a[:2]
            assert ft @ np.array(vel[:2]) < 0 and np.linalg.norm(ft) <= MU[0] * MASS * (a[2] + G) * np.sqrt(2) + 1e-9


def test_free_tumbling_prop_follows_the_numpy_integrator_and_keeps_its_energy():
    """A prop thrown upward with a spin about a non-principal axis: no contact for 30 steps.  Semi-implicit Euler with the
    explicit gyroscopic term and the quaternion exponential map, transcribed here, must reproduce the oracle to rounding; the
    rotational energy and the world-frame angular momentum of the torque-free body may only drift at O(h)."""
    b = T.build_planar_push()
    ora = push_states(b, 1, 4, settle=False)
    t, h = b.task, b.model.timestep
    s = ora.states[0]
    qa, da = t.obj_qadr[1], t.obj_dadr[1]
    s.qpos[qa + 2] = 0.15
    v0 = np.array([0.1, -0.05, 1.0, 7.0, -3.0, 5.0])
    for i in range(6):
        s.qvel[da + i] = v0[i]
    I = np.array([MASS * (3 * RAD**2 + (2 * HH) ** 2) / 12] * 2 + [MASS * RAD**2 / 2])
    x = np.array([s.qpos[qa + i] for i in range(3)]); q = np.array([s.qpos[qa + 3 + i] for i in range(4)]); v = v0.copy()
    E0 = 0.5 * (I * v[3:] ** 2).sum(); L0 = quat_to_mat(q) @ (I * v[3:])
    n = 30
    for _ in range(n):
        w = v[3:]
        v = v + h * np.concatenate([[0, 0, -G], -np.cross(w, I * w) / I])
        x = x + h * v[:3]
        w = v[3:]; nw = np.linalg.norm(w); ang = nw * h
        dq = np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * w / nw])
        q = np.array([q[0]*dq[0]-q[1]*dq[1]-q[2]*dq[2]-q[3]*dq[3], q[0]*dq[1]+q[1]*dq[0]+q[2]*dq[3]-q[3]*dq[2],
                      q[0]*dq[2]-q[1]*dq[3]+q[2]*dq[0]+q[3]*dq[1], q[0]*dq[3]+q[1]*dq[2]-q[2]*dq[1]+q[3]*dq[0]])
        q /= np.linalg.norm(q)
    ora.finish_reset(n)
    np.testing.assert_allclose([s.qpos[qa + i] for i in range(3)], x, atol=1e-12)
    np.testing.assert_allclose([s.qpos[qa + 3 + i] for i in range(4)], q, atol=1e-12)
    np.testing.assert_allclose([s.qvel[da + i] for i in range(6)], v, atol=1e-11)
    w = np.array([s.qvel[da + 3 + i] for i in range(3)])
    E1 = 0.5 * (I * w**2).sum(); L1 = quat_to_mat([s.qpos[qa + 3 + i] for i in range(4)]) @ (I * w)
    assert abs(E1 - E0) / E0 < 2e-3, (E0, E1)          # bound with margin over 0.15 s at |w| = 9 rad/s
    assert np.linalg.norm(L1 - L0) / np.linalg.norm(L0) < 5e-3


def test_bias_force_matches_a_finite_difference_lagrangian():
    """qfrc_bias comes out of recursive Newton-Euler, M out of the composite-rigid-body pass: two different algorithms.  Lagrange's
    equations tie them: bias = C(q, v) v + dV/dq with (C v)_i = sum_jk (dM_ij/dq_k - 1/2 dM_jk/dq_i) v_j v_k.  Gravity
    compensation (robot.py:81-82: gravcomp = 1 on the UR5e's own bodies) appears in qfrc_passive = +dV_comp/dq."""
    b = T.build_robot_reach()
    from oracle.oracle import OracleEnvs

    ora = OracleEnvs(b, 1)
    m, s = b.model, ora.states[0]
    rng = np.random.RandomState(5)

    def dump(qq, vv):
        for i in range(6):
            s.qpos[i] = qq[i]; s.qvel[i] = vv[i]; s.ctrl[i] = qq[i]
        return ora.stage_dump(0)

    def potential(dd, comp_only):
        V = 0.0
        for body in range(1, m.nbody):
            if comp_only and m.body_gravcomp[body] == 0:
                continue
            com = dd["xpos"][body] + quat_to_mat(dd["xquat"][body]) @ np.array([m.body_ipos[3 * body + i] for i in range(3)])
            V += m.body_mass[body] * G * com[2] * (m.body_gravcomp[body] if comp_only else 1.0)
        return V

    eps = 1e-6
    for trial in range(4):
        q = rng.uniform(-1.5, 1.5, 6); v = rng.uniform(-1.0, 1.0, 6)
        d0 = dump(q, v)
        dM, dV, dVc = np.zeros((6, 6, 6)), np.zeros(6), np.zeros(6)
        for k in range(6):
            e = np.zeros(6); e[k] = eps
            dp, dm_ = dump(q + e, v), dump(q - e, v)
            dM[:, :, k] = (dp["M"] - dm_["M"]) / (2 * eps)
            dV[k] = (potential(dp, False) - potential(dm_, False)) / (2 * eps)
            dVc[k] = (potential(dp, True) - potential(dm_, True)) / (2 * eps)
        Cv = np.einsum("ijk,j,k->i", dM, v, v) - 0.5 * np.einsum("jki,j,k->i", dM, v, v)
        np.testing.assert_allclose(d0["qfrc_bias"], Cv + dV, atol=2e-6, rtol=1e-6)
        np.testing.assert_allclose(d0["qfrc_passive"], dVc, atol=2e-6, rtol=1e-6)


def test_pusher_impact_conserves_momentum_between_pusher_and_prop():
    """Contact forces are internal: what the struck prop gains in linear momentum over an impact, the rest of the system
    (arm through the pusher contact + ground friction on the prop) must account for.  Checked on the prop alone:
    m dv = h * sum of the contact forces acting on it, rebuilt from efc_force and the numpy Jacobian rows."""
    b, ora = _settled(400, seed=7)
    t, h = b.task, b.model.timestep
    s = ora.states[0]
    # drive the pusher through prop 0: command a TCP target on the far side of it
    tgt = np.array([[s.qpos[t.obj_qadr[0]], s.qpos[t.obj_qadr[0] + 1]]])
    for _ in range(6):  # the position-controlled arm closes ~20 % of the remaining distance per control step
        ora.step(tgt)
    da = t.obj_dadr[0]
    seen = 0
    from oracle.oracle import lib
    import ctypes
    for _ in range(60):
        d = ora.stage_dump(0)
        v0 = np.array([s.qvel[da + i] for i in range(3)])
        F = np.zeros(3)
        touching = False
        row = 0
        # rows: the oracle lists joint-limit rows first only when active (none here), then contacts in order
        for c in range(d["ncon"]):
            nrow = 6
            Jc = d["efc_J"][row:row + nrow, da:da + 3]
            if np.abs(Jc).max() > 0:
                F += Jc.T @ d["efc_force"][row:row + nrow]
                if abs(d["con_frame"][c][2]) < 0.5:
                    touching = True
            row += nrow
        assert row == d["nefc"]
        lib().oracle_substeps(ctypes.byref(b.model), ctypes.byref(t), ora.states, 1, 1, ctypes.c_double(1e-4))
        v1 = np.array([s.qvel[da + i] for i in range(3)])
        np.testing.assert_allclose(MASS * (v1 - v0), h * (F + np.array([0, 0, -MASS * G])), atol=1e-9)
        seen += touching
    assert seen > 3  # the pusher really hit the prop


def test_joint_limit_row_matches_the_formulas_written_out():
    """A limit row is the simplest constraint row MuJoCo builds: pos = q - lo (or hi - q), J = +-e_j, impedance d(|pos|) from
    solimp, R = (1 - d) / d * dof_invweight0, D = 1 / R, aref = -b (J v) - k d pos with k = 1 / (dmax^2 tc^2 dr^2), b = 2 / (dmax tc)
    (mj_makeImpedance / mj_referenceConstraint as documented).  Numbers recomputed here from the model constants only."""
    from mujoco_sim_b200.model import tasks as T
    from tests.helpers import arm_states

    b = T.build_robot_reach()
    m = b.model
    ora = arm_states(b, 1, 5)
    s = ora.states[0]
    j = 2                                   # elbow: range +-3.1415
    lo, hi = m.jnt_range[2 * j], m.jnt_range[2 * j + 1]
    for side, q, v in ((+1, hi + 0.013, 0.4), (-1, lo - 0.0004, -0.7)):   # past the upper / lower limit, moving further out
        s.qpos[j] = q; s.qvel[j] = v; s.ctrl[j] = q   # ctrl = q: the position actuator does not pull the joint back by itself
        d = ora.stage_dump(0)
        rows = [r for r in range(d["nefc"]) if abs(d["efc_J"][r][j]) == 1.0 and np.count_nonzero(d["efc_J"][r]) == 1]
        assert len(rows) == 1
        r = rows[0]
        J = -1.0 if side > 0 else 1.0       # pos = hi - q  ->  J = -1;  pos = q - lo  ->  J = +1
        assert d["efc_J"][r][j] == J
        pos = (hi - q) if side > 0 else (q - lo)
        d0, dw, width, mid, power = [m.jnt_solimp[5 * j + k] for k in range(5)]
        tc, dr = m.jnt_solref[2 * j], m.jnt_solref[2 * j + 1]
        x = min(abs(pos) / width, 1.0)
        y = x ** power / mid ** (power - 1) if x <= mid else 1 - (1 - x) ** power / (1 - mid) ** (power - 1)
        imp = d0 + y * (dw - d0)
        k_ = 1.0 / (dw * dw * tc * tc * dr * dr); b_ = 2.0 / (dw * tc)
        aref = -b_ * (J * v) - k_ * imp * pos
        D = 1.0 / ((1 - imp) / imp * m.dof_invweight0[j])
        np.testing.assert_allclose(d["efc_aref"][r], aref, rtol=1e-12)
        np.testing.assert_allclose(d["efc_D"][r], D, rtol=1e-12)
        # (the UR5e's position actuators clamp ctrl to the same range and pull a joint that is past its limit back harder than
        # aref asks for, so J a >= aref already at the smooth acceleration and the row carries no force: nothing more to check)
        assert d["efc_force"][r] == 0 and J * d["qacc"][j] >= d["efc_aref"][r]
