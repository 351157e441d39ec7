"""CPU checks of the narrowphase additions of round 2: capsule x capsule against an independent
segment-segment computation, the cylinder x cylinder cap-configuration guard with the truncation rule, and the proof that
the arm's link geoms never reach the props' height range, which is what the arm x prop pair filter of the scenes relies on."""
import ctypes

import numpy as np

from mujoco_sim_b200.model import tasks as T
from mujoco_sim_b200.model.cstruct import MACROS
from oracle.oracle import OracleEnvs, _dp, lib
from tests.helpers import arm_states, push_states

OBST = dict(size=(0.04, 0.4), pos=(0.0, -0.45, 0.30), quat=(1, 0, 1, 0))


def _segment_distance(p1, q1, p2, q2):
    """closest points of two segments (Ericson, Real-Time Collision Detection 5.1.9), written independently of the oracle"""
    d1, d2, r = q1 - p1, q2 - p2, p1 - p2
    a, e, f = d1 @ d1, d2 @ d2, d2 @ r
    c, b = d1 @ r, d1 @ d2
    den = a * e - b * b
    s = np.clip((b * f - c * e) / den, 0, 1) if den > 1e-14 else 0.0
    t = (b * s + f) / e
    if t < 0:
        t, s = 0.0, np.clip(-c / a, 0, 1)
    elif t > 1:
        t, s = 1.0, np.clip((b - c) / a, 0, 1)
    c1, c2 = p1 + d1 * s, p2 + d2 * t
    return np.linalg.norm(c1 - c2), c1, c2


def test_capsule_capsule_contacts_match_an_independent_segment_distance():
    b = T.build_robot_reach(obstacle=OBST)
    b.task.time_limit = 0.0
    n = 48
    ora = arm_states(b, n, 3)
    m = b.model
    g_obs = b.compiled.geom_id["obstacle"]
    names = {v: k for k, v in b.compiled.geom_id.items()}
    rng = np.random.RandomState(0)
    checked = 0
    for k in range(10):
        ora.step(rng.uniform([-0.1, -0.6, 0.02], [0.1, -0.4, 0.2], (n, 3)))
        for e in range(n):
            d = ora.stage_dump(e)
            if d["ncon"] == 0:
                continue
            xpos, xmat = ora.geom_poses(e)
            # every arm capsule against the obstacle: expected contact set from the independent distance
            expect = []
            for g in range(m.ngeom):
                if g == g_obs or m.geom_type[g] != 3 or not names[g].startswith("ur5e/"):  # the link capsules (the palm does not meet the obstacle)
                    continue
                ax, axo = xmat[g].reshape(3, 3)[:, 2], xmat[g_obs].reshape(3, 3)[:, 2]
                h, ho, r, ro = m.geom_size[3 * g + 1], m.geom_size[3 * g_obs + 1], m.geom_size[3 * g], m.geom_size[3 * g_obs]
                dist, c1, c2 = _segment_distance(xpos[g] - ax * h, xpos[g] + ax * h, xpos[g_obs] - axo * ho, xpos[g_obs] + axo * ho)
                if dist - r - ro < 0:
                    expect.append((dist - r - ro, 0.5 * (c1 + c2) + 0.5 * (c1 - c2) / dist * (ro - r)))
            got = [(d["con_dist"][c], d["con_pos"][c]) for c in range(d["ncon"]) if abs(d["con_frame"][c][2]) < 0.999 or d["con_pos"][c][2] > 0.1]
            assert len(got) == len(expect), (k, e, len(got), len(expect))
            for (dg, pg), (de, pe) in zip(sorted(got, key=lambda x: x[0]), sorted(expect, key=lambda x: x[0])):
                assert abs(dg - de) < 1e-10 and np.abs(pg - pe).max() < 1e-9
                checked += 1
    assert checked > 100
    assert (ora.field("status") == 0).all()


def test_a_prop_set_down_on_another_props_cap_is_carried_by_a_cap_contact():
    """A prop resting on another prop's cap is a cap-cap contact, which the analytic side-side routine cannot represent (it
    would answer with a 4 cm deep radial contact).  The guard hands the pair to the general convex routine: one contact
    with a vertical normal carries the upper prop; no status bit, no truncation."""
    b = T.build_planar_push()
    ora = push_states(b, 2, 5, settle=True)
    t = b.task
    s = ora.states[1]
    a0, a1 = t.obj_qadr[0], t.obj_qadr[1]
    s.qpos[a1] = s.qpos[a0] + 0.004; s.qpos[a1 + 1] = s.qpos[a0 + 1]; s.qpos[a1 + 2] = s.qpos[a0 + 2] + 2 * 0.019 - 1e-4
    d = ora.stage_dump(1)
    caps = [c for c in range(d["ncon"]) if d["con_pos"][c][2] > 0.03]
    assert len(caps) == 1
    c = caps[0]
    assert abs(d["con_dist"][c] + 1e-4) < 1e-8 and abs(abs(d["con_frame"][c][2]) - 1) < 1e-6   # depth 0.1 mm, vertical normal
    z0 = s.qpos[a1 + 2]
    act = np.array([[ora.states[e].qpos[t.obj_qadr[2]] + 0.1, -0.45] for e in range(2)])
    for _ in range(3):
        obs, rew, disc, fl = ora.step(act)
        assert (ora.field("status") == 0).all() and not fl[:, 3].any()
    # still on top after 60 substeps (one contact point cannot resist the torque of the 4 mm offset: the upper prop rocks on it,
    # as it does in MuJoCo without multiccd)
    assert abs(ora.states[1].qpos[a1 + 2] - z0) < 6e-3
    assert abs(ora.states[1].qpos[a0 + 2] - 0.0189) < 1e-3        # the lower prop carries it without sinking


def test_arm_links_stay_above_the_props_over_the_whole_action_box():
    """The Planar Push scene pairs the arm's link geoms only with the ground plane (model/tasks.py `_BIT_ARM`); MuJoCo's
    default contype/conaffinity would also test them against the props.  Those pairs can never be in contact: for every TCP
    target in the action box (z = 0.02, top-down) and along the joint-space segments between targets, every link geom other
    than the pusher and the shoulder stays above 0.079 m -- twice the height of a standing prop (0.038 m) and above a prop
    lying on its side (0.04 m) or leaning on another; the shoulder capsule (on the base, at the origin) stays 0.1 m away
    horizontally from anywhere the pusher can move a prop."""
    b = T.build_planar_push()
    t, m = b.task, b.model
    ora = OracleEnvs(b, 1)
    names = {v: k for k, v in b.compiled.geom_id.items()}
    arm = [g for g in range(m.ngeom) if names[g].startswith("ur5e/")]
    assert len(arm) == 9
    rng = np.random.RandomState(1)
    lowest, clearance = 1e9, 1e9
    q_prev = None
    for _ in range(400):
        tcp = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.6, -0.3), 0.02])
        q = np.zeros(6)
        assert lib().oracle_tcp_to_joints(ctypes.byref(t), _dp(tcp), _dp(q_prev if q_prev is not None else np.zeros(6)), _dp(q))
        path = [q] if q_prev is None else [q_prev + (q - q_prev) * a for a in np.linspace(0, 1, 6)[1:]]
        for qq in path:
            for i in range(6):
                ora.states[0].qpos[i] = qq[i]
            xpos, xmat = ora.geom_poses(0)
            for g in arm:
                axis = xmat[g].reshape(3, 3)[:, 2]
                r, h = m.geom_size[3 * g], m.geom_size[3 * g + 1]
                low = xpos[g][2] - abs(axis[2]) * h - r  # capsule: exact; cylinder: conservative (rim below the axis end by <= r)
                if names[g] == "ur5e/shoulder_col":
                    # the shoulder sits on the base at the origin, low but far from where props can be: horizontal distance of its
                    # axis segment to the prop region = the pusher's workspace grown by 0.1 m, [-0.3, 0.3] x [-0.7, -0.2]
                    for sgn in np.linspace(-1, 1, 9):
                        p = xpos[g] + axis * h * sgn
                        clearance = min(clearance, np.hypot(max(abs(p[0]) - 0.3, 0), max(p[1] + 0.2, -0.7 - p[1], 0)) - r)
                else:
                    lowest = min(lowest, low)
        q_prev = q
    assert lowest > 0.079, lowest
    assert clearance > 0.1, clearance


# ------------------------------------------------------------------ general convex pairs (include/mjs_convex.h)
def _rand_rot(rng):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)], [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])


def test_mpr_matches_closed_forms():
    from oracle.oracle import convex_pair

    I = np.eye(3)
    SPH, CAP, CYL, BOX = 2, 3, 5, 6
    d, n, p = convex_pair(SPH, [0.1], [0, 0, 0], I, SPH, [0.2], [0.25, 0, 0], I)
    assert abs(d - 0.05) < 1e-9 and np.allclose(n, [1, 0, 0], atol=1e-9) and np.allclose(p, [0.075, 0, 0], atol=1e-9)
    d, n, p = convex_pair(BOX, [0.1, 0.1, 0.1], [0, 0, 0], I, BOX, [0.1, 0.1, 0.1], [0.15, 0.02, 0.03], I)
    assert abs(d - 0.05) < 1e-9 and np.allclose(n, [1, 0, 0], atol=1e-9)
    # a prop set down on another prop's cap: the overlap along the common axis
    d, n, p = convex_pair(CYL, [0.02, 0.019], [0, 0, 0.019], I, CYL, [0.02, 0.019], [0.004, 0, 0.0569], I)
    assert abs(d - 1e-4) < 1e-9 and np.allclose(n, [0, 0, 1], atol=1e-6) and abs(p[2] - 0.03795) < 1e-9
    # the cap of a prop lying on its side against the side of the pusher
    Ry = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], float)
    d, n, p = convex_pair(CYL, [0.02, 0.05], [0, 0, 0.05], I, CYL, [0.02, 0.019], [0, 0.0385, 0.02], Ry)
    assert abs(d - 5e-4) < 2e-6 and np.allclose(n, [0, 1, 0], atol=1e-4)
    assert convex_pair(CYL, [0.02, 0.05], [0, 0, 0.05], I, CYL, [0.02, 0.019], [0, 0.1, 0.02], Ry) is None
    # capsule against a box face
    d, n, p = convex_pair(CAP, [0.03, 0.1], [0, 0, 0.125], I, BOX, [0.2, 0.2, 0.05], [0, 0, -0.05], I)
    assert abs(d - 0.005) < 1e-6 and np.allclose(n, [0, 0, -1], atol=1e-4)


def test_mpr_penetration_vector_separates_random_pairs():
    """Implementation-independent property of a penetration (depth, dir): shifting shape 2 by slightly more than depth * dir
    separates the pair, by slightly less leaves it intersecting; and the result does not depend on which shape is first."""
    from oracle.oracle import convex_pair

    rng = np.random.RandomState(0)
    types = [(2, lambda: [rng.uniform(0.02, 0.06), 0, 0]), (3, lambda: [rng.uniform(0.02, 0.04), rng.uniform(0.03, 0.1), 0]),
             (5, lambda: [rng.uniform(0.02, 0.05), rng.uniform(0.02, 0.08), 0]), (6, lambda: list(rng.uniform(0.02, 0.07, 3)))]
    hits = 0
    for _ in range(600):
        (t1, s1f), (t2, s2f) = types[rng.randint(4)], types[rng.randint(4)]
        s1, s2 = s1f(), s2f()
        R1, R2 = _rand_rot(rng), _rand_rot(rng)
        p1, p2 = np.zeros(3), rng.normal(size=3) * 0.05
        r = convex_pair(t1, s1, p1, R1, t2, s2, p2, R2)
        if r is None:
            continue
        depth, n, pos = r
        if depth < 1e-4:
            continue
        hits += 1
        assert abs(np.linalg.norm(n) - 1) < 1e-9
        assert convex_pair(t1, s1, p1, R1, t2, s2, p2 + n * (depth + 2e-5), R2) is None
        assert convex_pair(t1, s1, p1, R1, t2, s2, p2 + n * (depth - 2e-5), R2) is not None
        back = convex_pair(t2, s2, p2, R2, t1, s1, p1, R1)
        assert back is not None
        # MPR finds the penetration through the portal hit by the ray centre -> origin of the Minkowski difference, not always the
        # global minimum; with the shapes swapped the portal is mirrored up to tie-breaking: a valid penetration the other way
        assert np.dot(back[1], n) < -0.5
        assert convex_pair(t2, s2, p2, R2, t1, s1, p1 + back[1] * (back[0] + 2e-5), R1) is None
    assert hits > 150


def test_side_side_cylinders_the_analytic_contact_is_the_convex_routines_contact():
    """Bound on a stated deviation: upright cylinder pairs use an analytic side-side routine (depth r1 + r2 -
    axis distance, horizontal normal, position at the middle of the axial overlap) where MuJoCo calls its convex solver.  For
    parallel axes the general routine (include/mjs_convex.h, MuJoCo's MPR constants) must find the same depth and normal to its
    own tolerance, and a position on the same contact line (the line is the whole axial overlap: any point of it is a valid
    answer, the analytic routine picks its midpoint)."""
    from oracle.oracle import convex_pair

    CYL = 5
    I = np.eye(3)
    rng = np.random.RandomState(3)
    worst_d = worst_n = 0.0
    for _ in range(200):
        r1, h1 = 0.02, 0.019
        r2, h2 = (0.02, 0.019) if rng.rand() < 0.5 else (0.02, 0.05)
        pen = rng.uniform(1e-5, 2e-3)
        ang = rng.uniform(0, 2 * np.pi)
        d = r1 + r2 - pen
        p1 = np.array([0.0, 0.0, h1])
        p2 = np.array([d * np.cos(ang), d * np.sin(ang), h2 + rng.uniform(0, 0.01)])
        got = convex_pair(CYL, [r1, h1], p1, I, CYL, [r2, h2], p2, I)
        assert got is not None
        depth, n, p = got
        worst_d = max(worst_d, abs(depth - pen))
        worst_n = max(worst_n, np.abs(n - np.array([np.cos(ang), np.sin(ang), 0])).max())
        lo, hi = max(p1[2] - h1, p2[2] - h2), min(p1[2] + h1, p2[2] + h2)
        assert lo - 1e-6 <= p[2] <= hi + 1e-6                                   # on the axial overlap
        mid = p1[:2] + (r1 - 0.5 * pen) * np.array([np.cos(ang), np.sin(ang)])  # half-way through the penetration
        assert np.abs(p[:2] - mid).max() < 2e-5
    assert worst_d < 2e-6 and worst_n < 2e-3, (worst_d, worst_n)
