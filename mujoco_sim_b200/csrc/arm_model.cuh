// arm_model.cuh -- device-side model of the "UR arm + free props + static geoms"
// scene family (Robot Reach, Planar Push, Button Push) and the host code that
// specialises the flat MjsModel into it.  The warp-per-env kernel
// (arm_kernel.cuh) assumes this structure; anything else is rejected loudly.
#pragma once
#include <cmath>
#include <cstring>

#include "mjs_common.cuh"

#define ARM_NLINK 6
#define ARM_MAXAB 8      // inertial bodies riding on the chain: 6 links + <=2 fixed attachments
#define ARM_MAXOBJ 4
#define ARM_MAXGEOM 16
#define ARM_MAXPAIR 32   // one lane per candidate pair
#define ARM_MAXC 16      // contacts kept per env per substep
#define ARM_MAXCB 8      // contacts per dof block

// body slots: 0 static (world frame), 1..6 links (fixed children share the link's slot), 9..12 objects, 13 overridable static body
#define SLOT_STATIC 0
#define SLOT_LINK0 1
#define SLOT_ATT0 7
#define SLOT_OBJ0 9
#define SLOT_OVR 13
#define ARM_NSLOT 14

template <typename real>
struct ArmModel {
  int nobj, nv, nq, nab, ngeom, npair, n_sub, kind;
  int reward_type, max_steps, action_dim, obs_dim, integrator, max_iter, has_ovr, touch_pair_mask;
  real timestep, tol, scale, impratio;
  real gravity[3];
  // chain
  real link_pos[ARM_NLINK][3], link_quat[ARM_NLINK][4], link_axis[ARM_NLINK][3];
  real armature[ARM_NLINK];
  int ab_link[ARM_MAXAB];
  real ab_com[ARM_MAXAB][3];                          // com in the LINK frame
  real ab_I[ARM_MAXAB][6];                            // inertia about the com, LINK frame (xx yy zz xy xz yz)
  real ab_mass[ARM_MAXAB], ab_gfac[ARM_MAXAB];        // gfac = 1 - gravcomp
  // actuators (arm joints only), evaluated in double
  double kp[ARM_NLINK], kv[ARM_NLINK], ctrl_lo[ARM_NLINK], ctrl_hi[ARM_NLINK], frc_lo[ARM_NLINK], frc_hi[ARM_NLINK];
  int ctrllimited[ARM_NLINK], forcelimited[ARM_NLINK];
  // joint limits
  real jnt_lo[ARM_NLINK], jnt_hi[ARM_NLINK], jnt_margin[ARM_NLINK], jnt_solref[ARM_NLINK][2], jnt_solimp[ARM_NLINK][5], jnt_diag[ARM_NLINK];
  real jnt_kb[ARM_NLINK][2], pair_kb[ARM_MAXPAIR][2];  // reference stiffness / damping (mjs_kb of solref, solimp, timestep): constants of the model
  int jnt_limited[ARM_NLINK];
  // objects (free bodies, com at the body origin, principal axes = body axes)
  real obj_mass[ARM_MAXOBJ], obj_I[ARM_MAXOBJ][3];
  int obj_qadr[ARM_MAXOBJ];
  // static override body (switch)
  real ovr_mat[9];
  // geoms
  int geom_type[ARM_MAXGEOM], geom_slot[ARM_MAXGEOM];
  real geom_pos[ARM_MAXGEOM][3], geom_mat[ARM_MAXGEOM][9], geom_size[ARM_MAXGEOM][3], geom_rbound[ARM_MAXGEOM];
  // pairs (mj_contactParam pre-mixed on the host)
  int pair_g1[ARM_MAXPAIR], pair_g2[ARM_MAXPAIR], pair_dim[ARM_MAXPAIR];
  real pair_mu[ARM_MAXPAIR][3];  // tangential, torsional, rolling
  real pair_solref[ARM_MAXPAIR][2], pair_solimp[ARM_MAXPAIR][5], pair_margin[ARM_MAXPAIR], pair_incl[ARM_MAXPAIR];
  real pair_tran[ARM_MAXPAIR], pair_rot[ARM_MAXPAIR];
  // sites (flange, touch) in their slot frames
  int tcp_slot, touch_slot;
  real tcp_pos[3], tcp_mat[9], tcp_offset[3], touch_pos[3], touch_mat[9], touch_size[2];
  // task constants
  double control_dt, time_limit, goal_threshold, target_radius, nearest_coef, tcp_z;
  double spawn_lo[3], spawn_hi[3], target_lo[3], target_hi[3], obj_lo[3], obj_hi[3], action_lo[8], action_hi[8];
  double obj_radius, ik_dh[6], button_goal[3], button_goal_radius, touch_lo, touch_hi;
  int settle_steps, button_disturbances;
  double disturb_prob;
};

// ------------------------------------------------------------------ host build
namespace armhost {
inline void quat2mat(double* m, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
inline void mulquat(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
inline void matvec(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2], z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
inline void matmul(double* r, const double* a, const double* b) {
  double t[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) t[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
  memcpy(r, t, sizeof(t));
}
}  // namespace armhost

// Fills `out`; returns 0 or -1 (mjs_set_error has the reason).
template <typename real>
int build_arm_model(const MjsModel& m, const MjsTask& t, int precision_f64, ArmModel<real>& out) {
  using namespace armhost;
  ArmModel<real>& A = out;
  memset(&A, 0, sizeof(A));
#define FAIL(...) do { mjs_set_error("arm kernel: " __VA_ARGS__); return -1; } while (0)
  int nb = m.nbody;
  // classify bodies
  int slot_of[MJS_MAXBODY];
  double s_pos[MJS_MAXBODY][3], s_mat[MJS_MAXBODY][9], s_quat[MJS_MAXBODY][4];  // pose in the slot's reference frame
  int link_body[ARM_NLINK], nlink = 0, natt = 0, nobj = 0;
  for (int b = 0; b < nb; b++) slot_of[b] = -1;
  slot_of[0] = SLOT_STATIC;
  for (int k = 0; k < 3; k++) s_pos[0][k] = 0;
  s_quat[0][0] = 1; s_quat[0][1] = s_quat[0][2] = s_quat[0][3] = 0;
  quat2mat(s_mat[0], s_quat[0]);
  int att_body[2] = {-1, -1};
  for (int b = 1; b < nb; b++) {
    int p = m.body_parent[b];
    const double* bp = m.body_pos + 3 * b; const double* bq = m.body_quat + 4 * b;
    if (m.body_mocapid[b] >= 0) FAIL("mocap bodies unsupported");
    if (m.body_jntnum[b] == 0) {
      // fixed to its parent: compose the pose into the parent's slot frame
      int ps = slot_of[p];
      if (ps < 0) FAIL("body %d: unsupported parent", b);
      double v[3];
      matvec(v, s_mat[p], bp);
      for (int k = 0; k < 3; k++) s_pos[b][k] = s_pos[p][k] + v[k];
      mulquat(s_quat[b], s_quat[p], bq);
      quat2mat(s_mat[b], s_quat[b]);
      if (ps == SLOT_STATIC || ps == SLOT_OVR) {
        if (b == m.ovr_body) {
          if (p != 0) FAIL("override body must be a child of the world");
          slot_of[b] = SLOT_OVR;
          for (int k = 0; k < 3; k++) s_pos[b][k] = 0;  // position comes from the env state
          memcpy(s_quat[b], bq, 4 * sizeof(double));
          quat2mat(s_mat[b], s_quat[b]);
          for (int k = 0; k < 9; k++) A.ovr_mat[k] = (real)s_mat[b][k];
          A.has_ovr = 1;
        } else slot_of[b] = ps;
      } else if (ps >= SLOT_LINK0 && ps < SLOT_LINK0 + ARM_NLINK) {
        if (m.body_mass[b] > 0) {
          if (natt >= 2) FAIL("more than 2 massive attachments");
          att_body[natt++] = b;
        }
        slot_of[b] = ps;  // geometry of fixed children lives in the link frame
      } else FAIL("body %d fixed to an object is unsupported", b);
      continue;
    }
    int j = m.body_jntadr[b];
    if (m.body_jntnum[b] != 1) FAIL("body %d: exactly one joint per body", b);
    if (m.jnt_type[j] == MJS_JNT_HINGE) {
      if (nlink >= ARM_NLINK) FAIL("more than 6 hinge links");
      int ps = slot_of[p];
      if (nlink == 0 ? (ps != SLOT_STATIC) : (p != link_body[nlink - 1] )) FAIL("links must form one serial chain");
      if (m.jnt_dofadr[j] != nlink || m.jnt_qposadr[j] != nlink) FAIL("arm dofs must come first");
      const double* jp = m.jnt_pos + 3 * j;
      if (jp[0] != 0 || jp[1] != 0 || jp[2] != 0 || m.qpos0[nlink] != 0) FAIL("hinge anchors must sit at the body origin, qpos0 = 0");
      // pose relative to the previous link (or the static frame for link 0)
      double rp[3], rq[4];
      if (nlink == 0) {
        double v[3];
        matvec(v, s_mat[p], bp);
        for (int k = 0; k < 3; k++) rp[k] = s_pos[p][k] + v[k];
        mulquat(rq, s_quat[p], bq);
      } else {
        memcpy(rp, bp, sizeof(rp)); memcpy(rq, bq, sizeof(rq));
      }
      for (int k = 0; k < 3; k++) { A.link_pos[nlink][k] = (real)rp[k]; A.link_axis[nlink][k] = (real)m.jnt_axis[3 * j + k]; }
      for (int k = 0; k < 4; k++) A.link_quat[nlink][k] = (real)rq[k];
      A.armature[nlink] = (real)m.dof_armature[nlink];
      if (m.dof_damping[nlink] != 0) FAIL("arm joint damping unsupported");
      A.jnt_limited[nlink] = m.jnt_limited[j];
      A.jnt_lo[nlink] = (real)m.jnt_range[2 * j]; A.jnt_hi[nlink] = (real)m.jnt_range[2 * j + 1];
      A.jnt_margin[nlink] = (real)m.jnt_margin[j];
      for (int k = 0; k < 2; k++) A.jnt_solref[nlink][k] = (real)m.jnt_solref[2 * j + k];
      for (int k = 0; k < 5; k++) A.jnt_solimp[nlink][k] = (real)m.jnt_solimp[5 * j + k];
      A.jnt_diag[nlink] = (real)m.dof_invweight0[nlink];
      link_body[nlink] = b;
      slot_of[b] = SLOT_LINK0 + nlink;
      for (int k = 0; k < 3; k++) s_pos[b][k] = 0;
      s_quat[b][0] = 1; s_quat[b][1] = s_quat[b][2] = s_quat[b][3] = 0;
      quat2mat(s_mat[b], s_quat[b]);
      nlink++;
    } else if (m.jnt_type[j] == MJS_JNT_FREE) {
      if (p != 0) FAIL("free bodies must be children of the world");
      if (nobj >= ARM_MAXOBJ) FAIL("more than %d free bodies", ARM_MAXOBJ);
      const double* ip = m.body_ipos + 3 * b; const double* iq = m.body_iquat + 4 * b;
      if (ip[0] != 0 || ip[1] != 0 || ip[2] != 0 || iq[0] != 1) FAIL("free bodies need com at origin and principal axes = body axes");
      if (m.jnt_dofadr[j] != ARM_NLINK + 6 * nobj || m.jnt_qposadr[j] != ARM_NLINK + 7 * nobj) FAIL("free bodies must follow the arm in dof order");
      if (m.body_gravcomp[b] != 0) FAIL("gravcomp on free bodies unsupported");
      for (int k = 0; k < 6; k++) if (m.dof_damping[m.jnt_dofadr[j] + k] != 0 || m.dof_armature[m.jnt_dofadr[j] + k] != 0) FAIL("free-joint damping/armature unsupported");
      A.obj_mass[nobj] = (real)m.body_mass[b];
      for (int k = 0; k < 3; k++) A.obj_I[nobj][k] = (real)m.body_inertia[3 * b + k];
      A.obj_qadr[nobj] = m.jnt_qposadr[j];
      slot_of[b] = SLOT_OBJ0 + nobj;
      for (int k = 0; k < 3; k++) s_pos[b][k] = 0;
      s_quat[b][0] = 1; s_quat[b][1] = s_quat[b][2] = s_quat[b][3] = 0;
      quat2mat(s_mat[b], s_quat[b]);
      nobj++;
    } else FAIL("joint type %d unsupported", m.jnt_type[j]);
  }
  if (nlink != ARM_NLINK) FAIL("expected a 6-joint arm, found %d hinge links", nlink);
  if (m.nv != ARM_NLINK + 6 * nobj) FAIL("dof count mismatch");
  if (m.nu != ARM_NLINK) FAIL("expected 6 actuators");
  if (m.neq != 0) FAIL("equality constraints unsupported");
  A.nobj = nobj; A.nv = m.nv; A.nq = m.nq; A.nab = ARM_NLINK + natt;
  // inertial bodies on the chain
  for (int a = 0; a < A.nab; a++) {
    int b = a < ARM_NLINK ? link_body[a] : att_body[a - ARM_NLINK];
    int link = a < ARM_NLINK ? a : slot_of[b] - SLOT_LINK0;
    if (link < 0 || link >= ARM_NLINK) FAIL("attachment %d not on the chain", b);
    A.ab_link[a] = link;
    double com[3], v[3];
    matvec(v, s_mat[b], m.body_ipos + 3 * b);
    for (int k = 0; k < 3; k++) com[k] = s_pos[b][k] + v[k];
    for (int k = 0; k < 3; k++) A.ab_com[a][k] = (real)com[k];
    double qi[4], Ri[9], I[9];
    mulquat(qi, s_quat[b], m.body_iquat + 4 * b);
    quat2mat(Ri, qi);
    const double* di = m.body_inertia + 3 * b;
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) I[3 * r + c] = Ri[3 * r] * di[0] * Ri[3 * c] + Ri[3 * r + 1] * di[1] * Ri[3 * c + 1] + Ri[3 * r + 2] * di[2] * Ri[3 * c + 2];
    A.ab_I[a][0] = (real)I[0]; A.ab_I[a][1] = (real)I[4]; A.ab_I[a][2] = (real)I[8];
    A.ab_I[a][3] = (real)I[1]; A.ab_I[a][4] = (real)I[2]; A.ab_I[a][5] = (real)I[5];
    A.ab_mass[a] = (real)m.body_mass[b];
    A.ab_gfac[a] = (real)(1.0 - m.body_gravcomp[b]);
  }
  for (int u = 0; u < ARM_NLINK; u++) {
    if (m.act_dof[u] != u || m.act_gear[u] != 1.0 || m.act_bias[3 * u] != 0) FAIL("actuator %d: expected a unit-gear position actuator on joint %d", u, u);
    A.kp[u] = m.act_gain[u]; A.kv[u] = -m.act_bias[3 * u + 2];
    if (m.act_bias[3 * u + 1] != -m.act_gain[u]) FAIL("actuator %d: biasprm[1] must equal -kp", u);
    A.ctrllimited[u] = m.act_ctrllimited[u]; A.forcelimited[u] = m.act_forcelimited[u];
    A.ctrl_lo[u] = m.act_ctrlrange[2 * u]; A.ctrl_hi[u] = m.act_ctrlrange[2 * u + 1];
    A.frc_lo[u] = m.act_forcerange[2 * u]; A.frc_hi[u] = m.act_forcerange[2 * u + 1];
  }
  // geoms
  if (m.ngeom > ARM_MAXGEOM) FAIL("too many geoms (%d)", m.ngeom);
  A.ngeom = m.ngeom;
  for (int g = 0; g < m.ngeom; g++) {
    int b = m.geom_body[g];
    A.geom_type[g] = m.geom_type[g]; A.geom_slot[g] = slot_of[b];
    double v[3], gm[9], R[9];
    matvec(v, s_mat[b], m.geom_pos + 3 * g);
    quat2mat(gm, m.geom_quat + 4 * g);
    matmul(R, s_mat[b], gm);
    for (int k = 0; k < 3; k++) { A.geom_pos[g][k] = (real)(s_pos[b][k] + v[k]); A.geom_size[g][k] = (real)m.geom_size[3 * g + k]; }
    for (int k = 0; k < 9; k++) A.geom_mat[g][k] = (real)R[k];
    A.geom_rbound[g] = (real)m.geom_rbound[g];
  }
  if (m.npair > ARM_MAXPAIR) FAIL("too many collision pairs (%d)", m.npair);
  A.npair = m.npair;
  for (int p = 0; p < m.npair; p++) {
    int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];
    A.pair_g1[p] = g1; A.pair_g2[p] = g2;
    double fr[3], sr[2], si[5];
    int dim;
    if (m.geom_priority[g1] != m.geom_priority[g2]) {
      int g = m.geom_priority[g1] > m.geom_priority[g2] ? g1 : g2;
      dim = m.geom_condim[g];
      for (int i = 0; i < 2; i++) sr[i] = m.geom_solref[2 * g + i];
      for (int i = 0; i < 5; i++) si[i] = m.geom_solimp[5 * g + i];
      for (int i = 0; i < 3; i++) fr[i] = m.geom_friction[3 * g + i];
    } else {
      dim = m.geom_condim[g1] > m.geom_condim[g2] ? m.geom_condim[g1] : m.geom_condim[g2];
      double x1 = m.geom_solmix[g1], x2 = m.geom_solmix[g2];
      double mix = (x1 >= MJS_MINVAL && x2 >= MJS_MINVAL) ? x1 / (x1 + x2) : ((x1 < MJS_MINVAL && x2 < MJS_MINVAL) ? 0.5 : (x1 < MJS_MINVAL ? 0.0 : 1.0));
      const double *r1 = m.geom_solref + 2 * g1, *r2 = m.geom_solref + 2 * g2;
      for (int i = 0; i < 2; i++) sr[i] = (r1[0] > 0 && r2[0] > 0) ? mix * r1[i] + (1 - mix) * r2[i] : std::fmin(r1[i], r2[i]);
      for (int i = 0; i < 5; i++) si[i] = mix * m.geom_solimp[5 * g1 + i] + (1 - mix) * m.geom_solimp[5 * g2 + i];
      for (int i = 0; i < 3; i++) fr[i] = std::fmax(m.geom_friction[3 * g1 + i], m.geom_friction[3 * g2 + i]);
    }
    if (dim != 1 && dim != 3 && dim != 4) FAIL("condim %d unsupported (1, 3, 4 only)", dim);
    A.pair_dim[p] = dim;
    for (int i = 0; i < 3; i++) A.pair_mu[p][i] = (real)std::fmax(MJS_MINMU, fr[i]);
    for (int i = 0; i < 2; i++) A.pair_solref[p][i] = (real)sr[i];
    for (int i = 0; i < 5; i++) A.pair_solimp[p][i] = (real)si[i];
    double margin = std::fmax(m.geom_margin[g1], m.geom_margin[g2]), gap = std::fmax(m.geom_gap[g1], m.geom_gap[g2]);
    A.pair_margin[p] = (real)margin; A.pair_incl[p] = (real)(margin - gap);
    int b1 = m.geom_body[g1], b2 = m.geom_body[g2];
    A.pair_tran[p] = (real)(m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2]);
    A.pair_rot[p] = (real)(m.body_invweight0[2 * b1 + 1] + m.body_invweight0[2 * b2 + 1]);
  }
  // sites
  {
    int sb = m.site_body[t.tcp_site];
    A.tcp_slot = slot_of[sb];
    double v[3], sm[9], R[9];
    matvec(v, s_mat[sb], m.site_pos + 3 * t.tcp_site);
    quat2mat(sm, m.site_quat + 4 * t.tcp_site);
    matmul(R, s_mat[sb], sm);
    for (int k = 0; k < 3; k++) { A.tcp_pos[k] = (real)(s_pos[sb][k] + v[k]); A.tcp_offset[k] = (real)t.tcp_offset[k]; }
    for (int k = 0; k < 9; k++) A.tcp_mat[k] = (real)R[k];
    A.touch_slot = -1;
    if (m.touch_site >= 0) {
      int tb = m.site_body[m.touch_site];
      A.touch_slot = slot_of[tb];
      matvec(v, s_mat[tb], m.site_pos + 3 * m.touch_site);
      quat2mat(sm, m.site_quat + 4 * m.touch_site);
      matmul(R, s_mat[tb], sm);
      for (int k = 0; k < 3; k++) A.touch_pos[k] = (real)(s_pos[tb][k] + v[k]);
      for (int k = 0; k < 9; k++) A.touch_mat[k] = (real)R[k];
      A.touch_size[0] = (real)m.touch_size[0]; A.touch_size[1] = (real)m.touch_size[1];
      for (int p = 0; p < m.npair; p++) {
        int s1 = slot_of[m.geom_body[m.pair_geom1[p]]], s2 = slot_of[m.geom_body[m.pair_geom2[p]]];
        if (s1 == A.touch_slot || s2 == A.touch_slot) A.touch_pair_mask |= (1 << p);
      }
    }
  }
  A.n_sub = t.n_sub; A.kind = t.kind; A.reward_type = t.reward_type; A.max_steps = t.max_steps;
  A.action_dim = t.action_dim; A.obs_dim = t.obs_dim; A.integrator = m.integrator; A.max_iter = m.solver_iterations;
  A.timestep = (real)m.timestep; A.impratio = (real)m.impratio;
  for (int j = 0; j < ARM_NLINK; j++) mjs_kb<real>(A.jnt_solref[j], A.jnt_solimp[j], A.timestep, &A.jnt_kb[j][0], &A.jnt_kb[j][1]);
  for (int q = 0; q < A.npair; q++) mjs_kb<real>(A.pair_solref[q], A.pair_solimp[q], A.timestep, &A.pair_kb[q][0], &A.pair_kb[q][1]);
  A.scale = (real)(1.0 / (m.meaninertia * (m.nv > 1 ? m.nv : 1)));
  // mjOption.tolerance as given (MuJoCo default 1e-8; parity tests pass 1e-12 to both sides); FP32 cannot resolve below 1e-6
  A.tol = precision_f64 ? (real)m.tolerance : (real)(m.tolerance > 1e-6 ? m.tolerance : 1e-6);
  for (int k = 0; k < 3; k++) A.gravity[k] = (real)m.gravity[k];
  A.control_dt = t.control_dt; A.time_limit = t.time_limit; A.goal_threshold = t.goal_threshold; A.target_radius = t.target_radius;
  A.nearest_coef = t.nearest_coef; A.tcp_z = t.tcp_z; A.obj_radius = t.obj_radius;
  for (int k = 0; k < 3; k++) {
    A.spawn_lo[k] = t.spawn_lo[k]; A.spawn_hi[k] = t.spawn_hi[k]; A.target_lo[k] = t.target_lo[k]; A.target_hi[k] = t.target_hi[k];
    A.obj_lo[k] = t.obj_lo[k]; A.obj_hi[k] = t.obj_hi[k]; A.button_goal[k] = t.button_goal[k];
  }
  for (int k = 0; k < 8; k++) { A.action_lo[k] = t.action_lo[k]; A.action_hi[k] = t.action_hi[k]; }
  for (int k = 0; k < 6; k++) A.ik_dh[k] = t.ik_dh[k];
  A.button_goal_radius = t.button_goal_radius; A.touch_lo = t.touch_lo; A.touch_hi = t.touch_hi; A.settle_steps = t.settle_steps;
  A.button_disturbances = t.button_disturbances; A.disturb_prob = t.disturb_prob;
#undef FAIL
  return 0;
}
